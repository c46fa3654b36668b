//! `World` → the flat tables of `rtc_scene_desc` (include/rtc.h). Rust twin of the tested C++ flattener
//! (`host/rth_flatten.hpp` of the backend repository); NOT COMPILED in that repository's image (no rustc there).
//!
//! Pre-order walk of `World::objects()` in insertion order (world.rs:19): a `Group` is followed by its children in
//! order (group.rs:17), a `Csg` by its left subtree then its right subtree (csg.rs:47-48); every node carries a skip
//! index so the device walk needs no stack. Group transforms are already baked into the leaves by the reference
//! itself (`Group::add_child` / `set_transform`, group.rs:71-94,128-133) and `Group::divide` has already run, so the
//! walk only reads: `as_any()` downcasts for the kind, `get_base().transform_inverse`, `material()`, `get_bounds()`,
//! `has_shadow()`, and the accessors added by patches/raytracer-accessors.diff.
//!
//! Transforms, patterns and materials are de-duplicated by exact bit pattern. Value-equality classes (`impl
//! PartialEq for dyn Shape`, geometry/mod.rs:110-114, used by the containers walk of intersection.rs:63-90) are
//! assigned by exact-bit keys over everything the derived `PartialEq` compares; shapes equal only within the
//! reference's 1e-5 epsilon get distinct classes (documented deviation).
use std::collections::HashMap;

use anyhow::{bail, Result};
use raytracer::{
    geometry::{
        shape::{Cone, Csg, Cube, Cylinder, Group, Operation, Plane, SmoothTriangle, Sphere, Triangle},
        Shape,
    },
    material::Material,
    matrix::Matrix,
    point::Point,
    vector::cross,
    world::World,
};

use crate::ffi::*;

#[derive(Default)]
pub struct FlatWorld {
    pub nodes: Vec<rtc_node>,
    pub xforms: Vec<f64>,      // 12 per transform: rows 0..2 of transform_inverse
    pub boxes: Vec<f64>,       // 6 per Group / Csg
    pub limits: Vec<f64>,      // 2 per Cylinder / Cone
    pub tri_p1e1e2: Vec<f64>,  // 9 per triangle
    pub tri_normals: Vec<f64>, // 9 per triangle
    pub materials: Vec<rtc_material>,
    pub patterns: Vec<rtc_pattern>,
    pub lights: Vec<rtc_light>,
}

impl FlatWorld {
    /// The descriptor borrows `self`: keep the `FlatWorld` alive until `rtc_scene_create` has returned.
    pub fn desc(&self) -> rtc_scene_desc {
        rtc_scene_desc {
            abi_version: RTC_ABI_VERSION,
            n_nodes: self.nodes.len() as u32,
            n_xforms: (self.xforms.len() / 12) as u32,
            n_boxes: (self.boxes.len() / 6) as u32,
            n_limits: (self.limits.len() / 2) as u32,
            n_triangles: (self.tri_p1e1e2.len() / 9) as u32,
            n_materials: self.materials.len() as u32,
            n_patterns: self.patterns.len() as u32,
            n_lights: self.lights.len() as u32,
            nodes: self.nodes.as_ptr(),
            xforms: self.xforms.as_ptr(),
            boxes: self.boxes.as_ptr(),
            limits: self.limits.as_ptr(),
            tri_p1e1e2: self.tri_p1e1e2.as_ptr(),
            tri_normals: self.tri_normals.as_ptr(),
            materials: self.materials.as_ptr(),
            patterns: self.patterns.as_ptr(),
            lights: self.lights.as_ptr(),
        }
    }
}

fn rows3(m: &Matrix) -> [f64; 12] {
    let mut r = [0.0; 12];
    for i in 0..3 {
        for j in 0..4 {
            r[4 * i + j] = m[(i, j)];
        }
    }
    r
}

fn bits(values: &[f64]) -> Vec<u64> {
    values.iter().map(|v| v.to_bits()).collect()
}

#[derive(Default)]
struct Flattener {
    out: FlatWorld,
    xf_ids: HashMap<Vec<u64>, i32>,
    pat_ids: HashMap<Vec<u64>, i32>,
    mat_ids: HashMap<Vec<u64>, i32>,
    classes: HashMap<Vec<u64>, i32>,
}

impl Flattener {
    fn xform_id(&mut self, inverse: &Matrix) -> i32 {
        let r = rows3(inverse);
        let key = bits(&r);
        if let Some(&id) = self.xf_ids.get(&key) {
            return id;
        }
        let id = (self.out.xforms.len() / 12) as i32;
        self.out.xforms.extend_from_slice(&r);
        self.xf_ids.insert(key, id);
        id
    }

    fn pattern_id(&mut self, m: &Material) -> i32 {
        let p = match m.pattern() {
            None => return -1,
            Some(p) => p,
        };
        let (kind, a, b) = p.kind_and_colors();
        let rp = rtc_pattern {
            kind: kind as i32,
            _pad: 0,
            a: [a.red, a.green, a.blue],
            b: [b.red, b.green, b.blue],
            inverse: rows3(p.transform_inverse()),
        };
        let mut key = vec![kind as u64];
        key.extend(bits(&rp.a));
        key.extend(bits(&rp.b));
        key.extend(bits(&rp.inverse));
        if let Some(&id) = self.pat_ids.get(&key) {
            return id;
        }
        let id = self.out.patterns.len() as i32;
        self.out.patterns.push(rp);
        self.pat_ids.insert(key, id);
        id
    }

    fn material_id(&mut self, m: &Material) -> i32 {
        let pattern = self.pattern_id(m);
        let rm = rtc_material {
            color: [m.color.red, m.color.green, m.color.blue],
            ambient: m.ambient,
            diffuse: m.diffuse,
            specular: m.specular,
            shininess: m.shininess,
            reflective: m.reflective,
            transparency: m.transparency,
            refractive_index: m.refractive_index,
            pattern,
            _pad: 0,
        };
        let mut key = bits(&rm.color);
        key.extend(bits(&[
            rm.ambient,
            rm.diffuse,
            rm.specular,
            rm.shininess,
            rm.reflective,
            rm.transparency,
            rm.refractive_index,
        ]));
        key.push(pattern as u32 as u64);
        if let Some(&id) = self.mat_ids.get(&key) {
            return id;
        }
        let id = self.out.materials.len() as i32;
        self.out.materials.push(rm);
        self.mat_ids.insert(key, id);
        id
    }

    fn box_id(&mut self, s: &dyn Shape) -> i32 {
        let (lo, hi) = (s.get_bounds().get_min(), s.get_bounds().get_max());
        let id = (self.out.boxes.len() / 6) as i32;
        self.out.boxes.extend_from_slice(&[lo.x, lo.y, lo.z, hi.x, hi.y, hi.z]);
        id
    }

    /// everything `#[derive(PartialEq)]` on the shape structs compares (BaseShape: transform, inverse, material,
    /// bounding box, shadow; plus the kind's own fields), by exact bits
    fn class_id(&mut self, s: &dyn Shape, kind: i32, xform: i32, material: i32, extra: &[f64]) -> i32 {
        let (lo, hi) = (s.get_bounds().get_min(), s.get_bounds().get_max());
        let mut key = vec![kind as u64, xform as u64, material as u64, s.has_shadow() as u64];
        key.extend(bits(&rows3(s.transform())));
        key.extend(bits(&[lo.x, lo.y, lo.z, hi.x, hi.y, hi.z]));
        key.extend(bits(extra));
        let next = self.classes.len() as i32;
        *self.classes.entry(key).or_insert(next)
    }

    fn leaf(&mut self, s: &dyn Shape, kind: i32, aux: i32, closed: bool, extra: &[f64]) {
        let xform = self.xform_id(&s.get_base().transform_inverse);
        let material = self.material_id(s.material());
        let mut flags = 0u32;
        if s.has_shadow() {
            flags |= RTC_FLAG_SHADOW;
        }
        if closed {
            flags |= RTC_FLAG_CLOSED;
        }
        let eq_class = self.class_id(s, kind, xform, material, extra);
        let idx = self.out.nodes.len() as i32;
        self.out.nodes.push(rtc_node { kind, skip: idx + 1, xform, material, aux, flags, right: -1, eq_class });
    }

    fn triangle(&mut self, p: [Point; 3], normals: [f64; 9]) -> (i32, Vec<f64>) {
        // triangle.rs:31-44 / smooth_triangle.rs:34-50: e1 = p2 - p1, e2 = p3 - p1 (private fields, recomputed with
        // the constructor's own expressions: bit-identical)
        let (e1, e2) = (p[1] - p[0], p[2] - p[0]);
        let g = [p[0].x, p[0].y, p[0].z, e1.x, e1.y, e1.z, e2.x, e2.y, e2.z];
        let aux = (self.out.tri_p1e1e2.len() / 9) as i32;
        self.out.tri_p1e1e2.extend_from_slice(&g);
        self.out.tri_normals.extend_from_slice(&normals);
        let mut extra = g.to_vec();
        extra.extend_from_slice(&normals);
        (aux, extra)
    }

    fn emit(&mut self, s: &dyn Shape) -> Result<()> {
        let any = s.as_any();
        if any.is::<Sphere>() {
            self.leaf(s, RTC_SPHERE, -1, false, &[]);
        } else if any.is::<Plane>() {
            self.leaf(s, RTC_PLANE, -1, false, &[]);
        } else if any.is::<Cube>() {
            self.leaf(s, RTC_CUBE, -1, false, &[]);
        } else if let Some(c) = any.downcast_ref::<Cylinder>() {
            let aux = (self.out.limits.len() / 2) as i32;
            self.out.limits.extend_from_slice(&[c.minimum(), c.maximum()]);
            self.leaf(s, RTC_CYLINDER, aux, c.closed(), &[c.minimum(), c.maximum(), c.closed() as u8 as f64]);
        } else if let Some(c) = any.downcast_ref::<Cone>() {
            let aux = (self.out.limits.len() / 2) as i32;
            self.out.limits.extend_from_slice(&[c.minimum(), c.maximum()]);
            self.leaf(s, RTC_CONE, aux, c.closed(), &[c.minimum(), c.maximum(), c.closed() as u8 as f64]);
        } else if let Some(t) = any.downcast_ref::<Triangle>() {
            let n = cross(t.p3 - t.p1, t.p2 - t.p1).normalize(); // triangle.rs:43: cross(e2, e1).normalize()
            let (aux, extra) = self.triangle([t.p1, t.p2, t.p3], [n.x, n.y, n.z, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0]);
            self.leaf(s, RTC_TRIANGLE, aux, false, &extra);
        } else if let Some(t) = any.downcast_ref::<SmoothTriangle>() {
            let normals = [t.n1.x, t.n1.y, t.n1.z, t.n2.x, t.n2.y, t.n2.z, t.n3.x, t.n3.y, t.n3.z];
            let (aux, extra) = self.triangle([t.p1, t.p2, t.p3], normals);
            self.leaf(s, RTC_SMOOTH_TRIANGLE, aux, false, &extra);
        } else if let Some(g) = any.downcast_ref::<Group>() {
            let idx = self.out.nodes.len();
            let aux = self.box_id(s);
            self.out.nodes.push(rtc_node { kind: RTC_GROUP, skip: 0, xform: -1, material: -1, aux, flags: 0, right: -1, eq_class: -1 });
            for child in &g.children {
                self.emit(child.as_ref())?;
            }
            self.out.nodes[idx].skip = self.out.nodes.len() as i32;
        } else if let Some(c) = any.downcast_ref::<Csg>() {
            let idx = self.out.nodes.len();
            let aux = self.box_id(s);
            let xform = self.xform_id(&s.get_base().transform_inverse);
            let op = match c.operation() {
                Operation::Union => RTC_CSG_UNION,
                Operation::Intersection => RTC_CSG_INTERSECTION,
                Operation::Difference => RTC_CSG_DIFFERENCE,
            };
            self.out.nodes.push(rtc_node { kind: RTC_CSG, skip: 0, xform, material: -1, aux, flags: op << 8, right: -1, eq_class: -1 });
            self.emit(c.left.as_ref())?;
            self.out.nodes[idx].right = self.out.nodes.len() as i32;
            self.emit(c.right.as_ref())?;
            self.out.nodes[idx].skip = self.out.nodes.len() as i32;
        } else {
            bail!("raytracer-cuda: unsupported Shape implementation {:?} (TestShape and foreign shapes have no device kernel)", s);
        }
        Ok(())
    }
}

/// Flatten a `World` for `rtc_scene_create`.
pub fn flatten(world: &World) -> Result<FlatWorld> {
    let mut f = Flattener::default();
    for l in world.lights() {
        let (p, i) = (l.position(), l.intensity());
        f.out.lights.push(rtc_light { position: [p.x, p.y, p.z], intensity: [i.red, i.green, i.blue] });
    }
    for o in world.objects() {
        f.emit(o.as_ref())?;
    }
    Ok(f.out)
}
