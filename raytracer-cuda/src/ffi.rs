//! `extern "C"` bindings to `include/rtc.h` (RTC_ABI_VERSION 1), one to one: every struct in declaration order with
//! the header's field names, every exported function. tests/test_rust_shim.py of the backend repository parses this
//! file and the header and checks the export list, the field order and the struct sizes against each other.
#![allow(non_camel_case_types, dead_code)]
use std::os::raw::{c_char, c_int, c_uchar, c_void};

pub const RTC_ABI_VERSION: u32 = 1;

// rtc_status
pub const RTC_OK: c_int = 0;
pub const RTC_ERR_INVALID: c_int = -1;
pub const RTC_ERR_CUDA: c_int = -2;
pub const RTC_ERR_OVERFLOW: c_int = -3;
pub const RTC_ERR_NAN: c_int = -4;
pub const RTC_ERR_UNSUPPORTED: c_int = -5;

// rtc_kind
pub const RTC_SPHERE: i32 = 0;
pub const RTC_PLANE: i32 = 1;
pub const RTC_CUBE: i32 = 2;
pub const RTC_CYLINDER: i32 = 3;
pub const RTC_CONE: i32 = 4;
pub const RTC_TRIANGLE: i32 = 5;
pub const RTC_SMOOTH_TRIANGLE: i32 = 6;
pub const RTC_GROUP: i32 = 7;
pub const RTC_CSG: i32 = 8;

// rtc_pattern_kind
pub const RTC_PAT_TEST: i32 = 0;
pub const RTC_PAT_STRIPE: i32 = 1;
pub const RTC_PAT_GRADIENT: i32 = 2;
pub const RTC_PAT_RING: i32 = 3;
pub const RTC_PAT_CHECKERS: i32 = 4;

// rtc_csg_op
pub const RTC_CSG_UNION: u32 = 0;
pub const RTC_CSG_INTERSECTION: u32 = 1;
pub const RTC_CSG_DIFFERENCE: u32 = 2;

pub const RTC_FLAG_SHADOW: u32 = 1;
pub const RTC_FLAG_CLOSED: u32 = 2;

pub const RTC_OPT_FAITHFUL_TRAVERSAL: u32 = 1;
pub const RTC_OPT_AVERAGE: u32 = 2;
pub const RTC_OPT_WAVEFRONT: u32 = 4;
pub const RTC_OPT_RECURSIVE: u32 = 8;
pub const RTC_OPT_WAVE_STREAM: u32 = 16;
pub const RTC_STAT_WAVEFRONT: u32 = 1;
pub const RTC_STAT_WAVE_FALLBACK: u32 = 2;
pub const RTC_STAT_WAVE_OVERFLOWED: u32 = 4;

pub const RTC_TRACE_STATIC_MAX: usize = 4096;

#[repr(C)]
#[derive(Debug, Clone, Copy)]
pub struct rtc_node {
    pub kind: i32,
    pub skip: i32,
    pub xform: i32,
    pub material: i32,
    pub aux: i32,
    pub flags: u32,
    pub right: i32,
    pub eq_class: i32,
}

#[repr(C)]
#[derive(Debug, Clone, Copy)]
pub struct rtc_material {
    pub color: [f64; 3],
    pub ambient: f64,
    pub diffuse: f64,
    pub specular: f64,
    pub shininess: f64,
    pub reflective: f64,
    pub transparency: f64,
    pub refractive_index: f64,
    pub pattern: i32,
    pub _pad: i32,
}

#[repr(C)]
#[derive(Debug, Clone, Copy)]
pub struct rtc_pattern {
    pub kind: i32,
    pub _pad: i32,
    pub a: [f64; 3],
    pub b: [f64; 3],
    pub inverse: [f64; 12],
}

#[repr(C)]
#[derive(Debug, Clone, Copy)]
pub struct rtc_light {
    pub position: [f64; 3],
    pub intensity: [f64; 3],
}

#[repr(C)]
#[derive(Debug, Clone, Copy)]
pub struct rtc_scene_desc {
    pub abi_version: u32,
    pub n_nodes: u32,
    pub n_xforms: u32,
    pub n_boxes: u32,
    pub n_limits: u32,
    pub n_triangles: u32,
    pub n_materials: u32,
    pub n_patterns: u32,
    pub n_lights: u32,
    pub nodes: *const rtc_node,
    pub xforms: *const f64,
    pub boxes: *const f64,
    pub limits: *const f64,
    pub tri_p1e1e2: *const f64,
    pub tri_normals: *const f64,
    pub materials: *const rtc_material,
    pub patterns: *const rtc_pattern,
    pub lights: *const rtc_light,
}

#[repr(C)]
#[derive(Debug, Clone, Copy)]
pub struct rtc_camera {
    pub hsize: u32,
    pub vsize: u32,
    pub pixel_size: f64,
    pub half_width: f64,
    pub half_height: f64,
    pub inverse: [f64; 16],
}

#[repr(C)]
#[derive(Debug, Clone, Copy, Default)]
pub struct rtc_opts {
    pub aa_samples: u32,
    pub max_depth: u32,
    pub row_begin: u32,
    pub row_end: u32,
    pub tile_rows: u32,
    pub rank: u32,
    pub nranks: u32,
    pub flags: u32,
}

#[repr(C)]
#[derive(Debug, Clone, Copy, Default)]
pub struct rtc_stats {
    pub rays_primary: u64,
    pub rays_shadow: u64,
    pub rays_reflect: u64,
    pub rays_refract: u64,
    pub overflow_count: u64,
    pub nan_count: u64,
    pub kernel_ms: f64,
    pub total_ms: f64,
    pub launches: u32,
    pub flags: u32,
}

/// Opaque handles.
#[repr(C)]
pub struct rtc_scene {
    _private: [u8; 0],
}
#[repr(C)]
pub struct rtc_canvas {
    _private: [u8; 0],
}
#[repr(C)]
pub struct rtc_host_canvas {
    _private: [u8; 0],
}

extern "C" {
    pub fn rtc_last_error() -> *const c_char;
    pub fn rtc_abi_version() -> c_int;
    pub fn rtc_device_count() -> c_int;
    pub fn rtc_scene_create(desc: *const rtc_scene_desc, device: c_int, out: *mut *mut rtc_scene) -> c_int;
    pub fn rtc_scene_destroy(scene: *mut rtc_scene);
    pub fn rtc_render(scene: *mut rtc_scene, camera: *const rtc_camera, opts: *const rtc_opts, rgb_host: *mut f64, stats: *mut rtc_stats) -> c_int;
    pub fn rtc_render_device(scene: *mut rtc_scene, camera: *const rtc_camera, opts: *const rtc_opts, rgb_dev: *mut f64, stream: *mut c_void, stats: *mut rtc_stats) -> c_int;
    pub fn rtc_scene_finish(scene: *mut rtc_scene, stream: *mut c_void, stats: *mut rtc_stats) -> c_int;
    pub fn rtc_quantize_device(device: c_int, rgb_dev: *const f64, n: usize, rgb8_dev: *mut u8, stream: *mut c_void) -> c_int;
    pub fn rtc_render_rgb8(scene: *mut rtc_scene, camera: *const rtc_camera, opts: *const rtc_opts, rgb8_host: *mut u8, stats: *mut rtc_stats) -> c_int;
    pub fn rtc_trace_rays(scene: *mut rtc_scene, n: usize, origins: *const f64, directions: *const f64, remaining: u32, rgb_out: *mut f64, t_out: *mut f64) -> c_int;
    pub fn rtc_trace_rays_static(scene: *mut rtc_scene, n: usize, origins: *const f64, directions: *const f64, remaining: u32, rgb_out: *mut f64, t_out: *mut f64) -> c_int;
    pub fn rtc_probe_hit(scene: *mut rtc_scene, origin: *const f64, direction: *const f64, out20: *mut f64) -> c_int;
    pub fn rtc_is_shadowed(scene: *mut rtc_scene, point: *const f64, light_index: u32) -> c_int;
    pub fn rtc_canvas_create(device: c_int, n_doubles: usize, out: *mut *mut rtc_canvas) -> c_int;
    pub fn rtc_canvas_export(canvas: *mut rtc_canvas, handle: *mut c_uchar) -> c_int;
    pub fn rtc_canvas_open(device: c_int, handle: *const c_uchar, n_doubles: usize, out: *mut *mut rtc_canvas) -> c_int;
    pub fn rtc_canvas_ptr(canvas: *mut rtc_canvas) -> *mut f64;
    pub fn rtc_canvas_download(canvas: *mut rtc_canvas, host: *mut f64, offset_doubles: usize, n_doubles: usize) -> c_int;
    pub fn rtc_canvas_push_tiles(canvas: *mut rtc_canvas, local_dev: *const f64, hsize: u32, vsize: u32, tile_rows: u32, rank: u32, nranks: u32, stream: *mut c_void) -> c_int;
    pub fn rtc_canvas_destroy(canvas: *mut rtc_canvas);
    pub fn rtc_host_canvas_create(name: *const c_char, n_doubles: usize, device: c_int, create: c_int, out: *mut *mut rtc_host_canvas) -> c_int;
    pub fn rtc_host_canvas_place(canvas: *mut rtc_host_canvas, hsize: u32, vsize: u32, tile_rows: u32, rank: u32, nranks: u32) -> c_int;
    pub fn rtc_host_canvas_ptr(canvas: *mut rtc_host_canvas) -> *mut f64;
    pub fn rtc_host_canvas_arrive(canvas: *mut rtc_host_canvas, rank: u32, frame: u32) -> c_int;
    pub fn rtc_host_canvas_wait(canvas: *mut rtc_host_canvas, nranks: u32, frame: u32, timeout_s: f64) -> c_int;
    pub fn rtc_host_canvas_destroy(canvas: *mut rtc_host_canvas);
    pub fn rtc_debug_normalize(device: c_int, n: usize, xyz_host: *const f64, out_host: *mut f64) -> c_int;
    pub fn rtc_measure_fp64(device: c_int, fused: c_int, gflops_out: *mut f64) -> c_int;
}
