//! raytracer-cuda — drop-in B200 (sm_100a) backend for `Camera::render(&World) -> Canvas` (raytracer/src/camera.rs:133)
//! and `Camera::render_multithreaded` (:150). NOT COMPILED in the backend repository's build image (no rustc/cargo
//! there): shipped as source; the same C ABI is exercised there through ctypes and through a C++ host library.
//!
//! ```ignore
//! use raytracer_cuda::RenderCuda;
//! let canvas = camera.render_cuda(&world)?;          // instead of camera.render(&world)
//! ```
//! `scene-parser` switches with `--features cuda` (patches/raytracer-accessors.diff, scene-parser/src/lib.rs:278).
//!
//! The frame is the CPU render's: f64, the reference's operation order, no FMA contraction; there is no CPU
//! fallback — without a CUDA device every call returns an error.
pub mod ffi;
pub mod flatten;

use std::ffi::CStr;

use anyhow::{anyhow, Result};
use raytracer::{camera::Camera, canvas::Canvas, color::Color, world::World};

pub use flatten::{flatten, FlatWorld};

fn check(rc: i32) -> Result<i32> {
    if rc < 0 {
        let msg = unsafe { CStr::from_ptr(ffi::rtc_last_error()) }.to_string_lossy().into_owned();
        return Err(anyhow!("rtc error {}: {}", rc, msg));
    }
    Ok(rc)
}

/// A `World` resident on one GPU (tables + acceleration structure). Re-usable for any number of cameras / frames.
pub struct CudaScene {
    handle: *mut ffi::rtc_scene,
}

// one render per scene handle at a time (include/rtc.h); distinct handles are independent
unsafe impl Send for CudaScene {}

impl CudaScene {
    pub fn new(world: &World, device: i32) -> Result<Self> {
        let flat = flatten(world)?;
        Self::from_flat(&flat, device)
    }

    pub fn from_flat(flat: &FlatWorld, device: i32) -> Result<Self> {
        let desc = flat.desc();
        let mut handle = std::ptr::null_mut();
        check(unsafe { ffi::rtc_scene_create(&desc, device, &mut handle) })?;
        Ok(Self { handle })
    }

    /// `Camera::render` (aa == 1) / `render_multithreaded` (the camera's `render_opts.aa_samples`): row-major
    /// `y * hsize + x` doubles, the layout of `Canvas.pixels` (canvas.rs:44-48).
    pub fn render_rgb(&mut self, camera: &Camera, opts: &ffi::rtc_opts) -> Result<(Vec<f64>, ffi::rtc_stats)> {
        let cam = camera_desc(camera);
        let mut rgb = vec![0.0f64; camera.hsize() * camera.vsize() * 3];
        let mut stats = ffi::rtc_stats::default();
        check(unsafe { ffi::rtc_render(self.handle, &cam, opts, rgb.as_mut_ptr(), &mut stats) })?;
        Ok((rgb, stats))
    }

    /// 8-bit image as `image/png.rs:13-31` / `ppm.rs:73-75` quantise it, 3 bytes per pixel off the device.
    pub fn render_rgb8(&mut self, camera: &Camera, opts: &ffi::rtc_opts) -> Result<Vec<u8>> {
        let cam = camera_desc(camera);
        let mut rgb8 = vec![0u8; camera.hsize() * camera.vsize() * 3];
        check(unsafe { ffi::rtc_render_rgb8(self.handle, &cam, opts, rgb8.as_mut_ptr(), std::ptr::null_mut()) })?;
        Ok(rgb8)
    }

    pub fn render(&mut self, camera: &Camera) -> Result<Canvas> {
        let aa = camera.render_opts.get_aa_samples() as u32;
        // render_multithreaded averages even a single sample: (0 + c) * (1.0 / 1) (camera.rs:186, color.rs:26-33)
        let threaded = camera.render_opts.get_num_threads() > 1 || aa > 1;
        let opts = ffi::rtc_opts {
            aa_samples: aa,
            max_depth: raytracer::world::MAX_RECURSION_DEPTH as u32,
            flags: if threaded { ffi::RTC_OPT_AVERAGE } else { 0 },
            ..Default::default()
        };
        let (rgb, _) = self.render_rgb(camera, &opts)?;
        let (w, h) = (camera.hsize(), camera.vsize());
        let mut canvas = Canvas::new(w, h);
        for y in 0..h {
            for x in 0..w {
                let i = (y * w + x) * 3;
                canvas.set_pixel(x, y, Color::new(rgb[i], rgb[i + 1], rgb[i + 2]));
            }
        }
        Ok(canvas)
    }
}

impl Drop for CudaScene {
    fn drop(&mut self) {
        unsafe { ffi::rtc_scene_destroy(self.handle) }
    }
}

pub fn camera_desc(camera: &Camera) -> ffi::rtc_camera {
    let inv = camera.transform_inverse();
    let mut inverse = [0.0f64; 16];
    for r in 0..4 {
        for c in 0..4 {
            inverse[4 * r + c] = inv[(r, c)];
        }
    }
    ffi::rtc_camera {
        hsize: camera.hsize() as u32,
        vsize: camera.vsize() as u32,
        pixel_size: camera.pixel_size(),
        half_width: camera.half_width(),
        half_height: camera.half_height(),
        inverse,
    }
}

/// `camera.render(&world)` on GPU 0.
pub fn render(camera: &Camera, world: &World) -> Result<Canvas> {
    CudaScene::new(world, 0)?.render(camera)
}

pub trait RenderCuda {
    fn render_cuda(&self, world: &World) -> Result<Canvas>;
}

impl RenderCuda for Camera {
    fn render_cuda(&self, world: &World) -> Result<Canvas> {
        render(self, world)
    }
}

pub fn device_count() -> i32 {
    unsafe { ffi::rtc_device_count() }
}
