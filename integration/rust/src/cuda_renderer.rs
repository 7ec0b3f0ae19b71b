// cuda_renderer.rs — `impl Renderer for CudaRenderer`: the B200 backend behind the reference's render boundary
// (trait Renderer, src/renderer.rs:38-40).  It stands where TileRenderer stands (src/renderer.rs:232-240, :244-387): one
// synchronous `render` that owns all of its parallelism — TileRenderer spawns and joins worker threads (:272-296, :367-369),
// this one drives every GPU of the box through ONE FFI call (rtb200_multi_render, include/rtb200.h).
//
// The call sequence of `render` below — multi_create, multi_upload_scene, multi_render, multi_destroy — is replayed by a
// plain-C program, tests/abi_c_caller.c, which IS compiled and run on the GPU box (tests/test_c_caller.py); this file cannot
// be (no Rust toolchain there).
use std::ffi::CStr;
use std::sync::Arc;

use image::RgbImage;

use crate::objects::Object;
use crate::ray_tracing::{Camera, Scene};
use crate::renderer::{RenderStat, Renderer};
use crate::rtb200_sys::*;

pub struct CudaRenderer {
    multi: *mut rtb200_multi,
    seed: u64,
    row_groups: i32,
}

// The context is only touched from `render`, which the reference calls from one thread (src/main.rs:64).
unsafe impl Send for CudaRenderer {}

fn check(rc: i32) {
    // the reference panics on every error (src/main.rs:33,38; src/renderer.rs:264,288,292): so does its CUDA renderer
    if rc != RTB200_OK {
        let msg = unsafe { CStr::from_ptr(rtb200_last_error()) }.to_string_lossy().into_owned();
        panic!("rtb200 error {}: {}", rc, msg);
    }
}

impl CudaRenderer {
    /// `devices`: the CUDA devices to shard over (None = device 0 .. count-1 as far as `RTB200_GPUS`, default 1, says).
    /// `seed`: Philox key of the deterministic sample stream (the reference seeds from the OS, src/ray_tracing.rs:78).
    pub fn new(devices: Option<&[i32]>, seed: u64) -> Box<dyn Renderer> {
        let mut multi = std::ptr::null_mut();
        match devices {
            Some(d) => check(unsafe { rtb200_multi_create(d.as_ptr(), d.len() as i32, &mut multi) }),
            None => {
                let n: i32 = std::env::var("RTB200_GPUS").ok().and_then(|v| v.parse().ok()).unwrap_or(1);
                check(unsafe { rtb200_multi_create(std::ptr::null(), n, &mut multi) })
            }
        }
        Box::new(CudaRenderer { multi, seed, row_groups: 1 })   // 1 = sample-range sharding (perfectly balanced)
    }
}

impl Drop for CudaRenderer {
    fn drop(&mut self) {
        unsafe { rtb200_multi_destroy(self.multi) }
    }
}

impl Renderer for CudaRenderer {
    fn render(&self, max_bounces: usize, samples_per_pixel: usize, scene: &Arc<Scene>, camera: &Arc<Camera>) -> (RgbImage, RenderStat) {
        // Flatten Vec<Object> in list order (the order is significant: the last object wins closest-hit ties,
        // src/objects.rs:141).  One material record per object keeps the Arc<dyn Material> association without
        // pointer-identity games.
        let mut materials: Vec<rtb200_material> = Vec::with_capacity(scene.len());
        let spheres: Vec<rtb200_sphere> = scene
            .objects()
            .iter()
            .map(|o| match o {
                Object::Sphere(s) => {
                    materials.push(s.material().gpu_desc());
                    let c = s.center();
                    rtb200_sphere { center: [c.x(), c.y(), c.z()], radius: s.radius(), material: (materials.len() - 1) as i32, _pad: 0 }
                }
            })
            .collect();
        let sc = rtb200_scene {
            spheres: spheres.as_ptr(),
            n_spheres: spheres.len() as i32,
            materials: materials.as_ptr(),
            n_materials: materials.len() as i32,
        };
        check(unsafe { rtb200_multi_upload_scene(self.multi, &sc) });

        let cam = camera.gpu_desc();
        let params = rtb200_params {
            max_bounces: max_bounces as i32,
            samples_per_pixel: samples_per_pixel as i32,
            seed: self.seed,
            sample_begin: 0,
            sample_end: samples_per_pixel as i32,
            row_begin: 0,
            row_end: cam.image_height,
            flags: RTB200_FLAG_NONE,
            row_stride: 1,
        };
        let (w, h) = (cam.image_width as u32, cam.image_height as u32);
        let mut rgb = vec![0u8; (w as usize) * (h as usize) * 3];
        let mut st = rtb200_stats {
            seconds: 0.0, pixels_rendered: 0, pixels_per_second: 0.0, paths: 0, segments: 0, sphere_tests: 0,
            trace_ms: 0.0, resolve_ms: 0.0, h2d_ms: 0.0, d2h_ms: 0.0,
        };
        check(unsafe { rtb200_multi_render(self.multi, &cam, &params, self.row_groups, rgb.as_mut_ptr(), std::ptr::null_mut(), &mut st) });
        // RgbImage is row-major, top row first, 3 bytes per pixel: exactly what the resolve kernel wrote (src/renderer.rs:376-384)
        (RgbImage::from_raw(w, h, rgb).unwrap(), RenderStat::new(std::time::Duration::from_secs_f64(st.seconds), st.pixels_rendered as usize))
    }
}
