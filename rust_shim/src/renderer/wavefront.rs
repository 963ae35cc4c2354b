// src/renderer/wavefront.rs — the `Renderer` variant backed by the CUDA core (UNCOMPILED HERE).
//
// Wiring in src/renderer/mod.rs:
//   :21-24  add `WavefrontPathTracer` to the enum_dispatch `Renderer` enum
//   :32-35  add the arm   "pt_wavefront" => WavefrontPathTracer::new(max_depth, pixel_sampler, filter).into(),
use std::ffi::CStr;

use super::ffi::*;
use crate::{
    core::scene::Scene,
    filter::Filter,
    pixel_sampler::PixelSampler,
    renderer::{OutputConfig, RendererT},
};

pub struct WavefrontPathTracer {
    max_depth: u32,
    pixel_sampler: PixelSampler,
    filter: Filter,
}

impl WavefrontPathTracer {
    pub fn new(max_depth: u32, pixel_sampler: PixelSampler, filter: Filter) -> Self {
        Self { max_depth, pixel_sampler, filter }
    }
}

/// The reference panics / unwraps on the render path (pt.rs:287, scene.rs:29-41); so does the shim.
fn check(status: i32) {
    if status != SPT_OK {
        let msg = unsafe { CStr::from_ptr(spt_last_error()) }.to_string_lossy().into_owned();
        panic!("spt: {}", msg);
    }
}

impl RendererT for WavefrontPathTracer {
    fn render(&self, scene: &Scene, config: &OutputConfig) {
        let flat = scene.flatten(); // src/core/flatten.rs — owns every array `flat.desc()` points into
        let (kind, spp, div_x, div_y) = match &self.pixel_sampler {
            PixelSampler::RandomSampler(s) => (SPT_SAMPLER_RANDOM, s.spp() as i32, 1, 1),
            PixelSampler::JitteredSampler(s) => (SPT_SAMPLER_JITTERED, (s.division_x() * s.division_y()) as i32, s.division_x() as i32, s.division_y() as i32),
            PixelSampler::AdditiveRecurrenceSampler(s) => (SPT_SAMPLER_RECURRENCE, s.spp() as i32, 1, 1),
        };
        let desc = SptRenderDesc {
            camera_index: flat.camera_index(config.used_camera_name.as_deref()), // Scene::get_camera semantics
            width: config.width as i32,
            height: config.height as i32,
            max_depth: self.max_depth as i32,
            sampler_kind: kind,
            spp,
            div_x,
            div_y,
            filter_radius: self.filter.radius(),
            seed: rand::random(),
            tile_w: 0,
            tile_h: 0,
            rank: 0,
            world_size: 1,
            pool_size: 0,
            flags: 0,
            spp_first: 0, // the whole frame in one call (progressive passes: see SPT_RENDER_ACCUMULATE in spt.h)
            spp_count: 0,
        };
        let n = (config.width * config.height * 3) as usize;
        let (mut mean, mut rgb8) = (vec![0f32; n], vec![0u8; n]);
        unsafe {
            let mut ctx = std::ptr::null_mut();
            let mut sc = std::ptr::null_mut();
            check(spt_context_create(0, &mut ctx));
            check(spt_scene_create(ctx, &flat.desc(), &mut sc));
            check(spt_render(sc, &desc, mean.as_mut_ptr(), rgb8.as_mut_ptr(), std::ptr::null_mut()));
            spt_scene_destroy(sc);
            spt_context_destroy(ctx);
        }
        // film.rs:53-62 / 94-99: the library already applied x255 + clamp
        let image = image::RgbImage::from_raw(config.width, config.height, rgb8).unwrap();
        if let Err(err) = image.save(&config.output_filename) {
            println!("Failed to save image, err: {}", err);
        }
    }
}
