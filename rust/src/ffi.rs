//! `extern "C"` surface of include/rtc.h (ABI version 2).  New module `src/ffi.rs` of the `raytracer` crate.
//! UNBUILT IN THIS REPOSITORY'S IMAGE (no Rust toolchain); field order and types mirror include/rtc.h one to one —
//! tests/test_abi_host.py pins the C side's sizes against the same header.
#![allow(non_camel_case_types, dead_code)]
use std::os::raw::{c_char, c_int};

pub const RTC_ABI_VERSION: i32 = 2;
/// rtc_scene_desc.build_flags: the scene is rendered once and destroyed -> build the BVH on the device (LBVH)
pub const RTC_BUILD_FAST: u32 = 1;

#[repr(C)]
#[derive(Clone, Copy)]
pub struct rtc_prim {
    pub kind: i32, // 0 sphere, 1 plane, 2 cube, 3 cylinder, 4 cone, 5 triangle, 6 smooth triangle
    pub material: i32,
    pub casts_shadow: i32,
    pub closed: i32,
    pub transform_inv: [f64; 16],
    pub transform_inv_tsp: [f64; 16],
    pub material_inv: [f64; 16],
    pub min: f64,
    pub max: f64,
    pub p1: [f64; 3],
    pub e1: [f64; 3],
    pub e2: [f64; 3],
    pub n1: [f64; 3],
    pub n2: [f64; 3],
    pub n3: [f64; 3],
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct rtc_node {
    pub kind: i32, // 0 union, 1 intersection, 2 difference, 3 aggregation
    pub first_child: i32,
    pub n_children: i32,
    pub reserved: i32,
    pub bbox_min: [f64; 3],
    pub bbox_max: [f64; 3],
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct rtc_material {
    pub ambient: f64,
    pub diffuse: f64,
    pub specular: f64,
    pub shininess: f64,
    pub reflective: f64,
    pub transparency: f64,
    pub refractive_index: f64,
    pub pattern: i32,
    pub reserved: i32,
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct rtc_pattern {
    pub kind: i32, // 0 debug, 1 plain, 2 colour jitter, 3 point jitter, 4 blend, 5 checkers, 6 ring gradient, 7 ring, 8 gradient, 9 stripes
    pub left: i32,
    pub right: i32,
    pub noise_kind: i32, // 0 simplex, 1 fractal
    pub octaves: i32,
    pub reserved: i32,
    pub noise_scale: f64,
    pub color: [f64; 3],
    pub transform_inv: [f64; 16],
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct rtc_light {
    pub intensity: [f64; 3],
    pub origin: [f64; 3],
}

#[repr(C)]
pub struct rtc_scene_desc {
    pub abi_version: i32,
    pub n_prims: i32,
    pub n_nodes: i32,
    pub n_children: i32,
    pub n_materials: i32,
    pub n_patterns: i32,
    pub n_lights: i32,
    /// RTC_BUILD_*: 0 = default (host SAH build); occupies the former alignment padding
    pub build_flags: u32,
    pub prims: *const rtc_prim,
    pub nodes: *const rtc_node,
    pub children: *const i32,
    pub materials: *const rtc_material,
    pub patterns: *const rtc_pattern,
    pub lights: *const rtc_light,
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct rtc_camera {
    pub hsize: i32,
    pub vsize: i32,
    pub transform_inv: [f64; 16],
    pub pixel_size: f64,
    pub half_width: f64,
    pub half_height: f64,
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct rtc_opts {
    pub fuel: i32,
    pub spp_x: i32,
    pub spp_y: i32,
    pub shard_index: i32,
    pub shard_count: i32,
    pub tile_w: i32,
    pub tile_h: i32,
    pub flags: u32,
    pub win_x0: i32,
    pub win_y0: i32,
    pub win_w: i32,
    pub win_h: i32,
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct rtc_stats {
    pub rays_primary: u64,
    pub rays_shadow: u64,
    pub rays_reflected: u64,
    pub rays_refracted: u64,
    pub rays_ref: u64,
    pub kernel_launches: u64,
    pub pixels: u64,
    pub ms_total: f64,
    pub ms_kernels: f64,
    pub ms_trace_shade: f64,
    pub ms_d2h: f64,
    pub overflow: i32,
    pub reserved: i32,
    pub surface_hits: u64,
    pub bvh_node_visits: u64,
    pub triangle_tests: u64,
    pub primitive_tests: u64,
    pub ray_transforms: u64,
    pub ms_trace: f64,
    pub ms_shadow: f64,
    pub ms_shade: f64,
    pub max_node_visits: u64,
    pub shadow_queries_culled: u64,
}

/// Tuning knobs and test hooks, process-wide (0 = library default); none changes a pixel.
#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct rtc_tuning {
    pub batches: i32,
    pub batch_samples: i32,
    pub pipes: i32,
    pub blocks_per_sm: i32,
    pub no_taper: i32,
    pub no_copy_stream: i32,
    pub no_stream_priorities: i32,
    pub no_graphs: i32,
    pub no_fast_planes: i32,
    pub no_plain_shade: i32,
    pub two_pass: i32,
    pub lbvh_leaf: i32,
    pub bvh_builder: i32,
    pub build_threads: i32,
    pub sah_leaf_weight_x100: i32,
    pub sah_bins: i32,
    pub sah_small_node: i32,
    pub host_paced_copies: i32,
    pub reserved: [i32; 14],
}

#[repr(C)]
pub struct rtc_scene {
    _private: [u8; 0],
}

extern "C" {
    pub fn rtc_device_count() -> c_int;
    pub fn rtc_scene_create(desc: *const rtc_scene_desc, device: c_int, out: *mut *mut rtc_scene) -> c_int;
    pub fn rtc_scene_create_multi(desc: *const rtc_scene_desc, devices: *const c_int, n: c_int, out: *mut *mut rtc_scene) -> c_int;
    pub fn rtc_scene_destroy(scene: *mut rtc_scene);
    pub fn rtc_render(scene: *mut rtc_scene, camera: *const rtc_camera, opts: *const rtc_opts, out_rgb: *mut f64, stats: *mut rtc_stats) -> c_int;
    pub fn rtc_render_multi(scenes: *const *mut rtc_scene, n: c_int, camera: *const rtc_camera, opts: *const rtc_opts, out_rgb: *mut f64, stats: *mut rtc_stats) -> c_int;
    pub fn rtc_render_noalloc(scene: *mut rtc_scene, camera: *const rtc_camera, opts: *const rtc_opts, out_rgb: *mut f64, stats: *mut rtc_stats) -> c_int;
    pub fn rtc_alloc_host(bytes: u64) -> *mut std::ffi::c_void;
    pub fn rtc_free_host(p: *mut std::ffi::c_void);
    pub fn rtc_release_host_memory();
    pub fn rtc_set_tuning(tuning: *const rtc_tuning);
    pub fn rtc_get_tuning(out: *mut rtc_tuning);
    pub fn rtc_last_error() -> *const c_char;
}
