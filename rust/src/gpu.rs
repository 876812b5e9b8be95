//! World -> flat arrays -> librtc (new module `src/gpu.rs` of the `raytracer` crate).
//! UNBUILT IN THIS REPOSITORY'S IMAGE (no Rust toolchain).  The C++ mirror of exactly this walk is
//! raytracer_challenge_b200/host/raytracer.hpp (FlatScene), which IS exercised by the tests.
//!
//! Needs from the rest of the crate (all additive, nothing else changes):
//!   camera.rs : `pub(crate) fn ffi_parts(&self) -> (Matrix, f64, f64, f64)` returning the private
//!               (transform_inv, pixel_size, half_width, half_height) — src/camera.rs:9-12
//!   lib.rs    : `pub mod ffi; pub mod gpu;`
use crate::camera::Camera;
use crate::color::Color;
use crate::config::FUEL;
use crate::ffi::*;
use crate::linalg::{Matrix, Vector};
use crate::material::{JitterKind, Material, MixtureKind, Pattern};
use crate::noise::Noise;
use crate::shape::{Element, Geometry, GroupKind, Shape};
use crate::world::World;
use std::ffi::CStr;

fn m16(m: &Matrix) -> [f64; 16] {
    let mut out = [0.0; 16];
    for r in 0..4 {
        for c in 0..4 {
            out[r * 4 + c] = m[r][c]; // ops::Index<usize>, src/linalg/matrix.rs:225-231
        }
    }
    out
}
fn v3(v: &Vector) -> [f64; 3] {
    [v.x, v.y, v.z]
}

#[derive(Default)]
pub struct FlatScene {
    prims: Vec<rtc_prim>,
    nodes: Vec<rtc_node>,
    children: Vec<i32>,
    materials: Vec<rtc_material>,
    patterns: Vec<rtc_pattern>,
    lights: Vec<rtc_light>,
}

impl FlatScene {
    /// Depth-first over World.elements — the order World::intersect visits them (src/world.rs:21-23).
    pub fn new(world: &World) -> FlatScene {
        let mut f = FlatScene::default();
        for l in &world.lights {
            f.lights.push(rtc_light { intensity: [l.intensity.r, l.intensity.g, l.intensity.b], origin: v3(&l.origin) });
        }
        let inf = f64::INFINITY;
        f.nodes.push(rtc_node { kind: 3, first_child: 0, n_children: 0, reserved: 0, bbox_min: [-inf; 3], bbox_max: [inf; 3] });
        f.emit_children(0, &world.elements);
        f
    }

    fn emit_pattern(&mut self, p: &Pattern) -> i32 {
        let blank = rtc_pattern { kind: 1, left: -1, right: -1, noise_kind: 0, octaves: 0, reserved: 0, noise_scale: 0.0, color: [0.0; 3], transform_inv: m16(&Matrix::id()) };
        let rec = match p {
            Pattern::Debug => rtc_pattern { kind: 0, ..blank },
            Pattern::Plain { color } => rtc_pattern { kind: 1, color: [color.r, color.g, color.b], ..blank },
            Pattern::Jitter { kind, noise, pattern } => {
                let inner = self.emit_pattern(pattern);
                let (noise_kind, noise_scale, octaves) = match noise {
                    Noise::Simplex { scale } => (0, *scale, 0),
                    Noise::Fractal { scale, octaves } => (1, *scale, *octaves as i32),
                };
                let k = match kind { JitterKind::Color => 2, JitterKind::Point => 3 };
                rtc_pattern { kind: k, left: inner, noise_kind, noise_scale, octaves, ..blank }
            }
            Pattern::Mixture { kind, transform_inv, left, right } => {
                let l = self.emit_pattern(left);
                let r = self.emit_pattern(right);
                let k = match kind {
                    MixtureKind::Blend => 4, MixtureKind::Checkers => 5, MixtureKind::RingGradient => 6,
                    MixtureKind::Ring => 7, MixtureKind::Gradient => 8, MixtureKind::Stripes => 9,
                };
                rtc_pattern { kind: k, left: l, right: r, transform_inv: m16(transform_inv), ..blank }
            }
        };
        self.patterns.push(rec); // children precede their parent (post-order), as rtc_scene_create requires
        self.patterns.len() as i32 - 1
    }

    fn emit_material(&mut self, m: &Material) -> i32 {
        let pattern = self.emit_pattern(&m.pattern);
        self.materials.push(rtc_material {
            ambient: m.ambient, diffuse: m.diffuse, specular: m.specular, shininess: m.shininess, reflective: m.reflective,
            transparency: m.transparency, refractive_index: m.refractive_index, pattern, reserved: 0,
        });
        self.materials.len() as i32 - 1
    }

    fn emit_prim(&mut self, s: &Shape) -> i32 {
        let z = [0.0; 3];
        let mut p = rtc_prim {
            kind: 0, material: self.emit_material(&s.material), casts_shadow: s.casts_shadow as i32, closed: 0,
            transform_inv: m16(&s.transform_inv), transform_inv_tsp: m16(&s.transform_inv_tsp), material_inv: m16(&s.material_inv),
            min: 0.0, max: 0.0, p1: z, e1: z, e2: z, n1: z, n2: z, n3: z,
        };
        match s.geometry {
            Geometry::Sphere => p.kind = 0,
            Geometry::Plane => p.kind = 1,
            Geometry::Cube => p.kind = 2,
            Geometry::Cylinder { min, max, closed } => { p.kind = 3; p.min = min; p.max = max; p.closed = closed as i32 }
            Geometry::Cone { min, max, closed } => { p.kind = 4; p.min = min; p.max = max; p.closed = closed as i32 }
            Geometry::Triangle { p1, e1, e2, n, .. } => { p.kind = 5; p.p1 = v3(&p1); p.e1 = v3(&e1); p.e2 = v3(&e2); p.n1 = v3(&n) }
            Geometry::SmoothTriangle { p1, e1, e2, n1, n2, n3, .. } => {
                p.kind = 6; p.p1 = v3(&p1); p.e1 = v3(&e1); p.e2 = v3(&e2); p.n1 = v3(&n1); p.n2 = v3(&n2); p.n3 = v3(&n3)
            }
        }
        self.prims.push(p);
        self.prims.len() as i32 - 1
    }

    fn emit_children(&mut self, node: usize, elems: &[Element]) {
        let first = self.children.len();
        self.children.resize(first + elems.len(), 0); // contiguous child slots, filled depth-first
        self.nodes[node].first_child = first as i32;
        self.nodes[node].n_children = elems.len() as i32;
        for (i, e) in elems.iter().enumerate() {
            match e {
                Element::Primitive(shape) => self.children[first + i] = self.emit_prim(shape),
                Element::Composite(group) => {
                    let kind = match group.kind {
                        GroupKind::Union => 0, GroupKind::Intersection => 1, GroupKind::Difference => 2, GroupKind::Aggregation => 3,
                    };
                    self.nodes.push(rtc_node { kind, first_child: 0, n_children: 0, reserved: 0, bbox_min: v3(&group.bbox.min), bbox_max: v3(&group.bbox.max) });
                    let idx = self.nodes.len() - 1;
                    self.children[first + i] = !(idx as i32);
                    self.emit_children(idx, &group.children);
                }
            }
        }
    }

    fn desc(&self) -> rtc_scene_desc {
        rtc_scene_desc {
            abi_version: RTC_ABI_VERSION,
            n_prims: self.prims.len() as i32, n_nodes: self.nodes.len() as i32, n_children: self.children.len() as i32,
            n_materials: self.materials.len() as i32, n_patterns: self.patterns.len() as i32, n_lights: self.lights.len() as i32,
            build_flags: RTC_BUILD_FAST, // Image::par_render builds, renders once and destroys
            prims: self.prims.as_ptr(), nodes: self.nodes.as_ptr(), children: self.children.as_ptr(),
            materials: self.materials.as_ptr(), patterns: self.patterns.as_ptr(), lights: self.lights.as_ptr(),
        }
    }
}

fn last_error() -> String {
    unsafe { CStr::from_ptr(rtc_last_error()).to_string_lossy().into_owned() }
}

/// Pixel store of `Image` (replaces the private field `pixels: Vec<Color>`, src/image.rs:14; `Image::new`, `read`, `write`
/// and `ppm` keep working through `Deref<Target = [Color]>`).  A frame rendered by the library lands in page-locked host
/// memory as f64 (r, g, b) triples — with `#[repr(C)]` on `Color` (src/color.rs:6-11; additive) that IS a `[Color]`, so
/// the buffer is adopted as it is instead of being copied (199 MB at 3840x2160) and goes back to the library's pool on drop.
pub enum Pixels {
    Owned(Vec<Color>),
    Pinned { ptr: *mut Color, len: usize },
}
impl std::ops::Deref for Pixels {
    type Target = [Color];
    fn deref(&self) -> &[Color] {
        match self {
            Pixels::Owned(v) => v,
            Pixels::Pinned { ptr, len } => unsafe { std::slice::from_raw_parts(*ptr, *len) },
        }
    }
}
impl std::ops::DerefMut for Pixels {
    fn deref_mut(&mut self) -> &mut [Color] {
        match self {
            Pixels::Owned(v) => v,
            Pixels::Pinned { ptr, len } => unsafe { std::slice::from_raw_parts_mut(*ptr, *len) },
        }
    }
}
impl Drop for Pixels {
    fn drop(&mut self) {
        if let Pixels::Pinned { ptr, .. } = self {
            unsafe { rtc_free_host(*ptr as *mut _) } // parked in the library's pool: the next frame reuses it
        }
    }
}
impl std::fmt::Debug for Pixels {
    fn fmt(&self, f: &mut std::fmt::Formatter<'_>) -> std::fmt::Result {
        write!(f, "Pixels[{}]", self.len())
    }
}
impl From<Vec<Color>> for Pixels {
    fn from(v: Vec<Color>) -> Pixels {
        Pixels::Owned(v)
    }
}
unsafe impl Send for Pixels {}

/// The new body of Image::par_render (src/image.rs:65-81): every visible GPU renders its bands of rows.
/// Infallible signature in the reference, so failures panic (no CPU fallback exists).
pub fn render(camera: &Camera, world: &World) -> Pixels {
    let flat = FlatScene::new(world);
    let desc = flat.desc();
    let n_dev = unsafe { rtc_device_count() };
    assert!(n_dev > 0, "raytracer: no CUDA device: {}", last_error());
    // one host build (flattening, BVH), one upload per GPU
    let devices: Vec<i32> = (0..n_dev).collect();
    let mut scenes: Vec<*mut rtc_scene> = vec![std::ptr::null_mut(); devices.len()];
    let rc = unsafe { rtc_scene_create_multi(&desc, devices.as_ptr(), n_dev, scenes.as_mut_ptr()) };
    if rc != 0 {
        panic!("rtc_scene_create: {}", last_error());
    }
    let (transform_inv, pixel_size, half_width, half_height) = camera.ffi_parts();
    let cam = rtc_camera { hsize: camera.hsize as i32, vsize: camera.vsize as i32, transform_inv: m16(&transform_inv), pixel_size, half_width, half_height };
    let opts = rtc_opts { fuel: FUEL, spp_x: 1, spp_y: 1, shard_count: 1, ..Default::default() };
    let n = camera.hsize * camera.vsize;
    let bytes = (n * 3 * std::mem::size_of::<f64>()) as u64;
    let buf = unsafe { rtc_alloc_host(bytes) } as *mut f64; // page-locked, pooled: copies overlap the kernels
    assert!(!buf.is_null(), "rtc_alloc_host: {}", last_error());
    let mut stats = rtc_stats::default();
    let rc = unsafe { rtc_render_multi(scenes.as_ptr(), scenes.len() as i32, &cam, &opts, buf, &mut stats) };
    let msg = if rc != 0 { last_error() } else { String::new() };
    for s in &scenes { unsafe { rtc_scene_destroy(*s) } }
    if rc != 0 {
        unsafe { rtc_free_host(buf as *mut _) };
        panic!("rtc_render_multi: {}", msg);
    }
    Pixels::Pinned { ptr: buf as *mut Color, len: n } // the frame becomes the image: nothing is copied
}
