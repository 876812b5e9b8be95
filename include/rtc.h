/*
 * rtc.h — C ABI of the B200-native render path (drop-in boundary for the reference's
 * `Image::par_render(camera: &Camera, world: &World) -> Image`, /root/reference/src/image.rs:65-81).
 *
 * The reference (a pure-Rust crate) has no FFI today; these are the entry points its
 * `src/image.rs` would bind through `extern "C"` (see INTEGRATION.md for the Rust stub).
 * Everything crossing the boundary is plain-old-data in host memory: the caller flattens its
 * `World` (already holding baked per-primitive matrices, src/shape.rs:47-72) into the arrays
 * below; the library copies what it needs during rtc_scene_create and builds its own device
 * structures (BVH, SoA tables).  No CPU fallback exists: without a CUDA device every call fails.
 *
 * Numeric contract: all reference arithmetic is f64 and so is every parity-critical device
 * computation; matrices are row-major 4x4 (src/linalg/matrix.rs:7-10) and must be affine
 * (last row 0 0 0 1), which is all the reference's constructors can produce.
 */
#ifndef RTC_H
#define RTC_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTC_ABI_VERSION 2

/* Geometry kinds: src/shape.rs:465-498 */
enum rtc_prim_kind {
    RTC_SPHERE = 0,
    RTC_PLANE = 1,
    RTC_CUBE = 2,
    RTC_CYLINDER = 3,
    RTC_CONE = 4,
    RTC_TRIANGLE = 5,
    RTC_SMOOTH_TRIANGLE = 6
};

/* Group kinds: src/shape.rs:161-167 */
enum rtc_group_kind { RTC_UNION = 0, RTC_INTERSECTION = 1, RTC_DIFFERENCE = 2, RTC_AGGREGATION = 3 };

/* Pattern node kinds: src/material.rs:58-65,190-194,226-234 */
enum rtc_pattern_kind {
    RTC_PAT_DEBUG = 0,
    RTC_PAT_PLAIN = 1,
    RTC_PAT_JITTER_COLOR = 2,
    RTC_PAT_JITTER_POINT = 3,
    RTC_PAT_BLEND = 4,
    RTC_PAT_CHECKERS = 5,
    RTC_PAT_RING_GRADIENT = 6,
    RTC_PAT_RING = 7,
    RTC_PAT_GRADIENT = 8,
    RTC_PAT_STRIPES = 9
};

/* Noise kinds: src/noise.rs:4-7 */
enum rtc_noise_kind { RTC_NOISE_SIMPLEX = 0, RTC_NOISE_FRACTAL = 1 };

/* One primitive = the render-time content of `Shape` (src/shape.rs:296-306) after
 * propagate_inverses.  Primitives are listed in depth-first order of World.elements. */
typedef struct rtc_prim {
    int32_t kind;          /* rtc_prim_kind */
    int32_t material;      /* index into materials[] */
    int32_t casts_shadow;  /* src/shape.rs:304 */
    int32_t closed;        /* cylinder / cone caps */
    double transform_inv[16];     /* world -> object */
    double transform_inv_tsp[16]; /* normal matrix */
    double material_inv[16];      /* world -> pattern space (src/shape.rs:61-70) */
    double min, max;              /* cylinder / cone y-range */
    double p1[3], e1[3], e2[3];   /* triangles: as precomputed at src/shape.rs:369-412 */
    double n1[3], n2[3], n3[3];   /* smooth triangle normals; flat triangle: n in n1 */
} rtc_prim;

/* One group = `Group` (src/shape.rs:180-185) with its world-space box.  nodes[0] is a virtual
 * root (kind AGGREGATION, no box gate: World::intersect, src/world.rs:18-24) whose children are
 * World.elements.  children[first_child .. first_child+n_children): value >= 0 is a primitive
 * index, value < 0 is ~node_index. */
typedef struct rtc_node {
    int32_t kind; /* rtc_group_kind */
    int32_t first_child;
    int32_t n_children;
    int32_t reserved;
    double bbox_min[3];
    double bbox_max[3];
} rtc_node;

/* `Material` (src/material.rs:18-28) */
typedef struct rtc_material {
    double ambient, diffuse, specular, shininess, reflective, transparency, refractive_index;
    int32_t pattern; /* root index into patterns[] */
    int32_t reserved;
} rtc_material;

/* One node of a `Pattern` tree (src/material.rs:58-65) */
typedef struct rtc_pattern {
    int32_t kind;       /* rtc_pattern_kind */
    int32_t left;       /* Mixture: left child; Jitter: the inner pattern */
    int32_t right;      /* Mixture: right child */
    int32_t noise_kind; /* Jitter: rtc_noise_kind */
    int32_t octaves;    /* Jitter with fractal noise */
    int32_t reserved;
    double noise_scale;
    double color[3];          /* Plain */
    double transform_inv[16]; /* Mixture */
} rtc_pattern;

/* `PointLight` (src/light.rs:4-8) */
typedef struct rtc_light {
    double intensity[3];
    double origin[3];
} rtc_light;

/* rtc_scene_desc.build_flags */
#define RTC_BUILD_FAST 1u /* the scene is rendered once and destroyed (the call pattern of Image::par_render, src/image.rs:65-81):
                           * build the BVH on the device (LBVH, tens of microseconds) instead of the host's binned-SAH build
                           * (1-2 ms for the teapot).  The tree is of lower quality — a frame traces a few per cent slower — and
                           * the rendered pixels are identical: the BVH only culls. */

typedef struct rtc_scene_desc {
    int32_t abi_version; /* RTC_ABI_VERSION */
    int32_t n_prims, n_nodes, n_children, n_materials, n_patterns, n_lights;
    uint32_t build_flags; /* RTC_BUILD_*; 0 = default.  (Occupies what was alignment padding: the layout is unchanged.) */
    const rtc_prim* prims;
    const rtc_node* nodes;
    const int32_t* children;
    const rtc_material* materials;
    const rtc_pattern* patterns;
    const rtc_light* lights;
} rtc_scene_desc;

/* The six values `Camera::ray_at_pixel` reads (src/camera.rs:5-13,39-55) */
typedef struct rtc_camera {
    int32_t hsize, vsize;
    double transform_inv[16];
    double pixel_size, half_width, half_height;
} rtc_camera;

#define RTC_FLAG_BRUTE_FORCE 1u /* test every primitive for every ray (no BVH culling); validation only */
#define RTC_FLAG_SCATTER_ROWS 4u /* sharded render with full-width tiles (tile_w == window width): out_rgb is the FULL row-major
                                  * window buffer and the shard's pixel rows are written at their row-major positions (the other
                                  * rows are left untouched), so N shards fill one shared host image with N parallel copies */
#define RTC_FLAG_PROFILE 2u     /* one pipeline, CUDA events around every launch: fills rtc_stats.ms_trace/ms_shadow/ms_shade */

typedef struct rtc_opts {
    int32_t fuel;          /* src/config.rs:2 (FUEL = 5) */
    int32_t spp_x, spp_y;  /* supersampling grid; 1x1 = the reference's single centre ray */
    int32_t shard_index;   /* this call renders tiles t with t % shard_count == shard_index */
    int32_t shard_count;   /* 1 = whole image */
    int32_t tile_w, tile_h;/* shard tile size in pixels (0 -> 8x4) */
    uint32_t flags;
    int32_t win_x0, win_y0, win_w, win_h; /* render only this window of the camera image (0,0,0,0 = all) */
} rtc_opts;

typedef struct rtc_stats {
    uint64_t rays_primary, rays_shadow, rays_reflected, rays_refracted; /* R_uniq: the reference's World::intersect queries answered, secondary
                                                                         * sub-trees deduplicated across lights.  rays_shadow = one is_shadowed query per
                                                                         * (hit, light); shadow_queries_culled of them were answered without a traversal */
    uint64_t rays_ref;       /* R_ref: World::intersect queries the reference would issue for the same image */
    uint64_t kernel_launches;
    uint64_t pixels;         /* pixels produced by this call (shard) */
    double ms_total;         /* rtc_render entry -> last pixel in host memory */
    double ms_kernels;       /* CUDA-event time of all kernels of this call */
    double ms_trace_shade;   /* CUDA-event time of the wavefront kernels only */
    double ms_d2h;           /* reserved (0): rtc_render copies each finished batch out while later batches are traced */
    int32_t overflow;        /* non-zero: an on-chip list overflowed; result invalid (call fails) */
    int32_t reserved;
    /* work actually done by the traversal kernels (inputs of the roofline model) */
    uint64_t surface_hits;    /* rays that hit something = shaded surface points */
    /* work tallies, RTC_FLAG_PROFILE only (0 otherwise: the counters cost registers in the traversal loop) */
    uint64_t bvh_node_visits; /* each visit = two fp32 slab tests */
    uint64_t triangle_tests, primitive_tests; /* exact f64 intersections: triangles / every other kind */
    uint64_t ray_transforms;  /* world -> object ray transforms (src/ray.rs:14-19) actually computed */
    /* RTC_FLAG_PROFILE only: CUDA-event time per kernel class, launches serialised on one stream */
    double ms_trace, ms_shadow, ms_shade;
    uint64_t max_node_visits; /* the longest single closest-hit traversal of the call (BVH node visits of one ray) */
    uint64_t shadow_queries_culled; /* is_shadowed queries whose light lies behind the surface: Shape::lighting ignores their answer
                                     * (src/shape.rs:449), so nothing is traced for them; included in rays_shadow */
} rtc_stats;

typedef struct rtc_scene rtc_scene;

/* Number of CUDA devices visible (0 or negative status on failure -> rtc_last_error). */
int rtc_device_count(void);

/* Copies and uploads the scene to CUDA device `device`, building the BVH (on the host — binned SAH — or, with
 * desc->build_flags & RTC_BUILD_FAST and 64 .. 4 M bounded items, on the device after the upload).  0 = ok. */
int rtc_scene_create(const rtc_scene_desc* desc, int device, rtc_scene** out);

/* The same for `n` distinct devices at once: ONE host build (validation, flattening, BVH), one upload per device
 * (in parallel); out[i] is the handle for devices[i].  What Image::par_render does before rtc_render_multi.
 * All or nothing: on failure no handle is returned. */
int rtc_scene_create_multi(const rtc_scene_desc* desc, const int* devices, int n, rtc_scene** out);

void rtc_scene_destroy(rtc_scene* scene);

/* One render at a time per device: the library keeps one render context per GPU (streams, scratch pools, cached CUDA graphs)
 * behind a mutex, so concurrent calls on the same device serialise; calls on different devices run concurrently.
 *
 * Blocking render into caller-allocated host memory: out_rgb holds one (r,g,b) f64 triple per
 * output slot (see rtc_shard_pixels; shard_count == 1: the whole window row-major, `y*w + x`, like
 * Image.pixels, src/image.rs:114-116).
 * Replaces the body of Image::par_render (src/image.rs:65-81).  0 = ok. */
int rtc_render(rtc_scene* scene, const rtc_camera* camera, const rtc_opts* opts, double* out_rgb, rtc_stats* stats);

/* One frame on n GPUs of this box from ONE host thread (the form Image::par_render needs): scenes[i] are uploads of
 * the same scene on n distinct devices; bands of tile_h pixel rows are dealt round-robin to the devices, each device
 * renders its bands on its own worker thread and copies them straight to their row-major place in out_rgb over its
 * own PCIe link (RTC_FLAG_SCATTER_ROWS).  No collective is involved: pixels are independent (src/image.rs:66-74).
 * stats: ray counts summed over devices, ms_kernels = the slowest device. */
int rtc_render_multi(rtc_scene* const* scenes, int n, const rtc_camera* camera, const rtc_opts* opts, double* out_rgb,
                     rtc_stats* stats);

/* Same as rtc_render, but leaves the result in device memory owned by the library's per-device render context (valid until
 * the next render on that device or rtc_release_device_memory); *out_dev_rgb receives the device pointer (f64
 * triples), *out_pixels the count. */
int rtc_render_device(rtc_scene* scene, const rtc_camera* camera, const rtc_opts* opts, double** out_dev_rgb,
                      uint64_t* out_pixels, rtc_stats* stats);

/* No counterpart in the reference: added for callers that cannot allocate (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * rtc_render for a small window with NO allocation: no cudaMalloc / cudaFree / cudaHostAlloc, no stream or event is
 * created; every byte of per-render scratch lives in one static device array of RTC_NOALLOC_SCRATCH_BYTES that each GPU
 * holds from the moment it loads the library until the process exits.  Same kernels, same pixels as rtc_render.
 *   - `scene` is an existing handle (rtc_scene_create allocates; this call does not); out_rgb receives win_w * win_h (or
 *     hsize * vsize) f64 rgb triples, row-major, and may be HOST or DEVICE memory of the caller.
 *   - opts: fuel, spp, window and RTC_FLAG_BRUTE_FORCE are honoured; sharding and the other flags are rejected.
 *   - A window whose queues do not fit the array is rejected with an error that names the bytes needed and the limit.
 *     Rule of thumb for fuel 5: ~19 000 samples (pixels x samples per pixel) when no material both reflects and refracts,
 *     ~1 100 samples when one does (the binary path tree is sized for its worst case).
 *   - Every call on a device shares the array, which keeps the previous call's contents; the call clears what must start at
 *     zero.  Calls that overlap on one GPU would race: the library serialises the calls of one process (per-device
 *     mutex, and the call blocks until the pixels have landed).  Each process has its own copy of the array.
 *   - Work runs on the legacy default stream and the call ends with a stream synchronisation.
 * 0 = ok; otherwise rtc_last_error(). */
#define RTC_NOALLOC_SCRATCH_BYTES (16u << 20)
int rtc_render_noalloc(rtc_scene* scene, const rtc_camera* camera, const rtc_opts* opts, double* out_rgb, rtc_stats* stats);

/* Output slots of shard (shard_index, shard_count) of a w x h window.  The window is cut into tile_w x tile_h
 * tiles (default 8x4, numbered row-major); shard i owns tiles i, i+shard_count, ...; its output lists the owned
 * tiles in ascending order, each tile row-major and FULL size: slots that fall outside the window (partial edge
 * tiles) are padding.  Returns the slot count; if out_indices != NULL it receives, per slot, the linear window
 * index y*w+x of the pixel or 0xffffffff for padding.  With shard_count == 1 rtc_render instead returns the plain
 * row-major window (w*h pixels, no padding).  Host-side helper; needs no GPU. */
uint64_t rtc_shard_pixels(int32_t w, int32_t h, int32_t tile_w, int32_t tile_h, int32_t shard_index, int32_t shard_count,
                          uint32_t* out_indices);

/* `Color::clamp` (src/color.rs:42-46) applied to n doubles -> n bytes, on the host. */
void rtc_quantize(const double* rgb, uint8_t* out, uint64_t n);

/* Scene build facts (rtc_scene_stat keys). */
enum rtc_stat_key {
    RTC_STAT_PRIMS = 0, RTC_STAT_BVH_NODES = 1, RTC_STAT_BVH_DEPTH = 2, RTC_STAT_UNBOUNDED = 3,
    RTC_STAT_GATES_EXACT = 4, RTC_STAT_GATES_CONSERVATIVE = 5, RTC_STAT_GATES_ALWAYS_PASS = 6,
    RTC_STAT_CSG_ROOTS = 7, RTC_STAT_XFORMS = 8, RTC_STAT_FEATURES = 9, RTC_STAT_BVH_ITEMS = 10, RTC_STAT_DEVICE_BYTES = 11,
    RTC_STAT_CONE_GUARDS = 12
};
int64_t rtc_scene_stat(const rtc_scene* scene, int key);

/* Runs the host half of scene upload only (validation, flattening, BVH build) — no GPU needed.  0 = the
 * scene would be accepted; stats (optional, 16 entries indexed by rtc_stat_key) are filled on success. */
int rtc_scene_validate(const rtc_scene_desc* desc, int64_t* stats16);

/* The library keeps its render scratch (ray queues, hit lists, frame buffer) per device across scenes, so that
 * create-render-destroy per frame (what Image::par_render does) does not reallocate it.  This frees it. */
void rtc_release_device_memory(int device);

/* Page-locked host memory for out_rgb (optional: any host buffer works, pinned ones copy at full PCIe rate).  Pooled:
 * rtc_free_host parks the buffer inside the library and the next rtc_alloc_host of a similar size reuses it, so that one
 * frame-sized buffer per Image::par_render call (src/image.rs:65-81) costs nothing after the first call (page-locking
 * 199 MB takes tens of milliseconds).  rtc_release_host_memory hands the parked buffers back to the driver. */
void* rtc_alloc_host(uint64_t bytes);
void rtc_free_host(void* p);
void rtc_release_host_memory(void);
/* Page-locks / unlocks memory the caller already owns (e.g. a shared-memory frame several processes render into). */
int rtc_host_register(void* p, uint64_t bytes);
int rtc_host_unregister(void* p);

/* Measured roofline denominators on the current device: TFLOP/s (two flops per multiply-add pair) or GB/s. */
enum rtc_microbench_kind { RTC_MB_FP64_FMA = 0, RTC_MB_FP64_MUL_ADD = 1, RTC_MB_FP32_FMA = 2, RTC_MB_FP32_MUL_ADD = 3, RTC_MB_L2_READ = 4 };
int rtc_microbench(int what, double* result);

/* Tuning knobs and test hooks of the library, process-wide, set explicitly (the library reads no environment variable for
 * them; RTC_TRACE / RTC_TIMELINE print debugging timelines and change nothing else).  Every field: 0 = the library's
 * default.  None of them changes a pixel — they move work between batches, streams, builders and kernel variants. */
typedef struct rtc_tuning {
    int32_t batches;          /* aim for this many batches per rtc_render call (default 4 device-resident, 14 when copying out) */
    int32_t batch_samples;    /* ... or fix the samples per batch */
    int32_t pipes;            /* batch pipelines (CUDA streams) in flight, 1..8 (default 4) */
    int32_t blocks_per_sm;    /* grid size of the wavefront kernels = SMs x this (default 16) */
    int32_t no_taper;         /* 1: equal batches even when each finished batch is copied to the host */
    int32_t no_copy_stream;   /* 1: copy a finished batch out on its own pipeline's stream */
    int32_t no_stream_priorities; /* 1: equal stream priorities for host-output renders */
    int32_t no_graphs;        /* 1: enqueue every frame launch by launch instead of replaying a captured CUDA graph */
    int32_t no_fast_planes;   /* 1: always-tested planes go through the general primitive list, not the kernel parameters */
    int32_t no_plain_shade;   /* 1: scenes of plain-colour materials still use the pattern evaluator's kernel */
    int32_t two_pass;         /* 1: transparent hits re-walk the scene for their n1 / n2 crossing list (no one-pass collection) */
    int32_t lbvh_leaf;        /* device-built BVH: items per leaf, 1..4 (default 1; 2 when every item is a triangle) */
    int32_t bvh_builder;      /* 0: as rtc_scene_desc.build_flags says; 1: device (LBVH) whenever possible; 2: host (binned SAH) always */
    int32_t build_threads;    /* host BVH build: threads it may use (default: all cores; one process per GPU should pass cores / ranks) */
    int32_t sah_leaf_weight_x100, sah_bins, sah_small_node; /* host SAH build: leaf cost x 100 (default 300), bins 2..16 (16), exact-sweep size 2..8 (8) */
    int32_t host_paced_copies; /* 1: rtc_render into a host buffer issues each batch's copy from the calling thread once the batch's event has fired, instead of making the copy stream wait for it on the device (no CUDA-graph replay in this mode) */
    int32_t reserved[14];
} rtc_tuning;
void rtc_set_tuning(const rtc_tuning* tuning); /* NULL: back to the defaults */
void rtc_get_tuning(rtc_tuning* out);

/* Thread-local message for the last non-zero status. */
const char* rtc_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* RTC_H */
