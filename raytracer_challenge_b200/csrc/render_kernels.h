// Launch interface between the host driver (rtc_api.cpp) and the kernels (render_kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "device_scene.h"

#define RTC_BLOCK 128
// resident blocks per SM the compiler must allow for (register cap = 65536 / (128 * blocks))
#ifndef RTC_TRACE_MIN_BLOCKS
#define RTC_TRACE_MIN_BLOCKS 7
#endif
#ifndef RTC_SHADOW_MIN_BLOCKS
#define RTC_SHADOW_MIN_BLOCKS 8
#endif
#ifndef RTC_TRACE_MIN_BLOCKS_TRANSPARENT  // the n1/n2 replay keeps more state alive: trade occupancy for no spills
#define RTC_TRACE_MIN_BLOCKS_TRANSPARENT 4  // (r02i sweep with the one-pass walk: 4 blocks = 128 registers beats 3 and 5 on the glass scenes)
#endif
#ifndef RTC_TRACE_MIN_BLOCKS_CSG  // CSG evaluation on top of that
#define RTC_TRACE_MIN_BLOCKS_CSG 3
#endif
#ifndef RTC_SHADOW_MIN_BLOCKS_CSG
#define RTC_SHADOW_MIN_BLOCKS_CSG 4  // (r02n sweep: 128 registers instead of 64 — the CSG lists spill less: chapter14 shadows 4.17 -> 3.73 ms)
#endif
#ifndef RTC_SHADE_MIN_BLOCKS
#define RTC_SHADE_MIN_BLOCKS 8  // (r02s sweep: 64 registers; the lighting kernels stream records and want occupancy: csg_noise 6.40 -> 6.19 ms, hexagon 0.93 -> 0.89)
#endif
#ifndef RTC_SHADE_PLAIN_MIN_BLOCKS  // scenes without procedural patterns: no evaluator, no permutation table, small register budget
#define RTC_SHADE_PLAIN_MIN_BLOCKS 8
#endif

namespace rtcdev {

constexpr int MAX_LEVELS = 13;  // fuel <= 12

// Per-frame constants (camera = the six values Camera::ray_at_pixel reads, src/camera.rs:39-55).
struct FrameParams {
    double cam_inv[12];
    double pixel_size, half_width, half_height;
    int32_t win_x0, win_y0, win_w, win_h;  // window of the camera image being rendered
    int32_t tile_w, tile_h, tiles_x, tiles_y;
    int32_t shard_index, shard_count;
    int32_t spp_x, spp_y, spp;
    int32_t fuel;
    uint32_t flags;
    int32_t two_bins;  // compact rays / hits that leave BVH items apart from those that leave always-tested primitives (render_kernels.cu)
};

// Hit list of one batch: structure-of-arrays of 16-byte chunks (a warp's accesses are whole 512-byte lines), one
// segment per bounce level at a static offset, each holding two bins that grow towards each other: hits on BVH items
// from the segment's first slot upwards (counts), hits on always-tested primitives from its last slot downwards
// (counts_back).  A segment is sized for "every ray of the level hits", so the bins cannot meet.
// Written by trace_spawn, read by trace_shadow (a, b, c) and light:
//     a = (over_point.x, over_point.y)   b = (over_point.z, normal.x)   c = (normal.y, normal.z)
//     d = (eye.x, eye.y)                 e = (eye.z, reflectance)       f = (primitive | parent << 32, HIT_* bits)
// parent: level 0 -> the sample of the batch; level k > 0 -> index of the parent's record in the whole list (its level's
// seg_off + slot).
// child_rgb: the colour World::color_at returned for the reflected (kind 0) and refracted (kind 1) ray a hit spawned,
// 3 doubles per slot; slot = record * 2 + kind if some material both reflects and refracts (fan == 2), else = record.
// Every spawned ray stores its colour exactly once: black when it misses (trace_spawn), the shaded colour when it
// hits (light) — so the array needs no clearing, and no atomics.
constexpr uint32_t HIT_KIND_REFRACTED = 1u;  // this hit's ray is the refracted child of its parent (else the reflected one)
constexpr uint32_t HIT_SPAWNED_REFLECTED = 2u, HIT_SPAWNED_REFRACTED = 4u;  // the children this hit spawned (src/world.rs:84-132)
struct HitList {
    double2 *a, *b, *c, *d, *e, *f;
    double* child_rgb;
    const unsigned long long* counts;       // hits per level, front bin (device counters)
    const unsigned long long* counts_back;  // ... back bin
    uint32_t seg_off[MAX_LEVELS];      // first record of each level's segment
    uint32_t seg_cap[MAX_LEVELS];
    int32_t n_levels;
    uint32_t total_cap;                // all segments together = stride of the per-light shadow flag planes
    int32_t fan;                       // child slots per record (1 or 2)
    int32_t pad;
};

// One bounce level of one batch.
//   ray queue (64 B per ray):  q0 = (origin.x, origin.y)  q1 = (origin.z, dir.x)  q2 = (dir.y, dir.z)
//                              q3 = (parent record | fuel << 32 | child kind << 40, unused)
//   two bins like the hit list: children of hits on BVH items in slots 0, 1, ... (count), children of hits on
//   always-tested primitives in slots capacity-1, capacity-2, ... (count_back)
struct LevelParams {
    const double2 *in_q0, *in_q1, *in_q2, *in_q3;
    double2 *out_q0, *out_q1, *out_q2, *out_q3;
    const unsigned long long *in_count, *in_count_back;
    unsigned long long *out_count, *out_count_back;
    uint64_t in_capacity, out_capacity;
    uint64_t first_sample;  // primary level: first sample ordinal of this batch
    uint64_t n_primary;     // primary level: samples in this batch
    int32_t level;
    int32_t pad;
    unsigned long long *hit_count, *hit_count_back;  // this level's counters (= hits.counts + level, hits.counts_back + level)
    uint8_t* shadow;            // [light][hits.total_cap]: 1 = occluded
    double* accum;              // 3 doubles per sample of the batch: the colour of its primary ray
    unsigned long long* stats;  // [0] active primary rays, [1] reflected pushed, [2] refracted pushed, [3..6] Tally: nodes, tris, prims, xforms,
                                // [7] most node visits of one closest-hit query, [8] shadow queries answered without tracing (light behind the surface)
    int* overflow;
};

void upload_perm(const uint8_t* perm512);
// Per batch: trace_spawn once per level (a chain); then over every hit of the batch trace_shadow (one thread per
// (hit, light)); then light per level, deepest first.
void launch_trace(int feat, bool primary, int grid, cudaStream_t st, const Scene& sc, const FrameParams& f, const LevelParams& lv, const HitList& hits);
void launch_shadow(int feat, int grid, cudaStream_t st, const Scene& sc, const FrameParams& f, const LevelParams& lv, const HitList& hits);
void launch_light(int grid, cudaStream_t st, const Scene& sc, const LevelParams& lv, const HitList& hits);
// Device-side LBVH build (bvh_build.cu): everything lives in the scene arena.
struct LbvhArgs {
    int n;                      // BVH items
    int leaf_max;               // subtrees of at most this many items (1..4) collapse into one leaf
    const float4* box;          // 2 per item: (lo.xyz, -), (hi.xyz, -): the padded, outward-rounded fp32 item boxes
    const int32_t* item_ids;    // per item: primitive index, or ~csg root
    const Prim* prims;          // the uploaded primitive table (kind, transform id, geometry slot of an item)
    const double* tri_pe;       // the uploaded triangle records, by geometry slot
    LeafItem* leaf_items;       // out: leaf order
    double* leaf_tri;           // out: leaf order
    BvhNode* bvh;               // out: n - 1 slots
    unsigned long long* keys;   // n
    unsigned long long* keys_alt;  // n > LBVH_ONE_BLOCK_ITEMS: the radix sort's second buffer
    uint32_t* radix_hist;          // ... and its (digit, tile) histogram: 16 * ceil(n / 256)
    int2* child;                // n - 1
    int* parent;                // 2 n
    int2* range;                // n - 1
    float4* nbox;               // 2 (n - 1)
    int* counter;               // n
    int* dense;                 // n
    int* result;                // 4: root link, radix-tree depth, node count
    float root_lo[3], root_inv_extent[3];
};
int build_lbvh(const LbvhArgs& a, cudaStream_t st);

void launch_resolve(cudaStream_t st, const FrameParams& f, const double* accum, uint64_t first_pixel, uint64_t n_pixels, double* out, int row_major);

}  // namespace rtcdev
