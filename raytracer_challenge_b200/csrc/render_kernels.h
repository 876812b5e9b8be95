// Launch interface between the host driver (rtc_api.cpp) and the kernels (render_kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "device_scene.h"

#define RTC_BLOCK 128
// resident blocks per SM the compiler must allow for (register cap = 65536 / (128 * blocks))
#ifndef RTC_TRACE_MIN_BLOCKS
#define RTC_TRACE_MIN_BLOCKS 5
#endif
#ifndef RTC_SHADE_MIN_BLOCKS
#define RTC_SHADE_MIN_BLOCKS 4
#endif

namespace rtcdev {

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
    int32_t pad;
};

// One bounce level of one batch.  All arrays are structure-of-arrays of 16-byte chunks, so a warp's loads and
// stores are fully coalesced 512-byte transactions.
//   ray queue (64 B per ray):  q0 = (origin.x, origin.y)  q1 = (origin.z, dir.x)  q2 = (dir.y, dir.z)
//                              q3 = (throughput, fuel<<32 | sample)
//   hit list (compacted: one entry per ray that hit something; 36 / 52 B per hit):
//                              hit_ray = index of the ray in the level's input,  hit0 = (t, primitive index)
//                              hit1 = (u, v)   hit2 = (n1, n2) [scenes with transparent materials only]
//   shadow[li * n_hits + h]    1 = light li is occluded at hit h
struct LevelParams {
    const double2 *in_q0, *in_q1, *in_q2, *in_q3;
    double2 *out_q0, *out_q1, *out_q2, *out_q3;
    const unsigned long long* in_count;
    unsigned long long* out_count;
    uint64_t in_capacity, out_capacity;
    uint64_t first_sample;  // primary level: first sample ordinal of this batch
    uint64_t n_primary;     // primary level: samples in this batch
    uint32_t* hit_ray;
    double2 *hit0, *hit1, *hit2;
    unsigned long long* hit_count;
    uint8_t* shadow;
    double* accum;              // 3 doubles per sample of the batch
    unsigned long long* stats;  // [0] active primary rays, [1] reflected pushed, [2] refracted pushed, [3..6] Tally: nodes, tris, prims, xforms, [7] most node visits of one closest-hit query
    int* overflow;
};

void upload_perm(const uint8_t* perm512);
// The three kernels of one level, in launch order.
void launch_trace(int feat, bool primary, int grid, cudaStream_t st, const Scene& sc, const FrameParams& f, const LevelParams& lv);
void launch_shadow(int feat, bool primary, int grid, cudaStream_t st, const Scene& sc, const FrameParams& f, const LevelParams& lv);
void launch_shade(int feat, bool primary, int grid, cudaStream_t st, const Scene& sc, const FrameParams& f, const LevelParams& lv);
void launch_resolve(cudaStream_t st, const FrameParams& f, const double* accum, uint64_t first_pixel, uint64_t n_pixels, double* out, int row_major);

}  // namespace rtcdev
