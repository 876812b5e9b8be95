// Launch interface between the host driver (rtc_api.cpp) and the kernels (render_kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "device_scene.h"

#define RTC_BLOCK 128

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

// Ray queue: structure-of-arrays of 16-byte chunks, so a warp's loads/stores are four fully coalesced
// 512-byte transactions per ray record (64 B per ray):
//   q0 = (origin.x, origin.y)  q1 = (origin.z, dir.x)  q2 = (dir.y, dir.z)  q3 = (throughput, fuel<<32 | sample)
struct LevelParams {
    const double2 *in_q0, *in_q1, *in_q2, *in_q3;
    double2 *out_q0, *out_q1, *out_q2, *out_q3;
    const unsigned long long* in_count;
    unsigned long long* out_count;
    uint64_t in_capacity, out_capacity;
    uint64_t first_sample;  // primary level: first sample ordinal of this batch
    uint64_t n_primary;     // primary level: samples in this batch
    double* accum;          // 3 doubles per sample of the batch
    unsigned long long* stats;  // [0] rays in, [1] shadow, [2] reflected pushed, [3] refracted pushed, [4] R_ref
    int* overflow;
};

void upload_perm(const uint8_t* perm512);
void launch_level(int feat, bool primary, int grid, cudaStream_t st, const Scene& sc, const FrameParams& f, const LevelParams& lv);
void launch_resolve(cudaStream_t st, const FrameParams& f, const double* accum, uint64_t first_pixel, uint64_t n_pixels, double* out, int row_major);

}  // namespace rtcdev
