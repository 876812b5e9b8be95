// rtc_render_noalloc: the wavefront of rtc_render for a small window, with every byte of per-render scratch (ray queues,
// hit list, child colours, shadow bytes, sample accumulators, the frame, the counters) in ONE file-scope device array
// instead of the cudaMalloc'ed pools of the per-device render context (rtc_api.cpp).  For callers that must not
// allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  The call makes no
// cudaMalloc / cudaFree / cudaHostAlloc and creates no stream or event; it launches the same kernels as rtc_render on
// the legacy default stream and ends with a stream synchronisation.
//
// No counterpart in the reference (include/rtc.h says the same).
//
// The array is shared by every call on a device and keeps the last call's contents: what must start at zero (the
// counters, the overflow flag, the sample accumulators) is cleared by each call; everything else is written before it is
// read.  Calls that overlap on one GPU would race on it: a per-device mutex serialises the calls of this process, and
// the call is blocking, so two host threads cannot overlap; processes do not share the array.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <mutex>
#include <string>

#include "../../include/rtc.h"
#include "render_kernels.h"
#include "rtc_internal.hpp"

namespace {

// Every GPU that loads the library holds this array until the process exits.
static __device__ __align__(256) unsigned char g_noalloc_scratch[RTC_NOALLOC_SCRATCH_BYTES];

constexpr int MAX_DEVICES = 64;
// counters section (unsigned long long each): per level queue / hit-list fill of both bins, then LevelParams::stats
constexpr int CNT_QUEUE = 0, CNT_HITS = 16, CNT_QUEUE_BACK = 32, CNT_HITS_BACK = 48, CNT_STATS = 64, CNT_TOTAL = 80;

struct PerDevice {
    std::mutex mu;
    bool ready = false;
    unsigned char* base = nullptr;  // DEVICE address of g_noalloc_scratch on this device
    int sm_count = 0;
};
PerDevice g_dev[MAX_DEVICES];

// In host code `g_noalloc_scratch` is a host-side stand-in for the symbol, not a device address: the address the kernels
// get comes from cudaGetSymbolAddress, asked once per device with that device current.
int setup_device(PerDevice& d, int device) {  // caller holds d.mu and has made `device` current
    if (d.ready) return 0;
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) return rtcapi::fail(std::string("rtc_render_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
    int sms = 0;
    e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    if (e != cudaSuccess || sms <= 0) return rtcapi::fail(std::string("rtc_render_noalloc: cudaDeviceGetAttribute: ") + cudaGetErrorString(e));
    d.base = (unsigned char*)p;
    d.sm_count = sms;
    d.ready = true;
    return 0;
}

// Sections of the scratch array, 256-byte aligned.
struct Carver {
    size_t at = 0;
    size_t take(size_t bytes) {
        const size_t off = (at + 255) & ~(size_t)255;
        at = off + std::max<size_t>(bytes, 16);
        return off;
    }
};

#define CU(call)                                                                                                       \
    do {                                                                                                               \
        cudaError_t e_ = (call);                                                                                       \
        if (e_ != cudaSuccess) return rtcapi::fail(std::string("rtc_render_noalloc: " #call ": ") + cudaGetErrorString(e_)); \
    } while (0)

}  // namespace

extern "C" int rtc_render_noalloc(rtc_scene* s, const rtc_camera* cam, const rtc_opts* o_in, double* out_rgb, rtc_stats* stats) {
    if (!s || !cam || !out_rgb) return rtcapi::fail("rtc_render_noalloc: null argument");
    rtc_opts o{};
    if (o_in) o = *o_in;
    else o.fuel = 5;
    if (o.spp_x <= 0) o.spp_x = 1;
    if (o.spp_y <= 0) o.spp_y = 1;
    if (o.tile_w <= 0) o.tile_w = 8;
    if (o.tile_h <= 0) o.tile_h = 4;
    if (o.shard_count > 1) return rtcapi::fail("rtc_render_noalloc: sharding is not supported (render a window instead)");
    if (o.flags & ~RTC_FLAG_BRUTE_FORCE) return rtcapi::fail("rtc_render_noalloc: only RTC_FLAG_BRUTE_FORCE is supported");
    if (o.fuel < 0 || o.fuel > 12) return rtcapi::fail("rtc_render_noalloc: fuel must be in 0..12");
    if (cam->hsize <= 0 || cam->vsize <= 0) return rtcapi::fail("rtc_render_noalloc: empty camera image");
    if ((int64_t)o.spp_x * o.spp_y > 4096) return rtcapi::fail("rtc_render_noalloc: more than 4096 samples per pixel");
    const bool whole = o.win_w <= 0 || o.win_h <= 0;
    const int32_t wx0 = whole ? 0 : o.win_x0, wy0 = whole ? 0 : o.win_y0;
    const int32_t ww = whole ? cam->hsize : o.win_w, wh = whole ? cam->vsize : o.win_h;
    if (wx0 < 0 || wy0 < 0 || (int64_t)wx0 + ww > cam->hsize || (int64_t)wy0 + wh > cam->vsize) return rtcapi::fail("rtc_render_noalloc: window outside the image");
    for (int i = 12; i < 16; ++i)
        if (!(std::fabs(cam->transform_inv[i] - (i == 15 ? 1.0 : 0.0)) <= 1e-9)) return rtcapi::fail("rtc_render_noalloc: camera transform is not affine");
    if (s->device < 0 || s->device >= MAX_DEVICES) return rtcapi::fail("rtc_render_noalloc: device ordinal out of range");

    rtcdev::FrameParams f{};
    for (int i = 0; i < 12; ++i) f.cam_inv[i] = cam->transform_inv[i];
    f.pixel_size = cam->pixel_size;
    f.half_width = cam->half_width;
    f.half_height = cam->half_height;
    f.win_x0 = wx0; f.win_y0 = wy0; f.win_w = ww; f.win_h = wh;
    f.tile_w = o.tile_w; f.tile_h = o.tile_h;
    f.tiles_x = (ww + o.tile_w - 1) / o.tile_w;
    f.tiles_y = (wh + o.tile_h - 1) / o.tile_h;
    f.shard_index = 0; f.shard_count = 1;
    f.spp_x = o.spp_x; f.spp_y = o.spp_y; f.spp = o.spp_x * o.spp_y;
    f.fuel = o.fuel;
    f.flags = o.flags;
    f.two_bins = (s->feat & rtcdev::FEAT_CSG) != 0;

    // ---- how much scratch this window needs: one batch holds the whole window ----
    const int L = s->dev.n_lights;
    const int fan = s->dev.max_children >= 2 ? 2 : 1;
    const uint64_t n_slots = (uint64_t)f.tiles_x * f.tiles_y * (uint64_t)(o.tile_w * o.tile_h);  // pixel slots incl. tile padding
    const uint64_t samples = n_slots * (uint64_t)f.spp;
    const uint64_t n_pixels = (uint64_t)ww * wh;
    // (sizes in double so that an absurd window cannot wrap: anything near the limit is rejected long before 2^53)
    double worst = 1.0;
    for (int k = 0; k < o.fuel && fan == 2; ++k) worst *= 2.0;
    double hit_cap_d = 0.0, level_cap_d = (double)samples;
    for (int k = 0; k <= o.fuel; ++k) { hit_cap_d += level_cap_d; if (fan == 2) level_cap_d *= 2.0; }
    const double need_estimate = 128.0 * (double)samples * worst + hit_cap_d * (96.0 + 24.0 * fan + std::max(L, 1)) + 24.0 * (double)samples + 24.0 * (double)n_pixels + 16384.0;
    if (need_estimate > (double)RTC_NOALLOC_SCRATCH_BYTES)
        return rtcapi::fail("rtc_render_noalloc: this window needs about " + std::to_string((uint64_t)need_estimate) + " bytes of scratch, the static array holds " +
                            std::to_string((uint64_t)RTC_NOALLOC_SCRATCH_BYTES) + " (RTC_NOALLOC_SCRATCH_BYTES): render a smaller window, fewer samples per pixel or less fuel, or use rtc_render");
    const uint64_t q_cap = samples * (uint64_t)worst;
    rtcdev::HitList hits{};
    uint64_t hit_cap = 0, level_cap = samples;
    hits.n_levels = o.fuel + 1;
    for (int k = 0; k <= o.fuel; ++k) {
        hits.seg_off[k] = (uint32_t)hit_cap;
        hits.seg_cap[k] = (uint32_t)level_cap;
        hit_cap += level_cap;
        if (fan == 2) level_cap *= 2;
    }
    hits.total_cap = (uint32_t)hit_cap;
    hits.fan = fan;
    Carver cv;
    size_t o_q[2][4], o_hit[6];
    const size_t o_cnt = cv.take(CNT_TOTAL * sizeof(unsigned long long));
    const size_t o_ovf = cv.take(sizeof(int));
    for (auto& side : o_q)
        for (auto& off : side) off = cv.take(q_cap * sizeof(double2));
    for (auto& off : o_hit) off = cv.take(hit_cap * sizeof(double2));
    const size_t o_child = cv.take(hit_cap * (uint64_t)fan * 3 * sizeof(double));
    const size_t o_shadow = cv.take(hit_cap * (uint64_t)std::max(L, 1));
    const size_t o_accum = cv.take(samples * 3 * sizeof(double));
    const size_t o_frame = cv.take(n_pixels * 3 * sizeof(double));
    if (cv.at > (size_t)RTC_NOALLOC_SCRATCH_BYTES)
        return rtcapi::fail("rtc_render_noalloc: this window needs " + std::to_string(cv.at) + " bytes of scratch, the static array holds " +
                            std::to_string((uint64_t)RTC_NOALLOC_SCRATCH_BYTES) + " (RTC_NOALLOC_SCRATCH_BYTES)");

    CU(cudaSetDevice(s->device));
    PerDevice& dev = g_dev[s->device];
    std::lock_guard<std::mutex> lock(dev.mu);  // one call at a time per device: the calls share the array
    if (int rc = setup_device(dev, s->device)) return rc;
    unsigned char* base = dev.base;
    cudaStream_t st = nullptr;  // legacy default stream: nothing to create

    unsigned long long* cnt = (unsigned long long*)(base + o_cnt);
    int* overflow = (int*)(base + o_ovf);
    double* accum = (double*)(base + o_accum);
    double* frame = (double*)(base + o_frame);
    // what must start at zero; the rest of the array is written before it is read
    CU(cudaMemsetAsync(cnt, 0, CNT_TOTAL * sizeof(unsigned long long), st));
    CU(cudaMemsetAsync(overflow, 0, sizeof(int), st));
    CU(cudaMemsetAsync(accum, 0, samples * 3 * sizeof(double), st));
    hits.a = (double2*)(base + o_hit[0]); hits.b = (double2*)(base + o_hit[1]); hits.c = (double2*)(base + o_hit[2]);
    hits.d = (double2*)(base + o_hit[3]); hits.e = (double2*)(base + o_hit[4]); hits.f = (double2*)(base + o_hit[5]);
    hits.child_rgb = (double*)(base + o_child);
    hits.counts = cnt + CNT_HITS;
    hits.counts_back = cnt + CNT_HITS_BACK;
    auto level_params = [&](int level) {
        rtcdev::LevelParams lv{};
        const int in = level & 1, outq = in ^ 1;
        lv.in_q0 = (const double2*)(base + o_q[in][0]); lv.in_q1 = (const double2*)(base + o_q[in][1]);
        lv.in_q2 = (const double2*)(base + o_q[in][2]); lv.in_q3 = (const double2*)(base + o_q[in][3]);
        lv.out_q0 = (double2*)(base + o_q[outq][0]); lv.out_q1 = (double2*)(base + o_q[outq][1]);
        lv.out_q2 = (double2*)(base + o_q[outq][2]); lv.out_q3 = (double2*)(base + o_q[outq][3]);
        lv.in_count = cnt + CNT_QUEUE + level;
        lv.in_count_back = cnt + CNT_QUEUE_BACK + level;
        lv.out_count = cnt + CNT_QUEUE + level + 1;
        lv.out_count_back = cnt + CNT_QUEUE_BACK + level + 1;
        lv.in_capacity = q_cap;
        lv.out_capacity = q_cap;
        lv.first_sample = 0;
        lv.n_primary = samples;
        lv.level = level;
        lv.hit_count = cnt + CNT_HITS + level;
        lv.hit_count_back = cnt + CNT_HITS_BACK + level;
        lv.shadow = base + o_shadow;
        lv.accum = accum;
        lv.stats = cnt + CNT_STATS;
        lv.overflow = overflow;
        return lv;
    };
    const int grid_full = dev.sm_count * 16;
    const int grid0 = (int)std::max<uint64_t>(std::min<uint64_t>((uint64_t)grid_full, (samples + RTC_BLOCK - 1) / RTC_BLOCK), 1);
    const int feat = s->feat;
    for (int level = 0; level <= o.fuel; ++level)
        rtcdev::launch_trace(feat, level == 0, level == 0 ? grid0 : grid_full, st, s->dev, f, level_params(level), hits);
    rtcdev::launch_shadow(feat, grid_full, st, s->dev, f, level_params(0), hits);
    for (int level = o.fuel; level >= 0; --level)
        rtcdev::launch_light(level == 0 ? grid0 : grid_full, st, s->dev, level_params(level), hits);
    rtcdev::launch_resolve(st, f, accum, 0, n_slots, frame, 1);
    CU(cudaGetLastError());
    // out_rgb may be host or device memory (unified addressing tells them apart)
    CU(cudaMemcpyAsync(out_rgb, frame, n_pixels * 3 * sizeof(double), cudaMemcpyDefault, st));
    unsigned long long c[CNT_TOTAL];
    int h_overflow = 0;
    CU(cudaMemcpyAsync(c, cnt, sizeof c, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(&h_overflow, overflow, sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (stats) {
        std::memset(stats, 0, sizeof *stats);
        auto level_rays = [&](int level) { return std::min<uint64_t>(c[CNT_QUEUE + level] + c[CNT_QUEUE_BACK + level], q_cap); };
        uint64_t refl = 0, refr = 0, hits_all = 0;
        if (s->feat & rtcdev::FEAT_TRANSPARENT) { refl = c[CNT_STATS + 1]; refr = c[CNT_STATS + 2]; }
        else for (int level = 1; level <= o.fuel; ++level) refl += level_rays(level);
        double ref = 0.0, mult = 1.0;  // R_ref: see rtc_render
        for (int level = 0; level <= o.fuel; ++level) {
            const uint64_t n_in = level == 0 ? c[CNT_STATS] : level_rays(level);
            const uint64_t h = c[CNT_HITS + level] + c[CNT_HITS_BACK + level];
            hits_all += h;
            ref += mult * ((double)n_in + (double)h * L);
            mult *= (double)L;
        }
        stats->rays_primary = c[CNT_STATS];
        stats->rays_shadow = hits_all * (uint64_t)L;
        stats->rays_reflected = refl;
        stats->rays_refracted = refr;
        stats->rays_ref = (uint64_t)ref;
        stats->surface_hits = hits_all;
        stats->shadow_queries_culled = c[CNT_STATS + 8];
        stats->kernel_launches = 2 * (uint64_t)(o.fuel + 1) + (L > 0 ? 1 : 0) + 1;
        stats->pixels = n_pixels;
        stats->overflow = h_overflow;
    }
    if (h_overflow) return rtcapi::fail(h_overflow == 2 ? "rtc_render_noalloc: ray queue overflow" : "rtc_render_noalloc: on-chip crossing list overflow (scene too deep for the device evaluator)");
    return 0;
}
