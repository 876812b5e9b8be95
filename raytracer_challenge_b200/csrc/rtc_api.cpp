// C ABI of include/rtc.h: scene upload and the wavefront driver (host side, CUDA runtime API only).
// rtc_render is the drop-in for the body of Image::par_render (/root/reference/src/image.rs:65-81).
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/rtc.h"
#include "render_kernels.h"
#include "scene_build.hpp"

namespace {

thread_local std::string g_error;
int fail(const std::string& msg) {
    g_error = msg;
    return 1;
}
#define CU(call)                                                                                             \
    do {                                                                                                     \
        cudaError_t e_ = (call);                                                                             \
        if (e_ != cudaSuccess) return fail(std::string(#call) + ": " + cudaGetErrorString(e_));              \
    } while (0)

// src/noise.rs:56-84 — Ken Perlin's reference permutation, stored twice
const uint8_t kPerm[256] = {
    151, 160, 137, 91,  90,  15,  131, 13,  201, 95,  96,  53,  194, 233, 7,   225, 140, 36,  103, 30,  69,  142, 8,   99,  37,  240,
    21,  10,  23,  190, 6,   148, 247, 120, 234, 75,  0,   26,  197, 62,  94,  252, 219, 203, 117, 35,  11,  32,  57,  177, 33,  88,
    237, 149, 56,  87,  174, 20,  125, 136, 171, 168, 68,  175, 74,  165, 71,  134, 139, 48,  27,  166, 77,  146, 158, 231, 83,  111,
    229, 122, 60,  211, 133, 230, 220, 105, 92,  41,  55,  46,  245, 40,  244, 102, 143, 54,  65,  25,  63,  161, 1,   216, 80,  73,
    209, 76,  132, 187, 208, 89,  18,  169, 200, 196, 135, 130, 116, 188, 159, 86,  164, 100, 109, 198, 173, 186, 3,   64,  52,  217,
    226, 250, 124, 123, 5,   202, 38,  147, 118, 126, 255, 82,  85,  212, 207, 206, 59,  227, 47,  16,  58,  17,  182, 189, 28,  42,
    223, 183, 170, 213, 119, 248, 152, 2,   44,  154, 163, 70,  221, 153, 101, 155, 167, 43,  172, 9,   129, 22,  39,  253, 19,  98,
    108, 110, 79,  113, 224, 232, 178, 185, 112, 104, 218, 246, 97,  228, 251, 34,  242, 193, 238, 210, 144, 12,  191, 179, 162, 241,
    81,  51,  145, 235, 249, 14,  239, 107, 49,  192, 214, 31,  181, 199, 106, 157, 184, 84,  204, 176, 115, 121, 50,  45,  127, 4,
    150, 254, 138, 236, 205, 93,  222, 114, 67,  29,  24,  72,  243, 141, 128, 195, 78,  66,  215, 61,  156, 180};

template <class T>
cudaError_t upload(const std::vector<T>& host, T** dev) {
    *dev = nullptr;
    const size_t bytes = std::max<size_t>(host.size(), 1) * sizeof(T);
    cudaError_t e = cudaMalloc((void**)dev, bytes);
    if (e != cudaSuccess) return e;
    if (!host.empty()) e = cudaMemcpy(*dev, host.data(), host.size() * sizeof(T), cudaMemcpyHostToDevice);
    return e;
}

struct Buffers {  // per-scene render scratch, grown on demand and reused across renders
    double2* q[2][4] = {{nullptr, nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr, nullptr}};
    uint64_t q_capacity = 0;
    double* accum = nullptr;
    uint64_t accum_samples = 0;
    double* out = nullptr;
    uint64_t out_pixels = 0;
    unsigned long long* counters = nullptr;  // [0..15] queue counts, [16..23] stats
    int* overflow = nullptr;
};

}  // namespace

struct rtc_scene {
    int device = 0;
    int sm_count = 0;
    int feat = 0;
    rtchost::HostScene host;
    rtcdev::Scene dev{};
    std::vector<void*> allocations;
    Buffers buf;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
};

extern "C" {

const char* rtc_last_error(void) { return g_error.c_str(); }

int rtc_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        g_error = std::string("cudaGetDeviceCount: ") + cudaGetErrorString(e);
        return 0;
    }
    return n;
}

void* rtc_alloc_host(uint64_t bytes) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) {
        g_error = "rtc_alloc_host: cudaHostAlloc failed";
        return nullptr;
    }
    return p;
}
void rtc_free_host(void* p) {
    if (p) cudaFreeHost(p);
}

void rtc_quantize(const double* rgb, uint8_t* out, uint64_t n) {
    // Color::clamp, src/color.rs:42-46: min(1).max(0) * 255, round half away from zero, saturating cast
    for (uint64_t i = 0; i < n; ++i) {
        const double v = std::round(std::fmax(std::fmin(rgb[i], 1.0), 0.0) * 255.0);
        out[i] = v != v ? 0 : (v >= 255.0 ? 255 : (v <= 0.0 ? 0 : (uint8_t)v));
    }
}

uint64_t rtc_shard_pixels(int32_t w, int32_t h, int32_t tile_w, int32_t tile_h, int32_t shard_index, int32_t shard_count,
                          uint32_t* out_indices) {
    if (tile_w <= 0) tile_w = 8;
    if (tile_h <= 0) tile_h = 4;
    if (shard_count <= 0) shard_count = 1;
    if (w <= 0 || h <= 0 || shard_index < 0 || shard_index >= shard_count) return 0;
    const uint64_t tiles_x = (uint64_t)(w + tile_w - 1) / tile_w, tiles_y = (uint64_t)(h + tile_h - 1) / tile_h;
    const uint64_t tiles = tiles_x * tiles_y;
    const uint64_t owned = tiles > (uint64_t)shard_index ? (tiles - shard_index + shard_count - 1) / shard_count : 0;
    const uint64_t per_tile = (uint64_t)tile_w * tile_h;
    if (out_indices) {
        for (uint64_t k = 0; k < owned; ++k) {
            const uint64_t tile = (uint64_t)shard_index + k * shard_count;
            const uint64_t tx = tile % tiles_x, ty = tile / tiles_x;
            for (uint64_t r = 0; r < per_tile; ++r) {
                const uint64_t px = tx * tile_w + r % tile_w, py = ty * tile_h + r / tile_w;
                out_indices[k * per_tile + r] = (px < (uint64_t)w && py < (uint64_t)h) ? (uint32_t)(py * w + px) : 0xffffffffu;
            }
        }
    }
    return owned * per_tile;
}

static int64_t host_stat(const rtchost::HostScene& h, int key) {
    switch (key) {
        case RTC_STAT_PRIMS: return (int64_t)h.prims.size();
        case RTC_STAT_BVH_NODES: return h.n_bvh_items ? (int64_t)h.bvh.size() : 0;
        case RTC_STAT_BVH_DEPTH: return h.bvh_depth;
        case RTC_STAT_UNBOUNDED: return (int64_t)h.unbounded.size();
        case RTC_STAT_GATES_EXACT: return h.n_gates_exact;
        case RTC_STAT_GATES_CONSERVATIVE: return h.n_gates_conservative;
        case RTC_STAT_GATES_ALWAYS_PASS: return h.n_gates_always_pass;
        case RTC_STAT_CSG_ROOTS: return (int64_t)h.csg_roots.size();
        case RTC_STAT_XFORMS: return (int64_t)h.xforms.size();
        case RTC_STAT_BVH_ITEMS: return h.n_bvh_items;
        case RTC_STAT_FEATURES: {
            bool transparent = false;
            for (auto& m : h.materials) transparent = transparent || m.transparency != 0.0;
            return (transparent ? rtcdev::FEAT_TRANSPARENT : 0) | (!h.csg_roots.empty() ? rtcdev::FEAT_CSG : 0) |
                   (!h.gates.empty() ? rtcdev::FEAT_GATES : 0);
        }
        default: return -1;
    }
}

int64_t rtc_scene_stat(const rtc_scene* s, int key) { return s ? host_stat(s->host, key) : -1; }

int rtc_scene_validate(const rtc_scene_desc* desc, int64_t* stats16) {
    if (!desc) return fail("rtc_scene_validate: null argument");
    rtchost::HostScene h;
    const std::string err = rtchost::build_host_scene(*desc, h);
    if (!err.empty()) return fail("rtc_scene_validate: " + err);
    if (stats16)
        for (int k = 0; k < 16; ++k) stats16[k] = host_stat(h, k);
    return 0;
}

void rtc_scene_destroy(rtc_scene* s) {
    if (!s) return;
    cudaSetDevice(s->device);
    for (void* p : s->allocations) cudaFree(p);
    for (int a = 0; a < 2; ++a)
        for (int b = 0; b < 4; ++b) cudaFree(s->buf.q[a][b]);
    cudaFree(s->buf.accum);
    cudaFree(s->buf.out);
    cudaFree(s->buf.counters);
    cudaFree(s->buf.overflow);
    for (auto& e : s->ev)
        if (e) cudaEventDestroy(e);
    if (s->stream) cudaStreamDestroy(s->stream);
    delete s;
}

int rtc_scene_create(const rtc_scene_desc* desc, int device, rtc_scene** out) {
    if (!desc || !out) return fail("rtc_scene_create: null argument");
    *out = nullptr;
    int n_dev = 0;
    CU(cudaGetDeviceCount(&n_dev));
    if (n_dev <= 0) return fail("rtc_scene_create: no CUDA device (this library has no CPU fallback)");
    if (device < 0 || device >= n_dev) return fail("rtc_scene_create: device ordinal out of range");
    rtc_scene* s = new rtc_scene;
    s->device = device;
    const std::string err = rtchost::build_host_scene(*desc, s->host);
    if (!err.empty()) {
        delete s;
        return fail("rtc_scene_create: " + err);
    }
    auto bail = [&](const std::string& m) {
        rtc_scene_destroy(s);
        return fail(m);
    };
    if (cudaSetDevice(device) != cudaSuccess) return bail("cudaSetDevice failed");
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return bail("cudaGetDeviceProperties failed");
    s->sm_count = prop.multiProcessorCount;

    const rtchost::HostScene& h = s->host;
    rtcdev::Scene& d = s->dev;
#define UP(field, vec, type)                                                                  \
    {                                                                                         \
        type* p_ = nullptr;                                                                   \
        cudaError_t e_ = upload(vec, &p_);                                                    \
        if (p_) s->allocations.push_back(p_);                                                 \
        if (e_ != cudaSuccess) return bail(std::string("upload " #field ": ") + cudaGetErrorString(e_)); \
        d.field = p_;                                                                         \
    }
    UP(prims, h.prims, rtcdev::Prim)
    UP(xforms, h.xforms, rtcdev::Xform)
    UP(matinvs, h.matinvs, rtcdev::MatInv)
    UP(tri_pe, h.tri_pe, double)
    UP(tri_n, h.tri_n, double)
    UP(ranges, h.ranges, double)
    UP(materials, h.materials, rtcdev::Material)
    UP(patterns, h.patterns, rtcdev::Pattern)
    UP(lights, h.lights, rtcdev::Light)
    UP(gates, h.gates, rtcdev::Gate)
    UP(bvh, h.bvh, rtcdev::BvhNode)
    UP(leaf_items, h.leaf_items, int32_t)
    UP(unbounded, h.unbounded, int32_t)
    UP(csg_ops, h.csg_ops, rtcdev::CsgOp)
    UP(csg_roots, h.csg_roots, rtcdev::CsgRoot)
#undef UP
    d.n_prims = (int32_t)h.prims.size();
    d.n_lights = (int32_t)h.lights.size();
    d.n_gates = (int32_t)h.gates.size();
    d.n_unbounded = (int32_t)h.unbounded.size();
    d.n_bvh_items = h.n_bvh_items;
    d.n_csg_roots = (int32_t)h.csg_roots.size();
    d.bvh_root = h.bvh_root;
    d.all_cast_shadow = h.all_cast_shadow;
    d.max_children = h.max_children;
    for (int k = 0; k < 3; ++k) { d.root_lo[k] = h.root_lo[k]; d.root_hi[k] = h.root_hi[k]; }

    s->feat = (int)host_stat(h, RTC_STAT_FEATURES);

    uint8_t perm[512];
    for (int i = 0; i < 512; ++i) perm[i] = kPerm[i & 255];
    rtcdev::upload_perm(perm);
    if (cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking) != cudaSuccess) return bail("cudaStreamCreate failed");
    for (auto& e : s->ev)
        if (cudaEventCreate(&e) != cudaSuccess) return bail("cudaEventCreate failed");
    if (cudaMalloc((void**)&s->buf.counters, 32 * sizeof(unsigned long long)) != cudaSuccess) return bail("cudaMalloc counters failed");
    if (cudaMalloc((void**)&s->buf.overflow, sizeof(int)) != cudaSuccess) return bail("cudaMalloc flag failed");
    if (cudaDeviceSynchronize() != cudaSuccess) return bail(std::string("scene upload: ") + cudaGetErrorString(cudaGetLastError()));
    *out = s;
    return 0;
}

static int ensure_buffers(rtc_scene* s, uint64_t q_capacity, uint64_t accum_samples, uint64_t out_pixels) {
    Buffers& b = s->buf;
    if (q_capacity > b.q_capacity) {
        for (int a = 0; a < 2; ++a)
            for (int c = 0; c < 4; ++c) {
                cudaFree(b.q[a][c]);
                b.q[a][c] = nullptr;
            }
        b.q_capacity = 0;
        for (int a = 0; a < 2; ++a)
            for (int c = 0; c < 4; ++c) CU(cudaMalloc((void**)&b.q[a][c], q_capacity * sizeof(double2)));
        b.q_capacity = q_capacity;
    }
    if (accum_samples > b.accum_samples) {
        cudaFree(b.accum);
        b.accum = nullptr;
        b.accum_samples = 0;
        CU(cudaMalloc((void**)&b.accum, accum_samples * 3 * sizeof(double)));
        b.accum_samples = accum_samples;
    }
    if (out_pixels > b.out_pixels) {
        cudaFree(b.out);
        b.out = nullptr;
        b.out_pixels = 0;
        CU(cudaMalloc((void**)&b.out, out_pixels * 3 * sizeof(double)));
        b.out_pixels = out_pixels;
    }
    return 0;
}

int rtc_render_device(rtc_scene* s, const rtc_camera* cam, const rtc_opts* o_in, double** out_dev_rgb, uint64_t* out_pixels,
                      rtc_stats* stats) {
    if (!s || !cam) return fail("rtc_render: null argument");
    rtc_opts o{};
    if (o_in) o = *o_in;
    else { o.fuel = 5; }
    if (o.spp_x <= 0) o.spp_x = 1;
    if (o.spp_y <= 0) o.spp_y = 1;
    if (o.shard_count <= 0) o.shard_count = 1;
    if (o.tile_w <= 0) o.tile_w = 8;
    if (o.tile_h <= 0) o.tile_h = 4;
    if (o.fuel < 0) return fail("rtc_render: negative fuel");
    if (o.fuel > 12) return fail("rtc_render: fuel > 12 not supported");
    if (o.shard_index < 0 || o.shard_index >= o.shard_count) return fail("rtc_render: shard_index out of range");
    if (cam->hsize <= 0 || cam->vsize <= 0) return fail("rtc_render: empty camera image");
    const bool whole = o.win_w <= 0 || o.win_h <= 0;
    const int32_t wx0 = whole ? 0 : o.win_x0, wy0 = whole ? 0 : o.win_y0;
    const int32_t ww = whole ? cam->hsize : o.win_w, wh = whole ? cam->vsize : o.win_h;
    if (wx0 < 0 || wy0 < 0 || (int64_t)wx0 + ww > cam->hsize || (int64_t)wy0 + wh > cam->vsize) return fail("rtc_render: window outside the image");
    if ((int64_t)ww * wh >= 0xffffffffll) return fail("rtc_render: window too large");
    for (int i = 12; i < 16; ++i)  // last row equals (0,0,0,1) only up to the rounding of the reference's inverse
        if (!(std::fabs(cam->transform_inv[i] - (i == 15 ? 1.0 : 0.0)) <= 1e-9)) return fail("rtc_render: camera transform is not affine");

    CU(cudaSetDevice(s->device));
    const auto t_begin = std::chrono::steady_clock::now();

    rtcdev::FrameParams f{};
    for (int i = 0; i < 12; ++i) f.cam_inv[i] = cam->transform_inv[i];
    f.pixel_size = cam->pixel_size;
    f.half_width = cam->half_width;
    f.half_height = cam->half_height;
    f.win_x0 = wx0; f.win_y0 = wy0; f.win_w = ww; f.win_h = wh;
    f.tile_w = o.tile_w; f.tile_h = o.tile_h;
    f.tiles_x = (ww + o.tile_w - 1) / o.tile_w;
    f.tiles_y = (wh + o.tile_h - 1) / o.tile_h;
    f.shard_index = o.shard_index; f.shard_count = o.shard_count;
    f.spp_x = o.spp_x; f.spp_y = o.spp_y; f.spp = o.spp_x * o.spp_y;
    f.fuel = o.fuel;
    f.flags = o.flags;

    const uint64_t per_tile = (uint64_t)o.tile_w * o.tile_h;
    const uint64_t n_slots = rtc_shard_pixels(ww, wh, o.tile_w, o.tile_h, o.shard_index, o.shard_count, nullptr);  // incl. tile padding
    const bool row_major = o.shard_count == 1;
    const uint64_t n_out = row_major ? (uint64_t)ww * wh : n_slots;

    // Batch size: whole tiles; queues sized for the worst-case fan-out (binary tree only if some material both
    // reflects and refracts), capped so both ping-pong queues stay within ~4 GiB.
    const int fan = s->dev.max_children >= 2 ? 2 : 1;
    uint64_t worst = 1;
    for (int k = 0; k < o.fuel && fan == 2; ++k) worst *= 2;
    const uint64_t max_records = (uint64_t)32 << 20;  // 32 Mi records * 64 B = 2 GiB per queue
    uint64_t batch_samples = std::max<uint64_t>(max_records / worst, per_tile * f.spp);
    batch_samples = std::min<uint64_t>(batch_samples, (uint64_t)8 << 20);
    uint64_t batch_pixels = std::max<uint64_t>(batch_samples / f.spp / per_tile, 1) * per_tile;
    batch_pixels = std::min(batch_pixels, std::max<uint64_t>(n_slots, per_tile));
    batch_samples = batch_pixels * f.spp;
    if (batch_samples >= 0xffffffffull) return fail("rtc_render: batch too large");
    const uint64_t q_cap = batch_samples * worst;
    if (int rc = ensure_buffers(s, q_cap, batch_samples, n_out)) return rc;
    Buffers& b = s->buf;

    cudaStream_t st = s->stream;
    CU(cudaMemsetAsync(b.counters, 0, 32 * sizeof(unsigned long long), st));
    CU(cudaMemsetAsync(b.overflow, 0, sizeof(int), st));
    CU(cudaEventRecord(s->ev[0], st));
    const int grid = s->sm_count * 8;
    uint64_t launches = 0;
    for (uint64_t first = 0; first < n_slots; first += batch_pixels) {
        const uint64_t npix = std::min(batch_pixels, n_slots - first);
        CU(cudaMemsetAsync(b.accum, 0, npix * f.spp * 3 * sizeof(double), st));
        CU(cudaMemsetAsync(b.counters, 0, 16 * sizeof(unsigned long long), st));
        for (int level = 0; level <= o.fuel; ++level) {
            rtcdev::LevelParams lv{};
            const int in = level & 1, outq = in ^ 1;
            lv.in_q0 = b.q[in][0]; lv.in_q1 = b.q[in][1]; lv.in_q2 = b.q[in][2]; lv.in_q3 = b.q[in][3];
            lv.out_q0 = b.q[outq][0]; lv.out_q1 = b.q[outq][1]; lv.out_q2 = b.q[outq][2]; lv.out_q3 = b.q[outq][3];
            lv.in_count = b.counters + level;
            lv.out_count = b.counters + level + 1;
            lv.in_capacity = b.q_capacity;
            lv.out_capacity = b.q_capacity;
            lv.first_sample = first * f.spp;
            lv.n_primary = npix * f.spp;
            lv.accum = b.accum;
            lv.stats = b.counters + 16;
            lv.overflow = b.overflow;
            rtcdev::launch_level(s->feat, level == 0, grid, st, s->dev, f, lv);
            launches++;
        }
        rtcdev::launch_resolve(st, f, b.accum, first, npix, b.out, row_major ? 1 : 0);
        launches++;
    }
    CU(cudaEventRecord(s->ev[1], st));
    CU(cudaGetLastError());
    unsigned long long h_counters[32];
    int h_overflow = 0;
    CU(cudaMemcpyAsync(h_counters, b.counters, sizeof h_counters, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(&h_overflow, b.overflow, sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, s->ev[0], s->ev[1]));
    if (stats) {
        std::memset(stats, 0, sizeof *stats);
        const uint64_t pushed_refl = h_counters[18], pushed_refr = h_counters[19];
        stats->rays_primary = h_counters[16] - std::min<uint64_t>(h_counters[16], pushed_refl + pushed_refr);
        stats->rays_shadow = h_counters[17];
        stats->rays_reflected = pushed_refl;
        stats->rays_refracted = pushed_refr;
        stats->rays_ref = h_counters[20];
        stats->kernel_launches = launches;
        stats->pixels = n_out;
        stats->ms_kernels = ms;
        stats->ms_trace_shade = ms;
        stats->overflow = h_overflow;
        stats->ms_total = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count();
    }
    if (h_overflow) return fail(h_overflow == 2 ? "rtc_render: ray queue overflow" : "rtc_render: on-chip crossing list overflow (scene too deep for the device evaluator)");
    if (out_dev_rgb) *out_dev_rgb = b.out;
    if (out_pixels) *out_pixels = n_out;
    return 0;
}

int rtc_render(rtc_scene* s, const rtc_camera* cam, const rtc_opts* o, double* out_rgb, rtc_stats* stats) {
    if (!out_rgb) return fail("rtc_render: null output buffer");
    const auto t_begin = std::chrono::steady_clock::now();
    double* dev = nullptr;
    uint64_t n = 0;
    rtc_stats local{};
    if (int rc = rtc_render_device(s, cam, o, &dev, &n, &local)) {
        if (stats) *stats = local;
        return rc;
    }
    CU(cudaEventRecord(s->ev[2], s->stream));
    CU(cudaMemcpyAsync(out_rgb, dev, n * 3 * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
    CU(cudaEventRecord(s->ev[3], s->stream));
    CU(cudaStreamSynchronize(s->stream));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, s->ev[2], s->ev[3]));
    local.ms_d2h = ms;
    local.ms_total = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count();
    if (stats) *stats = local;
    return 0;
}

}  // extern "C"
