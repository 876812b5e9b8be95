// C ABI of include/rtc.h: scene upload and the wavefront driver (host side, CUDA runtime API only).
// rtc_render is the drop-in for the body of Image::par_render (/root/reference/src/image.rs:65-81).
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/rtc.h"
#include "render_kernels.h"
#include "rtc_internal.hpp"
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

constexpr int MAX_PIPES = 8;
constexpr int MAX_DEVICES = 64;
// Device counters of one batch (unsigned long long each):
constexpr int CNT_QUEUE = 0, CNT_HITS = 16, CNT_QUEUE_BACK = 32, CNT_HITS_BACK = 48;  // per level: queue / hit-list fill, front and back bins
constexpr int CNT_STATS = 64;                                                           // LevelParams::stats
constexpr int COUNTERS_PER_BATCH = 80;

// Scratch of one pipeline (= one CUDA stream working through batches): ray queues (ping-pong), hit list, shadow
// bytes, sample accumulators.
struct Pipe {
    cudaStream_t stream = nullptr;  // the one in use by the current render: stream_flat or stream_staggered
    cudaStream_t stream_flat = nullptr, stream_staggered = nullptr;
    cudaEvent_t done = nullptr;
    double2* q[2][4] = {{nullptr, nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr, nullptr}};
    double2* hit[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    double* child_rgb = nullptr;  // colours the children of a hit returned (HitList::child_rgb)
    uint8_t* shadow = nullptr;
    double* accum = nullptr;
    uint64_t q_capacity = 0, hit_capacity = 0, child_doubles = 0, shadow_bytes = 0, accum_samples = 0;
};

// Per-device render context: streams, events and scratch memory.  It outlives scenes, so that the drop-in call
// sequence of Image::par_render (create scene, render once, destroy scene) does not pay for gigabytes of
// cudaMalloc / cudaFree per frame.  One render at a time per device (mutex); released by rtc_release_device_memory.
struct DeviceCtx {
    std::mutex mu;
    bool ready = false;
    int device = 0;
    int sm_count = 0;
    Pipe pipe[MAX_PIPES];
    cudaEvent_t ev[2] = {nullptr, nullptr};
    double* out = nullptr;
    uint64_t out_pixels = 0;
    unsigned long long* counters = nullptr;
    unsigned long long* counters_host = nullptr;  // pinned mirror
    uint64_t counter_batches = 0;
    int* overflow = nullptr;
    std::vector<cudaEvent_t> prof_events;  // RTC_FLAG_PROFILE: one event per launch boundary
    cudaStream_t copy_stream[MAX_PIPES] = {};  // device -> host copies of finished batches, one per pipeline: a pipeline's next
                                               // batch does not queue behind its previous copy, nor a copy behind another pipeline's
    cudaEvent_t copy_done = nullptr;
    std::vector<cudaEvent_t> batch_events; // "batch b is resolved": what its copy waits for
    void* spare_arena = nullptr;           // scene arena of the last destroyed scene, reused by the next create
    size_t spare_arena_bytes = 0;
    cudaEvent_t fork = nullptr;            // "the counters are cleared": what the other pipelines wait for
    // Frames that repeat — same scene layout, camera, options, output buffer, scratch — replay a captured CUDA graph: one
    // launch instead of ~60 kernel launches, 30 event operations and 10 copies enqueued one by one (the enqueueing alone
    // kept the last pipeline idle for the first ~150 us of a frame; at 8 GPUs a rank's share of the frame takes 1 ms).
    struct CachedGraph {
        std::vector<unsigned char> key;
        cudaGraphExec_t exec = nullptr;
        uint64_t last_used = 0;
    };
    CachedGraph graphs[4];
    uint64_t graph_clock = 0;
};
DeviceCtx g_ctx[MAX_DEVICES];

void free_pipe(Pipe& p) {
    for (int a = 0; a < 2; ++a)
        for (int b = 0; b < 4; ++b) { cudaFree(p.q[a][b]); p.q[a][b] = nullptr; }
    for (auto& h : p.hit) { cudaFree(h); h = nullptr; }
    cudaFree(p.child_rgb); p.child_rgb = nullptr;
    cudaFree(p.shadow); p.shadow = nullptr;
    cudaFree(p.accum); p.accum = nullptr;
    p.q_capacity = p.hit_capacity = p.child_doubles = p.shadow_bytes = p.accum_samples = 0;
}

// rtc_tuning (include/rtc.h): process-wide, set through rtc_set_tuning only
std::mutex g_tuning_mu;
rtc_tuning g_tuning{};
rtc_tuning tuning_snapshot() {
    std::lock_guard<std::mutex> lock(g_tuning_mu);
    return g_tuning;
}
int knob(int32_t value, int fallback, int lo, int hi) { return value > 0 ? std::min(std::max((int)value, lo), hi) : fallback; }

int ctx_init(DeviceCtx& c, int device) {  // caller holds c.mu and has set the device
    if (c.ready) return 0;
    c.device = device;
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    c.sm_count = prop.multiProcessorCount;
    // Two streams per pipeline.  Renders that leave the frame in HBM use equal priorities: the batches share the SMs
    // evenly and finish together, which is the fastest way through the kernels.  Renders that copy each finished
    // batch to the host use descending priorities: batch 0 is then done (and on its way over PCIe) while the other
    // pipelines are still tracing, instead of every copy queueing up behind the last kernel.
    int prio_least = 0, prio_greatest = 0;
    CU(cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest));  // numerically lower = higher priority
    int pipe_index = 0;
    for (Pipe& p : c.pipe) {
        CU(cudaStreamCreateWithPriority(&p.stream_flat, cudaStreamNonBlocking, prio_least));
        CU(cudaStreamCreateWithPriority(&p.stream_staggered, cudaStreamNonBlocking, std::min(prio_least, prio_greatest + pipe_index)));
        p.stream = p.stream_flat;
        CU(cudaEventCreateWithFlags(&p.done, cudaEventDisableTiming));
        pipe_index++;
    }
    for (auto& e : c.ev) CU(cudaEventCreate(&e));
    for (auto& cs : c.copy_stream) CU(cudaStreamCreateWithPriority(&cs, cudaStreamNonBlocking, prio_greatest));
    CU(cudaEventCreateWithFlags(&c.copy_done, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&c.fork, cudaEventDisableTiming));
    CU(cudaMalloc((void**)&c.overflow, sizeof(int)));
    uint8_t perm[512];
    for (int i = 0; i < 512; ++i) perm[i] = kPerm[i & 255];
    rtcdev::upload_perm(perm);
    CU(cudaGetLastError());
    c.ready = true;
    return 0;
}

}  // namespace

namespace rtcapi {
int fail(const std::string& msg) { return ::fail(msg); }
}  // namespace rtcapi

extern "C" {

const char* rtc_last_error(void) { return g_error.c_str(); }

void rtc_set_tuning(const rtc_tuning* t) {
    std::lock_guard<std::mutex> lock(g_tuning_mu);
    g_tuning = t ? *t : rtc_tuning{};
}
void rtc_get_tuning(rtc_tuning* out) {
    if (out) *out = tuning_snapshot();
}

int rtc_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        g_error = std::string("cudaGetDeviceCount: ") + cudaGetErrorString(e);
        return 0;
    }
    return n;
}

// Page-locked host buffers are pooled: cudaHostAlloc / cudaFreeHost of a 199 MB frame cost tens of milliseconds each, and
// Image::par_render (src/image.rs:65-81) needs one frame-sized buffer per call.  rtc_free_host parks a buffer, the next
// rtc_alloc_host of a similar size takes it back; rtc_release_host_memory returns the parked ones to the driver.
namespace {
struct HostBlock { void* p; uint64_t bytes; };
std::mutex g_host_mu;
std::vector<HostBlock> g_host_live, g_host_free;
constexpr uint64_t HOST_POOL_MAX_BYTES = 4ull << 30;  // parked bytes beyond this are freed at once
}  // namespace

void* rtc_alloc_host(uint64_t bytes) {
    if (!bytes) bytes = 1;
    {
        std::lock_guard<std::mutex> lock(g_host_mu);
        size_t best = g_host_free.size();
        for (size_t i = 0; i < g_host_free.size(); ++i)  // smallest parked block that fits without wasting more than half of itself
            if (g_host_free[i].bytes >= bytes && g_host_free[i].bytes <= 2 * bytes + 4096 && (best == g_host_free.size() || g_host_free[i].bytes < g_host_free[best].bytes)) best = i;
        if (best != g_host_free.size()) {
            const HostBlock b = g_host_free[best];
            g_host_free.erase(g_host_free.begin() + (long)best);
            g_host_live.push_back(b);
            return b.p;
        }
    }
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes, cudaHostAllocPortable) != cudaSuccess) {
        g_error = "rtc_alloc_host: cudaHostAlloc failed";
        return nullptr;
    }
    std::lock_guard<std::mutex> lock(g_host_mu);
    g_host_live.push_back({p, bytes});
    return p;
}
void rtc_free_host(void* p) {
    if (!p) return;
    std::lock_guard<std::mutex> lock(g_host_mu);
    for (size_t i = 0; i < g_host_live.size(); ++i)
        if (g_host_live[i].p == p) {
            const HostBlock b = g_host_live[i];
            g_host_live.erase(g_host_live.begin() + (long)i);
            uint64_t parked = b.bytes;
            for (const HostBlock& f : g_host_free) parked += f.bytes;
            if (parked <= HOST_POOL_MAX_BYTES) g_host_free.push_back(b);
            else cudaFreeHost(b.p);
            return;
        }
    cudaFreeHost(p);  // (not one of ours: e.g. allocated before a rtc_release_host_memory raced; let the driver decide)
}
void rtc_release_host_memory(void) {
    std::lock_guard<std::mutex> lock(g_host_mu);
    for (const HostBlock& f : g_host_free) cudaFreeHost(f.p);
    g_host_free.clear();
}
int rtc_host_register(void* p, uint64_t bytes) {
    if (!p || !bytes) return fail("rtc_host_register: null argument");
    CU(cudaHostRegister(p, bytes, cudaHostRegisterPortable));
    return 0;
}
int rtc_host_unregister(void* p) {
    if (!p) return fail("rtc_host_unregister: null argument");
    CU(cudaHostUnregister(p));
    return 0;
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
                   (!h.gates.empty() || !h.guards.empty() ? rtcdev::FEAT_GATES : 0);
        }
        case RTC_STAT_DEVICE_BYTES: return h.device_bytes;
        case RTC_STAT_CONE_GUARDS: return (int64_t)h.guards.size();
        default: return -1;
    }
}

int64_t rtc_scene_stat(const rtc_scene* s, int key) {
    if (!s || !s->host) return -1;
    if (s->bvh_depth >= 0 && key == RTC_STAT_BVH_DEPTH) return s->bvh_depth;  // device-side build: what the device reported
    if (s->bvh_nodes >= 0 && key == RTC_STAT_BVH_NODES) return s->bvh_nodes;
    if (key == RTC_STAT_FEATURES) return s->feat;  // (an uploaded scene: incl. the kernel variant chosen, e.g. FEAT_ONEPASS)
    return host_stat(*s->host, key);
}

int rtc_scene_validate(const rtc_scene_desc* desc, int64_t* stats16) {
    if (!desc) return fail("rtc_scene_validate: null argument");
    rtchost::HostScene h;
    const rtc_tuning tune = tuning_snapshot();
    const std::string err = rtchost::build_host_scene(*desc, h, &tune);
    if (!err.empty()) return fail("rtc_scene_validate: " + err);
    if (stats16)
        for (int k = 0; k < 16; ++k) stats16[k] = host_stat(h, k);
    return 0;
}

void rtc_scene_destroy(rtc_scene* s) {
    if (!s) return;
    cudaSetDevice(s->device);
    if (s->arena) {
        // cudaFree synchronises the device and was seen to take hundreds of ms now and then; the arena is parked in
        // the device context instead, where the next rtc_scene_create (the next frame of a par_render loop) picks it up
        DeviceCtx& c = g_ctx[s->device];
        std::lock_guard<std::mutex> lock(c.mu);
        void* drop = s->arena;
        if (s->arena_bytes > c.spare_arena_bytes) {
            drop = c.spare_arena;
            c.spare_arena = s->arena;
            c.spare_arena_bytes = s->arena_bytes;
        }
        if (drop) cudaFree(drop);
    }
    delete s;
}

void rtc_release_device_memory(int device) {
    if (device < 0 || device >= MAX_DEVICES) return;
    DeviceCtx& c = g_ctx[device];
    std::lock_guard<std::mutex> lock(c.mu);
    if (!c.ready) return;
    cudaSetDevice(device);
    cudaDeviceSynchronize();
    for (Pipe& p : c.pipe) free_pipe(p);
    for (auto& g : c.graphs) {
        if (g.exec) cudaGraphExecDestroy(g.exec);
        g.exec = nullptr;
        g.key.clear();
    }
    cudaFree(c.out); c.out = nullptr; c.out_pixels = 0;
    cudaFree(c.counters); c.counters = nullptr;
    cudaFreeHost(c.counters_host); c.counters_host = nullptr;
    c.counter_batches = 0;
    cudaFree(c.spare_arena); c.spare_arena = nullptr; c.spare_arena_bytes = 0;
}

}  // extern "C"

// Every array of the flattened scene in one staging blob (256-byte aligned sections): one cudaMalloc + one copy per device.
struct PackedScene {
    // Page-locked staging from the library's pool (rtc_alloc_host parks and reuses it): the upload is one DMA straight from
    // it instead of the driver's bounce-buffer path for pageable memory, and nothing is zero-filled first.
    struct Blob {
        unsigned char* p = nullptr;
        size_t used = 0, cap = 0;
        bool pooled = false;
        Blob() = default;
        Blob(const Blob&) = delete;
        Blob& operator=(const Blob&) = delete;
        ~Blob() { if (pooled) rtc_free_host(p); else std::free(p); }
        void allocate(size_t bytes) {
            p = (unsigned char*)rtc_alloc_host(bytes);
            pooled = p != nullptr;
            if (!p) p = (unsigned char*)std::malloc(bytes);
            cap = p ? bytes : 0;
        }
        unsigned char* data() const { return p; }
        size_t size() const { return used; }
    } blob;
    size_t o_prims, o_xforms, o_matinvs, o_tri_pe, o_tri_n, o_ranges, o_materials, o_patterns, o_lights, o_gates, o_bvh, o_leaf_items,
        o_leaf_tri, o_unbounded, o_csg_ops, o_csg_roots, o_guards;
    // device-side BVH build only (HostScene::device_build): staging + scratch sections
    size_t upload_bytes = 0, total_bytes = 0;  // host -> device copy / arena size
    size_t o_item_ids = 0, o_box32 = 0, o_keys = 0, o_keys_alt = 0, o_radix_hist = 0, o_child = 0, o_parent = 0, o_range = 0, o_nbox = 0, o_counter = 0, o_dense = 0, o_result = 0;
};
static bool pack_scene(rtchost::HostScene& h, PackedScene& p) {
    bool fits = true;
    auto place = [&](const void* src, size_t bytes) -> size_t {
        const size_t off = (p.blob.used + 255) & ~(size_t)255;
        const size_t end = off + std::max<size_t>(bytes, 16);
        if (end > p.blob.cap) { fits = false; return off; }  // (the capacity below is the sum of the padded sections)
        if (bytes) std::memcpy(p.blob.p + off, src, bytes);
        else std::memset(p.blob.p + off, 0, 16);
        p.blob.used = end;
        return off;
    };
#define PLACE(vec) place((vec).data(), (vec).size() * sizeof((vec)[0]))
#define BYTES(vec) (((vec).size() * sizeof((vec)[0]) + 16 + 255) & ~(size_t)255)
    p.blob.allocate(BYTES(h.prims) + BYTES(h.xforms) + BYTES(h.matinvs) + BYTES(h.tri_pe) + BYTES(h.tri_n) + BYTES(h.ranges) + BYTES(h.materials) +
                   BYTES(h.patterns) + BYTES(h.lights) + BYTES(h.gates) + BYTES(h.bvh) + BYTES(h.leaf_items) + BYTES(h.leaf_tri) + BYTES(h.unbounded) +
                   BYTES(h.csg_ops) + BYTES(h.csg_roots) + BYTES(h.guards) + 4096 +
                   (h.device_build ? BYTES(h.item_ids) + BYTES(h.item_box32) + 4096 : 0));  // one allocation for every section
#undef BYTES
    if (!p.blob.p) return false;
    p.o_prims = PLACE(h.prims); p.o_xforms = PLACE(h.xforms); p.o_matinvs = PLACE(h.matinvs); p.o_tri_pe = PLACE(h.tri_pe);
    p.o_tri_n = PLACE(h.tri_n); p.o_ranges = PLACE(h.ranges); p.o_materials = PLACE(h.materials); p.o_patterns = PLACE(h.patterns);
    p.o_lights = PLACE(h.lights); p.o_gates = PLACE(h.gates); p.o_bvh = PLACE(h.bvh); p.o_leaf_items = PLACE(h.leaf_items);
    p.o_leaf_tri = PLACE(h.leaf_tri); p.o_unbounded = PLACE(h.unbounded); p.o_csg_ops = PLACE(h.csg_ops); p.o_csg_roots = PLACE(h.csg_roots);
    p.o_guards = PLACE(h.guards);
    p.upload_bytes = p.blob.size();
    p.total_bytes = p.blob.size();
    if (h.device_build) {  // staging for bvh_build.cu, then space the device fills itself (never uploaded)
        const size_t n = (size_t)h.n_bvh_items;
        p.o_item_ids = PLACE(h.item_ids);
        p.o_box32 = PLACE(h.item_box32);
        p.upload_bytes = p.blob.size();
        size_t at = p.upload_bytes;
        auto reserve = [&](size_t bytes) -> size_t {
            const size_t off = (at + 255) & ~(size_t)255;
            at = off + std::max<size_t>(bytes, 16);
            return off;
        };
        p.o_bvh = reserve(n * sizeof(rtcdev::BvhNode));
        p.o_leaf_items = reserve(n * sizeof(rtcdev::LeafItem));
        p.o_leaf_tri = reserve(n * 80);
        p.o_keys = reserve(n * 8);
        if (n > (size_t)rtcdev::LBVH_ONE_BLOCK_ITEMS) {
            p.o_keys_alt = reserve(n * 8);
            p.o_radix_hist = reserve(16 * ((n + 255) / 256) * 4);
        }
        p.o_child = reserve(n * 8);
        p.o_parent = reserve(2 * n * 4);
        p.o_range = reserve(n * 8);
        p.o_nbox = reserve(2 * n * 16);
        p.o_counter = reserve(n * 4);
        p.o_dense = reserve(n * 4);
        p.o_result = reserve(16);
        p.total_bytes = at;
    }
#undef PLACE
    h.device_bytes = (int64_t)p.total_bytes;
    return fits;
}

// Device half of rtc_scene_create: arena (reused from the device context when one is parked there), copy, pointers.
static int upload_scene(rtc_scene* s, const PackedScene& p, std::string& err) {
    const rtc_tuning tune = tuning_snapshot();
    const int device = s->device;
    if (cudaSetDevice(device) != cudaSuccess) { err = "cudaSetDevice failed"; return 1; }
    DeviceCtx& c = g_ctx[device];
    {
        std::lock_guard<std::mutex> lock(c.mu);
        if (ctx_init(c, device)) { err = g_error; return 1; }
        if (c.spare_arena && c.spare_arena_bytes >= p.total_bytes) {
            s->arena = c.spare_arena;
            s->arena_bytes = c.spare_arena_bytes;
            c.spare_arena = nullptr;
            c.spare_arena_bytes = 0;
        }
    }
    if (!s->arena) {
        if (cudaMalloc(&s->arena, p.total_bytes) != cudaSuccess) { err = "cudaMalloc of the scene arena failed"; return 1; }
        s->arena_bytes = p.total_bytes;
    }
    if (cudaMemcpy(s->arena, p.blob.data(), p.upload_bytes, cudaMemcpyHostToDevice) != cudaSuccess) {
        err = std::string("scene upload failed: ") + cudaGetErrorString(cudaGetLastError());
        cudaFree(s->arena);
        s->arena = nullptr;
        return 1;
    }
    const rtchost::HostScene& h = *s->host;
    rtcdev::Scene& d = s->dev;
    unsigned char* base = (unsigned char*)s->arena;
    d.prims = (const rtcdev::Prim*)(base + p.o_prims);
    d.xforms = (const rtcdev::Xform*)(base + p.o_xforms);
    d.matinvs = (const rtcdev::MatInv*)(base + p.o_matinvs);
    d.tri_pe = (const double*)(base + p.o_tri_pe);
    d.tri_n = (const double*)(base + p.o_tri_n);
    d.ranges = (const double*)(base + p.o_ranges);
    d.materials = (const rtcdev::Material*)(base + p.o_materials);
    d.patterns = (const rtcdev::Pattern*)(base + p.o_patterns);
    d.lights = (const rtcdev::Light*)(base + p.o_lights);
    d.gates = (const rtcdev::Gate*)(base + p.o_gates);
    d.bvh = (const rtcdev::BvhNode*)(base + p.o_bvh);
    d.leaf_items = (const rtcdev::LeafItem*)(base + p.o_leaf_items);
    d.leaf_tri = (const double*)(base + p.o_leaf_tri);
    d.unbounded = (const int32_t*)(base + p.o_unbounded);
    d.csg_ops = (const rtcdev::CsgOp*)(base + p.o_csg_ops);
    d.csg_roots = (const rtcdev::CsgRoot*)(base + p.o_csg_roots);
    d.guards = (const rtcdev::ConeGuard*)(base + p.o_guards);
    d.n_guards = (int32_t)h.guards.size();
    d.n_prims = (int32_t)h.prims.size();
    d.n_lights = (int32_t)h.lights.size();
    d.n_gates = (int32_t)h.gates.size();
    d.n_unbounded = (int32_t)h.unbounded.size();
    d.n_bvh_items = h.n_bvh_items;
    d.n_csg_roots = (int32_t)h.csg_roots.size();
    d.bvh_root = h.bvh_root;
    // host-built trees are numbered breadth-first at the top (Builder::compact): the kernels stage that prefix in shared memory
    d.n_top = (!h.device_build && h.n_bvh_items > 0 && h.bvh_root >= 0) ? (int32_t)std::min<size_t>(h.bvh.size(), (size_t)rtcdev::MAX_TOP_NODES) : 0;
    d.all_cast_shadow = h.all_cast_shadow;
    d.max_children = h.max_children;
    for (int k = 0; k < 3; ++k) { d.root_lo[k] = h.root_lo[k]; d.root_hi[k] = h.root_hi[k]; }
    if (h.device_build) {  // build the BVH where the scene now lives (bvh_build.cu); the host only staged the item boxes
        rtcdev::LbvhArgs a{};
        a.n = h.n_bvh_items;
        // (sweeps: one item per leaf where an item test is a general f64 primitive intersection, dearer than a node visit; two for
        // meshes since the triangles-only leaf path — teapot 5.08 -> 4.93 ms, 101 120-triangle teapot 6.85 -> 6.36 ms, r02z)
        a.leaf_max = knob(tune.lbvh_leaf, h.all_items_triangles ? 2 : 1, 1, 4);
        a.box = (const float4*)(base + p.o_box32);
        a.item_ids = (const int32_t*)(base + p.o_item_ids);
        a.prims = d.prims;
        a.tri_pe = d.tri_pe;
        a.leaf_items = (rtcdev::LeafItem*)(base + p.o_leaf_items);
        a.leaf_tri = (double*)(base + p.o_leaf_tri);
        a.bvh = (rtcdev::BvhNode*)(base + p.o_bvh);
        a.keys = (unsigned long long*)(base + p.o_keys);
        a.keys_alt = (unsigned long long*)(base + p.o_keys_alt);
        a.radix_hist = (uint32_t*)(base + p.o_radix_hist);
        a.child = (int2*)(base + p.o_child);
        a.parent = (int*)(base + p.o_parent);
        a.range = (int2*)(base + p.o_range);
        a.nbox = (float4*)(base + p.o_nbox);
        a.counter = (int*)(base + p.o_counter);
        a.dense = (int*)(base + p.o_dense);
        a.result = (int*)(base + p.o_result);
        for (int k = 0; k < 3; ++k) {
            a.root_lo[k] = (float)h.root_lo[k];
            const double ext = h.root_hi[k] - h.root_lo[k];
            a.root_inv_extent[k] = ext > 0.0 ? (float)(1.0 / ext) : 0.0f;
        }
        int result[4] = {0, 0, 0, 0};
        const int rc = rtcdev::build_lbvh(a, nullptr);
        if (rc != 0 || cudaMemcpy(result, a.result, sizeof result, cudaMemcpyDeviceToHost) != cudaSuccess) {
            err = std::string("device BVH build failed: ") + cudaGetErrorString(rc != 0 ? (cudaError_t)rc : cudaGetLastError());
            cudaFree(s->arena);
            s->arena = nullptr;
            return 1;
        }
        if (result[1] + 2 > rtcdev::BVH_STACK) {
            err = "device BVH deeper than the traversal stack";
            cudaFree(s->arena);
            s->arena = nullptr;
            return 1;
        }
        d.bvh_root = result[0];
        s->bvh_depth = result[1];
        s->bvh_nodes = result[2];
    }
    // always-tested list = ungated planes only -> keep them in the kernel parameters (device_scene.h, FastPlane)
    d.n_fast_planes = 0;
    d.plain_only = 1;
    for (const rtcdev::Material& m : h.materials) d.plain_only = d.plain_only && (m.flags & 1) != 0;
    if (tune.no_plain_shade) d.plain_only = 0;
    std::memset(d.fast_planes, 0, sizeof d.fast_planes);
    if (!h.unbounded.empty() && h.unbounded.size() <= (size_t)rtcdev::MAX_FAST_PLANES && !tune.no_fast_planes) {
        bool ok = true;
        for (int32_t id : h.unbounded) ok = ok && id >= 0 && (h.prims[id].kind_flags & 0xff) == RTC_PLANE && h.prims[id].gate_mask == 0u && h.prims[id].gate_mask_hi == 0u;
        if (ok) {
            for (int32_t id : h.unbounded) {
                rtcdev::FastPlane& fp = d.fast_planes[d.n_fast_planes++];
                for (int k = 0; k < 4; ++k) fp.row[k] = h.xforms[h.prims[id].xform].inv[4 + k];
                fp.prim = id;
            }
        }
    }
    s->feat = (int)host_stat(h, RTC_STAT_FEATURES);
    if ((s->feat & rtcdev::FEAT_TRANSPARENT) && h.n_bvh_items <= rtcdev::ONEPASS_MAX_ITEMS && !tune.two_pass) s->feat |= rtcdev::FEAT_ONEPASS;
    if (s->feat == 0 && h.n_bvh_items > 0 && h.all_items_triangles) s->feat |= rtcdev::FEAT_TRIS;  // (only the plain kernel has a triangles-only variant)
    return 0;
}

extern "C" {

// One host build (validation, flattening, BVH), one upload per device.  rtc_scene_create is the n = 1 case.
int rtc_scene_create_multi(const rtc_scene_desc* desc, const int* devices, int n, rtc_scene** out) {
    if (!desc || !devices || !out || n <= 0) return fail("rtc_scene_create: null argument");
    for (int i = 0; i < n; ++i) out[i] = nullptr;
    int n_dev = 0;
    CU(cudaGetDeviceCount(&n_dev));
    if (n_dev <= 0) return fail("rtc_scene_create: no CUDA device (this library has no CPU fallback)");
    for (int i = 0; i < n; ++i) {
        if (devices[i] < 0 || devices[i] >= n_dev || devices[i] >= MAX_DEVICES) return fail("rtc_scene_create: device ordinal out of range");
        for (int j = 0; j < i; ++j)
            if (devices[j] == devices[i]) return fail("rtc_scene_create_multi: device listed twice");
    }
    const auto t_build0 = std::chrono::steady_clock::now();
    auto host = std::make_shared<rtchost::HostScene>();
    const rtc_tuning tune = tuning_snapshot();
    const std::string err = rtchost::build_host_scene(*desc, *host, &tune);
    if (!err.empty()) return fail("rtc_scene_create: " + err);
    const auto t_build1 = std::chrono::steady_clock::now();
    PackedScene packed;
    int caller_device = -1;
    cudaGetDevice(&caller_device);
    cudaSetDevice(devices[0]);  // (the staging buffer is page-locked by the first device's context, portable to the others)
    const bool packed_ok = pack_scene(*host, packed);
    if (n > 1 && caller_device >= 0) cudaSetDevice(caller_device);  // (n = 1 renders from this thread: upload_scene selects the device anyway)
    if (!packed_ok) return fail("rtc_scene_create: out of host memory for the upload staging");
    std::vector<std::unique_ptr<rtc_scene>> made((size_t)n);
    std::vector<std::string> errors((size_t)n);
    std::vector<int> rcs((size_t)n, 0);
    auto one = [&](int i) {
        made[(size_t)i].reset(new rtc_scene);
        made[(size_t)i]->device = devices[i];
        made[(size_t)i]->host = host;
        rcs[(size_t)i] = upload_scene(made[(size_t)i].get(), packed, errors[(size_t)i]);
    };
    if (n == 1) one(0);
    else {
        std::vector<std::thread> workers;
        for (int i = 0; i < n; ++i) workers.emplace_back(one, i);
        for (auto& w : workers) w.join();
    }
    for (int i = 0; i < n; ++i)
        if (rcs[(size_t)i]) {
            const std::string why = errors[(size_t)i];
            for (auto& m : made)
                if (m && m->arena) { cudaSetDevice(m->device); cudaFree(m->arena); }
            return fail("rtc_scene_create: device " + std::to_string(devices[i]) + ": " + why);
        }
    if (std::getenv("RTC_TRACE")) {
        auto ms = [](auto a, auto b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
        std::fprintf(stderr, "rtc_scene_create: host build %.3f ms, pack + upload%s to %d device(s) %.3f ms (%zu bytes each)\n", ms(t_build0, t_build1),
                     host->device_build ? " + device BVH build" : "", n, ms(t_build1, std::chrono::steady_clock::now()), packed.upload_bytes);
    }
    for (int i = 0; i < n; ++i) out[i] = made[(size_t)i].release();
    return 0;
}

int rtc_scene_create(const rtc_scene_desc* desc, int device, rtc_scene** out) {
    if (!desc || !out) return fail("rtc_scene_create: null argument");
    return rtc_scene_create_multi(desc, &device, 1, out);
}

static int ensure_pipe(Pipe& p, uint64_t q_capacity, uint64_t hit_capacity, uint64_t child_doubles, uint64_t shadow_bytes, uint64_t accum_samples) {
    if (q_capacity > p.q_capacity) {
        for (int a = 0; a < 2; ++a)
            for (int c = 0; c < 4; ++c) {
                cudaFree(p.q[a][c]); p.q[a][c] = nullptr;
            }
        p.q_capacity = 0;
        for (int a = 0; a < 2; ++a)
            for (int c = 0; c < 4; ++c) CU(cudaMalloc((void**)&p.q[a][c], q_capacity * sizeof(double2)));
        p.q_capacity = q_capacity;
    }
    if (hit_capacity > p.hit_capacity) {
        for (auto& h : p.hit) { cudaFree(h); h = nullptr; }
        p.hit_capacity = 0;
        for (auto& h : p.hit) CU(cudaMalloc((void**)&h, hit_capacity * sizeof(double2)));
        p.hit_capacity = hit_capacity;
    }
    if (child_doubles > p.child_doubles) {
        cudaFree(p.child_rgb); p.child_rgb = nullptr; p.child_doubles = 0;
        CU(cudaMalloc((void**)&p.child_rgb, child_doubles * sizeof(double)));
        p.child_doubles = child_doubles;
    }
    if (shadow_bytes > p.shadow_bytes) {
        cudaFree(p.shadow); p.shadow = nullptr; p.shadow_bytes = 0;
        CU(cudaMalloc((void**)&p.shadow, shadow_bytes));
        p.shadow_bytes = shadow_bytes;
    }
    if (accum_samples > p.accum_samples) {
        cudaFree(p.accum); p.accum = nullptr; p.accum_samples = 0;
        CU(cudaMalloc((void**)&p.accum, accum_samples * 3 * sizeof(double)));
        p.accum_samples = accum_samples;
    }
    return 0;
}

// The body shared by rtc_render (host_out != null: each finished batch is copied to the host while later batches
// are still being traced) and rtc_render_device (host_out == null).
static int render_impl(rtc_scene* s, const rtc_camera* cam, const rtc_opts* o_in, double* host_out, double** out_dev_rgb,
                       uint64_t* out_pixels, rtc_stats* stats) {
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
    if ((int64_t)o.spp_x * o.spp_y > 4096) return fail("rtc_render: more than 4096 samples per pixel");
    const bool whole = o.win_w <= 0 || o.win_h <= 0;
    const int32_t wx0 = whole ? 0 : o.win_x0, wy0 = whole ? 0 : o.win_y0;
    const int32_t ww = whole ? cam->hsize : o.win_w, wh = whole ? cam->vsize : o.win_h;
    if (wx0 < 0 || wy0 < 0 || (int64_t)wx0 + ww > cam->hsize || (int64_t)wy0 + wh > cam->vsize) return fail("rtc_render: window outside the image");
    for (int i = 12; i < 16; ++i)  // last row equals (0,0,0,1) only up to the rounding of the reference's inverse
        if (!(std::fabs(cam->transform_inv[i] - (i == 15 ? 1.0 : 0.0)) <= 1e-9)) return fail("rtc_render: camera transform is not affine");

    const rtc_tuning tune = tuning_snapshot();
    CU(cudaSetDevice(s->device));
    DeviceCtx& ctx = g_ctx[s->device];
    std::lock_guard<std::mutex> lock(ctx.mu);
    if (int rc = ctx_init(ctx, s->device)) return rc;
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
    f.two_bins = (s->feat & rtcdev::FEAT_CSG) != 0;

    const uint64_t per_tile = (uint64_t)o.tile_w * o.tile_h;
    const uint64_t all_slots = (uint64_t)f.tiles_x * f.tiles_y * per_tile;  // padded pixel slots of the whole window
    if (all_slots * (uint64_t)f.spp >= 0x7fffffffull) return fail("rtc_render: window too large (pixel slots x samples must stay below 2^31)");
    const uint64_t n_slots = rtc_shard_pixels(ww, wh, o.tile_w, o.tile_h, o.shard_index, o.shard_count, nullptr);  // incl. tile padding
    const bool row_major = o.shard_count == 1;
    const bool scatter_rows = !row_major && (o.flags & RTC_FLAG_SCATTER_ROWS) != 0;
    if (scatter_rows && o.tile_w != ww) return fail("rtc_render: RTC_FLAG_SCATTER_ROWS needs full-width tiles (tile_w == window width)");
    const uint64_t n_out = row_major ? (uint64_t)ww * wh : n_slots;
    const int L = s->dev.n_lights;

    // Batches: whole tiles (row-major output: whole tile ROWS, so a batch is a contiguous band of the image and can
    // be copied out on its own).  Queues are sized for the worst-case fan-out: a binary tree only if some material
    // both reflects and refracts.
    const int fan = s->dev.max_children >= 2 ? 2 : 1;
    uint64_t worst = 1;
    for (int k = 0; k < o.fuel && fan == 2; ++k) worst *= 2;
    const uint64_t unit = row_major ? (uint64_t)f.tiles_x * per_tile : per_tile;  // batch granularity in pixel slots
    // default: about RTC_BATCHES batches per call so the pipelines have something to overlap, within [256 Ki, 2 Mi]
    // samples (measured: more, smaller batches only add launches; fewer leave the deep levels' tails exposed).
    // 4 when the frame stays in HBM; 14 (tapered at the end) when each finished batch is copied out, so that the copies start earlier
    // (r02 sweep with graph replay, profiles/r02m_e2e_batch_sweep.txt: 14 x 4 pipelines 6.11 ms, 10 x 4 6.28 ms for the teapot frame).
    const uint64_t auto_samples = std::min<uint64_t>(std::max<uint64_t>(n_slots * (uint64_t)f.spp / (uint64_t)knob(tune.batches, host_out ? 14 : 4, 1, 1024), 256u << 10), 2u << 20);
    const uint64_t target_samples = (uint64_t)knob(tune.batch_samples, (int)auto_samples, 1024, 1 << 28);
    const uint64_t max_records = (uint64_t)24 << 20;  // records per queue
    uint64_t batch_samples = std::min(target_samples, std::max<uint64_t>(max_records / worst, 1));
    uint64_t batch_pixels = std::max<uint64_t>(batch_samples / f.spp / unit, 1) * unit;
    batch_pixels = std::min(batch_pixels, std::max<uint64_t>((n_slots + unit - 1) / unit, 1) * unit);
    batch_samples = batch_pixels * f.spp;
    const uint64_t q_cap = batch_samples * worst;
    // hit list: one segment per level, sized for that level's worst case (fan^level rays per sample)
    rtcdev::HitList hits{};
    uint64_t hit_cap = 0, level_cap = batch_samples;
    hits.n_levels = o.fuel + 1;
    for (int k = 0; k <= o.fuel; ++k) {
        hits.seg_off[k] = (uint32_t)hit_cap;
        hits.seg_cap[k] = (uint32_t)level_cap;
        hit_cap += level_cap;
        if (fan == 2) level_cap *= 2;
    }
    if (hit_cap * (uint64_t)std::max(L, 1) >= 0xf0000000ull) return fail("rtc_render: batch too large");
    hits.total_cap = (uint32_t)hit_cap;
    hits.fan = fan;
    // Batch schedule: (first pixel slot, slots).  Device-resident frames use equal batches.  When each finished batch is
    // copied to the host, the copy of the LAST batch cannot overlap any kernel, so the tail of the schedule is tapered
    // (..., B, B/2, B/4, B/4): the exposed copy shrinks from 1/6 to 1/24 of the frame (teapot 4K: 0.6 -> 0.15 ms).
    std::vector<std::pair<uint64_t, uint64_t>> schedule;
    {
        const bool taper = host_out != nullptr && !tune.no_taper;
        uint64_t at = 0;
        while (at < n_slots) {
            const uint64_t left = n_slots - at;
            uint64_t take = std::min(batch_pixels, left);
            if (taper && left <= 2 * batch_pixels && left > 4 * unit) {
                // halve what is left, in whole units, but never below 1/4 of a full batch
                const uint64_t floor_px = std::max<uint64_t>(batch_pixels / 4 / unit, 1) * unit;
                uint64_t half = std::max<uint64_t>(left / 2 / unit, 1) * unit;
                if (half < floor_px) half = left;
                take = std::min(half, left);
            }
            schedule.emplace_back(at, take);
            at += take;
        }
        if (schedule.empty()) schedule.emplace_back(0, 0);
    }
    const uint64_t n_batches = schedule.size();
    const bool profile = (o.flags & RTC_FLAG_PROFILE) != 0;
    const int n_pipes = profile ? 1 : (int)std::min<uint64_t>((uint64_t)knob(tune.pipes, 4, 1, MAX_PIPES), n_batches);
    for (int p = 0; p < n_pipes; ++p)
        if (int rc = ensure_pipe(ctx.pipe[p], q_cap, hit_cap, hit_cap * (uint64_t)fan * 3, hit_cap * (uint64_t)std::max(L, 1), batch_samples)) return rc;
    if (n_out > ctx.out_pixels) {
        cudaFree(ctx.out); ctx.out = nullptr; ctx.out_pixels = 0;
        CU(cudaMalloc((void**)&ctx.out, std::max<uint64_t>(n_out, 1) * 3 * sizeof(double)));
        ctx.out_pixels = n_out;
    }
    if (n_batches > ctx.counter_batches) {
        cudaFree(ctx.counters); ctx.counters = nullptr;
        cudaFreeHost(ctx.counters_host); ctx.counters_host = nullptr;
        ctx.counter_batches = 0;
        const uint64_t nb = std::max<uint64_t>(n_batches, 64);
        CU(cudaMalloc((void**)&ctx.counters, nb * COUNTERS_PER_BATCH * sizeof(unsigned long long)));
        CU(cudaHostAlloc((void**)&ctx.counters_host, nb * COUNTERS_PER_BATCH * sizeof(unsigned long long) + 64, cudaHostAllocDefault));
        ctx.counter_batches = nb;
    }

    const bool copy_on_own_stream = host_out != nullptr && !tune.no_copy_stream;
    const bool staggered = host_out != nullptr && !tune.no_stream_priorities;
    for (Pipe& p : ctx.pipe) p.stream = staggered ? p.stream_staggered : p.stream_flat;
    cudaStream_t st0 = ctx.pipe[0].stream;
    const size_t counter_bytes = n_batches * COUNTERS_PER_BATCH * sizeof(unsigned long long);
    const int grid_full = ctx.sm_count * knob(tune.blocks_per_sm, 16, 1, 32);
    const int feat = s->feat | (profile ? rtcdev::FEAT_TALLY : 0);  // the work tallies are counted under RTC_FLAG_PROFILE only
    uint64_t n_live = 0;
    for (const auto& b : schedule) n_live += b.second != 0;
    const uint64_t launches = n_live * (2 * (uint64_t)(o.fuel + 1) + (L > 0 ? 1 : 0) + 1);
    while (host_out && copy_on_own_stream && ctx.batch_events.size() < n_batches) {  // (created here: nothing is allocated while a graph is being captured)
        cudaEvent_t e;
        CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        ctx.batch_events.push_back(e);
    }
    const bool timeline = std::getenv("RTC_TIMELINE") != nullptr;  // debugging aid: when each batch's kernels / copy finished
    std::vector<cudaEvent_t> tl_events;
    size_t n_prof = 0;  // profile mode: event k is recorded after launch k-1 (classes cycle trace, shadow, shade)
    auto prof_mark = [&](cudaStream_t st) -> int {
        if (!profile) return 0;
        if (n_prof == ctx.prof_events.size()) {
            cudaEvent_t e;
            CU(cudaEventCreate(&e));
            ctx.prof_events.push_back(e);
        }
        CU(cudaEventRecord(ctx.prof_events[n_prof++], st));
        return 0;
    };
    std::vector<int> prof_class;  // class of the interval that ENDS at event k (0 trace, 1 shadow, 2 shade, -1 other)
    int* h_overflow = (int*)(ctx.counters_host + n_batches * COUNTERS_PER_BATCH);
    // Everything one frame puts on the device's streams, from clearing the counters to reading them back: issued directly,
    // or once under stream capture and replayed as a graph afterwards.
    struct CopyOp { double* dst; size_t dpitch; const double* src; size_t spitch; size_t width, height; int batch; };  // height 1: contiguous
    auto issue_copy = [&](const CopyOp& c, cudaStream_t cs) -> int {
        if (c.height == 1) CU(cudaMemcpyAsync(c.dst, c.src, c.width, cudaMemcpyDeviceToHost, cs));
        else CU(cudaMemcpy2DAsync(c.dst, c.dpitch, c.src, c.spitch, c.width, c.height, cudaMemcpyDeviceToHost, cs));
        return 0;
    };
    // Host-paced copies (rtc_tuning.host_paced_copies): the copy of a batch is not made to wait for the batch's event ON
    // THE DEVICE; this thread polls the events and issues each copy once its batch is done.
    const bool host_paced = host_out != nullptr && copy_on_own_stream && tune.host_paced_copies && !profile && !timeline;
    std::vector<CopyOp> paced;
    auto enqueue = [&]() -> int {
    CU(cudaMemsetAsync(ctx.counters, 0, counter_bytes, st0));
    CU(cudaMemsetAsync(ctx.overflow, 0, sizeof(int), st0));
    CU(cudaEventRecord(ctx.fork, st0));
    for (int p = 1; p < n_pipes; ++p) CU(cudaStreamWaitEvent(ctx.pipe[p].stream, ctx.fork, 0));
    for (uint64_t b = 0; b < n_batches; ++b) {
        Pipe& p = ctx.pipe[b % (uint64_t)n_pipes];
        cudaStream_t st = p.stream;
        const uint64_t first = schedule[b].first;
        const uint64_t npix = schedule[b].second;
        if (!npix) break;
        unsigned long long* cnt = ctx.counters + b * COUNTERS_PER_BATCH;
        CU(cudaMemsetAsync(p.accum, 0, npix * f.spp * 3 * sizeof(double), st));  // (padding slots of partial tiles are never written)
        hits.child_rgb = p.child_rgb;
        hits.a = p.hit[0]; hits.b = p.hit[1]; hits.c = p.hit[2]; hits.d = p.hit[3]; hits.e = p.hit[4]; hits.f = p.hit[5];
        hits.counts = cnt + CNT_HITS;
        hits.counts_back = cnt + CNT_HITS_BACK;
        auto level_params = [&](int level) {
            rtcdev::LevelParams lv{};
            const int in = level & 1, outq = in ^ 1;
            lv.in_q0 = p.q[in][0]; lv.in_q1 = p.q[in][1]; lv.in_q2 = p.q[in][2]; lv.in_q3 = p.q[in][3];
            lv.out_q0 = p.q[outq][0]; lv.out_q1 = p.q[outq][1]; lv.out_q2 = p.q[outq][2]; lv.out_q3 = p.q[outq][3];
            lv.in_count = cnt + CNT_QUEUE + level;
            lv.in_count_back = cnt + CNT_QUEUE_BACK + level;
            lv.out_count = cnt + CNT_QUEUE + level + 1;
            lv.out_count_back = cnt + CNT_QUEUE_BACK + level + 1;
            lv.in_capacity = p.q_capacity;
            lv.out_capacity = p.q_capacity;
            lv.first_sample = first * f.spp;
            lv.n_primary = npix * f.spp;
            lv.level = level;
            lv.hit_count = cnt + CNT_HITS + level;
            lv.hit_count_back = cnt + CNT_HITS_BACK + level;
            lv.shadow = p.shadow;
            lv.accum = p.accum;
            lv.stats = cnt + CNT_STATS;
            lv.overflow = ctx.overflow;
            return lv;
        };
        const int grid0 = (int)std::min<uint64_t>((uint64_t)grid_full, (npix * f.spp + RTC_BLOCK - 1) / RTC_BLOCK);
        if (profile) { if (int rc = prof_mark(st)) return rc; prof_class.push_back(-1); }
        for (int level = 0; level <= o.fuel; ++level)  // the chain: level k+1 depends on level k only
            rtcdev::launch_trace(feat, level == 0, level == 0 ? grid0 : grid_full, st, s->dev, f, level_params(level), hits);
        if (profile) { if (int rc = prof_mark(st)) return rc; prof_class.push_back(0); }
        rtcdev::launch_shadow(feat, grid_full, st, s->dev, f, level_params(0), hits);  // every (hit, light) of the batch
        if (profile) { if (int rc = prof_mark(st)) return rc; prof_class.push_back(1); }
        for (int level = o.fuel; level >= 0; --level)  // deepest first: a hit is lit when its children's colours are final
            rtcdev::launch_light(level == 0 ? grid0 : grid_full, st, s->dev, level_params(level), hits);
        if (profile) { if (int rc = prof_mark(st)) return rc; prof_class.push_back(2); }
        rtcdev::launch_resolve(st, f, p.accum, first, npix, ctx.out, row_major ? 1 : 0);
        if (timeline) { cudaEvent_t e; CU(cudaEventCreate(&e)); CU(cudaEventRecord(e, st)); tl_events.push_back(e); }
        if (host_out) {  // this batch's pixels are final: copy them out while the other pipelines keep tracing
            cudaStream_t cst = st;
            if (copy_on_own_stream) {
                CU(cudaEventRecord(ctx.batch_events[b], st));
                cst = ctx.copy_stream[b % (uint64_t)n_pipes];
                if (!host_paced) CU(cudaStreamWaitEvent(cst, ctx.batch_events[b], 0));
            }
            // the batch's copies: one contiguous piece, or (bands of a sharded frame) one strided piece + a partial last band
            CopyOp ops[2];
            int n_ops = 0;
            if (row_major) {
                const uint64_t row0 = first / unit * (uint64_t)o.tile_h;
                const uint64_t row1 = std::min<uint64_t>((first + npix) / unit * (uint64_t)o.tile_h, (uint64_t)wh);
                const uint64_t off_px = row0 * (uint64_t)ww, len_px = (row1 - row0) * (uint64_t)ww;
                if (len_px) ops[n_ops++] = CopyOp{host_out + 3 * off_px, 0, ctx.out + 3 * off_px, 0, len_px * 3 * sizeof(double), 1, (int)b};
            } else if (scatter_rows) {
                // tiles are full-width rows of tile_h pixel rows: owned tile row k of this shard is global tile row
                // shard_index + k * shard_count, i.e. one strided 2-D copy puts the batch where it belongs in the frame
                const uint64_t k0 = first / per_tile, k1 = (first + npix) / per_tile;  // owned tile rows [k0, k1)
                const uint64_t row_bytes = (uint64_t)ww * 3 * sizeof(double);
                auto global_y = [&](uint64_t k) { return ((uint64_t)o.shard_index + k * (uint64_t)o.shard_count) * (uint64_t)o.tile_h; };
                // only the last tile row of the frame can hold fewer than tile_h pixel rows
                const bool partial_last = k1 > k0 && global_y(k1 - 1) + (uint64_t)o.tile_h > (uint64_t)wh;
                const uint64_t n_full = (k1 - k0) - (partial_last ? 1 : 0);
                if (n_full)
                    ops[n_ops++] = CopyOp{host_out + 3 * global_y(k0) * (uint64_t)ww, (uint64_t)o.shard_count * (uint64_t)o.tile_h * row_bytes,
                                          ctx.out + 3 * k0 * per_tile, per_tile * 3 * sizeof(double), (uint64_t)o.tile_h * row_bytes, n_full, (int)b};
                if (partial_last) {
                    const uint64_t y0 = global_y(k1 - 1);
                    ops[n_ops++] = CopyOp{host_out + 3 * y0 * (uint64_t)ww, 0, ctx.out + 3 * (k1 - 1) * per_tile, 0, ((uint64_t)wh - y0) * row_bytes, 1, (int)b};
                }
            } else if (npix) {
                ops[n_ops++] = CopyOp{host_out + 3 * first, 0, ctx.out + 3 * first, 0, npix * 3 * sizeof(double), 1, (int)b};
            }
            for (int k = 0; k < n_ops; ++k) {
                if (host_paced) paced.push_back(ops[k]);
                else if (int rc = issue_copy(ops[k], cst)) return rc;
            }
            if (timeline && !host_paced) { cudaEvent_t e; CU(cudaEventCreate(&e)); CU(cudaEventRecord(e, cst)); tl_events.push_back(e); }
        }
    }
    CU(cudaGetLastError());
    if (host_out && copy_on_own_stream && !host_paced) {
        for (int p = 0; p < n_pipes; ++p) {
            CU(cudaEventRecord(ctx.copy_done, ctx.copy_stream[p]));
            CU(cudaStreamWaitEvent(st0, ctx.copy_done, 0));
        }
    }
    for (int p = 1; p < n_pipes; ++p) {
        CU(cudaEventRecord(ctx.pipe[p].done, ctx.pipe[p].stream));
        CU(cudaStreamWaitEvent(st0, ctx.pipe[p].done, 0));
    }
    if (host_paced) return 0;  // (the read-back follows the paced copies, below)
    CU(cudaMemcpyAsync(ctx.counters_host, ctx.counters, counter_bytes, cudaMemcpyDeviceToHost, st0));
    CU(cudaMemcpyAsync(h_overflow, ctx.overflow, sizeof(int), cudaMemcpyDeviceToHost, st0));
    return 0;
    };  // enqueue

    // ---- replay a captured graph when the frame repeats ----
    bool use_graph = !profile && !timeline && !tune.no_graphs && !host_paced;
    if (use_graph && host_out) {  // (copies into pageable memory cannot be captured)
        cudaPointerAttributes attr{};
        if (cudaPointerGetAttributes(&attr, host_out) != cudaSuccess || attr.type != cudaMemoryTypeHost) { use_graph = false; cudaGetLastError(); }
    }
    if (use_graph) {
        std::vector<unsigned char> key;
        auto put = [&key](const void* p, size_t n) { key.insert(key.end(), (const unsigned char*)p, (const unsigned char*)p + n); };
        put(&s->dev, sizeof s->dev); put(&f, sizeof f); put(&o, sizeof o); put(&host_out, sizeof host_out); put(&feat, sizeof feat);
        put(&n_pipes, sizeof n_pipes); put(&staggered, sizeof staggered); put(&copy_on_own_stream, sizeof copy_on_own_stream); put(&grid_full, sizeof grid_full);
        put(&ctx.out, sizeof ctx.out); put(&ctx.counters, sizeof ctx.counters); put(&ctx.counters_host, sizeof ctx.counters_host);
        put(schedule.data(), schedule.size() * sizeof schedule[0]);
        for (int p = 0; p < n_pipes; ++p) {
            const Pipe& pp = ctx.pipe[p];
            put(pp.q, sizeof pp.q); put(pp.hit, sizeof pp.hit); put(&pp.child_rgb, sizeof pp.child_rgb); put(&pp.shadow, sizeof pp.shadow); put(&pp.accum, sizeof pp.accum);
            put(&pp.q_capacity, sizeof pp.q_capacity);
        }
        DeviceCtx::CachedGraph* slot = nullptr;
        for (auto& g : ctx.graphs)
            if (g.exec && g.key == key) slot = &g;
        if (!slot) {
            slot = &ctx.graphs[0];
            for (auto& g : ctx.graphs)
                if (!g.exec || g.last_used < slot->last_used) { slot = &g; if (!g.exec) break; }
            if (slot->exec) { cudaGraphExecDestroy(slot->exec); slot->exec = nullptr; }
            slot->key.clear();
            cudaGraph_t graph = nullptr;
            CU(cudaStreamBeginCapture(st0, cudaStreamCaptureModeThreadLocal));
            const int rc = enqueue();
            const cudaError_t ce = cudaStreamEndCapture(st0, &graph);
            if (rc != 0 || ce != cudaSuccess || !graph) {
                if (graph) cudaGraphDestroy(graph);
                cudaGetLastError();
                return rc != 0 ? rc : fail(std::string("rtc_render: cudaStreamEndCapture: ") + cudaGetErrorString(ce));
            }
            const cudaError_t ie = cudaGraphInstantiate(&slot->exec, graph, 0);
            cudaGraphDestroy(graph);
            if (ie != cudaSuccess) { slot->exec = nullptr; return fail(std::string("rtc_render: cudaGraphInstantiate: ") + cudaGetErrorString(ie)); }
            slot->key = std::move(key);
        }
        slot->last_used = ++ctx.graph_clock;
        CU(cudaEventRecord(ctx.ev[0], st0));
        CU(cudaGraphLaunch(slot->exec, st0));
        CU(cudaEventRecord(ctx.ev[1], st0));
    } else {
        CU(cudaEventRecord(ctx.ev[0], st0));
        if (int rc = enqueue()) return rc;
        if (host_paced) {
            std::vector<char> issued(paced.size(), 0);
            size_t left = paced.size();
            while (left) {
                for (size_t k = 0; k < paced.size(); ++k) {
                    if (issued[k]) continue;
                    const cudaError_t q = cudaEventQuery(ctx.batch_events[paced[k].batch]);
                    if (q == cudaErrorNotReady) continue;
                    if (q != cudaSuccess) return fail(std::string("rtc_render: cudaEventQuery: ") + cudaGetErrorString(q));
                    if (int rc = issue_copy(paced[k], ctx.copy_stream[(size_t)paced[k].batch % (size_t)n_pipes])) return rc;
                    issued[k] = 1;
                    --left;
                }
            }
            for (int p = 0; p < n_pipes; ++p) {
                CU(cudaEventRecord(ctx.copy_done, ctx.copy_stream[p]));
                CU(cudaStreamWaitEvent(st0, ctx.copy_done, 0));
            }
            CU(cudaMemcpyAsync(ctx.counters_host, ctx.counters, counter_bytes, cudaMemcpyDeviceToHost, st0));
            CU(cudaMemcpyAsync(h_overflow, ctx.overflow, sizeof(int), cudaMemcpyDeviceToHost, st0));
        }
        CU(cudaEventRecord(ctx.ev[1], st0));
    }
    CU(cudaStreamSynchronize(st0));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, ctx.ev[0], ctx.ev[1]));
    const int overflow = *h_overflow;
    if (timeline) {
        std::fprintf(stderr, "rtc_render timeline (ms after start; per batch: kernels done%s):", host_out ? ", copy done" : "");
        for (cudaEvent_t e : tl_events) {
            float t = 0.f;
            cudaEventElapsedTime(&t, ctx.ev[0], e);
            std::fprintf(stderr, " %.2f", t);
            cudaEventDestroy(e);
        }
        std::fprintf(stderr, " | all %.2f\n", ms);
    }
    if (stats) {
        std::memset(stats, 0, sizeof *stats);
        // R_ref: the reference re-traces every secondary sub-tree once per light, so a ray at
        // bounce level k exists L^k times there, and each of its hits issues L shadow queries.
        double mult = 1.0;
        uint64_t primary = 0, shadow = 0, refl = 0, refr = 0, hits_all = 0, tally[4] = {0, 0, 0, 0}, longest = 0, skipped = 0;
        double ref = 0.0;
        for (uint64_t b = 0; b < n_batches; ++b) {
            const unsigned long long* c = ctx.counters_host + b * COUNTERS_PER_BATCH;
            auto level_rays = [&](int level) { return std::min<uint64_t>(c[CNT_QUEUE + level] + c[CNT_QUEUE_BACK + level], q_cap); };
            primary += c[CNT_STATS];
            if (s->feat & rtcdev::FEAT_TRANSPARENT) { refl += c[CNT_STATS + 1]; refr += c[CNT_STATS + 2]; }
            else for (int level = 1; level <= o.fuel; ++level) refl += level_rays(level);  // every child is a reflection
            for (int k = 0; k < 4; ++k) tally[k] += c[CNT_STATS + 3 + k];
            longest = std::max<uint64_t>(longest, c[CNT_STATS + 7]);
            skipped += c[CNT_STATS + 8];
            mult = 1.0;
            for (int level = 0; level <= o.fuel; ++level) {
                const uint64_t n_in = level == 0 ? c[CNT_STATS] : level_rays(level);
                const uint64_t hits = c[CNT_HITS + level] + c[CNT_HITS_BACK + level];
                shadow += hits * (uint64_t)L;
                hits_all += hits;
                ref += mult * ((double)n_in + (double)hits * L);
                mult *= (double)L;
            }
        }
        stats->rays_primary = primary;
        stats->rays_shadow = shadow;
        stats->rays_reflected = refl;
        stats->rays_refracted = refr;
        stats->rays_ref = (uint64_t)ref;
        stats->kernel_launches = launches;
        stats->pixels = n_out;
        stats->ms_kernels = ms;
        stats->ms_trace_shade = ms;
        stats->overflow = overflow;
        stats->surface_hits = hits_all;
        stats->bvh_node_visits = tally[0];
        stats->triangle_tests = tally[1];
        stats->primitive_tests = tally[2];
        stats->ray_transforms = tally[3];
        stats->max_node_visits = longest;
        stats->shadow_queries_culled = skipped;
        for (size_t k = 1; k < n_prof; ++k) {
            float t = 0.f;
            if (prof_class[k] < 0 || cudaEventElapsedTime(&t, ctx.prof_events[k - 1], ctx.prof_events[k]) != cudaSuccess) continue;
            (prof_class[k] == 0 ? stats->ms_trace : prof_class[k] == 1 ? stats->ms_shadow : stats->ms_shade) += t;
        }
        stats->ms_total = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count();
    }
    if (overflow) return fail(overflow == 2 ? "rtc_render: ray queue overflow" : "rtc_render: on-chip crossing list overflow (scene too deep for the device evaluator)");
    if (out_dev_rgb) *out_dev_rgb = ctx.out;
    if (out_pixels) *out_pixels = n_out;
    return 0;
}

int rtc_render_device(rtc_scene* s, const rtc_camera* cam, const rtc_opts* o, double** out_dev_rgb, uint64_t* out_pixels, rtc_stats* stats) {
    return render_impl(s, cam, o, nullptr, out_dev_rgb, out_pixels, stats);
}

int rtc_render(rtc_scene* s, const rtc_camera* cam, const rtc_opts* o, double* out_rgb, rtc_stats* stats) {
    if (!out_rgb) return fail("rtc_render: null output buffer");
    return render_impl(s, cam, o, out_rgb, nullptr, nullptr, stats);
}

int rtc_render_multi(rtc_scene* const* scenes, int n, const rtc_camera* cam, const rtc_opts* o_in, double* out_rgb, rtc_stats* stats) {
    if (!scenes || n <= 0 || !cam || !out_rgb) return fail("rtc_render_multi: null argument");
    for (int i = 0; i < n; ++i) {
        if (!scenes[i]) return fail("rtc_render_multi: null scene");
        for (int j = 0; j < i; ++j)
            if (scenes[j]->device == scenes[i]->device) return fail("rtc_render_multi: two scenes live on the same device");
    }
    if (n == 1) return rtc_render(scenes[0], cam, o_in, out_rgb, stats);
    rtc_opts base{};
    if (o_in) base = *o_in;
    else base.fuel = 5;
    if (base.shard_count > 1) return fail("rtc_render_multi: the call shards the frame itself; pass shard_count <= 1");
    const bool whole = base.win_w <= 0 || base.win_h <= 0;
    base.tile_w = whole ? cam->hsize : base.win_w;  // full-width tiles: a shard owns every n-th band of tile_h pixel rows
    if (base.tile_h <= 0) base.tile_h = 4;
    base.shard_count = n;
    base.flags |= RTC_FLAG_SCATTER_ROWS;
    const auto t_begin = std::chrono::steady_clock::now();
    std::vector<rtc_stats> st((size_t)n);
    std::vector<std::string> errs((size_t)n);
    std::vector<int> rcs((size_t)n, 0);
    std::vector<std::thread> threads;
    for (int i = 0; i < n; ++i)
        threads.emplace_back([&, i] {
            rtc_opts o = base;
            o.shard_index = i;
            rcs[(size_t)i] = render_impl(scenes[i], cam, &o, out_rgb, nullptr, nullptr, &st[(size_t)i]);
            if (rcs[(size_t)i]) errs[(size_t)i] = g_error;  // g_error is thread-local
        });
    for (auto& t : threads) t.join();
    for (int i = 0; i < n; ++i)
        if (rcs[(size_t)i]) return fail("rtc_render_multi: device " + std::to_string(scenes[i]->device) + ": " + errs[(size_t)i]);
    if (stats) {
        std::memset(stats, 0, sizeof *stats);
        for (const rtc_stats& a : st) {
            stats->rays_primary += a.rays_primary; stats->rays_shadow += a.rays_shadow;
            stats->rays_reflected += a.rays_reflected; stats->rays_refracted += a.rays_refracted;
            stats->rays_ref += a.rays_ref; stats->kernel_launches += a.kernel_launches;
            stats->surface_hits += a.surface_hits; stats->bvh_node_visits += a.bvh_node_visits;
            stats->triangle_tests += a.triangle_tests; stats->primitive_tests += a.primitive_tests;
            stats->ray_transforms += a.ray_transforms;
            stats->max_node_visits = std::max(stats->max_node_visits, a.max_node_visits);
            stats->ms_kernels = std::max(stats->ms_kernels, a.ms_kernels);  // devices run concurrently: the slowest one counts
            stats->ms_trace_shade = std::max(stats->ms_trace_shade, a.ms_trace_shade);
        }
        const bool whole_w = o_in == nullptr || o_in->win_w <= 0 || o_in->win_h <= 0;
        stats->pixels = whole_w ? (uint64_t)cam->hsize * cam->vsize : (uint64_t)o_in->win_w * o_in->win_h;
        stats->ms_total = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count();
    }
    return 0;
}

}  // extern "C"
