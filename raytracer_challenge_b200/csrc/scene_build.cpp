#include "scene_build.hpp"

#include <emmintrin.h>

#include <algorithm>
#include <atomic>
#include <thread>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstddef>
#include <cstring>
#include <functional>
#include <limits>
#include <map>

namespace rtchost {
namespace {

using rtcdev::BvhNode;
constexpr double INF = std::numeric_limits<double>::infinity();

struct Box {
    double lo[3] = {INF, INF, INF}, hi[3] = {-INF, -INF, -INF};
    bool bounded = true;  // false: must be tested for every ray
    bool exact = true;    // false: holds a finite cone, whose single-root crossing may lie outside (see rtcdev::ConeGuard)
    bool closed = true;   // everything below is a closed solid: a line crosses its surface an even number of times (sphere, cube,
                          // capped finite cylinder), so "inside" as filter_by_group counts it ends where the solid ends
    void add(const double p[3]) {
        for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], p[k]); hi[k] = std::max(hi[k], p[k]); }
    }
    void add(const Box& b) {
        if (!b.bounded) bounded = false;
        if (!b.exact) exact = false;
        if (!b.closed) closed = false;
        for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], b.lo[k]); hi[k] = std::max(hi[k], b.hi[k]); }
    }
    bool empty() const { return lo[0] > hi[0]; }
};

// The reference's cofactor inverse (src/linalg/matrix.rs:162-207) returns a last row that is (0,0,0,1) only up to
// rounding (element [3][3] is det3/det computed along two different paths), so "affine" is tested with a tolerance.
bool affine(const double m[16]) {
    return std::fabs(m[12]) <= 1e-9 && std::fabs(m[13]) <= 1e-9 && std::fabs(m[14]) <= 1e-9 && std::fabs(m[15] - 1.0) <= 1e-9;
}
bool all_finite(const double* v, int n) {
    for (int i = 0; i < n; ++i)
        if (!std::isfinite(v[i])) return false;
    return true;
}

// General 3x4 affine inverse (object -> world from world -> object); only used for bounds, which are padded.
bool invert_affine(const double m[16], double out[12]) {
    const double a = m[0], b = m[1], c = m[2], d = m[4], e = m[5], f = m[6], g = m[8], h = m[9], i = m[10];
    const double det = a * (e * i - f * h) - b * (d * i - f * g) + c * (d * h - e * g);
    if (det == 0.0 || !std::isfinite(det)) return false;
    const double r[9] = {(e * i - f * h) / det, (c * h - b * i) / det, (b * f - c * e) / det,
                         (f * g - d * i) / det, (a * i - c * g) / det, (c * d - a * f) / det,
                         (d * h - e * g) / det, (b * g - a * h) / det, (a * e - b * d) / det};
    for (int row = 0; row < 3; ++row) {
        for (int col = 0; col < 3; ++col) out[row * 4 + col] = r[row * 3 + col];
        out[row * 4 + 3] = -(r[row * 3] * m[3] + r[row * 3 + 1] * m[7] + r[row * 3 + 2] * m[11]);
    }
    return true;
}
void xf_point(const double m[12], const double p[3], double out[3]) {
    for (int r = 0; r < 3; ++r) out[r] = m[r * 4] * p[0] + m[r * 4 + 1] * p[1] + m[r * 4 + 2] * p[2] + m[r * 4 + 3];
}

// World-space bounds of one primitive, or unbounded.  Planes (src/shape.rs:954-957) and cylinders / cones with an
// infinite range are unbounded.  A cone with a finite range gets the box of its walls and caps, marked inexact: its
// single-root branch (src/shape.rs:811-818) has no y-range check, so for the rare rays that take it the crossing is
// not confined to [min, max] — those rays are detected per query (rtcdev::ConeGuard) and test every item.
// fwd: object -> world (invert_affine of p.transform_inv), or null if that matrix is singular
Box prim_bounds(const rtc_prim& p, const double* fwd) {
    Box b;
    b.closed = p.kind == RTC_SPHERE || p.kind == RTC_CUBE || (p.kind == RTC_CYLINDER && p.closed && std::isfinite(p.min) && std::isfinite(p.max));
    if (!fwd) { b.bounded = false; return b; }
    double olo[3], ohi[3];
    switch (p.kind) {
        case RTC_SPHERE:
        case RTC_CUBE:
            for (int k = 0; k < 3; ++k) { olo[k] = -1.0; ohi[k] = 1.0; }
            break;
        case RTC_CYLINDER:
            if (!std::isfinite(p.min) || !std::isfinite(p.max)) { b.bounded = false; return b; }
            olo[0] = -1.0; ohi[0] = 1.0; olo[2] = -1.0; ohi[2] = 1.0;
            olo[1] = std::min(p.min, p.max); ohi[1] = std::max(p.min, p.max);
            break;
        case RTC_CONE: {  // walls: x^2 + z^2 = y^2 with min < y < max; caps: radius |min|, |max| (src/shape.rs:767-822)
            if (!std::isfinite(p.min) || !std::isfinite(p.max)) { b.bounded = false; return b; }
            const double lim = std::max(std::fabs(p.min), std::fabs(p.max));
            olo[0] = -lim; ohi[0] = lim; olo[2] = -lim; ohi[2] = lim;
            olo[1] = std::min(p.min, p.max); ohi[1] = std::max(p.min, p.max);
            b.exact = false;
            break;
        }
        case RTC_TRIANGLE:
        case RTC_SMOOTH_TRIANGLE: {
            double v[3][3];
            for (int k = 0; k < 3; ++k) { v[0][k] = p.p1[k]; v[1][k] = p.p1[k] + p.e1[k]; v[2][k] = p.p1[k] + p.e2[k]; }
            for (auto& q : v) { double w[3]; xf_point(fwd, q, w); b.add(w); }
            goto pad;
        }
        default:  // plane
            b.bounded = false;
            return b;
    }
    for (int c = 0; c < 8; ++c) {
        const double q[3] = {(c & 4) ? ohi[0] : olo[0], (c & 2) ? ohi[1] : olo[1], (c & 1) ? ohi[2] : olo[2]};
        double w[3];
        xf_point(fwd, q, w);
        b.add(w);
    }
pad:
    for (int k = 0; k < 3; ++k) {
        if (!std::isfinite(b.lo[k]) || !std::isfinite(b.hi[k])) { b.bounded = false; return b; }
        const double e = 1e-9 * std::max({1.0, std::fabs(b.lo[k]), std::fabs(b.hi[k])});
        b.lo[k] -= e;
        b.hi[k] += e;
    }
    return b;
}

// Nearest float not above / not below x.  The neighbouring float is reached by stepping the bit pattern (what
// nextafterf computes, without the libm call: this runs six times per BVH item).
inline float step_float(float f, bool up) {
    if (!(std::fabs(f) <= std::numeric_limits<float>::max())) {  // NaN stays; an infinity steps back to the largest finite value
        if (f > 0.0f && !up) return std::numeric_limits<float>::max();
        if (f < 0.0f && up) return -std::numeric_limits<float>::max();
        return f;
    }
    if (f == 0.0f) return up ? std::numeric_limits<float>::denorm_min() : -std::numeric_limits<float>::denorm_min();
    uint32_t u;
    std::memcpy(&u, &f, 4);
    u += ((f > 0.0f) == up) ? 1u : (uint32_t)-1;
    std::memcpy(&f, &u, 4);
    return f;
}
float f32_down(double x) {
    float f = (float)x;
    if ((double)f > x) f = step_float(f, false);
    return f;
}
float f32_up(double x) {
    float f = (float)x;
    if ((double)f < x) f = step_float(f, true);
    return f;
}

// n records of 8 doubles (lo.xyz, 0, hi.xyz, 0) -> 8 floats each, the first four rounded towards -inf and the last four
// towards +inf by the conversion instruction itself (MXCSR rounding control).  The compiler does not model MXCSR as an input
// of cvtpd2ps, so the conversions sit between two compiler barriers and only touch memory.
__attribute__((noinline)) void convert_directed(const double* in, float* out, size_t n) {
    const unsigned saved = _mm_getcsr();
    for (int pass = 0; pass < 2; ++pass) {
        _mm_setcsr((saved & ~0x6000u) | (pass == 0 ? 0x2000u : 0x4000u));  // round down, then round up
        __asm__ volatile("" ::: "memory");
        for (size_t i = 0; i < n; ++i) {
            const double* q = in + i * 8 + pass * 4;
            const __m128 a = _mm_cvtpd_ps(_mm_loadu_pd(q)), b = _mm_cvtpd_ps(_mm_loadu_pd(q + 2));
            _mm_storeu_ps(out + i * 8 + pass * 4, _mm_movelh_ps(a, b));
        }
        __asm__ volatile("" ::: "memory");
    }
    _mm_setcsr(saved);
}

// Checked once per process against the scalar f32_down / f32_up on values that round, values that are exact, both signs,
// the float range's ends and the denormal range: should a compiler ever move the conversions out of their rounding-mode
// bracket, the staging falls back to the scalar form instead of handing the device inward-rounded boxes.
bool directed_conversion_is_exact() {
    static const bool ok = [] {
        const double v[16] = {0.1, -0.1, 1.0, -1.0, 1.0 + 1e-12, -(1.0 + 1e-12), 3.4028234663852886e38, 3.5e38,
                              -3.5e38, 1e-46, -1e-46, 0.0, 123456.789, -98765.4321, 1e-39, -1e-39};
        double in[32];
        float got[32];
        for (int r = 0; r < 4; ++r)
            for (int k = 0; k < 4; ++k) { in[r * 8 + k] = v[r * 4 + k]; in[r * 8 + 4 + k] = v[r * 4 + k]; }
        convert_directed(in, got, 4);
        for (int r = 0; r < 4; ++r)
            for (int k = 0; k < 4; ++k) {
                const float lo = f32_down(v[r * 4 + k]), hi = f32_up(v[r * 4 + k]);
                if (std::memcmp(&got[r * 8 + k], &lo, 4) != 0 || std::memcmp(&got[r * 8 + 4 + k], &hi, 4) != 0) return false;
            }
        return true;
    }();
    return ok;
}

// ---- binned-SAH BVH2 over item boxes ------------------------------------------------------
// The build runs on the boxes the device will test: per item the fp32 box, padded and rounded outwards.  Rounding
// outwards is monotone, so the union of the rounded item boxes IS the rounded union — a node's stored box is the
// plain min / max of its items' boxes, and the whole build is four-wide SSE min / max (baseline x86-64, no -march).
struct Builder {
    struct alignas(16) Item {  // kept in leaf order (partitioned in place: sequential passes)
        __m128 lo, hi;  // lanes x, y, z; lane w unused
        __m128 c;       // centroid; lane w carries the item's index (bit pattern)
    };
    struct Bounds {
        __m128 lo, hi;
        void reset() { lo = _mm_set1_ps(std::numeric_limits<float>::infinity()); hi = _mm_set1_ps(-std::numeric_limits<float>::infinity()); }
        void add(__m128 l, __m128 h) { lo = _mm_min_ps(lo, l); hi = _mm_max_ps(hi, h); }
        void add(const Bounds& b) { add(b.lo, b.hi); }
        float area() const {
            alignas(16) float d[4];
            _mm_store_ps(d, _mm_sub_ps(hi, lo));
            return 2.0f * (d[0] * d[1] + d[1] * d[2] + d[2] * d[0]);
        }
    };
    std::vector<Item> items;
    std::vector<BvhNode> nodes;  // build(): slot mid - 1 of the node that splits at `mid` (sparse); compact(): final
    float leaf_weight = 3.0f;    // cost of one exact f64 primitive test relative to one fp32 node visit
    int small_node = 8;
    int max_bins = 16;
    int fork_levels = 4;         // the top levels of the tree fork a thread per left subtree (RTC_BUILD_FORK_LEVELS, 0 = sequential)
    std::atomic<int> depth{0};

    Builder(const std::vector<Box>& boxes, double pad, const rtc_tuning& tune) {
        if (tune.sah_leaf_weight_x100 > 0) leaf_weight = std::max(0.1f, (float)tune.sah_leaf_weight_x100 / 100.0f);
        {   // thread budget: all cores unless the caller says otherwise (one process per GPU: the ranks of a node build at the same time)
            int cores = (int)std::thread::hardware_concurrency();
            if (tune.build_threads > 0) cores = tune.build_threads;
            fork_levels = 0;
            while (fork_levels < 4 && (2 << fork_levels) <= cores) fork_levels++;
        }
        if (tune.sah_bins > 0) max_bins = std::min(16, std::max(2, (int)tune.sah_bins));
        if (tune.sah_small_node > 0) small_node = std::min(8, std::max(2, (int)tune.sah_small_node));
        items.resize(boxes.size());
        for (size_t i = 0; i < boxes.size(); ++i) {
            alignas(16) float lo[4] = {0.f, 0.f, 0.f, 0.f}, hi[4] = {0.f, 0.f, 0.f, 0.f}, c[4];
            for (int k = 0; k < 3; ++k) {
                lo[k] = f32_down(boxes[i].lo[k] - pad);
                hi[k] = f32_up(boxes[i].hi[k] + pad);
                c[k] = 0.5f * lo[k] + 0.5f * hi[k];
            }
            const int32_t index = (int32_t)i;
            std::memcpy(&c[3], &index, 4);
            items[i].lo = _mm_load_ps(lo);
            items[i].hi = _mm_load_ps(hi);
            items[i].c = _mm_load_ps(c);
        }
        nodes.resize(boxes.size());
    }
    int32_t index_of(size_t slot) const {
        alignas(16) float c[4];
        _mm_store_ps(c, items[slot].c);
        int32_t index;
        std::memcpy(&index, &c[3], 4);
        return index;
    }
    Bounds range_bounds(int begin, int end) const {
        Bounds b;
        b.reset();
        for (int i = begin; i < end; ++i) b.add(items[(size_t)i].lo, items[(size_t)i].hi);
        return b;
    }
    static int32_t leaf_link(int begin, int count) { return ~((begin << 2) | (count - 1)); }

    int32_t finish(int begin, int mid, int end, int level, const Bounds& lb, const Bounds& rb) {
        const int32_t self = mid - 1;
        int32_t l, r;
        if (level < fork_levels && end - begin >= 1024) {
            std::thread left([&] { l = build(begin, mid, level + 1, lb); });
            r = build(mid, end, level + 1, rb);
            left.join();
        } else {
            l = build(begin, mid, level + 1, lb);
            r = build(mid, end, level + 1, rb);
        }
        BvhNode& nd = nodes[(size_t)self];
        std::memset(&nd, 0, sizeof nd);
        alignas(16) float v[4];
        _mm_store_ps(v, lb.lo); for (int k = 0; k < 3; ++k) nd.lo0[k] = v[k];
        _mm_store_ps(v, lb.hi); for (int k = 0; k < 3; ++k) nd.hi0[k] = v[k];
        _mm_store_ps(v, rb.lo); for (int k = 0; k < 3; ++k) nd.lo1[k] = v[k];
        _mm_store_ps(v, rb.hi); for (int k = 0; k < 3; ++k) nd.hi1[k] = v[k];
        nd.c0 = l;
        nd.c1 = r;
        return self;
    }

    // 2..8 items: every split of the items sorted by centroid, per axis (no bins: the fixed cost of 48 bins is what
    // the thousands of small nodes at the bottom of the tree would pay).
    int32_t build_small(int begin, int end, int level, const Bounds& whole) {
        const int n = end - begin;
        float best_cost = std::numeric_limits<float>::infinity();
        int best_split = -1;
        int best_order[8];
        Bounds best_left, best_right;
        alignas(16) float cen[8][4];
        for (int i = 0; i < n; ++i) _mm_store_ps(cen[i], items[(size_t)(begin + i)].c);
        for (int a = 0; a < 3; ++a) {
            int order[8];
            for (int i = 0; i < n; ++i) {  // stable insertion sort by centroid
                int j = i;
                while (j > 0 && cen[order[j - 1]][a] > cen[i][a]) { order[j] = order[j - 1]; --j; }
                order[j] = i;
            }
            if (!(cen[order[0]][a] < cen[order[n - 1]][a])) continue;  // no extent along this axis
            Bounds suffix[8];
            Bounds acc;
            acc.reset();
            for (int i = n - 1; i > 0; --i) {
                const Item& it = items[(size_t)(begin + order[i])];
                acc.add(it.lo, it.hi);
                suffix[i] = acc;
            }
            acc.reset();
            for (int sp = 1; sp < n; ++sp) {
                const Item& it = items[(size_t)(begin + order[sp - 1])];
                acc.add(it.lo, it.hi);
                const float cost = acc.area() * (float)sp + suffix[sp].area() * (float)(n - sp);
                if (cost < best_cost) {
                    best_cost = cost; best_split = sp; best_left = acc; best_right = suffix[sp];
                    for (int i = 0; i < n; ++i) best_order[i] = order[i];
                }
            }
        }
        int mid;
        Bounds lb, rb;
        if (best_split < 0) {
            if (n <= 4) return leaf_link(begin, n);
            mid = begin + n / 2;  // identical centroids: split by index
            lb = range_bounds(begin, mid);
            rb = range_bounds(mid, end);
        } else {
            // leaf cost: every item test is an f64 primitive intersection (expensive) -> weight 3 vs 1 per node
            const float leaf_cost = leaf_weight * (float)n;
            const float split_cost = 1.0f + leaf_weight * best_cost / std::max(whole.area(), 1e-30f);
            if (n <= 4 && leaf_cost <= split_cost) return leaf_link(begin, n);
            Item tmp[8];
            for (int i = 0; i < n; ++i) tmp[i] = items[(size_t)(begin + best_order[i])];
            for (int i = 0; i < n; ++i) items[(size_t)(begin + i)] = tmp[i];
            mid = begin + best_split;
            lb = best_left;
            rb = best_right;
        }
        return finish(begin, mid, end, level, lb, rb);
    }

    // Binned-SAH split of items [begin, end) whose union is `whole`; returns the link for the range.
    // Subtrees over disjoint item ranges are independent, so the top levels fork threads.  A node's slot is derived
    // from its split position (unique per inner node), which makes the result independent of thread timing;
    // compact() then renumbers the nodes in depth-first order, the layout a sequential build would produce.
    int32_t build(int begin, int end, int level, const Bounds& whole) {
        for (int seen = depth.load(); level > seen && !depth.compare_exchange_weak(seen, level);) {}
        const int n = end - begin;
        if (n <= 1) return leaf_link(begin, n);
        if (n <= small_node) return build_small(begin, end, level, whole);
        const int NB = std::min(max_bins, n);
        Bounds cb;  // centroid bounds: the split candidates
        cb.reset();
        for (int i = begin; i < end; ++i) cb.add(items[(size_t)i].c, items[(size_t)i].c);
        alignas(16) float ext[4], scale[4];
        _mm_store_ps(ext, _mm_sub_ps(cb.hi, cb.lo));
        for (int a = 0; a < 3; ++a) scale[a] = ext[a] > 0.0f ? (float)NB / ext[a] : 0.0f;
        scale[3] = 0.0f;
        const __m128 vscale = _mm_load_ps(scale), vtop = _mm_set1_ps((float)(NB - 1)), vzero = _mm_setzero_ps();
        auto bins_of = [&](const Item& it, int k[4]) {
            const __m128 f = _mm_min_ps(_mm_max_ps(_mm_mul_ps(_mm_sub_ps(it.c, cb.lo), vscale), vzero), vtop);
            _mm_store_si128(reinterpret_cast<__m128i*>(k), _mm_cvttps_epi32(f));
        };
        Bounds bin_box[3][16];
        int bin_n[3][16];
        for (int a = 0; a < 3; ++a)
            for (int k = 0; k < NB; ++k) { bin_box[a][k].reset(); bin_n[a][k] = 0; }
        for (int i = begin; i < end; ++i) {  // one pass fills the bins of all three axes
            const Item& it = items[(size_t)i];
            alignas(16) int k[4];
            bins_of(it, k);
            for (int a = 0; a < 3; ++a) {
                bin_box[a][k[a]].add(it.lo, it.hi);
                bin_n[a][k[a]]++;
            }
        }
        float best_cost = std::numeric_limits<float>::infinity();
        int best_axis = -1, best_bin = -1;
        Bounds best_left, best_right;
        for (int a = 0; a < 3; ++a) {
            if (!(ext[a] > 0.0f)) continue;
            Bounds right_box[16];
            int right_n[16];
            Bounds acc;
            acc.reset();
            int cnt = 0;
            for (int k = NB - 1; k > 0; --k) {
                acc.add(bin_box[a][k]);
                cnt += bin_n[a][k];
                right_box[k] = acc;
                right_n[k] = cnt;
            }
            Bounds left;
            left.reset();
            int ln = 0;
            for (int k = 0; k < NB - 1; ++k) {
                left.add(bin_box[a][k]);
                ln += bin_n[a][k];
                if (ln == 0 || right_n[k + 1] == 0) continue;
                const float cost = left.area() * (float)ln + right_box[k + 1].area() * (float)right_n[k + 1];
                if (cost < best_cost) { best_cost = cost; best_axis = a; best_bin = k; best_left = left; best_right = right_box[k + 1]; }
            }
        }
        int mid;
        Bounds lb, rb;
        if (best_axis < 0) {
            mid = begin + n / 2;  // identical centroids: split by index
            lb = range_bounds(begin, mid);
            rb = range_bounds(mid, end);
        } else {
            auto it = std::partition(items.begin() + begin, items.begin() + end, [&](const Item& x) {
                alignas(16) int k[4];
                bins_of(x, k);
                return k[best_axis] <= best_bin;
            });
            mid = (int)(it - items.begin());
            lb = best_left;
            rb = best_right;
        }
        return finish(begin, mid, end, level, lb, rb);
    }
    // depth-first renumbering of the sparse slots into a dense array; returns the new root link
    // Renumbering of the sparse slots into a dense array: the top of the tree breadth-first (the first MAX_TOP_NODES
    // indices are what the kernels stage in shared memory: every walk starts there), the subtrees below depth-first (a
    // parent is followed by its left subtree).  Returns the new root link.
    int32_t compact(int32_t root) {
        if (root < 0) { nodes.clear(); return root; }
        std::vector<BvhNode> dense;
        dense.reserve(nodes.size());
        std::vector<std::pair<int32_t, int32_t>> work;  // (sparse slot, dense index): a queue while the top is numbered, then a stack
        size_t head = 0;
        dense.emplace_back();
        work.push_back({root, 0});
        auto emit = [&](int32_t slot, int32_t at, bool breadth_first) {
            BvhNode nd = nodes[(size_t)slot];
            // children are numbered when their parent is written
            int32_t kids[2] = {nd.c0, nd.c1};
            int32_t dense_kids[2];
            for (int k = 0; k < 2; ++k) {
                if (kids[k] >= 0) { dense_kids[k] = (int32_t)dense.size(); dense.emplace_back(); }
                else dense_kids[k] = kids[k];
            }
            if (breadth_first) {
                if (kids[0] >= 0) work.push_back({kids[0], dense_kids[0]});
                if (kids[1] >= 0) work.push_back({kids[1], dense_kids[1]});
            } else {
                if (kids[1] >= 0) work.push_back({kids[1], dense_kids[1]});
                if (kids[0] >= 0) work.push_back({kids[0], dense_kids[0]});
            }
            nd.c0 = dense_kids[0];
            nd.c1 = dense_kids[1];
            dense[(size_t)at] = nd;
        };
        while (head < work.size() && dense.size() < (size_t)rtcdev::MAX_TOP_NODES) {
            const auto [slot, at] = work[head++];
            emit(slot, at, true);
        }
        // what is still queued are the roots of the subtrees below the top: depth-first from here, leftmost first
        std::vector<std::pair<int32_t, int32_t>> rest(work.begin() + (long)head, work.end());
        work.assign(rest.rbegin(), rest.rend());
        while (!work.empty()) {
            const auto [slot, at] = work.back();
            work.pop_back();
            emit(slot, at, false);
        }
        nodes.swap(dense);
        return 0;
    }
};

template <class T>
int32_t dedupe(std::map<std::string, int32_t>& index, std::vector<T>& store, const T& value) {
    std::string key((const char*)&value, sizeof(T));
    auto it = index.find(key);
    if (it != index.end()) return it->second;
    store.push_back(value);
    const int32_t id = (int32_t)store.size() - 1;
    index.emplace(std::move(key), id);
    return id;
}

}  // namespace

std::string build_host_scene(const rtc_scene_desc& d, HostScene& out, const rtc_tuning* tuning) {
    out = HostScene{};
    rtc_tuning tune{};
    if (tuning) tune = *tuning;
    if (d.abi_version != RTC_ABI_VERSION) return "abi_version mismatch";
    if (d.n_prims < 0 || d.n_nodes < 1 || d.n_children < 0 || d.n_materials < 0 || d.n_patterns < 0 || d.n_lights < 0)
        return "negative array length (or no root node)";
    if (d.n_lights > rtcdev::MAX_LIGHTS) return "too many lights (max " + std::to_string(rtcdev::MAX_LIGHTS) + ")";
    if ((d.n_prims && !d.prims) || !d.nodes || (d.n_children && !d.children) || (d.n_materials && !d.materials) ||
        (d.n_patterns && !d.patterns) || (d.n_lights && !d.lights))
        return "null array pointer";

    // ---- patterns ----
    for (int i = 0; i < d.n_patterns; ++i) {
        const rtc_pattern& p = d.patterns[i];
        if (p.kind < RTC_PAT_DEBUG || p.kind > RTC_PAT_STRIPES) return "pattern " + std::to_string(i) + ": unknown kind";
        const bool jitter = p.kind == RTC_PAT_JITTER_COLOR || p.kind == RTC_PAT_JITTER_POINT;
        const bool mixture = p.kind >= RTC_PAT_BLEND;
        // children must precede their parent (post-order), which also rules out cycles
        if ((jitter || mixture) && (p.left < 0 || p.left >= i)) return "pattern " + std::to_string(i) + ": bad left/inner index";
        if (mixture && (p.right < 0 || p.right >= i)) return "pattern " + std::to_string(i) + ": bad right index";
        if (mixture && !affine(p.transform_inv)) return "pattern " + std::to_string(i) + ": transform is not affine";
        if (jitter && p.noise_kind != RTC_NOISE_SIMPLEX && p.noise_kind != RTC_NOISE_FRACTAL) return "pattern: unknown noise kind";
        if (jitter && p.octaves < 0) return "pattern: negative octaves";
        rtcdev::Pattern q{};
        q.kind = p.kind; q.left = p.left; q.right = p.right; q.noise_kind = p.noise_kind; q.octaves = p.octaves;
        q.noise_scale = p.noise_scale;
        for (int k = 0; k < 3; ++k) q.color[k] = p.color[k];
        for (int k = 0; k < 16; ++k) q.m[k] = p.transform_inv[k];
        out.patterns.push_back(q);
    }
    {  // evaluation-frame depth: Blend / Gradient / RingGradient / ColorJitter keep a frame while a child is evaluated
        std::vector<int> frames((size_t)d.n_patterns, 0);
        for (int i = 0; i < d.n_patterns; ++i) {
            const rtc_pattern& p = d.patterns[i];
            int child = 0;
            if (p.kind == RTC_PAT_JITTER_COLOR || p.kind == RTC_PAT_JITTER_POINT) child = frames[(size_t)p.left];
            else if (p.kind >= RTC_PAT_BLEND) child = std::max(frames[(size_t)p.left], frames[(size_t)p.right]);
            const bool needs = p.kind == RTC_PAT_BLEND || p.kind == RTC_PAT_GRADIENT || p.kind == RTC_PAT_RING_GRADIENT || p.kind == RTC_PAT_JITTER_COLOR;
            frames[(size_t)i] = child + (needs ? 1 : 0);
            if (frames[(size_t)i] > rtcdev::MAX_PATTERN_FRAMES) return "pattern nesting too deep for the device evaluator";
        }
    }
    // ---- materials ----
    for (int i = 0; i < d.n_materials; ++i) {
        const rtc_material& m = d.materials[i];
        if (m.pattern < 0 || m.pattern >= d.n_patterns) return "material " + std::to_string(i) + ": bad pattern index";
        rtcdev::Material q{};
        q.ambient = m.ambient; q.diffuse = m.diffuse; q.specular = m.specular; q.shininess = m.shininess;
        q.reflective = m.reflective; q.transparency = m.transparency; q.refractive_index = m.refractive_index;
        q.pattern = m.pattern;
        if (d.patterns[m.pattern].kind == RTC_PAT_PLAIN) {
            q.flags = 1;
            for (int k = 0; k < 3; ++k) q.plain[k] = d.patterns[m.pattern].color[k];
        }
        out.materials.push_back(q);
        out.max_children = std::max(out.max_children, (int)(m.reflective != 0.0) + (int)(m.transparency != 0.0));
    }
    for (int i = 0; i < d.n_lights; ++i) {
        rtcdev::Light l{};
        for (int k = 0; k < 3; ++k) { l.intensity[k] = d.lights[i].intensity[k]; l.origin[k] = d.lights[i].origin[k]; }
        out.lights.push_back(l);
    }

    const bool trace = std::getenv("RTC_TRACE") != nullptr;
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
    const auto t0 = now();
    // ---- primitives ----
    std::map<std::string, int32_t> xf_index, mi_index;
    std::vector<Box> prim_box((size_t)d.n_prims);
    out.prims.reserve((size_t)d.n_prims);
    {
        size_t n_tri = 0;
        for (int i = 0; i < d.n_prims; ++i) n_tri += d.prims[i].kind == RTC_TRIANGLE || d.prims[i].kind == RTC_SMOOTH_TRIANGLE;
        out.tri_pe.reserve(n_tri * 10);
        out.tri_n.reserve(n_tri * 10);
    }
    // The triangles of a mesh share their three matrices (Element::composite bakes the group transform into every
    // child, src/shape.rs:47-72): a primitive whose matrices are byte-equal to its predecessor's reuses that one's
    // validation, transform ids and forward transform — one 384-byte compare instead of 48 finiteness tests, two
    // record compares and a 3x3 inversion per triangle.
    int32_t last_xf = -1, last_mi = -1;
    const rtc_prim* last = nullptr;
    double fwd[12];
    bool fwd_ok = false;
    for (int i = 0; i < d.n_prims; ++i) {
        const rtc_prim& p = d.prims[i];
        if (p.kind < RTC_SPHERE || p.kind > RTC_SMOOTH_TRIANGLE) return "prim " + std::to_string(i) + ": unknown kind";
        if (p.material < 0 || p.material >= d.n_materials) return "prim " + std::to_string(i) + ": bad material index";
        static_assert(offsetof(rtc_prim, material_inv) + sizeof(p.material_inv) - offsetof(rtc_prim, transform_inv) == 3 * 16 * sizeof(double), "the three matrices are contiguous");
        if (!last || std::memcmp(p.transform_inv, last->transform_inv, 3 * 16 * sizeof(double)) != 0) {
            if (!affine(p.transform_inv) || !affine(p.material_inv)) return "prim " + std::to_string(i) + ": non-affine matrix";
            if (!all_finite(p.transform_inv, 16) || !all_finite(p.transform_inv_tsp, 16) || !all_finite(p.material_inv, 16))
                return "prim " + std::to_string(i) + ": non-finite matrix";
            rtcdev::Xform xf{};
            for (int k = 0; k < 12; ++k) xf.inv[k] = p.transform_inv[k];
            for (int r = 0; r < 3; ++r)
                for (int c = 0; c < 3; ++c) xf.tsp[r * 3 + c] = p.transform_inv_tsp[r * 4 + c];
            rtcdev::MatInv mi{};
            for (int k = 0; k < 16; ++k) mi.m[k] = p.material_inv[k];
            if (last_xf < 0 || std::memcmp(&out.xforms[(size_t)last_xf], &xf, sizeof xf) != 0) last_xf = dedupe(xf_index, out.xforms, xf);
            if (last_mi < 0 || std::memcmp(&out.matinvs[(size_t)last_mi], &mi, sizeof mi) != 0) last_mi = dedupe(mi_index, out.matinvs, mi);
            fwd_ok = invert_affine(p.transform_inv, fwd);
            last = &p;
        }
        rtcdev::Prim q{};
        q.kind_flags = p.kind | (p.casts_shadow ? 0x100 : 0) | (p.closed ? 0x200 : 0);
        q.material = p.material;
        q.xform = last_xf;
        q.matinv = last_mi;
        q.geom = -1;
        if (p.kind == RTC_TRIANGLE || p.kind == RTC_SMOOTH_TRIANGLE) {
            q.geom = (int32_t)(out.tri_pe.size() / 10);
            const double pe[10] = {p.p1[0], p.p1[1], p.p1[2], p.e1[0], p.e1[1], p.e1[2], p.e2[0], p.e2[1], p.e2[2], 0.0};
            const double nn[10] = {p.n1[0], p.n1[1], p.n1[2], p.n2[0], p.n2[1], p.n2[2], p.n3[0], p.n3[1], p.n3[2], 0.0};
            out.tri_pe.insert(out.tri_pe.end(), pe, pe + 10);
            out.tri_n.insert(out.tri_n.end(), nn, nn + 10);
        } else if (p.kind == RTC_CYLINDER || p.kind == RTC_CONE) {
            q.geom = (int32_t)(out.ranges.size() / 2);
            out.ranges.push_back(p.min);
            out.ranges.push_back(p.max);
        }
        if (!p.casts_shadow) out.all_cast_shadow = 0;
        out.prims.push_back(q);
        prim_box[(size_t)i] = prim_bounds(p, fwd_ok ? fwd : nullptr);
        if (p.kind == RTC_CONE && prim_box[(size_t)i].bounded) {  // one guard per distinct linear part
            rtcdev::ConeGuard g{};
            for (int r = 0; r < 3; ++r)
                for (int c = 0; c < 3; ++c) g.lin[r * 3 + c] = p.transform_inv[r * 4 + c];
            bool seen = false;
            for (const auto& o : out.guards) seen = seen || std::memcmp(&o, &g, sizeof g) == 0;
            if (!seen) out.guards.push_back(g);
        }
    }

    const auto t1 = now();
    // ---- group tree: gates, CSG roots, items --------------------------------------------------
    std::vector<int> visited_node((size_t)d.n_nodes, 0), visited_prim((size_t)d.n_prims, 0);
    std::vector<Box> item_boxes;   // per BVH item
    std::vector<int32_t> item_ids; // per BVH item: prim index or ~csg index
    item_boxes.reserve((size_t)d.n_prims);
    item_ids.reserve((size_t)d.n_prims);
    int next_prim = 0;             // primitives must appear in depth-first order
    std::string err;

    // Returns the bounds of the subtree as the reference computes them (the union of everything below); fills prim range
    // [pb, pe) and `eff`, the box that holds every crossing the subtree can REPORT: for a Difference that is the left
    // operand's box, for an Intersection the overlap of the two (a crossing of the right operand survives the filter of
    // src/shape.rs:230-269 only inside the left solid, and vice versa) — for CLOSED operands only, see below.  A top-level CSG item enters the BVH with `eff`:
    // a ray that misses it but hits the reference's box evaluates, in the reference, to the empty list.
    std::function<Box(int, uint32_t, int, int, int&, int&, Box&)> walk;
    // csg_root: index into out.csg_roots of the enclosing top-level CSG item, or -1.
    walk = [&](int ni, uint32_t mask, int csg_root, int csg_depth, int& pb, int& pe, Box& eff) -> Box {
        Box sub;
        eff = Box();
        Box child_eff[2];
        (void)mask;
        pb = next_prim;
        pe = next_prim;
        if (!err.empty()) return sub;
        if (ni < 0 || ni >= d.n_nodes || visited_node[(size_t)ni]) { err = "group tree: bad or repeated node index"; return sub; }
        visited_node[(size_t)ni] = 1;
        const rtc_node& nd = d.nodes[ni];
        if (nd.kind < RTC_UNION || nd.kind > RTC_AGGREGATION) { err = "group: unknown kind"; return sub; }
        if (nd.first_child < 0 || nd.n_children < 0 || nd.first_child + nd.n_children > d.n_children) { err = "group: child range out of bounds"; return sub; }
        const bool is_csg = nd.kind != RTC_AGGREGATION;
        if (is_csg && nd.n_children != 2) { err = "CSG group needs exactly two children (src/shape.rs:80-83)"; return sub; }
        if (ni == 0 && is_csg) { err = "the root node must be an aggregation"; return sub; }

        // Gate of this group (the virtual root has none: src/world.rs:21-23).  Whether the box must be evaluated per ray
        // is only known once the subtree's bounds are: the slot is assigned AFTER the walk (below), so boxes that turn
        // out to be conservative, and boxes that can never fail, cost no slot.
        bool gate_candidate = false;
        if (ni != 0) {
            int constraining = 0;
            for (int k = 0; k < 3; ++k) {
                const double lo = nd.bbox_min[k], hi = nd.bbox_max[k];
                if (lo != lo || hi != hi) { err = "group box holds NaN"; return sub; }
                const bool free_axis = (lo == -INF && hi == INF) || (lo == INF && hi == -INF);
                if (!free_axis) constraining++;
            }
            if (constraining <= 1) out.n_gates_always_pass++;  // the slab test cannot fail ("group boxes")
            else gate_candidate = true;
        }

        int my_root = csg_root;
        int my_depth = csg_depth;
        if (is_csg) {
            if (csg_root < 0) {
                my_root = (int)out.csg_roots.size();
                rtcdev::CsgRoot r{};
                r.first_op = (int32_t)out.csg_ops.size();
                out.csg_roots.push_back(r);
            }
            my_depth = csg_depth + 1;
            if (my_depth > rtcdev::MAX_CSG_MARKS) { err = "CSG nesting too deep"; return sub; }
            out.csg_ops.push_back({rtcdev::OP_BEGIN, 0, 0, 0});
        }
        int left_b = 0, left_e = 0;
        size_t skip_at = 0;
        const bool skippable = is_csg && (nd.kind == RTC_DIFFERENCE || nd.kind == RTC_INTERSECTION);
        for (int c = 0; c < nd.n_children; ++c) {
            const int32_t ch = d.children[nd.first_child + c];
            int cb = next_prim, ce = next_prim;
            if (ch >= 0) {
                if (ch != next_prim || ch >= d.n_prims || visited_prim[(size_t)ch]) { err = "primitives must be listed in depth-first order, each exactly once"; return sub; }
                visited_prim[(size_t)ch] = 1;
                next_prim++;
                ce = next_prim;
                sub.add(prim_box[(size_t)ch]);
                eff.add(prim_box[(size_t)ch]);
                if (c < 2) child_eff[c] = prim_box[(size_t)ch];
                if (my_root >= 0) {
                    out.csg_ops.push_back({rtcdev::OP_PRIM, ch, 0, 0});
                } else if (prim_box[(size_t)ch].bounded) {
                    item_boxes.push_back(prim_box[(size_t)ch]);
                    item_ids.push_back(ch);
                    out.prims[(size_t)ch].kind_flags |= rtcdev::PRIM_IN_BVH;
                } else {
                    out.unbounded.push_back(ch);
                }
            } else {
                Box ceff;
                Box b = walk(~ch, 0u, my_root, my_depth, cb, ce, ceff);
                if (!err.empty()) return sub;
                sub.add(b);
                eff.add(ceff);
                if (c < 2) child_eff[c] = ceff;
            }
            if (c == 0) {
                left_b = cb; left_e = ce;
                if (skippable) { skip_at = out.csg_ops.size(); out.csg_ops.push_back({rtcdev::OP_SKIP_IF_EMPTY, 0, 0, 0}); }
            }
        }
        pe = next_prim;
        if (is_csg) {
            out.csg_ops.push_back({rtcdev::OP_END, nd.kind, left_b, left_e});
            if (skippable) {
                out.csg_ops[skip_at].a = (int32_t)(out.csg_ops.size() - 1 - skip_at);  // from the skip op to this group's OP_END
                // Tightening needs CLOSED operands.  filter_by_group keeps a right crossing while "inside the left operand", and
                // inside is a parity count: a left operand a line can cross ONCE (a plane, a triangle, an open cylinder — entered
                // through the wall, left through the open end) stays "entered" to the end of the line, so right crossings far
                // outside its box are reported (fuzz:20, 38, 57 ... differed from brute force before this rule).  With a closed
                // left operand everything a Difference reports lies in the left box; an Intersection reports inside every
                // closed operand's box.  (Boxes of finite cones are not exact — see ConeGuard — so a subtree that holds one keeps
                // the union too.)
                const bool c0 = child_eff[0].closed && child_eff[0].bounded, c1 = child_eff[1].closed && child_eff[1].bounded;
                if (child_eff[0].exact && child_eff[1].exact && (c0 || (nd.kind == RTC_INTERSECTION && c1))) {
                    Box t = c0 ? child_eff[0] : child_eff[1];
                    if (nd.kind == RTC_INTERSECTION && c0 && c1) {
                        for (int k = 0; k < 3; ++k) { t.lo[k] = std::max(t.lo[k], child_eff[1].lo[k]); t.hi[k] = std::min(t.hi[k], child_eff[1].hi[k]); }
                        bool overlap = true;
                        for (int k = 0; k < 3; ++k) overlap = overlap && t.lo[k] <= t.hi[k];
                        if (!overlap) t = child_eff[0];  // disjoint operands: nothing can be reported; any valid box will do
                    }
                    t.closed = child_eff[0].closed && child_eff[1].closed;
                    eff = t;
                }
            }
            if (csg_root < 0) {  // this node is a top-level CSG item
                rtcdev::CsgRoot& r = out.csg_roots[(size_t)my_root];
                r.n_ops = (int32_t)out.csg_ops.size() - r.first_op;
                r.prim_begin = pb;
                r.prim_end = pe;
                if (eff.bounded && !eff.empty()) {
                    item_boxes.push_back(eff);
                    item_ids.push_back(~my_root);
                    for (int q = pb; q < pe; ++q) out.prims[(size_t)q].kind_flags |= rtcdev::PRIM_IN_BVH;
                } else if (pe > pb) {
                    out.unbounded.push_back(~my_root);
                }
            }
        }
        // a finite box that encloses everything below it can only reject rays that miss all of it: no gate needed
        if (gate_candidate) {
            bool conservative = sub.bounded && sub.exact;  // (a finite cone's single-root crossing may lie outside its box)
            for (int k = 0; k < 3 && conservative; ++k) {
                const double lo = nd.bbox_min[k], hi = nd.bbox_max[k];
                if (!std::isfinite(lo) || !std::isfinite(hi)) { conservative = false; break; }
                if (sub.empty()) continue;
                const double tol = 4e-9 * std::max({1.0, std::fabs(sub.lo[k]), std::fabs(sub.hi[k])});
                if (lo > sub.lo[k] + tol || hi < sub.hi[k] - tol) conservative = false;
            }
            if (conservative || pe == pb) {
                out.n_gates_conservative++;
            } else {
                const int slot = (int)out.gates.size();
                if (slot >= rtcdev::MAX_GATES) { err = "more than " + std::to_string(rtcdev::MAX_GATES) + " group boxes need exact evaluation"; return sub; }
                rtcdev::Gate g{};
                for (int k = 0; k < 3; ++k) { g.lo[k] = nd.bbox_min[k]; g.hi[k] = nd.bbox_max[k]; }
                out.gates.push_back(g);
                for (int q = pb; q < pe; ++q) {
                    if (slot < 32) out.prims[(size_t)q].gate_mask |= 1u << slot;
                    else out.prims[(size_t)q].gate_mask_hi |= 1u << (slot - 32);
                }
                out.n_gates_exact++;
            }
        }
        return sub;
    };
    int pb = 0, pe = 0;
    Box root_eff;
    walk(0, 0u, -1, 0, pb, pe, root_eff);
    if (!err.empty()) return err;
    if (next_prim != d.n_prims) return "some primitives are not reachable from the root node";

    const auto t2 = now();
    auto t3 = t2;
    // ---- BVH over the bounded items -------------------------------------------------------------
    out.n_bvh_items = (int32_t)item_ids.size();
    for (int32_t id : item_ids)
        if (id < 0 || (out.prims[(size_t)id].kind_flags & 0xff) < RTC_TRIANGLE) { out.all_items_triangles = 0; break; }
    if (!item_ids.empty()) {
        Box root;
        for (auto& b : item_boxes) root.add(b);
        double S = 0.0;
        for (int k = 0; k < 3; ++k) S = std::max({S, std::fabs(root.lo[k]), std::fabs(root.hi[k])});
        // fp32 traversal error bound is ~9 * 2^-24 * S (DESIGN.md "conservative fp32 culling"); pad with margin
        const double pad = 4e-6 * std::max(S, 1e-30);
        for (int k = 0; k < 3; ++k) { out.root_lo[k] = root.lo[k] - pad; out.root_hi[k] = root.hi[k] + pad; }
        const int n_items = (int)item_ids.size();
        const bool fast = ((d.build_flags & RTC_BUILD_FAST) != 0 || tune.bvh_builder == 1) && tune.bvh_builder != 2;
        if (fast && n_items >= rtcdev::LBVH_MIN_ITEMS && n_items <= rtcdev::LBVH_MAX_ITEMS) {
            // The device builds the tree after the upload (bvh_build.cu): hand it the same padded fp32 item boxes the
            // host build would use, and the leaf records in item order.
            out.device_build = 1;
            out.item_box32.resize((size_t)n_items * 8);
            // f32_down / f32_up of every padded bound, two at a time: the double -> float conversion itself rounds towards
            // -inf / +inf when MXCSR says so (identical values: test_staged_item_boxes_round_outwards)
            std::vector<double> padded((size_t)n_items * 8);
            for (int i = 0; i < n_items; ++i) {
                double* q = &padded[(size_t)i * 8];
                for (int k = 0; k < 3; ++k) { q[k] = item_boxes[(size_t)i].lo[k] - pad; q[4 + k] = item_boxes[(size_t)i].hi[k] + pad; }
                q[3] = 0.0;
                q[7] = 0.0;
            }
            if (directed_conversion_is_exact()) {
                convert_directed(padded.data(), out.item_box32.data(), (size_t)n_items);
            } else {
                for (size_t i = 0; i < (size_t)n_items * 8; ++i) out.item_box32[i] = (i & 4) ? f32_up(padded[i]) : f32_down(padded[i]);
            }
            out.item_ids = std::move(item_ids);
            out.bvh_root = 0;
            out.bvh_depth = 0;  // known after the device build
            if (trace) std::fprintf(stderr, "build_host_scene: prims %.3f ms, group tree %.3f ms, device-build staging %.3f ms%s\n", ms(t0, t1), ms(t1, t2), ms(t2, now()),
                                    directed_conversion_is_exact() ? "" : " (scalar rounding: the directed conversion failed its self-check)");
            return "";
        }
        Builder bld(item_boxes, pad, tune);
        out.bvh_root = bld.compact(bld.build(0, (int)item_ids.size(), 0, bld.range_bounds(0, (int)item_ids.size())));
        out.bvh = std::move(bld.nodes);
        out.bvh_depth = bld.depth;
        t3 = now();
        if (out.bvh_depth + 2 > rtcdev::BVH_STACK) return "BVH deeper than the traversal stack";
        out.leaf_items.resize(item_ids.size());
        out.leaf_tri.assign(item_ids.size() * 10, 0.0);
        for (size_t i = 0; i < item_ids.size(); ++i) {
            const int32_t id = item_ids[(size_t)bld.index_of(i)];
            rtcdev::LeafItem li{id, 0, 0, -1};
            if (id >= 0) {
                const rtcdev::Prim& p = out.prims[(size_t)id];
                li.kind_flags = p.kind_flags;
                li.xform = p.xform;
                li.geom = p.geom;
                if ((p.kind_flags & 0xff) >= RTC_TRIANGLE)
                    for (int k = 0; k < 10; ++k) out.leaf_tri[i * 10 + (size_t)k] = out.tri_pe[(size_t)p.geom * 10 + (size_t)k];
            }
            out.leaf_items[i] = li;
        }
        for (int k = 0; k < 3; ++k) { out.root_lo[k] = root.lo[k] - pad; out.root_hi[k] = root.hi[k] + pad; }
    }
    if (out.bvh.empty()) out.bvh.emplace_back();  // keep device pointers non-null
    if (trace)
        std::fprintf(stderr, "build_host_scene: prims %.3f ms, group tree %.3f ms, bvh %.3f ms, leaf items %.3f ms\n", ms(t0, t1), ms(t1, t2), ms(t2, t3), ms(t3, now()));
    return "";
}

}  // namespace rtchost
