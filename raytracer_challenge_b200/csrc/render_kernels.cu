// Wavefront kernels of the render path.  The recursion of World::color_at -> shade_hit -> reflected_color /
// refracted_color (src/world.rs:50-149) becomes, per batch of samples:
//
//   trace_spawn  x (fuel + 1), a chain down the bounce levels: one thread per ray of the level (level 0:
//                  Camera::ray_at_pixel, src/camera.rs:39-55, generated in-kernel; level k > 0: the queue level k-1
//                  wrote) -> closest hit (World::intersect + sort + hit, src/world.rs:18-24,
//                  src/intersection.rs:123-132), n1/n2 for transparent hits (src/intersection.rs:70-103), the surface
//                  frame of prepare_state (src/intersection.rs:50-68) -> hit record appended to the level's segment of
//                  the hit list, and the reflected / refracted children (src/world.rs:84-132) pushed to the next
//                  level's queue.  A ray that misses stores black where its parent will look for its colour.
//   trace_shadow x 1: one thread per (hit, light) over the hits of ALL levels, light-major: World::is_shadowed
//                  (src/world.rs:26-48) -> one byte per (hit, light)
//   light        x (fuel + 1), a chain UP the levels (deepest first): World::shade_hit for every hit of the level —
//                  pattern colour, Shape::lighting per light (src/shape.rs:429-463), plus the colours its children
//                  stored, combined in the reference's own order (src/world.rs:58-79) — stored where the parent
//                  will look for it (level 0: the sample).  No atomics, no weights: every colour has one writer, so
//                  frames are bit-reproducible, and the sum has the reference's association.
//
// Each kernel holds one BVH traversal (or none), so its instruction footprint stays small (profiles/README.md, r01a);
// neither the shadow nor the lighting kernels gather from the ray queue or recompute the surface frame: everything
// they need is in the hit record (96 B, coalesced).
//
// Ray queues and hit lists are compacted with ballot + popc and one atomic per warp and counter.  Nothing depends on
// the order of a queue — every ray carries the record that waits for its colour — so a scene may ask for TWO bins
// (FrameParams::two_bins): rays that leave a BVH item and rays that leave an always-tested primitive (the planes).
// The former start inside the BVH's root box and walk it, most of the latter pass it by.  Measured (profiles/README.md,
// r02c): -10 % frame time where a BVH item is expensive to test (CSG programs), +-2 % on the mesh and sphere scenes
// (their walkers' path lengths differ too much for a warp of walkers to stay full), so only CSG scenes use it.
//
// The reference evaluates the secondary rays once per light inside the light loop (src/world.rs:58-79); here each
// child is traced once and its colour is added once per light.
// Built with -fmad=false: see device_math.cuh.
#include "device_pattern.cuh"
#include "device_trace.cuh"
#include "render_kernels.h"

#include <algorithm>

namespace rtcdev {

__device__ __forceinline__ double powi5(double x) {  // Rust powi(5): x * ((x*x) * (x*x))
    const double x2 = x * x;
    return x * (x2 * x2);
}

// src/intersection.rs:24-39
__device__ __forceinline__ double schlick(V3 eye, V3 normal, double n1, double n2) {
    double c = dot(eye, normal);
    if (n1 > n2) {
        const double n = n1 / n2;
        const double sin2_t = (n * n) * (1.0 - c * c);
        if (sin2_t > 1.0) return 1.0;
        c = sqrt(1.0 - sin2_t);
    }
    const double q = (n1 - n2) / (n1 + n2);
    const double r0 = q * q;
    return r0 + (1.0 - r0) * powi5(1.0 - c);
}

// Pixel of output slot `ordinal` of this shard (tiles dealt round-robin to shards; see rtc_shard_pixels).
__device__ __forceinline__ void tile_pixel(const FrameParams& f, uint32_t ordinal, int& px, int& py) {
    const uint32_t per_tile = (uint32_t)(f.tile_w * f.tile_h);
    const uint32_t k = ordinal / per_tile;
    const uint32_t r = ordinal - k * per_tile;
    const uint32_t tile = (uint32_t)f.shard_index + k * (uint32_t)f.shard_count;
    const uint32_t ty = tile / (uint32_t)f.tiles_x, tx = tile - ty * (uint32_t)f.tiles_x;
    const uint32_t ry = r / (uint32_t)f.tile_w, rx = r - ry * (uint32_t)f.tile_w;
    px = (int)(tx * (uint32_t)f.tile_w + rx);
    py = (int)(ty * (uint32_t)f.tile_h + ry);
}

// Camera::ray_at_pixel (src/camera.rs:39-55) with the sub-sample extension of SURVEY.md §8(d)
__device__ __forceinline__ Ray camera_ray(const FrameParams& f, int px, int py, int sub) {
    double sx = 0.5, sy = 0.5;
    if (f.spp > 1) {
        const int j = sub / f.spp_x, i = sub - j * f.spp_x;
        sx = ((double)i + 0.5) / (double)f.spp_x;
        sy = ((double)j + 0.5) / (double)f.spp_y;
    }
    const double fx = (double)(f.win_x0 + px) + sx;
    const double fy = (double)(f.win_y0 + py) + sy;
    const double xoffset = fx * f.pixel_size;
    const double yoffset = fy * f.pixel_size;
    const double world_x = f.half_width - xoffset;
    const double world_y = f.half_height - yoffset;
    const V3 pixel = xf_point(f.cam_inv, mk(world_x, world_y, -1.0));
    Ray r;
    r.o = mk(f.cam_inv[3], f.cam_inv[7], f.cam_inv[11]);
    r.d = normalize(pixel - r.o);
    return r;
}

// Ray `idx` of the level's input.  Returns false for the padding slots of a partial tile (primary level only).
// parent: the record (level 0: the sample) that waits for this ray's colour; kind: HIT_KIND_REFRACTED or 0.
template <bool PRIMARY>
__device__ __forceinline__ bool load_ray(const FrameParams& f, const LevelParams& lv, uint32_t idx, uint32_t n_front, Ray& ray, uint32_t& parent, uint32_t& kind, int& fuel) {
    if (PRIMARY) {
        const uint32_t s = (uint32_t)lv.first_sample + idx;
        uint32_t pix = s, sub = 0;
        if (f.spp > 1) { pix = s / (uint32_t)f.spp; sub = s - pix * (uint32_t)f.spp; }
        int px, py;
        tile_pixel(f, pix, px, py);
        parent = idx;
        kind = 0;
        fuel = f.fuel;
        if (px >= f.win_w || py >= f.win_h) return false;
        ray = camera_ray(f, px, py, (int)sub);
        return true;
    }
    // front bin: slots 0 .. n_front-1; back bin: the last slots of the queue, downwards
    const size_t at = idx < n_front ? (size_t)idx : (size_t)lv.in_capacity - 1u - (idx - n_front);
    const double2 a = lv.in_q0[at], b = lv.in_q1[at], c = lv.in_q2[at], e = lv.in_q3[at];
    ray.o = mk(a.x, a.y, b.x);
    ray.d = mk(b.y, c.x, c.y);
    const uint64_t bits = (uint64_t)__double_as_longlong(e.x);
    parent = (uint32_t)(bits & 0xffffffffu);
    fuel = (int)((bits >> 32) & 0xffu);
    kind = (uint32_t)((bits >> 40) & 1u);
    return true;
}

// Rays of the level's input: both bins of the queue (n_front of them in the front bin).
template <bool PRIMARY>
__device__ __forceinline__ uint32_t level_rays(const LevelParams& lv, uint32_t& n_front) {
    if (PRIMARY) { n_front = (uint32_t)lv.n_primary; return n_front; }
    const unsigned long long nf = *lv.in_count, nb = *lv.in_count_back;
    n_front = (uint32_t)(nf < lv.in_capacity ? nf : lv.in_capacity);
    const uint64_t room = lv.in_capacity - n_front;
    return n_front + (uint32_t)(nb < room ? nb : room);
}

// Adds a thread's tally to the batch counters: warp reduce, one atomic per counter per warp.
__device__ __forceinline__ void flush_tally(const Tally& t, unsigned long long* stats, unsigned lane) {
    uint32_t v[4] = {t.nodes, t.tris, t.prims, t.xforms};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        unsigned long long x = v[k];
        for (int off = 16; off; off >>= 1) x += __shfl_down_sync(0xffffffffu, x, off);
        if (lane == 0 && x) atomicAdd(stats + 3 + k, x);
    }
}

// Geometric half of Intersection::prepare_state (src/intersection.rs:50-68)
struct Surface {
    V3 point, eye, normal, over_point;
};
__device__ __forceinline__ Surface surface_at(const Scene& sc, const Prim& prim, const Ray& ray, double t, double u, double v) {
    Surface s;
    s.point = ray.o + ray.d * t;
    s.eye = -ray.d;
    s.normal = prim_normal(sc, prim, s.point, u, v);
    if (dot(s.normal, s.eye) < 0.0) s.normal = -s.normal;
    s.over_point = s.point + (s.normal * RTC_EPSILON);
    return s;
}

// The hits of a batch, enumerated bin by bin: the front bins (hits on BVH items) of all levels, then the back bins
// (hits on always-tested primitives).  Range r covers packed indices [first[r], first[r + 1]); its records sit at
// base[r] + step[r] * k (front bins grow up from the start of the level's segment, back bins down from its end).
struct HitRanges {
    uint32_t first[2 * MAX_LEVELS + 1];
    uint32_t base[2 * MAX_LEVELS];
    int n;
};
__device__ __forceinline__ void level_bins(const HitList& hl, int level, uint32_t& n_front, uint32_t& n_back) {
    const uint32_t cap = hl.seg_cap[level];
    const unsigned long long cf = hl.counts[level], cb = hl.counts_back[level];
    n_front = (uint32_t)(cf < cap ? cf : cap);
    n_back = (uint32_t)(cb < cap - n_front ? cb : cap - n_front);
}
__device__ __forceinline__ uint32_t hit_ranges(const HitList& hl, HitRanges& hr) {
    uint32_t total = 0;
    hr.n = 2 * hl.n_levels;
    for (int k = 0; k < hl.n_levels; ++k) {
        uint32_t nf, nb;
        level_bins(hl, k, nf, nb);
        hr.first[k] = nf;                  // (counts for now: prefix sums below)
        hr.first[hl.n_levels + k] = nb;
        hr.base[k] = hl.seg_off[k];
        hr.base[hl.n_levels + k] = hl.seg_off[k] + hl.seg_cap[k] - 1u;
    }
    for (int r = 0; r < hr.n; ++r) {
        const uint32_t n = hr.first[r];
        hr.first[r] = total;
        total += n;
    }
    hr.first[hr.n] = total;
    return total;
}
// r: the range the caller's previous (smaller) index fell into — a thread's indices only grow, so the search resumes there
__device__ __forceinline__ size_t hit_at(const HitRanges& hr, uint32_t j, int& r) {
    while (r + 1 < hr.n && j >= hr.first[r + 1]) ++r;
    const uint32_t k = j - hr.first[r];
    return r < (hr.n >> 1) ? (size_t)hr.base[r] + k : (size_t)hr.base[r] - k;
}
// Record of hit h of one level (light): front bin first.
__device__ __forceinline__ size_t level_hit_at(const HitList& hl, int level, uint32_t n_front, uint32_t h) {
    return h < n_front ? (size_t)hl.seg_off[level] + h : (size_t)hl.seg_off[level] + hl.seg_cap[level] - 1u - (h - n_front);
}

// Where the colour of the `kind` child of record `rec` is stored (HitList::child_rgb).
__device__ __forceinline__ double* child_slot(const HitList& hl, uint32_t rec, uint32_t kind) {
    return hl.child_rgb + 3 * (hl.fan == 2 ? 2 * (size_t)rec + kind : (size_t)rec);
}
// A ray's colour goes to its parent's child slot, or — primary rays — to the sample.
__device__ __forceinline__ void store_ray_color(const HitList& hl, const LevelParams& lv, int level, uint32_t parent, uint32_t kind, double r, double g, double b) {
    double* dst = level == 0 ? lv.accum + 3 * (size_t)parent : child_slot(hl, parent, kind);
    dst[0] = r;
    dst[1] = g;
    dst[2] = b;
}

// ---- kernel 1: closest hit + surface frame + children --------------------------------------------------------
// One thread per ray of the level.  For a hit it computes the geometric half of prepare_state once, appends the hit
// record to this level's segment of the hit list and pushes the reflected / refracted children to the next level's
// queue (src/world.rs:66-78,84-132) — so the next level depends on this kernel only, and the shadow / lighting work
// of all levels is done afterwards.
template <int FEAT, bool PRIMARY>
__global__ void __launch_bounds__(RTC_BLOCK, (FEAT & FEAT_CSG) ? RTC_TRACE_MIN_BLOCKS_CSG : (FEAT & FEAT_TRANSPARENT) ? RTC_TRACE_MIN_BLOCKS_TRANSPARENT : RTC_TRACE_MIN_BLOCKS) trace_spawn(const __grid_constant__ Scene sc, const __grid_constant__ FrameParams f, const __grid_constant__ LevelParams lv, const __grid_constant__ HitList hl) {
    const bool brute = (f.flags & 1u) != 0;
    uint32_t n_front_in = 0;
    const uint32_t n_in = level_rays<PRIMARY>(lv, n_front_in);
    const unsigned lane = threadIdx.x & 31u;
    const uint32_t stride = gridDim.x * blockDim.x;
    const int L = sc.n_lights;
    const uint32_t seg = hl.seg_off[lv.level], seg_cap = hl.seg_cap[lv.level];
    // every lane of a warp runs the same number of iterations so the warp-aggregated pushes stay convergent
    const uint32_t n_round = (n_in + 31u) & ~31u;
    // Counters cost registers in the hottest loop of the program: the work tallies exist only under RTC_FLAG_PROFILE
    // (FEAT_TALLY), the reflected / refracted split only where both kinds occur (otherwise the host reads the number
    // of children off the queue counters).
    constexpr bool TALLY = (FEAT & FEAT_TALLY) != 0;
    constexpr bool SPLIT = (FEAT & FEAT_TRANSPARENT) != 0;
#ifdef RTC_STAGE_TOP
    __shared__ __align__(128) BvhNode s_top[MAX_TOP_NODES + 1];
    __shared__ unsigned long long s_top_bar;
    const TopNodes top = stage_top_nodes(sc, s_top, &s_top_bar);
#else
    const TopNodes top{};
#endif
    constexpr bool ONEPASS = (FEAT & FEAT_ONEPASS) != 0;  // the closest-hit walk gathers the crossings of the n1 / n2 replay (device_trace.cuh)
    Cross xs[ONEPASS ? MAX_CROSSINGS : 1];
    int n_xs = 0;
    uint32_t c_active = 0, longest = 0, c_refl = 0, c_refr = 0;
    Tally tally = {0u, 0u, 0u, 0u};
    for (uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x; idx < n_round; idx += stride) {
        Ray ray;
        uint32_t parent = 0, kind = 0;
        int fuel = 0;
        const bool active = idx < n_in && load_ray<PRIMARY>(f, lv, idx, n_front_in, ray, parent, kind, fuel);
        Hit hit;
        hit.prim = -1;
        bool push_refl = false, push_refr = false, front = false, front_q = false;
        Ray r_refl, r_refr;
        double reflectance = 0.0;
        Surface s;
        if (active) {
            if (PRIMARY) c_active++;
            const uint32_t nodes_before = tally.nodes;
            if (ONEPASS) hit = closest_hit_collect<FEAT>(sc, ray, brute, xs, n_xs, tally, lv.overflow, top);
            else hit = closest_hit<FEAT>(sc, ray, CUDART_INF, brute, tally, lv.overflow, top);
            if (TALLY) longest = max(longest, tally.nodes - nodes_before);
            if (hit.prim >= 0) {
                const Prim prim = load_prim(sc, hit.prim);
                front = f.two_bins != 0 && (prim.kind_flags & PRIM_IN_BVH) != 0;
                front_q = front;
                const Material& mat = sc.materials[prim.material];
                const double reflective = mat.reflective, transparency = mat.transparency;
                double n1 = 1.0, n2 = 1.0;
                if ((FEAT & FEAT_TRANSPARENT) && transparency != 0.0) {
                    if (ONEPASS) {
                        IndexResult res;
                        replay_containers(sc, xs, n_xs, hit.t, hit.prim, lv.overflow, &res);
                        n1 = res.n1;
                        n2 = res.n2;
                    } else refractive_indices<FEAT>(sc, ray, hit.t, hit.prim, brute, n1, n2, tally, lv.overflow);
                }
                s = surface_at(sc, prim, ray, hit.t, hit.u, hit.v);
                if ((FEAT & FEAT_TRANSPARENT) && transparency != 0.0) reflectance = schlick(s.eye, s.normal, n1, n2);
                // With no light World::shade_hit adds nothing at all (src/world.rs:58): the children would be traced for nothing.
                if (fuel > 0 && L > 0) {
                    if (reflective != 0.0) {
                        push_refl = true;
                        r_refl.o = s.over_point;
                        r_refl.d = reflect(ray.d, s.normal);
                    }
                    if ((FEAT & FEAT_TRANSPARENT) && transparency != 0.0) {
                        const double n_ratio = n1 / n2;
                        const double cos_i = dot(s.eye, s.normal);
                        const double sin2_t = (n_ratio * n_ratio) * (1.0 - cos_i * cos_i);
                        if (!(sin2_t > 1.0)) {
                            const double cos_t = sqrt(1.0 - sin2_t);
                            push_refr = true;
                            r_refr.o = s.point - (s.normal * RTC_EPSILON);
                            r_refr.d = s.normal * (n_ratio * cos_i - cos_t) - s.eye * n_ratio;
                        }
                    }
                }
            } else {
                store_ray_color(hl, lv, lv.level, parent, kind, 0.0, 0.0, 0.0);  // World::color_at: a miss is black (src/world.rs:146-148)
            }
        }
        const unsigned below = (1u << lane) - 1u;
        // Slots for the hit records and for the children, per bin (front: the hit primitive is a BVH item): one atomic
        // per warp, bin and counter, all issued by lane 0 back to back so that their round trips to L2 overlap (r01f:
        // dependent round trips were 6-12 % of the kernel's stall samples).
        const bool is_hit = active && hit.prim >= 0;
        const unsigned m = __ballot_sync(0xffffffffu, is_hit);
        const unsigned m_front = __ballot_sync(0xffffffffu, is_hit && front);
        const unsigned q_front = __ballot_sync(0xffffffffu, is_hit && front_q);
        const unsigned m_refl = __ballot_sync(0xffffffffu, push_refl);
        const unsigned m_refr = (FEAT & FEAT_TRANSPARENT) ? __ballot_sync(0xffffffffu, push_refr) : 0u;
        const unsigned hf = m & m_front, hb = m & ~m_front;
        const int n_qf = __popc(m_refl & q_front) + __popc(m_refr & q_front), n_qb = __popc(m_refl & ~q_front) + __popc(m_refr & ~q_front);
        unsigned long long base_hf = 0, base_hb = 0, base_qf = 0, base_qb = 0;
        if (lane == 0) {
            if (hf) base_hf = atomicAdd(lv.hit_count, (unsigned long long)__popc(hf));
            if (hb) base_hb = atomicAdd(lv.hit_count_back, (unsigned long long)__popc(hb));
            if (n_qf) base_qf = atomicAdd(lv.out_count, (unsigned long long)n_qf);
            if (n_qb) base_qb = atomicAdd(lv.out_count_back, (unsigned long long)n_qb);
        }
        base_hf = __shfl_sync(0xffffffffu, base_hf, 0);
        base_hb = __shfl_sync(0xffffffffu, base_hb, 0);
        base_qf = __shfl_sync(0xffffffffu, base_qf, 0);
        base_qb = __shfl_sync(0xffffffffu, base_qb, 0);
        const unsigned mine = front ? m_front : ~m_front;  // the lanes of this lane's bin
        // ---- hit record ----
        uint32_t record = 0;
        if (is_hit) {
            // the two bins grow towards each other inside the level's segment, which is sized for "every ray hits"
            const uint64_t k = (front ? base_hf : base_hb) + (uint64_t)__popc(m & mine & below);
            if (k < seg_cap) {
                const size_t at = front ? (size_t)seg + k : (size_t)seg + seg_cap - 1u - k;
                record = (uint32_t)at;
                const uint32_t bits = kind | (push_refl ? HIT_SPAWNED_REFLECTED : 0u) | (push_refr ? HIT_SPAWNED_REFRACTED : 0u);
                hl.a[at] = make_double2(s.over_point.x, s.over_point.y);
                hl.b[at] = make_double2(s.over_point.z, s.normal.x);
                hl.c[at] = make_double2(s.normal.y, s.normal.z);
                hl.d[at] = make_double2(s.eye.x, s.eye.y);
                hl.e[at] = make_double2(s.eye.z, reflectance);
                hl.f[at] = make_double2(__longlong_as_double((long long)((uint64_t)parent << 32 | (uint32_t)hit.prim)), __longlong_as_double((long long)bits));
            } else {
                *lv.overflow = 2;
                push_refl = false;
                push_refr = false;
            }
        }
        // ---- children: warp-aggregated compaction into the next level's queue, same bin as their parent's record ----
        if (push_refl || push_refr) {
            const uint64_t link = (uint64_t)(uint32_t)(fuel - 1) << 32 | record;
            const uint64_t base = front_q ? base_qf : base_qb;
            const unsigned mine_q = front_q ? q_front : ~q_front;
            if (push_refl) {
                const uint64_t k = base + (uint64_t)__popc(m_refl & mine_q & below);
                if (k < lv.out_capacity) {
                    const size_t slot = front_q ? (size_t)k : (size_t)lv.out_capacity - 1u - k;
                    lv.out_q0[slot] = make_double2(r_refl.o.x, r_refl.o.y);
                    lv.out_q1[slot] = make_double2(r_refl.o.z, r_refl.d.x);
                    lv.out_q2[slot] = make_double2(r_refl.d.y, r_refl.d.z);
                    lv.out_q3[slot] = make_double2(__longlong_as_double((long long)link), 0.0);
                } else {
                    *lv.overflow = 2;
                    store_ray_color(hl, lv, lv.level + 1, record, 0u, 0.0, 0.0, 0.0);
                }
                if (SPLIT) c_refl++;
            }
            if ((FEAT & FEAT_TRANSPARENT) && push_refr) {
                const uint64_t k = base + (uint64_t)__popc(m_refl & mine_q) + (uint64_t)__popc(m_refr & mine_q & below);
                if (k < lv.out_capacity) {
                    const size_t slot = front_q ? (size_t)k : (size_t)lv.out_capacity - 1u - k;
                    lv.out_q0[slot] = make_double2(r_refr.o.x, r_refr.o.y);
                    lv.out_q1[slot] = make_double2(r_refr.o.z, r_refr.d.x);
                    lv.out_q2[slot] = make_double2(r_refr.d.y, r_refr.d.z);
                    lv.out_q3[slot] = make_double2(__longlong_as_double((long long)(link | (uint64_t)HIT_KIND_REFRACTED << 40)), 0.0);
                } else {
                    *lv.overflow = 2;
                    store_ray_color(hl, lv, lv.level + 1, record, HIT_KIND_REFRACTED, 0.0, 0.0, 0.0);
                }
                if (SPLIT) c_refr++;
            }
        }
    }
    if (TALLY) flush_tally(tally, lv.stats, lane);
    for (int off = 16; off; off >>= 1) {
        if (TALLY) longest = max(longest, __shfl_down_sync(0xffffffffu, longest, off));
        if (SPLIT) c_refl += __shfl_down_sync(0xffffffffu, c_refl, off);
        if (SPLIT) c_refr += __shfl_down_sync(0xffffffffu, c_refr, off);
        if (PRIMARY) c_active += __shfl_down_sync(0xffffffffu, c_active, off);
    }
    if (lane == 0) {
        if (TALLY && longest) atomicMax(lv.stats + 7, (unsigned long long)longest);
        if (SPLIT && c_refl) atomicAdd(lv.stats + 1, (unsigned long long)c_refl);
        if (SPLIT && c_refr) atomicAdd(lv.stats + 2, (unsigned long long)c_refr);
        if (PRIMARY && c_active) atomicAdd(lv.stats + 0, (unsigned long long)c_active);
    }
}

// ---- kernel 2: World::is_shadowed for every (hit, light) of the batch, all levels in one launch ----------------
// blockIdx.y = light, threads stride over the hits: consecutive threads take consecutive hits towards the SAME light, so
// a warp's rays start next to each other and head the same way.  The hits are enumerated bin by bin (hits on BVH items
// of every level first, then hits on always-tested primitives: HitList), so a warp's shadow rays either all start
// inside the BVH's root box — and walk — or mostly pass it by.
// A light behind the surface (light . normal < 0, the same expression Shape::lighting evaluates, src/shape.rs:449) gets
// ambient light only whatever is_shadowed says: such a query is answered without tracing anything.
template <int FEAT, bool ANYHIT>
__global__ void __launch_bounds__(RTC_BLOCK, (FEAT & FEAT_CSG) ? RTC_SHADOW_MIN_BLOCKS_CSG : RTC_SHADOW_MIN_BLOCKS) trace_shadow(const __grid_constant__ Scene sc, const __grid_constant__ FrameParams f, const __grid_constant__ LevelParams lv, const __grid_constant__ HitList hl) {
    const bool brute = (f.flags & 1u) != 0;
#ifdef RTC_STAGE_TOP
    __shared__ __align__(128) BvhNode s_top[MAX_TOP_NODES + 1];
    __shared__ unsigned long long s_top_bar;
    const TopNodes top = stage_top_nodes(sc, s_top, &s_top_bar);
#else
    const TopNodes top{};
#endif
    __shared__ HitRanges hr;  // (one copy per block: per-thread copies of the two small tables sat in local memory)
    __shared__ uint32_t s_hits;
    if (threadIdx.x == 0) s_hits = hit_ranges(hl, hr);
    __syncthreads();
    const uint32_t n_hits = s_hits;
    const uint32_t li = blockIdx.y;
    const uint32_t stride = gridDim.x * blockDim.x;
    Tally tally = {0u, 0u, 0u, 0u};
    uint32_t skipped = 0;
    const Light& light = sc.lights[li];
    const V3 lpos = mk(light.origin[0], light.origin[1], light.origin[2]);
    int range = 0;
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n_hits; j += stride) {
        const size_t at = hit_at(hr, j, range);
        const double2 a = hl.a[at], b = hl.b[at], c = hl.c[at];
        const V3 over_point = mk(a.x, a.y, b.x), normal = mk(b.y, c.x, c.y);
        const V3 to_light = lpos - over_point;
        const double distance = magnitude(to_light);
        Ray sray;
        sray.o = over_point;
        sray.d = normalize(to_light);
        bool shadowed = true;
        if (dot(sray.d, normal) >= 0.0) {
            if (ANYHIT) {
                shadowed = any_hit<FEAT>(sc, sray, distance, brute, tally, lv.overflow, top);
            } else {  // some shape does not cast shadows: the nearest crossing decides (src/world.rs:43-47)
                const Hit sh = closest_hit<FEAT>(sc, sray, distance, brute, tally, lv.overflow, top);
                shadowed = sh.prim >= 0 && (__ldg(&sc.prims[sh.prim].kind_flags) & 0x100) != 0;
            }
        } else skipped++;
        lv.shadow[(size_t)li * hl.total_cap + at] = shadowed ? 1 : 0;
    }
    const unsigned lane = threadIdx.x & 31u;
    if (FEAT & FEAT_TALLY) flush_tally(tally, lv.stats, lane);
    for (int off = 16; off; off >>= 1) skipped += __shfl_down_sync(0xffffffffu, skipped, off);
    if (lane == 0 && skipped) atomicAdd(lv.stats + 8, (unsigned long long)skipped);
}

// ---- kernel 4: World::shade_hit for the hits of one level ---------------------------------------------------------
// One launch per level, deepest level first, so the colours of a hit's children (HitList::child_rgb) are final when the
// hit is lit.  Per hit: the pattern colour at over_point, Shape::lighting per light (src/shape.rs:429-463) with the
// shadow bytes, and the light loop of World::shade_hit with its own association (src/world.rs:56-79):
//     color = black;  for each light:  color += surface + (both ? reflected * R + refracted * (1 - R) : reflected + refracted)
// where reflected = color_at(child) * reflective and refracted = color_at(child) * transparency (src/world.rs:100,129).
// The result is stored where the parent looks for it (level 0: the sample): one writer per colour, no atomics.
// PLAIN: every material of the scene has a Plain pattern (its colour is inlined in the material record), so the
// pattern evaluator — its frames, the noise code and the shared permutation table — is not compiled in.
template <bool PLAIN>
__global__ void __launch_bounds__(RTC_BLOCK, PLAIN ? RTC_SHADE_PLAIN_MIN_BLOCKS : RTC_SHADE_MIN_BLOCKS) light(const __grid_constant__ Scene sc, const __grid_constant__ LevelParams lv, const __grid_constant__ HitList hl) {
    __shared__ uint8_t s_perm[PLAIN ? 1 : 512];
    if (!PLAIN) {
        for (int i = threadIdx.x; i < 512; i += blockDim.x) s_perm[i] = c_perm[i];
        __syncthreads();
    }
    uint32_t n_front, n_back;
    level_bins(hl, lv.level, n_front, n_back);
    const uint32_t n_hits = n_front + n_back;
    const int L = sc.n_lights;
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t h = blockIdx.x * blockDim.x + threadIdx.x; h < n_hits; h += stride) {
        const size_t at = level_hit_at(hl, lv.level, n_front, h);
        const double2 a = hl.a[at], b = hl.b[at], c = hl.c[at], d = hl.d[at], e = hl.e[at], fr = hl.f[at];
        const uint64_t link = (uint64_t)__double_as_longlong(fr.x);
        const uint32_t bits = (uint32_t)(uint64_t)__double_as_longlong(fr.y);
        const V3 over_point = mk(a.x, a.y, b.x), normal = mk(b.y, c.x, c.y), eye = mk(d.x, d.y, e.x);
        const double reflectance = e.y;
        const uint32_t parent = (uint32_t)(link >> 32);
        const Prim prim = load_prim(sc, (int)(uint32_t)(link & 0xffffffffu));
        const Material& mat = sc.materials[prim.material];
        // ---- what the children returned (src/world.rs:84-132) ----
        Rgb reflected = rgb(0.0, 0.0, 0.0), refracted = rgb(0.0, 0.0, 0.0);
        if (bits & HIT_SPAWNED_REFLECTED) {
            const double* q = child_slot(hl, (uint32_t)at, 0u);
            reflected = rgb(q[0], q[1], q[2]) * mat.reflective;
        }
        if (bits & HIT_SPAWNED_REFRACTED) {
            const double* q = child_slot(hl, (uint32_t)at, HIT_KIND_REFRACTED);
            refracted = rgb(q[0], q[1], q[2]) * mat.transparency;
        }
        Rgb bounce;
        if (mat.reflective > 0.0 && mat.transparency > 0.0) bounce = reflected * reflectance + refracted * (1.0 - reflectance);
        else bounce = reflected + refracted;
        // ---- the light loop ----
        Rgb color;
        if (PLAIN || (mat.flags & 1)) color = rgb(mat.plain[0], mat.plain[1], mat.plain[2]);
        else {
            V3 pp = over_point;
            double pw = 1.0;
            xf_point4(sc.matinvs[prim.matinv].m, pp, pw);
            color = pattern_color(sc, s_perm, mat.pattern, pp, pw);
        }
        const double specular_k = mat.specular, shininess = mat.shininess;
        Rgb sum = rgb(0.0, 0.0, 0.0);
        for (int li = 0; li < L; ++li) {
            const Light& light = sc.lights[li];
            const V3 lpos = mk(light.origin[0], light.origin[1], light.origin[2]);
            const Rgb intensity = rgb(light.intensity[0], light.intensity[1], light.intensity[2]);
            const V3 lightv = normalize(lpos - over_point);
            const bool shadowed = lv.shadow[(size_t)li * hl.total_cap + at] != 0;
            const Rgb effective = color * intensity;
            Rgb surface = effective * mat.ambient;
            const double light_dot_normal = dot(lightv, normal);
            if (!shadowed && light_dot_normal >= 0.0) {
                const Rgb diffuse = effective * mat.diffuse * light_dot_normal;
                Rgb specular = rgb(0.0, 0.0, 0.0);
                const double reflect_dot_eye = dot(reflect(-lightv, normal), eye);
                // intensity * 0 * powf(..) is +0 for any finite power: skip the pow when the material has no highlight
                if (reflect_dot_eye > 0.0 && specular_k != 0.0) specular = intensity * specular_k * pow(reflect_dot_eye, shininess);
                surface = surface + diffuse + specular;
            } else {
                surface = surface + rgb(0.0, 0.0, 0.0) + rgb(0.0, 0.0, 0.0);
            }
            sum = sum + (surface + bounce);
        }
        store_ray_color(hl, lv, lv.level, parent, bits & HIT_KIND_REFRACTED, sum.r, sum.g, sum.b);
    }
}

// Average the spp sub-samples of each pixel (fixed j-major, i-minor order) and store the pixel.
__global__ void resolve_pixels(const FrameParams f, const double* __restrict__ accum, uint64_t first_pixel, uint64_t n_pixels,
                               double* __restrict__ out, int row_major) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pixels) return;
    const double* a = accum + 3 * (size_t)(i * (uint64_t)f.spp);
    double r = 0.0, g = 0.0, b = 0.0;
    for (int s = 0; s < f.spp; ++s) { r = r + a[3 * s]; g = g + a[3 * s + 1]; b = b + a[3 * s + 2]; }
    const double inv = 1.0 / (double)f.spp;
    r = r * inv; g = g * inv; b = b * inv;
    const uint64_t ordinal = first_pixel + i;
    uint64_t dst = ordinal;
    if (row_major) {
        int px, py;
        tile_pixel(f, (uint32_t)ordinal, px, py);
        if (px >= f.win_w || py >= f.win_h) return;
        dst = (uint64_t)py * (uint64_t)f.win_w + (uint64_t)px;
    }
    out[3 * dst] = r;
    out[3 * dst + 1] = g;
    out[3 * dst + 2] = b;
}

void upload_perm(const uint8_t* perm512) { cudaMemcpyToSymbol(c_perm, perm512, 512); }

template <int FEAT>
static void launch_trace_feat(bool primary, int grid, cudaStream_t st, const Scene& sc, const FrameParams& f, const LevelParams& lv, const HitList& hl) {
    if (primary) trace_spawn<FEAT, true><<<grid, RTC_BLOCK, 0, st>>>(sc, f, lv, hl);
    else trace_spawn<FEAT, false><<<grid, RTC_BLOCK, 0, st>>>(sc, f, lv, hl);
}
template <int TALLY>
static void launch_trace_tally(int feat, bool primary, int grid, cudaStream_t st, const Scene& sc, const FrameParams& f, const LevelParams& lv, const HitList& hl) {
    const bool onepass = (feat & FEAT_ONEPASS) != 0;
    if (feat & (FEAT_CSG | FEAT_GATES)) {
        if (onepass) launch_trace_feat<FEAT_TRANSPARENT | FEAT_CSG | FEAT_GATES | FEAT_ONEPASS | TALLY>(primary, grid, st, sc, f, lv, hl);
        else launch_trace_feat<FEAT_TRANSPARENT | FEAT_CSG | FEAT_GATES | TALLY>(primary, grid, st, sc, f, lv, hl);
    } else if (feat & FEAT_TRANSPARENT) {
        if (onepass) launch_trace_feat<FEAT_TRANSPARENT | FEAT_ONEPASS | TALLY>(primary, grid, st, sc, f, lv, hl);
        else launch_trace_feat<FEAT_TRANSPARENT | TALLY>(primary, grid, st, sc, f, lv, hl);
    } else if (feat & FEAT_TRIS) launch_trace_feat<FEAT_TRIS | TALLY>(primary, grid, st, sc, f, lv, hl);
    else launch_trace_feat<TALLY>(primary, grid, st, sc, f, lv, hl);
}
// feat: the scene's FEAT_* bits, plus FEAT_TALLY when the caller wants the work counted
void launch_trace(int feat, bool primary, int grid, cudaStream_t st, const Scene& sc, const FrameParams& f, const LevelParams& lv, const HitList& hl) {
    if (feat & FEAT_TALLY) launch_trace_tally<FEAT_TALLY>(feat, primary, grid, st, sc, f, lv, hl);
    else launch_trace_tally<0>(feat, primary, grid, st, sc, f, lv, hl);
}

template <int TALLY>
static void launch_shadow_tally(int feat, dim3 grid, cudaStream_t st, const Scene& sc, const FrameParams& f, const LevelParams& lv, const HitList& hl) {
    const bool full = (feat & (FEAT_CSG | FEAT_GATES)) != 0;
    const bool tris = !full && (feat & FEAT_TRIS) != 0;
    if (sc.all_cast_shadow) {
        if (full) trace_shadow<FEAT_CSG | FEAT_GATES | TALLY, true><<<grid, RTC_BLOCK, 0, st>>>(sc, f, lv, hl);
        else if (tris) trace_shadow<FEAT_TRIS | TALLY, true><<<grid, RTC_BLOCK, 0, st>>>(sc, f, lv, hl);
        else trace_shadow<TALLY, true><<<grid, RTC_BLOCK, 0, st>>>(sc, f, lv, hl);
    } else {
        if (full) trace_shadow<FEAT_CSG | FEAT_GATES | TALLY, false><<<grid, RTC_BLOCK, 0, st>>>(sc, f, lv, hl);
        else if (tris) trace_shadow<FEAT_TRIS | TALLY, false><<<grid, RTC_BLOCK, 0, st>>>(sc, f, lv, hl);
        else trace_shadow<TALLY, false><<<grid, RTC_BLOCK, 0, st>>>(sc, f, lv, hl);
    }
}
void launch_shadow(int feat, int grid, cudaStream_t st, const Scene& sc, const FrameParams& f, const LevelParams& lv, const HitList& hl) {
    if (sc.n_lights == 0) return;
    const dim3 g((unsigned)std::max(grid / sc.n_lights, 1), (unsigned)sc.n_lights);  // blockIdx.y = light
    if (feat & FEAT_TALLY) launch_shadow_tally<FEAT_TALLY>(feat, g, st, sc, f, lv, hl);
    else launch_shadow_tally<0>(feat, g, st, sc, f, lv, hl);
}

void launch_light(int grid, cudaStream_t st, const Scene& sc, const LevelParams& lv, const HitList& hl) {
    if (sc.plain_only) light<true><<<grid, RTC_BLOCK, 0, st>>>(sc, lv, hl);
    else light<false><<<grid, RTC_BLOCK, 0, st>>>(sc, lv, hl);
}

void launch_resolve(cudaStream_t st, const FrameParams& f, const double* accum, uint64_t first_pixel, uint64_t n_pixels, double* out, int row_major) {
    if (!n_pixels) return;
    const int block = 256;
    const uint64_t grid = (n_pixels + block - 1) / block;
    resolve_pixels<<<(unsigned)grid, block, 0, st>>>(f, accum, first_pixel, n_pixels, out, row_major);
}

}  // namespace rtcdev
