// Wavefront kernels of the render path.  Three launches per bounce level replace the recursion of
// World::color_at -> shade_hit -> reflected_color / refracted_color (src/world.rs:50-149):
//
//   trace_closest : one thread per ray of the level (level 0: Camera::ray_at_pixel, src/camera.rs:39-55, generated
//                   in-kernel; level k>0: the compacted queue the previous level wrote) -> closest hit
//                   (World::intersect + sort + hit, src/world.rs:18-24, src/intersection.rs:123-132) and, for hits
//                   on transparent materials, n1/n2 (src/intersection.rs:70-103) -> compacted hit list
//   trace_shadow  : one thread per (hit, light), light-major so a warp's rays run towards the same light:
//                   World::is_shadowed (src/world.rs:26-48) -> one byte per (hit, light)
//   shade         : one thread per hit: prepare_state (src/intersection.rs:50-121), pattern colour, Phong term per
//                   light (src/shape.rs:429-463), accumulate throughput * surface colour, push reflected / refracted
//                   children (scalar throughput, remaining fuel) to the next level's queue (src/world.rs:66-132)
//
// Each kernel holds exactly one BVH traversal (or none), so its instruction footprint stays small and the lanes of a
// warp run the same code; the hit list removes the lanes that would idle after a miss (profiles/README.md, r01a).
// The geometric part of prepare_state (point, normal, over_point) is recomputed by the shadow and the shade kernel
// from (ray, t, primitive, u, v) instead of being stored: ~100 flops against ~150 bytes of queue traffic.
//
// The reference evaluates the secondary rays once per light inside the light loop (src/world.rs:58-79), i.e. a
// child's colour is added L times; here each child is traced once with its weight multiplied by L.
// Built with -fmad=false: see device_math.cuh.
#include "device_pattern.cuh"
#include "device_trace.cuh"
#include "render_kernels.h"

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
template <bool PRIMARY>
__device__ __forceinline__ bool load_ray(const FrameParams& f, const LevelParams& lv, uint32_t idx, Ray& ray, double& w, uint32_t& sample, int& fuel) {
    if (PRIMARY) {
        const uint32_t s = (uint32_t)lv.first_sample + idx;
        uint32_t pix = s, sub = 0;
        if (f.spp > 1) { pix = s / (uint32_t)f.spp; sub = s - pix * (uint32_t)f.spp; }
        int px, py;
        tile_pixel(f, pix, px, py);
        if (px >= f.win_w || py >= f.win_h) return false;
        ray = camera_ray(f, px, py, (int)sub);
        w = 1.0;
        sample = idx;
        fuel = f.fuel;
        return true;
    }
    const double2 a = lv.in_q0[idx], b = lv.in_q1[idx], c = lv.in_q2[idx], e = lv.in_q3[idx];
    ray.o = mk(a.x, a.y, b.x);
    ray.d = mk(b.y, c.x, c.y);
    w = e.x;
    const uint64_t bits = (uint64_t)__double_as_longlong(e.y);
    sample = (uint32_t)(bits & 0xffffffffu);
    fuel = (int)(bits >> 32);
    return true;
}

template <bool PRIMARY>
__device__ __forceinline__ uint32_t level_rays(const LevelParams& lv) {
    if (PRIMARY) return (uint32_t)lv.n_primary;
    const unsigned long long n = *lv.in_count;
    return (uint32_t)(n < lv.in_capacity ? n : lv.in_capacity);
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

// ---- kernel 1: closest hit -> compacted hit list ---------------------------------------------------------------
template <int FEAT, bool PRIMARY>
__global__ void __launch_bounds__(RTC_BLOCK, RTC_TRACE_MIN_BLOCKS) trace_closest(const Scene sc, const FrameParams f, const LevelParams lv) {
    const bool brute = (f.flags & 1u) != 0;
    const uint32_t n_in = level_rays<PRIMARY>(lv);
    const unsigned lane = threadIdx.x & 31u;
    const uint32_t stride = gridDim.x * blockDim.x;
    // every lane of a warp runs the same number of iterations so the warp-aggregated push stays convergent
    const uint32_t n_round = (n_in + 31u) & ~31u;
    uint32_t c_active = 0, longest = 0;
    Tally tally = {0u, 0u, 0u, 0u};
    for (uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x; idx < n_round; idx += stride) {
        Ray ray;
        double w;
        uint32_t sample;
        int fuel;
        const bool active = idx < n_in && load_ray<PRIMARY>(f, lv, idx, ray, w, sample, fuel);
        Hit hit;
        hit.prim = -1;
        double n1 = 1.0, n2 = 1.0;
        if (active) {
            c_active++;
            const uint32_t nodes_before = tally.nodes;
            hit = closest_hit<FEAT>(sc, ray, CUDART_INF, brute, tally, lv.overflow);
            longest = max(longest, tally.nodes - nodes_before);
            if ((FEAT & FEAT_TRANSPARENT) && hit.prim >= 0) {
                const int material = __ldg(&sc.prims[hit.prim].material);
                if (sc.materials[material].transparency != 0.0)
                    refractive_indices<FEAT>(sc, ray, hit.t, hit.prim, brute, n1, n2, tally, lv.overflow);
            }
        }
        const bool is_hit = active && hit.prim >= 0;
        const unsigned m = __ballot_sync(0xffffffffu, is_hit);
        if (m) {
            unsigned long long base = 0;
            if (lane == 0) base = atomicAdd(lv.hit_count, (unsigned long long)__popc(m));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (is_hit) {
                const uint64_t slot = base + (uint64_t)__popc(m & ((1u << lane) - 1u));
                lv.hit_ray[slot] = idx;
                lv.hit0[slot] = make_double2(hit.t, __longlong_as_double((long long)hit.prim));
                lv.hit1[slot] = make_double2(hit.u, hit.v);
                if (FEAT & FEAT_TRANSPARENT) lv.hit2[slot] = make_double2(n1, n2);
            }
        }
    }
    flush_tally(tally, lv.stats, lane);
    for (int off = 16; off; off >>= 1) longest = max(longest, __shfl_down_sync(0xffffffffu, longest, off));
    if (lane == 0 && longest) atomicMax(lv.stats + 7, (unsigned long long)longest);
    if (PRIMARY) {
        for (int off = 16; off; off >>= 1) c_active += __shfl_down_sync(0xffffffffu, c_active, off);
        if (lane == 0 && c_active) atomicAdd(lv.stats + 0, (unsigned long long)c_active);
    }
}

// ---- kernel 2: World::is_shadowed for every (hit, light) ------------------------------------------------------
template <int FEAT, bool PRIMARY, bool ANYHIT>
__global__ void __launch_bounds__(RTC_BLOCK, RTC_TRACE_MIN_BLOCKS) trace_shadow(const Scene sc, const FrameParams f, const LevelParams lv) {
    const bool brute = (f.flags & 1u) != 0;
    const uint32_t n_hits = (uint32_t)*lv.hit_count;
    const uint32_t L = (uint32_t)sc.n_lights;
    const uint32_t total = n_hits * L;  // < 2^32: checked by the host (queue capacity * lights)
    const uint32_t stride = gridDim.x * blockDim.x;
    Tally tally = {0u, 0u, 0u, 0u};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const uint32_t li = i / n_hits;
        const uint32_t h = i - li * n_hits;
        const double2 h0 = lv.hit0[h], h1 = lv.hit1[h];
        const int prim_id = (int)__double_as_longlong(h0.y);
        Ray ray;
        double w;
        uint32_t sample;
        int fuel;
        load_ray<PRIMARY>(f, lv, lv.hit_ray[h], ray, w, sample, fuel);
        const Prim prim = load_prim(sc, prim_id);
        const Surface s = surface_at(sc, prim, ray, h0.x, h1.x, h1.y);
        const Light& light = sc.lights[li];
        const V3 to_light = mk(light.origin[0], light.origin[1], light.origin[2]) - s.over_point;
        const double distance = magnitude(to_light);
        Ray sray;
        sray.o = s.over_point;
        sray.d = normalize(to_light);
        bool shadowed;
        if (ANYHIT) {
            shadowed = any_hit<FEAT>(sc, sray, distance, brute, tally, lv.overflow);
        } else {  // some shape does not cast shadows: the nearest crossing decides (src/world.rs:43-47)
            const Hit sh = closest_hit<FEAT>(sc, sray, distance, brute, tally, lv.overflow);
            shadowed = sh.prim >= 0 && (__ldg(&sc.prims[sh.prim].kind_flags) & 0x100) != 0;
        }
        lv.shadow[i] = shadowed ? 1 : 0;
    }
    flush_tally(tally, lv.stats, threadIdx.x & 31u);
}

// ---- kernel 3: shading + children -----------------------------------------------------------------------------
template <bool TRANSPARENT, bool PRIMARY>
__global__ void __launch_bounds__(RTC_BLOCK, RTC_SHADE_MIN_BLOCKS) shade(const Scene sc, const FrameParams f, const LevelParams lv) {
    __shared__ uint8_t s_perm[512];
    for (int i = threadIdx.x; i < 512; i += blockDim.x) s_perm[i] = c_perm[i];
    __syncthreads();

    const uint32_t n_hits = (uint32_t)*lv.hit_count;
    const int L = sc.n_lights;
    const unsigned lane = threadIdx.x & 31u;
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t n_round = (n_hits + 31u) & ~31u;
    uint32_t c_refl = 0, c_refr = 0;
    for (uint32_t h = blockIdx.x * blockDim.x + threadIdx.x; h < n_round; h += stride) {
        bool push_refl = false, push_refr = false;
        Ray r_refl, r_refr;
        double w_refl = 0.0, w_refr = 0.0;
        uint32_t sample = 0;
        int fuel = 0;
        if (h < n_hits) {
            const double2 h0 = lv.hit0[h], h1 = lv.hit1[h];
            const int prim_id = (int)__double_as_longlong(h0.y);
            Ray ray;
            double w;
            load_ray<PRIMARY>(f, lv, lv.hit_ray[h], ray, w, sample, fuel);
            const Prim prim = load_prim(sc, prim_id);
            const Material& mat = sc.materials[prim.material];
            const Surface s = surface_at(sc, prim, ray, h0.x, h1.x, h1.y);
            const double reflective = mat.reflective, transparency = mat.transparency;
            double n1 = 1.0, n2 = 1.0, reflectance = 0.0;
            if (TRANSPARENT && transparency != 0.0) {
                const double2 h2 = lv.hit2[h];
                n1 = h2.x;
                n2 = h2.y;
                reflectance = schlick(s.eye, s.normal, n1, n2);
            }
            // ---- World::shade_hit light loop, src/world.rs:58-64 + Shape::lighting, src/shape.rs:429-463 ----
            Rgb color;
            if (mat.flags & 1) color = rgb(mat.plain[0], mat.plain[1], mat.plain[2]);
            else {
                V3 pp = s.over_point;
                double pw = 1.0;
                xf_point4(sc.matinvs[prim.matinv].m, pp, pw);
                color = pattern_color(sc, s_perm, mat.pattern, pp, pw);
            }
            Rgb sum = rgb(0.0, 0.0, 0.0);
            for (int li = 0; li < L; ++li) {
                const Light& light = sc.lights[li];
                const V3 lpos = mk(light.origin[0], light.origin[1], light.origin[2]);
                const Rgb intensity = rgb(light.intensity[0], light.intensity[1], light.intensity[2]);
                const V3 lightv = normalize(lpos - s.over_point);
                const bool shadowed = lv.shadow[(uint32_t)li * n_hits + h] != 0;
                const Rgb effective = color * intensity;
                Rgb surface = effective * mat.ambient;
                const double light_dot_normal = dot(lightv, s.normal);
                if (!shadowed && light_dot_normal >= 0.0) {
                    const Rgb diffuse = effective * mat.diffuse * light_dot_normal;
                    Rgb specular = rgb(0.0, 0.0, 0.0);
                    const double reflect_dot_eye = dot(reflect(-lightv, s.normal), s.eye);
                    if (reflect_dot_eye > 0.0) specular = intensity * mat.specular * pow(reflect_dot_eye, mat.shininess);
                    surface = surface + diffuse + specular;
                } else {
                    surface = surface + rgb(0.0, 0.0, 0.0) + rgb(0.0, 0.0, 0.0);
                }
                sum = sum + surface;
            }
            double* acc = lv.accum + 3 * (size_t)sample;
            atomicAdd(acc + 0, sum.r * w);
            atomicAdd(acc + 1, sum.g * w);
            atomicAdd(acc + 2, sum.b * w);
            // ---- children: src/world.rs:66-78,84-132 ----
            if (fuel > 0 && L > 0) {
                const bool both = reflective > 0.0 && transparency > 0.0;
                if (reflective != 0.0) {
                    push_refl = true;
                    r_refl.o = s.over_point;
                    r_refl.d = reflect(ray.d, s.normal);
                    w_refl = w * (double)L * reflective * (both ? reflectance : 1.0);
                }
                if (TRANSPARENT && transparency != 0.0) {
                    const double n_ratio = n1 / n2;
                    const double cos_i = dot(s.eye, s.normal);
                    const double sin2_t = (n_ratio * n_ratio) * (1.0 - cos_i * cos_i);
                    if (!(sin2_t > 1.0)) {
                        const double cos_t = sqrt(1.0 - sin2_t);
                        push_refr = true;
                        r_refr.o = s.point - (s.normal * RTC_EPSILON);
                        r_refr.d = s.normal * (n_ratio * cos_i - cos_t) - s.eye * n_ratio;
                        w_refr = w * (double)L * transparency * (both ? (1.0 - reflectance) : 1.0);
                    }
                }
            }
        }
        // ---- warp-aggregated compaction into the next level's queue: one atomic per warp ----
        const unsigned m_refl = __ballot_sync(0xffffffffu, push_refl);
        const unsigned m_refr = TRANSPARENT ? __ballot_sync(0xffffffffu, push_refr) : 0u;
        const int n_push = __popc(m_refl) + __popc(m_refr);
        if (n_push) {
            unsigned long long base = 0;
            if (lane == 0) base = atomicAdd(lv.out_count, (unsigned long long)n_push);
            base = __shfl_sync(0xffffffffu, base, 0);
            const unsigned below = (1u << lane) - 1u;
            const uint64_t packed_fuel = (uint64_t)(uint32_t)(fuel - 1) << 32 | sample;
            if (push_refl) {
                const uint64_t slot = base + (uint64_t)__popc(m_refl & below);
                if (slot < lv.out_capacity) {
                    lv.out_q0[slot] = make_double2(r_refl.o.x, r_refl.o.y);
                    lv.out_q1[slot] = make_double2(r_refl.o.z, r_refl.d.x);
                    lv.out_q2[slot] = make_double2(r_refl.d.y, r_refl.d.z);
                    lv.out_q3[slot] = make_double2(w_refl, __longlong_as_double((long long)packed_fuel));
                } else *lv.overflow = 2;
                c_refl++;
            }
            if (TRANSPARENT && push_refr) {
                const uint64_t slot = base + (uint64_t)__popc(m_refl) + (uint64_t)__popc(m_refr & below);
                if (slot < lv.out_capacity) {
                    lv.out_q0[slot] = make_double2(r_refr.o.x, r_refr.o.y);
                    lv.out_q1[slot] = make_double2(r_refr.o.z, r_refr.d.x);
                    lv.out_q2[slot] = make_double2(r_refr.d.y, r_refr.d.z);
                    lv.out_q3[slot] = make_double2(w_refr, __longlong_as_double((long long)packed_fuel));
                } else *lv.overflow = 2;
                c_refr++;
            }
        }
    }
    for (int off = 16; off; off >>= 1) {
        c_refl += __shfl_down_sync(0xffffffffu, c_refl, off);
        c_refr += __shfl_down_sync(0xffffffffu, c_refr, off);
    }
    if (lane == 0) {
        if (c_refl) atomicAdd(lv.stats + 1, (unsigned long long)c_refl);
        if (c_refr) atomicAdd(lv.stats + 2, (unsigned long long)c_refr);
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
static void launch_trace_feat(bool primary, int grid, cudaStream_t st, const Scene& sc, const FrameParams& f, const LevelParams& lv) {
    if (primary) trace_closest<FEAT, true><<<grid, RTC_BLOCK, 0, st>>>(sc, f, lv);
    else trace_closest<FEAT, false><<<grid, RTC_BLOCK, 0, st>>>(sc, f, lv);
}
void launch_trace(int feat, bool primary, int grid, cudaStream_t st, const Scene& sc, const FrameParams& f, const LevelParams& lv) {
    if (feat & (FEAT_CSG | FEAT_GATES)) launch_trace_feat<FEAT_TRANSPARENT | FEAT_CSG | FEAT_GATES>(primary, grid, st, sc, f, lv);
    else if (feat & FEAT_TRANSPARENT) launch_trace_feat<FEAT_TRANSPARENT>(primary, grid, st, sc, f, lv);
    else launch_trace_feat<0>(primary, grid, st, sc, f, lv);
}

template <int FEAT, bool ANYHIT>
static void launch_shadow_feat(bool primary, int grid, cudaStream_t st, const Scene& sc, const FrameParams& f, const LevelParams& lv) {
    if (primary) trace_shadow<FEAT, true, ANYHIT><<<grid, RTC_BLOCK, 0, st>>>(sc, f, lv);
    else trace_shadow<FEAT, false, ANYHIT><<<grid, RTC_BLOCK, 0, st>>>(sc, f, lv);
}
void launch_shadow(int feat, bool primary, int grid, cudaStream_t st, const Scene& sc, const FrameParams& f, const LevelParams& lv) {
    if (sc.n_lights == 0) return;
    const bool full = (feat & (FEAT_CSG | FEAT_GATES)) != 0;
    if (sc.all_cast_shadow) {
        if (full) launch_shadow_feat<FEAT_CSG | FEAT_GATES, true>(primary, grid, st, sc, f, lv);
        else launch_shadow_feat<0, true>(primary, grid, st, sc, f, lv);
    } else {
        if (full) launch_shadow_feat<FEAT_CSG | FEAT_GATES, false>(primary, grid, st, sc, f, lv);
        else launch_shadow_feat<0, false>(primary, grid, st, sc, f, lv);
    }
}

void launch_shade(int feat, bool primary, int grid, cudaStream_t st, const Scene& sc, const FrameParams& f, const LevelParams& lv) {
    const bool tr = (feat & FEAT_TRANSPARENT) != 0;
    if (tr) {
        if (primary) shade<true, true><<<grid, RTC_BLOCK, 0, st>>>(sc, f, lv);
        else shade<true, false><<<grid, RTC_BLOCK, 0, st>>>(sc, f, lv);
    } else {
        if (primary) shade<false, true><<<grid, RTC_BLOCK, 0, st>>>(sc, f, lv);
        else shade<false, false><<<grid, RTC_BLOCK, 0, st>>>(sc, f, lv);
    }
}

void launch_resolve(cudaStream_t st, const FrameParams& f, const double* accum, uint64_t first_pixel, uint64_t n_pixels, double* out, int row_major) {
    if (!n_pixels) return;
    const int block = 256;
    const uint64_t grid = (n_pixels + block - 1) / block;
    resolve_pixels<<<(unsigned)grid, block, 0, st>>>(f, accum, first_pixel, n_pixels, out, row_major);
}

}  // namespace rtcdev
