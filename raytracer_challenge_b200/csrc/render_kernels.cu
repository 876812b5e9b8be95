// Wavefront kernels of the render path.  One launch per bounce level replaces the recursion of
// World::color_at -> shade_hit -> reflected_color / refracted_color (src/world.rs:50-149):
//
//   level 0   : rays come from Camera::ray_at_pixel (src/camera.rs:39-55), generated in-kernel
//   level k>0 : rays come from the compacted queue the previous level wrote
//   per ray   : closest hit -> prepare_state (src/intersection.rs:50-121) -> one shadow query + Phong term per
//               light (src/world.rs:58-64, src/shape.rs:429-463) -> accumulate throughput * surface colour
//               -> push reflected / refracted children (scalar throughput, remaining fuel) to the next queue
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

__device__ __forceinline__ void tile_pixel(const FrameParams& f, uint64_t ordinal, int& px, int& py) {
    const uint32_t per_tile = (uint32_t)(f.tile_w * f.tile_h);
    const uint64_t k = ordinal / per_tile;
    const uint32_t r = (uint32_t)(ordinal % per_tile);
    const uint64_t tile = (uint64_t)f.shard_index + k * (uint64_t)f.shard_count;
    const int tx = (int)(tile % (uint64_t)f.tiles_x), ty = (int)(tile / (uint64_t)f.tiles_x);
    px = tx * f.tile_w + (int)(r % (uint32_t)f.tile_w);
    py = ty * f.tile_h + (int)(r / (uint32_t)f.tile_w);
}

// Camera::ray_at_pixel (src/camera.rs:39-55) with the sub-sample extension of SURVEY.md §8(d)
__device__ __forceinline__ Ray camera_ray(const FrameParams& f, int px, int py, int sub) {
    const int i = sub % f.spp_x, j = sub / f.spp_x;
    const double fx = (double)(f.win_x0 + px) + ((double)i + 0.5) / (double)f.spp_x;
    const double fy = (double)(f.win_y0 + py) + ((double)j + 0.5) / (double)f.spp_y;
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

template <int FEAT, bool PRIMARY>
__global__ void __launch_bounds__(RTC_BLOCK) wavefront_level(const Scene sc, const FrameParams f, const LevelParams lv) {
    __shared__ uint8_t s_perm[512];
    for (int i = threadIdx.x; i < 512; i += blockDim.x) s_perm[i] = c_perm[i];
    __syncthreads();

    const bool brute = (f.flags & 1u) != 0;
    const int L = sc.n_lights;
    const uint64_t n_in = PRIMARY ? lv.n_primary : (uint64_t)(*lv.in_count < lv.in_capacity ? *lv.in_count : lv.in_capacity);
    uint64_t c_shadow = 0, c_refl = 0, c_refr = 0, c_ref = 0, c_in = 0;
    const unsigned lane = threadIdx.x & 31u;

    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    // every lane of a warp runs the same number of iterations so the warp-aggregated queue push stays convergent
    const uint64_t n_round = (n_in + 31u) & ~(uint64_t)31u;
    for (uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n_round; idx += stride) {
        bool active = idx < n_in;
        Ray ray;
        double w = 1.0;
        uint32_t sample = 0;
        int fuel = f.fuel;
        if (active) {
            if (PRIMARY) {
                const uint64_t s = lv.first_sample + idx;
                int px, py;
                tile_pixel(f, s / (uint64_t)f.spp, px, py);
                if (px >= f.win_w || py >= f.win_h) active = false;  // padding of a partial tile
                else ray = camera_ray(f, px, py, (int)(s % (uint64_t)f.spp));
                sample = (uint32_t)idx;
            } else {
                const double2 a = lv.in_q0[idx], b = lv.in_q1[idx], c = lv.in_q2[idx], e = lv.in_q3[idx];
                ray.o = mk(a.x, a.y, b.x);
                ray.d = mk(b.y, c.x, c.y);
                w = e.x;
                const uint64_t bits = (uint64_t)__double_as_longlong(e.y);
                sample = (uint32_t)(bits & 0xffffffffu);
                fuel = (int)(bits >> 32);
            }
        }
        bool push_refl = false, push_refr = false;
        Ray r_refl, r_refr;
        double w_refl = 0.0, w_refr = 0.0;
        if (active) {
            c_in++;
            // L^depth copies of this ray exist in the reference (one per light at every bounce above it)
            uint64_t mult = 1;
            for (int k = fuel; k < f.fuel; ++k) mult *= (uint64_t)L;
            c_ref += mult;
            const Hit hit = closest_hit<FEAT>(sc, ray, CUDART_INF, brute, lv.overflow);
            if (hit.prim >= 0) {
                c_ref += mult * (uint64_t)L;
                const Prim prim = sc.prims[hit.prim];
                const Material& mat = sc.materials[prim.material];
                // ---- Intersection::prepare_state, src/intersection.rs:50-68 ----
                const V3 point = ray.o + ray.d * hit.t;
                const V3 eye = -ray.d;
                V3 normal = prim_normal(sc, prim, point, hit.u, hit.v);
                if (dot(normal, eye) < 0.0) normal = -normal;
                const V3 over_point = point + (normal * RTC_EPSILON);
                const double reflective = mat.reflective, transparency = mat.transparency;
                double n1 = 1.0, n2 = 1.0, reflectance = 0.0;
                if ((FEAT & FEAT_TRANSPARENT) && transparency != 0.0) {
                    refractive_indices<FEAT>(sc, ray, hit, brute, n1, n2, lv.overflow);
                    reflectance = schlick(eye, normal, n1, n2);
                }
                // ---- World::shade_hit light loop, src/world.rs:58-64 + Shape::lighting, src/shape.rs:429-463 ----
                Rgb color;
                if (mat.flags & 1) color = rgb(mat.plain[0], mat.plain[1], mat.plain[2]);
                else {
                    V3 pp = over_point;
                    double pw = 1.0;
                    xf_point4(sc.matinvs[prim.matinv].m, pp, pw);
                    color = pattern_color(sc, s_perm, mat.pattern, pp, pw);
                }
                Rgb sum = rgb(0.0, 0.0, 0.0);
                for (int li = 0; li < L; ++li) {
                    const Light& light = sc.lights[li];
                    const V3 lpos = mk(light.origin[0], light.origin[1], light.origin[2]);
                    const Rgb intensity = rgb(light.intensity[0], light.intensity[1], light.intensity[2]);
                    const V3 to_light = lpos - over_point;
                    const double distance = magnitude(to_light);
                    const V3 lightv = normalize(to_light);
                    Ray sray;
                    sray.o = over_point;
                    sray.d = lightv;
                    c_shadow++;
                    const bool shadowed = is_shadowed<FEAT>(sc, sray, distance, brute, lv.overflow);
                    const Rgb effective = color * intensity;
                    Rgb surface = effective * mat.ambient;
                    const double light_dot_normal = dot(lightv, normal);
                    if (!shadowed && light_dot_normal >= 0.0) {
                        const Rgb diffuse = effective * mat.diffuse * light_dot_normal;
                        Rgb specular = rgb(0.0, 0.0, 0.0);
                        const double reflect_dot_eye = dot(reflect(-lightv, normal), eye);
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
                        r_refl.o = over_point;
                        r_refl.d = reflect(ray.d, normal);
                        w_refl = w * (double)L * reflective * (both ? reflectance : 1.0);
                    }
                    if ((FEAT & FEAT_TRANSPARENT) && transparency != 0.0) {
                        const double n_ratio = n1 / n2;
                        const double cos_i = dot(eye, normal);
                        const double sin2_t = (n_ratio * n_ratio) * (1.0 - cos_i * cos_i);
                        if (!(sin2_t > 1.0)) {
                            const double cos_t = sqrt(1.0 - sin2_t);
                            push_refr = true;
                            r_refr.o = point - (normal * RTC_EPSILON);
                            r_refr.d = normal * (n_ratio * cos_i - cos_t) - eye * n_ratio;
                            w_refr = w * (double)L * transparency * (both ? (1.0 - reflectance) : 1.0);
                        }
                    }
                }
            }
        }
        // ---- warp-aggregated compaction into the next level's queue: one atomic per warp ----
        const unsigned m_refl = __ballot_sync(0xffffffffu, push_refl);
        const unsigned m_refr = __ballot_sync(0xffffffffu, push_refr);
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
            if (push_refr) {
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
    // ---- ray statistics: warp reduce, one atomic per counter per warp ----
    uint64_t vals[5] = {c_in, c_shadow, c_refl, c_refr, c_ref};
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        uint64_t v = vals[k];
        for (int off = 16; off; off >>= 1) v += __shfl_down_sync(0xffffffffu, v, off);
        if (lane == 0 && v) atomicAdd(lv.stats + k, (unsigned long long)v);
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
        tile_pixel(f, ordinal, px, py);
        if (px >= f.win_w || py >= f.win_h) return;
        dst = (uint64_t)py * (uint64_t)f.win_w + (uint64_t)px;
    }
    out[3 * dst] = r;
    out[3 * dst + 1] = g;
    out[3 * dst + 2] = b;
}

void upload_perm(const uint8_t* perm512) { cudaMemcpyToSymbol(c_perm, perm512, 512); }

template <int FEAT>
static void launch_feat(bool primary, int grid, cudaStream_t st, const Scene& sc, const FrameParams& f, const LevelParams& lv) {
    if (primary) wavefront_level<FEAT, true><<<grid, RTC_BLOCK, 0, st>>>(sc, f, lv);
    else wavefront_level<FEAT, false><<<grid, RTC_BLOCK, 0, st>>>(sc, f, lv);
}

void launch_level(int feat, bool primary, int grid, cudaStream_t st, const Scene& sc, const FrameParams& f, const LevelParams& lv) {
    if (feat & (FEAT_CSG | FEAT_GATES)) launch_feat<FEAT_TRANSPARENT | FEAT_CSG | FEAT_GATES>(primary, grid, st, sc, f, lv);
    else if (feat & FEAT_TRANSPARENT) launch_feat<FEAT_TRANSPARENT>(primary, grid, st, sc, f, lv);
    else launch_feat<0>(primary, grid, st, sc, f, lv);
}

void launch_resolve(cudaStream_t st, const FrameParams& f, const double* accum, uint64_t first_pixel, uint64_t n_pixels, double* out, int row_major) {
    if (!n_pixels) return;
    const int block = 256;
    const uint64_t grid = (n_pixels + block - 1) / block;
    resolve_pixels<<<(unsigned)grid, block, 0, st>>>(f, accum, first_pixel, n_pixels, out, row_major);
}

}  // namespace rtcdev
