// Procedural colour: src/material.rs:164-301 (Pattern / Mixture / Jitter) and src/noise.rs (3-D simplex,
// fractal sum, jitter_3d) as device functions.  The recursive pattern tree is evaluated iteratively: kinds
// that merely select a child (checkers, ring, stripes, point jitter) are tail calls; kinds that combine child
// colours (blend, gradient, ring-gradient, colour jitter) push a frame.  The permutation table lives in
// __constant__ memory and is staged into shared memory per block, because the nine dependent lookups of one
// simplex evaluation are lane-divergent (constant-cache reads of different addresses serialise).
#pragma once
#include "device_math.cuh"

namespace rtcdev {

__constant__ uint8_t c_perm[512];  // src/noise.rs:56-84

struct Rgb {
    double r, g, b;
};
__device__ __forceinline__ Rgb rgb(double r, double g, double b) { Rgb c; c.r = r; c.g = g; c.b = b; return c; }
__device__ __forceinline__ Rgb operator+(Rgb a, Rgb b) { return rgb(a.r + b.r, a.g + b.g, a.b + b.b); }
__device__ __forceinline__ Rgb operator-(Rgb a, Rgb b) { return rgb(a.r - b.r, a.g - b.g, a.b - b.b); }
__device__ __forceinline__ Rgb operator*(Rgb a, double s) { return rgb(a.r * s, a.g * s, a.b * s); }
__device__ __forceinline__ Rgb operator*(Rgb a, Rgb b) { return rgb(a.r * b.r, a.g * b.g, a.b * b.b); }

// Rust `f64 as i32` saturates and maps NaN to 0; cvt.rzi.s32.f64 does the same.
__device__ __forceinline__ int f2i_sat(double x) { return __double2int_rz(x); }

// src/noise.rs:94-114
__device__ __forceinline__ double noise_grad(int h, double x, double y, double z) {
    switch (h & 0xF) {
        case 0x0: return x + y;
        case 0x1: return -x + y;
        case 0x2: return x - y;
        case 0x3: return -x - y;
        case 0x4: return x + z;
        case 0x5: return -x + z;
        case 0x6: return x - z;
        case 0x7: return -x - z;
        case 0x8: return y + z;
        case 0x9: return -y + z;
        case 0xA: return y - z;
        case 0xB: return -y - z;
        case 0xC: return y + x;
        case 0xD: return -y + z;
        case 0xE: return y - x;
        default: return -y - z;
    }
}
// src/noise.rs:116-122 (off by one for non-positive integers; i32 arithmetic wraps)
__device__ __forceinline__ int noise_floor(double x) { return x > 0.0 ? f2i_sat(x) : (int)((unsigned)f2i_sat(x) - 1u); }
// src/noise.rs:124-131 with m = 256
__device__ __forceinline__ int noise_mod256(int x) {
    const int a = x % 256;
    return a < 0 ? a + 256 : a;
}
__device__ __forceinline__ double noise_corner(double x, double y, double z, int gi) {
    double t = 0.6 - x * x - y * y - z * z;
    if (t < 0.0) return 0.0;
    t *= t;
    return t * t * noise_grad(gi, x, y, z);
}
// src/noise.rs:134-219
__device__ __noinline__ double simplex(const uint8_t* __restrict__ P, double x, double y, double z) {
    const double F3 = 1.0 / 3.0, G3 = 1.0 / 6.0;
    const double s = (x + y + z) * F3;
    const int i = noise_floor(x + s), j = noise_floor(y + s), k = noise_floor(z + s);
    const double t = (double)(int)((unsigned)i + (unsigned)j + (unsigned)k) * G3;
    const double x0 = x - ((double)i - t);
    const double y0 = y - ((double)j - t);
    const double z0 = z - ((double)k - t);
    int i1, j1, k1, i2, j2, k2;
    if (x0 >= y0) {
        if (y0 >= z0) { i1 = 1; j1 = 0; k1 = 0; i2 = 1; j2 = 1; k2 = 0; }
        else if (x0 >= z0) { i1 = 1; j1 = 0; k1 = 0; i2 = 1; j2 = 0; k2 = 1; }
        else { i1 = 0; j1 = 0; k1 = 1; i2 = 1; j2 = 0; k2 = 1; }
    } else {
        if (y0 < z0) { i1 = 0; j1 = 0; k1 = 1; i2 = 0; j2 = 1; k2 = 1; }
        else if (x0 < z0) { i1 = 0; j1 = 1; k1 = 0; i2 = 0; j2 = 1; k2 = 1; }
        else { i1 = 0; j1 = 1; k1 = 0; i2 = 1; j2 = 1; k2 = 0; }
    }
    const double x1 = x0 - (double)i1 + G3, y1 = y0 - (double)j1 + G3, z1 = z0 - (double)k1 + G3;
    const double x2 = x0 - (double)i2 + 2.0 * G3, y2 = y0 - (double)j2 + 2.0 * G3, z2 = z0 - (double)k2 + 2.0 * G3;
    const double x3 = x0 - 1.0 + 3.0 * G3, y3 = y0 - 1.0 + 3.0 * G3, z3 = z0 - 1.0 + 3.0 * G3;
    const int ii = noise_mod256(i), jj = noise_mod256(j), kk = noise_mod256(k);
    const int gi0 = P[ii + P[jj + P[kk]]];
    const int gi1 = P[ii + i1 + P[jj + j1 + P[kk + k1]]];
    const int gi2 = P[ii + i2 + P[jj + j2 + P[kk + k2]]];
    const int gi3 = P[ii + 1 + P[jj + 1 + P[kk + 1]]];
    const double n0 = noise_corner(x0, y0, z0, gi0);
    const double n1 = noise_corner(x1, y1, z1, gi1);
    const double n2 = noise_corner(x2, y2, z2, gi2);
    const double n3 = noise_corner(x3, y3, z3, gi3);
    return 32.0 * (n0 + n1 + n2 + n3);
}
// src/noise.rs:221-237 (lacunarity 2, persistence 0.5; 0 octaves -> 0/0 = NaN)
__device__ __forceinline__ double fractal(const uint8_t* __restrict__ P, double x, double y, double z, int octaves) {
    double output = 0.0, denom = 0.0, frequency = 1.0, amplitude = 1.0;
    for (int o = 0; o < octaves; ++o) {
        output += amplitude * simplex(P, x * frequency, y * frequency, z * frequency);
        denom += amplitude;
        frequency *= 2.0;
        amplitude *= 0.5;
    }
    return output / denom;
}
// src/noise.rs:31-52
__device__ __forceinline__ void jitter_3d(const uint8_t* __restrict__ P, const Pattern& pt, double& x, double& y, double& z) {
    double nx, ny, nz;
    if (pt.noise_kind == 0) {
        nx = simplex(P, x, y, z) * pt.noise_scale;
        ny = simplex(P, x, y, z + 1.0) * pt.noise_scale;
        nz = simplex(P, x, y, z + 2.0) * pt.noise_scale;
    } else {
        nx = fractal(P, x, y, z, pt.octaves) * pt.noise_scale;
        ny = fractal(P, x, y, z + 1.0, pt.octaves) * pt.noise_scale;
        nz = fractal(P, x, y, z + 2.0, pt.octaves) * pt.noise_scale;
    }
    x = x + nx;
    y = y + ny;
    z = z + nz;
}

struct PatFrame {
    int node, stage;  // stage 0: left pending, 1: right pending, 2: colour-jitter post-op
    V3 p;
    double w;
    Rgb left;
};

// Full 4x4 product incl. the w row (src/linalg/matrix.rs:261-284).  Pattern space is the one place where matrix
// products chain (material_inv, then each nested Mixture's transform_inv), and the reference's inverses carry a
// last row that is (0,0,0,1) only up to rounding — so w is carried exactly as the reference does.
__device__ __forceinline__ void xf_point4(const double* __restrict__ m, V3& p, double& w) {
    const double x = m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3] * w;
    const double y = m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7] * w;
    const double z = m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11] * w;
    const double ww = m[12] * p.x + m[13] * p.y + m[14] * p.z + m[15] * w;
    p = mk(x, y, z);
    w = ww;
}

// src/material.rs:164-301
// `p`, `w`: the pattern-space point material_inv * over_point (src/shape.rs:437)
__device__ __noinline__ Rgb pattern_color(const Scene& sc, const uint8_t* __restrict__ P, int root, V3 p, double w) {
    PatFrame stack[MAX_PATTERN_FRAMES];
    int sp = 0;
    int cur = root;
    Rgb result = rgb(0.0, 0.0, 0.0);
    for (;;) {
        // descend
        for (;;) {
            const Pattern& pt = sc.patterns[cur];
            const int kind = pt.kind;
            if (kind == 1) { result = rgb(pt.color[0], pt.color[1], pt.color[2]); break; }
            if (kind == 0) { result = rgb(p.x, p.y, p.z); break; }
            if (kind == 3) {  // point jitter: perturb the point, continue into the inner pattern
                jitter_3d(P, pt, p.x, p.y, p.z);
                w = 1.0;  // Vector::point(nx, ny, nz), src/material.rs:220
                cur = pt.left;
                continue;
            }
            if (kind == 2) {  // colour jitter: evaluate the inner pattern first
                stack[sp].node = cur; stack[sp].stage = 2; sp++;
                cur = pt.left;
                continue;
            }
            xf_point4(pt.m, p, w);  // Mixture: into its own space (src/material.rs:183)
            bool pick_left = true;
            if (kind == 5) {  // checkers :258-268 (i32 sum wraps; `% 2 == 0` keeps the sign)
                const int s = (int)((unsigned)f2i_sat(floor(p.x)) + (unsigned)f2i_sat(floor(p.y)) + (unsigned)f2i_sat(floor(p.z)));
                pick_left = (s % 2) == 0;
            } else if (kind == 7) {  // ring :278-284
                pick_left = (f2i_sat(floor(sqrt(p.x * p.x + p.z * p.z))) % 2) == 0;
            } else if (kind == 9) {  // stripes :293-299
                pick_left = (f2i_sat(floor(p.x)) % 2) == 0;
            } else {  // blend / ring gradient / gradient: both children
                stack[sp].node = cur; stack[sp].stage = 0; stack[sp].p = p; stack[sp].w = w; sp++;
                cur = pt.left;
                continue;
            }
            cur = pick_left ? pt.left : pt.right;
        }
        // ascend
        for (;;) {
            if (sp == 0) return result;
            PatFrame& f = stack[sp - 1];
            const Pattern& pt = sc.patterns[f.node];
            if (f.stage == 0) {
                f.left = result;
                f.stage = 1;
                cur = pt.right;
                p = f.p;
                w = f.w;
                break;  // descend into the right child
            }
            if (f.stage == 1) {
                const Rgb l = f.left, r = result;
                if (pt.kind == 4) {
                    result = (l + r) * 0.5;  // Color::avg, src/color.rs:48-50
                } else {
                    double fraction;
                    if (pt.kind == 6) {  // ring gradient uses the 3-D distance :270
                        const double dist = sqrt(f.p.x * f.p.x + f.p.y * f.p.y + f.p.z * f.p.z);
                        fraction = dist - floor(dist);
                    } else {
                        fraction = f.p.x - floor(f.p.x);
                    }
                    result = l + ((r - l) * fraction);
                }
            } else {  // colour jitter post-op, src/material.rs:209-217
                jitter_3d(P, pt, result.r, result.g, result.b);
            }
            sp--;
        }
    }
}

}  // namespace rtcdev
