// Device maths of the render path: f64 vectors, affine transforms and the per-primitive intersectors
// and normals.  Every parity-critical expression keeps the reference's operation order and the
// translation unit is compiled with -fmad=false, so no multiply-add is contracted and t, u, v, hit
// points and normals are bit-identical to the reference's scalar f64 code (division and sqrt are IEEE
// correctly rounded on the device).
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

#include "device_scene.h"

namespace rtcdev {

#define RTC_EPSILON 0.00001  // src/config.rs:1

struct V3 {
    double x, y, z;
};
__device__ __forceinline__ V3 mk(double x, double y, double z) { V3 v; v.x = x; v.y = y; v.z = z; return v; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator-(V3 a) { return mk(-a.x, -a.y, -a.z); }
__device__ __forceinline__ V3 operator*(V3 a, double s) { return mk(a.x * s, a.y * s, a.z * s); }
// src/linalg/vector.rs:39-49
__device__ __forceinline__ double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 cross(V3 a, V3 b) { return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
// src/linalg/vector.rs:24-37
__device__ __forceinline__ double magnitude(V3 a) { return sqrt(a.x * a.x + a.y * a.y + a.z * a.z); }
__device__ __forceinline__ V3 normalize(V3 a) {
    const double m = magnitude(a);
    return mk(a.x / m, a.y / m, a.z / m);
}
// src/linalg/vector.rs:75-77
__device__ __forceinline__ V3 reflect(V3 v, V3 n) { return v - n * (2.0 * dot(v, n)); }

// src/linalg/matrix.rs:261-284 restricted to affine matrices: for a point the w-term is m[r][3] * 1.0, for a
// direction it is m[r][3] * 0.0 (adds a zero), so three products and two / three sums reproduce the 4x4 result.
__device__ __forceinline__ V3 xf_point(const double* __restrict__ m, V3 p) {
    return mk(m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3], m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7],
              m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11]);
}
__device__ __forceinline__ V3 xf_dir(const double* __restrict__ m, V3 d) {
    return mk(m[0] * d.x + m[1] * d.y + m[2] * d.z, m[4] * d.x + m[5] * d.y + m[6] * d.z, m[8] * d.x + m[9] * d.y + m[10] * d.z);
}
__device__ __forceinline__ V3 xf_normal(const double* __restrict__ t, V3 n) {  // 3x3, row-major
    return mk(t[0] * n.x + t[1] * n.y + t[2] * n.z, t[3] * n.x + t[4] * n.y + t[5] * n.z, t[6] * n.x + t[7] * n.y + t[8] * n.z);
}

struct Ray {
    V3 o, d;
};

// src/approx.rs:19-23 against 0.0
__device__ __forceinline__ bool approx0(double a) { return fabs(a) < RTC_EPSILON; }

// src/shape.rs:635-653.  fmin/fmax drop a NaN operand exactly like f64::min/max.
__device__ __forceinline__ void cube_axis(double origin, double direction, double lo, double hi, double& t0, double& t1) {
    const double n0 = lo - origin, n1 = hi - origin;
    double a, b;
    if (fabs(direction) >= RTC_EPSILON) {
        a = n0 / direction;
        b = n1 / direction;
    } else {
        a = n0 * CUDART_INF;
        b = n1 * CUDART_INF;
    }
    if (a > b) { t0 = b; t1 = a; } else { t0 = a; t1 = b; }
}
// src/bounding_box.rs:80-92 (no t-range clipping)
__device__ __forceinline__ bool box_gate(const Gate& g, const Ray& r) {
    double x0, x1, y0, y1, z0, z1;
    cube_axis(r.o.x, r.d.x, g.lo[0], g.hi[0], x0, x1);
    cube_axis(r.o.y, r.d.y, g.lo[1], g.hi[1], y0, y1);
    cube_axis(r.o.z, r.d.z, g.lo[2], g.hi[2], z0, z1);
    return fmax(fmax(x0, y0), z0) <= fmin(fmin(x1, y1), z1);
}

// Crossings of one primitive with the object-space ray, in the reference's emission order.
struct Roots {
    double t[4];
    int n;
    double u, v;  // triangles only
};

// src/shape.rs:592-620
__device__ __forceinline__ void isect_sphere(const Ray& r, Roots& out) {
    const double a = dot(r.d, r.d);
    const double b = 2.0 * dot(r.d, r.o);
    const double c = dot(r.o, r.o) - 1.0;
    const double disc = b * b - 4.0 * a * c;
    if (disc < 0.0) return;
    const double sq = sqrt(disc);
    out.t[0] = (-b - sq) / (2.0 * a);
    out.t[1] = (-b + sq) / (2.0 * a);
    out.n = 2;
}
// The same for a query that only wants crossings with t >= 0 and only the nearest of them (closest hit, shadow):
// 2a > 0 and sq >= 0, so t0 <= t1 and the sign of t0 is the sign of its numerator.  If t0 >= 0 it is the answer and
// t1 cannot beat it; if t0 < 0 only t1 can count.  One IEEE division instead of two; the quotient kept is the
// reference's own.  (A numerator so small that the quotient could round to -0.0, which `t >= 0.0` accepts, and a
// degenerate a take the two-root path.)
__device__ __forceinline__ void isect_sphere_nearest(const Ray& r, Roots& out) {
    const double a = dot(r.d, r.d);
    const double b = 2.0 * dot(r.d, r.o);
    const double c = dot(r.o, r.o) - 1.0;
    const double disc = b * b - 4.0 * a * c;
    if (disc < 0.0) return;
    const double sq = sqrt(disc);
    const double n0 = -b - sq;
    if (!(a > 0.0) || (n0 < 0.0 && n0 > -1e-280)) {
        out.t[0] = n0 / (2.0 * a);
        out.t[1] = (-b + sq) / (2.0 * a);
        out.n = 2;
        return;
    }
    out.t[0] = (n0 >= 0.0 ? n0 : -b + sq) / (2.0 * a);
    out.n = 1;
}
// src/shape.rs:622-633
__device__ __forceinline__ void isect_plane(const Ray& r, Roots& out) {
    if (approx0(r.d.y)) return;
    out.t[0] = -r.o.y / r.d.y;
    out.n = 1;
}
// src/shape.rs:655-680
__device__ __forceinline__ void isect_cube(const Ray& r, Roots& out) {
    double x0, x1, y0, y1, z0, z1;
    cube_axis(r.o.x, r.d.x, -1.0, 1.0, x0, x1);
    cube_axis(r.o.y, r.d.y, -1.0, 1.0, y0, y1);
    cube_axis(r.o.z, r.d.z, -1.0, 1.0, z0, z1);
    const double tmin = fmax(fmax(x0, y0), z0);
    const double tmax = fmin(fmin(x1, y1), z1);
    if (tmin <= tmax) {
        out.t[0] = tmin;
        out.t[1] = tmax;
        out.n = 2;
    }
}
// src/shape.rs:682-722
__device__ __forceinline__ void isect_caps(const Ray& r, double lo, double hi, double lo_radius, double hi_radius, bool closed, Roots& out) {
    if (!closed || approx0(r.d.y)) return;
    double t = (lo - r.o.y) / r.d.y;
    double x = r.o.x + t * r.d.x, z = r.o.z + t * r.d.z;
    if ((x * x + z * z) <= lo_radius * lo_radius) out.t[out.n++] = t;
    t = (hi - r.o.y) / r.d.y;
    x = r.o.x + t * r.d.x;
    z = r.o.z + t * r.d.z;
    if ((x * x + z * z) <= hi_radius * hi_radius) out.t[out.n++] = t;
}
// src/shape.rs:724-768
__device__ __forceinline__ void isect_cylinder(const Ray& r, double lo, double hi, bool closed, Roots& out) {
    const double a = r.d.x * r.d.x + r.d.z * r.d.z;
    if (!approx0(a)) {
        const double b = 2.0 * r.o.x * r.d.x + 2.0 * r.o.z * r.d.z;
        const double c = r.o.x * r.o.x + r.o.z * r.o.z - 1.0;
        const double disc = b * b - 4.0 * a * c;
        if (disc >= 0.0) {
            const double sq = sqrt(disc);
            const double t0 = (-b - sq) / (2.0 * a);
            const double y0 = r.o.y + t0 * r.d.y;
            if (lo < y0 && y0 < hi) out.t[out.n++] = t0;
            const double t1 = (-b + sq) / (2.0 * a);
            const double y1 = r.o.y + t1 * r.d.y;
            if (lo < y1 && y1 < hi) out.t[out.n++] = t1;
        }
    }
    isect_caps(r, lo, hi, 1.0, 1.0, closed, out);
}
// src/shape.rs:770-822 (the single-root branch has no y-range check)
__device__ __forceinline__ void isect_cone(const Ray& r, double lo, double hi, bool closed, Roots& out) {
    const double a = r.d.x * r.d.x - r.d.y * r.d.y + r.d.z * r.d.z;
    const double b = 2.0 * r.o.x * r.d.x - 2.0 * r.o.y * r.d.y + 2.0 * r.o.z * r.d.z;
    if (!approx0(a) || !approx0(b)) {
        const double c = r.o.x * r.o.x - r.o.y * r.o.y + r.o.z * r.o.z;
        if (!approx0(a)) {
            const double disc = b * b - 4.0 * a * c;
            if (disc >= 0.0) {
                const double sq = sqrt(disc);
                const double t0 = (-b - sq) / (2.0 * a);
                const double y0 = r.o.y + t0 * r.d.y;
                if (lo < y0 && y0 < hi) out.t[out.n++] = t0;
                const double t1 = (-b + sq) / (2.0 * a);
                const double y1 = r.o.y + t1 * r.d.y;
                if (lo < y1 && y1 < hi) out.t[out.n++] = t1;
            }
        } else {
            out.t[out.n++] = -c / (2.0 * b);
        }
    }
    isect_caps(r, lo, hi, lo, hi, closed, out);
}
// src/shape.rs:824-860 (Moller-Trumbore, no test on t)
__device__ __forceinline__ void isect_triangle(const Ray& r, V3 p1, V3 e1, V3 e2, Roots& out) {
    const V3 dir_cross_e2 = cross(r.d, e2);
    const double det = dot(e1, dir_cross_e2);
    if (fabs(det) < RTC_EPSILON) return;
    const double f = 1.0 / det;
    const V3 p1_to_origin = r.o - p1;
    const double u = f * dot(p1_to_origin, dir_cross_e2);
    if (u < 0.0 || u > 1.0) return;
    const V3 origin_cross_e1 = cross(p1_to_origin, e1);
    const double v = f * dot(r.d, origin_cross_e1);
    if (v < 0.0 || u + v > 1.0) return;
    out.t[0] = f * dot(e2, origin_cross_e1);
    out.u = u;
    out.v = v;
    out.n = 1;
}

__device__ __forceinline__ void load_tri(const double* __restrict__ rec, V3& a, V3& b, V3& c) {
    // 80-byte record, 16-byte aligned: five 128-bit loads
    const double2* q = reinterpret_cast<const double2*>(rec);
    const double2 q0 = __ldg(q), q1 = __ldg(q + 1), q2 = __ldg(q + 2), q3 = __ldg(q + 3), q4 = __ldg(q + 4);
    a = mk(q0.x, q0.y, q1.x);
    b = mk(q1.y, q2.x, q2.y);
    c = mk(q3.x, q3.y, q4.x);
}

// World ray -> object space (Ray::transform, src/ray.rs:14-19), cached per transform id: consecutive items that
// share a transform (all triangles of a mesh do) reuse the transformed ray, bit-identical to recomputing it.
struct ObjRay {
    int xform;
    Ray r;
};
// Work actually done by a thread's queries (feeds the algorithmic FLOP / byte model of the roofline):
// BVH node visits (two slab tests each), triangle tests, other primitive tests, ray transforms.
struct Tally {
    uint32_t nodes, tris, prims, xforms;
};
__device__ __forceinline__ void load_xform_inv(const Scene& sc, int xform, double* __restrict__ m) {
    const double2* q = reinterpret_cast<const double2*>(sc.xforms[xform].inv);  // 192-byte records, 16-byte aligned
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        const double2 v = __ldg(q + k);
        m[2 * k] = v.x;
        m[2 * k + 1] = v.y;
    }
}
__device__ __forceinline__ const Ray& object_ray(const Scene& sc, const Ray& world, int xform, ObjRay& c, Tally& tally) {
    if (xform != c.xform) {
        tally.xforms++;
        double m[12];
        load_xform_inv(sc, xform, m);
        c.r.o = xf_point(m, world.o);
        c.r.d = xf_dir(m, world.d);
        c.xform = xform;
    }
    return c.r;
}

// All crossings of one primitive with the object-space ray r (src/shape.rs:862-885).  Out of line on purpose: the
// seven intersectors with their f64 divisions and square roots are ~2 000 instructions, and every kernel that can
// meet an arbitrary primitive shares this one copy (the hot kinds are inlined at their call sites instead).
// tri_rec: the primitive's (p1, e1, e2) record if it is a triangle.
//
// The out-of-line body only sees COPIES (roots_generic below): a pointer into the caller's query state handed to a
// real call would pin that state — the cached object ray, the tallies, the whole visitor and through it the scene
// descriptor — in local memory for the entire kernel (r01d: 4x more local than global requests in the walk).
struct GenericCall {
    Ray r;
    Roots out;
};
__device__ __forceinline__ void roots_by_kind(const double* __restrict__ ranges, int kind_flags, int geom, const double* __restrict__ tri_rec, const Ray& r, Roots& out) {
    out.n = 0;
    out.u = 0.0;
    out.v = 0.0;
    switch (kind_flags & 0xff) {
        case 0: isect_sphere(r, out); break;
        case 1: isect_plane(r, out); break;
        case 2: isect_cube(r, out); break;
        case 3: isect_cylinder(r, __ldg(ranges + 2 * geom), __ldg(ranges + 2 * geom + 1), (kind_flags & 0x200) != 0, out); break;
        case 4: isect_cone(r, __ldg(ranges + 2 * geom), __ldg(ranges + 2 * geom + 1), (kind_flags & 0x200) != 0, out); break;
        default: {
            V3 p1, e1, e2;
            load_tri(tri_rec, p1, e1, e2);
            isect_triangle(r, p1, e1, e2, out);
        }
    }
}
__device__ __noinline__ void roots_generic_body(const double* __restrict__ ranges, int kind_flags, int geom, const double* __restrict__ tri_rec, GenericCall* c) {
    const Ray r = c->r;
    Roots out;
    roots_by_kind(ranges, kind_flags, geom, tri_rec, r, out);
    c->out = out;
}
__device__ __forceinline__ void roots_generic(const Scene& sc, int kind_flags, int geom, const double* __restrict__ tri_rec, const Ray& r, Roots& out) {
    GenericCall c;
    c.r = r;
    roots_generic_body(sc.ranges, kind_flags, geom, tri_rec, &c);
    out.n = c.out.n;
#pragma unroll
    for (int k = 0; k < 4; ++k) out.t[k] = c.out.t[k];
    out.u = c.out.u;
    out.v = c.out.v;
}

// A plane only reads the y components of the object-space ray (src/shape.rs:622-633), i.e. row 1 of the transform:
// two dot products instead of the full Ray::transform, with the same expression order as xf_point / xf_dir, so o.y
// and d.y are bit-identical.  skip_negative: the caller only wants crossings with t >= 0 (closest hit, shadow), so
// the division is skipped when the signs of o.y and d.y already say t < 0.
__device__ __forceinline__ void plane_roots_row(double r0, double r1, double r2, double r3, const Ray& world, bool skip_negative, Roots& out) {
    const double oy = r0 * world.o.x + r1 * world.o.y + r2 * world.o.z + r3;
    const double dy = r0 * world.d.x + r1 * world.d.y + r2 * world.d.z;
    if (approx0(dy)) return;
    if (skip_negative && ((oy > 0.0 && dy > 0.0) || (oy < 0.0 && dy < 0.0))) return;  // -oy / dy < 0
    out.t[0] = -oy / dy;
    out.n = 1;
}
// The same crossing left as a fraction, t = -oy / dy, for queries that can often decide without the division (below).
// Returns false when there is no crossing to consider.
__device__ __forceinline__ bool plane_fraction(double r0, double r1, double r2, double r3, const Ray& world, bool skip_negative, double& oy, double& dy) {
    oy = r0 * world.o.x + r1 * world.o.y + r2 * world.o.z + r3;
    dy = r0 * world.d.x + r1 * world.d.y + r2 * world.d.z;
    if (approx0(dy)) return false;
    if (skip_negative && ((oy > 0.0 && dy > 0.0) || (oy < 0.0 && dy < 0.0))) return false;
    return true;
}
// Is t = -oy / dy (as the reference rounds it) certainly above `bound`?  |oy| > |bound * dy| with a margin of 1e-14
// (the product and the quotient are each rounded once, 1.1e-16) means the exact quotient exceeds bound by more than any
// rounding can undo.  Only meaningful for crossings with t >= 0 and bound >= 0; NaN operands answer false.
__device__ __forceinline__ bool plane_certainly_above(double oy, double dy, double bound) { return fabs(oy) > fabs(bound * dy) * (1.0 + 1e-14); }
__device__ __forceinline__ bool plane_certainly_below(double oy, double dy, double bound) { return fabs(oy) < fabs(bound * dy) * (1.0 - 1e-14); }
__device__ __forceinline__ void plane_roots(const Scene& sc, int xform, const Ray& world, bool skip_negative, Roots& out) {
    const double2* q = reinterpret_cast<const double2*>(sc.xforms[xform].inv + 4);
    const double2 a = __ldg(q), b = __ldg(q + 1);
    plane_roots_row(a.x, a.y, b.x, b.y, world, skip_negative, out);
}

// HOT selects which kinds are intersected inline at this call site (bit k = rtc_prim_kind k; triangles = bits 5|6);
// bit 8: every primitive that reaches this call site is a triangle.
template <int HOT>
__device__ __forceinline__ void prim_roots(const Scene& sc, int kind_flags, int xform, int geom, const double* __restrict__ tri_rec,
                                           const Ray& world, ObjRay& cache, Tally& tally, bool skip_negative, Roots& out) {
    out.n = 0;
    const int kind = kind_flags & 0xff;
    if ((HOT & 0x02) && kind == 1) {
        tally.prims++;
        plane_roots(sc, xform, world, skip_negative, out);
        return;
    }
    const Ray& r = object_ray(sc, world, xform, cache, tally);
    if (HOT & 0x100) {  // the caller knows the primitive is a triangle (FEAT_TRIS): no dispatch, no out-of-line fallback in this kernel
        tally.tris++;
        V3 p1, e1, e2;
        load_tri(tri_rec, p1, e1, e2);
        isect_triangle(r, p1, e1, e2, out);
        return;
    }
    if (kind >= 5) tally.tris++; else tally.prims++;
    if ((HOT & 0x60) && kind >= 5) {
        V3 p1, e1, e2;
        load_tri(tri_rec, p1, e1, e2);
        isect_triangle(r, p1, e1, e2, out);
    } else if ((HOT & 0x01) && kind == 0) {
        if (skip_negative) isect_sphere_nearest(r, out);
        else isect_sphere(r, out);
    } else if (HOT & 0x400) {  // the caller is itself an out-of-line routine (a CSG program): the dispatch is inlined into it
        roots_by_kind(sc.ranges, kind_flags, geom, tri_rec, r, out);
    } else {
        roots_generic(sc, kind_flags, geom, tri_rec, r, out);
    }
}
template <int HOT>
__device__ __forceinline__ void prim_roots(const Scene& sc, const Prim& p, const Ray& world, ObjRay& cache, Tally& tally, bool skip_negative, Roots& out) {
    prim_roots<HOT>(sc, p.kind_flags, p.xform, p.geom, sc.tri_pe + 10 * (size_t)(p.geom < 0 ? 0 : p.geom), world, cache, tally, skip_negative, out);
}

// src/shape.rs:419-427 + :887-946 — world normal at a world point (not yet flipped towards the eye)
__device__ __forceinline__ V3 prim_normal(const Scene& sc, const Prim& p, V3 world_point, double u, double v) {
    const Xform& xf = sc.xforms[p.xform];
    const V3 q = xf_point(xf.inv, world_point);
    const int kind = p.kind_flags & 0xff;
    V3 n;
    switch (kind) {
        case 0: n = q; break;
        case 1: n = mk(0.0, 1.0, 0.0); break;
        case 2: {
            const double ax = fabs(q.x), ay = fabs(q.y), az = fabs(q.z);
            const double mx = fmax(fmax(ax, ay), az);
            if (mx == ax) n = mk(q.x, 0.0, 0.0);
            else if (mx == ay) n = mk(0.0, q.y, 0.0);
            else n = mk(0.0, 0.0, q.z);
            break;
        }
        case 3:
        case 4: {
            const double lo = sc.ranges[2 * p.geom], hi = sc.ranges[2 * p.geom + 1];
            const double dist = q.x * q.x + q.z * q.z;
            if (dist < 1.0 && q.y >= hi - RTC_EPSILON) n = mk(0.0, 1.0, 0.0);
            else if (dist < 1.0 && q.y <= lo + RTC_EPSILON) n = mk(0.0, -1.0, 0.0);
            else if (kind == 3) n = mk(q.x, 0.0, q.z);
            else {
                double y = sqrt(dist);
                if (q.y > 0.0) y = -y;
                n = mk(q.x, y, q.z);
            }
            break;
        }
        case 5: {
            const double* rec = sc.tri_n + 10 * (size_t)p.geom;
            n = mk(rec[0], rec[1], rec[2]);
            break;
        }
        default: {
            V3 n1, n2, n3;
            load_tri(sc.tri_n + 10 * (size_t)p.geom, n1, n2, n3);
            n = n2 * u + n3 * v + n1 * (1.0 - u - v);  // src/shape.rs:943
        }
    }
    return normalize(xf_normal(xf.tsp, n));
}

}  // namespace rtcdev
