// Ray queries against the flattened scene: closest hit, shadow test and the crossing list that feeds n1/n2.
// They replace World::intersect + Intersection::sort + Intersection::hit (src/world.rs:18-24,
// src/intersection.rs:123-132), which test every primitive and sort every crossing for every ray.
//
// Culling is done on a BVH2 whose boxes are fp32, outward-rounded and padded by 4e-6 * (scene extent): the
// computed slab interval of a padded box always contains the exact interval of the true box (DESIGN.md,
// "conservative fp32 culling"), so the set of primitives that reach the exact f64 intersectors is a superset
// of those the ray really crosses and the result equals the brute-force scan.  Accepted primitives are
// intersected in f64 with the reference's own arithmetic.  Ties in t resolve to the lower primitive index
// (= depth-first order = what the reference's stable sort yields).
//
// Each query is instantiated once per kernel (the kernels of render_kernels.cu hold one traversal each), which
// keeps the instruction footprint of a kernel small: profiles/README.md, "Reading of r01a".
#pragma once
#include "device_math.cuh"

// The loops that hand an intersector's roots (1 for a triangle or a plane, 2 for a sphere, up to 4 for a capped cylinder) to a
// visitor are inlined at every place a primitive is tested; unrolled four times each they were ~1 200 of the 6-13 K
// instructions of a trace kernel (tools/code_footprint.py).  Kept rolled: the footprint is what the kernels with the most
// code (CSG, gates) stall on (`no_instruction`), a rolled loop of <= 4 trips costs nothing measurable.
#ifndef RTC_ROOTS_UNROLL
#define RTC_ROOTS_UNROLL 1
#endif
#define RTC_PRAGMA_(x) _Pragma(#x)
#define RTC_PRAGMA(x) RTC_PRAGMA_(x)
#define RTC_ROOTS_LOOP RTC_PRAGMA(unroll RTC_ROOTS_UNROLL)

namespace rtcdev {

struct Hit {
    double t;
    int prim;
    double u, v;
};
struct Cross {
    double t;
    int prim;
    int pad;
};

__device__ __forceinline__ Prim load_prim(const Scene& sc, int id) {
    const int4* q = reinterpret_cast<const int4*>(sc.prims + id);
    const int4 a = __ldg(q), b = __ldg(q + 1);
    Prim p;
    p.kind_flags = a.x; p.material = a.y; p.xform = a.z; p.matinv = a.w;
    p.geom = b.x; p.pad1 = 0; p.gate_mask = (uint32_t)b.z; p.gate_mask_hi = (uint32_t)b.w;
    return p;
}
__device__ __forceinline__ LeafItem load_leaf(const Scene& sc, int slot) {
    const int4 a = __ldg(reinterpret_cast<const int4*>(sc.leaf_items + slot));
    LeafItem li;
    li.id = a.x; li.kind_flags = a.y; li.xform = a.z; li.geom = a.w;
    return li;
}

__device__ __forceinline__ bool allows_intersection(int kind, bool left_hit, bool in_left, bool in_right) {  // src/shape.rs:169-178
    if (kind == 0) return (left_hit && !in_right) || (!left_hit && !in_left);
    if (kind == 1) return (left_hit && in_right) || (!left_hit && in_left);
    return (left_hit && !in_right) || (!left_hit && in_left);
}

template <int FEAT>
__device__ __forceinline__ uint64_t gates_passed(const Scene& sc, const Ray& ray) {
    uint64_t m = 0;
    if (FEAT & FEAT_GATES)
        for (int g = 0; g < sc.n_gates; ++g)
            if (box_gate(sc.gates[g], ray)) m |= 1ull << g;
    return m;
}
__device__ __forceinline__ uint64_t prim_gates(const Prim& p) { return (uint64_t)p.gate_mask_hi << 32 | p.gate_mask; }

// Finite cones are BVH items under finite boxes; the one crossing such a box may not hold belongs to the cone's
// single-root branch, taken iff |a| < EPSILON with a = dx^2 - dy^2 + dz^2 of the OBJECT-space direction
// (src/shape.rs:774-818).  This evaluates exactly what the reference evaluates — the direction half of Ray::transform
// (src/ray.rs:14-19, same operation order as xf_dir) and then a — for every guarded cone transform; the result is
// bit-identical, so "no guard fires" <=> no cone of the scene takes that branch for this ray, and the BVH result
// equals the brute-force scan.  A ray that fires a guard (a band ~1e-6 rad wide around each cone's slope) simply
// tests every item.
template <int FEAT>
__device__ __forceinline__ bool near_cone_slope(const Scene& sc, const Ray& ray) {
    if (!(FEAT & FEAT_GATES)) return false;
    bool fired = false;
    for (int g = 0; g < sc.n_guards; ++g) {
        const double2* q = reinterpret_cast<const double2*>(sc.guards + g);
        const double2 q0 = __ldg(q), q1 = __ldg(q + 1), q2 = __ldg(q + 2), q3 = __ldg(q + 3), q4 = __ldg(q + 4);
        const double dx = q0.x * ray.d.x + q0.y * ray.d.y + q1.x * ray.d.z;
        const double dy = q1.y * ray.d.x + q2.x * ray.d.y + q2.y * ray.d.z;
        const double dz = q3.x * ray.d.x + q3.y * ray.d.y + q4.x * ray.d.z;
        const double a = dx * dx - dy * dy + dz * dz;
        fired = fired || approx0(a);
    }
    return fired;
}

// Shared state of one query: the world ray, the object-space ray cache and the group-box gates it passed.
template <int FEAT>
struct Query {
    const Scene& sc;
    const Ray& ray;
    uint64_t passed;
    ObjRay cache;
    Tally& tally;
    Tally scratch;  // FEAT_TALLY off: the increments land here, are never read, and the compiler drops them
    int* overflow;
    bool skip_negative;  // the query only wants crossings with t >= 0 (set by the closest-hit and shadow visitors)
    __device__ __forceinline__ Query(const Scene& s, const Ray& r, Tally& tl, int* ovf, bool skip_neg)
        : sc(s), ray(r), passed(gates_passed<FEAT>(s, r)), tally(tl), scratch{0u, 0u, 0u, 0u}, overflow(ovf), skip_negative(skip_neg) {
        cache.xform = -1;
    }
    // copy of a query for an out-of-line evaluation: the gates are already known
    __device__ __forceinline__ Query(const Scene& s, const Ray& r, Tally& tl, int* ovf, bool skip_neg, uint64_t passed_gates)
        : sc(s), ray(r), passed(passed_gates), tally(tl), scratch{0u, 0u, 0u, 0u}, overflow(ovf), skip_negative(skip_neg) {
        cache.xform = -1;
    }
    static constexpr bool kTally = (FEAT & FEAT_TALLY) != 0;
    // a persistent kernel hands the query object its next ray (`ray` refers to the caller's variable, already reassigned)
    __device__ __forceinline__ void restart() {
        passed = gates_passed<FEAT>(sc, ray);
        cache.xform = -1;
    }
    __device__ __forceinline__ Tally& counters() { return kTally ? tally : scratch; }
    __device__ __forceinline__ bool gate_ok(int prim_id) const {
        if (!(FEAT & FEAT_GATES)) return true;
        const uint2 m = __ldg(reinterpret_cast<const uint2*>(&sc.prims[prim_id].gate_mask));
        return (((uint64_t)m.y << 32 | m.x) & ~passed) == 0ull;
    }
    // crossings of primitive `id` (by index: unbounded list, CSG programs, brute force)
    template <int HOT = 0x02>  // planes are the usual always-tested primitive
    __device__ __forceinline__ void roots_of_prim(int id, Roots& r) {
        r.n = 0;
        const Prim p = load_prim(sc, id);
        if ((FEAT & FEAT_GATES) && (prim_gates(p) & ~passed) != 0ull) return;
        prim_roots<HOT>(sc, p, ray, cache, counters(), skip_negative, r);
    }
    // crossing of a plane of the always-tested list held in the kernel parameters (ungated by construction)
    __device__ __forceinline__ void roots_of_fast_plane(const FastPlane& p, Roots& r) {
        r.n = 0;
        counters().prims++;
        plane_roots_row(p.row[0], p.row[1], p.row[2], p.row[3], ray, skip_negative, r);
    }
    // crossings of a BVH leaf item that is a primitive
    __device__ __forceinline__ void roots_of_leaf(const LeafItem& li, int slot, Roots& r) {
        r.n = 0;
        if (!gate_ok(li.id)) return;
        prim_roots<(FEAT & FEAT_TRIS) ? 0x160 : 0x61>(sc, li.kind_flags, li.xform, li.geom, sc.leaf_tri + 10 * (size_t)slot, ray, cache, counters(), skip_negative, r);  // triangles, spheres
    }
};

// Group::intersect for a CSG subtree (src/shape.rs:230-269): gather, stable-sort by t, filter, recursively.
// Returns the number of surviving crossings in `list` (sorted by t), or -1 on overflow.
template <int FEAT>
__device__ __forceinline__ int csg_eval(Query<FEAT>& q, const CsgRoot& root, Cross* list) {
    int marks[MAX_CSG_MARKS];
    int nm = 0, n = 0;
    for (int k = 0; k < root.n_ops; ++k) {
        const CsgOp op = q.sc.csg_ops[root.first_op + k];
        if (op.op == OP_PRIM) {
            Roots r;
            q.template roots_of_prim<0x402>(op.a, r);  // (r02w: the intersector dispatch inlined into this out-of-line routine instead of a second call: csg_noise 5.62 -> 5.20 ms)
            if (n + r.n > MAX_CSG_LIST) return -1;
            for (int j = 0; j < r.n; ++j) { list[n].t = r.t[j]; list[n].prim = op.a; n++; }
        } else if (op.op == OP_BEGIN) {
            marks[nm++] = n;
        } else if (op.op == OP_SKIP_IF_EMPTY) {
            if (n == marks[nm - 1]) { k += op.a; --nm; }  // lands on the group's OP_END; its (empty) segment is closed
        } else {
            const int m = marks[--nm];
            for (int i = m + 1; i < n; ++i) {  // stable insertion sort by t
                const Cross c = list[i];
                int j = i - 1;
                while (j >= m && list[j].t > c.t) { list[j + 1] = list[j]; --j; }
                list[j + 1] = c;
            }
            bool in_left = false, in_right = false;
            int out = m;
            for (int i = m; i < n; ++i) {
                const bool left_hit = list[i].prim >= op.b && list[i].prim < op.c;
                const bool keep = allows_intersection(op.a, left_hit, in_left, in_right);
                if (left_hit) in_left = !in_left; else in_right = !in_right;
                if (keep) list[out++] = list[i];
            }
            n = out;
        }
    }
    return n;
}

// Out-of-line evaluation of one CSG item.  It works on COPIES of the query state (see roots_generic in
// device_math.cuh: a pointer into the caller's state handed to a real call would pin that state in local memory).
struct CsgCall {
    Ray ray;
    uint64_t passed;
    int skip_negative;
    Tally tally;
    int n;  // surviving crossings in `list`, sorted by t; -1 = overflow
    Cross list[MAX_CSG_LIST];
};
template <int FEAT>
__device__ __noinline__ void csg_call_body(const Scene& sc, int root, int* overflow, CsgCall* c) {
    const Ray ray = c->ray;
    Tally t = {0u, 0u, 0u, 0u};
    Query<FEAT> q(sc, ray, t, overflow, c->skip_negative != 0, c->passed);
    c->n = csg_eval<FEAT>(q, sc.csg_roots[root], c->list);
    c->tally = t;
}
template <int FEAT>
__device__ __forceinline__ void csg_call(Query<FEAT>& q, int root, CsgCall& c) {
    c.ray = q.ray;
    c.passed = q.passed;
    c.skip_negative = q.skip_negative ? 1 : 0;
    csg_call_body<FEAT>(q.sc, root, q.overflow, &c);
    if (FEAT & FEAT_TALLY) { q.tally.nodes += c.tally.nodes; q.tally.tris += c.tally.tris; q.tally.prims += c.tally.prims; q.tally.xforms += c.tally.xforms; }
    if (c.n < 0) *q.overflow = 1;
}

// ---- fp32 BVH walk ---------------------------------------------------------------------------
struct Walk {
    float px, py, pz, ix, iy, iz;  // p = origin * inverse direction: a slab plane is at fma(plane, i, -p)
    double te;   // world t of the re-origin
    float tlo;   // lower end of the admissible interval, relative to te
};

// Inverse direction component of the fp32 walk.  A slab plane sits at fma(plane, i, -o * i): with an INFINITE i (a
// direction component that is zero, or below ~1e-38 in fp32) the two planes of that axis become inf - inf = NaN when
// the plane and the origin have the same sign and -+inf otherwise, and the min / max chain then culls boxes that
// straddle 0 on that axis although the ray runs inside the slab.  So |i| is clamped to 2^60: multiplying by a power
// of two is exact, hence fma(plane, i, -o * i) = (plane - o) * 2^60 rounded once, its sign is exactly the sign of
// plane - o, and its magnitude — at least 2^60 * (the padding of the boxes, 4e-6 x scene extent) for a ray inside the
// true box — is beyond any t a query can ask for.  Geometrically the clamped axis is a ray tilted by < 2^-60: over the
// scene's extent that is far inside the padding, so the walk stays conservative ("conservative fp32 culling").
__device__ __forceinline__ float walk_inverse(double d) {
    // rcp.approx (one MUFU, <= 1 ulp) instead of an IEEE division (r02b: the three divisions were 10 % of the uniform
    // shadow stage): a relative error of 2^-23 in i moves a slab plane by <= 2^-23 x (distance travelled), far inside
    // the padding of the boxes, and culling — all this feeds — stays conservative.  ftz: a denormal d gives inf -> clamped.
    float i;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(i) : "f"((float)d));
    const float big = 1152921504606846976.0f;  // 2^60
    return fabsf(i) <= big ? i : copysignf(big, i);  // (a NaN direction lands here too: any finite value will do)
}

// mode 0: hits with t >= 0 (closest / shadow); mode 1: every crossing up to tmax (also negative t).
// Returns false if the ray cannot touch any BVH item.
__device__ __forceinline__ bool walk_setup(const Scene& sc, const Ray& ray, int mode, Walk& w) {
    // The root-box interval only decides where the fp32 walk is re-originated and whether the ray can be dropped
    // without a walk — culling, not parity — so it is computed in fp32 (three f64 divisions less per query) and a ray
    // is dropped only when it misses by more than the fp32 error of the interval ends (relative 2^-22 is ~10x the
    // error of a subtract-convert-multiply chain).  The root box is padded like every other box.  NaN ends (a NaN
    // ray) never reject: comparisons with NaN are false, fminf / fmaxf ignore NaN.
    const double o[3] = {ray.o.x, ray.o.y, ray.o.z}, d[3] = {ray.d.x, ray.d.y, ray.d.z};
    w.ix = walk_inverse(d[0]);
    w.iy = walk_inverse(d[1]);
    w.iz = walk_inverse(d[2]);
    const float inv[3] = {w.ix, w.iy, w.iz};
    float tn = -CUDART_INF_F, tf = CUDART_INF_F;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float a = (float)(sc.root_lo[k] - o[k]) * inv[k], b = (float)(sc.root_hi[k] - o[k]) * inv[k];
        tn = fmaxf(tn, fminf(a, b));
        tf = fminf(tf, fmaxf(a, b));
    }
    const float slack = 2.4e-7f * (fminf(fabsf(tn), 1e30f) + fminf(fabsf(tf), 1e30f));  // (an infinite end still rejects)
    if (tn - tf > slack) return false;
    if (mode == 0) {
        if (tf < -slack) return false;
        w.te = (double)fmaxf(tn, 0.0f);
        w.tlo = __double2float_rd(-w.te);
    } else {
        w.te = (double)fminf(fmaxf(0.0f, tn), tf);
        w.tlo = -CUDART_INF_F;
    }
    if (!(fabs(w.te) <= 1.7e38)) { w.te = 0.0; w.tlo = mode == 0 ? 0.0f : -CUDART_INF_F; }
    // One fused multiply-add per plane instead of a subtract and a multiply (explicit: the build has -fmad=false for
    // the f64 code).  The extra rounding of p is <= 2^-24 |o| in distance, far inside the padding of the boxes
    // ("conservative fp32 culling").  Zero direction components: see walk_inverse.
    w.px = (float)(o[0] + w.te * d[0]) * w.ix;
    w.py = (float)(o[1] + w.te * d[1]) * w.iy;
    w.pz = (float)(o[2] + w.te * d[2]) * w.iz;
    return true;
}

__device__ __forceinline__ bool slab(const Walk& w, float lx, float ly, float lz, float hx, float hy, float hz, float thi, float& tnear) {
    const float ax = __fmaf_rn(lx, w.ix, -w.px), bx = __fmaf_rn(hx, w.ix, -w.px);
    const float ay = __fmaf_rn(ly, w.iy, -w.py), by = __fmaf_rn(hy, w.iy, -w.py);
    const float az = __fmaf_rn(lz, w.iz, -w.pz), bz = __fmaf_rn(hz, w.iz, -w.pz);
    const float t0 = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), w.tlo));
    const float t1 = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), thi));
    tnear = t0;
    return t0 <= t1;
}

// upper end of the admissible interval relative to the re-origin, rounded up (conservative)
__device__ __forceinline__ float walk_thi(double tmax, double te) { return __double2float_ru(__dsub_ru(tmax, te)); }

// Visitor protocol:  bool V::leaf(int slot) -> true stops the walk;  double V::tmax() -> current upper bound on t
// (re-read only after a leaf was visited: it cannot change in between).
//
// The walk is resumable: a cursor holds everything between two steps (walk_begin / walk_descend / walk_leaf); walk_bvh
// runs a query to its end.  (r02b tried persistent warps that hand a lane whose query has ended the next ray of a queue:
// 14 instead of 10 of 32 lanes in the node loop, no fewer instructions in total, worse overlap between the batch
// pipelines — profiles/README.md.)
constexpr int WALK_DONE = (int)0x80000000;  // never a leaf link: ~enc with enc >= 0 is > INT_MIN
struct WalkCursor {
    Walk w;
    float thi;       // upper end of the admissible interval, relative to the re-origin
    int link;        // >= 0: inner node; WALK_DONE: the query is over; any other negative value: a leaf link
    int sp;
    int brute_next;  // >= 0: brute-force scan (RTC_FLAG_BRUTE_FORCE, cone guards), the next item to test; -1: BVH walk
};
// The pending subtrees live in a separate array (local memory: it is indexed dynamically); keeping it out of the cursor
// lets the cursor's scalars stay in registers (r02b: with the array inside, the node loop reloaded the walk constants
// from local memory on every iteration).
struct WalkStack {
    int at[BVH_STACK];
};

// Starts a query.  Returns false (link = WALK_DONE) when the ray cannot touch any BVH item.
__device__ __forceinline__ bool walk_begin(const Scene& sc, const Ray& ray, int mode, bool brute, double tmax, WalkCursor& c) {
    c.sp = 0;
    c.brute_next = -1;
    c.link = WALK_DONE;
    if (sc.n_bvh_items == 0) return false;
    if (brute) {
        c.link = ~0;  // the leaf holding item 0 alone
        c.brute_next = 1;
        return true;
    }
    if (!walk_setup(sc, ray, mode, c.w)) return false;
    c.link = sc.bvh_root;
    c.thi = walk_thi(tmax, c.w.te);
    return true;
}

// Node phase: descends through inner nodes until this lane stands on a leaf (or its query is over).
// "while-while" (Aila & Laine): every lane first descends until it stands on a leaf, THEN the warp runs the leaf code
// once for all of them.  With a single if/else loop almost every iteration has some lane on a leaf, so every
// iteration pays for the (4-5x more expensive, f64) leaf code.
// top / n_top: the first n_top nodes staged in shared memory by the caller (persistent kernels), or (nullptr, 0).
template <bool TALLY>
__device__ __forceinline__ void walk_descend(const Scene& sc, WalkCursor& c, WalkStack& stack, Tally& tally, const BvhNode* __restrict__ top, int n_top) {
    int link = c.link, sp = c.sp;
    const float thi = c.thi;
    const Walk w = c.w;
    while (link >= 0) {
        if (TALLY) tally.nodes++;
#ifdef RTC_STAGE_TOP
        // one load path for both homes of a node (generic addressing reaches shared and global memory alike; two separate
        // paths cost the walk 12 % in registers and code)
        const float4* q = reinterpret_cast<const float4*>(link < n_top ? top + link : sc.bvh + link);
        const float4 q0 = q[0], q1 = q[1], q2 = q[2], q3 = q[3];
#else
        const float4* q = reinterpret_cast<const float4*>(sc.bvh + link);
        const float4 q0 = __ldg(q), q1 = __ldg(q + 1), q2 = __ldg(q + 2), q3 = __ldg(q + 3);
#endif
        float n0, n1;
        const bool h0 = slab(w, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, thi, n0);
        const bool h1 = slab(w, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, thi, n1);
        const int c0 = __float_as_int(q3.x), c1 = __float_as_int(q3.y);
        if (h0 && h1) {
            const bool first0 = n0 <= n1;
            stack.at[sp++] = first0 ? c1 : c0;  // (prefetching the deferred child's line here changed nothing: r02i)
            link = first0 ? c0 : c1;
        } else if (h0) {
            link = c0;
        } else if (h1) {
            link = c1;
        } else {
            link = sp ? stack.at[--sp] : WALK_DONE;
        }
    }
    c.link = link;
    c.sp = sp;
}

// Leaf phase: tests the items of the leaf this lane stands on, then moves to the next pending subtree.
// Returns true when the query is over (the visitor stopped it, or nothing is left to visit).
template <class V>
__device__ __forceinline__ bool walk_leaf(const Scene& sc, WalkCursor& c, WalkStack& stack, V& v) {
    const int enc = ~c.link;
    const int first = enc >> 2, count = (enc & 3) + 1;
    for (int i = 0; i < count; ++i)
        if (v.leaf(first + i)) { c.link = WALK_DONE; return true; }
    if (c.brute_next >= 0) {
        if (c.brute_next >= sc.n_bvh_items) { c.link = WALK_DONE; return true; }
        c.link = ~(c.brute_next << 2);
        c.brute_next++;
        return false;
    }
    c.thi = walk_thi(v.tmax(), c.w.te);
    if (c.sp == 0) { c.link = WALK_DONE; return true; }
    c.link = stack.at[--c.sp];
    return false;
}

// The first nodes of the tree (breadth-first: every walk starts there), staged in shared memory by the kernel.
struct TopNodes {
    const BvhNode* nodes = nullptr;
    int n = 0;
};

// Built, measured, and off by default (RTC_STAGE_TOP): the top levels of a tree every walk starts in are L1-resident
// anyway, and the staged walk loses the read-only path for the rest of the tree — teapot 3840x2160 5.03 -> 5.34 ms with
// generic loads, 5.54 ms with separate shared / global load paths (profiles/README.md, r02i).
// Stages the first sc.n_top nodes into `dst` (shared memory, 128-byte aligned, room for MAX_TOP_NODES) with ONE bulk
// asynchronous copy (cp.async.bulk, the TMA's 1-D form; UBLKCP in SASS) completed through an mbarrier: thread 0 arms the
// barrier with the byte count and issues the copy, every thread then waits for the barrier's phase to flip.
// Call from all threads of the block, once, before the first query.
__device__ __forceinline__ TopNodes stage_top_nodes(const Scene& sc, BvhNode* dst, unsigned long long* bar) {
    TopNodes top;
    if (sc.n_top <= 0) return top;  // (uniform over the grid)
    const uint32_t bar_s = (uint32_t)__cvta_generic_to_shared(bar), dst_s = (uint32_t)__cvta_generic_to_shared(dst);
    const uint32_t bytes = (uint32_t)sc.n_top * (uint32_t)sizeof(BvhNode);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_s));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_s), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_s), "l"(sc.bvh), "r"(bytes), "r"(bar_s) : "memory");
    }
    uint32_t done = 0;
    while (!done)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(bar_s) : "memory");
    top.nodes = dst;
    top.n = sc.n_top;
    return top;
}

template <class V>
__device__ __forceinline__ void walk_bvh(const Scene& sc, const Ray& ray, int mode, bool brute, V& v, Tally& tally, TopNodes top = TopNodes{}) {
    WalkCursor c;
    WalkStack stack;
    if (!walk_begin(sc, ray, mode, brute, v.tmax(), c)) return;
    for (;;) {
        walk_descend<V::kTally>(sc, c, stack, tally, top.nodes, top.n);
        if (c.link == WALK_DONE) return;
        if (walk_leaf(sc, c, stack, v)) return;
    }
}

// The always-tested primitives, before the walk.  Returns true if the visitor stopped the query.
template <class V>
__device__ __forceinline__ bool visit_unbounded(const Scene& sc, V& v) {
    if (sc.n_fast_planes > 0) {
        for (int i = 0; i < sc.n_fast_planes; ++i)
            if (v.plane(sc.fast_planes[i])) return true;
        return false;
    }
    for (int i = 0; i < sc.n_unbounded; ++i)
        if (v.item(__ldg(sc.unbounded + i))) return true;
    return false;
}

// ---- closest hit ---------------------------------------------------------------------------------
template <int FEAT>
struct ClosestVisitor {
    static constexpr bool kTally = (FEAT & FEAT_TALLY) != 0;
    Query<FEAT> q;
    Hit best;
    double limit;   // exclusive upper bound on t (shadow queries on scenes with non-casting shapes), else +inf
    __device__ __forceinline__ ClosestVisitor(const Scene& sc, const Ray& ray, double lim, Tally& tl, int* ovf) : q(sc, ray, tl, ovf, !(FEAT & FEAT_CSG)), limit(lim) {
        best.t = 0.0; best.prim = -1; best.u = 0.0; best.v = 0.0;
    }
    __device__ __forceinline__ double tmax() const { return best.prim >= 0 ? best.t : limit; }
    // next ray of a persistent kernel: its limit and the nearest crossing already known (prim0 < 0: none)
    __device__ __forceinline__ void restart(double lim, double t0, int prim0) {
        q.restart();
        limit = lim;
        best.t = t0; best.prim = prim0; best.u = 0.0; best.v = 0.0;
    }
    __device__ __forceinline__ void offer(double t, int prim, double u, double v) {
        if (t >= 0.0 && t < limit && (best.prim < 0 || t < best.t || (t == best.t && prim < best.prim))) {
            best.t = t; best.prim = prim; best.u = u; best.v = v;
        }
    }
    __device__ __forceinline__ void offer_roots(const Roots& r, int id) {
        RTC_ROOTS_LOOP
        for (int k = 0; k < r.n; ++k) offer(r.t[k], id, r.u, r.v);
    }
    __device__ __forceinline__ bool csg(int root) {
        CsgCall c;
        csg_call<FEAT>(q, root, c);
        const int n = c.n;
        const Cross* list = c.list;
        if (n < 0) return true;
        for (int k = 0; k < n; ++k)
            if (list[k].t >= 0.0) {  // list is sorted: the first non-negative entry is this item's hit
                double u = 0.0, v = 0.0;
                const Prim p = load_prim(q.sc, list[k].prim);
                if ((p.kind_flags & 0xff) >= 5) {  // triangle: recover (u, v)
                    Roots r;
                    prim_roots<0>(q.sc, p, q.ray, q.cache, q.counters(), false, r);
                    u = r.u; v = r.v;
                }
                offer(list[k].t, list[k].prim, u, v);
                break;
            }
        return false;
    }
    __device__ __forceinline__ bool item(int id) {  // unbounded list entry
        if ((FEAT & FEAT_CSG) && id < 0) return csg(~id);
        Roots r;
        q.roots_of_prim(id, r);
        offer_roots(r, id);
        return false;
    }
    __device__ __forceinline__ bool plane(const FastPlane& p) {
        // t = -oy / dy costs an IEEE division (~30 instructions); a plane that certainly lies beyond the nearest crossing
        // found so far (or beyond the query's limit) is dropped on a product instead — the result is unchanged.
        double oy, dy;
        q.counters().prims++;
        if (!plane_fraction(p.row[0], p.row[1], p.row[2], p.row[3], q.ray, q.skip_negative, oy, dy)) return false;
        if (plane_certainly_above(oy, dy, tmax())) return false;
        offer(-oy / dy, p.prim, 0.0, 0.0);
        return false;
    }
    __device__ __forceinline__ bool leaf(int slot) {
        const LeafItem li = load_leaf(q.sc, slot);
        if ((FEAT & FEAT_CSG) && li.id < 0) return csg(~li.id);
        Roots r;
        q.roots_of_leaf(li, slot, r);
        offer_roots(r, li.id);
        return false;
    }
};

template <int FEAT>
__device__ __forceinline__ Hit closest_hit(const Scene& sc, const Ray& ray, double limit, bool brute, Tally& tally, int* overflow, TopNodes top = TopNodes{}) {
    brute = brute || near_cone_slope<FEAT>(sc, ray);
    ClosestVisitor<FEAT> v(sc, ray, limit, tally, overflow);
    visit_unbounded(sc, v);
    walk_bvh(sc, ray, 0, brute, v, tally, top);
    return v.best;
}

// ---- closest hit AND the crossings before it, in one walk (FEAT_ONEPASS) ------------------------------------------
// Intersection::prepare_state derives n1 / n2 from EVERY crossing of the ray's line up to the hit, those behind the origin
// included (src/intersection.rs:70-103).  The two-pass form walks once for the closest hit (t >= 0 only) and, for hits on
// transparent shapes, a second time for that list.  Here one walk without a lower bound on t does both: every crossing
// not beyond the nearest hit found so far is kept in `out`, and whenever the list fills up the entries the hit has since
// overtaken are dropped.  The hit cannot move further away, so a crossing at t <= t_hit is never discarded or culled.
// Drops the crossings the nearest hit has overtaken (t > t_hit) from a full list.  `take` is inlined at every place a
// crossing is offered, and with n == MAX_CROSSINGS known the compiler unrolls this rarely-run loop completely at each of
// them: 5 588 of the 13 648 instructions of trace_spawn<23,0> (tools/code_footprint.py, r02x) — in the CSG kernels, whose
// dominant stall is the instruction fetch, the loop stays ROLLED.  The small transparent-only kernels measure the other way
// (glass bubble 4.06 ms unrolled, 4.39 ms rolled, 4.14-4.35 ms as an out-of-line call), so they keep the unrolled form.
template <bool ROLLED>
__device__ __forceinline__ int drop_overtaken(Cross* list, int n, double t_hit) {
    int k = 0;
    if (ROLLED) {
#pragma unroll 1
        for (int i = 0; i < n; ++i)
            if (list[i].t <= t_hit) list[k++] = list[i];
    } else {
        for (int i = 0; i < n; ++i)
            if (list[i].t <= t_hit) list[k++] = list[i];
    }
    return k;
}
template <int FEAT>
struct ClosestCollectVisitor {
    static constexpr bool kTally = (FEAT & FEAT_TALLY) != 0;
    Query<FEAT> q;
    Hit best;
    Cross* out;
    int n;
    bool full;
    __device__ __forceinline__ ClosestCollectVisitor(const Scene& sc, const Ray& ray, Cross* list, Tally& tl, int* ovf) : q(sc, ray, tl, ovf, false), out(list), n(0), full(false) {
        best.t = 0.0; best.prim = -1; best.u = 0.0; best.v = 0.0;
    }
    __device__ __forceinline__ double tmax() const { return best.prim >= 0 ? best.t : CUDART_INF; }
    __device__ __forceinline__ void take(double t, int prim, double u, double v) {
        if (best.prim < 0 || t <= best.t) {
            if (n == MAX_CROSSINGS && best.prim >= 0) n = drop_overtaken<(FEAT & FEAT_CSG) != 0>(out, n, best.t);
            if (n == MAX_CROSSINGS) full = true;
            else { out[n].t = t; out[n].prim = prim; n++; }
        }
        if (t >= 0.0 && (best.prim < 0 || t < best.t || (t == best.t && prim < best.prim))) {
            best.t = t; best.prim = prim; best.u = u; best.v = v;
        }
    }
    __device__ __forceinline__ bool take_roots(const Roots& r, int id) {
        RTC_ROOTS_LOOP
        for (int k = 0; k < r.n; ++k) take(r.t[k], id, r.u, r.v);
        return full;
    }
    __device__ __forceinline__ bool csg(int root) {
        CsgCall c;
        csg_call<FEAT>(q, root, c);
        if (c.n < 0) return true;
        bool first = true;
        for (int k = 0; k < c.n; ++k) {
            double u = 0.0, v = 0.0;
            if (first && c.list[k].t >= 0.0) {  // the list is sorted: the first non-negative entry is the only one that can be this item's hit
                first = false;
                const Prim p = load_prim(q.sc, c.list[k].prim);
                if ((p.kind_flags & 0xff) >= 5) {  // triangle: recover (u, v)
                    Roots r;
                    prim_roots<0>(q.sc, p, q.ray, q.cache, q.counters(), false, r);
                    u = r.u; v = r.v;
                }
            }
            take(c.list[k].t, c.list[k].prim, u, v);
        }
        return full;
    }
    __device__ __forceinline__ bool item(int id) {
        if ((FEAT & FEAT_CSG) && id < 0) return csg(~id);
        Roots r;
        q.roots_of_prim(id, r);
        return take_roots(r, id);
    }
    __device__ __forceinline__ bool plane(const FastPlane& p) {
        double oy, dy;
        q.counters().prims++;
        if (!plane_fraction(p.row[0], p.row[1], p.row[2], p.row[3], q.ray, false, oy, dy)) return false;
        const bool behind = (oy > 0.0 && dy > 0.0) || (oy < 0.0 && dy < 0.0);  // t < 0: always wanted
        if (!behind && plane_certainly_above(oy, dy, tmax())) return false;
        take(-oy / dy, p.prim, 0.0, 0.0);
        return full;
    }
    __device__ __forceinline__ bool leaf(int slot) {
        const LeafItem li = load_leaf(q.sc, slot);
        if ((FEAT & FEAT_CSG) && li.id < 0) return csg(~li.id);
        Roots r;
        q.roots_of_leaf(li, slot, r);
        return take_roots(r, li.id);
    }
};

// Returns the closest hit; xs[0..n_xs) holds (a superset of) the crossings with t <= hit.t, unsorted.
template <int FEAT>
__device__ __forceinline__ Hit closest_hit_collect(const Scene& sc, const Ray& ray, bool brute, Cross* xs, int& n_xs, Tally& tally, int* overflow, TopNodes top = TopNodes{}) {
    brute = brute || near_cone_slope<FEAT>(sc, ray);
    ClosestCollectVisitor<FEAT> v(sc, ray, xs, tally, overflow);
    if (!visit_unbounded(sc, v)) walk_bvh(sc, ray, 1, brute, v, tally, top);
    if (v.full) *overflow = 1;
    n_xs = v.n;
    return v.best;
}

// ---- shadow query: World::is_shadowed (src/world.rs:26-48) -------------------------------------------
template <int FEAT>
struct AnyVisitor {
    static constexpr bool kTally = (FEAT & FEAT_TALLY) != 0;
    Query<FEAT> q;
    double dist;
    bool found;
    __device__ __forceinline__ AnyVisitor(const Scene& sc, const Ray& ray, double d, Tally& tl, int* ovf) : q(sc, ray, tl, ovf, !(FEAT & FEAT_CSG)), dist(d), found(false) {}
    __device__ __forceinline__ double tmax() const { return dist; }
    __device__ __forceinline__ void restart(double d) {  // next ray of a persistent kernel
        q.restart();
        dist = d;
        found = false;
    }
    __device__ __forceinline__ bool test(const Roots& r) {
        for (int k = 0; k < r.n; ++k)
            if (r.t[k] >= 0.0 && r.t[k] < dist) { found = true; return true; }
        return false;
    }
    __device__ __forceinline__ bool csg(int root) {
        CsgCall c;
        csg_call<FEAT>(q, root, c);
        if (c.n < 0) return true;
        for (int k = 0; k < c.n; ++k)
            if (c.list[k].t >= 0.0 && c.list[k].t < dist) { found = true; return true; }
        return false;
    }
    __device__ __forceinline__ bool item(int id) {
        if ((FEAT & FEAT_CSG) && id < 0) return csg(~id);
        Roots r;
        q.roots_of_prim(id, r);
        return test(r);
    }
    __device__ __forceinline__ bool plane(const FastPlane& p) {
        // "0 <= -oy / dy < dist" without the division unless the crossing is within 1e-14 (relative) of the light
        double oy, dy;
        q.counters().prims++;
        if (!plane_fraction(p.row[0], p.row[1], p.row[2], p.row[3], q.ray, true, oy, dy)) return false;  // (no crossing, or t < 0)
        if (plane_certainly_above(oy, dy, dist)) return false;
        if (!plane_certainly_below(oy, dy, dist)) {
            const double t = -oy / dy;
            if (!(t >= 0.0 && t < dist)) return false;
        }
        found = true;
        return true;
    }
    __device__ __forceinline__ bool leaf(int slot) {
        const LeafItem li = load_leaf(q.sc, slot);
        if ((FEAT & FEAT_CSG) && li.id < 0) return csg(~li.id);
        Roots r;
        q.roots_of_leaf(li, slot, r);
        return test(r);
    }
};

// shadowed <=> the nearest crossing with t >= 0 belongs to a casting shape and has t < distance.
// Only instantiated for scenes where every primitive casts shadows: then "nearest t < dist" <=> "any t in
// [0, dist)" and the walk may stop at the first such crossing.  (Scenes with non-casting shapes use closest_hit.)
template <int FEAT>
__device__ __forceinline__ bool any_hit(const Scene& sc, const Ray& ray, double dist, bool brute, Tally& tally, int* overflow, TopNodes top = TopNodes{}) {
    brute = brute || near_cone_slope<FEAT>(sc, ray);
    AnyVisitor<FEAT> v(sc, ray, dist, tally, overflow);
    if (visit_unbounded(sc, v)) return v.found;
    walk_bvh(sc, ray, 0, brute, v, tally, top);
    return v.found;
}

// The query a shadow ray runs (World::is_shadowed, src/world.rs:26-48): any crossing in [0, distance) when every shape
// casts shadows; otherwise the nearest crossing below `distance`, which shadows iff its shape casts shadows.
template <int FEAT, bool ANYHIT>
struct ShadowVisitor;
template <int FEAT>
struct ShadowVisitor<FEAT, true> {
    using type = AnyVisitor<FEAT>;
    __device__ __forceinline__ static void restart(type& v, double dist, double, int) { v.restart(dist); }
    __device__ __forceinline__ static bool shadowed(const Scene&, const type& v) { return v.found; }
};
template <int FEAT>
struct ShadowVisitor<FEAT, false> {
    using type = ClosestVisitor<FEAT>;
    __device__ __forceinline__ static void restart(type& v, double dist, double t0, int prim0) { v.restart(dist, t0, prim0); }
    __device__ __forceinline__ static bool shadowed(const Scene& sc, const type& v) {
        return v.best.prim >= 0 && (__ldg(&sc.prims[v.best.prim].kind_flags) & 0x100) != 0;
    }
};

// ---- n1 / n2: replay of the container walk of Intersection::prepare_state (src/intersection.rs:70-103) ----
template <int FEAT>
struct CollectVisitor {
    static constexpr bool kTally = (FEAT & FEAT_TALLY) != 0;
    Query<FEAT> q;
    double thit;
    Cross* out;
    int n;
    __device__ __forceinline__ CollectVisitor(const Scene& sc, const Ray& ray, double th, Cross* o, Tally& tl, int* ovf) : q(sc, ray, tl, ovf, false), thit(th), out(o), n(0) {}
    __device__ __forceinline__ double tmax() const { return thit; }
    __device__ __forceinline__ bool push(double t, int prim) {
        if (!(t <= thit)) return false;
        if (n >= MAX_CROSSINGS) { *q.overflow = 1; return true; }
        out[n].t = t; out[n].prim = prim; n++;
        return false;
    }
    __device__ __forceinline__ bool push_roots(const Roots& r, int id) {
        RTC_ROOTS_LOOP
        for (int k = 0; k < r.n; ++k)
            if (push(r.t[k], id)) return true;
        return false;
    }
    __device__ __forceinline__ bool csg(int root) {
        CsgCall c;
        csg_call<FEAT>(q, root, c);
        if (c.n < 0) return true;
        for (int k = 0; k < c.n; ++k)
            if (push(c.list[k].t, c.list[k].prim)) return true;
        return false;
    }
    __device__ __forceinline__ bool item(int id) {
        if ((FEAT & FEAT_CSG) && id < 0) return csg(~id);
        Roots r;
        q.roots_of_prim(id, r);
        return push_roots(r, id);
    }
    __device__ __forceinline__ bool plane(const FastPlane& p) {
        Roots r;
        q.roots_of_fast_plane(p, r);
        return push_roots(r, p.prim);
    }
    __device__ __forceinline__ bool leaf(int slot) {
        const LeafItem li = load_leaf(q.sc, slot);
        if ((FEAT & FEAT_CSG) && li.id < 0) return csg(~li.id);
        Roots r;
        q.roots_of_leaf(li, slot, r);
        return push_roots(r, li.id);
    }
};

// The container walk of Intersection::prepare_state (src/intersection.rs:70-103) over xs[0..n): sorts the crossings the way
// the reference's stable sort leaves them (by t, ties in depth-first primitive order) and replays the list.
struct IndexResult {
    double n1, n2;
};
__device__ __noinline__ void replay_containers(const Scene& sc, Cross* xs, int n, double hit_t, int hit_prim, int* overflow, IndexResult* res) {
    int m = 0;
    for (int i = 0; i < n; ++i)  // (a one-pass list may still hold crossings the hit has overtaken)
        if (xs[i].t <= hit_t) xs[m++] = xs[i];
    n = m;
    for (int i = 1; i < n; ++i) {
        const Cross c = xs[i];
        int j = i - 1;
        while (j >= 0 && (xs[j].t > c.t || (xs[j].t == c.t && xs[j].prim > c.prim))) { xs[j + 1] = xs[j]; --j; }
        xs[j + 1] = c;
    }
    int open[MAX_CONTAINERS];
    int no = 0;
    double n1 = 1.0, n2 = 1.0;
    for (int i = 0; i < n; ++i) {
        const bool is_hit = xs[i].t == hit_t && xs[i].prim == hit_prim;  // Intersection::eq, src/intersection.rs:135-139
        if (is_hit) n1 = no ? sc.materials[sc.prims[open[no - 1]].material].refractive_index : 1.0;
        int at = -1;
        for (int k = 0; k < no; ++k)
            if (open[k] == xs[i].prim) at = k;
        if (at >= 0) {
            for (int k = at; k + 1 < no; ++k) open[k] = open[k + 1];
            no--;
        } else {
            if (no >= MAX_CONTAINERS) { *overflow = 1; break; }
            open[no++] = xs[i].prim;
        }
        if (is_hit) n2 = no ? sc.materials[sc.prims[open[no - 1]].material].refractive_index : 1.0;
    }
    res->n1 = n1;
    res->n2 = n2;
}

struct IndexCall {
    Ray ray;
    double n1, n2;
    Tally tally;
};
template <int FEAT>
__device__ __noinline__ void refractive_indices_body(const Scene& sc, double hit_t, int hit_prim, bool brute, int* overflow, IndexCall* call) {
    const Ray ray = call->ray;
    brute = brute || near_cone_slope<FEAT>(sc, ray);
    Tally tally = {0u, 0u, 0u, 0u};
    Cross xs[MAX_CROSSINGS];
    CollectVisitor<FEAT> v(sc, ray, hit_t, xs, tally, overflow);
    visit_unbounded(sc, v);
    walk_bvh(sc, ray, 1, brute, v, tally);
    IndexResult res;
    replay_containers(sc, xs, v.n, hit_t, hit_prim, overflow, &res);
    call->n1 = res.n1;
    call->n2 = res.n2;
    call->tally = tally;
}
// Out of line, over copies of the ray and the tallies (see roots_generic in device_math.cuh).
template <int FEAT>
__device__ __forceinline__ void refractive_indices(const Scene& sc, const Ray& ray, double hit_t, int hit_prim, bool brute, double& n1, double& n2, Tally& tally, int* overflow) {
    IndexCall c;
    c.ray = ray;
    refractive_indices_body<FEAT>(sc, hit_t, hit_prim, brute, overflow, &c);
    n1 = c.n1;
    n2 = c.n2;
    if (FEAT & FEAT_TALLY) { tally.nodes += c.tally.nodes; tally.tris += c.tally.tris; tally.prims += c.tally.prims; tally.xforms += c.tally.xforms; }
}

}  // namespace rtcdev
