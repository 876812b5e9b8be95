// Device-side scene layout (HBM-resident, read-only during a render; everything fits in L2).
// Built by scene_build.cpp from the rtc.h arrays; consumed by render_kernels.cu.
#pragma once
#include <stdint.h>

namespace rtcdev {

constexpr int MAX_LIGHTS = 16;
constexpr int MAX_GATES = 64;           // group boxes that must be evaluated exactly per ray (one bit each in Prim::gate_mask / gate_mask_hi)
constexpr int MAX_PATTERN_FRAMES = 8;   // nesting of Blend/Gradient/RingGradient/ColorJitter along one path
constexpr int MAX_CSG_LIST = 48;        // crossings held while evaluating one top-level CSG item
constexpr int MAX_CSG_MARKS = 8;        // CSG nesting depth
constexpr int MAX_CROSSINGS = 64;       // crossings (t <= t_hit) replayed for n1/n2 (src/intersection.rs:70-103)
constexpr int MAX_CONTAINERS = 24;      // simultaneously open shapes during that replay
constexpr int MAX_TOP_NODES = 127;      // nodes 0 .. n_top-1 are the top of the tree in breadth-first order (7 levels): staged in shared memory (8 KB)
constexpr int BVH_STACK = 64;           // a radix tree over unique 62-bit keys (30-bit Morton code, 32-bit item) is at most 62 deep
constexpr int LBVH_ONE_BLOCK_ITEMS = 8192;  // device-side build (bvh_build.cu): up to here one block sorts the keys in shared memory
constexpr int LBVH_MAX_ITEMS = 1 << 22;     // ... beyond, a multi-block radix sort (4 M items: scratch ~100 B per item)
constexpr int LBVH_MIN_ITEMS = 64;      // below this the host build costs microseconds

// Kernel feature bits (compile-time variants of the wavefront kernel)
constexpr int FEAT_TRANSPARENT = 1;  // some material has transparency != 0: n1/n2 replay needed
constexpr int FEAT_CSG = 2;          // CSG items present
constexpr int FEAT_GATES = 4;        // per-ray exact pre-tests: group boxes that must be evaluated exactly, cone guards
constexpr int FEAT_ONEPASS = 16;     // transparent scene with a small BVH: the closest-hit walk also gathers the crossings the n1/n2 replay needs
constexpr int ONEPASS_MAX_ITEMS = 512;  // (with more BVH items the walk over t < 0 that this needs for EVERY ray costs more than a second walk for the transparent hits)
constexpr int FEAT_TRIS = 32;        // every BVH item is a triangle (a mesh scene): the leaf code is Moller-Trumbore only — no kind dispatch and no
                                     // out-of-line generic intersector in the walk (teapot: shadows -12 %, frame -3 %)
constexpr int FEAT_TALLY = 8;        // RTC_FLAG_PROFILE: count the work done (node visits, primitive tests, ...); off, the counters and
                                     // their registers do not exist in the kernel

// Per distinct (transform_inv, transform_inv_tsp) pair.  Affine: rows 0..2 only.
struct Xform {
    double inv[12];  // world -> object, row-major 3x4
    double tsp[9];   // upper 3x3 of transform_inv_tsp (normals)
    double pad[3];
};  // 192 B

struct MatInv {
    double m[16];  // world -> pattern space, full row-major 4x4: pattern space carries w (see device_pattern.cuh)
};

// kind_flags: bits 0..7 kind (rtc_prim_kind), bit 8 casts_shadow, bit 9 closed, bit 10 PRIM_IN_BVH
constexpr int32_t PRIM_IN_BVH = 0x400;  // reached through the BVH (an item, or part of a CSG item), not through the always-tested list
struct Prim {
    int32_t kind_flags;
    int32_t material;
    int32_t xform;
    int32_t matinv;
    int32_t geom;        // triangles: slot in tri_pe / tri_n; cylinder / cone: slot in ranges
    int32_t pad1;
    uint32_t gate_mask;     // non-trivial ancestor group boxes that must pass (src/shape.rs:251): gates 0..31 ... (8-byte aligned pair)
    uint32_t gate_mask_hi;  // ... and gates 32..63
};  // 32 B

struct Material {
    double ambient, diffuse, specular, shininess, reflective, transparency, refractive_index;
    int32_t pattern;
    int32_t flags;  // bit 0: pattern is Plain (colour inlined below)
    double plain[3];
    double pad;
};  // 96 B

struct Pattern {
    int32_t kind, left, right, noise_kind;
    int32_t octaves, pad;
    double noise_scale;
    double color[3];
    double m[16];  // mixture transform_inv, full row-major 4x4
    double pad2;
};  // 192 B

struct Light {
    double intensity[3];
    double origin[3];
};

struct Gate {  // a group box that is neither always-passing nor provably conservative: evaluated exactly per ray
    double lo[3], hi[3];
};

// A cone with a finite range sits in the BVH under its finite box.  That box holds every crossing the reference can
// emit EXCEPT the one of the single-root branch (|a| < EPSILON, src/shape.rs:811-818), which has no y-range check.
// `a` depends only on the object-space direction, so a query first evaluates the reference's own expression per
// guarded cone (linear part of Ray::transform, then a) and tests every item if one of them is below EPSILON.
struct ConeGuard {
    double lin[9];  // upper 3x3 of transform_inv, row-major
    double pad;
};  // 80 B, 16-byte aligned

// BVH2 node, 64 B, two float4-aligned child boxes (outward-rounded, padded) + two child links.
// link >= 0: inner node index; link < 0: leaf, ~link = (first_item << 2) | (count - 1), count <= 4.
struct BvhNode {
    float lo0[3], hi0[3];
    float lo1[3], hi1[3];
    int32_t c0, c1;
    int32_t pad0, pad1;
};

// One BVH item in LEAF ORDER: everything the traversal needs to test it without touching prims[].
// Triangles additionally have their (p1, e1, e2) record at the same slot of Scene::leaf_tri, so the items of
// one leaf are contiguous in memory and a leaf visit is two independent loads instead of a dependent chain.
struct LeafItem {
    int32_t id;          // >= 0 primitive index, < 0: ~csg_root index
    int32_t kind_flags;  // as Prim::kind_flags (unused for CSG roots)
    int32_t xform;
    int32_t geom;        // cylinder / cone: slot in ranges
};  // 16 B

// CSG evaluation program of one top-level CSG item (post-order over its subtree).
//   OP_PRIM  a = primitive index
//   OP_BEGIN                     (push list length)
//   OP_END   a = group kind, b/c = primitive index range [b, c) of the LEFT child's subtree
//   OP_SKIP_IF_EMPTY  a = ops to jump over.  Sits between the left and the right operand of a Difference or an
//            Intersection: when the left operand produced no crossing the filter (src/shape.rs:230-269) keeps nothing
//            of the right one either (a right crossing survives only inside the left solid), so the right operand's
//            primitives are not intersected at all and the group's result is the empty list it would have been.
struct CsgOp {
    int32_t op, a, b, c;
};
enum { OP_PRIM = 0, OP_BEGIN = 1, OP_END = 2, OP_SKIP_IF_EMPTY = 3 };
struct CsgRoot {
    int32_t first_op, n_ops;
    int32_t prim_begin, prim_end;  // all primitives of the subtree are [prim_begin, prim_end)
};

// A plane of the always-tested list, held in the kernel parameters (constant bank): row 1 of its transform_inv — all
// a plane intersection reads (src/shape.rs:622-633) — and its primitive index.  Filled when the always-tested list
// consists of ungated planes only (every bin of the reference): the per-ray plane loop then issues no global loads,
// where the general list walks unbounded[] -> prims[] -> xforms[] (three dependent loads per plane).
constexpr int MAX_FAST_PLANES = 8;
struct FastPlane {
    double row[4];
    int32_t prim;
    int32_t pad;
};  // 40 B

struct Scene {
    const Prim* prims;
    const Xform* xforms;
    const MatInv* matinvs;
    const double* tri_pe;   // 10 doubles per triangle: p1, e1, e2, pad   (16-byte aligned records)
    const double* tri_n;    // 10 doubles per triangle: n1, n2, n3, pad
    const double* ranges;   // 2 doubles per cylinder / cone: min, max
    const Material* materials;
    const Pattern* patterns;
    const Light* lights;
    const Gate* gates;
    const BvhNode* bvh;
    const LeafItem* leaf_items; // BVH items in leaf order
    const double* leaf_tri;     // 10 doubles per BVH item (triangles only: p1, e1, e2, pad), leaf order
    const int32_t* unbounded;   // items tested for every ray (planes, unbounded cones / cylinders, CSG holding one)
    const ConeGuard* guards;    // one per distinct linear transform of the finite cones (see ConeGuard)
    const CsgOp* csg_ops;
    const CsgRoot* csg_roots;
    int32_t n_prims, n_lights, n_gates, n_unbounded, n_bvh_items, n_csg_roots;
    int32_t bvh_root;        // link of the root (may be a leaf); valid only if n_bvh_items > 0
    int32_t all_cast_shadow; // every primitive casts shadows -> shadow queries may stop at the first hit
    int32_t max_children;    // max over materials of (reflective != 0) + (transparency != 0)
    int32_t n_guards;
    double root_lo[3], root_hi[3];  // f64 bounds of all BVH items (already padded)
    int32_t n_top;                  // nodes [0, n_top) of bvh[] are numbered breadth-first from the root (0: unknown order, nothing is staged)
    int32_t pad_top;
    int32_t n_fast_planes;          // > 0: fast_planes[] IS the always-tested list (unbounded[] is not read)
    int32_t plain_only;             // every material's pattern is Plain: shading needs no pattern evaluator
    FastPlane fast_planes[MAX_FAST_PLANES];
};

}  // namespace rtcdev
