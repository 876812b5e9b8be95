// Scene upload, host half: validates the rtc.h arrays and derives the device layout of device_scene.h
// (deduplicated transforms, SoA triangle records, group-box gates, CSG programs, padded fp32 BVH).
// Replaces what the reference does per ray — walking the Element tree (src/shape.rs:139-144,248-269) —
// with structures built once per scene.  Pure host C++ (no CUDA calls), so it is unit-testable on CPU.
#pragma once

#include <cstdint>
#include <string>
#include <vector>

#include "../../include/rtc.h"
#include "device_scene.h"

namespace rtchost {

struct HostScene {
    std::vector<rtcdev::Prim> prims;
    std::vector<rtcdev::Xform> xforms;
    std::vector<rtcdev::MatInv> matinvs;
    std::vector<double> tri_pe, tri_n, ranges;
    std::vector<rtcdev::Material> materials;
    std::vector<rtcdev::Pattern> patterns;
    std::vector<rtcdev::Light> lights;
    std::vector<rtcdev::Gate> gates;
    std::vector<rtcdev::BvhNode> bvh;
    std::vector<rtcdev::LeafItem> leaf_items;
    std::vector<double> leaf_tri;
    std::vector<float> item_box32;      // device_build: 8 floats per BVH item (lo.xyz, 0, hi.xyz, 0), padded and rounded outwards
    std::vector<int32_t> item_ids;      // device_build: primitive index (or ~csg root) per BVH item
    int32_t device_build = 0;           // the BVH is built on the device after upload (rtc_scene_desc.build_flags & RTC_BUILD_FAST):
                                        // bvh / leaf_items / leaf_tri stay empty here, the device fills its own copies
    std::vector<int32_t> unbounded;
    std::vector<rtcdev::ConeGuard> guards;
    std::vector<rtcdev::CsgOp> csg_ops;
    std::vector<rtcdev::CsgRoot> csg_roots;
    int32_t bvh_root = 0;
    int32_t n_bvh_items = 0;
    int32_t all_cast_shadow = 1;
    int32_t all_items_triangles = 1;    // every BVH item is a triangle primitive (no sphere, cube, ..., no CSG item)
    int32_t max_children = 0;
    double root_lo[3] = {0, 0, 0}, root_hi[3] = {0, 0, 0};
    // build statistics (reported by tests / DESIGN.md)
    int32_t n_gates_always_pass = 0, n_gates_conservative = 0, n_gates_exact = 0;
    int32_t bvh_depth = 0;
    int64_t device_bytes = 0;  // size of the uploaded scene arena (set by rtc_scene_create)
};

// Returns "" on success, otherwise a description of why the scene was rejected.
// tuning: rtc_tuning (builder choice, thread budget, SAH parameters); null = defaults
std::string build_host_scene(const rtc_scene_desc& desc, HostScene& out, const rtc_tuning* tuning = nullptr);

}  // namespace rtchost
