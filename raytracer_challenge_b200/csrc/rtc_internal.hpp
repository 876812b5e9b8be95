// What the translation units of librtc.so share about a scene handle (not part of the C ABI: include/rtc.h only
// declares `struct rtc_scene`).
#pragma once

#include <memory>
#include <string>

#include "device_scene.h"
#include "scene_build.hpp"

struct rtc_scene {
    int device = 0;
    int feat = 0;
    std::shared_ptr<rtchost::HostScene> host;  // shared by the handles of one rtc_scene_create_multi call
    rtcdev::Scene dev{};
    void* arena = nullptr;  // one device allocation holding every scene array
    size_t arena_bytes = 0;
    int bvh_depth = -1, bvh_nodes = -1;  // device-side build: what the device reported (host build: see HostScene)
};

namespace rtcapi {
// Sets the calling thread's rtc_last_error() message; returns 1 (the non-zero status of the C ABI).
int fail(const std::string& msg);
}  // namespace rtcapi
