// Harness glue (not part of the C ABI): builds the named scenes of scenes/scenes.hpp through the product's host
// builder API (raytracer_challenge_b200/host/raytracer.hpp), flattens them and drives include/rtc.h.
// Loaded from Python (ctypes) by tests/, bench.py and __graft_entry__.py.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <sstream>
#include <memory>
#include <string>

#include "../raytracer_challenge_b200/host/raytracer.hpp"
#include "../scenes/scenes.hpp"

namespace {
struct HostApi {
    using Matrix = rtc::Matrix;
    using Vector = rtc::Vector;
    using Color = rtc::Color;
    using Noise = rtc::Noise;
    using Pattern = rtc::Pattern;
    using Material = rtc::Material;
    using ShapeArgs = rtc::ShapeArgs;
    using Element = rtc::Element;
    using GroupKind = rtc::GroupKind;
    using PointLight = rtc::PointLight;
    using World = rtc::World;
    using Camera = rtc::Camera;
    using ObjData = rtc::ObjData;
    using ObjParser = rtc::ObjParser;
};

struct Session {
    std::unique_ptr<scenes::Scene<HostApi>> built;  // the World and the Camera, as a bin's construct_world() leaves them
    std::unique_ptr<rtc::FlatScene> flat;
    rtc_camera camera{};
    rtc_scene* scene = nullptr;
};

void set_err(char* err, int errlen, const std::string& m) {
    if (err && errlen > 0) {
        std::strncpy(err, m.c_str(), (size_t)errlen - 1);
        err[errlen - 1] = 0;
    }
}
}  // namespace

extern "C" {

// Builds + flattens the scene on the host.  device >= 0: also uploads it to that GPU.  device < 0: host only
// (validation through rtc_scene_validate; no GPU touched).
int drv_open(const char* name, const char* data_dir, int hsize, int vsize, int device, void** out, int64_t* stats16, char* err, int errlen) {
    try {
        if (data_dir && *data_dir) scenes::data_dir() = data_dir;
        auto s = std::make_unique<Session>();
        s->built = std::make_unique<scenes::Scene<HostApi>>(scenes::build_scene<HostApi>(name, (size_t)hsize, (size_t)vsize));
        const auto& sc = *s->built;
        s->flat = std::make_unique<rtc::FlatScene>(sc.world);
        s->camera = sc.camera.to_c();
        const rtc_scene_desc d = s->flat->desc();
        if (device >= 0) {
            if (rtc_scene_create(&d, device, &s->scene) != 0) { set_err(err, errlen, rtc_last_error()); return 1; }
            if (stats16) for (int k = 0; k < 16; ++k) stats16[k] = rtc_scene_stat(s->scene, k);
        } else {
            if (rtc_scene_validate(&d, stats16) != 0) { set_err(err, errlen, rtc_last_error()); return 1; }
        }
        *out = s.release();
        return 0;
    } catch (const std::exception& e) {
        set_err(err, errlen, e.what());
        return 1;
    }
}

void drv_close(void* h) {
    Session* s = (Session*)h;
    if (!s) return;
    if (s->scene) rtc_scene_destroy(s->scene);
    delete s;
}

const rtc_scene_desc* drv_desc_tmp(void* h, rtc_scene_desc* storage) {
    *storage = ((Session*)h)->flat->desc();
    return storage;
}
const rtc_camera* drv_camera(void* h) { return &((Session*)h)->camera; }
rtc_scene* drv_scene(void* h) { return ((Session*)h)->scene; }

int drv_render(void* h, const rtc_opts* opts, double* out_rgb, rtc_stats* stats, char* err, int errlen) {
    Session* s = (Session*)h;
    if (!s || !s->scene) { set_err(err, errlen, "drv_render: scene was opened without a device"); return 1; }
    if (rtc_render(s->scene, &s->camera, opts, out_rgb, stats) != 0) { set_err(err, errlen, rtc_last_error()); return 1; }
    return 0;
}

int drv_render_device(void* h, const rtc_opts* opts, double** dev, uint64_t* n, rtc_stats* stats, char* err, int errlen) {
    Session* s = (Session*)h;
    if (!s || !s->scene) { set_err(err, errlen, "drv_render_device: scene was opened without a device"); return 1; }
    if (rtc_render_device(s->scene, &s->camera, opts, dev, n, stats) != 0) { set_err(err, errlen, rtc_last_error()); return 1; }
    return 0;
}

// One end-to-end step = the body of Image::par_render on an already-flattened World: upload (rtc_scene_create),
// render into the caller's host buffer (rtc_render: kernels + device->host copy), release (rtc_scene_destroy).
int drv_e2e_step(void* h, int device, const rtc_opts* opts, double* out_rgb, rtc_stats* stats, char* err, int errlen) {
    Session* s = (Session*)h;
    const rtc_scene_desc d = s->flat->desc(RTC_BUILD_FAST);  // as Image::par_render does (rtc_tuning.bvh_builder = 2 forces the host build)
    rtc_scene* scene = nullptr;
    const auto t0 = std::chrono::steady_clock::now();
    if (rtc_scene_create(&d, device, &scene) != 0) { set_err(err, errlen, rtc_last_error()); return 1; }
    const auto t1 = std::chrono::steady_clock::now();
    const int rc = rtc_render(scene, &s->camera, opts, out_rgb, stats);
    if (rc != 0) set_err(err, errlen, rtc_last_error());
    const auto t2 = std::chrono::steady_clock::now();
    rtc_scene_destroy(scene);
    const auto t3 = std::chrono::steady_clock::now();
    if (std::getenv("RTC_E2E_TRACE")) {
        auto ms = [](auto a, auto b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
        std::fprintf(stderr, "e2e step: create %.3f ms, render %.3f ms (kernels %.3f), destroy %.3f ms\n", ms(t0, t1), ms(t1, t2), stats ? stats->ms_kernels : 0.0, ms(t2, t3));
    }
    return rc;
}

// One end-to-end step of the REAL drop-in call: Image::par_render(camera, world) of the host builder API on the World
// drv_open built (mesh parsing excluded, like the bins' "Rendering image" timer): flatten the World, create the scene on
// `n_devices` GPUs starting at `first_device`, render, bring the frame to the host, destroy the scene, return the Image.
// probe3 (optional) receives the centre pixel, so that the caller can see the pixels arrived; phase_ms (optional, 2
// doubles): time spent flattening the World / in the rest of the call.
int drv_par_render_step(void* h, int first_device, int n_devices, const rtc_opts* opts, double* probe3, rtc_stats* stats, double* phase_ms, char* err, int errlen) {
    Session* s = (Session*)h;
    try {
        if (!s || !s->built) { set_err(err, errlen, "drv_par_render_step: no scene"); return 1; }
        rtc::RenderOptions ro;
        if (opts) { ro.opts.fuel = opts->fuel; ro.opts.spp_x = opts->spp_x; ro.opts.spp_y = opts->spp_y; ro.opts.flags = opts->flags & RTC_FLAG_BRUTE_FORCE; }
        for (int i = 0; i < n_devices; ++i) ro.devices.push_back(first_device + i);
        const auto t0 = std::chrono::steady_clock::now();
        rtc::Image img = rtc::Image::par_render(s->built->camera, s->built->world, ro);
        const auto t1 = std::chrono::steady_clock::now();
        if (probe3) {
            const rtc::Color c = img.read(img.hsize / 2, img.vsize / 2);
            probe3[0] = c.r; probe3[1] = c.g; probe3[2] = c.b;
        }
        if (stats) *stats = img.stats;
        if (phase_ms) { phase_ms[0] = img.flatten_ms; phase_ms[1] = std::chrono::duration<double, std::milli>(t1 - t0).count() - img.flatten_ms; }
        return 0;
    } catch (const std::exception& e) {
        set_err(err, errlen, e.what());
        return 1;
    }
}

// Host-only timing aid: milliseconds per FlatScene(world) (best of `reps`), and per host half of rtc_scene_create.
double drv_flatten_ms(void* h, int reps, double* validate_ms) {
    Session* s = (Session*)h;
    double best = 1e30, best_v = 1e30;
    for (int r = 0; r < reps; ++r) {
        const auto t0 = std::chrono::steady_clock::now();
        rtc::FlatScene flat(s->built->world);
        const auto t1 = std::chrono::steady_clock::now();
        const rtc_scene_desc d = flat.desc(RTC_BUILD_FAST);
        rtc_scene_validate(&d, nullptr);
        const auto t2 = std::chrono::steady_clock::now();
        best = std::min(best, std::chrono::duration<double, std::milli>(t1 - t0).count());
        best_v = std::min(best_v, std::chrono::duration<double, std::milli>(t2 - t1).count());
    }
    if (validate_ms) *validate_ms = best_v;
    return best;
}

// The single-process multi-GPU form (rtc_render_multi) through the host API: Image::par_render over the first
// n_devices devices (0 = all).  out_rgb: hsize*vsize*3 doubles.
int drv_par_render_multi(const char* name, const char* data_dir, int hsize, int vsize, int n_devices, double* out_rgb, rtc_stats* stats, char* err, int errlen) {
    try {
        if (data_dir && *data_dir) scenes::data_dir() = data_dir;
        auto sc = scenes::build_scene<HostApi>(name, (size_t)hsize, (size_t)vsize);
        rtc::RenderOptions ro;
        for (int i = 0; i < n_devices; ++i) ro.devices.push_back(i);
        rtc::Image img = rtc::Image::par_render(sc.camera, sc.world, ro);
        for (size_t i = 0; i < img.pixels.size(); ++i) { out_rgb[3 * i] = img.pixels[i].r; out_rgb[3 * i + 1] = img.pixels[i].g; out_rgb[3 * i + 2] = img.pixels[i].b; }
        if (stats) *stats = img.stats;
        return 0;
    } catch (const std::exception& e) {
        set_err(err, errlen, e.what());
        return 1;
    }
}

// Image::par_render + Image::ppm through the public host API (the path a bin would take): returns the PPM text length.
int64_t drv_par_render_ppm(const char* name, const char* data_dir, int hsize, int vsize, int device, char* out, int64_t out_cap, char* err, int errlen) {
    try {
        if (data_dir && *data_dir) scenes::data_dir() = data_dir;
        auto sc = scenes::build_scene<HostApi>(name, (size_t)hsize, (size_t)vsize);
        rtc::RenderOptions ro;
        if (device >= 0) ro.devices = {device};  // device < 0: shard over every visible device (= Camera::render, north_star's spelling)
        rtc::Image img = device >= 0 ? rtc::Image::par_render(sc.camera, sc.world, ro) : sc.camera.render(sc.world);
        const std::string ppm = img.ppm();
        if ((int64_t)ppm.size() <= out_cap) std::memcpy(out, ppm.data(), ppm.size());
        return (int64_t)ppm.size();
    } catch (const std::exception& e) {
        set_err(err, errlen, e.what());
        return -1;
    }
}

// ---- host-only helpers for the CPU tests of the steps either side of the path ----

// Image::ppm (kind 0) / Image::png (kind 1) of a caller-supplied f64 rgb image; returns the encoded length.
int64_t drv_image_encode(const double* rgb, int hsize, int vsize, int kind, char* out, int64_t out_cap) {
    rtc::Image img;
    img.hsize = (size_t)hsize;
    img.vsize = (size_t)vsize;
    img.pixels.resize((size_t)hsize * (size_t)vsize);
    for (size_t i = 0; i < img.pixels.size(); ++i) img.pixels[i] = rtc::Color{rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]};
    const std::string s = kind == 0 ? img.ppm() : img.png();
    if ((int64_t)s.size() <= out_cap) std::memcpy(out, s.data(), s.size());
    return (int64_t)s.size();
}

// ObjParser::parse_text on a text buffer -> a JSON description of what the reference's `parse_lines` yields
// (src/obj.rs:154-184): vertices, normals, fan-triangulated faces (1-based indices, normal indices or null, group),
// group names in first-seen order, ignored (line number, text) pairs; plus the shape of the Element `parse_obj`
// builds with the identity transform (src/obj.rs:186-258): triangles per non-empty group and whether they are wrapped.
int64_t drv_obj_parse(const char* text, int64_t len, char* out, int64_t out_cap, char* err, int errlen) {
    try {
        std::istringstream in(std::string(text, (size_t)len));
        const rtc::ObjData d = rtc::ObjParser::parse_text(in);
        std::ostringstream o;
        o.precision(17);
        auto esc = [](const std::string& s) {
            std::string r;
            for (char c : s) {
                if (c == '"' || c == '\\') { r.push_back('\\'); r.push_back(c); }
                else if ((unsigned char)c < 0x20) { char b[8]; std::snprintf(b, sizeof b, "\\u%04x", c); r += b; }
                else r.push_back(c);
            }
            return r;
        };
        o << "{\"vertices\":[";
        for (size_t i = 0; i < d.vertices.size(); ++i) o << (i ? "," : "") << "[" << d.vertices[i].x << "," << d.vertices[i].y << "," << d.vertices[i].z << "," << d.vertices[i].w << "]";
        o << "],\"normals\":[";
        for (size_t i = 0; i < d.normals.size(); ++i) o << (i ? "," : "") << "[" << d.normals[i].x << "," << d.normals[i].y << "," << d.normals[i].z << "," << d.normals[i].w << "]";
        o << "],\"triangles\":[";
        for (size_t i = 0; i < d.triangles.size(); ++i) {
            const auto& t = d.triangles[i];
            o << (i ? "," : "") << "{\"p\":[" << t.p[0] << "," << t.p[1] << "," << t.p[2] << "],\"n\":";
            if (t.has_n) o << "[" << t.n[0] << "," << t.n[1] << "," << t.n[2] << "]"; else o << "null";
            o << ",\"group\":" << t.group << "}";
        }
        o << "],\"groups\":[";
        for (size_t i = 0; i < d.groups.size(); ++i) o << (i ? "," : "") << "\"" << esc(d.groups[i]) << "\"";
        o << "],\"ignored\":[";
        for (size_t i = 0; i < d.ignored.size(); ++i) o << (i ? "," : "") << "[" << d.ignored[i].first << ",\"" << esc(d.ignored[i].second) << "\"]";
        o << "],\"element\":";
        if (d.triangles.empty()) o << "null";
        else {
            const rtc::Element e = rtc::ObjParser::build(d, rtc::Matrix::id(), rtc::Material{});
            std::function<void(const rtc::Element&)> dump = [&](const rtc::Element& el) {
                if (!el.is_group()) {
                    const rtc::Shape& s = std::get<rtc::Shape>(el.node);
                    o << "{\"kind\":" << s.kind << ",\"p\":[[" << s.p1.x << "," << s.p1.y << "," << s.p1.z << "],[" << s.p2.x << "," << s.p2.y << "," << s.p2.z << "],[" << s.p3.x << "," << s.p3.y << "," << s.p3.z << "]]}";
                    return;
                }
                const rtc::Group& g = std::get<rtc::Group>(el.node);
                o << "{\"group\":" << (int)g.kind << ",\"children\":[";
                for (size_t i = 0; i < g.children.size(); ++i) { if (i) o << ","; dump(g.children[i]); }
                o << "]}";
            };
            dump(e);
        }
        o << "}";
        const std::string s = o.str();
        if ((int64_t)s.size() <= out_cap) std::memcpy(out, s.data(), s.size());
        return (int64_t)s.size();
    } catch (const std::exception& e) {
        set_err(err, errlen, e.what());
        return -1;
    }
}

}  // extern "C"
