// ORACLE — TEST INFRASTRUCTURE ONLY.  C entry points (ctypes) around the CPU restatement.
// Used by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
#include "oracle_api.hpp"

#include <chrono>
#include <cstring>

namespace {
void set_err(char* err, int errlen, const std::string& m) {
    if (err && errlen > 0) {
        std::strncpy(err, m.c_str(), (size_t)errlen - 1);
        err[errlen - 1] = 0;
    }
}
}  // namespace

extern "C" {

// Renders window (x0,y0,w,h) of scene `name` at camera resolution hsize x vsize with the reference
// algorithm (src/image.rs:65-81).  out_rgb: w*h*3 doubles, row-major.  counts: primary, shadow,
// reflected, refracted World::intersect queries (R_ref).  seconds: wall time of the render phase only.
int orc_render(const char* name, const char* data_dir, int hsize, int vsize, int x0, int y0, int w, int h, int fuel,
               int spp_x, int spp_y, int threads, double* out_rgb, uint64_t* counts, double* seconds, char* err, int errlen) {
    try {
        if (data_dir && *data_dir) scenes::data_dir() = data_dir;
        auto sc = scenes::build_scene<OracleApi>(name, (size_t)hsize, (size_t)vsize);
        orc::RenderOpts o;
        o.fuel = fuel;
        o.spp_x = spp_x;
        o.spp_y = spp_y;
        o.x0 = (size_t)x0; o.y0 = (size_t)y0; o.w = (size_t)w; o.h = (size_t)h;
        o.threads = (unsigned)threads;
        orc::RayCounts rc;
        auto t0 = std::chrono::steady_clock::now();
        orc::Image img = orc::Image::par_render(sc.camera, sc.world, o, &rc);
        auto t1 = std::chrono::steady_clock::now();
        if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
        for (size_t i = 0; i < img.pixels.size(); i++) {
            out_rgb[3 * i + 0] = img.pixels[i].r;
            out_rgb[3 * i + 1] = img.pixels[i].g;
            out_rgb[3 * i + 2] = img.pixels[i].b;
        }
        if (counts) { counts[0] = rc.primary; counts[1] = rc.shadow; counts[2] = rc.reflected; counts[3] = rc.refracted; }
        return 0;
    } catch (const std::exception& e) {
        set_err(err, errlen, e.what());
        return 1;
    }
}

int orc_hardware_threads(void) { return (int)std::max(1u, std::thread::hardware_concurrency()); }

// src/color.rs:42-46 quantisation of n doubles -> n bytes (the comparator's units).
void orc_quantize(const double* in, uint8_t* out, size_t n) {
    for (size_t i = 0; i < n; i++) out[i] = orc::clamp_channel(in[i]);
}

// Single-function probes for the not-gpu tests.
double orc_simplex(double x, double y, double z) { return orc::noise_detail::simplex(x, y, z); }
double orc_fractal(double x, double y, double z, int octaves) { return orc::noise_detail::fractal(x, y, z, (size_t)octaves); }

}  // extern "C"
