// Scene definitions shared by the product host API (rtc::) and the oracle (orc::).
//
// Each function is a template over an API bundle `A` exposing the reference crate's builder
// names (Matrix, Vector, Color, Material, Pattern, Noise, Element, ShapeArgs, GroupKind,
// PointLight, World, Camera, ObjData, ObjParser).  The bodies transcribe the PARAMETERS of
// the reference's example binaries' construct_world() (cited per scene) — scene data, not
// algorithm — so that the CUDA path and the oracle are fed the same scene through two
// independent builder implementations.  Resolution is an argument because every bin
// hard-codes its own.
#pragma once

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <initializer_list>
#include <optional>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace scenes {

constexpr double PI = 3.14159265358979323846264338327950288;  // std::f64::consts::PI

inline std::string& data_dir() {
    static std::string d = "data";
    return d;
}

// Binary mesh fixture written by oracle/obj2mesh.cpp (`make -C oracle fixtures`) from an OBJ file:
//   "RTCMESH1" | u32 nv, nvn, ntri, ngroups | nv*3 f64 | nvn*3 f64 |
//   ntri * (u32 p1,p2,p3,n1,n2,n3 (1-based; n = 0 when absent), u32 group) | ngroups * (u32 len, bytes)
template <class A>
typename A::ObjData load_mesh_fixture(const std::string& path) {
    std::ifstream in(path, std::ios::binary);
    if (!in) throw std::runtime_error("scenes: cannot open mesh fixture " + path);
    char magic[8];
    in.read(magic, 8);
    if (std::memcmp(magic, "RTCMESH1", 8) != 0) throw std::runtime_error("scenes: bad mesh fixture magic in " + path);
    uint32_t hdr[4];
    in.read((char*)hdr, sizeof hdr);
    typename A::ObjData d;
    std::vector<double> buf(3 * (size_t)hdr[0]);
    in.read((char*)buf.data(), buf.size() * 8);
    for (uint32_t i = 0; i < hdr[0]; i++) d.vertices.push_back(A::Vector::point(buf[3 * i], buf[3 * i + 1], buf[3 * i + 2]));
    buf.resize(3 * (size_t)hdr[1]);
    in.read((char*)buf.data(), buf.size() * 8);
    for (uint32_t i = 0; i < hdr[1]; i++) d.normals.push_back(A::Vector::vector(buf[3 * i], buf[3 * i + 1], buf[3 * i + 2]));
    for (uint32_t i = 0; i < hdr[2]; i++) {
        uint32_t r[7];
        in.read((char*)r, sizeof r);
        typename A::ObjData::Tri t;
        for (int c = 0; c < 3; c++) { t.p[c] = r[c]; t.n[c] = r[3 + c]; }
        t.has_n = r[3] && r[4] && r[5];
        t.group = r[6];
        d.triangles.push_back(t);
    }
    for (uint32_t g = 0; g < hdr[3]; g++) {
        uint32_t len;
        in.read((char*)&len, 4);
        std::string name(len, '\0');
        in.read(name.data(), len);
        d.groups.push_back(name);
    }
    if (!in) throw std::runtime_error("scenes: truncated mesh fixture " + path);
    return d;
}

template <class A>
struct Scene {
    typename A::Camera camera;
    typename A::World world;
};

// --- C1a: World::default() with the camera of the reference's only end-to-end test ---------
// src/world.rs:152-183, src/image.rs:130-137 (11x11, fov pi/2, from (0,0,-5) to origin, up +y)
template <class A>
Scene<A> default_world(size_t hsize, size_t vsize) {
    using V = typename A::Vector;
    Scene<A> s;
    s.world = A::World::default_world();
    s.camera = A::Camera::make(hsize, vsize, PI / 2.0,
                               A::Camera::transform(V::point(0.0, 0.0, -5.0), V::point(0.0, 0.0, 0.0), V::vector(0.0, 1.0, 0.0)));
    return s;
}

// --- C1b: authored "three spheres + plane floor + one light" (no chapter-7 bin exists) ------
template <class A>
Scene<A> three_spheres(size_t hsize, size_t vsize) {
    using M = typename A::Matrix; using V = typename A::Vector; using C = typename A::Color;
    using P = typename A::Pattern; using E = typename A::Element;
    Scene<A> s;
    typename A::ShapeArgs floor;
    floor.material.pattern = P::plain(C::make(1.0, 0.9, 0.9));
    floor.material.specular = 0.0;
    typename A::ShapeArgs middle;
    middle.transform = M::translation(-0.5, 1.0, 0.5);
    middle.material.pattern = P::plain(C::make(0.1, 1.0, 0.5));
    middle.material.diffuse = 0.7;
    middle.material.specular = 0.3;
    typename A::ShapeArgs right;
    right.transform = M::translation(1.5, 0.5, -0.5) * M::scaling(0.5, 0.5, 0.5);
    right.material.pattern = P::plain(C::make(0.5, 1.0, 0.1));
    right.material.diffuse = 0.7;
    right.material.specular = 0.3;
    typename A::ShapeArgs left;
    left.transform = M::translation(-1.5, 0.33, -0.75) * M::scaling(0.33, 0.33, 0.33);
    left.material.pattern = P::plain(C::make(1.0, 0.8, 0.1));
    left.material.diffuse = 0.7;
    left.material.specular = 0.3;
    s.world.lights.push_back({C::white(), V::point(-10.0, 10.0, -10.0)});
    s.world.elements.push_back(E::plane(floor));
    s.world.elements.push_back(E::sphere(middle));
    s.world.elements.push_back(E::sphere(right));
    s.world.elements.push_back(E::sphere(left));
    s.camera = A::Camera::make(hsize, vsize, PI / 3.0,
                               A::Camera::transform(V::point(0.0, 1.5, -5.0), V::point(0.0, 1.0, 0.0), V::vector(0.0, 1.0, 0.0)));
    return s;
}

// --- C2: src/bin/chapter11_glass_air_bubble.rs:15-77 (native 4096x4096) ---------------------
template <class A>
Scene<A> glass_air_bubble(size_t hsize, size_t vsize) {
    using M = typename A::Matrix; using V = typename A::Vector; using C = typename A::Color;
    using P = typename A::Pattern; using E = typename A::Element;
    Scene<A> s;
    typename A::ShapeArgs floor;
    floor.transform = M::translation(0.0, -10.0, 0.0);
    floor.material.pattern = P::checkers(M::id(), P::plain(C::white()), P::plain(C::black()));
    floor.material.specular = 0.0;
    typename A::ShapeArgs glass;
    glass.material.pattern = P::plain(C::black());
    glass.material.diffuse = 0.1;
    glass.material.shininess = 300.0;
    glass.material.reflective = 1.0;
    glass.material.transparency = 1.0;
    glass.material.refractive_index = 1.52;
    typename A::ShapeArgs air = glass;
    air.transform = M::scaling(0.5, 0.5, 0.5);
    air.material.refractive_index = 1.0;
    s.world.lights.push_back({C::make(0.7, 0.7, 0.7), V::point(20.0, 10.0, 0.0)});
    s.world.elements.push_back(E::plane(floor));
    s.world.elements.push_back(E::sphere(glass));
    s.world.elements.push_back(E::sphere(air));
    s.camera = A::Camera::make(hsize, vsize, PI / 3.0,
                               A::Camera::transform(V::point(0.0, 2.5, 0.0), V::point(0.0, 0.0, 0.0), V::vector(0.0, 0.0, 1.0)));
    return s;
}

// --- C3a: src/bin/chapter14_hexagon.rs:16-89 (native 4096x2160) -----------------------------
template <class A>
Scene<A> hexagon(size_t hsize, size_t vsize) {
    using M = typename A::Matrix; using V = typename A::Vector; using C = typename A::Color;
    using P = typename A::Pattern; using E = typename A::Element; using G = typename A::GroupKind;
    Scene<A> s;
    typename A::Material material;
    material.pattern = P::point_jitter(
        A::Noise::Simplex(0.3),
        P::stripes(M::rotation_z(PI / 2.0) * M::scaling(0.05, 0.05, 0.05), P::plain(C::white()), P::plain(C::black())));
    std::vector<E> sides;
    for (int n = 0; n < 6; n++) {
        typename A::ShapeArgs corner;
        corner.transform = M::translation(0.0, 0.0, -1.0) * M::scaling(0.25, 0.25, 0.25);
        typename A::ShapeArgs edge;
        edge.transform = M::translation(0.0, 0.0, -1.0) * M::rotation_y(-PI / 6.0) * M::rotation_z(-PI / 2.0) *
                         M::scaling(0.25, 1.0, 0.25);
        std::vector<E> side;
        side.push_back(E::sphere(corner));
        side.push_back(E::cylinder(edge, 0.0, 1.0, false));
        sides.push_back(E::composite(M::rotation_y((double)n * PI / 3.0), std::nullopt, G::Aggregation, std::move(side)));
    }
    s.world.lights.push_back({C::white(), V::point(1.0, 6.9, -4.9)});
    s.world.elements.push_back(E::composite(M::id(), material, G::Aggregation, std::move(sides)));
    s.camera = A::Camera::make(hsize, vsize, 0.314,
                               A::Camera::transform(V::point(8.0, 5.0, 8.0), V::point(0.0, 0.0, 0.0), V::vector(0.0, 1.0, 0.0)));
    return s;
}

// --- C3b: src/bin/chapter14_benchmark.rs:15-109 (native 4096x2160) --------------------------
template <class A>
Scene<A> benchmark_with(size_t hsize, size_t vsize, bool glass);
template <class A>
Scene<A> benchmark(size_t hsize, size_t vsize) { return benchmark_with<A>(hsize, vsize, false); }
// The same 1 000 spheres made of glass: a line through a row of the lattice crosses up to 2 x 5 x 2 x ... spheres — dozens to
// hundreds of crossings per n1 / n2 replay, beyond the 64 the on-chip list holds (device_scene.h: BIG_CROSSINGS).
template <class A>
Scene<A> benchmark_glass(size_t hsize, size_t vsize) { return benchmark_with<A>(hsize, vsize, true); }
template <class A>
Scene<A> benchmark_with(size_t hsize, size_t vsize, bool glass) {
    using M = typename A::Matrix; using V = typename A::Vector; using C = typename A::Color;
    using P = typename A::Pattern; using E = typename A::Element; using G = typename A::GroupKind;
    Scene<A> s;
    auto sphere_block = [](size_t n, double h) {
        std::vector<E> elements;
        for (size_t x = 0; x < n; x++)
            for (size_t z = 0; z < n; z++)
                for (size_t y = 0; y < n; y++) {
                    typename A::ShapeArgs a;
                    a.transform = M::translation(0.1 + (double)x * h, 0.1 + (double)y * h, 0.1 + (double)z * h) *
                                  M::scaling(0.1, 0.1, 0.1);
                    a.material.pattern = P::plain(C::black());
                    elements.push_back(E::sphere(a));
                }
        return elements;
    };
    typename A::ShapeArgs backdrop;
    backdrop.material.pattern = P::plain(C::white());
    backdrop.material.specular = 0.0;
    const size_t n = 5;
    const double h = 0.3;
    std::vector<E> spheres;
    const double offset = (double)(n - 1) * (0.1 + h) - 0.1;
    for (int x = 0; x < 2; x++)
        for (int z = 0; z < 2; z++)
            for (int y = 0; y < 2; y++)
                spheres.push_back(E::composite(M::translation((double)x * offset, (double)y * offset, (double)z * offset),
                                               std::nullopt, G::Aggregation, sphere_block(n, h)));
    typename A::Material material;
    material.pattern = P::gradient(M::scaling(1.0, 2.9, 1.0) * M::rotation_z(PI / 2.0), P::plain(C::make(1.0, 0.0, 0.0)),
                                   P::plain(C::make(0.0, 0.0, 1.0)));
    if (glass) { material.transparency = 0.8; material.refractive_index = 1.0; material.diffuse = 0.3; }  // index 1: rays go straight through the rows
    s.world.lights.push_back({C::white(), V::point(-5.0, 7.0, -1.0)});
    s.world.elements.push_back(E::plane(backdrop));
    s.world.elements.push_back(E::composite(M::id(), material, G::Aggregation, std::move(spheres)));
    s.camera = A::Camera::make(hsize, vsize, 1.0,
                               A::Camera::transform(V::point(-8.0, 8.0, -8.0), V::point(2.0, 2.0, 2.0), V::vector(0.0, 1.0, 0.0)));
    return s;
}

// --- C4: src/bin/chapter15_teapot.rs:16-92 (native 4096x2160; mesh = obj/teapot_high.obj) ---
template <class A>
Scene<A> teapot_from(const typename A::ObjData& mesh, size_t hsize, size_t vsize) {
    using M = typename A::Matrix; using V = typename A::Vector; using C = typename A::Color;
    using P = typename A::Pattern; using E = typename A::Element;
    Scene<A> s;
    typename A::Material pot;
    pot.pattern = P::plain(C::make(0.7, 0.7, 1.0));
    pot.ambient = 0.1;
    pot.diffuse = 0.6;
    pot.specular = 0.4;
    pot.reflective = 0.1;
    pot.shininess = 5.0;
    E teapot = A::ObjParser::build(mesh, M::rotation_x(-PI / 2.0), pot);
    typename A::Material material;
    material.pattern = P::plain(C::black());
    material.ambient = 0.02;
    material.diffuse = 0.7;
    material.specular = 0.0;
    material.reflective = 0.5;
    typename A::ShapeArgs floor;
    floor.material = material;
    typename A::ShapeArgs left_wall;
    left_wall.material = material;
    left_wall.transform = M::rotation_y(-PI / 4.0) * M::translation(0.0, 0.0, 30.0) * M::rotation_x(PI / 2.0);
    typename A::ShapeArgs right_wall;
    right_wall.material = material;
    right_wall.transform = M::rotation_y(PI / 4.0) * M::translation(0.0, 0.0, 30.0) * M::rotation_x(PI / 2.0);
    s.world.elements.push_back(E::plane(floor));
    s.world.elements.push_back(E::plane(left_wall));
    s.world.elements.push_back(E::plane(right_wall));
    s.world.elements.push_back(std::move(teapot));
    s.world.lights.push_back({C::make(0.7, 0.7, 0.7), V::point(-100.0, 100.0, -100.0)});
    s.world.lights.push_back({C::make(0.7, 0.7, 0.7), V::point(100.0, 100.0, -100.0)});
    s.camera = A::Camera::make(hsize, vsize, 1.4,
                               A::Camera::transform(V::point(0.0, 30.0, -50.0), V::point(0.0, 0.0, 10.0), V::vector(0.0, 1.0, 0.0)));
    return s;
}
template <class A>
Scene<A> teapot(size_t hsize, size_t vsize) {
    return teapot_from<A>(load_mesh_fixture<A>(data_dir() + "/teapot_high.mesh"), hsize, vsize);
}
template <class A>
Scene<A> teapot_low(size_t hsize, size_t vsize) {
    return teapot_from<A>(load_mesh_fixture<A>(data_dir() + "/teapot_low.mesh"), hsize, vsize);
}

// --- C5: authored CSG + Simplex/fractal-noise scene (no such bin exists) --------
// Nested CSG (difference of (cube ∩ sphere) and a union of three cylinders), a glass CSG lens,
// a cone∪sphere, every Mixture kind and both Jitter kinds with Noise::Simplex and Noise::Fractal.
template <class A>
Scene<A> csg_noise(size_t hsize, size_t vsize) {
    using M = typename A::Matrix; using V = typename A::Vector; using C = typename A::Color;
    using P = typename A::Pattern; using E = typename A::Element; using G = typename A::GroupKind;
    using N = typename A::Noise;
    Scene<A> s;

    // floor: checkers of (jittered rings, fractal-jittered gradient)
    typename A::ShapeArgs floor;
    floor.material.pattern = P::checkers(
        M::scaling(2.0, 2.0, 2.0),
        P::point_jitter(N::Simplex(0.15), P::ring(M::scaling(0.2, 0.2, 0.2), P::plain(C::make(0.9, 0.85, 0.7)), P::plain(C::make(0.3, 0.25, 0.2)))),
        P::point_jitter(N::Fractal(0.4, 4), P::gradient(M::rotation_y(0.5), P::plain(C::make(0.2, 0.3, 0.6)), P::plain(C::make(0.8, 0.9, 1.0)))));
    floor.material.specular = 0.1;
    floor.material.reflective = 0.15;

    // back wall: blend of stripes and ring-gradient, colour-jittered
    typename A::ShapeArgs wall;
    wall.transform = M::translation(0.0, 0.0, 9.0) * M::rotation_x(PI / 2.0);
    wall.material.pattern = P::color_jitter(
        N::Fractal(0.08, 3),
        P::blend(M::id(), P::stripes(M::scaling(0.5, 0.5, 0.5) * M::rotation_y(0.3), P::plain(C::make(0.7, 0.2, 0.2)), P::plain(C::make(0.9, 0.8, 0.3))),
                 P::ring_gradient(M::scaling(1.5, 1.5, 1.5), P::plain(C::make(0.1, 0.4, 0.3)), P::plain(C::make(0.9, 0.9, 0.9)))));
    wall.material.specular = 0.0;

    // classic CSG: (cube ∩ sphere) − (cylX ∪ cylY ∪ cylZ)
    typename A::ShapeArgs cube;
    cube.material.pattern = P::color_jitter(N::Simplex(0.1), P::stripes(M::scaling(0.25, 0.25, 0.25), P::plain(C::make(0.85, 0.3, 0.2)), P::plain(C::make(0.95, 0.8, 0.6))));
    cube.material.reflective = 0.2;
    typename A::ShapeArgs ball;
    ball.transform = M::scaling(1.35, 1.35, 1.35);
    ball.material.pattern = P::plain(C::make(0.2, 0.3, 0.8));
    ball.material.shininess = 50.0;
    auto cyl = [](const typename A::Matrix& t, const typename A::Color& c) {
        typename A::ShapeArgs a;
        a.transform = t * M::scaling(0.6, 1.0, 0.6);
        a.material.pattern = P::plain(c);
        return E::cylinder(a, -1.5, 1.5, true);
    };
    std::vector<E> cxy;
    cxy.push_back(cyl(M::rotation_z(PI / 2.0), C::make(0.9, 0.1, 0.1)));
    cxy.push_back(cyl(M::id(), C::make(0.1, 0.9, 0.1)));
    std::vector<E> cxyz;
    cxyz.push_back(E::composite(M::id(), std::nullopt, G::Union, std::move(cxy)));
    cxyz.push_back(cyl(M::rotation_x(PI / 2.0), C::make(0.1, 0.1, 0.9)));
    std::vector<E> body;
    body.push_back(E::cube(cube));
    body.push_back(E::sphere(ball));
    std::vector<E> diff;
    diff.push_back(E::composite(M::id(), std::nullopt, G::Intersection, std::move(body)));
    diff.push_back(E::composite(M::id(), std::nullopt, G::Union, std::move(cxyz)));
    E classic = E::composite(M::translation(-2.2, 1.35, 0.5) * M::rotation_y(0.6) * M::rotation_x(-0.3), std::nullopt, G::Difference, std::move(diff));

    // glass lens: intersection of two offset spheres (transparent + reflective -> Schlick + containers)
    typename A::Material glass;
    glass.pattern = P::plain(C::make(0.05, 0.05, 0.08));
    glass.diffuse = 0.1;
    glass.ambient = 0.05;
    glass.shininess = 300.0;
    glass.reflective = 0.5;
    glass.transparency = 0.5;
    glass.refractive_index = 1.52;
    typename A::ShapeArgs l1;
    l1.transform = M::translation(0.0, 0.0, -0.6);
    typename A::ShapeArgs l2;
    l2.transform = M::translation(0.0, 0.0, 0.6);
    std::vector<E> lens_parts;
    lens_parts.push_back(E::sphere(l1));
    lens_parts.push_back(E::sphere(l2));
    E lens = E::composite(M::translation(1.0, 1.3, -2.0) * M::rotation_y(-0.4) * M::scaling(1.3, 1.3, 1.3), glass, G::Intersection, std::move(lens_parts));

    // cone ∪ sphere inside a plain group with its own transform (CSG under Aggregation)
    typename A::ShapeArgs cone;
    cone.transform = M::translation(0.0, 1.0, 0.0);
    cone.material.pattern = P::point_jitter(N::Fractal(0.25, 5), P::checkers(M::scaling(0.3, 0.3, 0.3), P::plain(C::make(0.9, 0.6, 0.1)), P::plain(C::make(0.2, 0.1, 0.4))));
    typename A::ShapeArgs top;
    top.transform = M::translation(0.0, 1.2, 0.0) * M::scaling(0.55, 0.55, 0.55);
    top.material.pattern = P::ring_gradient(M::scaling(0.2, 0.2, 0.2), P::plain(C::make(0.9, 0.9, 0.2)), P::plain(C::make(0.1, 0.5, 0.5)));
    top.material.reflective = 0.3;
    std::vector<E> cs;
    cs.push_back(E::cone(cone, -1.0, 0.0, true));
    cs.push_back(E::sphere(top));
    std::vector<E> grp;
    grp.push_back(E::composite(M::id(), std::nullopt, G::Union, std::move(cs)));
    typename A::ShapeArgs pedestal;
    pedestal.transform = M::translation(0.0, -0.25, 0.0) * M::scaling(0.9, 0.25, 0.9);
    pedestal.material.pattern = P::plain(C::make(0.6, 0.6, 0.65));
    grp.push_back(E::cube(pedestal));
    E tower = E::composite(M::translation(3.4, 0.5, 1.5) * M::rotation_y(0.8) * M::scaling(1.2, 1.2, 1.2), std::nullopt, G::Aggregation, std::move(grp));

    s.world.lights.push_back({C::make(0.8, 0.8, 0.8), V::point(-6.0, 9.0, -8.0)});
    s.world.lights.push_back({C::make(0.35, 0.35, 0.4), V::point(7.0, 5.0, -6.0)});
    s.world.elements.push_back(E::plane(floor));
    s.world.elements.push_back(E::plane(wall));
    s.world.elements.push_back(std::move(classic));
    s.world.elements.push_back(std::move(lens));
    s.world.elements.push_back(std::move(tower));
    s.camera = A::Camera::make(hsize, vsize, 0.9,
                               A::Camera::transform(V::point(0.5, 3.2, -9.0), V::point(0.3, 1.0, 0.0), V::vector(0.0, 1.0, 0.0)));
    return s;
}

// --- fuzz: seeded random scenes covering every primitive / pattern / group kind --------------
struct Rng {  // splitmix64 — identical stream for both API instantiations
    uint64_t s;
    explicit Rng(uint64_t seed) : s(seed) {}
    uint64_t next() {
        uint64_t z = (s += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    double uni() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }
    double range(double a, double b) { return a + (b - a) * uni(); }
    int below(int n) { return (int)(next() % (uint64_t)n); }
    bool chance(double p) { return uni() < p; }
    // three draws in a guaranteed order (function-argument evaluation order is unspecified in C++)
    struct D3 { double a, b, c; };
    D3 d3(double a0, double a1, double b0, double b1, double c0, double c1) {
        D3 d;
        d.a = range(a0, a1);
        d.b = range(b0, b1);
        d.c = range(c0, c1);
        return d;
    }
};

template <class A>
typename A::Color fuzz_color(Rng& r) {
    Rng::D3 d = r.d3(0.05, 1.0, 0.05, 1.0, 0.05, 1.0);
    return A::Color::make(d.a, d.b, d.c);
}

template <class A>
typename A::Matrix fuzz_small_transform(Rng& r) {
    using M = typename A::Matrix;
    Rng::D3 sc = r.d3(0.15, 1.5, 0.15, 1.5, 0.15, 1.5);
    M m = M::scaling(sc.a, sc.b, sc.c);
    if (r.chance(0.5)) m = M::rotation_y(r.range(-3.0, 3.0)) * m;
    if (r.chance(0.3)) m = M::rotation_z(r.range(-1.0, 1.0)) * m;
    if (r.chance(0.2)) {
        Rng::D3 sh = r.d3(-0.3, 0.3, -0.3, 0.3, 0.0, 0.0);
        m = M::shearing(sh.a, 0.0, 0.0, sh.b, 0.0, 0.0) * m;
    }
    return m;
}

template <class A>
typename A::Pattern fuzz_pattern(Rng& r, int depth) {
    using P = typename A::Pattern; using N = typename A::Noise;
    if (depth <= 0 || r.chance(0.35)) return P::plain(fuzz_color<A>(r));
    int k = r.below(8);
    if (k < 6) {
        typename A::Matrix t = fuzz_small_transform<A>(r);
        P l = fuzz_pattern<A>(r, depth - 1);
        P rr = fuzz_pattern<A>(r, depth - 1);
        switch (k) {
            case 0: return P::blend(t, l, rr);
            case 1: return P::checkers(t, l, rr);
            case 2: return P::ring_gradient(t, l, rr);
            case 3: return P::ring(t, l, rr);
            case 4: return P::gradient(t, l, rr);
            default: return P::stripes(t, l, rr);
        }
    }
    bool simplex = r.chance(0.5);
    double nscale = r.range(0.05, 0.5);
    size_t octaves = (size_t)(1 + r.below(4));
    typename A::Noise n = simplex ? N::Simplex(nscale) : N::Fractal(nscale, octaves);
    if (k == 6) return P::color_jitter(n, fuzz_pattern<A>(r, depth - 1));
    return P::point_jitter(n, fuzz_pattern<A>(r, depth - 1));
}

template <class A>
typename A::Material fuzz_material(Rng& r) {
    typename A::Material m;
    m.pattern = fuzz_pattern<A>(r, 3);
    m.ambient = r.range(0.0, 0.3);
    m.diffuse = r.range(0.2, 0.9);
    m.specular = r.chance(0.3) ? 0.0 : r.range(0.1, 0.9);
    m.shininess = r.range(5.0, 300.0);
    if (r.chance(0.35)) m.reflective = r.range(0.1, 1.0);
    if (r.chance(0.25)) {
        m.transparency = r.range(0.2, 1.0);
        m.refractive_index = r.range(1.0, 2.4);
    }
    return m;
}

template <class A>
typename A::Element fuzz_primitive(Rng& r, double spread) {
    using M = typename A::Matrix; using V = typename A::Vector; using E = typename A::Element;
    typename A::ShapeArgs a;
    Rng::D3 tr = r.d3(-spread, spread, 0.0, spread, -spread, spread);
    typename A::Matrix local = fuzz_small_transform<A>(r);
    a.transform = M::translation(tr.a, tr.b, tr.c) * local;
    a.material = fuzz_material<A>(r);
    a.casts_shadow = !r.chance(0.1);
    switch (r.below(7)) {
        case 0: return E::sphere(a);
        case 1: return E::cube(a);
        case 2: { double lo = r.range(-1.5, 0.0); double len = r.range(0.3, 2.0); bool closed = r.chance(0.6); return E::cylinder(a, lo, lo + len, closed); }
        case 3: { double lo = r.range(-1.2, 0.0); double len = r.range(0.3, 1.5); bool closed = r.chance(0.6); return E::cone(a, lo, lo + len, closed); }
        case 4: {
            Rng::D3 q1 = r.d3(-1, 1, -1, 1, -1, 1), q2 = r.d3(-1, 1, -1, 1, -1, 1), q3 = r.d3(-1, 1, -1, 1, -1, 1);
            V p1 = V::point(q1.a, q1.b, q1.c);
            V p2 = V::point(q2.a, q2.b, q2.c);
            V p3 = V::point(q3.a, q3.b, q3.c);
            return E::triangle(a, p1, p2, p3);
        }
        case 5: {
            Rng::D3 q1 = r.d3(-1, 1, -1, 1, -1, 1), q2 = r.d3(-1, 1, -1, 1, -1, 1), q3 = r.d3(-1, 1, -1, 1, -1, 1);
            V p1 = V::point(q1.a, q1.b, q1.c);
            V p2 = V::point(q2.a, q2.b, q2.c);
            V p3 = V::point(q3.a, q3.b, q3.c);
            Rng::D3 m1 = r.d3(-1, 1, 0.2, 1, -1, 1), m2 = r.d3(-1, 1, 0.2, 1, -1, 1), m3 = r.d3(-1, 1, 0.2, 1, -1, 1);
            V n1 = V::vector(m1.a, m1.b, m1.c).normalize();
            V n2 = V::vector(m2.a, m2.b, m2.c).normalize();
            V n3 = V::vector(m3.a, m3.b, m3.c).normalize();
            return E::smooth_triangle(a, p1, p2, p3, n1, n2, n3);
        }
        default: return E::sphere(a);
    }
}

template <class A>
typename A::Element fuzz_element(Rng& r, int depth, double spread) {
    using M = typename A::Matrix; using E = typename A::Element; using G = typename A::GroupKind;
    if (depth <= 0 || r.chance(0.55)) return fuzz_primitive<A>(r, spread);
    Rng::D3 gt = r.d3(-1.0, 1.0, 0.0, 1.0, -1.0, 1.0);
    double gr = r.range(-3.0, 3.0);
    M t = M::translation(gt.a, gt.b, gt.c) * M::rotation_y(gr);
    std::optional<typename A::Material> mat;
    if (r.chance(0.3)) mat = fuzz_material<A>(r);
    if (r.chance(0.4)) {  // CSG: exactly two children
        std::vector<E> c;
        c.push_back(fuzz_element<A>(r, depth - 1, spread * 0.5));
        c.push_back(fuzz_element<A>(r, depth - 1, spread * 0.5));
        G kinds[3] = {G::Union, G::Intersection, G::Difference};
        int which = r.below(3);
        return E::composite(t, mat, kinds[which], std::move(c));
    }
    std::vector<E> c;
    int n = 1 + r.below(4);
    for (int i = 0; i < n; i++) c.push_back(fuzz_element<A>(r, depth - 1, spread * 0.7));
    return E::composite(t, mat, G::Aggregation, std::move(c));
}

template <class A>
Scene<A> fuzz(uint64_t seed, size_t hsize, size_t vsize) {
    using M = typename A::Matrix; using V = typename A::Vector; using E = typename A::Element;
    Rng r(seed * 0x2545F4914F6CDD1Dull + 12345);
    Scene<A> s;
    int nl = 1 + r.below(3);
    for (int i = 0; i < nl; i++) {
        Rng::D3 lc = r.d3(0.3, 1.0, 0.3, 1.0, 0.3, 1.0);
        Rng::D3 lp = r.d3(-10, 10, 4, 12, -10, -2);
        s.world.lights.push_back({A::Color::make(lc.a, lc.b, lc.c), V::point(lp.a, lp.b, lp.c)});
    }
    if (r.chance(0.8)) {
        typename A::ShapeArgs floor;
        floor.material = fuzz_material<A>(r);
        floor.material.transparency = 0.0;
        if (r.chance(0.3)) floor.transform = M::translation(0.0, -0.5, 0.0) * M::rotation_x(r.range(-0.1, 0.1));
        s.world.elements.push_back(E::plane(floor));
    }
    int n = 2 + r.below(6);
    for (int i = 0; i < n; i++) s.world.elements.push_back(fuzz_element<A>(r, 3, 3.0));
    double fov = r.range(0.6, 1.3);
    Rng::D3 from = r.d3(-3, 3, 1.5, 5, -9, -6);
    Rng::D3 to = r.d3(-0.5, 0.5, 0.5, 1.5, 0.0, 0.0);
    s.camera = A::Camera::make(hsize, vsize, fov,
                               A::Camera::transform(V::point(from.a, from.b, from.c), V::point(to.a, to.b, 0.0), V::vector(0.0, 1.0, 0.0)));
    return s;
}

// --- §8(f)-4: the remaining example binaries (title images and the cover) ----------------------
// `Material { pattern, <fields>, ..Material::default() }` of the bins as one expression.
// keys: a ambient, d diffuse, s specular, h shininess, r reflective, t transparency, i refractive_index
template <class A>
typename A::Material material_with(const typename A::Pattern& pattern, std::initializer_list<std::pair<char, double>> fields) {
    typename A::Material m;
    m.pattern = pattern;
    for (const auto& f : fields) {
        switch (f.first) {
            case 'a': m.ambient = f.second; break;
            case 'd': m.diffuse = f.second; break;
            case 's': m.specular = f.second; break;
            case 'h': m.shininess = f.second; break;
            case 'r': m.reflective = f.second; break;
            case 't': m.transparency = f.second; break;
            case 'i': m.refractive_index = f.second; break;
            default: throw std::runtime_error("scenes: unknown material field key");
        }
    }
    return m;
}
template <class A>
typename A::ShapeArgs shape_args(const typename A::Matrix& transform, const typename A::Material& material) {
    typename A::ShapeArgs a;
    a.transform = transform;
    a.material = material;
    return a;
}

// src/bin/chapter11_title.rs:15-198 (native 4096x2160): six planes, four matte + one shiny + two glass spheres
template <class A>
Scene<A> chapter11_title(size_t hsize, size_t vsize) {
    using M = typename A::Matrix; using V = typename A::Vector; using C = typename A::Color;
    using P = typename A::Pattern; using E = typename A::Element;
    Scene<A> s;
    auto plain = [](double r, double g, double b) { return P::plain(C::make(r, g, b)); };
    const auto wall_material = material_with<A>(
        P::stripes(M::scaling(0.25, 0.25, 0.25) * M::rotation_y(PI / 2.0), plain(0.45, 0.45, 0.45), plain(0.55, 0.55, 0.55)),
        {{'a', 0.0}, {'d', 0.4}, {'s', 0.0}, {'r', 0.3}});
    auto& el = s.world.elements;
    el.push_back(E::plane(shape_args<A>(M::rotation_y(0.31415),
                                        material_with<A>(P::checkers(M::id(), plain(0.35, 0.35, 0.35), plain(0.65, 0.65, 0.65)), {{'s', 0.0}, {'r', 0.4}}))));
    el.push_back(E::plane(shape_args<A>(M::translation(0.0, 5.0, 0.0), material_with<A>(plain(0.8, 0.8, 0.8), {{'a', 0.3}, {'s', 0.0}}))));
    el.push_back(E::plane(shape_args<A>(M::translation(-5.0, 0.0, 0.0) * M::rotation_z(PI / 2.0) * M::rotation_y(PI / 2.0), wall_material)));
    el.push_back(E::plane(shape_args<A>(M::translation(5.0, 0.0, 0.0) * M::rotation_z(PI / 2.0) * M::rotation_y(PI / 2.0), wall_material)));
    el.push_back(E::plane(shape_args<A>(M::translation(0.0, 0.0, 5.0) * M::rotation_x(PI / 2.0), wall_material)));
    el.push_back(E::plane(shape_args<A>(M::translation(0.0, 0.0, -5.0) * M::rotation_x(PI / 2.0), wall_material)));
    el.push_back(E::sphere(shape_args<A>(M::translation(4.6, 0.4, 1.0) * M::scaling(0.4, 0.4, 0.4), material_with<A>(plain(0.8, 0.5, 0.3), {{'h', 50.0}}))));
    el.push_back(E::sphere(shape_args<A>(M::translation(4.7, 0.3, 0.4) * M::scaling(0.3, 0.3, 0.3), material_with<A>(plain(0.9, 0.4, 0.5), {{'h', 50.0}}))));
    el.push_back(E::sphere(shape_args<A>(M::translation(-1.0, 0.5, 4.5) * M::scaling(0.5, 0.5, 0.5), material_with<A>(plain(0.4, 0.9, 0.6), {{'h', 50.0}}))));
    el.push_back(E::sphere(shape_args<A>(M::translation(-1.7, 0.3, 4.7) * M::scaling(0.3, 0.3, 0.3), material_with<A>(plain(0.4, 0.6, 0.9), {{'h', 50.0}}))));
    el.push_back(E::sphere(shape_args<A>(M::translation(-0.6, 1.0, 0.6), material_with<A>(plain(1.0, 0.3, 0.2), {{'s', 0.4}, {'h', 5.0}}))));
    el.push_back(E::sphere(shape_args<A>(M::translation(0.6, 0.7, -0.6) * M::scaling(0.7, 0.7, 0.7),
                                         material_with<A>(plain(0.0, 0.0, 0.2), {{'a', 0.0}, {'d', 0.4}, {'s', 0.9}, {'h', 300.0}, {'r', 0.9}, {'t', 0.9}, {'i', 1.5}}))));
    el.push_back(E::sphere(shape_args<A>(M::translation(-0.7, 0.5, -0.8) * M::scaling(0.5, 0.5, 0.5),
                                         material_with<A>(plain(0.0, 0.2, 0.0), {{'a', 0.0}, {'d', 0.4}, {'s', 0.9}, {'h', 300.0}, {'r', 0.9}, {'t', 0.9}, {'i', 1.5}}))));
    s.world.lights.push_back({C::white(), V::point(-4.9, 4.9, -1.0)});
    s.camera = A::Camera::make(hsize, vsize, 1.152,
                               A::Camera::transform(V::point(-2.6, 1.5, -3.9), V::point(-0.6, 1.0, -0.8), V::vector(0.0, 1.0, 0.0)));
    return s;
}

// src/bin/chapter12_title.rs:15-281 (native 4096x2160): a room made of two huge cubes, a table, a glass cube, a mirror
template <class A>
Scene<A> chapter12_title(size_t hsize, size_t vsize) {
    using M = typename A::Matrix; using V = typename A::Vector; using C = typename A::Color;
    using P = typename A::Pattern; using E = typename A::Element;
    Scene<A> s;
    auto plain = [](double r, double g, double b) { return P::plain(C::make(r, g, b)); };
    auto& el = s.world.elements;
    el.push_back(E::cube(shape_args<A>(M::scaling(20.0, 7.0, 20.0) * M::translation(0.0, 1.0, 0.0),
                                       material_with<A>(P::checkers(M::scaling(0.07, 0.07, 0.07), P::plain(C::black()), plain(0.25, 0.25, 0.25)),
                                                        {{'a', 0.25}, {'d', 0.7}, {'s', 0.9}, {'h', 300.0}, {'r', 0.1}}))));
    el.push_back(E::cube(shape_args<A>(M::scaling(10.0, 10.0, 10.0),
                                       material_with<A>(P::checkers(M::scaling(0.05, 20.0, 0.05), plain(0.4863, 0.3765, 0.2941), plain(0.3725, 0.2902, 0.2275)),
                                                        {{'a', 0.1}, {'d', 0.7}, {'s', 0.9}, {'h', 300.0}, {'r', 0.1}}))));
    el.push_back(E::cube(shape_args<A>(M::translation(0.0, 3.1, 0.0) * M::scaling(3.0, 0.1, 2.0),
                                       material_with<A>(P::stripes(M::rotation_y(0.1) * M::scaling(0.05, 0.05, 0.05), plain(0.5529, 0.4235, 0.3255), plain(0.6588, 0.5098, 0.4000)),
                                                        {{'a', 0.1}, {'d', 0.7}, {'s', 0.9}, {'h', 300.0}, {'r', 0.2}}))));
    const auto leg_material = material_with<A>(plain(0.5529, 0.4235, 0.3255), {{'a', 0.2}, {'d', 0.7}});
    const double legs[4][2] = {{2.7, -1.7}, {2.7, 1.7}, {-2.7, -1.7}, {-2.7, 1.7}};
    for (const auto& l : legs) el.push_back(E::cube(shape_args<A>(M::translation(l[0], 1.5, l[1]) * M::scaling(0.1, 1.5, 0.1), leg_material)));
    el.push_back(E::cube(shape_args<A>(M::translation(0.0, 3.45001, 0.0) * M::rotation_y(0.2) * M::scaling(0.25, 0.25, 0.25),
                                       material_with<A>(plain(1.0, 1.0, 0.8), {{'a', 0.0}, {'d', 0.3}, {'s', 0.9}, {'h', 300.0}, {'r', 0.7}, {'t', 0.7}, {'i', 1.5}}))));
    el.push_back(E::cube(shape_args<A>(M::translation(1.0, 3.35, -0.9) * M::rotation_y(-0.4) * M::scaling(0.15, 0.15, 0.15),
                                       material_with<A>(plain(1.0, 0.5, 0.5), {{'r', 0.6}, {'d', 0.4}}))));
    el.push_back(E::cube(shape_args<A>(M::translation(-1.5, 3.27, 0.3) * M::rotation_y(0.4) * M::scaling(0.15, 0.07, 0.15), material_with<A>(plain(1.0, 1.0, 0.5), {}))));
    el.push_back(E::cube(shape_args<A>(M::translation(0.0, 3.25, 1.0) * M::rotation_y(0.4) * M::scaling(0.2, 0.05, 0.05), material_with<A>(plain(0.5, 1.0, 0.5), {}))));
    el.push_back(E::cube(shape_args<A>(M::translation(-0.6, 3.4, -1.0) * M::rotation_y(0.8) * M::scaling(0.05, 0.2, 0.05), material_with<A>(plain(0.5, 0.5, 1.0), {}))));
    el.push_back(E::cube(shape_args<A>(M::translation(2.0, 3.4, 1.0) * M::rotation_y(0.8) * M::scaling(0.05, 0.2, 0.05), material_with<A>(plain(0.5, 1.0, 1.0), {}))));
    el.push_back(E::cube(shape_args<A>(M::translation(-10.0, 4.0, 1.0) * M::scaling(0.05, 1.0, 1.0), material_with<A>(plain(0.7098, 0.2471, 0.2196), {{'d', 0.6}}))));
    el.push_back(E::cube(shape_args<A>(M::translation(-10.0, 3.4, 2.7) * M::scaling(0.05, 0.4, 0.4), material_with<A>(plain(0.2667, 0.2706, 0.6902), {{'d', 0.6}}))));
    el.push_back(E::cube(shape_args<A>(M::translation(-10.0, 4.6, 2.7) * M::scaling(0.05, 0.4, 0.4), material_with<A>(plain(0.3098, 0.5961, 0.3098), {{'d', 0.6}}))));
    el.push_back(E::cube(shape_args<A>(M::translation(-2.0, 3.5, 9.95) * M::scaling(5.0, 1.5, 0.05), material_with<A>(plain(0.3882, 0.2627, 0.1882), {{'d', 0.6}}))));
    el.push_back(E::cube(shape_args<A>(M::translation(-2.0, 3.5, 9.95) * M::scaling(4.8, 1.4, 0.06),
                                       material_with<A>(P::plain(C::black()), {{'a', 0.0}, {'d', 0.0}, {'s', 1.0}, {'h', 300.0}, {'r', 1.0}}))));
    s.world.lights.push_back({C::make(1.0, 1.0, 0.9), V::point(0.0, 6.9, -5.0)});
    s.camera = A::Camera::make(hsize, vsize, 0.785,
                               A::Camera::transform(V::point(8.0, 6.0, -8.0), V::point(0.0, 3.0, 0.0), V::vector(0.0, 1.0, 0.0)));
    return s;
}

// src/bin/chapter13_title.rs:14-248 (native 4096x2160): open / closed / concentric cylinders and a glass one
template <class A>
Scene<A> chapter13_title(size_t hsize, size_t vsize) {
    using M = typename A::Matrix; using V = typename A::Vector; using C = typename A::Color;
    using P = typename A::Pattern; using E = typename A::Element;
    Scene<A> s;
    auto plain = [](double r, double g, double b) { return P::plain(C::make(r, g, b)); };
    auto& el = s.world.elements;
    el.push_back(E::plane(shape_args<A>(M::id(), material_with<A>(P::checkers(M::rotation_y(0.3) * M::scaling(0.25, 0.25, 0.25), plain(0.5, 0.5, 0.5), plain(0.75, 0.75, 0.75)),
                                                                  {{'a', 0.2}, {'d', 0.9}, {'s', 0.0}}))));
    el.push_back(E::cylinder(shape_args<A>(M::translation(-1.0, 0.0, 1.0) * M::scaling(0.5, 1.0, 0.5),
                                           material_with<A>(plain(0.0, 0.0, 0.6), {{'d', 0.1}, {'s', 0.9}, {'h', 300.0}, {'r', 0.9}})),
                             0.0, 0.75, true));
    const double concentric[4][5] = {{0.8, 1.0, 1.0, 0.3, 0.2}, {0.6, 1.0, 0.9, 0.4, 0.3}, {0.4, 1.0, 0.8, 0.5, 0.4}, {0.2, 1.0, 0.7, 0.6, 0.5}};
    for (int k = 0; k < 4; ++k) {
        const auto& c = concentric[k];
        el.push_back(E::cylinder(shape_args<A>(M::translation(1.0, 0.0, 0.0) * M::scaling(c[0], 1.0, c[0]),
                                               material_with<A>(plain(c[1], c[2], c[3]), {{'a', 0.1}, {'d', 0.8}, {'s', 0.9}, {'h', 300.0}})),
                                 0.0, c[4], k == 3));
    }
    const auto deco_fields = std::initializer_list<std::pair<char, double>>{{'a', 0.1}, {'d', 0.9}, {'s', 0.9}, {'h', 300.0}};
    el.push_back(E::cylinder(shape_args<A>(M::translation(0.0, 0.0, -0.75) * M::scaling(0.05, 1.0, 0.05), material_with<A>(plain(1.0, 0.0, 0.0), deco_fields)), 0.0, 0.3, true));
    el.push_back(E::cylinder(shape_args<A>(M::translation(0.0, 0.0, -2.25) * M::rotation_y(-0.15) * M::translation(0.0, 0.0, 1.5) * M::scaling(0.05, 1.0, 0.05),
                                           material_with<A>(plain(1.0, 1.0, 0.0), deco_fields)), 0.0, 0.3, true));
    el.push_back(E::cylinder(shape_args<A>(M::translation(0.0, 0.0, -2.25) * M::rotation_y(-0.3) * M::translation(0.0, 0.0, 1.5) * M::scaling(0.05, 1.0, 0.05),
                                           material_with<A>(plain(0.0, 1.0, 0.0), deco_fields)), 0.0, 0.3, true));
    el.push_back(E::cylinder(shape_args<A>(M::translation(0.0, 0.0, -2.25) * M::rotation_y(-0.45) * M::translation(0.0, 0.0, 1.5) * M::scaling(0.05, 1.0, 0.05),
                                           material_with<A>(plain(0.0, 1.0, 1.0), deco_fields)), 0.0, 0.3, true));
    el.push_back(E::cylinder(shape_args<A>(M::translation(0.0, 0.0, -1.5) * M::scaling(0.33, 1.0, 0.33),
                                           material_with<A>(plain(0.25, 0.0, 0.0), {{'d', 0.1}, {'s', 0.9}, {'h', 300.0}, {'r', 0.9}, {'t', 0.9}, {'i', 1.5}})),
                             0.0001, 0.5, true));
    s.world.lights.push_back({C::white(), V::point(1.0, 6.9, -4.9)});
    s.camera = A::Camera::make(hsize, vsize, 0.314,
                               A::Camera::transform(V::point(8.0, 3.5, -9.0), V::point(0.0, 0.3, 0.0), V::vector(0.0, 1.0, 0.0)));
    return s;
}

// src/bin/chapter14_title.rs:14-165 (native 4096x2160): three "wacky" groups (6 legs + 2 caps of 6 open cones), FOUR lights
template <class A>
Scene<A> chapter14_title(size_t hsize, size_t vsize) {
    using M = typename A::Matrix; using V = typename A::Vector; using C = typename A::Color;
    using P = typename A::Pattern; using E = typename A::Element; using G = typename A::GroupKind;
    Scene<A> s;
    auto leg = [](const typename A::Matrix& transform) {
        typename A::ShapeArgs sphere;
        sphere.transform = M::translation(0.0, 0.0, -1.0) * M::scaling(0.25, 0.25, 0.25);
        typename A::ShapeArgs cylinder;
        cylinder.transform = M::translation(0.0, 0.0, -1.0) * M::rotation_y(-0.5236) * M::rotation_z(-1.5708) * M::scaling(0.25, 1.0, 0.25);
        std::vector<E> parts;
        parts.push_back(E::sphere(sphere));
        parts.push_back(E::cylinder(cylinder, 0.0, 1.0, false));
        return E::composite(transform, std::nullopt, G::Aggregation, std::move(parts));
    };
    auto cap = [](const typename A::Matrix& transform) {
        std::vector<E> cones;
        for (int i = 0; i < 6; i++) {
            typename A::ShapeArgs a;
            a.transform = M::rotation_y(PI / 3.0 * (double)i) * M::rotation_x(-0.7854) * M::scaling(0.24606, 1.37002, 0.24606);
            cones.push_back(E::cone(a, -1.0, 0.0, false));
        }
        return E::composite(transform, std::nullopt, G::Aggregation, std::move(cones));
    };
    auto wacky = [&](const typename A::Matrix& transform, const typename A::Material& material) {
        std::vector<E> elements;
        for (int i = 0; i < 6; i++) elements.push_back(leg(M::rotation_y(PI / 3.0 * (double)i)));
        elements.push_back(cap(M::translation(0.0, 1.0, 0.0)));
        elements.push_back(cap(M::rotation_x(PI) * M::translation(0.0, 1.0, 0.0)));
        return E::composite(transform, material, G::Aggregation, std::move(elements));
    };
    auto body = [](double r, double g, double b) {
        return material_with<A>(P::plain(C::make(r, g, b)), {{'a', 0.2}, {'d', 0.8}, {'s', 0.7}, {'h', 20.0}});
    };
    auto& el = s.world.elements;
    el.push_back(E::plane(shape_args<A>(M::translation(0.0, 0.0, 100.0) * M::rotation_x(1.5708),
                                        material_with<A>(P::plain(C::white()), {{'a', 1.0}, {'d', 0.0}, {'s', 0.0}}))));
    el.push_back(wacky(M::translation(-2.8, 0.0, 0.0) * M::rotation_x(0.4363) * M::rotation_y(0.1745), body(0.9, 0.2, 0.4)));
    el.push_back(wacky(M::rotation_y(0.1745), body(0.2, 0.9, 0.6)));
    el.push_back(wacky(M::translation(2.8, 0.0, 0.0) * M::rotation_x(-0.4363) * M::rotation_y(-0.1745), body(0.2, 0.3, 1.0)));
    const double sx[4] = {1.0, -1.0, 1.0, -1.0}, sy[4] = {1.0, 1.0, -1.0, -1.0};
    for (int k = 0; k < 4; ++k) s.world.lights.push_back({C::make(0.25, 0.25, 0.25), V::point(sx[k] * 10000.0, sy[k] * 10000.0, -10000.0)});
    s.camera = A::Camera::make(hsize, vsize, 0.9,
                               A::Camera::transform(V::point(0.0, 0.0, -9.0), V::point(0.0, 0.0, 0.0), V::vector(0.0, 1.0, 0.0)));
    return s;
}

// src/bin/cover.rs:14-223 (native 4096x4096): a glass sphere and 17 cubes in nested groups, two lights
template <class A>
Scene<A> cover(size_t hsize, size_t vsize) {
    using M = typename A::Matrix; using V = typename A::Vector; using C = typename A::Color;
    using P = typename A::Pattern; using E = typename A::Element; using G = typename A::GroupKind;
    Scene<A> s;
    auto tinted = [](const typename A::Pattern& p) { return material_with<A>(p, {{'d', 0.7}, {'a', 0.1}, {'s', 0.0}, {'r', 0.1}}); };
    const auto white = tinted(P::plain(C::white()));
    const auto blue = tinted(P::plain(C::make(0.537, 0.831, 0.914)));
    const auto red = tinted(P::plain(C::make(0.941, 0.322, 0.388)));
    const auto purple = tinted(P::plain(C::make(0.373, 0.404, 0.550)));
    const auto standard_transform = M::scaling(0.5, 0.5, 0.5) * M::translation(1.0, -1.0, 1.0);
    const auto large_object = M::scaling(3.5, 3.5, 3.5) * standard_transform;
    const auto medium_object = M::scaling(3.0, 3.0, 3.0) * standard_transform;
    const auto small_object = M::scaling(2.0, 2.0, 2.0) * standard_transform;
    auto cube = [](double x, double y, double z, const typename A::Matrix& size, const typename A::Material& m) {
        return E::cube(shape_args<A>(M::translation(x, y, z) * size, m));
    };
    std::vector<E> top;
    top.push_back(E::sphere(shape_args<A>(large_object, material_with<A>(P::plain(C::make(0.373, 0.404, 0.550)),
                                                                          {{'d', 0.2}, {'a', 0.0}, {'s', 1.0}, {'h', 200.0}, {'r', 0.7}, {'t', 0.7}, {'i', 1.5}}))));
    top.push_back(cube(4.0, 0.0, 0.0, medium_object, white));
    top.push_back(cube(8.5, 1.5, -0.5, large_object, blue));
    top.push_back(cube(0.0, 0.0, 4.0, large_object, red));
    top.push_back(cube(4.0, 0.0, 4.0, small_object, white));
    top.push_back(cube(7.5, 0.5, 4.0, medium_object, purple));
    top.push_back(cube(-0.25, 0.25, 8.0, medium_object, white));
    top.push_back(cube(4.0, 1.0, 7.5, large_object, blue));
    top.push_back(cube(10.0, 2.0, 7.5, medium_object, red));
    top.push_back(cube(8.0, 2.0, 12.0, small_object, white));
    top.push_back(cube(20.0, 1.0, 9.0, small_object, white));
    std::vector<E> bot;
    bot.push_back(cube(-0.5, -5.0, 0.25, large_object, blue));
    bot.push_back(cube(4.0, -4.0, 0.0, large_object, red));
    bot.push_back(cube(8.5, -4.0, 0.0, large_object, white));
    bot.push_back(cube(0.0, -4.0, 4.0, large_object, white));
    bot.push_back(cube(-0.5, -4.5, 8.0, large_object, purple));
    bot.push_back(cube(0.0, -8.0, 4.0, large_object, white));
    bot.push_back(cube(-0.5, -8.5, 8.0, large_object, white));
    std::vector<E> all;
    all.push_back(E::composite(M::id(), std::nullopt, G::Aggregation, std::move(top)));
    all.push_back(E::composite(M::id(), std::nullopt, G::Aggregation, std::move(bot)));
    s.world.elements.push_back(E::plane(shape_args<A>(M::translation(0.0, 0.0, 500.0) * M::rotation_x(PI / 2.0),
                                                      material_with<A>(P::plain(C::white()), {{'a', 1.0}, {'d', 0.0}, {'s', 0.0}}))));
    s.world.elements.push_back(E::composite(M::id(), std::nullopt, G::Aggregation, std::move(all)));
    s.world.lights.push_back({C::white(), V::point(50.0, 100.0, -50.0)});
    s.world.lights.push_back({C::make(0.2, 0.2, 0.2), V::point(-400.0, 50.0, -10.0)});
    s.camera = A::Camera::make(hsize, vsize, 0.785,
                               A::Camera::transform(V::point(-6.0, 6.0, -10.0), V::point(6.0, 0.0, 6.0), V::vector(-0.45, 1.0, 0.0)));
    return s;
}

// --- authored: rays parallel to a cone's slope (src/shape.rs:811-818) ---------------------------------
// A narrow camera (use 33x33) looks exactly along the slope of finite cones, so the rows around the image centre take
// the cone's single-root branch (|a| < EPSILON), whose crossing has no y-range check: it lies on the infinite double
// cone, far outside [min, max].  Cone A is a top-level element: its phantom crossing is visible.  Cone B sits in a
// group whose (finite) box the rays miss: the reference rejects it at the group box although its phantom crossing
// would be nearer.  Rows further from the centre have |a| >= EPSILON and see the floor.
template <class A>
Scene<A> cone_slope(size_t hsize, size_t vsize) {
    using M = typename A::Matrix; using V = typename A::Vector; using C = typename A::Color;
    using P = typename A::Pattern; using E = typename A::Element; using G = typename A::GroupKind;
    Scene<A> s;
    auto& el = s.world.elements;
    el.push_back(E::plane(shape_args<A>(M::translation(0.0, -10.0, 0.0),
                                        material_with<A>(P::checkers(M::scaling(0.01, 0.01, 0.01), P::plain(C::make(0.9, 0.9, 0.9)), P::plain(C::make(0.2, 0.2, 0.3))), {{'s', 0.0}}))));
    el.push_back(E::cone(shape_args<A>(M::id(), material_with<A>(P::plain(C::make(0.9, 0.3, 0.2)), {{'r', 0.3}})), -1.0, 0.0, true));
    std::vector<E> grouped;
    grouped.push_back(E::cone(shape_args<A>(M::id(), material_with<A>(P::plain(C::make(0.2, 0.9, 0.3)), {})), -1.0, 0.0, true));
    el.push_back(E::composite(M::translation(0.0, 0.0, -3.0), std::nullopt, G::Aggregation, std::move(grouped)));
    std::vector<E> csg;
    csg.push_back(E::cone(shape_args<A>(M::id(), material_with<A>(P::plain(C::make(0.2, 0.3, 0.9)), {})), -1.0, 0.0, true));
    csg.push_back(E::sphere(shape_args<A>(M::translation(0.0, -0.5, 0.0) * M::scaling(0.6, 0.6, 0.6), material_with<A>(P::plain(C::make(0.9, 0.9, 0.2)), {}))));
    el.push_back(E::composite(M::translation(6.0, -8.0, 0.3), std::nullopt, G::Difference, std::move(csg)));
    s.world.lights.push_back({C::white(), V::point(-10.0, 20.0, -10.0)});
    s.camera = A::Camera::make(hsize, vsize, 1e-4,
                               A::Camera::transform(V::point(-4.0, 5.5, 0.3), V::point(-3.0, 4.5, 0.3), V::vector(0.0, 1.0, 0.0)));
    return s;
}

// --- authored: rays with direction components that are exactly zero -------------------------------------------
// The camera sits OFF the origin and looks exactly along +z, so with hsize == 1 every primary ray has d.x == 0.0
// (world_x = half_width - 0.5 * (2 * half_width) = 0 exactly, src/camera.rs:40-50) and with vsize == 1 every ray has
// d.y == 0.0; a light sits exactly above the camera column, so the shadow rays of hits with x == 1 also have
// d.x == 0.0.  The 27 items form a BVH whose inner boxes straddle 0 on x and y while the ray origins do not: the case
// an fp32 slab test with infinite inverse directions gets wrong (walk_inverse in device_trace.cuh).
template <class A>
Scene<A> axis_aligned(size_t hsize, size_t vsize) {
    using M = typename A::Matrix; using V = typename A::Vector; using C = typename A::Color;
    using P = typename A::Pattern; using E = typename A::Element;
    Scene<A> s;
    auto& el = s.world.elements;
    el.push_back(E::plane(shape_args<A>(M::translation(0.0, -0.25, 0.0),
                                        material_with<A>(P::checkers(M::id(), P::plain(C::make(0.8, 0.8, 0.8)), P::plain(C::make(0.3, 0.3, 0.35))), {{'s', 0.0}, {'r', 0.3}}))));
    int k = 0;
    for (int ix = 0; ix < 3; ix++)
        for (int iy = 0; iy < 3; iy++)
            for (int iz = 0; iz < 3; iz++, k++) {
                const auto mat = material_with<A>(P::plain(C::make(0.2 + 0.3 * ix, 0.9 - 0.3 * iy, 0.3 + 0.25 * iz)), {{'r', (k % 3 == 0) ? 0.4 : 0.0}});
                const auto args = shape_args<A>(M::translation(-0.5 + 1.5 * ix, 0.5 + 1.5 * iy, 1.5 * iz) * M::scaling(0.45, 0.45, 0.45), mat);
                el.push_back((k & 1) ? E::cube(args) : E::sphere(args));
            }
    s.world.lights.push_back({C::make(0.8, 0.8, 0.8), V::point(1.0, 12.0, -6.0)});
    s.world.lights.push_back({C::make(0.3, 0.3, 0.3), V::point(-7.0, 2.0, -6.0)});
    s.camera = A::Camera::make(hsize, vsize, 0.8,
                               A::Camera::transform(V::point(1.0, 2.0, -6.0), V::point(1.0, 2.0, 0.0), V::vector(0.0, 1.0, 0.0)));
    return s;
}

// --- authored: shapes that do not cast shadows, and flat triangles ---------------------------------------------
// World::is_shadowed looks at the NEAREST crossing only (src/world.rs:40-47): a shape with casts_shadow == false in
// front of a casting one removes the shadow, so shadow queries cannot stop at the first crossing they meet.  The
// pyramid and the tilted quad are flat triangles (Geometry::Triangle, one precomputed normal, src/shape.rs:369-385),
// which no example binary uses.
template <class A>
Scene<A> no_shadow_flat(size_t hsize, size_t vsize) {
    using M = typename A::Matrix; using V = typename A::Vector; using C = typename A::Color;
    using P = typename A::Pattern; using E = typename A::Element; using G = typename A::GroupKind;
    Scene<A> s;
    auto& el = s.world.elements;
    el.push_back(E::plane(shape_args<A>(M::id(), material_with<A>(P::plain(C::make(0.9, 0.9, 0.85)), {{'s', 0.0}}))));
    // a casting cube with a non-casting sphere between it and the light, and a non-casting sphere on its own
    el.push_back(E::cube(shape_args<A>(M::translation(-1.5, 1.0, 0.5) * M::scaling(0.6, 1.0, 0.6), material_with<A>(P::plain(C::make(0.8, 0.3, 0.2)), {}))));
    auto ghost = shape_args<A>(M::translation(-2.6, 2.6, -0.9) * M::scaling(0.7, 0.7, 0.7), material_with<A>(P::plain(C::make(0.3, 0.8, 0.9)), {{'r', 0.2}}));
    ghost.casts_shadow = false;
    el.push_back(E::sphere(ghost));
    auto ghost2 = shape_args<A>(M::translation(2.4, 0.6, -1.0) * M::scaling(0.6, 0.6, 0.6), material_with<A>(P::plain(C::make(0.9, 0.8, 0.2)), {}));
    ghost2.casts_shadow = false;
    el.push_back(E::sphere(ghost2));
    // flat-triangle pyramid (4 sides) inside a group, and a tilted quad of two flat triangles
    const V apex = V::point(0.0, 1.6, 0.0);
    const V base[4] = {V::point(-0.8, 0.0, -0.8), V::point(0.8, 0.0, -0.8), V::point(0.8, 0.0, 0.8), V::point(-0.8, 0.0, 0.8)};
    std::vector<E> sides;
    for (int i = 0; i < 4; i++)
        sides.push_back(E::triangle(shape_args<A>(M::id(), material_with<A>(P::stripes(M::scaling(0.2, 0.2, 0.2), P::plain(C::make(0.9, 0.6, 0.1)), P::plain(C::make(0.3, 0.2, 0.5))), {})),
                                    base[i], base[(i + 1) % 4], apex));
    el.push_back(E::composite(M::translation(0.6, 0.0, 0.4) * M::rotation_y(0.5), std::nullopt, G::Aggregation, std::move(sides)));
    auto quad = shape_args<A>(M::translation(0.5, 0.0, 2.5) * M::rotation_x(-0.4), material_with<A>(P::plain(C::make(0.5, 0.5, 0.9)), {{'r', 0.3}}));
    el.push_back(E::triangle(quad, V::point(-2.0, 0.0, 0.0), V::point(2.0, 0.0, 0.0), V::point(2.0, 2.5, 0.0)));
    el.push_back(E::triangle(quad, V::point(-2.0, 0.0, 0.0), V::point(2.0, 2.5, 0.0), V::point(-2.0, 2.5, 0.0)));
    s.world.lights.push_back({C::make(0.9, 0.9, 0.9), V::point(-6.0, 7.0, -5.0)});
    s.camera = A::Camera::make(hsize, vsize, 1.0,
                               A::Camera::transform(V::point(0.5, 3.0, -7.0), V::point(0.0, 1.0, 0.0), V::vector(0.0, 1.0, 0.0)));
    return s;
}

// --- authored: more than 32 group boxes that must be evaluated exactly ------------------------------------------------
// 40 groups, each holding one finite cone: a cone's single-root crossing may lie outside its box (src/shape.rs:811-818), so
// none of these group boxes is provably conservative and each one stays an exact per-ray gate (gate slots).
template <class A>
Scene<A> many_gates(size_t hsize, size_t vsize) {
    using M = typename A::Matrix; using V = typename A::Vector; using C = typename A::Color;
    using P = typename A::Pattern; using E = typename A::Element; using G = typename A::GroupKind;
    Scene<A> s;
    auto& el = s.world.elements;
    el.push_back(E::plane(shape_args<A>(M::id(), material_with<A>(P::plain(C::make(0.8, 0.8, 0.75)), {{'s', 0.0}, {'r', 0.2}}))));
    for (int i = 0; i < 40; i++) {
        const int gx = i % 8, gz = i / 8;
        std::vector<E> inner;
        inner.push_back(E::cone(shape_args<A>(M::translation(0.0, 1.0, 0.0) * M::scaling(0.4, 1.0, 0.4),
                                              material_with<A>(P::plain(C::make(0.2 + 0.1 * gx, 0.9 - 0.15 * gz, 0.3 + 0.05 * i / 4)), {{'r', (i % 5 == 0) ? 0.3 : 0.0}})),
                                -1.0, 0.0, (i & 1) != 0));
        std::vector<E> outer;  // nested: the inner group's box is a gate too
        outer.push_back(E::composite(M::rotation_y(0.2 * i), std::nullopt, G::Aggregation, std::move(inner)));
        if (i % 4 == 0) outer.push_back(E::sphere(shape_args<A>(M::translation(0.0, 1.3, 0.0) * M::scaling(0.2, 0.2, 0.2), material_with<A>(P::plain(C::make(0.9, 0.9, 0.2)), {}))));
        el.push_back(E::composite(M::translation(-5.25 + 1.5 * gx, 0.0, 1.5 * gz), std::nullopt, G::Aggregation, std::move(outer)));
    }
    s.world.lights.push_back({C::make(0.9, 0.9, 0.9), V::point(-6.0, 9.0, -8.0)});
    s.camera = A::Camera::make(hsize, vsize, 0.9,
                               A::Camera::transform(V::point(0.0, 6.0, -9.0), V::point(0.0, 0.5, 3.0), V::vector(0.0, 1.0, 0.0)));
    return s;
}

// --- authored: a mesh beyond the one-block device sort (obj.rs has no size limit) -------------------
// teapot_high.obj with every triangle split into 16 (two rounds of edge midpoints; normals of the midpoints = normalised
// sums): 101 120 smooth triangles in the chapter-15 scene.  Exercises the multi-block radix sort of bvh_build.cu.
template <class A>
typename A::ObjData tessellate(const typename A::ObjData& in, int rounds) {
    using V = typename A::Vector;
    typename A::ObjData d = in;
    for (int r = 0; r < rounds; r++) {
        std::vector<typename A::ObjData::Tri> next;
        next.reserve(d.triangles.size() * 4);
        for (const auto& t : d.triangles) {
            size_t mp[3], mn[3] = {0, 0, 0};
            for (int e = 0; e < 3; e++) {
                const V& a = d.vertices[t.p[e] - 1];
                const V& b = d.vertices[t.p[(e + 1) % 3] - 1];
                d.vertices.push_back(V::point(0.5 * (a.x + b.x), 0.5 * (a.y + b.y), 0.5 * (a.z + b.z)));
                mp[e] = d.vertices.size();
                if (t.has_n) {
                    const V& na = d.normals[t.n[e] - 1];
                    const V& nb = d.normals[t.n[(e + 1) % 3] - 1];
                    d.normals.push_back(V::vector(na.x + nb.x, na.y + nb.y, na.z + nb.z).normalize());
                    mn[e] = d.normals.size();
                }
            }
            auto tri = [&](size_t p0, size_t p1, size_t p2, size_t n0, size_t n1, size_t n2) {
                typename A::ObjData::Tri q = t;
                q.p[0] = p0; q.p[1] = p1; q.p[2] = p2;
                q.n[0] = n0; q.n[1] = n1; q.n[2] = n2;
                next.push_back(q);
            };
            tri(t.p[0], mp[0], mp[2], t.n[0], mn[0], mn[2]);
            tri(mp[0], t.p[1], mp[1], mn[0], t.n[1], mn[1]);
            tri(mp[2], mp[1], t.p[2], mn[2], mn[1], t.n[2]);
            tri(mp[0], mp[1], mp[2], mn[0], mn[1], mn[2]);
        }
        d.triangles.swap(next);
    }
    return d;
}
template <class A>
Scene<A> teapot_tess(size_t hsize, size_t vsize) {
    return teapot_from<A>(tessellate<A>(load_mesh_fixture<A>(data_dir() + "/teapot_high.mesh"), 2), hsize, vsize);
}

// --- authored: crossing lists beyond the usual on-chip sizes -------------------------------------------------------------
// The camera sits inside 70 concentric glass shells: the line of every ray crosses all of them behind the camera, so the
// n1 / n2 replay of a hit (src/intersection.rs:70-103) sees 70+ crossings and 70 open containers — beyond the 64 / 24 the
// on-chip lists hold, inside what a slot of global scratch holds (device_scene.h: BIG_CROSSINGS, BIG_CONTAINERS).
template <class A>
Scene<A> glass_shells(size_t hsize, size_t vsize) {
    using M = typename A::Matrix; using V = typename A::Vector; using C = typename A::Color;
    using P = typename A::Pattern; using E = typename A::Element;
    Scene<A> s;
    auto& el = s.world.elements;
    el.push_back(E::plane(shape_args<A>(M::translation(0.0, -8.0, 0.0),
                                        material_with<A>(P::checkers(M::scaling(2.0, 2.0, 2.0), P::plain(C::make(0.9, 0.9, 0.9)), P::plain(C::make(0.2, 0.3, 0.2))), {{'s', 0.0}}))));
    for (int k = 0; k < 70; k++) {
        const double r = 1.0 + 0.05 * k;
        el.push_back(E::sphere(shape_args<A>(M::scaling(r, r, r), material_with<A>(P::plain(C::make(0.2 + 0.15 * (k % 5), 0.9 - 0.1 * (k % 7), 0.5)),
                                                                                      {{'a', 0.3}, {'d', 0.3}, {'s', 0.3}, {'h', 200.0}, {'r', (k % 9 == 0) ? 0.2 : 0.0}, {'t', 0.95}, {'i', 1.0 + 0.01 * (k % 13)}}))));
    }
    s.world.lights.push_back({C::make(0.9, 0.9, 0.9), V::point(10.0, 12.0, -10.0)});
    s.camera = A::Camera::make(hsize, vsize, 1.2,
                               A::Camera::transform(V::point(0.1, 0.05, -0.2), V::point(0.3, -0.6, 1.0), V::vector(0.0, 1.0, 0.0)));
    return s;
}

// --- authored: a triangles-only BVH (FEAT_TRIS kernels) with a shape that casts no shadow ---------------------------------
// Two fans of flat triangles and a strip of smooth ones over a plane; one fan does not cast shadows, so the shadow queries are
// nearest-hit queries (src/world.rs:43-47) in the triangles-only kernel variant.
template <class A>
Scene<A> tri_ghost(size_t hsize, size_t vsize) {
    using M = typename A::Matrix; using V = typename A::Vector; using C = typename A::Color;
    using P = typename A::Pattern; using E = typename A::Element; using G = typename A::GroupKind;
    Scene<A> s;
    auto& el = s.world.elements;
    el.push_back(E::plane(shape_args<A>(M::id(), material_with<A>(P::plain(C::make(0.85, 0.85, 0.8)), {{'s', 0.0}, {'r', 0.2}}))));
    for (int fan = 0; fan < 2; fan++) {
        std::vector<E> tris;
        const V apex = V::point(0.0, 1.8, 0.0);
        for (int i = 0; i < 24; i++) {
            const double a0 = 2.0 * PI * i / 24.0, a1 = 2.0 * PI * (i + 1) / 24.0;
            auto args = shape_args<A>(M::id(), material_with<A>(P::plain(C::make(0.3 + 0.6 * fan, 0.7 - 0.3 * fan, 0.2 + 0.03 * i)), {{'r', (i % 3 == 0) ? 0.3 : 0.0}}));
            args.casts_shadow = fan == 0;
            tris.push_back(E::triangle(args, V::point(std::cos(a0), 0.2, std::sin(a0)), V::point(std::cos(a1), 0.2, std::sin(a1)), apex));
        }
        el.push_back(E::composite(M::translation(-1.6 + 3.2 * fan, 0.0, 0.5 * fan) * M::rotation_z(0.15 * fan), std::nullopt, G::Aggregation, std::move(tris)));
    }
    std::vector<E> strip;
    for (int i = 0; i < 40; i++) {
        const double x0 = -2.0 + 0.1 * i, x1 = x0 + 0.1;
        auto args = shape_args<A>(M::id(), material_with<A>(P::plain(C::make(0.5, 0.5, 0.9)), {{'r', 0.2}}));
        const V n0 = V::vector(std::sin(3.0 * x0), 1.0, -0.3).normalize(), n1 = V::vector(std::sin(3.0 * x1), 1.0, -0.3).normalize();
        strip.push_back(E::smooth_triangle(args, V::point(x0, 0.9, 2.0), V::point(x1, 0.9, 2.0), V::point(x0, 1.6, 2.6), n0, n1, n0));
        strip.push_back(E::smooth_triangle(args, V::point(x1, 0.9, 2.0), V::point(x1, 1.6, 2.6), V::point(x0, 1.6, 2.6), n1, n1, n0));
    }
    el.push_back(E::composite(M::id(), std::nullopt, G::Aggregation, std::move(strip)));
    s.world.lights.push_back({C::make(0.9, 0.9, 0.9), V::point(-5.0, 8.0, -6.0)});
    s.world.lights.push_back({C::make(0.3, 0.3, 0.35), V::point(6.0, 4.0, -5.0)});
    s.camera = A::Camera::make(hsize, vsize, 1.0,
                               A::Camera::transform(V::point(0.3, 3.2, -6.5), V::point(0.0, 0.9, 0.5), V::vector(0.0, 1.0, 0.0)));
    return s;
}

// --- authored: CSG operands whose boxes differ a lot from the box of what the group can report ---------------------------
// The BVH holds a CSG item under the box of its reportable crossings (left operand of a Difference, overlap of an
// Intersection) and the CSG program skips the right operand when the left one reports nothing; this scene is made of the
// cases where that matters: a right operand sticking far out of the left one, a half-space (plane) as the right operand,
// an unbounded left operand of an Intersection, disjoint operands, a finite cone (inexact box) on either side, nesting.
template <class A>
Scene<A> csg_cuts(size_t hsize, size_t vsize) {
    using M = typename A::Matrix; using V = typename A::Vector; using C = typename A::Color;
    using P = typename A::Pattern; using E = typename A::Element; using G = typename A::GroupKind;
    Scene<A> s;
    auto& el = s.world.elements;
    auto plain = [](double r, double g, double b) { return P::plain(C::make(r, g, b)); };
    auto pair = [](E a, E b) { std::vector<E> v; v.push_back(std::move(a)); v.push_back(std::move(b)); return v; };
    el.push_back(E::plane(shape_args<A>(M::id(), material_with<A>(P::checkers(M::id(), plain(0.8, 0.8, 0.8), plain(0.35, 0.35, 0.4)), {{'s', 0.0}, {'r', 0.25}}))));
    // (1) small sphere minus a long thin cylinder through it: the cylinder sticks 6 units out on both sides
    el.push_back(E::composite(M::translation(-3.0, 1.0, 0.0) * M::rotation_z(0.4), std::nullopt, G::Difference,
                              pair(E::sphere(shape_args<A>(M::id(), material_with<A>(plain(0.9, 0.3, 0.2), {{'r', 0.1}}))),
                                   E::cylinder(shape_args<A>(M::scaling(0.45, 1.0, 0.45), material_with<A>(plain(0.95, 0.9, 0.3), {})), -6.0, 6.0, true))));
    // (2) cube cut by a half-space: the right operand is an unbounded plane
    el.push_back(E::composite(M::translation(-0.6, 0.8, 0.8) * M::rotation_y(0.5), std::nullopt, G::Difference,
                              pair(E::cube(shape_args<A>(M::scaling(0.8, 0.8, 0.8), material_with<A>(plain(0.2, 0.6, 0.9), {{'r', 0.15}}))),
                                   E::plane(shape_args<A>(M::translation(0.0, 0.3, 0.0) * M::rotation_z(0.6), material_with<A>(plain(0.9, 0.9, 0.9), {}))))));
    // (3) half-space AND sphere: the left operand is unbounded, the group can only report inside the sphere
    el.push_back(E::composite(M::translation(1.6, 0.9, -0.3), std::nullopt, G::Intersection,
                              pair(E::plane(shape_args<A>(M::rotation_x(0.5), material_with<A>(plain(0.9, 0.7, 0.2), {}))),
                                   E::sphere(shape_args<A>(M::scaling(0.9, 0.9, 0.9), material_with<A>(plain(0.3, 0.8, 0.4), {{'h', 80.0}}))))));
    // (4) disjoint operands: never reports anything, but its (large) reference box covers the cone group behind it
    el.push_back(E::composite(M::translation(0.0, 2.6, 2.5), std::nullopt, G::Intersection,
                              pair(E::sphere(shape_args<A>(M::translation(-3.0, 0.0, 0.0), material_with<A>(plain(1.0, 0.0, 1.0), {}))),
                                   E::sphere(shape_args<A>(M::translation(3.0, 0.0, 0.0), material_with<A>(plain(1.0, 0.0, 1.0), {}))))));
    // (5) a finite cone on the left (box not exact) minus a sphere; (6) a sphere minus a finite cone
    el.push_back(E::composite(M::translation(3.4, 1.6, 1.2), std::nullopt, G::Difference,
                              pair(E::cone(shape_args<A>(M::scaling(0.8, 1.4, 0.8), material_with<A>(plain(0.8, 0.5, 0.9), {})), -1.0, 0.0, true),
                                   E::sphere(shape_args<A>(M::translation(0.3, -0.9, -0.5) * M::scaling(0.6, 0.6, 0.6), material_with<A>(plain(0.2, 0.2, 0.2), {}))))));
    el.push_back(E::composite(M::translation(0.3, 1.0, 3.2), std::nullopt, G::Difference,
                              pair(E::sphere(shape_args<A>(M::id(), material_with<A>(plain(0.9, 0.6, 0.3), {{'r', 0.2}}))),
                                   E::cone(shape_args<A>(M::translation(0.0, 1.2, -0.6) * M::scaling(0.7, 1.6, 0.7), material_with<A>(plain(0.3, 0.3, 0.7), {})), -1.0, 0.0, false))));
    // (7) nested: (sphere AND cube) minus (cylinder OR (sphere minus cube)), glass, inside a plain group with a transform
    {
        std::vector<E> inner = pair(E::sphere(shape_args<A>(M::scaling(1.25, 1.25, 1.25), material_with<A>(plain(0.1, 0.1, 0.15), {{'t', 0.7}, {'i', 1.5}, {'r', 0.3}, {'d', 0.2}}))),
                                    E::cube(shape_args<A>(M::id(), material_with<A>(plain(0.1, 0.15, 0.1), {{'t', 0.7}, {'i', 1.5}, {'r', 0.3}, {'d', 0.2}}))));
        std::vector<E> dent = pair(E::sphere(shape_args<A>(M::translation(0.9, 0.9, -0.9) * M::scaling(0.7, 0.7, 0.7), material_with<A>(plain(0.9, 0.2, 0.2), {}))),
                                   E::cube(shape_args<A>(M::translation(1.4, 1.4, -1.4) * M::scaling(0.5, 0.5, 0.5), material_with<A>(plain(0.2, 0.9, 0.2), {}))));
        std::vector<E> holes = pair(E::cylinder(shape_args<A>(M::scaling(0.4, 1.0, 0.4), material_with<A>(plain(0.9, 0.9, 0.2), {})), -4.0, 4.0, false),
                                    E::composite(M::id(), std::nullopt, G::Difference, std::move(dent)));
        std::vector<E> whole = pair(E::composite(M::id(), std::nullopt, G::Intersection, std::move(inner)),
                                    E::composite(M::rotation_x(0.3), std::nullopt, G::Union, std::move(holes)));
        std::vector<E> grp;
        grp.push_back(E::composite(M::id(), std::nullopt, G::Difference, std::move(whole)));
        grp.push_back(E::sphere(shape_args<A>(M::translation(0.0, 1.7, 0.0) * M::scaling(0.3, 0.3, 0.3), material_with<A>(plain(0.9, 0.9, 0.9), {{'r', 0.4}}))));
        el.push_back(E::composite(M::translation(-2.0, 1.1, -2.6) * M::rotation_y(0.7) * M::scaling(0.85, 0.85, 0.85), std::nullopt, G::Aggregation, std::move(grp)));
    }
    s.world.lights.push_back({C::make(0.85, 0.85, 0.85), V::point(-6.0, 9.0, -7.0)});
    s.world.lights.push_back({C::make(0.3, 0.3, 0.3), V::point(7.0, 5.0, -4.0)});
    s.camera = A::Camera::make(hsize, vsize, 1.1,
                               A::Camera::transform(V::point(0.5, 4.2, -8.5), V::point(0.0, 1.0, 0.5), V::vector(0.0, 1.0, 0.0)));
    return s;
}

// --- registry ---------------------------------------------------------------------------------
inline const std::vector<std::string>& scene_names() {
    static const std::vector<std::string> n = {"default_world", "three_spheres", "glass_air_bubble", "hexagon",
                                               "benchmark",     "teapot",        "teapot_low",       "csg_noise",
                                               "chapter11_title", "chapter12_title", "chapter13_title", "chapter14_title", "cover", "cone_slope",
                                               "axis_aligned", "no_shadow_flat", "many_gates", "glass_shells", "tri_ghost", "csg_cuts"};
    return n;
}

// name is one of scene_names() or "fuzz:<seed>"
template <class A>
Scene<A> build_scene(const std::string& name, size_t hsize, size_t vsize) {
    if (name == "default_world") return default_world<A>(hsize, vsize);
    if (name == "three_spheres") return three_spheres<A>(hsize, vsize);
    if (name == "glass_air_bubble") return glass_air_bubble<A>(hsize, vsize);
    if (name == "hexagon") return hexagon<A>(hsize, vsize);
    if (name == "benchmark") return benchmark<A>(hsize, vsize);
    if (name == "teapot") return teapot<A>(hsize, vsize);
    if (name == "teapot_low") return teapot_low<A>(hsize, vsize);
    if (name == "csg_noise") return csg_noise<A>(hsize, vsize);
    if (name == "chapter11_title") return chapter11_title<A>(hsize, vsize);
    if (name == "chapter12_title") return chapter12_title<A>(hsize, vsize);
    if (name == "chapter13_title") return chapter13_title<A>(hsize, vsize);
    if (name == "chapter14_title") return chapter14_title<A>(hsize, vsize);
    if (name == "cover") return cover<A>(hsize, vsize);
    if (name == "cone_slope") return cone_slope<A>(hsize, vsize);
    if (name == "axis_aligned") return axis_aligned<A>(hsize, vsize);
    if (name == "no_shadow_flat") return no_shadow_flat<A>(hsize, vsize);
    if (name == "many_gates") return many_gates<A>(hsize, vsize);
    if (name == "teapot_tess") return teapot_tess<A>(hsize, vsize);
    if (name == "glass_shells") return glass_shells<A>(hsize, vsize);
    if (name == "tri_ghost") return tri_ghost<A>(hsize, vsize);
    if (name == "csg_cuts") return csg_cuts<A>(hsize, vsize);
    if (name == "benchmark_glass") return benchmark_glass<A>(hsize, vsize);
    if (name.rfind("fuzz:", 0) == 0) return fuzz<A>(std::strtoull(name.c_str() + 5, nullptr, 10), hsize, vsize);
    throw std::runtime_error("scenes: unknown scene '" + name + "'");
}

}  // namespace scenes
