// ORACLE — TEST INFRASTRUCTURE ONLY.
// Pins the C++ restatement (rt_oracle.hpp) to the reference's own known-answer unit tests.
// Each section transcribes the inputs/expected values (data) of the inline `#[cfg(test)]`
// modules of the cited reference file; assertions use the reference's own tolerance
// (Approx: |a-b| < EPSILON = 1e-5, src/approx.rs:19-23).  Run by tests/test_oracle_known_answers.py.
#include "oracle_api.hpp"

using namespace orc;

static int g_checks = 0, g_failed = 0;
static const char* g_section = "";
#define CHECK(cond)                                                                        \
    do {                                                                                   \
        g_checks++;                                                                        \
        if (!(cond)) {                                                                     \
            g_failed++;                                                                    \
            std::printf("FAIL [%s] %s:%d  %s\n", g_section, __FILE__, __LINE__, #cond);    \
        }                                                                                  \
    } while (0)
#define SECTION(name) g_section = name

static const double SQ2 = std::sqrt(2.0);
static Ray mkray(double ox, double oy, double oz, double dx, double dy, double dz) {
    return {Vector::point(ox, oy, oz), Vector::vector(dx, dy, dz)};
}
static Intersection mkx(double t, const Shape* s) { return {t, s, false, 0.0, 0.0}; }
static Material mat_color(double r, double g, double b) {
    Material m;
    m.pattern = Pattern::plain(Color::make(r, g, b));
    return m;
}

// ------------------------------------------------------------------------------------------
// src/world.rs:202-871
static World world_with_reflective_plane(double reflective, double transparency = 0.0, double index = 1.0) {
    World w = World::default_world();
    ShapeArgs a;
    a.transform = Matrix::translation(0.0, -1.0, 0.0);
    a.material.reflective = reflective;
    a.material.transparency = transparency;
    a.material.refractive_index = index;
    w.elements.push_back(Element::plane(a));
    return w;
}

static void test_world() {
    SECTION("world.rs");
    Intersections scratch;
    {  // intersect_default_world_ray :202-219
        World w = World::default_world();
        Intersections is;
        w.intersect(mkray(0, 0, -5, 0, 0, 1), is);
        sort_intersections(is);
        CHECK(is.size() == 4 && approx(is[0].t, 4.0) && approx(is[1].t, 4.5) && approx(is[2].t, 5.5) && approx(is[3].t, 6.0));
    }
    {  // shade_intersection_outside :221-240
        World w = World::default_world();
        Ray r = mkray(0, 0, -5, 0, 0, 1);
        State st = prepare_state(mkx(4.0, &w.elements[0].shape), r, {});
        CHECK(approx(w.shade_hit(st, FUEL, scratch), Color::make(0.38066, 0.47583, 0.28550)));
    }
    {  // shade_intersection_inside :242-267
        World w = World::default_world();
        w.lights = {{Color::white(), Vector::point(0.0, 0.25, 0.0)}};
        Ray r = mkray(0, 0, 0, 0, 0, 1);
        State st = prepare_state(mkx(0.5, &w.elements[1].shape), r, {});
        CHECK(approx(w.shade_hit(st, FUEL, scratch), Color::make(0.90498, 0.90498, 0.90498)));
    }
    {  // color_ray_miss / color_ray_hit :269-291
        World w = World::default_world();
        CHECK(approx(w.color_at(mkray(0, 0, -5, 0, 1, 0), FUEL, scratch), Color::black()));
        CHECK(approx(w.color_at(mkray(0, 0, -5, 0, 0, 1), FUEL, scratch), Color::make(0.38066, 0.47583, 0.28550)));
    }
    {  // color_intersection_behind_ray :293-331
        World w = World::default_world();
        w.elements.clear();
        ShapeArgs a1;
        a1.material = mat_color(0.8, 1.0, 0.6);
        a1.material.diffuse = 0.7;
        a1.material.specular = 0.2;
        a1.material.ambient = 1.0;
        ShapeArgs a2;
        a2.transform = Matrix::scaling(0.5, 0.5, 0.5);
        a2.material.ambient = 1.0;
        w.elements.push_back(Element::sphere(a1));
        w.elements.push_back(Element::sphere(a2));
        CHECK(approx(w.color_at(mkray(0, 0, 0.75, 0, 0, -1), FUEL, scratch), Color::white()));
    }
    {  // shadow_* :333-363
        World w = World::default_world();
        CHECK(!w.is_shadowed(w.lights[0], Vector::point(0, 10, 0), scratch));
        CHECK(w.is_shadowed(w.lights[0], Vector::point(10, -10, 10), scratch));
        CHECK(!w.is_shadowed(w.lights[0], Vector::point(-20, 20, -20), scratch));
        CHECK(!w.is_shadowed(w.lights[0], Vector::point(-2, 2, -2), scratch));
    }
    {  // color_intersection_in_shadow :365-399
        World w;
        w.lights = {{Color::white(), Vector::point(0, 0, -10)}};
        ShapeArgs a2;
        a2.transform = Matrix::translation(0, 0, 10);
        w.elements.push_back(Element::sphere(ShapeArgs{}));
        w.elements.push_back(Element::sphere(a2));
        Element lone = Element::sphere(a2);
        State st = prepare_state(mkx(4.0, &lone.shape), mkray(0, 0, 5, 0, 0, 1), {});
        CHECK(approx(w.shade_hit(st, FUEL, scratch), Color::make(0.1, 0.1, 0.1)));
    }
    {  // reflected_color_nonreflective_materiall :401-442
        World w = World::default_world();
        w.elements[1].shape.material.ambient = 1.0;
        State st = prepare_state(mkx(1.0, &w.elements[1].shape), mkray(0, 0, 0, 0, 0, 1), {});
        CHECK(approx(w.reflected_color(st, FUEL, scratch), Color::black()));
    }
    {  // reflected_color_reflective_material :444-471, shade_hit_reflective_material :473-500, max depth :538-565
        World w = world_with_reflective_plane(0.5);
        Ray r = mkray(0, 0, -3, 0, SQ2 / -2.0, SQ2 / 2.0);
        State st = prepare_state(mkx(SQ2, &w.elements[2].shape), r, {});
        CHECK(approx(w.reflected_color(st, FUEL, scratch), Color::make(0.190332, 0.237915, 0.142749)));
        CHECK(approx(w.shade_hit(st, FUEL, scratch), Color::make(0.876757, 0.924340, 0.829174)));
        CHECK(approx(w.reflected_color(st, 0, scratch), Color::black()));
    }
    {  // color_at_mutually_reflective_surfaces :502-536 (must terminate)
        World w;
        w.lights = {{Color::white(), Vector::point(0, 0, 0)}};
        ShapeArgs lo, up;
        lo.transform = Matrix::translation(0, -1, 0);
        lo.material.reflective = 1.0;
        up.transform = Matrix::translation(0, 1, 0);
        up.material.reflective = 1.0;
        w.elements.push_back(Element::plane(lo));
        w.elements.push_back(Element::plane(up));
        Color c = w.color_at(mkray(0, 0, 0, 0, 1, 0), FUEL, scratch);
        CHECK(c.r == c.r);
    }
    {  // refracted_color_opaque_surface :567-631
        World w = World::default_world();
        const Shape* s = &w.elements[0].shape;
        Intersections is = {mkx(4.0, s), mkx(6.0, s)};
        State st = prepare_state(is[0], mkray(0, 0, -5, 0, 0, 1), is);
        CHECK(approx(w.refracted_color(st, 5, scratch), Color::black()));
    }
    {  // refracted_color_maximum_recursive_depth :633-661, under_total_internal_reflection :663-691
        World w = World::default_world();
        w.elements[0].shape.material.transparency = 1.0;
        w.elements[0].shape.material.refractive_index = 1.5;
        const Shape* s = &w.elements[0].shape;
        Intersections is = {mkx(4.0, s), mkx(6.0, s)};
        State st = prepare_state(is[0], mkray(0, 0, -5, 0, 0, 1), is);
        CHECK(approx(w.refracted_color(st, 0, scratch), Color::black()));
        Intersections is2 = {mkx(SQ2 / -2.0, s), mkx(SQ2 / 2.0, s)};
        State st2 = prepare_state(is2[1], mkray(0, 0, SQ2 / 2.0, 0, 1, 0), is2);
        CHECK(approx(w.refracted_color(st2, 5, scratch), Color::black()));
    }
    {  // refracted_color_with_refracted_ray :693-739 (Pattern::Debug observes the refracted ray)
        World w = World::default_world();
        w.elements[0].shape.material.ambient = 1.0;
        w.elements[0].shape.material.pattern = Pattern::debug();
        w.elements[1].shape.material.transparency = 1.0;
        w.elements[1].shape.material.refractive_index = 1.5;
        const Shape* a = &w.elements[0].shape;
        const Shape* b = &w.elements[1].shape;
        Intersections is = {mkx(-0.9899, a), mkx(-0.4899, b), mkx(0.4899, b), mkx(0.9899, a)};
        State st = prepare_state(is[2], mkray(0, 0, 0.1, 0, 1, 0), is);
        CHECK(approx(w.refracted_color(st, 5, scratch), Color::make(0.0, 0.998874, 0.047218)));
    }
    for (int both = 0; both < 2; both++) {  // shade_hit_with_transparent_material :741-785, reflective+transparent :787-832
        World w = world_with_reflective_plane(both ? 0.5 : 0.0, 0.5, 1.5);
        ShapeArgs ball;
        ball.transform = Matrix::translation(0.0, -3.5, -0.5);
        ball.material = mat_color(1.0, 0.0, 0.0);
        ball.material.ambient = 0.5;
        w.elements.push_back(Element::sphere(ball));
        Intersections is = {mkx(SQ2, &w.elements[2].shape)};
        State st = prepare_state(is[0], mkray(0, 0, -3, 0, SQ2 / -2.0, SQ2 / 2.0), is);
        Color c = w.shade_hit(st, 5, scratch);
        if (both) CHECK(approx(c, Color::make(0.93391, 0.69643, 0.69243)));
        else CHECK(approx(c, Color::make(0.93642, 0.68642, 0.68642)));
    }
}

// ------------------------------------------------------------------------------------------
// src/intersection.rs:153-654
static Element sphere_with(const Matrix& t, double transparency, double index) {
    ShapeArgs a;
    a.transform = t;
    a.material.transparency = transparency;
    a.material.refractive_index = index;
    return Element::sphere(a);
}

static void test_intersection() {
    SECTION("intersection.rs");
    Element se = Element::sphere(ShapeArgs{});
    const Shape* sphere = &se.shape;
    {  // aggregating_intersections :153-180, hit_* :182-271
        Intersections is = {mkx(2.0, sphere), mkx(1.0, sphere)};
        sort_intersections(is);
        CHECK(is.size() == 2 && is[0].t == 1.0 && is[1].t == 2.0 && is[0].shape == sphere);
        CHECK(hit(is) && hit(is)->t == 1.0);
        Intersections b = {mkx(1.0, sphere), mkx(-1.0, sphere)};
        CHECK(hit(b) && hit(b)->t == 1.0);
        Intersections c = {mkx(5.0, sphere), mkx(7.0, sphere), mkx(-3.0, sphere), mkx(2.0, sphere)};
        sort_intersections(c);
        CHECK(hit(c) && hit(c)->t == 2.0);
        Intersections d = {mkx(-1.0, sphere), mkx(-2.0, sphere)};
        CHECK(hit(d) == nullptr);
    }
    {  // precompute_state_intersection :273-296, hit_intersection_outside/inside :298-339
        State st = prepare_state(mkx(4.0, sphere), mkray(0, 0, -5, 0, 0, 1), {});
        CHECK(approx(st.t, 4.0) && st.shape == sphere && approx(st.point, Vector::point(0, 0, -1)) &&
              approx(st.eye, Vector::vector(0, 0, -1)) && approx(st.normal, Vector::vector(0, 0, -1)));
        CHECK(!st.inside);
        State in = prepare_state(mkx(1.0, sphere), mkray(0, 0, 0, 0, 0, 1), {});
        CHECK(approx(in.point, Vector::point(0, 0, 1)) && approx(in.eye, Vector::vector(0, 0, -1)) &&
              approx(in.normal, Vector::vector(0, 0, -1)) && in.inside);
    }
    {  // hit_offset_over_point :341-360, under_point_offset_below_surface :484-503
        ShapeArgs a;
        a.transform = Matrix::translation(0, 0, 1);
        Element e = Element::sphere(a);
        State st = prepare_state(mkx(5.0, &e.shape), mkray(0, 0, -5, 0, 0, 1), {});
        CHECK(st.over_point.z < -EPSILON / 2.0 && st.over_point.z < st.point.z);
        CHECK(st.under_point.z > EPSILON / 2.0 && st.point.z < st.under_point.z);
    }
    {  // precomputing_reflection_vector :362-381
        Element p = Element::plane(ShapeArgs{});
        State st = prepare_state(mkx(SQ2, &p.shape), mkray(0, 1, -1, 0, SQ2 / -2.0, SQ2 / 2.0), {});
        CHECK(approx(st.reflect, Vector::vector(0, SQ2 / 2.0, SQ2 / 2.0)));
    }
    {  // finding_n1_and_n2_at_intersections :383-482 (three nested glass spheres, 6 crossings)
        Element a = sphere_with(Matrix::scaling(2, 2, 2), consts::transparency::GLASS, 1.5);
        Element b = sphere_with(Matrix::translation(0, 0, -0.25), consts::transparency::GLASS, 2.0);
        Element c = sphere_with(Matrix::translation(0, 0, 0.25), consts::transparency::GLASS, 2.5);
        Ray r = mkray(0, 0, -4, 0, 0, 1);
        Intersections is = {mkx(2.0, &a.shape), mkx(2.75, &b.shape), mkx(3.25, &c.shape),
                            mkx(4.75, &b.shape), mkx(5.25, &c.shape), mkx(6.0, &a.shape)};
        const double n1[6] = {1.0, 1.5, 2.0, 2.5, 2.5, 1.5};
        const double n2[6] = {1.5, 2.0, 2.5, 2.5, 1.5, 1.0};
        for (int i = 0; i < 6; i++) {
            State st = prepare_state(is[i], r, is);
            CHECK(approx(st.n1, n1[i]) && approx(st.n2, n2[i]));
        }
    }
    {  // schlick_* :505-654
        Element g = sphere_with(Matrix::id(), 1.0, 1.5);
        Intersections is = {mkx(SQ2 / -2.0, &g.shape), mkx(SQ2 / 2.0, &g.shape)};
        CHECK(approx(prepare_state(is[1], mkray(0, 0, SQ2 / 2.0, 0, 1, 0), is).reflectance, 1.0));
        Intersections is2 = {mkx(-1.0, &g.shape), mkx(1.0, &g.shape)};
        CHECK(approx(prepare_state(is2[1], mkray(0, 0, 0, 0, 1, 0), is2).reflectance, 0.04));
        Intersections is3 = {mkx(1.8589, &g.shape)};
        CHECK(approx(prepare_state(is3[0], mkray(0, 0.99, -2, 0, 0, 1), is3).reflectance, 0.48873));
    }
}

// ------------------------------------------------------------------------------------------
// src/shape.rs:1006-1450 — primitives
struct RayCase { double o[3], d[3]; };
static Intersections isect(const Element& e, const Ray& r) {
    Intersections is;
    e.intersect(r, is);
    return is;
}
static Ray ray_of(const RayCase& c, bool normalize = false) {
    Vector d = Vector::vector(c.d[0], c.d[1], c.d[2]);
    return {Vector::point(c.o[0], c.o[1], c.o[2]), normalize ? d.normalize() : d};
}
static ShapeArgs args_t(const Matrix& t) { ShapeArgs a; a.transform = t; return a; }

static void test_primitives() {
    SECTION("shape.rs primitives");
    const double inf = std::numeric_limits<double>::infinity();
    const double S3 = std::sqrt(3.0) / 3.0;
    {  // ray_sphere_hit :1008-1029
        struct { RayCase r; Matrix t; double t1, t2; } cs[] = {
            {{{0, 0, -5}, {0, 0, 1}}, Matrix::id(), 4.0, 6.0},   {{{0, 0, 0}, {0, 0, 1}}, Matrix::id(), -1.0, 1.0},
            {{{0, 0, 5}, {0, 0, 1}}, Matrix::id(), -6.0, -4.0},  {{{0, 0, -5}, {0, 0, 1}}, Matrix::scaling(2, 2, 2), 3.0, 7.0},
            {{{0, 1, -5}, {0, 0, 1}}, Matrix::id(), 5.0, 5.0}};
        for (auto& c : cs) {
            Element s = Element::sphere(args_t(c.t));
            Intersections is = isect(s, ray_of(c.r));
            CHECK(is.size() == 2 && approx(is[0].t, c.t1) && approx(is[1].t, c.t2) && is[0].shape == &s.shape);
        }
        // ray_sphere_miss :1031-1042
        CHECK(isect(Element::sphere(ShapeArgs{}), mkray(0, 2, -5, 0, 0, 1)).empty());
        CHECK(isect(Element::sphere(args_t(Matrix::translation(5, 0, 0))), mkray(0, 0, -5, 0, 0, 1)).empty());
    }
    {  // sphere_normal :1044-1072, sphere_normal_normalized :1074-1087, sphere_bbox :1089-1100
        struct { double p[3], n[3]; Matrix t; } cs[] = {
            {{1, 0, 0}, {1, 0, 0}, Matrix::id()}, {{0, 1, 0}, {0, 1, 0}, Matrix::id()}, {{0, 0, 1}, {0, 0, 1}, Matrix::id()},
            {{S3, S3, S3}, {S3, S3, S3}, Matrix::id()},
            {{0, 1.70711, -0.70711}, {0, 0.70711, -0.70711}, Matrix::translation(0, 1, 0)},
            {{0, SQ2 / 2.0, SQ2 / -2.0}, {0, 0.97014, -0.24254}, Matrix::scaling(1, 0.5, 1) * Matrix::rotation_z(scenes::PI / 5.0)}};
        for (auto& c : cs) {
            Element s = Element::sphere(args_t(c.t));
            CHECK(approx(s.shape.normal(Vector::point(c.p[0], c.p[1], c.p[2]), false, 0, 0), Vector::vector(c.n[0], c.n[1], c.n[2])));
        }
        Element s = Element::sphere(ShapeArgs{});
        Vector n = s.shape.normal(Vector::point(S3, S3, S3), false, 0, 0);
        CHECK(approx(n, n.normalize()));
        Element b = Element::sphere(args_t(Matrix::translation(1, -3, 5) * Matrix::scaling(0.5, 2, 4)));
        CHECK(approx(b.bbox().min, Vector::point(0.5, -5, 1)) && approx(b.bbox().max, Vector::point(1.5, -1, 9)));
    }
    {  // plane :1104-1133
        Element p = Element::plane(ShapeArgs{});
        Intersections a = isect(p, mkray(0, 1, 0, 0, -1, 0));
        CHECK(a.size() == 1 && approx(a[0].t, 1.0));
        Intersections b = isect(p, mkray(0, -1, 0, 0, 1, 0));
        CHECK(b.size() == 1 && approx(b[0].t, 1.0));
        CHECK(isect(p, mkray(0, 10, 0, 0, 0, 1)).empty());
        CHECK(isect(p, mkray(0, 0, 0, 0, 0, 1)).empty());
        const double pts[3][3] = {{0, 0, 0}, {10, 0, -10}, {-5, 0, 150}};
        for (auto& q : pts) CHECK(approx(p.shape.normal(Vector::point(q[0], q[1], q[2]), false, 0, 0), Vector::vector(0, 1, 0)));
    }
    {  // cube :1137-1190
        Element c = Element::cube(ShapeArgs{});
        struct { RayCase r; double t1, t2; } hits[] = {
            {{{5, 0.5, 0}, {-1, 0, 0}}, 4, 6}, {{{-5, 0.5, 0}, {1, 0, 0}}, 4, 6}, {{{0.5, 5, 0}, {0, -1, 0}}, 4, 6},
            {{{0.5, -5, 0}, {0, 1, 0}}, 4, 6}, {{{0.5, 0, 5}, {0, 0, -1}}, 4, 6}, {{{0.5, 0, -5}, {0, 0, 1}}, 4, 6},
            {{{0, 0.5, 0}, {0, 0, 1}}, -1, 1}};
        for (auto& h : hits) {
            Intersections is = isect(c, ray_of(h.r));
            CHECK(is.size() == 2 && approx(is[0].t, h.t1) && approx(is[1].t, h.t2));
        }
        RayCase miss[] = {{{-2, 0, 0}, {0.2673, 0.5345, 0.8018}}, {{0, -2, 0}, {0.8018, 0.2673, 0.5345}},
                          {{0, 0, -2}, {0.5345, 0.8018, 0.2673}}, {{2, 0, 2}, {0, 0, -1}}, {{0, 2, 2}, {0, -1, 0}}, {{2, 2, 0}, {-1, 0, 0}}};
        for (auto& m : miss) CHECK(isect(c, ray_of(m)).empty());
        const double nrm[8][6] = {{1, 0.5, -0.8, 1, 0, 0},  {-1, -0.2, 0.9, -1, 0, 0}, {-0.4, 1, -0.1, 0, 1, 0}, {0.3, -1, -0.7, 0, -1, 0},
                                  {-0.6, 0.3, 1, 0, 0, 1},  {0.4, 0.4, -1, 0, 0, -1},  {1, 1, 1, 1, 0, 0},       {-1, -1, -1, -1, 0, 0}};
        for (auto& q : nrm) CHECK(approx(c.shape.normal(Vector::point(q[0], q[1], q[2]), false, 0, 0), Vector::vector(q[3], q[4], q[5])));
    }
    {  // cylinder :1194-1310
        Element cyl = Element::cylinder(ShapeArgs{}, -inf, inf, false);
        struct { RayCase r; double t0, t1; } hits[] = {{{{0, 0, -5}, {0, 0, 1}}, 4.0, 6.0}, {{{0.5, 0, -5}, {0.1, 1, 1}}, 6.80798, 7.08872},
                                                       {{{1, 0, -5}, {0, 0, 1}}, 5.0, 5.0}};
        for (auto& h : hits) {
            Intersections is = isect(cyl, ray_of(h.r, true));
            CHECK(is.size() == 2 && approx(is[0].t, h.t0) && approx(is[1].t, h.t1));
        }
        RayCase miss[] = {{{1, 0, 0}, {0, 1, 0}}, {{0, 0, 0}, {0, 1, 0}}, {{1, 0, -5}, {1, 1, 1}}};
        for (auto& m : miss) CHECK(isect(cyl, ray_of(m, true)).empty());
        Element con = Element::cylinder(ShapeArgs{}, 1.0, 2.0, false);
        struct { RayCase r; size_t n; } cons[] = {{{{0, 1.5, 0}, {0.1, 1, 0}}, 0}, {{{0, 3, -5}, {0, 0, 1}}, 0}, {{{0, 0, -5}, {0, 0, 1}}, 0},
                                                  {{{0, 2, -5}, {0, 0, 1}}, 0},   {{{0, 1, -5}, {0, 0, 1}}, 0}, {{{0, 1.5, -5}, {0, 0, 1}}, 2}};
        for (auto& c : cons) CHECK(isect(con, ray_of(c.r, true)).size() == c.n);
        Element clo = Element::cylinder(ShapeArgs{}, 1.0, 2.0, true);
        RayCase caps[] = {{{0, 3, 0}, {0, -1, 0}}, {{0, 3, -2}, {0, -1, 2}}, {{0, 4, -2}, {0, -1, 1}}, {{0, 0, -2}, {0, 1, 2}}, {{0, -1, -2}, {0, 1, 1}}};
        for (auto& c : caps) CHECK(isect(clo, ray_of(c, true)).size() == 2);
        const double nrm[4][6] = {{1, 0, 0, 1, 0, 0}, {0, 5, -1, 0, 0, -1}, {0, -2, 1, 0, 0, 1}, {-1, 1, 0, -1, 0, 0}};
        for (auto& q : nrm) CHECK(approx(cyl.shape.normal(Vector::point(q[0], q[1], q[2]), false, 0, 0), Vector::vector(q[3], q[4], q[5])));
        const double capn[6][6] = {{0, 1, 0, 0, -1, 0}, {0.5, 1, 0, 0, -1, 0}, {0, 1, 0.5, 0, -1, 0},
                                   {0, 2, 0, 0, 1, 0},  {0.5, 2, 0, 0, 1, 0},  {0, 2, 0.5, 0, 1, 0}};
        for (auto& q : capn) CHECK(approx(clo.shape.normal(Vector::point(q[0], q[1], q[2]), false, 0, 0), Vector::vector(q[3], q[4], q[5])));
    }
}

static void test_primitives2() {
    SECTION("shape.rs cone/triangle");
    const double inf = std::numeric_limits<double>::infinity();
    {  // cone :1308-1364
        Element cone = Element::cone(ShapeArgs{}, -inf, inf, false);
        struct { RayCase r; double t0, t1; } hits[] = {{{{0, 0, -5}, {0, 0, 1}}, 5.0, 5.0}, {{{0, 0, -5}, {1, 1, 1}}, 8.66025, 8.66025},
                                                       {{{1, 1, -5}, {-0.5, -1, 1}}, 4.55006, 49.44994}};
        for (auto& h : hits) {
            Intersections is = isect(cone, ray_of(h.r, true));
            CHECK(is.size() == 2 && approx(is[0].t, h.t0) && approx(is[1].t, h.t1));
        }
        Intersections par = isect(cone, ray_of({{0, 0, -1}, {0, 1, 1}}, true));
        CHECK(par.size() == 1 && approx(par[0].t, 0.35355));
        Element closed = Element::cone(ShapeArgs{}, -0.5, 0.5, true);
        struct { RayCase r; size_t n; } cons[] = {{{{0, 0, -5}, {0, 1, 0}}, 0}, {{{0, 0, -0.25}, {0, 1, 1}}, 2}, {{{0, 0, -0.25}, {0, 1, 0}}, 4}};
        for (auto& c : cons) CHECK(isect(closed, ray_of(c.r, true)).size() == c.n);
        Element cn = Element::cone(ShapeArgs{}, -inf, inf, true);
        CHECK(approx(geom::normal(cn.shape.geometry, Vector::point(0, 0, 0), false, 0, 0), Vector::vector(0, 0, 0)));
        CHECK(approx(geom::normal(cn.shape.geometry, Vector::point(1, 1, 1), false, 0, 0), Vector::vector(1, -SQ2, 1)));
        CHECK(approx(geom::normal(cn.shape.geometry, Vector::point(-1, -1, 0), false, 0, 0), Vector::vector(-1, 1, 0)));
    }
    {  // triangle :1368-1400
        Element tri = Element::triangle(ShapeArgs{}, Vector::point(0, 1, 0), Vector::point(-1, 0, 0), Vector::point(1, 0, 0));
        Intersections is = isect(tri, mkray(0, 0.5, -2, 0, 0, 1));
        CHECK(is.size() == 1 && approx(is[0].t, 2.0) && is[0].shape == &tri.shape);
        CHECK(isect(tri, mkray(0, -1, -2, 0, 1, 0)).empty());
        CHECK(isect(tri, mkray(1, 1, -2, 0, 0, 1)).empty());
        CHECK(isect(tri, mkray(-1, 1, -2, 0, 0, 1)).empty());
        CHECK(isect(tri, mkray(0, -1, -2, 0, 0, 1)).empty());
    }
    {  // smooth triangle :1404-1436
        Element st = Element::smooth_triangle(ShapeArgs{}, Vector::point(0, 1, 0), Vector::point(-1, 0, 0), Vector::point(1, 0, 0),
                                              Vector::vector(0, 1, 0), Vector::vector(-1, 0, 0), Vector::vector(1, 0, 0));
        Intersections is = isect(st, mkray(-0.2, 0.3, -2, 0, 0, 1));
        CHECK(is.size() == 1 && is[0].has_uv && std::fabs(is[0].u - 0.45) < 2e-5 && std::fabs(is[0].v - 0.25) < 2e-5);
        CHECK(approx(st.shape.normal(Vector::point(0, 0, 0), true, 0.45, 0.25), Vector::vector(-0.5547, 0.83205, 0.0)));
    }
}

// ------------------------------------------------------------------------------------------
// src/shape.rs:1452-1818 — groups, CSG, lighting
static void test_groups_csg_lighting() {
    SECTION("shape.rs groups/CSG/lighting");
    {  // ray_group_miss :1454-1463
        Element g = Element::composite(Matrix::id(), std::nullopt, GroupKind::Aggregation, {});
        CHECK(isect(g, mkray(0, 0, 0, 0, 0, 1)).empty());
    }
    {  // ray_group_hit :1474-1510
        std::vector<Element> ch;
        ch.push_back(Element::sphere(ShapeArgs{}));
        ch.push_back(Element::sphere(args_t(Matrix::translation(0, 0, -3))));
        ch.push_back(Element::sphere(args_t(Matrix::translation(5, 0, 0))));
        Element g = Element::composite(Matrix::id(), std::nullopt, GroupKind::Aggregation, std::move(ch));
        Intersections is = isect(g, mkray(0, 0, -5, 0, 0, 1));
        sort_intersections(is);
        const Shape* c0 = &g.group.children[0].shape;
        const Shape* c1 = &g.group.children[1].shape;
        CHECK(is.size() == 4 && is[0].shape == c1 && is[1].shape == c1 && is[2].shape == c0 && is[3].shape == c0);
    }
    {  // ray_group_hit_transformed :1512-1531
        std::vector<Element> ch;
        ch.push_back(Element::sphere(args_t(Matrix::translation(5, 0, 0))));
        Element g = Element::composite(Matrix::scaling(2, 2, 2), std::nullopt, GroupKind::Aggregation, std::move(ch));
        CHECK(isect(g, mkray(10, 0, -10, 0, 0, 1)).size() == 2);
    }
    {  // group_normal :1533-1562 (nested groups: rotation_y(pi/2) { scaling(1,2,3) { sphere translated } })
        std::vector<Element> c2;
        c2.push_back(Element::sphere(args_t(Matrix::translation(5, 0, 0))));
        std::vector<Element> c1;
        c1.push_back(Element::composite(Matrix::scaling(1, 2, 3), std::nullopt, GroupKind::Aggregation, std::move(c2)));
        Element g1 = Element::composite(Matrix::rotation_y(scenes::PI / 2.0), std::nullopt, GroupKind::Aggregation, std::move(c1));
        const Shape& s = g1.group.children[0].group.children[0].shape;
        CHECK(approx(s.normal(Vector::point(1.7321, 1.1547, -5.5774), false, 0, 0), Vector::vector(0.285703, 0.428543, -0.857160)));
    }
    {  // group_bbox :1564-1590
        std::vector<Element> ch;
        ch.push_back(Element::sphere(args_t(Matrix::translation(2, 5, -3) * Matrix::scaling(2, 2, 2))));
        ch.push_back(Element::cylinder(args_t(Matrix::translation(-4, -1, 4) * Matrix::scaling(0.5, 1, 0.5)), -2.0, 2.0, true));
        Element g = Element::composite(Matrix::id(), std::nullopt, GroupKind::Aggregation, std::move(ch));
        CHECK(approx(g.bbox().min, Vector::point(-4.5, -3, -5)) && approx(g.bbox().max, Vector::point(4, 7, 4.5)));
    }
    {  // allowed_intersection :1592-1629 (24-row truth table: kind, left_hit, in_left, in_right -> expected)
        const bool T = true, F = false;
        struct { GroupKind k; bool lh, il, ir, ex; } tt[] = {
            {GroupKind::Union, T, T, T, F},        {GroupKind::Union, T, T, F, T},        {GroupKind::Union, T, F, T, F},
            {GroupKind::Union, T, F, F, T},        {GroupKind::Union, F, T, T, F},        {GroupKind::Union, F, T, F, F},
            {GroupKind::Union, F, F, T, T},        {GroupKind::Union, F, F, F, T},        {GroupKind::Intersection, T, T, T, T},
            {GroupKind::Intersection, T, T, F, F}, {GroupKind::Intersection, T, F, T, T}, {GroupKind::Intersection, T, F, F, F},
            {GroupKind::Intersection, F, T, T, T}, {GroupKind::Intersection, F, T, F, T}, {GroupKind::Intersection, F, F, T, F},
            {GroupKind::Intersection, F, F, F, F}, {GroupKind::Difference, T, T, T, F},   {GroupKind::Difference, T, T, F, T},
            {GroupKind::Difference, T, F, T, F},   {GroupKind::Difference, T, F, F, T},   {GroupKind::Difference, F, T, T, T},
            {GroupKind::Difference, F, T, F, T},   {GroupKind::Difference, F, F, T, F},   {GroupKind::Difference, F, F, F, F}};
        for (auto& r : tt) CHECK(allows_intersection(r.k, r.lh, r.il, r.ir) == r.ex);
    }
    {  // filter_by_group :1631-1675
        struct { GroupKind k; size_t i1, i2; } cs[] = {{GroupKind::Union, 0, 3}, {GroupKind::Intersection, 1, 2}, {GroupKind::Difference, 0, 1}};
        for (auto& c : cs) {
            Group g;
            g.kind = c.k;
            g.bbox = BoundingBox::empty();
            g.children.push_back(Element::sphere(ShapeArgs{}));
            g.children.push_back(Element::cube(ShapeArgs{}));
            const Shape* sp = &g.children[0].shape;
            const Shape* cu = &g.children[1].shape;
            Intersections is1 = {mkx(1, sp), mkx(2, cu), mkx(3, sp), mkx(4, cu)};
            Intersections is2 = is1;
            g.filter_by_group(is2);
            CHECK(is2.size() == 2 && same(is2[0], is1[c.i1]) && same(is2[1], is1[c.i2]));
        }
    }
    {  // ray_csg_miss :1677-1689, ray_csg_hit :1691-1722
        std::vector<Element> ch;
        ch.push_back(Element::sphere(ShapeArgs{}));
        ch.push_back(Element::cube(ShapeArgs{}));
        Element g = Element::composite(Matrix::id(), std::nullopt, GroupKind::Union, std::move(ch));
        CHECK(isect(g, mkray(0, 2, -5, 0, 0, 1)).empty());
        std::vector<Element> c2;
        c2.push_back(Element::sphere(ShapeArgs{}));
        c2.push_back(Element::sphere(args_t(Matrix::translation(0, 0, 0.5))));
        Element u = Element::composite(Matrix::id(), std::nullopt, GroupKind::Union, std::move(c2));
        Intersections is = isect(u, mkray(0, 0, -5, 0, 0, 1));
        CHECK(is.size() == 2 && approx(is[0].t, 4.0) && is[0].shape == &u.group.children[0].shape && approx(is[1].t, 6.5) &&
              is[1].shape == &u.group.children[1].shape);
    }
    {  // lightning :1726-1786 (6 Phong configurations)
        Element s = Element::sphere(ShapeArgs{});
        struct { double eye[3], lo[3]; bool shadowed; double ex; } cs[] = {
            {{0, 0, -1}, {0, 0, -10}, false, 1.9},  {{0, SQ2 / 2.0, SQ2 / -2.0}, {0, 0, -10}, false, 1.0},
            {{0, 0, -1}, {0, 10, -10}, false, 0.7364}, {{0, SQ2 / -2.0, SQ2 / -2.0}, {0, 10, -10}, false, 1.6364},
            {{0, 0, -1}, {0, 0, 10}, false, 0.1},   {{0, 0, -1}, {0, 0, -10}, true, 0.1}};
        for (auto& c : cs) {
            PointLight l{Color::white(), Vector::point(c.lo[0], c.lo[1], c.lo[2])};
            Color r = s.shape.lighting(l, Vector::point(0, 0, 0), Vector::vector(c.eye[0], c.eye[1], c.eye[2]), Vector::vector(0, 0, -1), c.shadowed);
            CHECK(approx(r, Color::make(c.ex, c.ex, c.ex)));
        }
    }
    {  // lighting_with_stripe_pattern :1788-1817
        ShapeArgs a;
        a.material.pattern = Pattern::stripes(Matrix::id(), Pattern::plain(Color::white()), Pattern::plain(Color::black()));
        a.material.ambient = 1.0;
        a.material.diffuse = 0.0;
        a.material.specular = 0.0;
        Element s = Element::sphere(a);
        PointLight l{Color::white(), Vector::point(0, 0, -10)};
        Vector eye = Vector::vector(0, 0, -1), n = Vector::vector(0, 0, -1);
        CHECK(approx(s.shape.lighting(l, Vector::point(0.9, 0, 0), eye, n, false), Color::white()));
        CHECK(approx(s.shape.lighting(l, Vector::point(1.1, 0, 0), eye, n, false), Color::black()));
    }
}

// ------------------------------------------------------------------------------------------
// src/material.rs:312-396 — patterns
static void test_patterns() {
    SECTION("material.rs");
    Pattern w = Pattern::plain(Color::white()), b = Pattern::plain(Color::black());
    {  // stripes :312-345
        Matrix id = Matrix::id(), s2 = Matrix::scaling(2, 2, 2);
        struct { Matrix t; Vector p; Color ex; } cs[] = {
            {id, Vector::point(0, 1, 0), Color::white()},    {id, Vector::point(0, 2, 0), Color::white()},
            {id, Vector::point(0, 0, 1), Color::white()},    {id, Vector::point(0, 0, 2), Color::white()},
            {id, Vector::point(0.9, 0, 0), Color::white()},  {id, Vector::point(1.0, 0, 0), Color::black()},
            {id, Vector::point(-0.1, 0, 0), Color::black()}, {id, Vector::point(-1.0, 0, 0), Color::black()},
            {id, Vector::point(-1.1, 0, 0), Color::white()}, {id, s2.inverse() * Vector::point(1.5, 0, 0), Color::white()},
            {s2, Vector::point(1.5, 0, 0), Color::white()},  {s2, s2.inverse() * Vector::point(2.5, 0, 0), Color::white()}};
        for (auto& c : cs) CHECK(approx(Pattern::stripes(c.t, w, b).color_at(c.p), c.ex));
    }
    {  // gradient :347-360
        Pattern g = Pattern::gradient(Matrix::id(), w, b);
        const double cs[4][2] = {{0.0, 1.0}, {0.25, 0.75}, {0.5, 0.5}, {0.75, 0.25}};
        for (auto& c : cs) CHECK(approx(g.color_at(Vector::point(c[0], 0, 0)), Color::make(c[1], c[1], c[1])));
    }
    {  // ring :362-375
        Pattern r = Pattern::ring(Matrix::id(), w, b);
        CHECK(approx(r.color_at(Vector::point(0, 0, 0)), Color::white()));
        CHECK(approx(r.color_at(Vector::point(1, 0, 0)), Color::black()));
        CHECK(approx(r.color_at(Vector::point(0, 0, 1)), Color::black()));
        CHECK(approx(r.color_at(Vector::point(0.708, 0, 0.708)), Color::black()));
    }
    {  // checkers :377-392
        Pattern c = Pattern::checkers(Matrix::id(), w, b);
        CHECK(approx(c.color_at(Vector::point(0.99, 0, 0)), Color::white()));
        CHECK(approx(c.color_at(Vector::point(1.01, 0, 0)), Color::black()));
        CHECK(approx(c.color_at(Vector::point(0, 0.99, 0)), Color::white()));
        CHECK(approx(c.color_at(Vector::point(0, 1.01, 0)), Color::black()));
        CHECK(approx(c.color_at(Vector::point(0, 0, 0.99)), Color::white()));
        CHECK(approx(c.color_at(Vector::point(0, 0, 1.01)), Color::black()));
    }
}

// ------------------------------------------------------------------------------------------
// src/camera.rs:85-161, src/image.rs:128-146
static void test_camera_image() {
    SECTION("camera.rs / image.rs");
    {  // view_transform :85-121
        CHECK(approx(Camera::transform(Vector::point(0, 0, 0), Vector::point(0, 0, -1), Vector::vector(0, 1, 0)), Matrix::id()));
        CHECK(approx(Camera::transform(Vector::point(0, 0, 0), Vector::point(0, 0, 1), Vector::vector(0, 1, 0)), Matrix::scaling(-1, 1, -1)));
        CHECK(approx(Camera::transform(Vector::point(0, 0, 8), Vector::point(0, 0, 0), Vector::vector(0, 1, 0)), Matrix::translation(0, 0, -8)));
        double m[4][4] = {{-0.50709, 0.50709, 0.67612, -2.36643}, {0.76772, 0.60609, 0.12122, -2.82843}, {-0.35857, 0.59761, -0.71714, 0.0}, {0, 0, 0, 1}};
        CHECK(approx(Camera::transform(Vector::point(1, 3, 2), Vector::point(4, -2, 8), Vector::vector(1, 1, 0)), Matrix::make(m)));
    }
    {  // pixel_size :123-128
        CHECK(approx(Camera::make(200, 125, scenes::PI / 2.0, Matrix::id()).pixel_size, 0.01));
        CHECK(approx(Camera::make(125, 200, scenes::PI / 2.0, Matrix::id()).pixel_size, 0.01));
    }
    {  // ray_at_pixel :130-160
        Camera c = Camera::make(201, 101, scenes::PI / 2.0, Matrix::id());
        Ray r = c.ray_at_pixel(100, 50);
        CHECK(approx(r.origin, Vector::point(0, 0, 0)) && approx(r.direction, Vector::vector(0, 0, -1)));
        r = c.ray_at_pixel(0, 0);
        CHECK(approx(r.origin, Vector::point(0, 0, 0)) && approx(r.direction, Vector::vector(0.66519, 0.33259, -0.66851)));
        Camera t = Camera::make(201, 101, scenes::PI / 2.0, Matrix::rotation_y(scenes::PI / 4.0) * Matrix::translation(0, -2, 5));
        r = t.ray_at_pixel(100, 50);
        CHECK(approx(r.origin, Vector::point(0, 2, -5)) && approx(r.direction, Vector::vector(SQ2 / 2.0, 0, SQ2 / -2.0)));
    }
    {  // rendering_default_world src/image.rs:128-146
        auto sc = scenes::default_world<OracleApi>(11, 11);
        Image img = Image::par_render(sc.camera, sc.world);
        CHECK(approx(img.read(5, 5), Color::make(0.38066, 0.47583, 0.28550)));
    }
}

// ------------------------------------------------------------------------------------------
// src/bounding_box.rs:103-225
static void test_bbox() {
    SECTION("bounding_box.rs");
    {
        BoundingBox b = BoundingBox::empty().insert(Vector::point(-5, 2, 0)).insert(Vector::point(7, 0, -3));
        CHECK(approx(b.min, Vector::point(-5, 0, -3)) && approx(b.max, Vector::point(7, 2, 0)));
        BoundingBox u = BoundingBox::make(Vector::point(-5, -2, 0), Vector::point(7, 4, 4)).union_(BoundingBox::make(Vector::point(8, -7, -2), Vector::point(14, 2, 8)));
        CHECK(approx(u.min, Vector::point(-5, -7, -2)) && approx(u.max, Vector::point(14, 4, 8)));
    }
    BoundingBox nb = BoundingBox::make(Vector::point(5, -2, 0), Vector::point(11, 4, 7));
    {
        struct { double p[3]; bool ex; } cs[] = {{{5, -2, 0}, true}, {{11, 4, 7}, true}, {{8, 1, 3}, true}, {{3, 0, 3}, false}, {{8, -4, 3}, false},
                                                 {{8, 1, -1}, false}, {{13, 1, 3}, false}, {{8, 5, 3}, false}, {{8, 1, 8}, false}};
        for (auto& c : cs) CHECK(nb.contains(Vector::point(c.p[0], c.p[1], c.p[2])) == c.ex);
        CHECK(nb.encloses(BoundingBox::make(Vector::point(5, -2, 0), Vector::point(11, 4, 7))));
        CHECK(nb.encloses(BoundingBox::make(Vector::point(6, -1, 1), Vector::point(10, 3, 6))));
        CHECK(!nb.encloses(BoundingBox::make(Vector::point(4, -3, -1), Vector::point(10, 3, 6))));
        CHECK(!nb.encloses(BoundingBox::make(Vector::point(6, -1, 1), Vector::point(12, 5, 8))));
    }
    {
        BoundingBox b = BoundingBox::make(Vector::point(-1, -1, -1), Vector::point(1, 1, 1))
                            .transform(Matrix::rotation_x(scenes::PI / 4.0) * Matrix::rotation_y(scenes::PI / 4.0));
        CHECK(approx(b.min, Vector::point(-1.414213, -1.707106, -1.707106)) && approx(b.max, Vector::point(1.414213, 1.707106, 1.707106)));
    }
    {
        BoundingBox cb = BoundingBox::make(Vector::point(-1, -1, -1), Vector::point(1, 1, 1));
        struct { RayCase r; bool ex; } cubic[] = {
            {{{5, 0.5, 0}, {-1, 0, 0}}, true},  {{{-5, 0.5, 0}, {1, 0, 0}}, true},  {{{0.5, 5, 0}, {0, -1, 0}}, true},  {{{0.5, -5, 0}, {0, 1, 0}}, true},
            {{{0.5, 0, 5}, {0, 0, -1}}, true},  {{{0.5, 0, -5}, {0, 0, 1}}, true},  {{{0, 0.5, 0}, {0, 0, 1}}, true},   {{{-2, 0, 0}, {2, 4, 6}}, false},
            {{{0, -2, 0}, {6, 2, 4}}, false},   {{{0, 0, -2}, {4, 6, 2}}, false},   {{{2, 0, 2}, {0, 0, -1}}, false},   {{{0, 2, 2}, {0, -1, 0}}, false},
            {{{2, 2, 0}, {-1, 0, 0}}, false}};
        for (auto& c : cubic) CHECK(cb.intersects(ray_of(c.r, true)) == c.ex);
        struct { RayCase r; bool ex; } non[] = {
            {{{15, 1, 2}, {-1, 0, 0}}, true},  {{{-5, -1, 4}, {1, 0, 0}}, true},  {{{7, 6, 5}, {0, -1, 0}}, true},   {{{9, -5, 6}, {0, 1, 0}}, true},
            {{{8, 2, 12}, {0, 0, -1}}, true},  {{{6, 0, -5}, {0, 0, 1}}, true},   {{{8, 1, 3.5}, {0, 0, 1}}, true},  {{{9, -1, -8}, {2, 4, 6}}, false},
            {{{8, 3, -4}, {6, 2, 4}}, false},  {{{9, -1, -2}, {4, 6, 2}}, false}, {{{4, 0, 9}, {0, 0, -1}}, false},  {{{8, 6, -1}, {0, -1, 0}}, false},
            {{{12, 5, 4}, {-1, 0, 0}}, false}};
        for (auto& c : non) CHECK(nb.intersects(ray_of(c.r, true)) == c.ex);
    }
}

// ------------------------------------------------------------------------------------------
// src/linalg/matrix.rs:486-750 (determinant, inverse, transforms), src/color.rs clamp, src/ray.rs
static void test_linalg() {
    SECTION("linalg / color / ray");
    {
        double d[4][4] = {{-2, -8, 3, 5}, {-3, 1, 7, 3}, {1, 2, -9, 6}, {-6, 7, 7, -9}};
        CHECK(approx(Matrix::make(d).determinant(), -4071.0));
        double m1[4][4] = {{-5, 2, 6, -8}, {1, -5, 1, 8}, {7, 7, -6, -7}, {1, -3, 7, 4}};
        double e1[4][4] = {{0.21805, 0.45113, 0.24060, -0.04511}, {-0.80827, -1.45677, -0.44361, 0.52068}, {-0.07895, -0.22368, -0.05263, 0.19737}, {-0.52256, -0.81391, -0.30075, 0.30639}};
        CHECK(approx(Matrix::make(m1).inverse(), Matrix::make(e1)));
        double m2[4][4] = {{8, -5, 9, 2}, {7, 5, 6, 1}, {-6, 0, 9, 6}, {-3, 0, -9, -4}};
        double e2[4][4] = {{-0.15385, -0.15385, -0.28205, -0.53846}, {-0.07692, 0.12308, 0.02564, 0.03077}, {0.35897, 0.35897, 0.43590, 0.92308}, {-0.69231, -0.69231, -0.76923, -1.92308}};
        CHECK(approx(Matrix::make(m2).inverse(), Matrix::make(e2)));
        double m3[4][4] = {{9, 3, 0, 9}, {-5, -2, -6, -3}, {-4, 9, 6, 4}, {-7, 6, 6, 2}};
        double e3[4][4] = {{-0.04074, -0.07778, 0.14444, -0.22222}, {-0.07778, 0.03333, 0.36667, -0.33333}, {-0.02901, -0.14630, -0.10926, 0.12963}, {0.17778, 0.06667, -0.26667, 0.33333}};
        CHECK(approx(Matrix::make(m3).inverse(), Matrix::make(e3)));
        CHECK(approx(Matrix::make(m1) * Matrix::make(m2) * Matrix::make(m2).inverse(), Matrix::make(m1)));
    }
    {  // transform_sequence / transform_chaining :720-750
        Vector p1 = Vector::point(1, 0, 1);
        Vector p2 = Matrix::rotation_x(scenes::PI / 2.0) * p1;
        CHECK(approx(p2, Vector::point(1, -1, 0)));
        Vector p3 = Matrix::scaling(5, 5, 5) * p2;
        CHECK(approx(p3, Vector::point(5, -5, 0)));
        CHECK(approx(Matrix::translation(10, 5, 7) * p3, Vector::point(15, 0, 7)));
        Matrix chain = Matrix::id().rotate_x(scenes::PI / 2.0).scale(5, 5, 5).translate(10, 5, 7);
        CHECK(approx(chain * p1, Vector::point(15, 0, 7)));
    }
    {  // src/ray.rs:28-53
        Ray r = mkray(2, 3, 4, 1, 0, 0);
        CHECK(approx(r.position(0.0), Vector::point(2, 3, 4)) && approx(r.position(1.0), Vector::point(3, 3, 4)) &&
              approx(r.position(-1.0), Vector::point(1, 3, 4)) && approx(r.position(2.5), Vector::point(4.5, 3, 4)));
        Ray q = mkray(1, 2, 3, 0, 1, 0);
        Ray a = q.transform(Matrix::translation(3, 4, 5));
        CHECK(approx(a.origin, Vector::point(4, 6, 8)) && approx(a.direction, Vector::vector(0, 1, 0)));
        Ray b = q.transform(Matrix::scaling(2, 3, 4));
        CHECK(approx(b.origin, Vector::point(2, 6, 12)) && approx(b.direction, Vector::vector(0, 3, 0)));
    }
    {  // src/color.rs:42-46 (round half away from zero, clamp)
        CHECK(clamp_channel(-0.5) == 0 && clamp_channel(1.5) == 255 && clamp_channel(0.5) == 128 && clamp_channel(1.0) == 255);
        CHECK(clamp_channel(0.1 / 255.0 + 0.4 / 255.0) == 1 || clamp_channel(0.5 / 255.0) == 1);
    }
}

int main() {
    test_world();
    test_intersection();
    test_primitives();
    test_primitives2();
    test_groups_csg_lighting();
    test_patterns();
    test_camera_image();
    test_bbox();
    test_linalg();
    std::printf("oracle known-answer checks: %d run, %d failed\n", g_checks, g_failed);
    return g_failed ? 1 : 0;
}
