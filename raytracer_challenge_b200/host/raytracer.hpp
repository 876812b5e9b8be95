// Host-side builder API of the B200 render path (product code; C++ because no Rust toolchain
// exists in this image).  It mirrors, name for name, the builder surface of the reference crate
// that its example binaries use:
//   Matrix / Vector / Color            src/linalg/matrix.rs, src/linalg/vector.rs, src/color.rs
//   Material / Pattern / Noise         src/material.rs:18-162, src/noise.rs:4-7
//   Element / ShapeArgs / GroupKind    src/shape.rs:74-137,161-167,272-286
//   PointLight / World / Camera        src/light.rs, src/world.rs:12-15,152-183, src/camera.rs:16-37,57-73
//   Image::par_render / read / ppm     src/image.rs:65-112
//   ObjParser                          src/obj.rs:152-258
// Build-time arithmetic (matrix products, cofactor inverse, box transforms, inverse propagation)
// follows the reference's operation order so the flattened matrices are bit-identical to what the
// Rust crate would hold; all per-pixel work happens on the GPU behind include/rtc.h.
#pragma once

#include <algorithm>
#include <array>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <fstream>
#include <limits>
#include <memory>
#include <optional>
#include <sstream>
#include <stdexcept>
#include <string>
#include <utility>
#include <variant>
#include <vector>

#include "../../include/rtc.h"

namespace rtc {

constexpr double EPSILON = 0.00001;  // src/config.rs:1
constexpr int FUEL = 5;              // src/config.rs:2

// ---- Vector (src/linalg/vector.rs) -------------------------------------------------------
struct Vector {
    double x = 0, y = 0, z = 0, w = 0;
    static Vector point(double x, double y, double z) { return Vector{x, y, z, 1.0}; }
    static Vector vector(double x, double y, double z) { return Vector{x, y, z, 0.0}; }
    double magnitude() const { return std::sqrt(x * x + y * y + z * z); }
    Vector normalize() const {
        const double len = magnitude();
        return Vector{x / len, y / len, z / len, 0.0};
    }
    double dot(const Vector& b) const { return x * b.x + y * b.y + z * b.z; }
    Vector cross(const Vector& b) const { return vector(y * b.z - z * b.y, z * b.x - x * b.z, x * b.y - y * b.x); }
    Vector operator+(const Vector& b) const { return Vector{x + b.x, y + b.y, z + b.z, w + b.w}; }
    Vector operator-(const Vector& b) const { return Vector{x - b.x, y - b.y, z - b.z, w - b.w}; }
    Vector operator-() const { return Vector{-x, -y, -z, -w}; }
    Vector operator*(double s) const { return Vector{x * s, y * s, z * s, w * s}; }
    Vector operator/(double s) const { return Vector{x / s, y / s, z / s, w / s}; }
    Vector reflect(const Vector& n) const { return *this - n * (2.0 * dot(n)); }
};

// ---- Matrix (src/linalg/matrix.rs) — row-major 4x4 stored flat --------------------------
struct Matrix {
    std::array<double, 16> m{};
    double& at(int r, int c) { return m[(size_t)(r * 4 + c)]; }
    double at(int r, int c) const { return m[(size_t)(r * 4 + c)]; }

    static Matrix make(const double (&rows)[4][4]) {
        Matrix out;
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) out.at(r, c) = rows[r][c];
        return out;
    }
    static Matrix id() {
        Matrix out;
        out.at(0, 0) = out.at(1, 1) = out.at(2, 2) = out.at(3, 3) = 1.0;
        return out;
    }
    static Matrix translation(double x, double y, double z) {
        Matrix out = id();
        out.at(0, 3) = x;
        out.at(1, 3) = y;
        out.at(2, 3) = z;
        return out;
    }
    static Matrix scaling(double x, double y, double z) {
        Matrix out = id();
        out.at(0, 0) = x;
        out.at(1, 1) = y;
        out.at(2, 2) = z;
        return out;
    }
    static Matrix rotation_x(double r) {
        Matrix out = id();
        out.at(1, 1) = std::cos(r);
        out.at(1, 2) = -std::sin(r);
        out.at(2, 1) = std::sin(r);
        out.at(2, 2) = std::cos(r);
        return out;
    }
    static Matrix rotation_y(double r) {
        Matrix out = id();
        out.at(0, 0) = std::cos(r);
        out.at(0, 2) = std::sin(r);
        out.at(2, 0) = -std::sin(r);
        out.at(2, 2) = std::cos(r);
        return out;
    }
    static Matrix rotation_z(double r) {
        Matrix out = id();
        out.at(0, 0) = std::cos(r);
        out.at(0, 1) = -std::sin(r);
        out.at(1, 0) = std::sin(r);
        out.at(1, 1) = std::cos(r);
        return out;
    }
    static Matrix shearing(double x_y, double x_z, double y_x, double y_z, double z_x, double z_y) {
        Matrix out = id();
        out.at(0, 1) = x_y;
        out.at(0, 2) = x_z;
        out.at(1, 0) = y_x;
        out.at(1, 2) = y_z;
        out.at(2, 0) = z_x;
        out.at(2, 1) = z_y;
        return out;
    }

    // src/linalg/matrix.rs:239-259: running sum from 0.0 over i = 0..3
    Matrix operator*(const Matrix& b) const {
        Matrix out;
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) {
                double acc = 0.0;
                for (int i = 0; i < 4; ++i) acc += at(r, i) * b.at(i, c);
                out.at(r, c) = acc;
            }
        return out;
    }
    // src/linalg/matrix.rs:261-284
    Vector operator*(const Vector& v) const {
        Vector out;
        out.x = at(0, 0) * v.x + at(0, 1) * v.y + at(0, 2) * v.z + at(0, 3) * v.w;
        out.y = at(1, 0) * v.x + at(1, 1) * v.y + at(1, 2) * v.z + at(1, 3) * v.w;
        out.z = at(2, 0) * v.x + at(2, 1) * v.y + at(2, 2) * v.z + at(2, 3) * v.w;
        out.w = at(3, 0) * v.x + at(3, 1) * v.y + at(3, 2) * v.z + at(3, 3) * v.w;
        return out;
    }
    // fluent forms pre-multiply (src/linalg/matrix.rs:102-124)
    Matrix translate(double x, double y, double z) const { return translation(x, y, z) * *this; }
    Matrix scale(double x, double y, double z) const { return scaling(x, y, z) * *this; }
    Matrix rotate_x(double r) const { return rotation_x(r) * *this; }
    Matrix rotate_y(double r) const { return rotation_y(r) * *this; }
    Matrix rotate_z(double r) const { return rotation_z(r) * *this; }
    Matrix shear(double a, double b, double c, double d, double e, double f) const { return shearing(a, b, c, d, e, f) * *this; }

    Matrix transpose() const {
        Matrix out;
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) out.at(c, r) = at(r, c);
        return out;
    }

    // The twelve 2x2 sub-determinants of the top and bottom row pairs (src/linalg/matrix.rs:141-154,165-178).
    struct Minors { double s[6], c[6]; };
    Minors minors() const {
        Minors k;
        k.s[0] = at(0, 0) * at(1, 1) - at(1, 0) * at(0, 1);
        k.s[1] = at(0, 0) * at(1, 2) - at(1, 0) * at(0, 2);
        k.s[2] = at(0, 0) * at(1, 3) - at(1, 0) * at(0, 3);
        k.s[3] = at(0, 1) * at(1, 2) - at(1, 1) * at(0, 2);
        k.s[4] = at(0, 1) * at(1, 3) - at(1, 1) * at(0, 3);
        k.s[5] = at(0, 2) * at(1, 3) - at(1, 2) * at(0, 3);
        k.c[5] = at(2, 2) * at(3, 3) - at(3, 2) * at(2, 3);
        k.c[4] = at(2, 1) * at(3, 3) - at(3, 1) * at(2, 3);
        k.c[3] = at(2, 1) * at(3, 2) - at(3, 1) * at(2, 2);
        k.c[2] = at(2, 0) * at(3, 3) - at(3, 0) * at(2, 3);
        k.c[1] = at(2, 0) * at(3, 2) - at(3, 0) * at(2, 2);
        k.c[0] = at(2, 0) * at(3, 1) - at(3, 0) * at(2, 1);
        return k;
    }
    static double det_of(const Minors& k) {
        return k.s[0] * k.c[5] - k.s[1] * k.c[4] + k.s[2] * k.c[3] + k.s[3] * k.c[2] - k.s[4] * k.c[1] + k.s[5] * k.c[0];
    }
    double determinant() const { return det_of(minors()); }

    // Cofactor inverse with the reference's term order (src/linalg/matrix.rs:162-207).
    Matrix inverse() const {
        const Minors k = minors();
        const double* s = k.s;
        const double* c = k.c;
        const double det = det_of(k);
        if (!(det != 0.0)) throw std::runtime_error("Matrix::inverse: determinant is zero");
        Matrix o;
        o.at(0, 0) = (at(1, 1) * c[5] - at(1, 2) * c[4] + at(1, 3) * c[3]) / det;
        o.at(0, 1) = (-at(0, 1) * c[5] + at(0, 2) * c[4] - at(0, 3) * c[3]) / det;
        o.at(0, 2) = (at(3, 1) * s[5] - at(3, 2) * s[4] + at(3, 3) * s[3]) / det;
        o.at(0, 3) = (-at(2, 1) * s[5] + at(2, 2) * s[4] - at(2, 3) * s[3]) / det;
        o.at(1, 0) = (-at(1, 0) * c[5] + at(1, 2) * c[2] - at(1, 3) * c[1]) / det;
        o.at(1, 1) = (at(0, 0) * c[5] - at(0, 2) * c[2] + at(0, 3) * c[1]) / det;
        o.at(1, 2) = (-at(3, 0) * s[5] + at(3, 2) * s[2] - at(3, 3) * s[1]) / det;
        o.at(1, 3) = (at(2, 0) * s[5] - at(2, 2) * s[2] + at(2, 3) * s[1]) / det;
        o.at(2, 0) = (at(1, 0) * c[4] - at(1, 1) * c[2] + at(1, 3) * c[0]) / det;
        o.at(2, 1) = (-at(0, 0) * c[4] + at(0, 1) * c[2] - at(0, 3) * c[0]) / det;
        o.at(2, 2) = (at(3, 0) * s[4] - at(3, 1) * s[2] + at(3, 3) * s[0]) / det;
        o.at(2, 3) = (-at(2, 0) * s[4] + at(2, 1) * s[2] - at(2, 3) * s[0]) / det;
        o.at(3, 0) = (-at(1, 0) * c[3] + at(1, 1) * c[1] - at(1, 2) * c[0]) / det;
        o.at(3, 1) = (at(0, 0) * c[3] - at(0, 1) * c[1] + at(0, 2) * c[0]) / det;
        o.at(3, 2) = (-at(3, 0) * s[3] + at(3, 1) * s[1] - at(3, 2) * s[0]) / det;
        o.at(3, 3) = (at(2, 0) * s[3] - at(2, 1) * s[1] + at(2, 2) * s[0]) / det;
        return o;
    }
};

// ---- Color (src/color.rs) ----------------------------------------------------------------
struct Color {
    double r = 0, g = 0, b = 0;
    static Color make(double r, double g, double b) { return Color{r, g, b}; }  // Color::new
    static Color rgb(uint8_t r, uint8_t g, uint8_t b) { return Color{r / 255.0, g / 255.0, b / 255.0}; }
    static Color white() { return Color{1.0, 1.0, 1.0}; }
    static Color black() { return Color{0.0, 0.0, 0.0}; }
    // src/color.rs:42-46
    std::array<uint8_t, 3> clamp() const {
        uint8_t q[3];
        const double v[3] = {r, g, b};
        rtc_quantize(v, q, 3);
        return {q[0], q[1], q[2]};
    }
    Color operator+(const Color& o) const { return Color{r + o.r, g + o.g, b + o.b}; }
    Color operator-(const Color& o) const { return Color{r - o.r, g - o.g, b - o.b}; }
    Color operator*(double s) const { return Color{r * s, g * s, b * s}; }
    Color operator*(const Color& o) const { return Color{r * o.r, g * o.g, b * o.b}; }
};

// ---- Noise / Pattern / Material (src/noise.rs:4-7, src/material.rs) -----------------------
struct Noise {
    int kind = RTC_NOISE_SIMPLEX;
    double scale = 0.0;
    size_t octaves = 0;
    static Noise Simplex(double scale) { return Noise{RTC_NOISE_SIMPLEX, scale, 0}; }
    static Noise Fractal(double scale, size_t octaves) { return Noise{RTC_NOISE_FRACTAL, scale, octaves}; }
};

// A pattern is an immutable tree (the reference boxes its children, src/material.rs:60-65).
class Pattern {
  public:
    struct Node {
        int kind = RTC_PAT_PLAIN;  // rtc_pattern_kind
        Color color;               // plain
        Noise noise;               // jitter
        Matrix transform_inv = Matrix::id();  // mixture: inverse of the given transform (src/material.rs:134)
        std::shared_ptr<const Node> left, right;
    };
    Pattern() : node_(leaf(RTC_PAT_PLAIN, Color::white())) {}
    static Pattern Debug() { return Pattern(leaf(RTC_PAT_DEBUG, Color::black())); }
    static Pattern plain(Color c) { return Pattern(leaf(RTC_PAT_PLAIN, c)); }
    static Pattern color_jitter(Noise n, Pattern inner) { return jitter(RTC_PAT_JITTER_COLOR, n, inner); }
    static Pattern point_jitter(Noise n, Pattern inner) { return jitter(RTC_PAT_JITTER_POINT, n, inner); }
    static Pattern blend(const Matrix& t, Pattern l, Pattern r) { return mixture(RTC_PAT_BLEND, t, l, r); }
    static Pattern checkers(const Matrix& t, Pattern l, Pattern r) { return mixture(RTC_PAT_CHECKERS, t, l, r); }
    static Pattern ring_gradient(const Matrix& t, Pattern l, Pattern r) { return mixture(RTC_PAT_RING_GRADIENT, t, l, r); }
    static Pattern ring(const Matrix& t, Pattern l, Pattern r) { return mixture(RTC_PAT_RING, t, l, r); }
    static Pattern gradient(const Matrix& t, Pattern l, Pattern r) { return mixture(RTC_PAT_GRADIENT, t, l, r); }
    static Pattern stripes(const Matrix& t, Pattern l, Pattern r) { return mixture(RTC_PAT_STRIPES, t, l, r); }
    const std::shared_ptr<const Node>& root() const { return node_; }

  private:
    explicit Pattern(std::shared_ptr<const Node> n) : node_(std::move(n)) {}
    static std::shared_ptr<const Node> leaf(int kind, Color c) {
        auto n = std::make_shared<Node>();
        n->kind = kind;
        n->color = c;
        return n;
    }
    static Pattern jitter(int kind, Noise noise, const Pattern& inner) {
        auto n = std::make_shared<Node>();
        n->kind = kind;
        n->noise = noise;
        n->left = inner.node_;
        return Pattern(n);
    }
    static Pattern mixture(int kind, const Matrix& transform, const Pattern& l, const Pattern& r) {
        auto n = std::make_shared<Node>();
        n->kind = kind;
        n->transform_inv = transform.inverse();
        n->left = l.node_;
        n->right = r.node_;
        return Pattern(n);
    }
    std::shared_ptr<const Node> node_;
};

namespace consts { namespace transparency {
constexpr double VACUUM = 1.0, AIR = 1.00029, WATER = 1.333, GLASS = 1.52, DIAMOND = 2.417;
}}

struct Material {  // defaults: src/material.rs:30-43
    Pattern pattern = Pattern::plain(Color::white());
    double ambient = 0.1;
    double diffuse = 0.9;
    double specular = 0.9;
    double shininess = 200.0;
    double reflective = 0.0;
    double transparency = 0.0;
    double refractive_index = 1.0;
};

struct PointLight {  // src/light.rs:4-8
    Color intensity;
    Vector origin;
};

// ---- BoundingBox (build-time half of src/bounding_box.rs:19-78) ---------------------------
// f64::min / f64::max return the other operand when one is NaN — std::fmin / std::fmax do too.
struct BoundingBox {
    Vector min = Vector::point(std::numeric_limits<double>::infinity(), std::numeric_limits<double>::infinity(),
                               std::numeric_limits<double>::infinity());
    Vector max = Vector::point(-std::numeric_limits<double>::infinity(), -std::numeric_limits<double>::infinity(),
                               -std::numeric_limits<double>::infinity());
    static BoundingBox empty() { return BoundingBox{}; }
    static BoundingBox make(const Vector& lo, const Vector& hi) { BoundingBox b; b.min = lo; b.max = hi; return b; }
    BoundingBox insert(const Vector& p) const {
        BoundingBox b;
        b.min = Vector::point(std::fmin(min.x, p.x), std::fmin(min.y, p.y), std::fmin(min.z, p.z));
        b.max = Vector::point(std::fmax(max.x, p.x), std::fmax(max.y, p.y), std::fmax(max.z, p.z));
        return b;
    }
    BoundingBox union_(const BoundingBox& o) const { return insert(o.min).insert(o.max); }
    // corners in the reference's order: z fastest, then y, then x (src/bounding_box.rs:63-70)
    BoundingBox transform(const Matrix& m) const {
        BoundingBox out;
        for (int i = 0; i < 8; ++i) {
            const Vector corner = Vector::point((i & 4) ? max.x : min.x, (i & 2) ? max.y : min.y, (i & 1) ? max.z : min.z);
            out = out.insert(m * corner);
        }
        return out;
    }
};

// ---- Shapes, groups, elements (src/shape.rs) ----------------------------------------------
enum class GroupKind { Union = RTC_UNION, Intersection = RTC_INTERSECTION, Difference = RTC_DIFFERENCE, Aggregation = RTC_AGGREGATION };

struct ShapeArgs {  // src/shape.rs:272-286
    Matrix transform = Matrix::id();
    Material material = Material{};
    bool casts_shadow = true;
};

struct Shape {  // src/shape.rs:296-306 (identity = position in the flattened scene)
    int kind = RTC_SPHERE;
    Matrix transform_inv, transform_inv_tsp, material_inv;
    BoundingBox bbox;
    Material material;
    bool casts_shadow = true;
    double min = 0.0, max = 0.0;
    bool closed = false;
    Vector p1, p2, p3, e1, e2, n1, n2, n3;  // flat triangle: n1 = face normal
};

class Element;
struct Group {  // src/shape.rs:180-185
    GroupKind kind = GroupKind::Aggregation;
    BoundingBox bbox;
    std::vector<Element> children;
};

class Element {
  public:
    std::variant<Shape, Group> node;

    bool is_group() const { return std::holds_alternative<Group>(node); }
    BoundingBox bbox() const { return is_group() ? std::get<Group>(node).bbox : std::get<Shape>(node).bbox; }

    static Element sphere(const ShapeArgs& a) { return prim(a, RTC_SPHERE, unit_box()); }
    static Element cube(const ShapeArgs& a) { return prim(a, RTC_CUBE, unit_box()); }
    static Element plane(const ShapeArgs& a) {
        const double inf = std::numeric_limits<double>::infinity();
        return prim(a, RTC_PLANE, BoundingBox::make(Vector::point(-inf, 0.0, -inf), Vector::point(inf, 0.0, inf)));
    }
    static Element cylinder(const ShapeArgs& a, double min, double max, bool closed) {
        const double inf = std::numeric_limits<double>::infinity();
        // an OPEN cylinder's reference box is infinite in y even for finite min/max (src/shape.rs:962-973)
        BoundingBox b = closed ? BoundingBox::make(Vector::point(-1.0, min, -1.0), Vector::point(1.0, max, 1.0))
                               : BoundingBox::make(Vector::point(-1.0, -inf, -1.0), Vector::point(1.0, inf, 1.0));
        return prim(a, RTC_CYLINDER, b, min, max, closed);
    }
    static Element cone(const ShapeArgs& a, double min, double max, bool closed) {
        const double inf = std::numeric_limits<double>::infinity();
        BoundingBox b;
        if (closed) {
            const double limit = std::fmax(std::fabs(min), std::fabs(max));
            b = BoundingBox::make(Vector::point(-limit, min, -limit), Vector::point(limit, max, limit));
        } else {
            b = BoundingBox::make(Vector::point(-inf, -inf, -inf), Vector::point(inf, inf, inf));
        }
        return prim(a, RTC_CONE, b, min, max, closed);
    }
    static Element triangle(const ShapeArgs& a, Vector p1, Vector p2, Vector p3) {
        Shape s;
        s.p1 = p1; s.p2 = p2; s.p3 = p3;
        s.e1 = p2 - p1;
        s.e2 = p3 - p1;
        s.n1 = s.e2.cross(s.e1).normalize();  // src/shape.rs:370-372
        return finish(a, RTC_TRIANGLE, BoundingBox::empty().insert(p1).insert(p2).insert(p3), std::move(s));
    }
    static Element smooth_triangle(const ShapeArgs& a, Vector p1, Vector p2, Vector p3, Vector n1, Vector n2, Vector n3) {
        Shape s;
        s.p1 = p1; s.p2 = p2; s.p3 = p3;
        s.e1 = p2 - p1;
        s.e2 = p3 - p1;
        s.n1 = n1; s.n2 = n2; s.n3 = n3;
        return finish(a, RTC_SMOOTH_TRIANGLE, BoundingBox::empty().insert(p1).insert(p2).insert(p3), std::move(s));
    }

    // src/shape.rs:74-101 — children are moved in; the group's transform (and material, if any) is baked
    // into every descendant primitive, and every descendant group box moves to the parent space.
    static Element composite(const Matrix& transform, const std::optional<Material>& material, GroupKind kind,
                             std::vector<Element> children) {
        if (kind != GroupKind::Aggregation && children.size() != 2)
            throw std::runtime_error("Element::composite: a CSG group takes exactly two children");
        const Matrix inv = transform.inverse();
        const Matrix inv_tsp = inv.transpose();
        Group g;
        g.kind = kind;
        for (const Element& c : children) g.bbox = g.bbox.union_(c.bbox());
        g.children = std::move(children);
        Element e;
        e.node = std::move(g);
        e.bake(transform, inv, inv_tsp, material ? &*material : nullptr);
        return e;
    }

  private:
    static BoundingBox unit_box() { return BoundingBox::make(Vector::point(-1.0, -1.0, -1.0), Vector::point(1.0, 1.0, 1.0)); }
    static Element prim(const ShapeArgs& a, int kind, const BoundingBox& local, double min = 0.0, double max = 0.0, bool closed = false) {
        Shape s;
        s.min = min;
        s.max = max;
        s.closed = closed;
        return finish(a, kind, local, std::move(s));
    }
    // src/shape.rs:335-347
    static Element finish(const ShapeArgs& a, int kind, const BoundingBox& local, Shape s) {
        s.kind = kind;
        s.transform_inv = a.transform.inverse();
        s.transform_inv_tsp = s.transform_inv.transpose();
        s.material_inv = s.transform_inv;
        s.bbox = local.transform(a.transform);
        s.material = a.material;
        s.casts_shadow = a.casts_shadow;
        Element e;
        e.node = std::move(s);
        return e;
    }
    // src/shape.rs:47-72
    void bake(const Matrix& transform, const Matrix& inv, const Matrix& inv_tsp, const Material* material) {
        if (Group* g = std::get_if<Group>(&node)) {
            for (Element& c : g->children) c.bake(transform, inv, inv_tsp, material);
            g->bbox = g->bbox.transform(transform);
        } else {
            Shape& s = std::get<Shape>(node);
            s.transform_inv = s.transform_inv * inv;
            s.transform_inv_tsp = inv_tsp * s.transform_inv_tsp;
            if (material) {
                s.material = *material;
                s.material_inv = inv;
            } else {
                s.material_inv = s.material_inv * inv;
            }
        }
    }
};

// ---- World / Camera (src/world.rs:12-15,152-183; src/camera.rs) ---------------------------
struct World {
    std::vector<PointLight> lights;
    std::vector<Element> elements;
    static World default_world() {
        World w;
        w.lights.push_back(PointLight{Color::white(), Vector::point(-10.0, 10.0, -10.0)});
        ShapeArgs outer;
        outer.material.pattern = Pattern::plain(Color::make(0.8, 1.0, 0.6));
        outer.material.diffuse = 0.7;
        outer.material.specular = 0.2;
        ShapeArgs inner;
        inner.transform = Matrix::scaling(0.5, 0.5, 0.5);
        w.elements.push_back(Element::sphere(outer));
        w.elements.push_back(Element::sphere(inner));
        return w;
    }
};

class Image;
struct Camera {
    // north_star's spelling of the drop-in call, `Camera::render(&World) -> Image`: an alias of Image::par_render(camera, world),
    // which is what the reference's bins call (src/image.rs:65; the reference has no Camera::render).  Defined below Image.
    Image render(const World& world) const;
    size_t hsize = 0, vsize = 0;
    double field_of_view = 0.0;
    Matrix transform_inv = Matrix::id();
    double pixel_size = 0.0, half_width = 0.0, half_height = 0.0;

    // Camera::new (src/camera.rs:16-37)
    static Camera make(size_t hsize, size_t vsize, double field_of_view, const Matrix& transform) {
        Camera c;
        c.hsize = hsize;
        c.vsize = vsize;
        c.field_of_view = field_of_view;
        const double half_view = std::tan(field_of_view / 2.0);
        const double aspect = (double)hsize / (double)vsize;
        if (aspect >= 1.0) {
            c.half_width = half_view;
            c.half_height = half_view / aspect;
        } else {
            c.half_width = half_view * aspect;
            c.half_height = half_view;
        }
        c.pixel_size = (c.half_width * 2.0) / (double)hsize;
        c.transform_inv = transform.inverse();
        return c;
    }
    // src/camera.rs:57-73
    static Matrix transform(const Vector& from, const Vector& to, const Vector& up) {
        const Vector forward = (to - from).normalize();
        const Vector left = forward.cross(up.normalize());
        const Vector true_up = left.cross(forward);
        Matrix o = Matrix::id();
        o.at(0, 0) = left.x;     o.at(0, 1) = left.y;     o.at(0, 2) = left.z;
        o.at(1, 0) = true_up.x;  o.at(1, 1) = true_up.y;  o.at(1, 2) = true_up.z;
        o.at(2, 0) = -forward.x; o.at(2, 1) = -forward.y; o.at(2, 2) = -forward.z;
        return o * Matrix::translation(-from.x, -from.y, -from.z);
    }
    rtc_camera to_c() const {
        rtc_camera c;
        c.hsize = (int32_t)hsize;
        c.vsize = (int32_t)vsize;
        for (int i = 0; i < 16; ++i) c.transform_inv[i] = transform_inv.m[(size_t)i];
        c.pixel_size = pixel_size;
        c.half_width = half_width;
        c.half_height = half_height;
        return c;
    }
};

// ---- Scene upload: flatten the Element tree into the rtc.h arrays ---------------------------
// Depth-first over World.elements (the order World::intersect visits them, src/world.rs:21-23), so a
// primitive's flat index orders equal-t crossings the way the reference's stable sort does.
// std::allocator whose construct() without arguments default-initialises (leaves a POD's bytes alone) instead of
// value-initialising (zeroing) it: emit_prim writes every one of a primitive's 560 bytes, so zeroing them first is a second
// pass over the 3.5 MB of the teapot's primitives per flatten.
template <class T>
struct DefaultInitAllocator : std::allocator<T> {
    template <class U> struct rebind { using other = DefaultInitAllocator<U>; };
    DefaultInitAllocator() = default;
    template <class U> DefaultInitAllocator(const DefaultInitAllocator<U>&) {}
    template <class U, class... Args>
    void construct(U* p, Args&&... args) {
        if constexpr (sizeof...(Args) == 0) ::new ((void*)p) U;
        else ::new ((void*)p) U(std::forward<Args>(args)...);
    }
};

class FlatScene {
  public:
    std::vector<rtc_prim, DefaultInitAllocator<rtc_prim>> prims;
    std::vector<rtc_node> nodes;
    std::vector<int32_t> children;
    std::vector<rtc_material> materials;
    std::vector<rtc_pattern> patterns;
    std::vector<rtc_light> lights;

    explicit FlatScene(const World& world) {
        for (const PointLight& l : world.lights) {
            rtc_light out;
            out.intensity[0] = l.intensity.r; out.intensity[1] = l.intensity.g; out.intensity[2] = l.intensity.b;
            out.origin[0] = l.origin.x; out.origin[1] = l.origin.y; out.origin[2] = l.origin.z;
            lights.push_back(out);
        }
        rtc_node root{};
        root.kind = RTC_AGGREGATION;
        const double inf = std::numeric_limits<double>::infinity();
        for (int k = 0; k < 3; ++k) { root.bbox_min[k] = -inf; root.bbox_max[k] = inf; }
        nodes.push_back(root);
        // one pass to size the arrays (growing a 600-byte-per-primitive vector by doubling copies it several times over)
        size_t n_prims = 0, n_groups = 0;
        count(world.elements, n_prims, n_groups);
        prims.reserve(n_prims);
        nodes.reserve(n_groups + 1);
        children.reserve(n_prims + n_groups);
        emit_children(0, world.elements);
    }

    // build_flags: RTC_BUILD_FAST for a scene that is rendered once and destroyed (Image::par_render)
    rtc_scene_desc desc(uint32_t build_flags = 0) const {
        rtc_scene_desc d{};
        d.abi_version = RTC_ABI_VERSION;
        d.build_flags = build_flags;
        d.n_prims = (int32_t)prims.size();
        d.n_nodes = (int32_t)nodes.size();
        d.n_children = (int32_t)children.size();
        d.n_materials = (int32_t)materials.size();
        d.n_patterns = (int32_t)patterns.size();
        d.n_lights = (int32_t)lights.size();
        d.prims = prims.data();
        d.nodes = nodes.data();
        d.children = children.data();
        d.materials = materials.data();
        d.patterns = patterns.data();
        d.lights = lights.data();
        return d;
    }

  private:
    std::vector<std::pair<const Pattern::Node*, int32_t>> pattern_cache_;   // identical subtrees are shared
    const Pattern::Node* last_pattern_ = nullptr;  // the previous primitive's material: the triangles of a mesh all carry the same one
    int32_t last_material_ = -1;

    static void count(const std::vector<Element>& elems, size_t& n_prims, size_t& n_groups) {
        for (const Element& e : elems) {
            if (e.is_group()) { n_groups++; count(std::get<Group>(e.node).children, n_prims, n_groups); }
            else n_prims++;
        }
    }

    static void copy16(double (&dst)[16], const Matrix& m) { for (int i = 0; i < 16; ++i) dst[i] = m.m[(size_t)i]; }
    static void copy3(double (&dst)[3], const Vector& v) { dst[0] = v.x; dst[1] = v.y; dst[2] = v.z; }

    int32_t emit_pattern(const std::shared_ptr<const Pattern::Node>& n) {
        for (auto& e : pattern_cache_) if (e.first == n.get()) return e.second;
        rtc_pattern p{};
        p.kind = n->kind;
        p.left = p.right = -1;
        p.noise_kind = n->noise.kind;
        p.octaves = (int32_t)n->noise.octaves;
        p.noise_scale = n->noise.scale;
        p.color[0] = n->color.r; p.color[1] = n->color.g; p.color[2] = n->color.b;
        copy16(p.transform_inv, n->transform_inv);
        if (n->left) p.left = emit_pattern(n->left);
        if (n->right) p.right = emit_pattern(n->right);
        patterns.push_back(p);
        const int32_t idx = (int32_t)patterns.size() - 1;
        pattern_cache_.push_back({n.get(), idx});
        return idx;
    }
    int32_t emit_material(const Material& m) {
        if (last_material_ >= 0 && m.pattern.root().get() == last_pattern_) {  // same pattern object as the previous primitive: compare the scalars only
            const rtc_material& o = materials[(size_t)last_material_];
            if (o.ambient == m.ambient && o.diffuse == m.diffuse && o.specular == m.specular && o.shininess == m.shininess &&
                o.reflective == m.reflective && o.transparency == m.transparency && o.refractive_index == m.refractive_index)
                return last_material_;
        }
        last_pattern_ = m.pattern.root().get();
        last_material_ = emit_material_slow(m);
        return last_material_;
    }
    int32_t emit_material_slow(const Material& m) {
        const int32_t pat = emit_pattern(m.pattern.root());
        for (size_t i = 0; i < materials.size(); ++i) {
            const rtc_material& o = materials[i];
            if (o.pattern == pat && o.ambient == m.ambient && o.diffuse == m.diffuse && o.specular == m.specular &&
                o.shininess == m.shininess && o.reflective == m.reflective && o.transparency == m.transparency &&
                o.refractive_index == m.refractive_index)
                return (int32_t)i;
        }
        rtc_material out{};
        out.ambient = m.ambient; out.diffuse = m.diffuse; out.specular = m.specular; out.shininess = m.shininess;
        out.reflective = m.reflective; out.transparency = m.transparency; out.refractive_index = m.refractive_index;
        out.pattern = pat;
        materials.push_back(out);
        return (int32_t)materials.size() - 1;
    }
    int32_t emit_prim(const Shape& s) {
        const int32_t material = emit_material(s.material);
        static_assert(sizeof(rtc_prim) == 4 * sizeof(int32_t) + (3 * 16 + 2 + 6 * 3) * sizeof(double), "rtc_prim has no padding: every byte is assigned below");
        prims.emplace_back();  // filled in place, NOT zeroed first (DefaultInitAllocator): one pass over its 560 bytes
        rtc_prim& p = prims.back();
        p.kind = s.kind;
        p.material = material;
        p.casts_shadow = s.casts_shadow ? 1 : 0;
        p.closed = s.closed ? 1 : 0;
        copy16(p.transform_inv, s.transform_inv);
        copy16(p.transform_inv_tsp, s.transform_inv_tsp);
        copy16(p.material_inv, s.material_inv);
        p.min = s.min;
        p.max = s.max;
        copy3(p.p1, s.p1); copy3(p.e1, s.e1); copy3(p.e2, s.e2);
        copy3(p.n1, s.n1); copy3(p.n2, s.n2); copy3(p.n3, s.n3);
        return (int32_t)prims.size() - 1;
    }
    void emit_children(int32_t node_index, const std::vector<Element>& elems) {
        // reserve this node's child slots first so they are contiguous, then fill them depth-first
        const int32_t first = (int32_t)children.size();
        children.resize(children.size() + elems.size(), 0);
        nodes[(size_t)node_index].first_child = first;
        nodes[(size_t)node_index].n_children = (int32_t)elems.size();
        for (size_t i = 0; i < elems.size(); ++i) {
            const Element& e = elems[i];
            if (e.is_group()) {
                const Group& g = std::get<Group>(e.node);
                rtc_node n{};
                n.kind = (int32_t)g.kind;
                copy3(n.bbox_min, g.bbox.min);
                copy3(n.bbox_max, g.bbox.max);
                nodes.push_back(n);
                const int32_t idx = (int32_t)nodes.size() - 1;
                children[(size_t)first + i] = ~idx;
                emit_children(idx, g.children);
            } else {
                children[(size_t)first + i] = emit_prim(std::get<Shape>(e.node));
            }
        }
    }
};

// ---- Image (src/image.rs) -----------------------------------------------------------------
// The pixel store of an Image: what `pixels: Vec<Color>` is in the reference (src/image.rs:11-15, a private field).  A
// Color is three f64 (src/color.rs:7-11) — exactly the (r, g, b) triples rtc_render writes — so the frame a render call
// lands in page-locked host memory IS the pixel array: Pixels adopts that buffer instead of copying 199 MB out of it,
// and hands it back to the library's pool (rtc_free_host) when the Image dies.  Otherwise it behaves like the vector it
// replaces (size, [], data, iteration, resize; copies are deep, into ordinary memory).
static_assert(sizeof(Color) == 3 * sizeof(double) && alignof(Color) == alignof(double), "Color must be layout-identical to an f64 rgb triple");
class Pixels {
  public:
    Pixels() = default;
    ~Pixels() { release(); }
    Pixels(const Pixels& o) : own_(o.begin(), o.end()) { p_ = own_.data(); n_ = own_.size(); }
    Pixels& operator=(const Pixels& o) {
        if (this != &o) { release(); own_.assign(o.begin(), o.end()); p_ = own_.data(); n_ = own_.size(); }
        return *this;
    }
    Pixels(Pixels&& o) noexcept { steal(o); }
    Pixels& operator=(Pixels&& o) noexcept {
        if (this != &o) { release(); steal(o); }
        return *this;
    }
    // takes ownership of a buffer obtained from rtc_alloc_host that holds n (r, g, b) triples
    static Pixels adopt_pinned(double* rgb, size_t n) {
        Pixels px;
        px.pinned_ = rgb;
        px.p_ = reinterpret_cast<Color*>(rgb);
        px.n_ = n;
        return px;
    }
    size_t size() const { return n_; }
    bool empty() const { return n_ == 0; }
    Color* data() { return p_; }
    const Color* data() const { return p_; }
    Color& operator[](size_t i) { return p_[i]; }
    const Color& operator[](size_t i) const { return p_[i]; }
    Color* begin() { return p_; }
    Color* end() { return p_ + n_; }
    const Color* begin() const { return p_; }
    const Color* end() const { return p_ + n_; }
    void resize(size_t n) {
        if (pinned_) { std::vector<Color> v(p_, p_ + std::min(n, n_)); release(); own_.swap(v); }
        own_.resize(n);
        p_ = own_.data();
        n_ = n;
    }

  private:
    void release() {
        if (pinned_) rtc_free_host(pinned_);
        pinned_ = nullptr;
        p_ = nullptr;
        n_ = 0;
    }
    void steal(Pixels& o) {
        own_ = std::move(o.own_);
        pinned_ = o.pinned_;
        n_ = o.n_;
        p_ = pinned_ ? reinterpret_cast<Color*>(pinned_) : own_.data();
        o.pinned_ = nullptr;
        o.p_ = nullptr;
        o.n_ = 0;
    }
    Color* p_ = nullptr;
    size_t n_ = 0;
    std::vector<Color> own_;
    double* pinned_ = nullptr;
};

struct RenderOptions {
    std::vector<int> devices;  // CUDA device ordinals to shard the frame over; empty = every visible device
    rtc_opts opts{};           // fuel / supersampling / window (sharding is chosen by par_render itself)
    RenderOptions() { opts.fuel = FUEL; opts.spp_x = 1; opts.spp_y = 1; opts.shard_count = 1; }
};

class Image {
  public:
    size_t hsize = 0, vsize = 0;
    Pixels pixels;
    rtc_stats stats{};
    double flatten_ms = 0.0;  // (diagnostic) time par_render spent turning the World into the flat arrays of the C ABI

    // Image::par_render(camera, world) (src/image.rs:65-81): flatten -> upload to every device -> CUDA wavefront on
    // each device's share of the rows -> rows land in one host image.  Infallible in the reference; here every
    // failure (no device, CUDA error, overflow) throws.
    static Image par_render(const Camera& camera, const World& world, const RenderOptions& ro = RenderOptions{}) {
        const auto t_flat = std::chrono::steady_clock::now();
        FlatScene flat(world);
        const double flatten_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_flat).count();
        const rtc_scene_desc d = flat.desc(RTC_BUILD_FAST);  // one frame, then the scene is released: BVH built on the device
        std::vector<int> devices = ro.devices;
        if (devices.empty()) {
            const int n = rtc_device_count();
            if (n <= 0) throw std::runtime_error(std::string("Image::par_render: no CUDA device: ") + rtc_last_error());
            for (int i = 0; i < n; ++i) devices.push_back(i);
        }
        std::vector<rtc_scene*> scenes(devices.size(), nullptr);
        auto release = [&] { for (rtc_scene* sc : scenes) rtc_scene_destroy(sc); };
        // one host build (flattening, BVH), one upload per device
        if (rtc_scene_create_multi(&d, devices.data(), (int)devices.size(), scenes.data()) != 0)
            throw std::runtime_error(std::string("rtc_scene_create: ") + rtc_last_error());
        const rtc_camera cam = camera.to_c();
        Image img;
        const bool windowed = ro.opts.win_w > 0 && ro.opts.win_h > 0;
        img.hsize = windowed ? (size_t)ro.opts.win_w : camera.hsize;
        img.vsize = windowed ? (size_t)ro.opts.win_h : camera.vsize;
        const size_t n = img.hsize * img.vsize;
        // page-locked frame (from the library's pool): the device -> host copies run at full PCIe rate, overlap the kernels, and
        // the buffer then becomes the Image's pixel array as it is
        double* buf = (double*)rtc_alloc_host((uint64_t)n * 3 * sizeof(double));
        if (!buf) { release(); throw std::runtime_error(std::string("rtc_alloc_host: ") + rtc_last_error()); }
        rtc_opts o = ro.opts;
        o.shard_index = 0;
        o.shard_count = 1;
        const int rc = rtc_render_multi(scenes.data(), (int)scenes.size(), &cam, &o, buf, &img.stats);
        const std::string err = rc ? rtc_last_error() : "";
        release();
        if (rc != 0) { rtc_free_host(buf); throw std::runtime_error("rtc_render: " + err); }
        img.pixels = Pixels::adopt_pinned(buf, n);
        img.flatten_ms = flatten_ms;
        return img;
    }
    Color read(size_t x, size_t y) const { return pixels[y * hsize + x]; }
    void write(size_t x, size_t y, Color c) { pixels[y * hsize + x] = c; }

    // src/image.rs:93-112 — "P3", "w h", "255", then at most five pixels per text line, restarting at each row.
    std::string ppm() const {
        std::ostringstream out;
        out << "P3\n" << hsize << " " << vsize << "\n255";
        size_t on_line = 0;
        for (size_t i = 0; i < pixels.size(); ++i) {
            const auto q = pixels[i].clamp();
            const bool new_line = (i % hsize == 0) || (on_line % 5 == 0);
            out << (new_line ? "\n" : " ") << (int)q[0] << " " << (int)q[1] << " " << (int)q[2];
            on_line = new_line ? 1 : on_line + 1;
        }
        out << "\n";
        return out.str();
    }

    // The reference's bins write `<name>.ppm` and then shell out to `zsh -c "pnmtopng <name>.ppm > <name>.png"`
    // (e.g. src/bin/chapter15_teapot.rs:96-120); neither zsh nor netpbm exists on the GPU box, so the PNG is encoded
    // here: 8-bit RGB (colour type 2) of the same Color::clamp triples the PPM holds, rows filtered with "None",
    // zlib stream of stored (uncompressed) deflate blocks — decodes to exactly the pixels of `pnmtopng`'s output.
    std::string png() const {
        std::string raw;
        raw.reserve(vsize * (1 + 3 * hsize));
        for (size_t y = 0; y < vsize; ++y) {
            raw.push_back('\0');
            for (size_t x = 0; x < hsize; ++x) {
                const auto q = pixels[y * hsize + x].clamp();
                raw.push_back((char)q[0]); raw.push_back((char)q[1]); raw.push_back((char)q[2]);
            }
        }
        std::string z;
        z.push_back((char)0x78); z.push_back((char)0x01);
        uint32_t a = 1, b = 0;  // Adler-32 of the raw scanlines
        for (size_t at = 0; at < raw.size() || at == 0;) {
            const size_t n = std::min<size_t>(65535, raw.size() - at);
            z.push_back((char)(at + n >= raw.size() ? 1 : 0));
            z.push_back((char)(n & 0xff)); z.push_back((char)(n >> 8));
            z.push_back((char)(~n & 0xff)); z.push_back((char)((~n >> 8) & 0xff));
            for (size_t i = 0; i < n; ++i) {
                a += (unsigned char)raw[at + i];
                if (a >= 65521) a -= 65521;
                b += a;
                if (b >= 65521) b -= 65521;
            }
            z.append(raw, at, n);
            at += n;
            if (n == 0) break;
        }
        put_be32(z, (b << 16) | a);
        std::string out("\x89PNG\r\n\x1a\n", 8);
        std::string ihdr;
        put_be32(ihdr, (uint32_t)hsize);
        put_be32(ihdr, (uint32_t)vsize);
        ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
        png_chunk(out, "IHDR", ihdr);
        png_chunk(out, "IDAT", z);
        png_chunk(out, "IEND", std::string());
        return out;
    }

    // What a bin does after par_render (src/bin/*.rs): `<path>.ppm` as text, `<path>.png` next to it.
    void save(const std::string& path) const {
        std::ofstream(path + ".ppm", std::ios::binary) << ppm();
        std::ofstream(path + ".png", std::ios::binary) << png();
    }

  private:
    static void put_be32(std::string& s, uint32_t v) {
        s.push_back((char)(v >> 24)); s.push_back((char)(v >> 16)); s.push_back((char)(v >> 8)); s.push_back((char)v);
    }
    static void png_chunk(std::string& out, const char* tag, const std::string& body) {
        put_be32(out, (uint32_t)body.size());
        std::string crc_in(tag, 4);
        crc_in += body;
        static const std::array<uint32_t, 256> table = [] {
            std::array<uint32_t, 256> t{};
            for (uint32_t i = 0; i < 256; ++i) {
                uint32_t c = i;
                for (int k = 0; k < 8; ++k) c = (c >> 1) ^ (0xedb88320u & (0u - (c & 1u)));
                t[i] = c;
            }
            return t;
        }();
        uint32_t c = 0xffffffffu;
        for (unsigned char ch : crc_in) c = table[(c ^ ch) & 0xffu] ^ (c >> 8);
        out += crc_in;
        put_be32(out, c ^ 0xffffffffu);
    }
};

inline Image Camera::render(const World& world) const { return Image::par_render(*this, world); }

// ---- ObjParser (src/obj.rs) ---------------------------------------------------------------
// Line grammar of the reference (nom combinators, src/obj.rs:53-149), first match wins, trailing text ignored:
//   "v"  ws x ws y ws z        vertex          "vn" ws x ws y ws z        normal
//   "f"  ws idx (ws idx)*      polygon, fan-triangulated (:115-123); idx = v | v/<anything>/n
//   "g"  ws name               group (leading alphanumerics of the name)
// anything else is reported back as an ignored (line number, text) pair.
struct ObjData {
    std::vector<Vector> vertices, normals;
    struct Tri { size_t p[3]; size_t n[3]; bool has_n; size_t group; };
    std::vector<Tri> triangles;
    std::vector<std::string> groups;  // groups[0] = "Default"
    std::vector<std::pair<uint32_t, std::string>> ignored;
};

class ObjParser {
  public:
    explicit ObjParser(std::string path) : path_(std::move(path)) {}

    static ObjData parse_text(std::istream& in) {
        ObjData d;
        d.groups.push_back("Default");
        size_t current = 0;
        std::string line;
        for (uint32_t lineno = 1; std::getline(in, line); ++lineno) {
            if (!line.empty() && line.back() == '\r') line.pop_back();
            Cursor c{line.c_str()};
            double x, y, z;
            bool ok = false;
            if (c.lit("v") && c.ws() && c.real(x) && c.ws() && c.real(y) && c.ws() && c.real(z)) {
                d.vertices.push_back(Vector::point(x, y, z));
                ok = true;
            } else if (c.reset(line.c_str()), c.lit("vn") && c.ws() && c.real(x) && c.ws() && c.real(y) && c.ws() && c.real(z)) {
                d.normals.push_back(Vector::vector(x, y, z));
                ok = true;
            } else if (c.reset(line.c_str()), c.lit("f") && c.ws()) {
                std::vector<Corner> poly;
                Corner k;
                if (c.corner(k)) {
                    poly.push_back(k);
                    for (;;) {
                        Cursor save = c;
                        if (c.ws() && c.corner(k)) poly.push_back(k);
                        else { c = save; break; }
                    }
                }
                if (poly.empty()) throw std::runtime_error("ObjParser: face without indices at line " + std::to_string(lineno));
                for (size_t i = 1; i + 1 < poly.size(); ++i) {
                    ObjData::Tri t{};
                    const Corner* cs[3] = {&poly[0], &poly[i], &poly[i + 1]};
                    t.has_n = true;
                    for (int j = 0; j < 3; ++j) {
                        t.p[j] = cs[j]->v;
                        t.n[j] = cs[j]->n;
                        t.has_n = t.has_n && cs[j]->has_n;
                        if (t.p[j] < 1 || t.p[j] > d.vertices.size()) throw std::runtime_error("ObjParser: vertex index out of range at line " + std::to_string(lineno));
                    }
                    for (int j = 0; j < 3 && t.has_n; ++j)
                        if (t.n[j] < 1 || t.n[j] > d.normals.size()) throw std::runtime_error("ObjParser: normal index out of range at line " + std::to_string(lineno));
                    t.group = current;
                    d.triangles.push_back(t);
                }
                ok = true;
            } else if (c.reset(line.c_str()), c.lit("g") && c.ws()) {
                std::string name = c.alnum();
                if (!name.empty()) {
                    size_t g = 0;
                    while (g < d.groups.size() && d.groups[g] != name) ++g;
                    if (g == d.groups.size()) d.groups.push_back(name);
                    current = g;
                    ok = true;
                }
            }
            if (!ok) d.ignored.push_back({lineno, line});
        }
        return d;
    }

    // One Aggregation group per non-empty `g` name, each built with the mesh transform + material
    // (src/obj.rs:239-248); a single group is returned as is, several are wrapped (src/obj.rs:250-257).
    // The reference drains a HashMap (unspecified order); first-seen order is used here.
    static Element build(const ObjData& d, const Matrix& transform, const Material& material) {
        std::vector<std::vector<Element>> buckets(d.groups.size());
        for (const ObjData::Tri& t : d.triangles) {
            const Vector &a = d.vertices[t.p[0] - 1], &b = d.vertices[t.p[1] - 1], &c = d.vertices[t.p[2] - 1];
            buckets[t.group].push_back(t.has_n ? Element::smooth_triangle(ShapeArgs{}, a, b, c, d.normals[t.n[0] - 1],
                                                                          d.normals[t.n[1] - 1], d.normals[t.n[2] - 1])
                                               : Element::triangle(ShapeArgs{}, a, b, c));
        }
        std::vector<Element> groups;
        for (auto& b : buckets)
            if (!b.empty()) groups.push_back(Element::composite(transform, material, GroupKind::Aggregation, std::move(b)));
        if (groups.size() == 1) return std::move(groups.front());
        return Element::composite(Matrix::id(), std::nullopt, GroupKind::Aggregation, std::move(groups));
    }

    // src/obj.rs:186-258
    std::pair<std::vector<std::pair<uint32_t, std::string>>, Element> parse_obj(const Matrix& transform, const Material& material) const {
        std::ifstream in(path_);
        if (!in) throw std::runtime_error("ObjParser: cannot open " + path_);
        ObjData d = parse_text(in);
        return {d.ignored, build(d, transform, material)};
    }

  private:
    struct Corner { size_t v = 0, n = 0; bool has_n = false; };
    struct Cursor {
        const char* p;
        void reset(const char* s) { p = s; }
        bool lit(const char* s) {
            const char* q = p;
            for (; *s; ++s, ++q) if (*q != *s) return false;
            p = q;
            return true;
        }
        bool ws() {
            const char* q = p;
            while (*q == ' ' || *q == '\t') ++q;
            const bool any = q != p;
            p = q;
            return any;
        }
        static bool digit(char ch) { return ch >= '0' && ch <= '9'; }
        bool uint(size_t& out) {
            if (!digit(*p)) return false;
            size_t v = 0;
            while (digit(*p)) v = v * 10 + (size_t)(*p++ - '0');
            out = v;
            return true;
        }
        // decimal float as nom's `double` recognises it; value = correctly rounded (strtod)
        bool real(double& out) {
            const char* q = p;
            if (*q == '+' || *q == '-') ++q;
            const char* int0 = q;
            while (digit(*q)) ++q;
            if (q != int0) {
                if (*q == '.') { ++q; while (digit(*q)) ++q; }
            } else {
                if (*q != '.' || !digit(q[1])) return false;
                ++q;
                while (digit(*q)) ++q;
            }
            if (*q == 'e' || *q == 'E') {
                const char* e = q + 1;
                if (*e == '+' || *e == '-') ++e;
                if (digit(*e)) { while (digit(*e)) ++e; q = e; }
            }
            out = std::strtod(std::string(p, q).c_str(), nullptr);
            p = q;
            return true;
        }
        bool corner(Corner& out) {
            Cursor t = *this;
            size_t v, n;
            if (t.uint(v) && *t.p == '/') {
                const char* q = t.p + 1;
                while (*q && *q != '/') ++q;  // many_till(anychar, '/')
                if (*q == '/') {
                    Cursor u{q + 1};
                    if (u.uint(n)) { out = Corner{v, n, true}; p = u.p; return true; }
                }
            }
            t = *this;
            if (t.uint(v)) { out = Corner{v, 0, false}; p = t.p; return true; }
            return false;
        }
        std::string alnum() {
            const char* q = p;
            while (digit(*q) || (*q >= 'a' && *q <= 'z') || (*q >= 'A' && *q <= 'Z')) ++q;
            std::string s(p, q);
            p = q;
            return s;
        }
    };
    std::string path_;
};

}  // namespace rtc
