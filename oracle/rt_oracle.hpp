// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path.
//
// CPU restatement (C++17, f64, no fast-math, compiled with -ffp-contract=off) of the
// reference ray tracer's per-pixel render path, following the reference file by file and
// operation by operation so that IEEE-754 results agree with the Rust build.  Only
// tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// use anything under oracle/.
//
// Parity pinning: oracle/test_oracle.cpp checks this restatement against the reference's
// own known-answer unit tests (transcribed as data), and tests/test_oracle_cpu.py checks
// windows of native-resolution renders against the reference's committed PNG outputs (tests/golden/).
// Unpinned by the reference's tests: noise.rs (Simplex is pinned only through the
// hexagon PNG; Fractal by nothing), MixtureKind::{Blend,RingGradient}, JitterKind::Color.
//
// Every function cites the reference file:line (relative to /root/reference) it follows.
#pragma once

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <limits>
#include <map>
#include <memory>
#include <optional>
#include <stdexcept>
#include <string>
#include <thread>
#include <utility>
#include <vector>

namespace orc {

// ---------------------------------------------------------------------------------------
// src/config.rs:1-2
constexpr double EPSILON = 0.00001;
constexpr int FUEL = 5;

// src/approx.rs:19-23
inline bool approx(double a, double b) { return std::fabs(a - b) < EPSILON; }

// Rust `x.powi(2)` lowers to x*x; `powi(5)` to x*((x*x)*(x*x)) (binary exponentiation).
inline double powi2(double x) { return x * x; }
inline double powi5(double x) {
    double x2 = x * x;
    double x4 = x2 * x2;
    return x * x4;
}
// Rust `f as i32`: saturating, NaN -> 0.
inline int32_t sat_i32(double x) {
    if (x != x) return 0;
    if (x >= 2147483647.0) return INT32_MAX;
    if (x <= -2147483648.0) return INT32_MIN;
    return (int32_t)x;
}
// i32 arithmetic in a release build wraps.
inline int32_t wrap_add(int32_t a, int32_t b) { return (int32_t)((uint32_t)a + (uint32_t)b); }
inline int32_t wrap_sub(int32_t a, int32_t b) { return (int32_t)((uint32_t)a - (uint32_t)b); }

// ---------------------------------------------------------------------------------------
// src/linalg/vector.rs
struct Vector {
    double x = 0, y = 0, z = 0, w = 0;
    static Vector point(double x, double y, double z) { return {x, y, z, 1.0}; }   // :16-18
    static Vector vector(double x, double y, double z) { return {x, y, z, 0.0}; }  // :20-22
    double magnitude() const { return std::sqrt(powi2(x) + powi2(y) + powi2(z)); }  // :24-26
    Vector normalize() const {                                                      // :28-37
        double m = magnitude();
        return {x / m, y / m, z / m, 0.0};
    }
    double dot(const Vector& o) const { return x * o.x + y * o.y + z * o.z; }       // :39-41
    Vector cross(const Vector& o) const {                                           // :43-49
        return vector(y * o.z - z * o.y, z * o.x - x * o.z, x * o.y - y * o.x);
    }
    Vector reflect(const Vector& normal) const;                                     // :75-77
};
inline Vector operator+(const Vector& a, const Vector& b) { return {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
inline Vector operator-(const Vector& a, const Vector& b) { return {a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w}; }
inline Vector operator-(const Vector& a) { return {-a.x, -a.y, -a.z, -a.w}; }
inline Vector operator*(const Vector& a, double s) { return {a.x * s, a.y * s, a.z * s, a.w * s}; }
inline Vector operator/(const Vector& a, double s) { return {a.x / s, a.y / s, a.z / s, a.w / s}; }
inline Vector Vector::reflect(const Vector& normal) const { return *this - normal * (2.0 * this->dot(normal)); }
inline bool approx(const Vector& a, const Vector& b) {
    return approx(a.x, b.x) && approx(a.y, b.y) && approx(a.z, b.z) && approx(a.w, b.w);
}

// ---------------------------------------------------------------------------------------
// src/linalg/matrix.rs
struct Matrix {
    double d[4][4];
    static Matrix make(const double (&m)[4][4]) {
        Matrix r;
        for (int i = 0; i < 4; i++)
            for (int j = 0; j < 4; j++) r.d[i][j] = m[i][j];
        return r;
    }
    static Matrix id() {  // :19-28
        double m[4][4] = {{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}};
        return make(m);
    }
    static Matrix translation(double x, double y, double z) {  // :30-40
        double m[4][4] = {{1, 0, 0, x}, {0, 1, 0, y}, {0, 0, 1, z}, {0, 0, 0, 1}};
        return make(m);
    }
    static Matrix scaling(double x, double y, double z) {  // :42-52
        double m[4][4] = {{x, 0, 0, 0}, {0, y, 0, 0}, {0, 0, z, 0}, {0, 0, 0, 1}};
        return make(m);
    }
    static Matrix rotation_x(double r) {  // :54-64
        double m[4][4] = {{1, 0, 0, 0}, {0, std::cos(r), -std::sin(r), 0}, {0, std::sin(r), std::cos(r), 0}, {0, 0, 0, 1}};
        return make(m);
    }
    static Matrix rotation_y(double r) {  // :66-76
        double m[4][4] = {{std::cos(r), 0, std::sin(r), 0}, {0, 1, 0, 0}, {-std::sin(r), 0, std::cos(r), 0}, {0, 0, 0, 1}};
        return make(m);
    }
    static Matrix rotation_z(double r) {  // :78-88
        double m[4][4] = {{std::cos(r), -std::sin(r), 0, 0}, {std::sin(r), std::cos(r), 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}};
        return make(m);
    }
    static Matrix shearing(double xy, double xz, double yx, double yz, double zx, double zy) {  // :90-100
        double m[4][4] = {{1, xy, xz, 0}, {yx, 1, yz, 0}, {zx, zy, 1, 0}, {0, 0, 0, 1}};
        return make(m);
    }
    Matrix transpose() const {  // :126-136
        Matrix r;
        for (int row = 0; row < 4; row++)
            for (int col = 0; col < 4; col++) r.d[col][row] = d[row][col];
        return r;
    }
    double determinant() const;  // :138-156
    Matrix inverse() const;      // :162-207
    // fluent forms pre-multiply (:102-124)
    Matrix translate(double x, double y, double z) const;
    Matrix scale(double x, double y, double z) const;
    Matrix rotate_x(double r) const;
    Matrix rotate_y(double r) const;
    Matrix rotate_z(double r) const;
    Matrix shear(double xy, double xz, double yx, double yz, double zx, double zy) const;
};

// src/linalg/matrix.rs:239-259 (accumulator starts at 0.0, i ascending)
inline Matrix operator*(const Matrix& a, const Matrix& b) {
    Matrix r;
    for (int row = 0; row < 4; row++)
        for (int col = 0; col < 4; col++) {
            double value = 0.0;
            for (int i = 0; i < 4; i++) value += a.d[row][i] * b.d[i][col];
            r.d[row][col] = value;
        }
    return r;
}
// src/linalg/matrix.rs:261-284 (full 4x4 incl. the w row, left-to-right sums)
inline Vector operator*(const Matrix& m, const Vector& v) {
    return {
        m.d[0][0] * v.x + m.d[0][1] * v.y + m.d[0][2] * v.z + m.d[0][3] * v.w,
        m.d[1][0] * v.x + m.d[1][1] * v.y + m.d[1][2] * v.z + m.d[1][3] * v.w,
        m.d[2][0] * v.x + m.d[2][1] * v.y + m.d[2][2] * v.z + m.d[2][3] * v.w,
        m.d[3][0] * v.x + m.d[3][1] * v.y + m.d[3][2] * v.z + m.d[3][3] * v.w,
    };
}
inline Matrix Matrix::translate(double x, double y, double z) const { return translation(x, y, z) * *this; }
inline Matrix Matrix::scale(double x, double y, double z) const { return scaling(x, y, z) * *this; }
inline Matrix Matrix::rotate_x(double r) const { return rotation_x(r) * *this; }
inline Matrix Matrix::rotate_y(double r) const { return rotation_y(r) * *this; }
inline Matrix Matrix::rotate_z(double r) const { return rotation_z(r) * *this; }
inline Matrix Matrix::shear(double xy, double xz, double yx, double yz, double zx, double zy) const {
    return shearing(xy, xz, yx, yz, zx, zy) * *this;
}
inline bool approx(const Matrix& a, const Matrix& b) {
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++)
            if (!approx(a.d[i][j], b.d[i][j])) return false;
    return true;
}

inline double Matrix::determinant() const {
    const auto& m = d;
    double s0 = m[0][0] * m[1][1] - m[1][0] * m[0][1];
    double s1 = m[0][0] * m[1][2] - m[1][0] * m[0][2];
    double s2 = m[0][0] * m[1][3] - m[1][0] * m[0][3];
    double s3 = m[0][1] * m[1][2] - m[1][1] * m[0][2];
    double s4 = m[0][1] * m[1][3] - m[1][1] * m[0][3];
    double s5 = m[0][2] * m[1][3] - m[1][2] * m[0][3];
    double c5 = m[2][2] * m[3][3] - m[3][2] * m[2][3];
    double c4 = m[2][1] * m[3][3] - m[3][1] * m[2][3];
    double c3 = m[2][1] * m[3][2] - m[3][1] * m[2][2];
    double c2 = m[2][0] * m[3][3] - m[3][0] * m[2][3];
    double c1 = m[2][0] * m[3][2] - m[3][0] * m[2][2];
    double c0 = m[2][0] * m[3][1] - m[3][0] * m[2][1];
    return s0 * c5 - s1 * c4 + s2 * c3 + s3 * c2 - s4 * c1 + s5 * c0;
}

// 2x2-subfactor cofactor inverse, term order exactly as src/linalg/matrix.rs:162-207.
inline Matrix Matrix::inverse() const {
    const auto& m = d;
    double s0 = m[0][0] * m[1][1] - m[1][0] * m[0][1];
    double s1 = m[0][0] * m[1][2] - m[1][0] * m[0][2];
    double s2 = m[0][0] * m[1][3] - m[1][0] * m[0][3];
    double s3 = m[0][1] * m[1][2] - m[1][1] * m[0][2];
    double s4 = m[0][1] * m[1][3] - m[1][1] * m[0][3];
    double s5 = m[0][2] * m[1][3] - m[1][2] * m[0][3];
    double c5 = m[2][2] * m[3][3] - m[3][2] * m[2][3];
    double c4 = m[2][1] * m[3][3] - m[3][1] * m[2][3];
    double c3 = m[2][1] * m[3][2] - m[3][1] * m[2][2];
    double c2 = m[2][0] * m[3][3] - m[3][0] * m[2][3];
    double c1 = m[2][0] * m[3][2] - m[3][0] * m[2][2];
    double c0 = m[2][0] * m[3][1] - m[3][0] * m[2][1];
    double det = s0 * c5 - s1 * c4 + s2 * c3 + s3 * c2 - s4 * c1 + s5 * c0;
    if (!(det != 0.0)) throw std::runtime_error("oracle: matrix not invertible (reference asserts det != 0)");

    Matrix r;
    r.d[0][0] = (m[1][1] * c5 - m[1][2] * c4 + m[1][3] * c3) / det;
    r.d[0][2] = (m[3][1] * s5 - m[3][2] * s4 + m[3][3] * s3) / det;
    r.d[1][1] = (m[0][0] * c5 - m[0][2] * c2 + m[0][3] * c1) / det;
    r.d[1][3] = (m[2][0] * s5 - m[2][2] * s2 + m[2][3] * s1) / det;
    r.d[2][0] = (m[1][0] * c4 - m[1][1] * c2 + m[1][3] * c0) / det;
    r.d[2][2] = (m[3][0] * s4 - m[3][1] * s2 + m[3][3] * s0) / det;
    r.d[3][1] = (m[0][0] * c3 - m[0][1] * c1 + m[0][2] * c0) / det;
    r.d[3][3] = (m[2][0] * s3 - m[2][1] * s1 + m[2][2] * s0) / det;

    r.d[0][1] = (-m[0][1] * c5 + m[0][2] * c4 - m[0][3] * c3) / det;
    r.d[0][3] = (-m[2][1] * s5 + m[2][2] * s4 - m[2][3] * s3) / det;
    r.d[1][0] = (-m[1][0] * c5 + m[1][2] * c2 - m[1][3] * c1) / det;
    r.d[1][2] = (-m[3][0] * s5 + m[3][2] * s2 - m[3][3] * s1) / det;
    r.d[2][1] = (-m[0][0] * c4 + m[0][1] * c2 - m[0][3] * c0) / det;
    r.d[2][3] = (-m[2][0] * s4 + m[2][1] * s2 - m[2][3] * s0) / det;
    r.d[3][0] = (-m[1][0] * c3 + m[1][1] * c1 - m[1][2] * c0) / det;
    r.d[3][2] = (-m[3][0] * s3 + m[3][1] * s1 - m[3][2] * s0) / det;
    return r;
}

// ---------------------------------------------------------------------------------------
// src/color.rs
struct Color {
    double r = 0, g = 0, b = 0;
    static Color make(double r, double g, double b) { return {r, g, b}; }                       // :14-16
    static Color rgb(uint8_t r, uint8_t g, uint8_t b) { return {r / 255.0, g / 255.0, b / 255.0}; }  // :18-24
    static Color white() { return {1.0, 1.0, 1.0}; }
    static Color black() { return {0.0, 0.0, 0.0}; }
    Color avg(const Color& o) const;  // :48-50
};
inline Color operator+(const Color& a, const Color& b) { return {a.r + b.r, a.g + b.g, a.b + b.b}; }
inline Color operator-(const Color& a, const Color& b) { return {a.r - b.r, a.g - b.g, a.b - b.b}; }
inline Color operator*(const Color& a, double s) { return {a.r * s, a.g * s, a.b * s}; }
inline Color operator*(const Color& a, const Color& b) { return {a.r * b.r, a.g * b.g, a.b * b.b}; }
inline Color& operator+=(Color& a, const Color& b) { a = a + b; return a; }
inline Color Color::avg(const Color& o) const { return (*this + o) * 0.5; }
inline bool approx(const Color& a, const Color& b) { return approx(a.r, b.r) && approx(a.g, b.g) && approx(a.b, b.b); }
// src/color.rs:42-46 — min(1).max(0)*255, round half away from zero, saturating u8 cast.
inline uint8_t clamp_channel(double x) {
    double v = std::round(std::fmax(std::fmin(x, 1.0), 0.0) * 255.0);
    if (v != v) return 0;
    if (v >= 255.0) return 255;
    if (v <= 0.0) return 0;
    return (uint8_t)v;
}

// ---------------------------------------------------------------------------------------
// src/ray.rs
struct Ray {
    Vector origin, direction;
    Vector position(double t) const { return origin + direction * t; }          // :10-12
    Ray transform(const Matrix& m) const { return {m * origin, m * direction}; }  // :14-19
};

// src/light.rs:4-8
struct PointLight {
    Color intensity;
    Vector origin;
};

// ---------------------------------------------------------------------------------------
// src/noise.rs
struct Noise {
    enum Kind { KSimplex, KFractal } kind = KSimplex;  // :4-7
    double scale = 0.0;
    size_t octaves = 0;
    static Noise Simplex(double scale) { return {KSimplex, scale, 0}; }
    static Noise Fractal(double scale, size_t octaves) { return {KFractal, scale, octaves}; }
    void jitter_3d(double x, double y, double z, double& ox, double& oy, double& oz) const;  // :31-52
};

namespace noise_detail {
// The classic 256-entry permutation, stored twice (src/noise.rs:56-84).
inline const uint8_t* perm() {
    static uint8_t P[512];
    static bool init = false;
    if (!init) {
        static const uint8_t base[256] = {
            151, 160, 137, 91,  90,  15,  131, 13,  201, 95,  96,  53,  194, 233, 7,   225, 140, 36,  103, 30,  69,  142,
            8,   99,  37,  240, 21,  10,  23,  190, 6,   148, 247, 120, 234, 75,  0,   26,  197, 62,  94,  252, 219, 203,
            117, 35,  11,  32,  57,  177, 33,  88,  237, 149, 56,  87,  174, 20,  125, 136, 171, 168, 68,  175, 74,  165,
            71,  134, 139, 48,  27,  166, 77,  146, 158, 231, 83,  111, 229, 122, 60,  211, 133, 230, 220, 105, 92,  41,
            55,  46,  245, 40,  244, 102, 143, 54,  65,  25,  63,  161, 1,   216, 80,  73,  209, 76,  132, 187, 208, 89,
            18,  169, 200, 196, 135, 130, 116, 188, 159, 86,  164, 100, 109, 198, 173, 186, 3,   64,  52,  217, 226, 250,
            124, 123, 5,   202, 38,  147, 118, 126, 255, 82,  85,  212, 207, 206, 59,  227, 47,  16,  58,  17,  182, 189,
            28,  42,  223, 183, 170, 213, 119, 248, 152, 2,   44,  154, 163, 70,  221, 153, 101, 155, 167, 43,  172, 9,
            129, 22,  39,  253, 19,  98,  108, 110, 79,  113, 224, 232, 178, 185, 112, 104, 218, 246, 97,  228, 251, 34,
            242, 193, 238, 210, 144, 12,  191, 179, 162, 241, 81,  51,  145, 235, 249, 14,  239, 107, 49,  192, 214, 31,
            181, 199, 106, 157, 184, 84,  204, 176, 115, 121, 50,  45,  127, 4,   150, 254, 138, 236, 205, 93,  222, 114,
            67,  29,  24,  72,  243, 141, 128, 195, 78,  66,  215, 61,  156, 180};
        for (int i = 0; i < 512; i++) P[i] = base[i & 255];
        init = true;
    }
    return P;
}
constexpr double F3 = 1.0 / 3.0;  // :86
constexpr double G3 = 1.0 / 6.0;  // :87
inline size_t hash(size_t i) { return perm()[i]; }  // :89-91

// :94-114 — note cases 0xC..0xF are (y+x, -y+z, y-x, -y-z)
inline double grad(size_t h, double x, double y, double z) {
    switch (h & 0xF) {
        case 0x0: return x + y;
        case 0x1: return -x + y;
        case 0x2: return x - y;
        case 0x3: return -x - y;
        case 0x4: return x + z;
        case 0x5: return -x + z;
        case 0x6: return x - z;
        case 0x7: return -x - z;
        case 0x8: return y + z;
        case 0x9: return -y + z;
        case 0xA: return y - z;
        case 0xB: return -y - z;
        case 0xC: return y + x;
        case 0xD: return -y + z;
        case 0xE: return y - x;
        default: return -y - z;
    }
}
// :116-122 — (i32)x for x > 0, else (i32)x - 1 (off by one for non-positive integers: kept)
inline int32_t fast_floor(double x) { return x > 0.0 ? sat_i32(x) : wrap_sub(sat_i32(x), 1); }
// :124-131
inline size_t modulus(int32_t x, int32_t m) {
    int32_t a = x % m;
    return a < 0 ? (size_t)(a + m) : (size_t)a;
}
// :134-219
inline double simplex(double x, double y, double z) {
    double s = (x + y + z) * F3;
    int32_t i = fast_floor(x + s);
    int32_t j = fast_floor(y + s);
    int32_t k = fast_floor(z + s);
    double t = (double)wrap_add(wrap_add(i, j), k) * G3;
    double x0 = x - ((double)i - t);
    double y0 = y - ((double)j - t);
    double z0 = z - ((double)k - t);

    int i1, j1, k1, i2, j2, k2;
    if (x0 >= y0) {
        if (y0 >= z0) { i1 = 1; j1 = 0; k1 = 0; i2 = 1; j2 = 1; k2 = 0; }
        else if (x0 >= z0) { i1 = 1; j1 = 0; k1 = 0; i2 = 1; j2 = 0; k2 = 1; }
        else { i1 = 0; j1 = 0; k1 = 1; i2 = 1; j2 = 0; k2 = 1; }
    } else {
        if (y0 < z0) { i1 = 0; j1 = 0; k1 = 1; i2 = 0; j2 = 1; k2 = 1; }
        else if (x0 < z0) { i1 = 0; j1 = 1; k1 = 0; i2 = 0; j2 = 1; k2 = 1; }
        else { i1 = 0; j1 = 1; k1 = 0; i2 = 1; j2 = 1; k2 = 0; }
    }

    double x1 = x0 - (double)i1 + G3;
    double y1 = y0 - (double)j1 + G3;
    double z1 = z0 - (double)k1 + G3;
    double x2 = x0 - (double)i2 + 2.0 * G3;
    double y2 = y0 - (double)j2 + 2.0 * G3;
    double z2 = z0 - (double)k2 + 2.0 * G3;
    double x3 = x0 - 1.0 + 3.0 * G3;
    double y3 = y0 - 1.0 + 3.0 * G3;
    double z3 = z0 - 1.0 + 3.0 * G3;

    size_t ii = modulus(i, 256), jj = modulus(j, 256), kk = modulus(k, 256);
    size_t gi0 = hash(ii + hash(jj + hash(kk)));
    size_t gi1 = hash(ii + i1 + hash(jj + j1 + hash(kk + k1)));
    size_t gi2 = hash(ii + i2 + hash(jj + j2 + hash(kk + k2)));
    size_t gi3 = hash(ii + 1 + hash(jj + 1 + hash(kk + 1)));

    double n0, n1, n2, n3;
    double t0 = 0.6 - x0 * x0 - y0 * y0 - z0 * z0;
    if (t0 < 0.0) n0 = 0.0; else { t0 *= t0; n0 = t0 * t0 * grad(gi0, x0, y0, z0); }
    double t1 = 0.6 - x1 * x1 - y1 * y1 - z1 * z1;
    if (t1 < 0.0) n1 = 0.0; else { t1 *= t1; n1 = t1 * t1 * grad(gi1, x1, y1, z1); }
    double t2 = 0.6 - x2 * x2 - y2 * y2 - z2 * z2;
    if (t2 < 0.0) n2 = 0.0; else { t2 *= t2; n2 = t2 * t2 * grad(gi2, x2, y2, z2); }
    double t3 = 0.6 - x3 * x3 - y3 * y3 - z3 * z3;
    if (t3 < 0.0) n3 = 0.0; else { t3 *= t3; n3 = t3 * t3 * grad(gi3, x3, y3, z3); }

    return 32.0 * (n0 + n1 + n2 + n3);
}
// :221-237 (0 octaves -> 0/0 = NaN, kept)
inline double fractal(double x, double y, double z, size_t octaves) {
    double output = 0.0, denom = 0.0, frequency = 1.0, amplitude = 1.0;
    const double lacunarity = 2.0, persistence = 0.5;
    for (size_t o = 0; o < octaves; o++) {
        output += amplitude * simplex(x * frequency, y * frequency, z * frequency);
        denom += amplitude;
        frequency *= lacunarity;
        amplitude *= persistence;
    }
    return output / denom;
}
}  // namespace noise_detail

inline void Noise::jitter_3d(double x, double y, double z, double& ox, double& oy, double& oz) const {
    using noise_detail::simplex;
    using noise_detail::fractal;
    double nx, ny, nz;
    if (kind == KSimplex) {
        nx = simplex(x, y, z) * scale;
        ny = simplex(x, y, z + 1.0) * scale;
        nz = simplex(x, y, z + 2.0) * scale;
    } else {
        nx = fractal(x, y, z, octaves) * scale;
        ny = fractal(x, y, z + 1.0, octaves) * scale;
        nz = fractal(x, y, z + 2.0, octaves) * scale;
    }
    ox = x + nx;
    oy = y + ny;
    oz = z + nz;
}

// ---------------------------------------------------------------------------------------
// src/material.rs
enum class JitterKind { Color, Point };                                              // :190-194
enum class MixtureKind { Blend, Checkers, RingGradient, Ring, Gradient, Stripes };  // :226-234

struct Pattern {  // :58-65 (Box<Pattern> children -> shared immutable nodes)
    enum Kind { Debug, Plain, Jitter, Mixture } kind = Plain;
    Color color;                           // Plain
    JitterKind jitter_kind = JitterKind::Color;
    Noise noise;                           // Jitter
    MixtureKind mixture_kind = MixtureKind::Blend;
    Matrix transform_inv = Matrix::id();   // Mixture
    std::shared_ptr<const Pattern> left, right;  // Mixture: left/right; Jitter: left = inner pattern

    static Pattern debug() { Pattern p; p.kind = Debug; return p; }
    static Pattern plain(Color c) { Pattern p; p.kind = Plain; p.color = c; return p; }  // :111-113
    static Pattern new_jitter(JitterKind k, Noise n, Pattern inner) {                  // :115-121
        Pattern p; p.kind = Jitter; p.jitter_kind = k; p.noise = n;
        p.left = std::make_shared<const Pattern>(std::move(inner));
        return p;
    }
    static Pattern color_jitter(Noise n, Pattern inner) { return new_jitter(JitterKind::Color, n, std::move(inner)); }
    static Pattern point_jitter(Noise n, Pattern inner) { return new_jitter(JitterKind::Point, n, std::move(inner)); }
    static Pattern new_mixture(MixtureKind k, const Matrix& transform, Pattern l, Pattern r) {  // :131-138
        Pattern p; p.kind = Mixture; p.mixture_kind = k; p.transform_inv = transform.inverse();
        p.left = std::make_shared<const Pattern>(std::move(l));
        p.right = std::make_shared<const Pattern>(std::move(r));
        return p;
    }
    static Pattern blend(const Matrix& t, Pattern l, Pattern r) { return new_mixture(MixtureKind::Blend, t, std::move(l), std::move(r)); }
    static Pattern checkers(const Matrix& t, Pattern l, Pattern r) { return new_mixture(MixtureKind::Checkers, t, std::move(l), std::move(r)); }
    static Pattern ring_gradient(const Matrix& t, Pattern l, Pattern r) { return new_mixture(MixtureKind::RingGradient, t, std::move(l), std::move(r)); }
    static Pattern ring(const Matrix& t, Pattern l, Pattern r) { return new_mixture(MixtureKind::Ring, t, std::move(l), std::move(r)); }
    static Pattern gradient(const Matrix& t, Pattern l, Pattern r) { return new_mixture(MixtureKind::Gradient, t, std::move(l), std::move(r)); }
    static Pattern stripes(const Matrix& t, Pattern l, Pattern r) { return new_mixture(MixtureKind::Stripes, t, std::move(l), std::move(r)); }

    Color color_at(const Vector& point) const;  // :164-187
};

// src/material.rs:250-301
inline Color mixture_color_at(MixtureKind kind, const Vector& point, const Pattern& left, const Pattern& right) {
    switch (kind) {
        case MixtureKind::Blend: {
            Color l = left.color_at(point);
            Color r = right.color_at(point);
            return l.avg(r);
        }
        case MixtureKind::Checkers: {
            int32_t x = sat_i32(std::floor(point.x));
            int32_t y = sat_i32(std::floor(point.y));
            int32_t z = sat_i32(std::floor(point.z));
            return (wrap_add(wrap_add(x, y), z) % 2 == 0) ? left.color_at(point) : right.color_at(point);
        }
        case MixtureKind::RingGradient: {
            double distance = (point - Vector::point(0.0, 0.0, 0.0)).magnitude();
            double fraction = distance - std::floor(distance);
            Color l = left.color_at(point);
            Color r = right.color_at(point);
            return l + ((r - l) * fraction);
        }
        case MixtureKind::Ring:
            return (sat_i32(std::floor(std::sqrt(powi2(point.x) + powi2(point.z)))) % 2 == 0) ? left.color_at(point)
                                                                                             : right.color_at(point);
        case MixtureKind::Gradient: {
            double fraction = point.x - std::floor(point.x);
            Color l = left.color_at(point);
            Color r = right.color_at(point);
            return l + ((r - l) * fraction);
        }
        case MixtureKind::Stripes:
        default:
            return (sat_i32(std::floor(point.x)) % 2 == 0) ? left.color_at(point) : right.color_at(point);
    }
}

inline Color Pattern::color_at(const Vector& point) const {
    switch (kind) {
        case Debug: return {point.x, point.y, point.z};
        case Plain: return color;
        case Jitter:  // src/material.rs:206-223
            if (jitter_kind == JitterKind::Color) {
                Color c = left->color_at(point);
                Color o;
                noise.jitter_3d(c.r, c.g, c.b, o.r, o.g, o.b);
                return o;
            } else {
                double nx, ny, nz;
                noise.jitter_3d(point.x, point.y, point.z, nx, ny, nz);
                return left->color_at(Vector::point(nx, ny, nz));
            }
        case Mixture:
        default: {
            Vector p = transform_inv * point;
            return mixture_color_at(mixture_kind, p, *left, *right);
        }
    }
}

namespace consts { namespace transparency {  // src/material.rs:8-16
constexpr double VACUUM = 1.0, AIR = 1.00029, WATER = 1.333, GLASS = 1.52, DIAMOND = 2.417;
}}

struct Material {  // src/material.rs:18-43
    Pattern pattern = Pattern::plain(Color::white());
    double ambient = 0.1;
    double diffuse = 0.9;
    double specular = 0.9;
    double shininess = 200.0;
    double reflective = 0.0;
    double transparency = 0.0;
    double refractive_index = 1.0;
};

// ---------------------------------------------------------------------------------------
// src/shape.rs:635-653 — slab interval on one axis (shared by cube and bounding box)
inline void intersect_cube_axis(double origin, double direction, double min, double max, double& t_lo, double& t_hi) {
    double t_min_numerator = min - origin;
    double t_max_numerator = max - origin;
    double t_min, t_max;
    if (std::fabs(direction) >= EPSILON) {
        t_min = t_min_numerator / direction;
        t_max = t_max_numerator / direction;
    } else {
        t_min = t_min_numerator * std::numeric_limits<double>::infinity();
        t_max = t_max_numerator * std::numeric_limits<double>::infinity();
    }
    if (t_min > t_max) { t_lo = t_max; t_hi = t_min; } else { t_lo = t_min; t_hi = t_max; }
}

// ---------------------------------------------------------------------------------------
// src/bounding_box.rs  (f64::min/max ignore a NaN operand, like fmin/fmax)
struct BoundingBox {
    Vector min, max;
    static BoundingBox empty() {  // :19-24
        const double inf = std::numeric_limits<double>::infinity();
        return {Vector::point(inf, inf, inf), Vector::point(-inf, -inf, -inf)};
    }
    static BoundingBox make(const Vector& mn, const Vector& mx) { return {mn, mx}; }  // :26-28
    BoundingBox insert(const Vector& p) const {                                       // :30-43
        return {Vector::point(std::fmin(min.x, p.x), std::fmin(min.y, p.y), std::fmin(min.z, p.z)),
                Vector::point(std::fmax(max.x, p.x), std::fmax(max.y, p.y), std::fmax(max.z, p.z))};
    }
    BoundingBox union_(const BoundingBox& o) const { return insert(o.min).insert(o.max); }  // :45-47
    bool contains(const Vector& p) const {                                                 // :49-56
        return min.x <= p.x && p.x <= max.x && min.y <= p.y && p.y <= max.y && min.z <= p.z && p.z <= max.z;
    }
    bool encloses(const BoundingBox& o) const { return contains(o.min) && contains(o.max); }  // :58-60
    BoundingBox transform(const Matrix& m) const {                                           // :62-78
        Vector p[8] = {min,
                       Vector::point(min.x, min.y, max.z),
                       Vector::point(min.x, max.y, min.z),
                       Vector::point(min.x, max.y, max.z),
                       Vector::point(max.x, min.y, min.z),
                       Vector::point(max.x, min.y, max.z),
                       Vector::point(max.x, max.y, min.z),
                       max};
        BoundingBox b = empty();
        for (auto& q : p) b = b.insert(m * q);
        return b;
    }
    bool intersects(const Ray& ray) const {  // :80-92 (no t-range clipping)
        double x0, x1, y0, y1, z0, z1;
        intersect_cube_axis(ray.origin.x, ray.direction.x, min.x, max.x, x0, x1);
        intersect_cube_axis(ray.origin.y, ray.direction.y, min.y, max.y, y0, y1);
        intersect_cube_axis(ray.origin.z, ray.direction.z, min.z, max.z, z0, z1);
        double t_min = std::fmax(std::fmax(x0, y0), z0);
        double t_max = std::fmin(std::fmin(x1, y1), z1);
        return t_min <= t_max;
    }
};

// ---------------------------------------------------------------------------------------
// src/shape.rs:465-498
struct Geometry {
    enum Kind { Sphere, Plane, Cube, Cylinder, Cone, Triangle, SmoothTriangle } kind = Sphere;
    double min = 0, max = 0;
    bool closed = false;
    Vector p1, p2, p3, e1, e2, n, n1, n2, n3;
    BoundingBox bbox() const;  // :948-996
};

struct Shape;

// src/intersection.rs:42-47
struct Intersection {
    double t;
    const Shape* shape;
    bool has_uv;
    double u, v;
};
using Intersections = std::vector<Intersection>;

// src/shape.rs:272-286
struct ShapeArgs {
    Matrix transform = Matrix::id();
    Material material = Material{};
    bool casts_shadow = true;
};

inline std::atomic<size_t>& id_counter() {  // src/shape.rs:18-28 (thread-local counter there)
    static std::atomic<size_t> c{0};
    return c;
}

// src/shape.rs:296-306
struct Shape {
    Matrix transform_inv, transform_inv_tsp;
    BoundingBox bbox;
    Matrix material_inv;
    Material material;
    Geometry geometry;
    bool casts_shadow = true;
    size_t id = 0;

    static Shape make(ShapeArgs args, const Geometry& g) {  // :335-347
        Shape s;
        Matrix inv = args.transform.inverse();
        s.transform_inv = inv;
        s.transform_inv_tsp = inv.transpose();
        s.bbox = g.bbox().transform(args.transform);
        s.material_inv = inv;
        s.material = std::move(args.material);
        s.geometry = g;
        s.casts_shadow = args.casts_shadow;
        s.id = id_counter()++;
        return s;
    }
    void intersect(const Ray& ray, Intersections& xs) const;                         // :414-417
    Vector normal(const Vector& point, bool has_uv, double u, double v) const;       // :419-427
    Color lighting(const PointLight& light, const Vector& point, const Vector& eye,  // :429-463
                   const Vector& normal, bool shadowed) const;
};

inline BoundingBox Geometry::bbox() const {
    const double inf = std::numeric_limits<double>::infinity();
    switch (kind) {
        case Sphere:
        case Cube: return BoundingBox::make(Vector::point(-1, -1, -1), Vector::point(1, 1, 1));
        case Plane: return BoundingBox::make(Vector::point(-inf, 0, -inf), Vector::point(inf, 0, inf));
        case Cylinder:
            if (closed) return BoundingBox::make(Vector::point(-1, min, -1), Vector::point(1, max, 1));
            return BoundingBox::make(Vector::point(-1, -inf, -1), Vector::point(1, inf, 1));
        case Cone:
            if (closed) {
                double limit = std::fmax(std::fabs(min), std::fabs(max));
                return BoundingBox::make(Vector::point(-limit, min, -limit), Vector::point(limit, max, limit));
            }
            return BoundingBox::make(Vector::point(-inf, -inf, -inf), Vector::point(inf, inf, inf));
        case Triangle:
        case SmoothTriangle:
        default: return BoundingBox::empty().insert(p1).insert(p2).insert(p3);
    }
}

namespace geom {
inline void push(Intersections& xs, double t, const Shape* s) { xs.push_back({t, s, false, 0.0, 0.0}); }

// src/shape.rs:592-620
inline void intersect_sphere(const Shape* shape, const Ray& ray, Intersections& xs) {
    Vector sphere_to_ray = ray.origin - Vector::point(0, 0, 0);
    double a = ray.direction.dot(ray.direction);
    double b = 2.0 * ray.direction.dot(sphere_to_ray);
    double c = sphere_to_ray.dot(sphere_to_ray) - 1.0;
    double discriminant = powi2(b) - 4.0 * a * c;
    if (discriminant < 0.0) return;
    double t0 = (-b - std::sqrt(discriminant)) / (2.0 * a);
    double t1 = (-b + std::sqrt(discriminant)) / (2.0 * a);
    push(xs, t0, shape);
    push(xs, t1, shape);
}
// src/shape.rs:622-633
inline void intersect_plane(const Shape* shape, const Ray& ray, Intersections& xs) {
    if (approx(ray.direction.y, 0.0)) return;
    push(xs, -ray.origin.y / ray.direction.y, shape);
}
// src/shape.rs:655-680
inline void intersect_cube(const Shape* shape, const Ray& ray, Intersections& xs) {
    double x0, x1, y0, y1, z0, z1;
    intersect_cube_axis(ray.origin.x, ray.direction.x, -1.0, 1.0, x0, x1);
    intersect_cube_axis(ray.origin.y, ray.direction.y, -1.0, 1.0, y0, y1);
    intersect_cube_axis(ray.origin.z, ray.direction.z, -1.0, 1.0, z0, z1);
    double t_min = std::fmax(std::fmax(x0, y0), z0);
    double t_max = std::fmin(std::fmin(x1, y1), z1);
    if (t_min <= t_max) {
        push(xs, t_min, shape);
        push(xs, t_max, shape);
    }
}
// src/shape.rs:682-687
inline bool intersects_cap(const Ray& ray, double t, double radius) {
    double x = ray.origin.x + t * ray.direction.x;
    double z = ray.origin.z + t * ray.direction.z;
    return (powi2(x) + powi2(z)) <= powi2(radius);
}
// src/shape.rs:689-722
inline void intersect_cap(const Shape* shape, const Ray& ray, double min, double max, double min_radius,
                          double max_radius, bool closed, Intersections& xs) {
    if (!closed || approx(ray.direction.y, 0.0)) return;
    double t = (min - ray.origin.y) / ray.direction.y;
    if (intersects_cap(ray, t, min_radius)) push(xs, t, shape);
    t = (max - ray.origin.y) / ray.direction.y;
    if (intersects_cap(ray, t, max_radius)) push(xs, t, shape);
}
// src/shape.rs:724-768
inline void intersect_cylinder(const Shape* shape, const Ray& ray, double min, double max, bool closed, Intersections& xs) {
    const Vector& o = ray.origin;
    const Vector& d = ray.direction;
    double a = powi2(d.x) + powi2(d.z);
    if (!approx(a, 0.0)) {
        double b = 2.0 * o.x * d.x + 2.0 * o.z * d.z;
        double c = powi2(o.x) + powi2(o.z) - 1.0;
        double discriminant = powi2(b) - 4.0 * a * c;
        if (discriminant >= 0.0) {
            double t0 = (-b - std::sqrt(discriminant)) / (2.0 * a);
            double y0 = ray.origin.y + t0 * ray.direction.y;
            if (min < y0 && y0 < max) push(xs, t0, shape);
            double t1 = (-b + std::sqrt(discriminant)) / (2.0 * a);
            double y1 = ray.origin.y + t1 * ray.direction.y;
            if (min < y1 && y1 < max) push(xs, t1, shape);
        }
    }
    intersect_cap(shape, ray, min, max, 1.0, 1.0, closed, xs);
}
// src/shape.rs:770-822
inline void intersect_cone(const Shape* shape, const Ray& ray, double min, double max, bool closed, Intersections& xs) {
    const Vector& o = ray.origin;
    const Vector& d = ray.direction;
    double a = powi2(d.x) - powi2(d.y) + powi2(d.z);
    double b = 2.0 * o.x * d.x - 2.0 * o.y * d.y + 2.0 * o.z * d.z;
    if (!approx(a, 0.0) || !approx(b, 0.0)) {
        double c = powi2(o.x) - powi2(o.y) + powi2(o.z);
        if (!approx(a, 0.0)) {
            double discriminant = powi2(b) - 4.0 * a * c;
            if (discriminant >= 0.0) {
                double t0 = (-b - std::sqrt(discriminant)) / (2.0 * a);
                double y0 = ray.origin.y + t0 * ray.direction.y;
                if (min < y0 && y0 < max) push(xs, t0, shape);
                double t1 = (-b + std::sqrt(discriminant)) / (2.0 * a);
                double y1 = ray.origin.y + t1 * ray.direction.y;
                if (min < y1 && y1 < max) push(xs, t1, shape);
            }
        } else {
            push(xs, -c / (2.0 * b), shape);  // no y-range check in the reference
        }
    }
    intersect_cap(shape, ray, min, max, min, max, closed, xs);
}
// src/shape.rs:824-860 (Moller-Trumbore; shared by flat and smooth triangles)
inline void intersect_triangle(const Shape* shape, const Ray& ray, const Vector& p1, const Vector& e1, const Vector& e2,
                               Intersections& xs) {
    Vector dir_cross_e2 = ray.direction.cross(e2);
    double determinant = e1.dot(dir_cross_e2);
    if (std::fabs(determinant) < EPSILON) return;
    double f = 1.0 / determinant;
    Vector p1_to_origin = ray.origin - p1;
    double u = f * p1_to_origin.dot(dir_cross_e2);
    if (u < 0.0 || u > 1.0) return;
    Vector origin_cross_e1 = p1_to_origin.cross(e1);
    double v = f * ray.direction.dot(origin_cross_e1);
    if (v < 0.0 || u + v > 1.0) return;
    xs.push_back({f * e2.dot(origin_cross_e1), shape, true, u, v});
}
// src/shape.rs:887-901
inline Vector normal_cube(const Vector& p) {
    double xa = std::fabs(p.x), ya = std::fabs(p.y), za = std::fabs(p.z);
    double mx = std::fmax(std::fmax(xa, ya), za);
    if (mx == xa) return Vector::vector(p.x, 0, 0);
    if (mx == ya) return Vector::vector(0, p.y, 0);
    return Vector::vector(0, 0, p.z);
}
// src/shape.rs:903-913
inline Vector normal_cylinder(const Vector& p, double min, double max) {
    double distance = powi2(p.x) + powi2(p.z);
    if (distance < 1.0 && p.y >= max - EPSILON) return Vector::vector(0, 1, 0);
    if (distance < 1.0 && p.y <= min + EPSILON) return Vector::vector(0, -1, 0);
    return Vector::vector(p.x, 0, p.z);
}
// src/shape.rs:915-929
inline Vector normal_cone(const Vector& p, double min, double max) {
    double distance = powi2(p.x) + powi2(p.z);
    if (distance < 1.0 && p.y >= max - EPSILON) return Vector::vector(0, 1, 0);
    if (distance < 1.0 && p.y <= min + EPSILON) return Vector::vector(0, -1, 0);
    double y = std::sqrt(distance);
    if (p.y > 0.0) y = -y;
    return Vector::vector(p.x, y, p.z);
}
// src/shape.rs:862-885
inline void intersect(const Geometry& g, const Shape* shape, const Ray& ray, Intersections& xs) {
    switch (g.kind) {
        case Geometry::Sphere: intersect_sphere(shape, ray, xs); break;
        case Geometry::Plane: intersect_plane(shape, ray, xs); break;
        case Geometry::Cube: intersect_cube(shape, ray, xs); break;
        case Geometry::Cylinder: intersect_cylinder(shape, ray, g.min, g.max, g.closed, xs); break;
        case Geometry::Cone: intersect_cone(shape, ray, g.min, g.max, g.closed, xs); break;
        case Geometry::Triangle:
        case Geometry::SmoothTriangle: intersect_triangle(shape, ray, g.p1, g.e1, g.e2, xs); break;
    }
}
// src/shape.rs:931-946
inline Vector normal(const Geometry& g, const Vector& p, bool has_uv, double u, double v) {
    switch (g.kind) {
        case Geometry::Sphere: return Vector::vector(p.x, p.y, p.z);
        case Geometry::Plane: return Vector::vector(0, 1, 0);
        case Geometry::Cube: return normal_cube(p);
        case Geometry::Cylinder: return normal_cylinder(p, g.min, g.max);
        case Geometry::Cone: return normal_cone(p, g.min, g.max);
        case Geometry::Triangle: return g.n;
        case Geometry::SmoothTriangle:
        default:
            if (!has_uv) throw std::runtime_error("oracle: smooth triangle normal without u/v (reference unwrap panics)");
            return g.n2 * u + g.n3 * v + g.n1 * (1.0 - u - v);
    }
}
}  // namespace geom

inline void Shape::intersect(const Ray& ray, Intersections& xs) const {
    Ray local = ray.transform(transform_inv);
    geom::intersect(geometry, this, local, xs);
}
inline Vector Shape::normal(const Vector& point, bool has_uv, double u, double v) const {
    Vector shape_point = transform_inv * point;
    Vector shape_normal = geom::normal(geometry, shape_point, has_uv, u, v);
    Vector world_normal = transform_inv_tsp * shape_normal;
    world_normal.w = 0.0;
    return world_normal.normalize();
}
inline Color Shape::lighting(const PointLight& light, const Vector& point, const Vector& eye, const Vector& normal,
                             bool shadowed) const {
    Color color = material.pattern.color_at(material_inv * point);
    Color effective_color = color * light.intensity;
    Color intensity = light.intensity;
    Vector lightv = (light.origin - point).normalize();
    Color ambient = effective_color * material.ambient;
    Color diffuse = Color::black();
    Color specular = Color::black();
    double light_dot_normal = lightv.dot(normal);
    if (!shadowed && light_dot_normal >= 0.0) {
        diffuse = effective_color * material.diffuse * light_dot_normal;
        Vector reflect = (-lightv).reflect(normal);
        double reflect_dot_eye = reflect.dot(eye);
        if (reflect_dot_eye > 0.0) specular = intensity * material.specular * std::pow(reflect_dot_eye, material.shininess);
    }
    return ambient + diffuse + specular;
}

// ---------------------------------------------------------------------------------------
// src/shape.rs:161-178
enum class GroupKind { Union, Intersection, Difference, Aggregation };
inline bool allows_intersection(GroupKind k, bool left_hit, bool in_left, bool in_right) {
    switch (k) {
        case GroupKind::Union: return (left_hit && !in_right) || (!left_hit && !in_left);
        case GroupKind::Intersection: return (left_hit && in_right) || (!left_hit && in_left);
        case GroupKind::Difference: return (left_hit && !in_right) || (!left_hit && in_left);
        default: return true;
    }
}

struct Element;
// src/shape.rs:180-185
struct Group {
    GroupKind kind = GroupKind::Aggregation;
    BoundingBox bbox;
    std::vector<Element> children;
    bool includes(const Shape& s) const;                          // :226-228
    void filter_by_group(Intersections& xs) const;                // :230-246
    void intersect(const Ray& ray, Intersections& xs) const;      // :248-269
};

inline void sort_intersections(Intersections& xs);  // src/intersection.rs:123-125

// src/shape.rs:30-34
struct Element {
    bool is_group = false;
    Shape shape;   // valid when !is_group
    Group group;   // valid when is_group

    static Element primitive(Shape s) { Element e; e.is_group = false; e.shape = std::move(s); return e; }

    // :47-72
    void propagate_inverses(const Matrix& transform, const Matrix& inv, const Matrix& inv_tsp, const Material* material) {
        if (is_group) {
            for (auto& child : group.children) child.propagate_inverses(transform, inv, inv_tsp, material);
            group.bbox = group.bbox.transform(transform);
        } else {
            shape.transform_inv = shape.transform_inv * inv;
            shape.transform_inv_tsp = inv_tsp * shape.transform_inv_tsp;
            if (material) {
                shape.material = *material;
                shape.material_inv = inv;
            } else {
                shape.material_inv = shape.material_inv * inv;
            }
        }
    }
    // :74-101
    static Element composite(const Matrix& transform, std::optional<Material> material, GroupKind kind,
                             std::vector<Element> children) {
        if (kind != GroupKind::Aggregation && children.size() != 2)
            throw std::runtime_error("oracle: CSG group needs exactly 2 children (reference asserts)");
        Matrix inv = transform.inverse();
        Matrix inv_tsp = inv.transpose();
        BoundingBox bbox = BoundingBox::empty();
        for (auto& child : children) bbox = bbox.union_(child.bbox());
        Element e;
        e.is_group = true;
        e.group.kind = kind;
        e.group.bbox = bbox;
        e.group.children = std::move(children);
        e.propagate_inverses(transform, inv, inv_tsp, material ? &*material : nullptr);
        return e;
    }
    // :103-137
    static Element sphere(ShapeArgs a) { Geometry g; g.kind = Geometry::Sphere; return primitive(Shape::make(std::move(a), g)); }
    static Element plane(ShapeArgs a) { Geometry g; g.kind = Geometry::Plane; return primitive(Shape::make(std::move(a), g)); }
    static Element cube(ShapeArgs a) { Geometry g; g.kind = Geometry::Cube; return primitive(Shape::make(std::move(a), g)); }
    static Element cylinder(ShapeArgs a, double min, double max, bool closed) {
        Geometry g; g.kind = Geometry::Cylinder; g.min = min; g.max = max; g.closed = closed;
        return primitive(Shape::make(std::move(a), g));
    }
    static Element cone(ShapeArgs a, double min, double max, bool closed) {
        Geometry g; g.kind = Geometry::Cone; g.min = min; g.max = max; g.closed = closed;
        return primitive(Shape::make(std::move(a), g));
    }
    static Element triangle(ShapeArgs a, Vector p1, Vector p2, Vector p3) {  // src/shape.rs:369-385
        Geometry g; g.kind = Geometry::Triangle;
        g.p1 = p1; g.p2 = p2; g.p3 = p3;
        g.e1 = p2 - p1;
        g.e2 = p3 - p1;
        g.n = g.e2.cross(g.e1).normalize();
        return primitive(Shape::make(std::move(a), g));
    }
    static Element smooth_triangle(ShapeArgs a, Vector p1, Vector p2, Vector p3, Vector n1, Vector n2, Vector n3) {  // :387-412
        Geometry g; g.kind = Geometry::SmoothTriangle;
        g.p1 = p1; g.p2 = p2; g.p3 = p3;
        g.e1 = p2 - p1;
        g.e2 = p3 - p1;
        g.n1 = n1; g.n2 = n2; g.n3 = n3;
        return primitive(Shape::make(std::move(a), g));
    }
    // :139-158
    void intersect(const Ray& ray, Intersections& xs) const {
        if (is_group) group.intersect(ray, xs); else shape.intersect(ray, xs);
    }
    BoundingBox bbox() const { return is_group ? group.bbox : shape.bbox; }
    bool includes(const Shape& s) const { return is_group ? group.includes(s) : shape.id == s.id; }
};

inline bool Group::includes(const Shape& s) const {
    for (auto& c : children)
        if (c.includes(s)) return true;
    return false;
}
inline void Group::filter_by_group(Intersections& xs) const {
    bool in_left = false, in_right = false;
    size_t out = 0;
    for (size_t i = 0; i < xs.size(); i++) {  // Vec::retain visits in order, keeps order
        bool left_hit = children[0].includes(*xs[i].shape);
        bool keep = allows_intersection(kind, left_hit, in_left, in_right);
        if (left_hit) in_left = !in_left; else in_right = !in_right;
        if (keep) xs[out++] = xs[i];
    }
    xs.resize(out);
}
inline void Group::intersect(const Ray& ray, Intersections& xs) const {
    if (bbox.intersects(ray)) {
        if (kind == GroupKind::Aggregation) {
            for (auto& child : children) child.intersect(ray, xs);
        } else {
            Intersections tmp;
            for (auto& child : children) child.intersect(ray, tmp);
            sort_intersections(tmp);
            filter_by_group(tmp);
            xs.insert(xs.end(), tmp.begin(), tmp.end());
        }
    }
}

// ---------------------------------------------------------------------------------------
// src/intersection.rs
struct State {  // :9-22
    double t;
    const Shape* shape;
    Vector point, over_point, under_point, eye, normal, reflect;
    bool inside;
    double n1, n2, reflectance;
};

// :24-39
inline double schlick(const Vector& eye, const Vector& normal, double n1, double n2) {
    double cos = eye.dot(normal);
    if (n1 > n2) {
        double n = n1 / n2;
        double sin2_t = powi2(n) * (1.0 - powi2(cos));
        if (sin2_t > 1.0) return 1.0;
        cos = std::sqrt(1.0 - sin2_t);
    }
    double r0 = powi2((n1 - n2) / (n1 + n2));
    return r0 + (1.0 - r0) * powi5(1.0 - cos);
}

// :123-125 — stable sort by t; the reference panics (partial_cmp().unwrap()) on a NaN t.
inline void sort_intersections(Intersections& xs) {
    for (auto& i : xs)
        if (i.t != i.t) throw std::runtime_error("oracle: NaN intersection t (reference panics in sort)");
    std::stable_sort(xs.begin(), xs.end(), [](const Intersection& a, const Intersection& b) { return a.t < b.t; });
}
// :127-132
inline const Intersection* hit(const Intersections& xs) {
    for (auto& i : xs)
        if (i.t >= 0.0) return &i;
    return nullptr;
}
// :135-139
inline bool same(const Intersection& a, const Intersection& b) { return a.t == b.t && a.shape->id == b.shape->id; }

// :50-121
inline State prepare_state(const Intersection& self, const Ray& ray, const Intersections& xs) {
    State s;
    s.t = self.t;
    s.shape = self.shape;
    s.point = ray.position(s.t);
    s.eye = -ray.direction;
    Vector normal = s.shape->normal(s.point, self.has_uv, self.u, self.v);
    s.inside = false;
    if (normal.dot(s.eye) < 0.0) {
        normal = -normal;
        s.inside = true;
    }
    s.normal = normal;
    s.over_point = s.point + (normal * EPSILON);
    s.under_point = s.point - (normal * EPSILON);
    s.reflect = ray.direction.reflect(normal);

    std::vector<const Shape*> shapes;  // ordered container list; the HashSet only mirrors membership
    double n1 = 1.0, n2 = 1.0;
    for (auto& i : xs) {
        if (same(self, i)) n1 = shapes.empty() ? 1.0 : shapes.back()->material.refractive_index;
        auto it = std::find_if(shapes.begin(), shapes.end(), [&](const Shape* q) { return q->id == i.shape->id; });
        if (it != shapes.end()) {
            shapes.erase(std::remove_if(shapes.begin(), shapes.end(), [&](const Shape* q) { return q->id == i.shape->id; }),
                         shapes.end());
        } else {
            shapes.push_back(i.shape);
        }
        if (same(self, i)) n2 = shapes.empty() ? 1.0 : shapes.back()->material.refractive_index;
    }
    s.n1 = n1;
    s.n2 = n2;
    s.reflectance = schlick(s.eye, normal, n1, n2);
    return s;
}

// ---------------------------------------------------------------------------------------
// Ray accounting (not in the reference): one count per World::intersect query.
struct RayCounts {
    uint64_t primary = 0, shadow = 0, reflected = 0, refracted = 0;
    uint64_t total() const { return primary + shadow + reflected + refracted; }
    void add(const RayCounts& o) { primary += o.primary; shadow += o.shadow; reflected += o.reflected; refracted += o.refracted; }
};

// ---------------------------------------------------------------------------------------
// src/world.rs
struct World {
    std::vector<PointLight> lights;
    std::vector<Element> elements;

    void intersect(const Ray& ray, Intersections& xs) const {  // :18-24
        xs.clear();
        for (auto& e : elements) e.intersect(ray, xs);
    }
    bool is_shadowed(const PointLight& light, const Vector& point, Intersections& xs, RayCounts* rc = nullptr) const {  // :26-48
        Vector vector = light.origin - point;
        double distance = vector.magnitude();
        Ray ray{point, vector.normalize()};
        if (rc) rc->shadow++;
        intersect(ray, xs);
        sort_intersections(xs);
        if (const Intersection* h = hit(xs)) return h->shape->casts_shadow && h->t < distance;
        return false;
    }
    Color shade_hit(const State& state, int fuel, Intersections& xs, RayCounts* rc = nullptr) const {  // :50-82
        Color color = Color::black();
        for (auto& light : lights) {
            bool shadowed = is_shadowed(light, state.over_point, xs, rc);
            Color surface_color = state.shape->lighting(light, state.over_point, state.eye, state.normal, shadowed);
            Color reflected = reflected_color(state, fuel, xs, rc);
            Color refracted = refracted_color(state, fuel, xs, rc);
            const Material& m = state.shape->material;
            if (m.reflective > 0.0 && m.transparency > 0.0)
                color += surface_color + (reflected * state.reflectance + refracted * (1.0 - state.reflectance));
            else
                color += surface_color + (reflected + refracted);
        }
        return color;
    }
    Color reflected_color(const State& state, int fuel, Intersections& xs, RayCounts* rc = nullptr) const {  // :84-102
        if (fuel <= 0 || state.shape->material.reflective == 0.0) return Color::black();
        Ray reflect_ray{state.over_point, state.reflect};
        if (rc) rc->reflected++;
        Color color = color_at(reflect_ray, fuel - 1, xs, rc);
        return color * state.shape->material.reflective;
    }
    Color refracted_color(const State& state, int fuel, Intersections& xs, RayCounts* rc = nullptr) const {  // :104-132
        if (fuel <= 0 || state.shape->material.transparency == 0.0) return Color::black();
        double n_ratio = state.n1 / state.n2;
        double cos_i = state.eye.dot(state.normal);
        double sin2_t = powi2(n_ratio) * (1.0 - powi2(cos_i));
        if (sin2_t > 1.0) return Color::black();
        double cos_t = std::sqrt(1.0 - sin2_t);
        Vector direction = state.normal * (n_ratio * cos_i - cos_t) - state.eye * n_ratio;
        Ray refract_ray{state.under_point, direction};
        if (rc) rc->refracted++;
        return color_at(refract_ray, fuel - 1, xs, rc) * state.shape->material.transparency;
    }
    Color color_at(const Ray& ray, int fuel, Intersections& xs, RayCounts* rc = nullptr) const {  // :134-149
        intersect(ray, xs);
        sort_intersections(xs);
        if (const Intersection* h = hit(xs)) {
            Intersection hc = *h;  // the list is clobbered by the shadow queries below
            State state = prepare_state(hc, ray, xs);
            return shade_hit(state, fuel, xs, rc);
        }
        return Color::black();
    }
    static World default_world() {  // :152-183
        World w;
        w.lights.push_back({Color::white(), Vector::point(-10.0, 10.0, -10.0)});
        ShapeArgs a1;
        a1.material.pattern = Pattern::plain(Color::make(0.8, 1.0, 0.6));
        a1.material.diffuse = 0.7;
        a1.material.specular = 0.2;
        w.elements.push_back(Element::sphere(a1));
        ShapeArgs a2;
        a2.transform = Matrix::scaling(0.5, 0.5, 0.5);
        w.elements.push_back(Element::sphere(a2));
        return w;
    }
};

// ---------------------------------------------------------------------------------------
// src/camera.rs
struct Camera {
    size_t hsize = 0, vsize = 0;
    double field_of_view = 0;
    Matrix transform_inv;
    double pixel_size = 0, half_width = 0, half_height = 0;

    static Camera make(size_t hsize, size_t vsize, double fov, const Matrix& transform) {  // :16-37
        Camera c;
        double half_view = std::tan(fov / 2.0);
        double aspect = (double)hsize / (double)vsize;
        if (aspect >= 1.0) { c.half_width = half_view; c.half_height = half_view / aspect; }
        else { c.half_width = half_view * aspect; c.half_height = half_view; }
        c.pixel_size = (c.half_width * 2.0) / (double)hsize;
        c.hsize = hsize; c.vsize = vsize; c.field_of_view = fov;
        c.transform_inv = transform.inverse();
        return c;
    }
    // :39-55; (sx, sy) = sub-pixel position in [0,1), 0.5 = the reference's pixel centre
    Ray ray_at(double px, double py) const {
        double xoffset = px * pixel_size;
        double yoffset = py * pixel_size;
        double world_x = half_width - xoffset;
        double world_y = half_height - yoffset;
        Vector pixel = transform_inv * Vector::point(world_x, world_y, -1.0);
        Vector origin = Vector::point(transform_inv.d[0][3], transform_inv.d[1][3], transform_inv.d[2][3]);
        Vector direction = (pixel - origin).normalize();
        return {origin, direction};
    }
    Ray ray_at_pixel(size_t x, size_t y) const { return ray_at((double)x + 0.5, (double)y + 0.5); }
    static Matrix transform(const Vector& from, const Vector& to, const Vector& up_in) {  // :57-73
        Vector forward = (to - from).normalize();
        Vector up = up_in.normalize();
        Vector left = forward.cross(up);
        Vector true_up = left.cross(forward);
        double m[4][4] = {{left.x, left.y, left.z, 0.0},
                          {true_up.x, true_up.y, true_up.z, 0.0},
                          {-forward.x, -forward.y, -forward.z, 0.0},
                          {0.0, 0.0, 0.0, 1.0}};
        return Matrix::make(m) * Matrix::translation(-from.x, -from.y, -from.z);
    }
};

// ---------------------------------------------------------------------------------------
// src/image.rs
struct RenderOpts {
    int fuel = FUEL;             // src/config.rs:2, passed at src/image.rs:72
    int spp_x = 1, spp_y = 1;    // supersampling grid (NOT in the reference; SURVEY §8(d) definition)
    size_t x0 = 0, y0 = 0;       // window (for bounded CPU samples); w = h = 0 -> whole image
    size_t w = 0, h = 0;
    unsigned threads = 0;        // 0 -> hardware_concurrency (rayon uses all logical cores)
};

struct Image {
    size_t hsize = 0, vsize = 0;
    std::vector<Color> pixels;  // row-major y*hsize+x (src/image.rs:114-116); window renders hold w*h

    Color read(size_t x, size_t y) const { return pixels[y * hsize + x]; }

    // One pixel of src/image.rs:68-73 (+ the supersampling extension: sub-sample (i,j) replaces
    // (x+0.5, y+0.5) by (x+(i+0.5)/spp_x, y+(j+0.5)/spp_y); colours averaged unclamped, j-major).
    static Color render_pixel(const Camera& camera, const World& world, size_t x, size_t y, const RenderOpts& o,
                              Intersections& scratch, RayCounts* rc) {
        Color acc = Color::black();
        for (int j = 0; j < o.spp_y; j++)
            for (int i = 0; i < o.spp_x; i++) {
                double px = (double)x + ((double)i + 0.5) / (double)o.spp_x;
                double py = (double)y + ((double)j + 0.5) / (double)o.spp_y;
                Ray ray = camera.ray_at(px, py);
                if (rc) rc->primary++;
                scratch.clear();
                acc += world.color_at(ray, o.fuel, scratch, rc);
            }
        return acc * (1.0 / (double)(o.spp_x * o.spp_y));
    }

    // src/image.rs:65-81 — rayon's parallel iterator becomes dynamic row chunks over std::threads.
    static Image par_render(const Camera& camera, const World& world, const RenderOpts& opts = RenderOpts{},
                            RayCounts* counts = nullptr) {
        RenderOpts o = opts;
        if (o.w == 0 || o.h == 0) { o.x0 = 0; o.y0 = 0; o.w = camera.hsize; o.h = camera.vsize; }
        unsigned nt = o.threads ? o.threads : std::max(1u, std::thread::hardware_concurrency());
        Image img;
        img.hsize = o.w;
        img.vsize = o.h;
        img.pixels.assign(o.w * o.h, Color::black());
        std::atomic<size_t> next_row{0};
        std::vector<RayCounts> per_thread(nt);
        std::vector<std::string> errors(nt);
        auto worker = [&](unsigned tid) {
            Intersections scratch;
            try {
                for (;;) {
                    size_t row = next_row.fetch_add(1);
                    if (row >= o.h) break;
                    for (size_t cx = 0; cx < o.w; cx++)
                        img.pixels[row * o.w + cx] =
                            render_pixel(camera, world, o.x0 + cx, o.y0 + row, o, scratch, counts ? &per_thread[tid] : nullptr);
                }
            } catch (const std::exception& e) { errors[tid] = e.what(); }
        };
        if (nt == 1) worker(0);
        else {
            std::vector<std::thread> th;
            for (unsigned t = 0; t < nt; t++) th.emplace_back(worker, t);
            for (auto& t : th) t.join();
        }
        for (auto& e : errors)
            if (!e.empty()) throw std::runtime_error(e);
        if (counts) for (auto& c : per_thread) counts->add(c);
        return img;
    }

    // src/image.rs:93-112 — P3 text, at most 5 pixels per line, a new line at every row start.
    std::string ppm() const {
        std::string out = "P3\n" + std::to_string(hsize) + " " + std::to_string(vsize) + "\n255";
        size_t j = 0;
        for (size_t i = 0; i < pixels.size(); i++) {
            const Color& c = pixels[i];
            std::string px = std::to_string((int)clamp_channel(c.r)) + " " + std::to_string((int)clamp_channel(c.g)) + " " +
                             std::to_string((int)clamp_channel(c.b));
            if (i % hsize == 0 || j % 5 == 0) { out += "\n" + px; j = 1; }
            else { out += " " + px; j += 1; }
        }
        out += "\n";
        return out;
    }
};

// ---------------------------------------------------------------------------------------
// src/obj.rs — record grammar restated by hand (the reference uses nom 6.1.0 macros; numbers
// via nom::number::complete::double = correctly rounded decimal -> f64, same as strtod).
struct ObjData {
    std::vector<Vector> vertices, normals;
    struct Tri { size_t p[3]; size_t n[3]; bool has_n; size_t group; };
    std::vector<Tri> triangles;           // already fan-triangulated (:115-123)
    std::vector<std::string> groups;      // first-seen order; groups[0] = "Default" (:191-196)
    std::vector<std::pair<uint32_t, std::string>> ignored;  // (line number, text) (:230-232)
};

namespace obj_detail {
inline bool space1(const char*& p) {
    const char* q = p;
    while (*q == ' ' || *q == '\t') q++;
    if (q == p) return false;
    p = q;
    return true;
}
// nom recognize_float: [+-]? (digit+ ('.' digit*)? | '.' digit+) ([eE] [+-]? digit+)?
inline bool parse_double(const char*& p, double& out) {
    const char* q = p;
    if (*q == '+' || *q == '-') q++;
    const char* d0 = q;
    while (*q >= '0' && *q <= '9') q++;
    if (q > d0) {
        if (*q == '.') { q++; while (*q >= '0' && *q <= '9') q++; }
    } else {
        if (*q != '.') return false;
        q++;
        const char* f0 = q;
        while (*q >= '0' && *q <= '9') q++;
        if (q == f0) return false;
    }
    if (*q == 'e' || *q == 'E') {
        const char* e = q + 1;
        if (*e == '+' || *e == '-') e++;
        const char* e0 = e;
        while (*e >= '0' && *e <= '9') e++;
        if (e > e0) q = e;
    }
    std::string tok(p, q);
    out = std::strtod(tok.c_str(), nullptr);
    p = q;
    return true;
}
inline bool parse_usize(const char*& p, size_t& out) {  // :53-56
    const char* q = p;
    size_t v = 0;
    while (*q >= '0' && *q <= '9') { v = v * 10 + (size_t)(*q - '0'); q++; }
    if (q == p) return false;
    out = v;
    p = q;
    return true;
}
inline bool parse_xyz(const char*& p, double& x, double& y, double& z) {
    return space1(p) && parse_double(p, x) && space1(p) && parse_double(p, y) && space1(p) && parse_double(p, z);
}
// :88-109 — "v/<anything>/n" (many_till(anychar, '/')), else plain "v"
inline bool parse_face(const char*& p, size_t& v, size_t& n, bool& has_n) {
    const char* q = p;
    size_t vv, nn;
    if (parse_usize(q, vv) && *q == '/') {
        const char* r = q + 1;
        while (*r && *r != '/') r++;
        if (*r == '/') {
            r++;
            if (parse_usize(r, nn)) { v = vv; n = nn; has_n = true; p = r; return true; }
        }
    }
    q = p;
    if (parse_usize(q, vv)) { v = vv; n = 0; has_n = false; p = q; return true; }
    return false;
}
}  // namespace obj_detail

struct ObjParser {
    std::string path;
    explicit ObjParser(std::string p) : path(std::move(p)) {}

    // :156-184 + the record half of :198-233
    static ObjData parse_text(std::istream& in) {
        using namespace obj_detail;
        ObjData data;
        data.groups.push_back("Default");
        size_t group = 0;
        std::string line;
        uint32_t cnt = 1;
        while (std::getline(in, line)) {
            if (!line.empty() && line.back() == '\r') { /* BufRead::lines strips "\r\n" */ line.pop_back(); }
            const char* p = line.c_str();
            double x, y, z;
            bool matched = false;
            if (p[0] == 'v') {  // parse_vertex :58-73, then parse_normal :75-90
                const char* q = p + 1;
                if (parse_xyz(q, x, y, z)) { data.vertices.push_back(Vector::point(x, y, z)); matched = true; }
                else if (p[1] == 'n') {
                    q = p + 2;
                    if (parse_xyz(q, x, y, z)) { data.normals.push_back(Vector::vector(x, y, z)); matched = true; }
                }
            } else if (p[0] == 'f') {  // parse_faces :125-136
                const char* q = p + 1;
                if (space1(q)) {
                    std::vector<std::pair<size_t, std::pair<size_t, bool>>> idx;
                    size_t v, n; bool hn;
                    if (parse_face(q, v, n, hn)) {
                        idx.push_back({v, {n, hn}});
                        for (;;) {
                            const char* r = q;
                            if (!space1(r) || !parse_face(r, v, n, hn)) break;
                            idx.push_back({v, {n, hn}});
                            q = r;
                        }
                    }
                    if (idx.empty()) throw std::runtime_error("oracle: face record without indices (reference panics in triangulate)");
                    for (size_t i = 1; i + 1 < idx.size(); i++) {  // :115-123
                        ObjData::Tri t;
                        const size_t k[3] = {0, i, i + 1};
                        bool all_n = true;
                        for (int c = 0; c < 3; c++) {
                            t.p[c] = idx[k[c]].first;
                            t.n[c] = idx[k[c]].second.first;
                            all_n = all_n && idx[k[c]].second.second;
                        }
                        t.has_n = all_n;  // :207-223: smooth only if all three carry a normal
                        t.group = group;
                        for (int c = 0; c < 3; c++) {
                            if (t.p[c] < 1 || t.p[c] > data.vertices.size()) throw std::runtime_error("oracle: vertex index out of range (reference panics)");
                            if (t.has_n && (t.n[c] < 1 || t.n[c] > data.normals.size())) throw std::runtime_error("oracle: normal index out of range (reference panics)");
                        }
                        data.triangles.push_back(t);
                    }
                    matched = true;
                }
            } else if (p[0] == 'g') {  // parse_group :138-149
                const char* q = p + 1;
                if (space1(q)) {
                    const char* r = q;
                    while ((*r >= '0' && *r <= '9') || (*r >= 'a' && *r <= 'z') || (*r >= 'A' && *r <= 'Z')) r++;
                    if (r > q) {
                        std::string name(q, r);
                        auto it = std::find(data.groups.begin(), data.groups.end(), name);
                        if (it == data.groups.end()) { data.groups.push_back(name); group = data.groups.size() - 1; }
                        else group = (size_t)(it - data.groups.begin());
                        matched = true;
                    }
                }
            }
            if (!matched) data.ignored.push_back({cnt, line});
            cnt++;
        }
        return data;
    }

    // the Element half of :198-258.  The reference drains a HashMap (arbitrary group order);
    // here groups come in first-seen order — tie order between groups is not part of the contract.
    static Element build(const ObjData& data, const Matrix& transform, const Material& material) {
        std::vector<std::vector<Element>> per_group(data.groups.size());
        for (auto& t : data.triangles) {
            const Vector& a = data.vertices[t.p[0] - 1];
            const Vector& b = data.vertices[t.p[1] - 1];
            const Vector& c = data.vertices[t.p[2] - 1];
            if (t.has_n)
                per_group[t.group].push_back(Element::smooth_triangle(ShapeArgs{}, a, b, c, data.normals[t.n[0] - 1],
                                                                      data.normals[t.n[1] - 1], data.normals[t.n[2] - 1]));
            else
                per_group[t.group].push_back(Element::triangle(ShapeArgs{}, a, b, c));
        }
        std::vector<Element> elements;
        for (auto& children : per_group)
            if (!children.empty())
                elements.push_back(Element::composite(transform, material, GroupKind::Aggregation, std::move(children)));
        if (elements.size() == 1) return std::move(elements[0]);
        return Element::composite(Matrix::id(), std::nullopt, GroupKind::Aggregation, std::move(elements));
    }

    // :186-258
    std::pair<std::vector<std::pair<uint32_t, std::string>>, Element> parse_obj(const Matrix& transform,
                                                                                const Material& material) const {
        std::ifstream in(path);
        if (!in) throw std::runtime_error("oracle: cannot open " + path);
        ObjData data = parse_text(in);
        return {data.ignored, build(data, transform, material)};
    }
};

}  // namespace orc
