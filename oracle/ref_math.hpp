// ORACLE — TEST INFRASTRUCTURE ONLY.  CPU restatement of the reference's f32 math layer.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// use anything under oracle/.  The product (raytracingoneweekend_b200/, include/) never does.
//
// PARITY UNPINNED: the reference (Rust) has no tests, golden vectors or fixtures and cannot be
// compiled in this environment (no rustc/cargo); this file follows the reference SOURCE
// operation by operation (same association order, reciprocal-multiply division, Kahan
// determinant) and must be compiled with -ffp-contract=off and without fast-math.
//
// Follows: src/math/vec3.rs, src/math/vec4.rs, src/math/mat3x3.rs, src/math/mat4x4.rs,
//          src/utils.rs:4-17,29-35.
#pragma once
#include <cmath>
#include <cstdint>
#include <limits>

namespace ref {

constexpr float INF = std::numeric_limits<float>::infinity();  // utils.rs:30
constexpr float PI = 3.1415926535897932385f;                   // utils.rs:29

// Rust f32::max / f32::min ignore a NaN operand; fmaxf / fminf do the same.
inline float fmax_(float a, float b) { return std::fmax(a, b); }
inline float fmin_(float a, float b) { return std::fmin(a, b); }
inline float clamp(float x, float lo, float hi) { return fmin_(fmax_(x, lo), hi); }  // utils.rs:4-6
inline float degrees_to_radians(float d) { return d * PI / 180.0f; }                   // utils.rs:33-35

struct Vec3 {  // vec3.rs:2-4
    float e[3];
    Vec3() : e{0.f, 0.f, 0.f} {}
    Vec3(float x, float y, float z) : e{x, y, z} {}
    float x() const { return e[0]; }
    float y() const { return e[1]; }
    float z() const { return e[2]; }
    float operator[](int i) const { return e[i]; }
    float& operator[](int i) { return e[i]; }
    float dot(const Vec3& o) const { return e[0] * o.e[0] + e[1] * o.e[1] + e[2] * o.e[2]; }  // vec3.rs:29-31
    float length_squared() const { return dot(*this); }                                        // :32-34
    float length() const { return std::sqrt(length_squared()); }                               // :35-37
    Vec3 abs() const { return Vec3(std::fabs(e[0]), std::fabs(e[1]), std::fabs(e[2])); }       // :44-46
    float max_val() const { return fmax_(e[0], fmax_(e[1], e[2])); }                           // :47-49
    float min_val() const { return fmin_(e[0], fmin_(e[1], e[2])); }                           // :50-52
    Vec3 max(const Vec3& v) const { return Vec3(fmax_(e[0], v.e[0]), fmax_(e[1], v.e[1]), fmax_(e[2], v.e[2])); }
    Vec3 min(const Vec3& v) const { return Vec3(fmin_(e[0], v.e[0]), fmin_(e[1], v.e[1]), fmin_(e[2], v.e[2])); }
    Vec3 cross(const Vec3& o) const {  // vec3.rs:62-68
        return Vec3(e[1] * o.e[2] - e[2] * o.e[1], e[2] * o.e[0] - e[0] * o.e[2], e[0] * o.e[1] - e[1] * o.e[0]);
    }
    bool near_zero() const {  // vec3.rs:69-72
        const float eps = 1e-8f;
        return (std::fabs(e[0]) < eps) && (std::fabs(e[1]) < eps) && (std::fabs(e[2]) < eps);
    }
    inline Vec3 unit() const;  // vec3.rs:38-40
};
inline Vec3 operator-(const Vec3& a) { return Vec3(-a.e[0], -a.e[1], -a.e[2]); }
inline Vec3 operator+(const Vec3& a, const Vec3& b) { return Vec3(a.e[0] + b.e[0], a.e[1] + b.e[1], a.e[2] + b.e[2]); }
// vec3.rs:194-199: a + (-b); bitwise the same as a - b in IEEE arithmetic.
inline Vec3 operator-(const Vec3& a, const Vec3& b) { return a + (-b); }
inline Vec3 operator*(const Vec3& a, const Vec3& b) { return Vec3(a.e[0] * b.e[0], a.e[1] * b.e[1], a.e[2] * b.e[2]); }
inline Vec3 operator*(const Vec3& a, float s) { return Vec3(a.e[0] * s, a.e[1] * s, a.e[2] * s); }
inline Vec3 operator*(float s, const Vec3& a) { return a * s; }               // vec3.rs:230-235
inline Vec3 operator/(const Vec3& a, float s) { return a * (1.f / s); }       // vec3.rs:236-240 (reciprocal multiply)
inline Vec3 Vec3::unit() const { return *this / length(); }
// vec3.rs:73-76: `as u8` saturates and maps NaN to 0.
inline uint8_t f32_as_u8(float f) {
    if (!(f == f)) return 0;
    if (f <= 0.f) return 0;
    if (f >= 255.f) return 255;
    return (uint8_t)f;
}

struct Vec4 {  // vec4.rs:2-4
    float e[4];
    Vec4() : e{0.f, 0.f, 0.f, 0.f} {}
    Vec4(float x, float y, float z, float w) : e{x, y, z, w} {}
    static Vec4 new_v3(const Vec3& v) { return Vec4(v.x(), v.y(), v.z(), 0.f); }  // vec4.rs:18
    static Vec4 new_p3(const Vec3& v) { return Vec4(v.x(), v.y(), v.z(), 1.f); }  // vec4.rs:19
    float x() const { return e[0]; }
    float y() const { return e[1]; }
    float z() const { return e[2]; }
    float w() const { return e[3]; }
    float operator[](int i) const { return e[i]; }
    float dot(const Vec4& o) const { return e[0] * o.e[0] + e[1] * o.e[1] + e[2] * o.e[2] + e[3] * o.e[3]; }  // vec4.rs:27-29
    Vec3 xyz() const { return Vec3(e[0], e[1], e[2]); }
};
inline Vec4 operator-(const Vec4& a) { return Vec4(-a.e[0], -a.e[1], -a.e[2], -a.e[3]); }
inline Vec4 operator+(const Vec4& a, const Vec4& b) { return Vec4(a.e[0] + b.e[0], a.e[1] + b.e[1], a.e[2] + b.e[2], a.e[3] + b.e[3]); }
inline Vec4 operator-(const Vec4& a, const Vec4& b) { return a + (-b); }
inline Vec4 operator*(const Vec4& a, const Vec4& b) { return Vec4(a.e[0] * b.e[0], a.e[1] * b.e[1], a.e[2] * b.e[2], a.e[3] * b.e[3]); }
inline Vec4 operator*(const Vec4& a, float s) { return Vec4(a.e[0] * s, a.e[1] * s, a.e[2] * s, a.e[3] * s); }
inline Vec4 operator*(float s, const Vec4& a) { return a * s; }

struct Mat3x3 {  // mat3x3.rs:6-8 (rows)
    Vec3 e[3];
    static Mat3x3 new_3vec(const Vec3& v0, const Vec3& v1, const Vec3& v2) { Mat3x3 m; m.e[0] = v0; m.e[1] = v1; m.e[2] = v2; return m; }
    static Mat3x3 new_3vec_vert(const Vec3& v0, const Vec3& v1, const Vec3& v2) {  // mat3x3.rs:22-26
        return new_3vec(Vec3(v0.x(), v1.x(), v2.x()), Vec3(v0.y(), v1.y(), v2.y()), Vec3(v0.z(), v1.z(), v2.z()));
    }
    Vec3 dot(const Vec3& v) const { return Vec3(e[0].dot(v), e[1].dot(v), e[2].dot(v)); }  // :27-29
    float at(int r, int c) const { return e[r].e[c]; }
    Vec3 at_col(int c) const { return Vec3(e[0][c], e[1][c], e[2][c]); }
    Mat3x3 transpose() const { return new_3vec(at_col(0), at_col(1), at_col(2)); }  // :49-51
    float determinant() const {  // mat3x3.rs:52-68 (Kahan summation of the six products)
        float sum = 0.f, c = 0.f;
        const float input[6] = {(at(0, 0) * at(1, 1) * at(2, 2)),  (at(0, 1) * at(1, 2) * at(2, 0)),
                                (at(0, 2) * at(1, 0) * at(2, 1)),  (-at(0, 0) * at(1, 2) * at(2, 1)),
                                (-at(0, 1) * at(1, 0) * at(2, 2)), (-at(0, 2) * at(1, 1) * at(2, 0))};
        for (float i : input) {
            // volatile keeps the compiler from re-associating the compensation away
            volatile float y = i - c;
            volatile float t = sum + y;
            c = (t - sum) - y;
            sum = t;
        }
        return sum;
    }
    Mat3x3 inverse() const {  // mat3x3.rs:69-81
        const float det = determinant();
        Vec3 row0(at(1, 1) * at(2, 2) - at(1, 2) * at(2, 1), at(0, 2) * at(2, 1) - at(0, 1) * at(2, 2), at(0, 1) * at(1, 2) - at(0, 2) * at(1, 1));
        Vec3 row1(at(1, 2) * at(2, 0) - at(1, 0) * at(2, 2), at(0, 0) * at(2, 2) - at(0, 2) * at(2, 0), at(0, 2) * at(1, 0) - at(0, 0) * at(1, 2));
        Vec3 row2(at(1, 0) * at(2, 1) - at(1, 1) * at(2, 0), at(0, 1) * at(2, 0) - at(0, 0) * at(2, 1), at(0, 0) * at(1, 1) - at(0, 1) * at(1, 0));
        const float s = 1.f / det;  // Div<f32> = multiply by reciprocal, mat3x3.rs:92-96
        return new_3vec(row0 * s, row1 * s, row2 * s);
    }
};

struct Mat4x4 {  // mat4x4.rs:8-10 (rows)
    Vec4 e[4];
    static Mat4x4 identity() { Mat4x4 m; m.e[0] = Vec4(1, 0, 0, 0); m.e[1] = Vec4(0, 1, 0, 0); m.e[2] = Vec4(0, 0, 1, 0); m.e[3] = Vec4(0, 0, 0, 1); return m; }
    static Mat4x4 new_4vec(const Vec4& a, const Vec4& b, const Vec4& c, const Vec4& d) { Mat4x4 m; m.e[0] = a; m.e[1] = b; m.e[2] = c; m.e[3] = d; return m; }
    static Mat4x4 new_4vec_vert(const Vec4& v0, const Vec4& v1, const Vec4& v2, const Vec4& v3) {  // :32-37
        return new_4vec(Vec4(v0.x(), v1.x(), v2.x(), v3.x()), Vec4(v0.y(), v1.y(), v2.y(), v3.y()),
                        Vec4(v0.z(), v1.z(), v2.z(), v3.z()), Vec4(v0.w(), v1.w(), v2.w(), v3.w()));
    }
    static Mat4x4 new_m3x3(const Mat3x3& m) {  // :38-43
        return new_4vec(Vec4::new_v3(m.e[0]), Vec4::new_v3(m.e[1]), Vec4::new_v3(m.e[2]), Vec4(0, 0, 0, 1));
    }
    static Mat4x4 from_rows16(const float* p) {
        Mat4x4 m;
        for (int r = 0; r < 4; r++) m.e[r] = Vec4(p[4 * r], p[4 * r + 1], p[4 * r + 2], p[4 * r + 3]);
        return m;
    }
    void to_rows16(float* p) const {
        for (int r = 0; r < 4; r++)
            for (int c = 0; c < 4; c++) p[4 * r + c] = e[r].e[c];
    }
    Vec4 dot(const Vec4& v) const { return Vec4(e[0].dot(v), e[1].dot(v), e[2].dot(v), e[3].dot(v)); }  // :45-47
    Vec3 dot_p3(const Vec3& v) const { Vec4 v4 = Vec4::new_p3(v); return Vec3(e[0].dot(v4), e[1].dot(v4), e[2].dot(v4)); }  // :49-52
    Vec3 dot_v3(const Vec3& v) const { Vec4 v4 = Vec4::new_v3(v); return Vec3(e[0].dot(v4), e[1].dot(v4), e[2].dot(v4)); }  // :54-57
    Vec4 at_row(int r) const { return e[r]; }
    Vec4 at_col(int c) const { return Vec4(e[0][c], e[1][c], e[2][c], e[3][c]); }
    Mat4x4 transpose() const { return new_4vec(at_col(0), at_col(1), at_col(2), at_col(3)); }
    Mat4x4 dot_mat(const Mat4x4& m) const {  // :66-72
        Mat4x4 t = m.transpose(), r;
        for (int i = 0; i < 4; i++) r.e[i] = Vec4(e[i].dot(t.e[0]), e[i].dot(t.e[1]), e[i].dot(t.e[2]), e[i].dot(t.e[3]));
        return r;
    }
    static Mat4x4 new_translate(const Vec3& v) {  // :88-93
        return new_4vec(Vec4(1, 0, 0, v.x()), Vec4(0, 1, 0, v.y()), Vec4(0, 0, 1, v.z()), Vec4(0, 0, 0, 1));
    }
    static Mat4x4 new_scale(const Vec3& v) {  // :94-99
        return new_4vec(Vec4(v.x(), 0, 0, 0), Vec4(0, v.y(), 0, 0), Vec4(0, 0, v.z(), 0), Vec4(0, 0, 0, 1));
    }
    static Mat4x4 new_rotate_x(float f) {  // :100-107
        float c = std::cos(f), s = std::sin(f);
        return new_4vec(Vec4(1, 0, 0, 0), Vec4(0, c, s, 0), Vec4(0, -s, c, 0), Vec4(0, 0, 0, 1));
    }
    static Mat4x4 new_rotate_y(float f) {  // :108-115
        float c = std::cos(f), s = std::sin(f);
        return new_4vec(Vec4(c, 0, -s, 0), Vec4(0, 1, 0, 0), Vec4(s, 0, c, 0), Vec4(0, 0, 0, 1));
    }
    static Mat4x4 new_rotate_z(float f) {  // :116-123
        float c = std::cos(f), s = std::sin(f);
        return new_4vec(Vec4(c, -s, 0, 0), Vec4(s, c, 0, 0), Vec4(0, 0, 1, 0), Vec4(0, 0, 0, 1));
    }
    Mat4x4 fast_homogenous_inverse() const {  // :59-64   M = T S  ->  S^-1 T^-1
        Mat4x4 s_inv = new_m3x3(Mat3x3::new_3vec(e[0].xyz(), e[1].xyz(), e[2].xyz()).inverse());
        Mat4x4 t_inv = new_translate(-at_col(3).xyz());
        return s_inv.dot_mat(t_inv);
    }
    // :125-147 — returns (translate, rotate, scale)
    void decompose_into_translate_rotate_scale(Mat4x4& translate, Mat4x4& rotate, Mat4x4& scale) const {
        Vec3 a = at_col(0).xyz(), b = at_col(1).xyz(), c = at_col(2).xyz(), d = at_col(3).xyz();
        float al = a.length(), bl = b.length(), cl = c.length();
        translate = new_4vec_vert(Vec4(1, 0, 0, 0), Vec4(0, 1, 0, 0), Vec4(0, 0, 1, 0), Vec4::new_p3(d));
        rotate = new_4vec_vert(Vec4::new_v3(a / al), Vec4::new_v3(b / bl), Vec4::new_v3(c / cl), Vec4::new_p3(Vec3()));
        scale = new_4vec(Vec4(al, 0, 0, 0), Vec4(0, bl, 0, 0), Vec4(0, 0, cl, 0), Vec4(0, 0, 0, 1));
    }
    Vec4 diag() const { return Vec4(e[0].x(), e[1].y(), e[2].z(), e[3].w()); }  // :148-150
};

}  // namespace ref
