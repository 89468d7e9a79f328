// qt5_math.h — TEST INFRASTRUCTURE (oracle only; never included by the product).
//
// Restatement of the arithmetic semantics of the Qt value classes the reference's hot path
// uses (QVector3D, QVector4D, QQuaternion, QMatrix4x4).  Qt is a third-party dependency that
// is NOT vendored under /root/reference and is not installed in this image, so its source
// cannot be compiled here.  Pinned version: Qt 5.15 (CMakeLists.txt:13 `find_package(Qt5 ...)`,
// README.md); the shipped Windows binary was built with Qt 6.4.0 (gamer.pro.user:94-96).
// What is restated is Qt 5.15's published behaviour (qtbase src/gui/math3d/):
//   * all components are stored as float32; `double * QVector3D` narrows the scalar to float
//     first (the overload is operator*(float, const QVector3D&));
//   * dotProduct is float mul/add, left to right;
//   * length()/normalized() accumulate the squared length in double, take sqrt in double and
//     round to float; normalized() returns the input when |len2 - 1| <= 1e-12 and the zero
//     vector when len2 <= 1e-12;
//   * QQuaternion product is the 9-multiplication form; rotatedVector(v) = (q * (0,v) * q~).vector();
//   * fromAxisAndAngle(axis, deg): a = radians(deg/2) in float, (cos a, axis_n * sin a).normalized();
//   * QMatrix4x4 is column-major m[col][row]; M*v4 is four float dot rows.
// Qt 6 differs in two places (length() via qHypot in float: <= 1 ulp; normalized() keeps the
// input when |len-1| <= 1e-5).  Both are far below the 1e-3 parity budget and, for everything
// but the per-ray normalisations, common-mode between oracle and device (same inputs handed to both).
#pragma once
#include <cmath>

namespace q5 {

static inline bool fuzzy_null_d(double d) { return std::fabs(d) <= 0.000000000001; }
static inline bool fuzzy_null_f(float f)  { return std::fabs(f) <= 0.00001f; }

struct V3 {
    float x = 0.f, y = 0.f, z = 0.f;
    V3() {}
    V3(float a, float b, float c) : x(a), y(b), z(c) {}
};
static inline V3 operator+(const V3 &a, const V3 &b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline V3 operator-(const V3 &a, const V3 &b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline V3 operator-(const V3 &a)              { return V3(-a.x, -a.y, -a.z); }
static inline V3 operator*(const V3 &a, float f)     { return V3(a.x * f, a.y * f, a.z * f); }
static inline V3 operator*(float f, const V3 &a)     { return V3(a.x * f, a.y * f, a.z * f); }
static inline V3 operator/(const V3 &a, float f)     { return V3(a.x / f, a.y / f, a.z / f); }
static inline float dot(const V3 &a, const V3 &b)    { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline V3 cross(const V3 &a, const V3 &b) {
    return V3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
static inline double len2_d(const V3 &a) {
    return double(a.x) * double(a.x) + double(a.y) * double(a.y) + double(a.z) * double(a.z);
}
static inline float length(const V3 &a) { return float(std::sqrt(len2_d(a))); }
static inline V3 normalized(const V3 &a) {
    double len = len2_d(a);
    if (fuzzy_null_d(len - 1.0f)) return a;
    if (!fuzzy_null_d(len)) {
        double s = std::sqrt(len);
        return V3(float(double(a.x) / s), float(double(a.y) / s), float(double(a.z) / s));
    }
    return V3();
}

struct Quat {
    float w = 1.f, x = 0.f, y = 0.f, z = 0.f;
    Quat() {}
    Quat(float ww, float xx, float yy, float zz) : w(ww), x(xx), y(yy), z(zz) {}
};
static inline Quat normalized(const Quat &q) {
    double len = double(q.x) * double(q.x) + double(q.y) * double(q.y) +
                 double(q.z) * double(q.z) + double(q.w) * double(q.w);
    if (fuzzy_null_d(len - 1.0f)) return q;
    if (!fuzzy_null_d(len)) {
        float s = float(std::sqrt(len));           // QQuaternion operator/(q, float)
        return Quat(q.w / s, q.x / s, q.y / s, q.z / s);
    }
    return Quat(0.f, 0.f, 0.f, 0.f);
}
static inline Quat mul(const Quat &a, const Quat &b) {
    float yy = (a.w - a.y) * (b.w + b.z);
    float zz = (a.w + a.y) * (b.w - b.z);
    float ww = (a.z + a.x) * (b.x + b.y);
    float xx = ww + yy + zz;
    float qq = 0.5f * (xx + (a.z - a.x) * (b.x - b.y));
    float w = qq - ww + (a.z - a.y) * (b.y - b.z);
    float x = qq - xx + (a.x + a.w) * (b.x + b.w);
    float y = qq - yy + (a.w - a.x) * (b.y + b.z);
    float z = qq - zz + (a.z + a.y) * (b.w - b.x);
    return Quat(w, x, y, z);
}
static inline V3 rotate(const Quat &q, const V3 &v) {
    Quat r = mul(mul(q, Quat(0.f, v.x, v.y, v.z)), Quat(q.w, -q.x, -q.y, -q.z));
    return V3(r.x, r.y, r.z);
}
static inline float deg2rad_f(float deg) { return deg * float(3.14159265358979323846 / 180.0); }
static inline Quat from_axis_angle(const V3 &axis, float deg) {
    float a = deg2rad_f(deg / 2.0f);
    float s = std::sin(a);
    float c = std::cos(a);
    V3 ax = normalized(axis);
    return normalized(Quat(c, ax.x * s, ax.y * s, ax.z * s));
}
static inline Quat rotation_to(const V3 &from, const V3 &to) {
    V3 v0 = normalized(from), v1 = normalized(to);
    float d = dot(v0, v1) + 1.0f;
    if (fuzzy_null_f(d)) {
        V3 axis = cross(V3(1.f, 0.f, 0.f), v0);
        if (fuzzy_null_f(dot(axis, axis))) axis = cross(V3(0.f, 1.f, 0.f), v0);
        axis = normalized(axis);
        return Quat(0.f, axis.x, axis.y, axis.z);
    }
    d = std::sqrt(2.0f * d);
    V3 axis = cross(v0, v1) / d;
    return normalized(Quat(d * 0.5f, axis.x, axis.y, axis.z));
}

// column-major like QMatrix4x4: m[col][row]
struct M4 {
    float m[4][4];
    static M4 identity() {
        M4 r;
        for (int c = 0; c < 4; c++) for (int w = 0; w < 4; w++) r.m[c][w] = (c == w) ? 1.f : 0.f;
        return r;
    }
};
static inline M4 mul(const M4 &a, const M4 &b) {   // a*b
    M4 r;
    for (int c = 0; c < 4; c++)
        for (int w = 0; w < 4; w++)
            r.m[c][w] = a.m[0][w] * b.m[c][0] + a.m[1][w] * b.m[c][1] +
                        a.m[2][w] * b.m[c][2] + a.m[3][w] * b.m[c][3];
    return r;
}
static inline M4 perspective(float vdeg, float aspect, float zn, float zf) {
    M4 r = M4::identity();
    float rad = deg2rad_f(vdeg / 2.0f);
    float sine = std::sin(rad);
    if (zn == zf || aspect == 0.f || sine == 0.f) return r;
    float cotan = std::cos(rad) / sine;
    float clip = zf - zn;
    for (int c = 0; c < 4; c++) for (int w = 0; w < 4; w++) r.m[c][w] = 0.f;
    r.m[0][0] = cotan / aspect;
    r.m[1][1] = cotan;
    r.m[2][2] = -(zn + zf) / clip;
    r.m[2][3] = -1.0f;
    r.m[3][2] = -(2.0f * zn * zf) / clip;
    return r;
}
static inline M4 look_at(const V3 &eye, const V3 &center, const V3 &up) {
    V3 f = center - eye;
    if (fuzzy_null_f(f.x) && fuzzy_null_f(f.y) && fuzzy_null_f(f.z)) return M4::identity();
    f = normalized(f);
    V3 s = normalized(cross(f, up));
    V3 u = cross(s, f);
    M4 r = M4::identity();
    r.m[0][0] = s.x;  r.m[1][0] = s.y;  r.m[2][0] = s.z;
    r.m[0][1] = u.x;  r.m[1][1] = u.y;  r.m[2][1] = u.z;
    r.m[0][2] = -f.x; r.m[1][2] = -f.y; r.m[2][2] = -f.z;
    // translate(-eye): last column += M * (-eye)
    M4 t = M4::identity();
    t.m[3][0] = -eye.x; t.m[3][1] = -eye.y; t.m[3][2] = -eye.z;
    return mul(r, t);
}
// general inverse: Qt copies to double, forms cofactors / determinant in double, stores float.
static inline M4 inverted(const M4 &a) {
    double s[4][4];
    for (int c = 0; c < 4; c++) for (int w = 0; w < 4; w++) s[w][c] = a.m[c][w];   // s[row][col]
    auto det3 = [&](int r0, int r1, int r2, int c0, int c1, int c2) {
        return s[r0][c0] * (s[r1][c1] * s[r2][c2] - s[r1][c2] * s[r2][c1]) -
               s[r0][c1] * (s[r1][c0] * s[r2][c2] - s[r1][c2] * s[r2][c0]) +
               s[r0][c2] * (s[r1][c0] * s[r2][c1] - s[r1][c1] * s[r2][c0]);
    };
    int idx[4][3] = {{1, 2, 3}, {0, 2, 3}, {0, 1, 3}, {0, 1, 2}};
    double cof[4][4];
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++) {
            double d = det3(idx[r][0], idx[r][1], idx[r][2], idx[c][0], idx[c][1], idx[c][2]);
            cof[r][c] = ((r + c) & 1) ? -d : d;
        }
    double det = s[0][0] * cof[0][0] + s[0][1] * cof[0][1] + s[0][2] * cof[0][2] + s[0][3] * cof[0][3];
    M4 inv = M4::identity();
    if (det == 0.0) return inv;
    double id = 1.0 / det;
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++) inv.m[c][r] = float(cof[c][r] * id);   // inverse(row r, col c) = cof(c,r)/det
    return inv;
}

}  // namespace q5
