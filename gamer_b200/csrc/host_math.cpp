// host_math.cpp — host-side pieces of the parameter model that produce kernel inputs:
//   GamerCamera::setupViewmatrix   source/galaxy/gamercamera.cpp:185-208   -> gamer_b200_camera_inv_vp
//   GalaxyInstance::setupQuaternions  source/galaxy/galaxyinstance.cpp:69-71 -> gamer_b200_rotation_to_orientation
// The reference does these once per frame with Qt's float matrix/quaternion classes; the same
// float32 steps are taken here (projection, look-at and their product in float; the inverse via
// double cofactors stored as float, as QMatrix4x4::inverted does).
#include <cmath>
#include "../../include/gamer_b200.h"

namespace {

struct Mat4 { float e[4][4]; };   // e[row][col]

Mat4 ident() {
    Mat4 m;
    for (int r = 0; r < 4; r++) for (int c = 0; c < 4; c++) m.e[r][c] = (r == c) ? 1.f : 0.f;
    return m;
}
Mat4 matmul(const Mat4 &a, const Mat4 &b) {
    Mat4 m;
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++)
            m.e[r][c] = a.e[r][0] * b.e[0][c] + a.e[r][1] * b.e[1][c] + a.e[r][2] * b.e[2][c] + a.e[r][3] * b.e[3][c];
    return m;
}
void unit3(float v[3]) {   // QVector3D::normalize, Qt 5.15: double length
    double l2 = double(v[0]) * v[0] + double(v[1]) * v[1] + double(v[2]) * v[2];
    if (std::fabs(l2 - 1.0) <= 1e-12) return;
    if (std::fabs(l2) <= 1e-12) { v[0] = v[1] = v[2] = 0.f; return; }
    double l = std::sqrt(l2);
    for (int i = 0; i < 3; i++) v[i] = float(v[i] / l);
}
void cross3(const float a[3], const float b[3], float o[3]) {
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}
float radians_f(float deg) { return deg * float(3.14159265358979323846 / 180.0); }

double det3(const double s[4][4], const int r[3], const int c[3]) {
    return s[r[0]][c[0]] * (s[r[1]][c[1]] * s[r[2]][c[2]] - s[r[1]][c[2]] * s[r[2]][c[1]]) -
           s[r[0]][c[1]] * (s[r[1]][c[0]] * s[r[2]][c[2]] - s[r[1]][c[2]] * s[r[2]][c[0]]) +
           s[r[0]][c[2]] * (s[r[1]][c[0]] * s[r[2]][c[1]] - s[r[1]][c[1]] * s[r[2]][c[0]]);
}

}  // namespace

extern "C" {

int gamer_b200_camera_inv_vp(const float cam[3], const float target[3], const float up[3], double fov_deg, float out[16]) {
    if (!cam || !target || !up || !out) return GAMER_E_ARG;
    // projection: perspective(fov, aspect 1, near 1, far 100)
    Mat4 proj = ident();
    {
        float half = radians_f(float(fov_deg) / 2.0f);
        float sine = std::sin(half);
        if (sine != 0.f) {
            float cot = std::cos(half) / sine, clip = 100.f - 1.f;
            for (int r = 0; r < 4; r++) for (int c = 0; c < 4; c++) proj.e[r][c] = 0.f;
            proj.e[0][0] = cot / 1.f;
            proj.e[1][1] = cot;
            proj.e[2][2] = -(1.f + 100.f) / clip;
            proj.e[2][3] = -(2.0f * 1.f * 100.f) / clip;
            proj.e[3][2] = -1.0f;
        }
    }
    // view: lookAt(eye = target, center = camera, up) — the swap is the reference's (gamercamera.cpp:200)
    Mat4 view = ident();
    {
        float f[3] = {cam[0] - target[0], cam[1] - target[1], cam[2] - target[2]};
        bool degenerate = std::fabs(f[0]) <= 0.00001f && std::fabs(f[1]) <= 0.00001f && std::fabs(f[2]) <= 0.00001f;
        if (!degenerate) {
            unit3(f);
            float s[3], u[3];
            cross3(f, up, s);
            unit3(s);
            cross3(s, f, u);
            Mat4 rot = ident();
            for (int c = 0; c < 3; c++) { rot.e[0][c] = s[c]; rot.e[1][c] = u[c]; rot.e[2][c] = -f[c]; }
            Mat4 tr = ident();
            for (int r = 0; r < 3; r++) tr.e[r][3] = -target[r];
            view = matmul(rot, tr);
        }
    }
    Mat4 pv = matmul(proj, view);
    double s[4][4];
    for (int r = 0; r < 4; r++) for (int c = 0; c < 4; c++) s[r][c] = pv.e[r][c];
    const int pick[4][3] = {{1, 2, 3}, {0, 2, 3}, {0, 1, 3}, {0, 1, 2}};
    double cof[4][4];
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++) {
            double d = det3(s, pick[r], pick[c]);
            cof[r][c] = ((r + c) & 1) ? -d : d;
        }
    double det = 0;
    for (int c = 0; c < 4; c++) det += s[0][c] * cof[0][c];
    if (det == 0.0) {
        Mat4 id = ident();
        for (int r = 0; r < 4; r++) for (int c = 0; c < 4; c++) out[r * 4 + c] = id.e[r][c];
        return GAMER_OK;
    }
    double inv = 1.0 / det;
    for (int r = 0; r < 4; r++) for (int c = 0; c < 4; c++) out[r * 4 + c] = float(cof[c][r] * inv);
    return GAMER_OK;
}

int gamer_b200_rotation_to_orientation(const float orient[3], float q[4]) {
    if (!orient || !q) return GAMER_E_ARG;
    float from[3] = {0.f, 1.f, 0.f}, to[3] = {orient[0], orient[1], orient[2]};
    unit3(to);
    float d = from[0] * to[0] + from[1] * to[1] + from[2] * to[2] + 1.0f;
    float ax[3];
    if (std::fabs(d) <= 0.00001f) {          // opposite: any perpendicular axis, half turn
        const float ex[3] = {1.f, 0.f, 0.f}, ey[3] = {0.f, 1.f, 0.f};
        cross3(ex, from, ax);
        if (std::fabs(ax[0] * ax[0] + ax[1] * ax[1] + ax[2] * ax[2]) <= 0.00001f) cross3(ey, from, ax);
        unit3(ax);
        q[0] = 0.f; q[1] = ax[0]; q[2] = ax[1]; q[3] = ax[2];
        return GAMER_OK;
    }
    d = std::sqrt(2.0f * d);
    cross3(from, to, ax);
    float w = d * 0.5f, x = ax[0] / d, y = ax[1] / d, z = ax[2] / d;
    double l2 = double(x) * x + double(y) * y + double(z) * z + double(w) * w;
    if (!(std::fabs(l2 - 1.0) <= 1e-12) && std::fabs(l2) > 1e-12) {
        float l = float(std::sqrt(l2));
        w /= l; x /= l; y /= l; z /= l;
    }
    q[0] = w; q[1] = x; q[2] = y; q[3] = z;
    return GAMER_OK;
}

// ---- star list: the expressions of Buffer2D::RenderStars (util/buffer2d.cpp:140-160)
// over a counter-based generator. Draw j of star i is splitmix64(seed, 16 i + j): no hidden state.
namespace {
inline uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
struct StarDraws {
    uint64_t seed, base;
    int j = 0;
    uint64_t next() { return splitmix64(splitmix64(seed) ^ (base + uint64_t(j++))); }
    double uniform() { return double(next() >> 11) * (1.0 / 9007199254740992.0); }            // [0,1)
    double gaussian(double mean, double sigma) {                                              // Box-Muller
        const double u1 = 1.0 - uniform(), u2 = uniform();                                    // u1 in (0,1]
        return mean + sigma * std::sqrt(-2.0 * std::log(u1)) * std::cos(6.283185307179586 * u2);
    }
};
}  // namespace

int gamer_b200_star_list(int32_t size, int32_t no_stars, double base_size_d, double size_spread_d, double strength,
                         uint64_t seed, GamerStar *out) {
    if (size <= 0 || no_stars < 0 || (no_stars > 0 && !out)) return GAMER_E_ARG;
    const int baseSize = int(base_size_d), sizeSpread = int(size_spread_d);   // RenderStars takes ints (buffer2d.cpp:140)
    for (int32_t i = 0; i < no_stars; i++) {
        StarDraws r{seed, uint64_t(i) * 16};
        GamerStar &s = out[i];
        s.x = int32_t(r.next() % uint64_t(size));
        s.y = int32_t(r.next() % uint64_t(size));
        const double cx = std::fmin(r.uniform() + 0.6, 1.0);
        const double cy = std::fmin(r.uniform() + double(0.6f), cx);
        const double cz = std::fmin(r.uniform() + 0.6, 1.0);
        const double sz = std::fmax(r.gaussian(baseSize, sizeSpread), baseSize / 3.0);
        s.w = int32_t(int(sz * size) / 245.0);
        const double ss = sz * std::fabs(r.gaussian(strength, strength));
        s.cs[0] = float(ss) * float(cx);        // double * QVector3D narrows the scalar first
        s.cs[1] = float(ss) * float(cy);
        s.cs[2] = float(ss) * float(cz);
    }
    return GAMER_OK;
}

}  // extern "C"
