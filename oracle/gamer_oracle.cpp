// gamer_oracle.cpp — TEST INFRASTRUCTURE.  CPU restatement of the reference's hot path.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library, and only as the checker or the timed CPU baseline — never as a product path.
//
// PARITY PINNED — against the reference's own code, compiled in place and run in the build container.
//   * whole path (rasterizer.cpp Prepare/RenderPixels/renderPixel/getIntensity, galaxycomponent.h/.cpp,
//     galaxycomponents.cpp, galaxy.cpp SetupComponents + .gax reader, galaxyinstance.cpp, gamercamera.cpp,
//     buffer2d.cpp ToColorBuffer/PostProcess, util.h IntersectSphere/clamp/smoothstep/floor, noise/*.cpp):
//     `make -C oracle ref` compiles those UNMODIFIED sources where they lie under /root/reference into
//     oracle/_ref/libgamer_ref_path*.so (door: oracle/ref_path_wrap.cpp). tests/golden/ref_path_frames.npz
//     holds 26 scenes rendered by it — all nine shipped galaxies, five views, preview mode, tilted / flipped /
//     displaced instances, synthetic galaxies covering the classes and branches no shipped file uses, 3- and
//     11-galaxy scenes — and tests/test_ref_path.py requires this oracle to reproduce EVERY FLOAT of every
//     frame bit-for-bit, every ARGB32 pixel of AssembleImage at four exposure/gamma/saturation settings,
//     and the camera rays. It does, in both readings of the return-less path of
//     GalaxyComponent::calculateIntensity (galaxycomponent.cpp:62-88; GAMER_FRAME_NO_RADIAL_GATE in
//     gamer_b200.h explains the two).
//   * what is NOT the reference's code in that build: Qt. Its headers are absent here, so the QVector3D /
//     QQuaternion / QMatrix4x4 arithmetic is oracle/qt5_math.h's restatement of Qt 5.15 (shared by this
//     oracle and by the stand-in classes of oracle/qtshim/qtshim.h the reference sources compile against),
//     QThread::start runs synchronously and OpenMP is off. The pin is therefore "modulo Qt value-class
//     arithmetic": control flow, formulas, evaluation order, casts and libm calls are the reference's own.
//   * noise layer: additionally pinned on its own (tests/golden/simplex_ref.npz, noise_class_ref.npz): 1009
//     raw values + fBm / ridged cases from the reference's simplexnoise.cpp / noise.cpp / simplex.cpp.
//   * "stars small" (galaxycomponents.cpp:159-170) draws from libc rand() per sample and is not
//     reproducible; it is dropped here, in the product and in the reference's tools/galaxy_repro.py:734-737.
//   * the star layer's random draws (buffer2d.cpp:128-165, libc rand()) are likewise not reproducible; only
//     the stamping arithmetic is restated (go_stamp_stars).
//
// Precision follows the reference exactly: scalars are double, vectors are float32 Qt value
// classes, `double * QVector3D` narrows the scalar to float first, M_PI is 3.14159265 in the
// arm code (source/util/util.h:20) — and nowhere else.
//
// Build: see oracle/Makefile (-O2 -ffp-contract=off strict build = the parity oracle;
// -Ofast -march=native build = the reference's own flags, gamer.pro:13,20,23 = the speed baseline).
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#include <algorithm>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "../include/gamer_b200.h"
#include "qt5_math.h"

using q5::V3;
using q5::Quat;

namespace {

// ---------------------------------------------------------------------------------------------
// L0 noise: 3-D simplex (source/noise/simplexnoise.cpp:243-333, tables simplexnoise.h:130-185)
// ---------------------------------------------------------------------------------------------
// Ken Perlin's reference permutation (the reference stores it twice, 512 entries; indices here
// are reduced with &255 which reads the same values).
const uint8_t kPerm[256] = {
    151, 160, 137, 91,  90,  15,  131, 13,  201, 95,  96,  53,  194, 233, 7,   225, 140, 36,  103, 30,
    69,  142, 8,   99,  37,  240, 21,  10,  23,  190, 6,   148, 247, 120, 234, 75,  0,   26,  197, 62,
    94,  252, 219, 203, 117, 35,  11,  32,  57,  177, 33,  88,  237, 149, 56,  87,  174, 20,  125, 136,
    171, 168, 68,  175, 74,  165, 71,  134, 139, 48,  27,  166, 77,  146, 158, 231, 83,  111, 229, 122,
    60,  211, 133, 230, 220, 105, 92,  41,  55,  46,  245, 40,  244, 102, 143, 54,  65,  25,  63,  161,
    1,   216, 80,  73,  209, 76,  132, 187, 208, 89,  18,  169, 200, 196, 135, 130, 116, 188, 159, 86,
    164, 100, 109, 198, 173, 186, 3,   64,  52,  217, 226, 250, 124, 123, 5,   202, 38,  147, 118, 126,
    255, 82,  85,  212, 207, 206, 59,  227, 47,  16,  58,  17,  182, 189, 28,  42,  223, 183, 170, 213,
    119, 248, 152, 2,   44,  154, 163, 70,  221, 153, 101, 155, 167, 43,  172, 9,   129, 22,  39,  253,
    19,  98,  108, 110, 79,  113, 224, 232, 178, 185, 112, 104, 218, 246, 97,  228, 251, 34,  242, 193,
    238, 210, 144, 12,  191, 179, 162, 241, 81,  51,  145, 235, 249, 14,  239, 107, 49,  192, 214, 31,
    181, 199, 106, 157, 184, 84,  204, 176, 115, 121, 50,  45,  127, 4,   150, 254, 138, 236, 205, 93,
    222, 114, 67,  29,  24,  72,  243, 141, 128, 195, 78,  66,  215, 61,  156, 180};
// midpoints of the edges of a cube (simplexnoise.h:142-146)
const int kGrad3[12][3] = {{1, 1, 0}, {-1, 1, 0}, {1, -1, 0}, {-1, -1, 0}, {1, 0, 1},  {-1, 0, 1},
                           {1, 0, -1}, {-1, 0, -1}, {0, 1, 1}, {0, -1, 1}, {0, 1, -1}, {0, -1, -1}};

inline int perm_at(int i) { return kPerm[i & 255]; }

// simplexnoise.h:130 — NOT floor(): off by one for exact non-positive integers. Kept as is.
inline int fast_floor(double v) { return v > 0 ? (int)v : (int)v - 1; }

struct SimplexLattice {   // what "bit-exact integer hashing" compares
    int i, j, k, c1, c2, g[4];
};

// simplexnoise.cpp:243-333. Evaluation order of every double expression follows the reference.
double simplex3(double x, double y, double z, SimplexLattice *lat = nullptr) {
    const double F3 = 1.0 / 3.0, G3 = 1.0 / 6.0;
    double s = (x + y + z) * F3;
    int i = fast_floor(x + s), j = fast_floor(y + s), k = fast_floor(z + s);
    double t = (i + j + k) * G3;
    double X0 = i - t, Y0 = j - t, Z0 = k - t;
    double x0 = x - X0, y0 = y - Y0, z0 = z - Z0;

    int i1, j1, k1, i2, j2, k2;   // the six orderings, :267-276
    if (x0 >= y0) {
        if (y0 >= z0)      { i1 = 1; j1 = 0; k1 = 0; i2 = 1; j2 = 1; k2 = 0; }
        else if (x0 >= z0) { i1 = 1; j1 = 0; k1 = 0; i2 = 1; j2 = 0; k2 = 1; }
        else               { i1 = 0; j1 = 0; k1 = 1; i2 = 1; j2 = 0; k2 = 1; }
    } else {
        if (y0 < z0)       { i1 = 0; j1 = 0; k1 = 1; i2 = 0; j2 = 1; k2 = 1; }
        else if (x0 < z0)  { i1 = 0; j1 = 1; k1 = 0; i2 = 0; j2 = 1; k2 = 1; }
        else               { i1 = 0; j1 = 1; k1 = 0; i2 = 1; j2 = 1; k2 = 0; }
    }
    double cx[4] = {x0, x0 - i1 + G3, x0 - i2 + 2.0 * G3, x0 - 1.0 + 3.0 * G3};
    double cy[4] = {y0, y0 - j1 + G3, y0 - j2 + 2.0 * G3, y0 - 1.0 + 3.0 * G3};
    double cz[4] = {z0, z0 - k1 + G3, z0 - k2 + 2.0 * G3, z0 - 1.0 + 3.0 * G3};
    int oi[4] = {0, i1, i2, 1}, oj[4] = {0, j1, j2, 1}, ok[4] = {0, k1, k2, 1};
    int ii = i & 255, jj = j & 255, kk = k & 255;
    double n[4];
    int g[4];
    for (int c = 0; c < 4; c++) {
        g[c] = perm_at(ii + oi[c] + perm_at(jj + oj[c] + perm_at(kk + ok[c]))) % 12;
        double tt = 0.6 - cx[c] * cx[c] - cy[c] * cy[c] - cz[c] * cz[c];
        if (tt < 0) n[c] = 0.0;
        else {
            tt *= tt;
            n[c] = tt * tt * (kGrad3[g[c]][0] * cx[c] + kGrad3[g[c]][1] * cy[c] + kGrad3[g[c]][2] * cz[c]);
        }
    }
    if (lat) {
        lat->i = i; lat->j = j; lat->k = k;
        lat->c1 = i1 + 2 * j1 + 4 * k1; lat->c2 = i2 + 2 * j2 + 4 * k2;
        for (int c = 0; c < 4; c++) lat->g[c] = g[c];
    }
    return 32.0 * (n[0] + n[1] + n[2] + n[3]);
}

struct Counters { uint64_t c[GAMER_CNT_N] = {0, 0, 0, 0, 0, 0, 0, 0}; };

// Noise::octave_noise_3d (noise.cpp:162-180). `octaves` is a double in the reference; loop is i < octaves.
double fbm3(double octaves, double persistence, double scale, double x, double y, double z, Counters *cn) {
    double total = 0, frequency = scale, amplitude = 1, max_amp = 0;
    for (int i = 0; i < octaves; i++) {
        total += simplex3(x * frequency, y * frequency, z * frequency) * amplitude;
        frequency *= 2;
        max_amp += amplitude;
        amplitude *= persistence;
        if (cn) cn->c[GAMER_CNT_OCTAVES]++;
    }
    return total / max_amp;
}

// Noise::getRidgedMf (noise.cpp:81-128). p is a float QVector3D and is scaled in float per octave
// (`vt = vt * lacunarity` narrows lacunarity to float). The lazily built `weights` table (:88-89)
// is never read by the reference and is left out.
double ridged3(V3 p, double frequency, int octaves, double lacunarity, double offset, double gain, Counters *cn) {
    double value = 0.0, weight = 1.0;
    double w = -0.05f;
    V3 vt = p;
    for (int o = 0; o < octaves; o++) {
        double signal = simplex3(vt.x, vt.y, vt.z);
        if (cn) cn->c[GAMER_CNT_OCTAVES]++;
        signal = std::fabs(signal);
        signal = offset - signal;
        signal *= signal;
        signal *= weight;
        weight = signal * gain;
        if (weight > 1.0) weight = 1.0;
        if (weight < 0.0) weight = 0.0;
        value += (signal * std::pow(frequency, w));
        vt = vt * float(lacunarity);
        frequency *= lacunarity;
    }
    return (value * 1.25) - 1.0;
}

// ---------------------------------------------------------------------------------------------
// util (source/util/util.h:62-98, util.cpp:76-80,114-120)
// ---------------------------------------------------------------------------------------------
const double kArmPi = 3.14159265;   // util.h:20 `#define M_PI 3.14159265`, seen by galaxycomponent.h:103-146

inline double u_clamp(double val, double mi, double ma) {   // std::min(ma,val) then std::max(mi,val)
    val = (val < ma) ? val : ma;
    val = (mi < val) ? val : mi;
    return val;
}
inline double u_smoothstep(double e0, double e1, double x) {
    x = u_clamp((x - e0) / (e1 - e0), 0.0, 1.0);   // e1==e0: x/0 -> +inf or NaN -> clamp gives 1
    return x * x * (3 - 2 * x);
}
inline V3 u_floor(const V3 &v) {   // Util::floor = per-channel fmax(0.0f, .)
    return V3(std::fmax(0.0f, v.x), std::fmax(0.0f, v.y), std::fmax(0.0f, v.z));
}

struct Hit { bool ok; V3 isp1, isp2; double t0, t1; };
// Util::IntersectSphere (util.h:66-98): ellipsoid with semi-axes r.
Hit intersect_ellipsoid(V3 o, V3 d, V3 r) {
    r.x = float(1.0 / (r.x * r.x));
    r.y = float(1.0 / (r.y * r.y));
    r.z = float(1.0 / (r.z * r.z));
    V3 rD(d.x * r.x, d.y * r.y, d.z * r.z);
    V3 rO(o.x * r.x, o.y * r.y, o.z * r.z);
    double A = q5::dot(d, rD);
    double B = 2.0 * (q5::dot(d, rO));
    double C = q5::dot(o, rO) - 1.0;
    double S = (B * B - 4.0f * A * C);
    Hit h;
    if (S <= 0) { h.ok = false; h.t0 = h.t1 = 0; return h; }
    h.t0 = (-B - std::sqrt(S)) / (2.0 * A);
    h.t1 = (-B + std::sqrt(S)) / (2.0 * A);
    h.isp1 = o + d * float(h.t0);
    h.isp2 = o + d * float(h.t1);
    h.ok = true;
    return h;
}

// ---------------------------------------------------------------------------------------------
// L1 components (galaxycomponent.h:57-165, galaxycomponent.cpp:45-88, galaxycomponents.cpp)
// ---------------------------------------------------------------------------------------------
struct Pixel {          // RasterPixel (rasterpixel.h:5-27)
    V3 I;
    double winding = 0, radius = 0, z = 0, tmp = 0, step = 0, scale = 1;
    V3 P, dir;
};

struct Ctx {            // what a GalaxyComponent sees
    const GamerGalaxy *gal;
    const GamerComponent *cp;
    V3 orient;
    Quat rotmat;
    Counters *cn;
    bool radial_gate = true;   // false: GAMER_FRAME_NO_RADIAL_GATE (the test g++ -O1+ deletes, see gamer_b200.h)
};

inline double c_cosh(double x) { return (std::exp(x) + std::exp(-x)) / 2.0; }   // galaxycomponent.h:57-59

// getHeightModulation (galaxycomponent.h:61-70); Bulge override returns 1 (galaxycomponents.cpp:27-29)
inline double height_modulation(const Ctx &c, double height) {
    if (c.cp->class_id == GAMER_CLASS_BULGE) return 1;
    double h = std::fabs(height / c.cp->z0);
    if (h > 2.0) return 0;
    double val = 1.0f / c_cosh(h);
    return val * val;
}
// getRadius — the base, non-virtual one is what the march calls (galaxycomponent.h:72-76)
inline double get_radius(const Ctx &c, const V3 &p, V3 &P, double &dott) {
    dott = q5::dot(p, c.orient);
    P = p - c.orient * float(dott);
    return q5::length(P) / c.gal->axis[0];
}
inline double radial_intensity(const Ctx &c, double rad) {   // :78-81
    double r = std::exp(-rad / (c.cp->r0 * 0.5f));
    return u_clamp(r - 0.01f, 0, 1);
}
double g_dbg_twirl_rel = 0.0;   // debug: relative perturbation of the twirl angle (sensitivity studies only)
inline V3 twirl(const Ctx &c, const V3 &p, double tw) {      // :86-90
    if (g_dbg_twirl_rel != 0.0) tw *= (1.0 + g_dbg_twirl_rel);
    Quat q = q5::from_axis_angle(c.orient, float(tw * 180.0));
    return q5::rotate(q, p);
}
inline double cloud_noise(const Ctx &c, const V3 &p, double t, int NN, double ks, double pers) {   // :93-98
    V3 r = twirl(c, p, t);
    return fbm3(NN, pers, ks * 0.1f, r.x, r.y, r.z, c.cn);
}
inline double find_difference(double t1, double t2) {        // :103-118
    double v1 = std::fabs(t1 - t2);
    double v2 = std::fabs(t1 - t2 - 2 * kArmPi);
    double v3 = std::fabs(t1 - t2 + 2 * kArmPi);
    double v4 = std::fabs(t1 - t2 - 2 * kArmPi * 2);
    double v5 = std::fabs(t1 - t2 + 2 * kArmPi * 2);
    double v = std::fmin(v1, v2);
    v = std::fmin(v, v3);
    v = std::fmin(v, v4);
    v = std::fmin(v, v5);
    return v;
}
inline double get_theta(const Ctx &c, const V3 &p) {         // :150-153
    V3 q = q5::rotate(c.rotmat, p);
    return std::atan2(q.x, q.z) + c.cp->delta;
}
inline double get_winding(const Ctx &c, double rad) {        // :156-165
    double r = rad + 0.05;
    double t = std::atan(std::exp(-0.25 / (0.5 * (r))) / c.gal->winding_b) * 2 * c.gal->winding_n;
    double scale = 1.0, t2 = 0.0;
    t = t * scale + t2 * (1 - scale);
    return t;
}
inline double get_arm(const Ctx &c, double rad, const V3 &p, double disp) {   // :140-146
    double ww = get_winding(c, rad);
    double th = -get_theta(c, p);
    double v = std::fabs(find_difference(ww, th + disp)) / kArmPi;
    return std::pow(1.0 - v, c.cp->arm * 15);
}
inline double arm_value(const Ctx &c, double rad, const V3 &P) {              // :120-137
    double v1 = get_arm(c, rad, P, c.gal->arm[0]);
    if (c.gal->no_arms == 1) return v1;
    double v = std::fmax(v1, get_arm(c, rad, P, c.gal->arm[1]));
    if (c.gal->no_arms == 2) return v;
    v = std::fmax(v, get_arm(c, rad, P, c.gal->arm[2]));
    if (c.gal->no_arms == 3) return v;
    v = std::fmax(v, get_arm(c, rad, P, c.gal->arm[3]));
    return v;
}

inline V3 spec_of(const Ctx &c) { return V3(c.cp->spec[0], c.cp->spec[1], c.cp->spec[2]); }

// dust-type absorption, shared by Dust (:78-82) and Dust2 (:105-109)
inline void absorb(const Ctx &c, Pixel &rp, double p2, double ival) {
    double s = 0.01;
    rp.I.x = float(rp.I.x * std::exp(-p2 * ival * c.cp->spec[0] * s));
    rp.I.y = float(rp.I.y * std::exp(-p2 * ival * c.cp->spec[1] * s));
    rp.I.z = float(rp.I.z * std::exp(-p2 * ival * c.cp->spec[2] * s));
}

// componentIntensity, one case per class (galaxycomponents.cpp)
void component_intensity(const Ctx &c, Pixel &rp, V3 &p, double ival) {
    const GamerComponent &cp = *c.cp;
    switch (cp.class_id) {
    case GAMER_CLASS_BULGE: {            // :5-18
        double rho_0 = cp.strength * ival;
        V3 pos = q5::rotate(c.rotmat, p);
        double rad = (q5::length(pos) + 0.01) * cp.r0;
        rad += 0.01;
        double i = rho_0 * (std::pow(rad, -0.855) * std::exp(-std::pow(rad, 1 / 4.0)) - 0.05) * c.gal->intensity_scale;
        if (i < 0) i = 0;
        rp.I = rp.I + float(i) * spec_of(c) * float(rp.scale);
        rp.tmp = i;
        c.cn->c[GAMER_CNT_COMP_EVALS]++;
        break;
    }
    case GAMER_CLASS_DISK: {             // :41-59
        if (ival < 0.0005) return;
        c.cn->c[GAMER_CNT_COMP_EVALS]++;
        double p2 = std::fabs(cloud_noise(c, p, rp.winding, 10, cp.scale, cp.ks));
        p2 = std::fmax(p2, 0.01f);
        p2 = std::pow(p2, cp.noise_tilt);
        p2 += cp.noise_offset;
        if (p2 < 0) return;
        rp.I = rp.I + float(ival * p2) * spec_of(c) * float(rp.scale);
        break;
    }
    case GAMER_CLASS_DUST: {             // :64-84
        if (ival < 0.0005) return;
        c.cn->c[GAMER_CNT_COMP_EVALS]++;
        double p2 = cloud_noise(c, p, rp.winding, 9, cp.scale, cp.ks);
        p2 = std::fmax(p2 - cp.noise_offset, 0.0f);
        p2 = u_clamp(std::pow(5 * p2, cp.noise_tilt), -10, 10);
        absorb(c, rp, p2, ival);
        break;
    }
    case GAMER_CLASS_DUST2:              // :85-114
    case GAMER_CLASS_DUST_POSITIVE: {    // :116-136
        if (ival < 0.0005) return;
        c.cn->c[GAMER_CNT_COMP_EVALS]++;
        double scale = cp.scale;
        V3 r = twirl(c, p, rp.winding);
        double p2 = ridged3(r * float(scale), cp.ks, 9, 2.5, cp.noise_offset, cp.noise_tilt, c.cn);
        p2 = std::fmax(p2, 0.0f);
        if (cp.class_id == GAMER_CLASS_DUST2) absorb(c, rp, p2, ival);
        else rp.I = rp.I + float(ival * p2) * spec_of(c) * float(rp.scale);
        break;
    }
    case GAMER_CLASS_STARS: {            // :139-155
        c.cn->c[GAMER_CNT_COMP_EVALS]++;
        double perlinnoise = std::fabs(fbm3(10, cp.ks, 0.01f * cp.scale * 100, p.x, p.y, p.z, c.cn));
        double add = 0;
        if (cp.noise_offset != 0) {
            add = (cp.noise_offset * cloud_noise(c, p, rp.winding, 4, 2, -2));
            add += 0.5f * (cp.noise_offset * cloud_noise(c, p, rp.winding * 0.5f, 4, 4, -2));
        }
        double val = std::fabs(std::pow(perlinnoise + 1 + add, cp.noise_tilt));
        rp.I = rp.I + float(ival * val) * spec_of(c) * float(rp.scale);
        break;
    }
    default:                             // "stars small": rand()-driven, dropped (see gamer_b200.h)
        break;
    }
}

// calculateIntensity: base (galaxycomponent.cpp:45-88) and the Bulge override (galaxycomponents.cpp:20-25)
void calculate_intensity(const Ctx &c, Pixel &rp, V3 &p, double weight) {
    const GamerComponent &cp = *c.cp;
    if (cp.class_id == GAMER_CLASS_BULGE) { component_intensity(c, rp, p, weight); return; }
    double armVal = 1, winding = 0;
    if (rp.z > 0.01) {
        double intensity = radial_intensity(c, rp.radius);
        if (intensity > 0.1) intensity = 0.1;
        if (intensity > 0.001 || !c.radial_gate) {
            double scale = std::pow(u_smoothstep(0, 1.0f * cp.inner, rp.radius), 4);
            if (cp.arm != 0) {
                armVal = arm_value(c, rp.radius, rp.P);
                if (cp.winding != 0) winding = get_winding(c, rp.radius) * cp.winding;
            }
            rp.winding = winding;
            double val = (cp.strength) * scale * armVal * rp.z * intensity * c.gal->intensity_scale;
            if (val * weight > 0.0005) component_intensity(c, rp, p, val * weight);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// L2 pixel loop (rasterizer.cpp:283-313,373-483; gamercamera.cpp:210-217)
// ---------------------------------------------------------------------------------------------
struct Scene {
    GamerFrame f;
    std::vector<GamerGalaxy> gal;
    std::vector<GamerComponent> comp;
};

// GamerCamera::coord2ray (gamercamera.cpp:210-217): QMatrix4x4 * QVector4D, then .toVector3D().normalized()
V3 coord2ray(const GamerFrame &f, double x, double y, double width) {
    double xx = x / (width * 0.5) - 1.0;
    double yy = y / (width * 0.5) - 1.0;
    float v[4] = {float(xx), float(-yy), 1.0f, 1.0f};
    const float *m = f.inv_vp;   // [row*4+col]
    float o[3];
    for (int r = 0; r < 3; r++) o[r] = v[0] * m[r * 4 + 0] + v[1] * m[r * 4 + 1] + v[2] * m[r * 4 + 2] + v[3] * m[r * 4 + 3];
    return q5::normalized(V3(o[0], o[1], o[2]));
}

// debug trace of one pixel's march (go_trace_pixel): (p.xyz, step, I.xyz) after every sample
static float *g_trace = nullptr;
static int g_trace_n = 0, g_trace_max = 0;

// Rasterizer::getIntensity (rasterizer.cpp:422-483)
void march(const Scene &sc, const GamerGalaxy &g, Pixel &rp, V3 isp1, V3 isp2, Counters &cn) {
    V3 origin = isp1;
    double length = q5::length(isp1 - isp2);
    V3 dir = q5::normalized(isp1 - isp2);
    double step = sc.f.ray_step;
    V3 p = origin;
    rp.scale = step;
    V3 camera = V3(sc.f.cam[0], sc.f.cam[1], sc.f.cam[2]) - V3(g.pos[0], g.pos[1], g.pos[2]);
    double minRayStep = sc.f.min_ray_step;
    rp.dir = dir;
    V3 ll = q5::normalized(isp2 - origin);
    Ctx c;
    c.gal = &g;
    c.orient = V3(g.orient[0], g.orient[1], g.orient[2]);
    c.rotmat = Quat(g.rotmat[0], g.rotmat[1], g.rotmat[2], g.rotmat[3]);
    c.cn = &cn;
    c.radial_gate = !(sc.f.flags & GAMER_FRAME_NO_RADIAL_GATE);
    cn.c[GAMER_CNT_RAYS]++;
    while (q5::dot(p - origin, ll) < length + step) {
        step = u_clamp(std::pow(q5::length(p - camera), 1) * sc.f.ray_step, minRayStep, 0.01);
        rp.step = step;
        for (int k = 0; k < g.comp_count; k++) {
            c.cp = &sc.comp[g.comp_begin + k];
            if (c.cp->active == 1) {
                cn.c[GAMER_CNT_GATES]++;
                rp.radius = get_radius(c, p, rp.P, rp.z);
                rp.z = height_modulation(c, rp.z);
                calculate_intensity(c, rp, p, step * 200.0);
            }
        }
        if (g_trace && g_trace_n < g_trace_max) {
            float *o = g_trace + (size_t)g_trace_n * 7;
            o[0] = p.x; o[1] = p.y; o[2] = p.z; o[3] = float(step); o[4] = rp.I.x; o[5] = rp.I.y; o[6] = rp.I.z;
        }
        if (g_trace) g_trace_n++;
        p = p - dir * float(step);
        cn.c[GAMER_CNT_SAMPLES]++;
        rp.I = u_floor(rp.I);
    }
}

// Rasterizer::renderPixel (rasterizer.cpp:379-416)
V3 render_pixel(const Scene &sc, V3 dir, Counters &cn) {
    Pixel rp;
    V3 cam(sc.f.cam[0], sc.f.cam[1], sc.f.cam[2]);
    V3 isp1, isp2;   // like the reference, these persist across galaxies when a galaxy is missed
    for (int gi = 0; gi < sc.f.n_galaxies; gi++) {
        const GamerGalaxy &g = sc.gal[gi];
        V3 o = cam - V3(g.pos[0], g.pos[1], g.pos[2]);
        Hit h = intersect_ellipsoid(o, dir, V3(g.axis[0], g.axis[1], g.axis[2]));
        bool intersects = h.ok;
        isp1 = h.ok ? h.isp1 : V3();
        isp2 = h.ok ? h.isp2 : V3();
        if (h.t1 > 0) isp2 = o;                       // camera inside: clip at the camera
        if (h.t0 > 0 && h.t1 > 0) intersects = false; // volume entirely on the far side of the swapped ray
        if (intersects) march(sc, g, rp, isp1, isp2, cn);
    }
    return (rp.I * float(0.01)) / float(sc.f.ray_step);   // rp.I()*0.01/rayStep : both scalars narrow to float
}

inline int image_row(const GamerFrame &f, int L) {
    if (f.band_rows > 0) return f.y0 + (L / f.band_rows) * f.band_stride + (L % f.band_rows);
    return f.y0 + L;
}

// Buffer2D::PostProcess (buffer2d.cpp:106-126) -> (r,g,b) of the saved image, each 0..255
inline void post_process(const float in[3], double exposure, double gamma, double saturation, int out_rgb[3]) {
    V3 v(in[0], in[1], in[2]);
    v = v * float(1.0 / exposure);
    v.x = float(std::pow(v.x, gamma));
    v.y = float(std::pow(v.y, gamma));
    v.z = float(std::pow(v.z, gamma));
    double center = (v.x + v.y + v.z) / 3.0;   // float sum, then /3.0 in double
    V3 tmp(float(center - v.x), float(center - v.y), float(center - v.z));
    v.x = float(center - saturation * tmp.x);
    v.y = float(center - saturation * tmp.y);
    v.z = float(center - saturation * tmp.z);
    double s = 10;
    V3 vs = v * float(s);
    V3 cc(float(u_clamp(vs.x, 0, 255)), float(u_clamp(vs.y, 0, 255)), float(u_clamp(vs.z, 0, 255)));
    out_rgb[0] = (int)cc.x;   // QColor(int,int,int) truncates
    out_rgb[1] = (int)cc.y;
    out_rgb[2] = (int)cc.z;
}

}  // namespace

// =============================================================================================
// C entry points (ctypes)
// =============================================================================================
extern "C" {

int go_abi_version(void) { return GAMER_B200_ABI_VERSION; }
void go_debug_twirl_perturb(double rel) { g_dbg_twirl_rel = rel; }

int go_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

// Rasterizer::RenderPixelsOMP (rasterizer.cpp:283-313). The shuffled render order and the DrawBox
// progressive splat (buffer2d.cpp:52-63) do not change the final image (every pixel is eventually
// rendered and written at its own location), so each pixel is rendered once and stored once.
// Debug aid for tests: march pixel (i, j) single-threaded and record every sample. Returns the sample count.
int go_trace_pixel(const GamerFrame *frame, const GamerGalaxy *gal, const GamerComponent *comp, int i, int j,
                   int max_samples, float *out7) {
    Scene sc;
    sc.f = *frame;
    sc.gal.assign(gal, gal + frame->n_galaxies);
    sc.comp.assign(comp, comp + frame->n_components);
    Counters cn;
    g_trace = out7; g_trace_n = 0; g_trace_max = max_samples;
    V3 dir = coord2ray(sc.f, i, j, frame->full_size);
    render_pixel(sc, dir, cn);
    g_trace = nullptr;
    return g_trace_n;
}

int go_render(const GamerFrame *frame, const GamerGalaxy *gal, const GamerComponent *comp,
              float *rgb_out, uint64_t counters_out[GAMER_CNT_N], int n_threads) {
    if (!frame || !gal || !comp || !rgb_out) return GAMER_E_ARG;
    Scene sc;
    sc.f = *frame;
    sc.gal.assign(gal, gal + frame->n_galaxies);
    sc.comp.assign(comp, comp + frame->n_components);
    const int W = frame->width, H = frame->height;
    Counters total;
#ifdef _OPENMP
    if (n_threads <= 0) n_threads = omp_get_max_threads();
#pragma omp parallel num_threads(n_threads)
#endif
    {
        Counters local;
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 64)
#endif
        for (int k = 0; k < W * H; k++) {
            int lx = k % W, ly = k / W;
            int i = frame->x0 + lx, j = image_row(*frame, ly);   // idx = i + j*size (rasterizer.cpp:304-305)
            V3 dir = coord2ray(sc.f, i, j, frame->full_size);
            V3 I = render_pixel(sc, dir, local);
            rgb_out[3 * (size_t)k + 0] = I.x;
            rgb_out[3 * (size_t)k + 1] = I.y;
            rgb_out[3 * (size_t)k + 2] = I.z;
        }
#ifdef _OPENMP
#pragma omp critical
#endif
        for (int q = 0; q < GAMER_CNT_N; q++) total.c[q] += local.c[q];
    }
    if (counters_out) for (int q = 0; q < GAMER_CNT_N; q++) counters_out[q] = total.c[q];
    return GAMER_OK;
}

// Rasterizer::AssembleImage (rasterizer.cpp:315-327): CopyTo, Add(stars), ToColorBuffer (buffer2d.cpp:65-85)
int go_tonemap(const float *rgb, int width, int height, double exposure, double gamma, double saturation,
               const float *stars_or_null, uint32_t *argb_out, uint32_t *argb_shadow_out) {
    if (!rgb) return GAMER_E_ARG;
    size_t n = (size_t)width * height;
#ifdef _OPENMP
#pragma omp parallel for
#endif
    for (long long k = 0; k < (long long)n; k++) {
        float v[3] = {rgb[3 * k], rgb[3 * k + 1], rgb[3 * k + 2]};
        if (stars_or_null) { v[0] += stars_or_null[3 * k]; v[1] += stars_or_null[3 * k + 1]; v[2] += stars_or_null[3 * k + 2]; }
        int c[3];
        post_process(v, exposure, gamma, saturation, c);
        // PostProcess returns QColor(c.z, c.y, c.x): image gets that; shadowImage gets it R/B swapped back.
        if (argb_out)        argb_out[k]        = 0xFF000000u | ((uint32_t)c[2] << 16) | ((uint32_t)c[1] << 8) | (uint32_t)c[0];
        if (argb_shadow_out) argb_shadow_out[k] = 0xFF000000u | ((uint32_t)c[0] << 16) | ((uint32_t)c[1] << 8) | (uint32_t)c[2];
    }
    return GAMER_OK;
}

// Buffer2D::RenderStars' drawing part (util/buffer2d.cpp:140-143,161) + Buffer2D::RenderGaussian
// (:224-243) + Util::maxQvector3D (util/util.h:175-177), in star order, sequentially like the reference.
// The star parameters themselves come from the caller (gamer_b200_star_list documents why).
int go_stamp_stars(int size, const GamerStar *stars, int n, float *layer_out) {
    if (size <= 0 || !layer_out || (n > 0 && !stars)) return GAMER_E_ARG;
    std::fill(layer_out, layer_out + (size_t)size * size * 3, 0.0f);            // fill(QVector3D(0,0,0))
    for (int q = 0; q < n; q++) {
        const int i = stars[q].x, j = stars[q].y, w = stars[q].w;
        const float *cs = stars[q].cs;
        for (int x = -w / 2; x < w / 2; x++)
            for (int y = -w / 2; y < w / 2; y++) {
                int xx = i + x, yy = j + y;
                double dx = x / (double)w, dy = y / (double)w;
                if (xx >= 0 && xx < size && yy >= 0 && yy < size) {
                    double d = (dx * dx + dy * dy);
                    double ss = 0.01f;
                    double v = exp(-d / ss);
                    float *px = layer_out + 3 * ((size_t)xx + (size_t)yy * size);
                    const float vf = (float)v;                                  // double * QVector3D narrows first
                    for (int c = 0; c < 3; c++) px[c] = (float)fmax(vf * cs[c], px[c]);
                }
            }
    }
    return GAMER_OK;
}

double go_raw3d(double x, double y, double z) { return simplex3(x, y, z); }

int go_simplex(const float *xyz, int n, double frequency, int32_t *lattice_out, double *value_out) {
    for (int q = 0; q < n; q++) {
        SimplexLattice L;
        double v = simplex3(xyz[3 * q] * frequency, xyz[3 * q + 1] * frequency, xyz[3 * q + 2] * frequency, &L);
        if (value_out) value_out[q] = v;
        if (lattice_out) {
            int32_t *o = lattice_out + 10 * (size_t)q;
            o[0] = L.i; o[1] = L.j; o[2] = L.k; o[3] = L.c1; o[4] = L.c2;
            o[5] = L.g[0]; o[6] = L.g[1]; o[7] = L.g[2]; o[8] = L.g[3]; o[9] = 0;
        }
    }
    return GAMER_OK;
}

int go_fbm(const float *xyz, int n, int octaves, double persistence, double scale, double *value_out) {
    for (int q = 0; q < n; q++)
        value_out[q] = fbm3(octaves, persistence, scale, xyz[3 * q], xyz[3 * q + 1], xyz[3 * q + 2], nullptr);
    return GAMER_OK;
}

int go_ridged(const float *xyz, int n, double frequency, int octaves, double lacunarity, double offset,
              double gain, double *value_out) {
    for (int q = 0; q < n; q++)
        value_out[q] = ridged3(V3(xyz[3 * q], xyz[3 * q + 1], xyz[3 * q + 2]), frequency, octaves, lacunarity, offset, gain, nullptr);
    return GAMER_OK;
}

// GamerCamera::setupViewmatrix (gamercamera.cpp:185-208) with m_rotMatrix = identity (its value in the
// CLI path). NOTE eye/centre are swapped in the reference: lookAt(target, camera, up).
int go_camera_inv_vp(const float cam[3], const float target[3], const float up[3], double fov_deg, float out16[16]) {
    q5::M4 proj = q5::perspective(float(fov_deg), 1.f, 1.f, 100.f);
    q5::M4 view = q5::look_at(V3(target[0], target[1], target[2]), V3(cam[0], cam[1], cam[2]), V3(up[0], up[1], up[2]));
    q5::M4 inv = q5::inverted(q5::mul(proj, view));
    for (int r = 0; r < 4; r++) for (int c = 0; c < 4; c++) out16[r * 4 + c] = inv.m[c][r];
    return GAMER_OK;
}

// GalaxyInstance::setupQuaternions (galaxyinstance.cpp:69-71)
int go_rotation_to_orientation(const float orient[3], float out_wxyz[4]) {
    Quat q = q5::rotation_to(V3(0.f, 1.f, 0.f), V3(orient[0], orient[1], orient[2]));
    out_wxyz[0] = q.w; out_wxyz[1] = q.x; out_wxyz[2] = q.y; out_wxyz[3] = q.z;
    return GAMER_OK;
}

// single-ray probe: direction for pixel (i,j) and its per-galaxy clip, for unit tests
int go_coord2ray(const GamerFrame *frame, int i, int j, float dir_out[3]) {
    V3 d = coord2ray(*frame, i, j, frame->full_size);
    dir_out[0] = d.x; dir_out[1] = d.y; dir_out[2] = d.z;
    return GAMER_OK;
}

}  // extern "C"
