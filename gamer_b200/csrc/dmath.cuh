// dmath.cuh — call-free FP64 elementary functions for the control arithmetic of K1 (sm_100a).
//
// The reference evaluates getWinding (galaxycomponent.h:156-165: exp, atan) and the twirl quaternion
// (galaxycomponent.h:86-90 via QQuaternion::fromAxisAndAngle: sin, cos) with the host libm. Their results are narrowed to
// float right away (the winding angle, the quaternion), so what has to agree is the float, and an FP64 value within
// ~1 ulp gives the same float except when it lands within 2^-52 of a rounding boundary (probability ~2^-29 per call) -
// the same guarantee CUDA's own exp/atan/sincos gave. What CUDA's versions also bring is out-of-line slow paths
// (Payne-Hanek reduction with a local array, division special cases): CALL sites that pin registers and made the
// allocator spill the march state. These are straight-line: Cody-Waite reduction + fdlibm's minimax kernels
// (e_exp.c, s_atan.c, k_sin.c, k_cos.c - Sun Microsystems' freely redistributable coefficients), arguments assumed
// finite and of moderate size (|x| < 700 for exp, |a| < 1e5 rad for sincos), which the callers guarantee.
// Compiles for the host as well (tests/ compare it with libm there).
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDA_ARCH__)
#define DM_FN __device__ __forceinline__
#define DM_FMA(a, b, c) __fma_rn((a), (b), (c))
#define DM_MUL(a, b) __dmul_rn((a), (b))
#define DM_ADD(a, b) __dadd_rn((a), (b))
#define DMF_MUL(a, b) __fmul_rn((a), (b))
#define DMF_ADD(a, b) __fadd_rn((a), (b))
#define DMF_DIV(a, b) __fdiv_rn((a), (b))
#elif defined(__CUDACC__)
#define DM_FN __host__ __device__ inline
#define DM_FMA(a, b, c) fma((a), (b), (c))
#define DM_MUL(a, b) ((a) * (b))
#define DM_ADD(a, b) ((a) + (b))
#else
#define DM_FN inline
#define DM_FMA(a, b, c) fma((a), (b), (c))
#define DM_MUL(a, b) ((a) * (b))
#define DM_ADD(a, b) ((a) + (b))
#endif
#if !defined(__CUDA_ARCH__)   // host: the translation units are built with -ffp-contract=off
#define DMF_MUL(a, b) ((a) * (b))
#define DMF_ADD(a, b) ((a) + (b))
#define DMF_DIV(a, b) ((a) / (b))
#endif

namespace dmath {

// a / b, correctly rounded for normal operands, without the library's slow-path subroutine
DM_FN double div_nc(double a, double b) {
#if defined(__CUDA_ARCH__)
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(b));             // ~2^-20
    double e = __fma_rn(-b, y, 1.0);
    e = __fma_rn(e, e, e);
    y = __fma_rn(y, e, y);                                             // ~2^-40
    e = __fma_rn(-b, y, 1.0);
    y = __fma_rn(y, e, y);                                             // full precision
    const double q = __dmul_rn(a, y);
    return __fma_rn(__fma_rn(-b, q, a), y, q);
#else
    return a / b;
#endif
}
// sqrt(x), correctly rounded, x > 0 and normal
DM_FN double sqrt_nc(double x) {
#if defined(__CUDA_ARCH__)
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));           // ~2^-20
    const double e = __fma_rn(-x, __dmul_rn(y, y), 1.0);               // 1 - x y^2
    y = __fma_rn(__fma_rn(e, 0.375, 0.5), __dmul_rn(y, e), y);         // third-order step: ~2^-58
    const double g = __dmul_rn(x, y), h = __dmul_rn(0.5, y);
    return __fma_rn(__fma_rn(-g, g, x), h, g);
#else
    return sqrt(x);
#endif
}

DM_FN double from_bits(uint64_t u) {
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long)u);
#else
    double d; memcpy(&d, &u, 8); return d;
#endif
}
DM_FN uint64_t to_bits(double d) {
#if defined(__CUDA_ARCH__)
    return (uint64_t)__double_as_longlong(d);
#else
    uint64_t u; memcpy(&u, &d, 8); return u;
#endif
}

// nearest integer of a moderate-size double as (double, int) through the 1.5 * 2^52 shifter
DM_FN double round_int(double v, int &k) {
    const double shifter = 6755399441055744.0;
    const double t = DM_ADD(v, shifter);
    k = (int)(uint32_t)to_bits(t);
    return DM_ADD(t, -shifter);
}

// exp(x), |x| < 700: x = k ln2 + r, |r| <= ln2/2; exp(r) by the degree-13 Taylor polynomial (remainder < 4e-18)
DM_FN double exp_nc(double x) {
    int k;
    const double kd = round_int(DM_MUL(x, 1.4426950408889634074), k);
    double r = DM_FMA(kd, -6.93147180369123816490e-01, x);            // ln2 hi (fdlibm e_exp.c)
    r = DM_FMA(kd, -1.90821492927058770002e-10, r);                   // ln2 lo
    double p = 1.0 / 6227020800.0;
    p = DM_FMA(p, r, 1.0 / 479001600.0);
    p = DM_FMA(p, r, 1.0 / 39916800.0);
    p = DM_FMA(p, r, 1.0 / 3628800.0);
    p = DM_FMA(p, r, 1.0 / 362880.0);
    p = DM_FMA(p, r, 1.0 / 40320.0);
    p = DM_FMA(p, r, 1.0 / 5040.0);
    p = DM_FMA(p, r, 1.0 / 720.0);
    p = DM_FMA(p, r, 1.0 / 120.0);
    p = DM_FMA(p, r, 1.0 / 24.0);
    p = DM_FMA(p, r, 1.0 / 6.0);
    p = DM_FMA(p, r, 0.5);
    p = DM_MUL(p, DM_MUL(r, r));
    p = DM_ADD(DM_ADD(p, r), 1.0);                                     // 1 + r + r^2 (...)
    return from_bits(to_bits(p) + ((uint64_t)(int64_t)k << 52));        // * 2^k (results stay normal for |x| < 700)
}

// atan(x) after fdlibm s_atan.c, branch-free: one division picks the reduced argument for all five intervals
DM_FN double atan_nc(double x) {
    const double ax = fabs(x);
    double num, den, hi, lo;
    if (ax < 0.4375)      { num = ax; den = 1.0; hi = 0.0; lo = 0.0; }
    else if (ax < 0.6875) { num = DM_FMA(2.0, ax, -1.0); den = DM_ADD(2.0, ax); hi = 4.63647609000806093515e-01; lo = 2.26987774529616870924e-17; }
    else if (ax < 1.1875) { num = DM_ADD(ax, -1.0); den = DM_ADD(ax, 1.0); hi = 7.85398163397448278999e-01; lo = 3.06161699786838301793e-17; }
    else if (ax < 2.4375) { num = DM_ADD(ax, -1.5); den = DM_FMA(1.5, ax, 1.0); hi = 9.82793723247329054082e-01; lo = 1.39033110312309984516e-17; }
    else                  { num = -1.0; den = ax; hi = 1.57079632679489655800e+00; lo = 6.12323399573676603587e-17; }
    const double t = (ax < 0.4375) ? ax : div_nc(num, den);
    const double z = DM_MUL(t, t), w = DM_MUL(z, z);
    double s1 = 1.62858201153657823623e-02;                             // aT[10]
    s1 = DM_FMA(s1, w, 4.97687799461593236017e-02);                     // aT[8]
    s1 = DM_FMA(s1, w, 6.66107313738753120669e-02);                     // aT[6]
    s1 = DM_FMA(s1, w, 9.09088713343650656196e-02);                     // aT[4]
    s1 = DM_FMA(s1, w, 1.42857142725034663711e-01);                     // aT[2]
    s1 = DM_FMA(s1, w, 3.33333333333329318027e-01);                     // aT[0]
    s1 = DM_MUL(s1, z);
    double s2 = -3.65315727442169155270e-02;                            // aT[9]
    s2 = DM_FMA(s2, w, -5.83357013379057348645e-02);                    // aT[7]
    s2 = DM_FMA(s2, w, -7.69187620504482999495e-02);                    // aT[5]
    s2 = DM_FMA(s2, w, -1.11111104054623557880e-01);                    // aT[3]
    s2 = DM_FMA(s2, w, -1.99999999998764832476e-01);                    // aT[1]
    s2 = DM_MUL(s2, w);
    const double corr = DM_MUL(t, DM_ADD(s1, s2));
    const double r = DM_ADD(hi, -DM_ADD(DM_ADD(corr, -lo), -t));        // hi - ((t (s1+s2) - lo) - t); hi = lo = 0 below 7/16
    return x < 0.0 ? -r : r;
}

// atanf(x) and atan2f(y, x) exactly as glibc < 2.41 computes them (sysdeps/ieee754/flt-32/s_atanf.c, e_atan2f.c: the
// fdlibm float versions, every operation a separately rounded float; constants as the DECIMAL literals of the source,
// which are what the compiler sees — the hex in its comments is one ulp off for aT[0]). Not correctly rounded: 16 % of
// arguments give the float next to the rounded double atan2. The reference's getTheta (galaxycomponent.h:150-153) calls
// atan2 on two floats, which in C++ is this float overload; with a NEGATIVE arm exponent (arm_value_neg, kernels.cu) a
// sample next to the pole of pow(1 - v/pi, e) turns one ulp of theta into a percent of the pixel, so the libm is part
// of the result there. Finite arguments only (the callers pass components of a finite point).
// tests/test_cpu.py compares both with the host's libm.
DM_FN float atanf_glibc(float x) {
    const float aT0 = 3.3333334327e-01f, aT1 = -2.0000000298e-01f, aT2 = 1.4285714924e-01f, aT3 = -1.1111110449e-01f,
                aT4 = 9.0908870101e-02f, aT5 = -7.6918758452e-02f, aT6 = 6.6610731184e-02f, aT7 = -5.8335702866e-02f,
                aT8 = 4.9768779427e-02f, aT9 = -3.6531571299e-02f, aT10 = 1.6285819933e-02f;
    const bool neg = x < 0.0f;
    const float ax = fabsf(x);
    float hi = 0.0f, lo = 0.0f, t = x;
    bool small = false;
    if (ax >= 33554432.0f) return neg ? -1.5707963267948966f : 1.5707963267948966f;        // |x| >= 2^25: hi[3] + lo[3]
    if (ax < 0.4375f) { if (ax < 0x1p-29f) return x; small = true; }
    else if (ax < 0.6875f) { t = DMF_DIV(DMF_ADD(DMF_MUL(2.0f, ax), -1.0f), DMF_ADD(2.0f, ax)); hi = 4.6364760399e-01f; lo = 5.0121582440e-09f; }
    else if (ax < 1.1875f) { t = DMF_DIV(DMF_ADD(ax, -1.0f), DMF_ADD(ax, 1.0f)); hi = 7.8539812565e-01f; lo = 3.7748947079e-08f; }
    else if (ax < 2.4375f) { t = DMF_DIV(DMF_ADD(ax, -1.5f), DMF_ADD(1.0f, DMF_MUL(1.5f, ax))); hi = 9.8279368877e-01f; lo = 3.4473217170e-08f; }
    else { t = DMF_DIV(-1.0f, ax); hi = 1.5707962513e+00f; lo = 7.5497894159e-08f; }
    const float z = DMF_MUL(t, t), w = DMF_MUL(z, z);
    float s1 = DMF_ADD(aT8, DMF_MUL(w, aT10));
    s1 = DMF_ADD(aT6, DMF_MUL(w, s1)); s1 = DMF_ADD(aT4, DMF_MUL(w, s1));
    s1 = DMF_ADD(aT2, DMF_MUL(w, s1)); s1 = DMF_ADD(aT0, DMF_MUL(w, s1));
    s1 = DMF_MUL(z, s1);
    float s2 = DMF_ADD(aT7, DMF_MUL(w, aT9));
    s2 = DMF_ADD(aT5, DMF_MUL(w, s2)); s2 = DMF_ADD(aT3, DMF_MUL(w, s2)); s2 = DMF_ADD(aT1, DMF_MUL(w, s2));
    s2 = DMF_MUL(w, s2);
    const float ts = DMF_MUL(t, DMF_ADD(s1, s2));
    if (small) return DMF_ADD(t, -ts);
    const float r = DMF_ADD(hi, -DMF_ADD(DMF_ADD(ts, -lo), -t));
    return neg ? -r : r;
}
DM_FN float atan2f_glibc(float y, float x) {
    const float pi_o_2 = 1.5707963705e+00f, pi = 3.1415927410e+00f, pi_lo = -8.7422776573e-08f;
    if (x == 1.0f) return atanf_glibc(y);
    const bool yneg = signbit(y), xneg = signbit(x);
    if (y == 0.0f) return xneg ? (yneg ? -pi : pi) : y;                  // pi + tiny rounds to pi
    if (x == 0.0f) return yneg ? -pi_o_2 : pi_o_2;
    uint32_t hx, hy;
    memcpy(&hx, &x, 4); memcpy(&hy, &y, 4);
    const int k = ((int)(hy & 0x7fffffffu) - (int)(hx & 0x7fffffffu)) >> 23;
    float z;
    if (k > 60) z = DMF_ADD(pi_o_2, DMF_MUL(0.5f, pi_lo));
    else if (xneg && k < -60) z = 0.0f;
    else z = atanf_glibc(fabsf(DMF_DIV(y, x)));
    if (!xneg) return yneg ? -z : z;
    return yneg ? DMF_ADD(DMF_ADD(z, -pi_lo), -pi) : DMF_ADD(pi, -DMF_ADD(z, -pi_lo));
}

// sin and cos of a (|a| < 1e5): a = k pi/2 + r by a three-term Cody-Waite reduction, fdlibm's k_sin / k_cos kernels
DM_FN void sincos_nc(double a, double &s, double &c) {
    int k;
    const double kd = round_int(DM_MUL(a, 6.36619772367581382433e-01), k);
    double r = DM_FMA(kd, -1.57079632673412561417e+00, a);            // pio2_1  (33 bits)
    r = DM_FMA(kd, -6.07710050630396597660e-11, r);                   // pio2_2  (33 bits)
    r = DM_FMA(kd, -2.02226624871116645580e-21, r);                   // pio2_3  (33 bits): residual k * 8e-32
    const double z = DM_MUL(r, r);
    double ps = 1.58969099521155010221e-10;                           // S6
    ps = DM_FMA(ps, z, -2.50507602534068634195e-08);                  // S5
    ps = DM_FMA(ps, z, 2.75573137070700676789e-06);                   // S4
    ps = DM_FMA(ps, z, -1.98412698298579493134e-04);                  // S3
    ps = DM_FMA(ps, z, 8.33333333332248946124e-03);                   // S2
    ps = DM_FMA(ps, z, -1.66666666666666324348e-01);                  // S1
    const double sn = DM_FMA(DM_MUL(z, r), ps, r);                     // r + r^3 (S1 + ...)
    double pc = -1.13596475577881948265e-11;                          // C6
    pc = DM_FMA(pc, z, 2.08757232129817482790e-09);                   // C5
    pc = DM_FMA(pc, z, -2.75573143513906633035e-07);                  // C4
    pc = DM_FMA(pc, z, 2.48015872894767294178e-05);                   // C3
    pc = DM_FMA(pc, z, -1.38888888888741095749e-03);                  // C2
    pc = DM_FMA(pc, z, 4.16666666666666019037e-02);                   // C1
    const double cs = DM_FMA(DM_MUL(z, z), pc, DM_FMA(-0.5, z, 1.0));  // 1 - z/2 + z^2 (C1 + ...)
    const bool swap = (k & 1) != 0;
    const double ss = swap ? cs : sn, cc = swap ? sn : cs;
    s = (k & 2) ? -ss : ss;
    c = ((k + 1) & 2) ? -cc : cc;
}

// sinf(y) and cosf(y) exactly as glibc >= 2.28 computes them (sysdeps/ieee754/flt-32/s_sincosf.h, s_sinf.c, s_cosf.c:
// reduce_fast + sinf_poly, the non-TOINT_INTRINSICS branch that x86-64 builds use) for |y| < 120: a double-precision
// quadrant reduction and two short polynomials, NOT correctly rounded (2.6 % of arguments give the neighbouring float).
// The reference's twirl quaternion is built from these two calls (QQuaternion::fromAxisAndAngle, galaxycomponent.h:86-90)
// and the ridged octaves magnify a 1-ulp change of it past the parity tolerance, so the libm the reference links is
// part of the result; the oracle and the reference build (oracle/_ref) link glibc. Branch-free: glibc's own shortcuts
// for |y| < 0.75 (no reduction) and |y| < 2^-12 (y, 1) return the same floats as the general path (n = 0; the
// polynomial rounds back to y and 1). tests/test_cpu.py checks every float in [0, 16) against the host's libm.
// Returns false for |y| >= 120 (glibc's large-argument path is not restated; callers fall back to sincos_nc).
DM_FN bool sincosf_glibc(float y, float &s, float &c) {
    double x = (double)y;
    const double r = DM_MUL(x, 0x1.45F306DC9C883p+23);                 // x * 2/pi * 2^24
    const int n = ((int)r + 0x800000) >> 24;                           // quadrant, round to nearest
    x = DM_FMA(-(double)n, 0x1.921FB54442D18p0, x);
    const double x2 = DM_MUL(x, x);
    const double xs = (n + 1) & 2 ? -x : x;                            // sign[n & 3] = {1, -1, -1, 1}
    // sine polynomial on xs
    const double x3 = DM_MUL(xs, x2);
    const double s1 = DM_FMA(x2, -0x1.994eb3774cf24p-13, 0x1.1107605230bc4p-7);
    const double x7 = DM_MUL(x3, x2);
    const double sp = DM_FMA(x7, s1, DM_FMA(x3, -0x1.555545995a603p-3, xs));
    // cosine polynomial (negated as a whole in quadrants 2 and 3: glibc's second table holds the negated coefficients)
    const double c2 = DM_FMA(x2, 0x1.99343027bf8c3p-16, -0x1.6c087e89a359dp-10);
    const double c1 = DM_FMA(x2, 0x1.55553e1068f19p-5, -0x1.ffffffd0c621cp-2);
    const double x6 = DM_MUL(DM_MUL(x2, x2), x2);
    double cp = DM_FMA(x6, c2, DM_FMA(x2, c1, 1.0));
    if (n & 2) cp = -cp;
    const float fs = (float)sp, fc = (float)cp;
    s = (n & 1) ? fc : fs;
    c = (n & 1) ? fs : fc;
    return fabsf(y) < 120.0f;
}

// atan2f(y, x) in float to ~1e-7 absolute (value arithmetic: the arm phase): quotient by reciprocal, one more
// reciprocal for the (a-1)/(a+1) reduction above tan(pi/8), degree-9 odd minimax polynomial (error < 1e-9)
DM_FN float atan2f_fast(float y, float x) {
    const float ax = fabsf(x), ay = fabsf(y);
    const float mx = fmaxf(ax, ay), mn = fminf(ax, ay);
#if defined(__CUDA_ARCH__)
    float r0, r1;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r0) : "f"(mx));
    float a = mn * r0;
    if (!(mx > 0.0f)) a = 0.0f;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r1) : "f"(a + 1.0f));
    const bool red = a > 0.41421356f;
    const float t = red ? (a - 1.0f) * r1 : a;
#else
    float a = mx > 0.0f ? mn / mx : 0.0f;
    const bool red = a > 0.41421356f;
    const float t = red ? (a - 1.0f) / (a + 1.0f) : a;
#endif
    const float u = t * t;
    float p = -0.06451928160230586f;
    p = fmaf(p, u, 0.1074373147094983f);
    p = fmaf(p, u, -0.14263955595196748f);
    p = fmaf(p, u, 0.19999540483500192f);
    p = fmaf(p, u, -0.33333331761166285f);
    float r = fmaf(t * u, p, t);
    if (red) r += 0.78539816339744831f;
    if (ay > ax) r = 1.57079632679489662f - r;
    if (x < 0.0f) r = 3.14159265358979324f - r;
    return y < 0.0f ? -r : r;
}

}  // namespace dmath
