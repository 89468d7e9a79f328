// kernels.cu — K1 (ray march), K2 (post-process) and the noise probes, plus their launchers.
// sm_100a only.  See march.cuh / noise.cuh for the design notes and the reference citations.
#include <mutex>
#include "march.cuh"
#include "tone.cuh"
#include "kernels.h"

using namespace gmarch;

namespace {

// ONE persistent CTA of K1_WARPS warps per SM (80 registers per thread); each warp draws 8 x 4 pixel tiles from a
// ticket counter. The noise tables (2 x 64 KB) and the per-thread ray records (84 KB) then exist once per SM.
// Measured: 24 x 1 = 12 x 2 warps; 10 x 2 (96 registers) and 8 x 3 are 2-3 % slower; 28 x 1 (72 registers, 48 B of spill
// stores) is 1.0 % faster on the exterior view and 0.8 % slower on the in-disk one.
#ifndef K1_WARPS
#define K1_WARPS 24
#endif
#define K1_THREADS (32 * K1_WARPS)
#ifndef K1_MINBLOCKS
#define K1_MINBLOCKS 1
#endif

struct NoiseCtx {
    uint32_t lane4;
};

// The per-thread ray record of march(): K1_REC_WORDS floats per thread in shared memory behind the noise tables,
// word k of thread t at [k][t] (conflict free). volatile: every access is a real load or store.
// Words: 0-2 p, 3-5 dir, 6-8 origin, 9 length, 10 loop bound (-inf: no ray), 11 step, 12 weight, 13 distance marched
// (+inf: no ray), 14 / 15 the fast-sample windows (marched < [14]: the step is certainly 0.01 and the ray alive;
// marched > [15]: certainly minRayStep), 16 whether the winding in 24-25 came from the table (1) or from the direct
// evaluation (0), 17 twirl group of the point in 18-20 (-1: none yet). The camera relative to the galaxy is NOT here: it
// is the same for every ray and is formed from the constant bank where it is needed (three subtractions, the same
// three the record's loads were). 28 words: 24 or 28 warps fit beside the noise tables.
#define REC_LEN 9
#define REC_THR 10
#define REC_STEP 11
#define REC_WEIGHT 12
#define REC_MARCHED 13
#define REC_HI_LIM 14
#define REC_LO_FROM 15
#define REC_WTAB 16
#define REC_TG 17
// values of ONE in-slab sample shared by every component of the galaxy. They live here and not in registers because
// they are needed on both sides of the noise calls: what ptxas cannot keep in a register across a call it keeps in
// LOCAL memory for its whole life, and with 768 threads a 96-byte stack is 72 KB per SM against 28 KB of L1 (ncu: 61 %
// local hit rate, long-scoreboard stalls on every reload). A shared-memory load costs one issue slot and ~25 cycles.
#define REC_DOTT 21      // height above the plane along the axis (exact)
#define REC_PLEN 22      // |P| fast float
#define REC_WT 23        // float(getWinding(radius)) + atan2(q.x, q.z), q = rotmat * P: the arm phase without delta and
                         // the arm's own offset (value arithmetic); NaN: not formed yet for this sample
#define REC_WD 24        // getWinding(radius) in double (it feeds the twirl angle): words 24, 25
#define REC_IVAL 26      // val * weight of the component being evaluated (live across its noise call)
#define REC_K 27         // index of the component being evaluated (the loop counter, as float bits)
#define K1_REC_WORDS 28
static_assert(GN_DYN_SMEM + K1_REC_WORDS * K1_THREADS * 4u <= 232448u, "tables + ray records exceed the 227 KB a CTA can have");
struct Rec {
    volatile float *base;
    __device__ __forceinline__ volatile float &operator[](int k) const { return base[k * K1_THREADS]; }
};
#define K1_DYN_SMEM (GN_DYN_SMEM + K1_REC_WORDS * K1_THREADS * 4u)

__device__ __forceinline__ double rec_get_d(const Rec rec, int k) {
    return __hiloint2double(__float_as_int(rec[k + 1]), __float_as_int(rec[k]));
}
__device__ __forceinline__ void rec_set_d(const Rec rec, int k, double v) {
    rec[k] = __int_as_float(__double2loint(v)); rec[k + 1] = __int_as_float(__double2hiint(v));
}

// getWinding (galaxycomponent.h:156-165), the reference's double expression atan(exp(-0.25/(0.5*r))/windingB)*2*windingN,
// evaluated directly: a serial FP64 chain of two divisions, an exp and an atan. Out of line: taken only when there is no
// table, or for the one sample in ~10^5 whose twirl angle the table cannot decide (eval_component).
__device__ __noinline__ void winding_direct(const DGalaxy &g, float radius, float *rec_base) {
    const Rec rec = {rec_base};
    const double r = __dadd_rn((double)radius, 0.05);
    const double e = dmath::exp_nc(dmath::div_nc(-0.25, __dmul_rn(0.5, r)));
    const double Wd = __dmul_rn(__dmul_rn(dmath::atan_nc(dmath::div_nc(e, g.winding_b)), 2.0), g.winding_n);
    rec_set_d(rec, REC_WD, Wd);
    rec[REC_WTAB] = 0.0f;
}
// radius exactly as getRadius returns it: float |P| (double sqrt) / float axis.x
__device__ __forceinline__ float exact_radius(const DGalaxy &g, f3 P) {
    return fdiv_nc(v_length(P), g.axis_x, frcp_nc(g.axis_x));
}
// getWinding and the atan2 of getTheta (:150-153), once per sample. With a table (scene_build.cpp, winding_table):
// s = 1 / (radius + 0.05) by a Newton reciprocal, the interval of s, one degree-9 Horner chain on coefficients read
// through the read-only path — certified on the host to 2^-46 of the reference's own expression.
__device__ __forceinline__ void arm_base(const DScene &sc, const DGalaxy &g, const Rec rec, f3 P) {
    const float radius = exact_radius(g, P);
    bool done = false;
    if (g.wt_n > 0) {
        const double r = __dadd_rn((double)radius, 0.05);
        double y;
        asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(r));
        double e = __fma_rn(-r, y, 1.0);
        e = __fma_rn(e, e, e);
        y = __fma_rn(y, e, y);
        y = __fma_rn(__fma_rn(-r, y, 1.0), y, y);                        // s = 1 / r to an ulp
        const double u = __dmul_rn(__dsub_rn(y, g.wt_s0), 1.0 / WT_H);
        const int i = __double2int_rz(u);
        if (u >= 0.0 && i < g.wt_n) {
            const double x = __fma_rn(2.0, __dsub_rn(u, (double)i), -1.0);
            const double2 *c = reinterpret_cast<const double2 *>(sc.wtab + g.wt_off + i * WT_COEFFS);
            const double2 c01 = __ldg(c), c23 = __ldg(c + 1), c45 = __ldg(c + 2), c67 = __ldg(c + 3), c89 = __ldg(c + 4);
            double v = __fma_rn(c89.y, x, c89.x);
            v = __fma_rn(v, x, c67.y); v = __fma_rn(v, x, c67.x);
            v = __fma_rn(v, x, c45.y); v = __fma_rn(v, x, c45.x);
            v = __fma_rn(v, x, c23.y); v = __fma_rn(v, x, c23.x);
            v = __fma_rn(v, x, c01.y); v = __fma_rn(v, x, c01.x);
            rec_set_d(rec, REC_WD, v);
            rec[REC_WTAB] = 1.0f;
            done = true;
        }
    }
    if (!done) winding_direct(g, radius, const_cast<float *>(rec.base));
    f3 q = g.rot_identity ? P : q_rotate(g.rot, P);
    rec[REC_WT] = (float)rec_get_d(rec, REC_WD) + dmath::atan2f_fast(q.x, q.z);
}

// findDifference (galaxycomponent.h:103-118) with M_PI = 3.14159265 (util.h:20)
__device__ __forceinline__ float find_difference(float d) {
    const float TP = 6.2831853f, FP = 12.5663706f;
    float v = fminf(fabsf(d), fabsf(d - TP));
    v = fminf(v, fabsf(d + TP));
    v = fminf(v, fabsf(d - FP));
    v = fminf(v, fabsf(d + FP));
    return v;
}

// calculateArmValue (galaxycomponent.h:120-146): max over the arms of pow(1 - v_j / pi, e), where an arm whose base is
// negative (wrapped phase difference beyond 3.14159265) gives NaN and is dropped by the reference's fmax; with every arm
// out of range, or a single one (no fmax then), the NaN reaches `val` and fails the `val * weight > 0.0005` test.
// e > 0: the maximum is pow(1 - min_j v_j / pi, e), and a base clamped to 0 gives 0, which also contributes nothing.
// e < 0 (no shipped galaxy; out of line): the maximum is at the LARGEST in-range v_j; none in range -> 0. Near
// v = pi the base goes to zero and the negative exponent amplifies any error of the phase, which the reference forms in
// double (getWinding, theta + delta, findDifference: galaxycomponent.h:100-153) — except atan2 itself, which on two
// floats is the libm's FLOAT atan2f. This path does the same: the winding in double from the record, glibc's atan2f
// restated bit for bit (dmath::atan2f_glibc), the differences in double with M_PI = 3.14159265; only the final pow of
// the (well-conditioned) base is float.
__device__ __noinline__ float arm_value_neg(const DGalaxy &g, float arm15, double delta_d, float *rec_base, bool orient_y) {
    const Rec rec = {rec_base};
    const f3 pa = {rec[0], rec[1], rec[2]};
    const f3 orient = {g.orient[0], g.orient[1], g.orient[2]};
    const f3 P = orient_y ? f3{pa.x, 0.0f, pa.z} : v_sub(pa, v_mul(orient, rec[REC_DOTT]));
    // rotmat * P exactly as QQuaternion rotates a vector (two float quaternion products): even the identity quaternion
    // perturbs z by an ulp there ((z + y) - y), and a sample within 1e-6 of the pole turns that into a percent
    const fq rq = {g.rot[0], g.rot[1], g.rot[2], g.rot[3]};
    const fq qv = q_mul_qt(q_mul_qt(rq, fq{0.0f, P.x, P.y, P.z}), fq{rq.w, -rq.x, -rq.y, -rq.z});
    const double base = __dadd_rn(__dadd_rn(rec_get_d(rec, REC_WD), (double)dmath::atan2f_glibc(qv.x, qv.z)), delta_d);
    const double PI_REF = 3.14159265, TP = 2 * 3.14159265, FP = 2 * 3.14159265 * 2;
    double best = -1.0;                 // the smallest non-negative base 1 - v / pi over the arms
    for (int j = 0; j < g.no_arms; j++) {
        const double d = __dsub_rn(base, g.arm_d[j]);
        double v = fmin(fabs(d), fabs(d - TP));
        v = fmin(v, fabs(d + TP));
        v = fmin(v, fabs(d - FP));
        v = fmin(v, fabs(d + FP));
        const double bj = 1.0 - dmath::div_nc(v, PI_REF);
        if (bj >= 0.0 && (best < 0.0 || bj < best)) best = bj;
    }
    if (best < 0.0) return 0.0f;
    return ex2a(arm15 * lg2a((float)best));
}
__device__ __forceinline__ float arm_value(const DGalaxy &g, float arm15, float base) {
    float v = find_difference(base - g.arm[0]);
    if (g.no_arms > 1) v = fminf(v, find_difference(base - g.arm[1]));
    if (g.no_arms > 2) v = fminf(v, find_difference(base - g.arm[2]));
    if (g.no_arms > 3) v = fminf(v, find_difference(base - g.arm[3]));
    return ex2a(arm15 * lg2a(fmaxf(fmaf(v, -1.0f / 3.14159265f, 1.0f), 0.0f)));
}

// Bulge emission (galaxycomponents.cpp:5-18): rad = (|p|+0.01) r0 + 0.01; i = k (rad^-0.855 exp(-rad^0.25) - 0.05), i >= 0
__device__ __forceinline__ void bulge_add(float k, float r0, float sr, float sg, float sb, float plen3,
                                          float &Ir, float &Ig, float &Ib) {
    float rad = fmaf(plen3 + 0.01f, r0, 0.01f);
    float L = lg2a(rad);
    float v = ex2a(fmaf(-1.4426950408889634f, ex2a(0.25f * L), -0.855f * L)) - 0.05f;
    float i = fmaxf(k * v, 0.0f);
    Ir = fmaf(i, sr, Ir); Ig = fmaf(i, sg, Ig); Ib = fmaf(i, sb, Ib);
}

// the scalar of bulge_add alone: max(k (rad^-0.855 exp(-rad^0.25) - 0.05), 0)
__device__ __forceinline__ float bulge_term(float k, float r0, float plen3) {
    float rad = fmaf(plen3 + 0.01f, r0, 0.01f);
    float L = lg2a(rad);
    float v = ex2a(fmaf(-1.4426950408889634f, ex2a(0.25f * L), -0.855f * L)) - 0.05f;
    return fmaxf(k * v, 0.0f);
}

// Stars with noiseOffset != 0 (galaxycomponents.cpp:144-148): two 4-octave helpers, persistence -2
// (amplitudes 1,-2,4,-8, sum -5). No shipped galaxy uses it: kept out of line so the hot loop stays small.
__device__ __noinline__ float stars_helper_octaves(const DGalaxy &g, float noise_offset, float px, float py, float pz,
                                                   uint32_t lane4, double winding) {
    const f3 p = {px, py, pz};          // by value: a reference would pin the caller's per-sample state in local memory
    f3 r1 = twirl_qt(g.axis_n, p, winding);
    f3 r2 = twirl_qt(g.axis_n, p, __dmul_rn(winding, 0.5));
    float add = noise_offset * gnoise::fbm(lane4, r1.x, r1.y, r1.z, 2.0 * (double)0.1f, 4, -2.0f, -0.2f);
    add += 0.5f * (noise_offset * gnoise::fbm(lane4, r2.x, r2.y, r2.z, 4.0 * (double)0.1f, 4, -2.0f, -0.2f));
    return add;
}

// twirl of the sample point (galaxycomponent.h:86-90). ORIENT_Y: the axis is exactly (0,1,0) — twirl_y of qtf.cuh,
// bit-identical to the general sequence, which stays out of line for the (never observed) case that the quaternion's
// length is outside the two values twirl_y knows.
__device__ __noinline__ void twirl_general(const float *axis_n, float px, float py, float pz, double winding, float *out) {
    const f3 r = twirl_qt(axis_n, {px, py, pz}, winding);
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
template <bool ORIENT_Y>
__device__ __forceinline__ f3 twirl_point(const DGalaxy &g, f3 p, double winding) {
    if (ORIENT_Y) {
        bool ok;
        f3 r = twirl_y(p, winding, &ok);
        if (!ok) { float o[3]; twirl_general(g.axis_n, p.x, p.y, p.z, winding, o); r = {o[0], o[1], o[2]}; }
        return r;
    }
    return twirl_qt(g.axis_n, p, winding);
}

template <bool COUNT, bool ORIENT_Y>
__device__ __forceinline__ void eval_component(const DScene &sc, const DGalaxy &g, const DComp &c, const Rec rec,
                                               const NoiseCtx &nz, float &Ir, float &Ig, float &Ib, Counts &cn) {
    const float rs = sc.ray_step_f;
    const int cls = c.cls;
    if (cls == GAMER_CLASS_BULGE) {   // galaxycomponents.cpp:5-25 — evaluated at every sample, no gates
        const f3 pb = {rec[0], rec[1], rec[2]};
        const float i = bulge_term(c.bk * rec[REC_WEIGHT], c.r0, sqrta(fmaf(pb.z, pb.z, fmaf(pb.y, pb.y, pb.x * pb.x))));
        Ir = fmaf(i, c.spec[0], Ir); Ig = fmaf(i, c.spec[1], Ig); Ib = fmaf(i, c.spec[2], Ib);
        if (COUNT) cn.evals++;
        return;
    }
    // ---- slab gate: h = |dott/z0| > 2 -> 0 (galaxycomponent.h:61-64), exact --------------------
    const float ad = fabsf(rec[REC_DOTT]);
    if (!(ad <= c.h_gate)) return;
    // ---- radial gate: intensity > 0.001 (galaxycomponent.cpp:62-65), exact -----------------------
    const float Plen = rec[REC_PLEN];
    bool pass;
    if (Plen < c.plen_lo) pass = true;
    else if (Plen > c.plen_hi) pass = false;
    else {      // rare: inside the band where only the exact |P| decides; P = p - orient * dott like getRadius
        const f3 orient = {g.orient[0], g.orient[1], g.orient[2]};
        const f3 pq = {rec[0], rec[1], rec[2]};
        const f3 Pq = ORIENT_Y ? f3{pq.x, 0.0f, pq.z} : v_sub(pq, v_mul(orient, rec[REC_DOTT]));
        pass = v_length(Pq) <= c.plen_thr;
    }
    if (!pass) return;
    // ---- val = strength * scale * armVal * z * intensity (galaxycomponent.cpp:58-79), value arithmetic: raw MUFU ------
    // z = sech^2(h) = 4 E / (1 + E)^2 with E = exp(-2h): one exponential, one reciprocal. h <= 2 behind the slab gate,
    // so z >= 0.07 and the reference's `z > 0.01` test (galaxycomponent.cpp:55) cannot fail here.
    const int flags = c.flags;
    const float E = ex2a(-ad * c.hz2), onepE = 1.0f + E;
    float val = c.strength4 * E * rcpa(onepE * onepE);
    val *= fminf(fmaxf(ex2a(Plen * c.k_radial) - 0.01f, 0.0f), 0.1f);
    if (!(flags & DC_INNER_ZERO)) {
        const float x = __saturatef(Plen * c.inner_k);
        float sm = x * x * fmaf(-2.0f, x, 3.0f);
        sm *= sm;
        val *= sm * sm;
    }
    double winding = 0.0;
    if (flags & DC_HAS_ARM) {
        // getWinding / getTheta of the sample, formed by the first component that gets this far (a NaN marks "not yet"):
        // a fifth of the in-slab samples of an exterior view lie beyond every armed component's radial cut-off
        if (rec[REC_WT] != rec[REC_WT]) {
            const f3 pa = {rec[0], rec[1], rec[2]};
            const f3 orient = {g.orient[0], g.orient[1], g.orient[2]};
            arm_base(sc, g, rec, ORIENT_Y ? f3{pa.x, 0.0f, pa.z} : v_sub(pa, v_mul(orient, rec[REC_DOTT])));
        }
        // winding - (-theta + disp): W + theta + delta - disp
        if (c.arm15 > 0.0f) val *= arm_value(g, c.arm15, rec[REC_WT] + c.delta);
        else val *= arm_value_neg(g, c.arm15, c.delta_d, const_cast<float *>(rec.base), ORIENT_Y);
        if (flags & DC_HAS_WINDING) {
            winding = __dmul_rn(rec_get_d(rec, REC_WD), c.winding_d);
            if (rec[REC_WTAB] != 0.0f) {
                // What the reference consumes is float(winding * 180), the twirl angle. The table is within 2^-46 of the
                // direct evaluation (certified, scene_build.cpp); if 2^-44 either way rounds to the same float the table
                // decides, otherwise (one sample in ~10^5) getWinding is evaluated directly for this sample.
                const double a = __dmul_rn(winding, 180.0), del = __dmul_rn(fabs(a), 0x1p-44);
                if (__double2float_rn(__dsub_rn(a, del)) != __double2float_rn(__dadd_rn(a, del))) {
                    const f3 pa = {rec[0], rec[1], rec[2]};
                    const f3 orient = {g.orient[0], g.orient[1], g.orient[2]};
                    winding_direct(g, exact_radius(g, ORIENT_Y ? f3{pa.x, 0.0f, pa.z} : v_sub(pa, v_mul(orient, rec[REC_DOTT]))),
                                   const_cast<float *>(rec.base));
                    winding = __dmul_rn(rec_get_d(rec, REC_WD), c.winding_d);
                }
            }
        }
    }
    {
        const float ival0 = val * rec[REC_WEIGHT];
        if (!(ival0 > 0.0005f)) return;
        rec[REC_IVAL] = ival0;
    }
    if (COUNT) cn.evals++;

    // One twirl and ONE noise evaluation site for every class (keeps the hot loop small enough for the
    // instruction cache); the class-specific use of the noise value follows.
    if (cls > GAMER_CLASS_STARS) return;              // "stars small": no-op (see gamer_b200.h)
    const bool is_ridged = (cls == GAMER_CLASS_DUST2) | (cls == GAMER_CLASS_DUST_POSITIVE);
    const f3 sp = {rec[0], rec[1], rec[2]};
    f3 r = sp;
#ifdef K1_NO_TWIRL_CACHE
    if (cls != GAMER_CLASS_STARS) r = twirl_point<ORIENT_Y>(g, sp, winding);
#else
    // components with the same winding factor twirl the sample to the same point (DComp::tgroup, scene_build.cpp):
    // the point of the last twirl of this sample sits in the ray record, words 17 (group) and 18-20
    if (cls != GAMER_CLASS_STARS) {
        // decided for the lanes that are here together: a lane-wise branch would send the two sides through the
        // noise call below one after the other (a lane that has the point recomputes the same bits)
        const float tg = (float)c.tgroup;
        const bool have = c.tgroup >= 0 && rec[REC_TG] == tg;
        if (__all_sync(__activemask(), have)) r = {rec[REC_TG + 1], rec[REC_TG + 2], rec[REC_TG + 3]};
        else {
            r = twirl_point<ORIENT_Y>(g, sp, winding);
            if (c.tgroup >= 0) { rec[REC_TG] = tg; rec[REC_TG + 1] = r.x; rec[REC_TG + 2] = r.y; rec[REC_TG + 3] = r.z; }
        }
    }
#endif
    float n;
#ifdef K1_DIAG_NONOISE                  // timing experiment only (results are wrong): everything but the noise octaves
    n = 0.37f + 1e-3f * r.x;
    if (COUNT) cn.octaves += is_ridged ? DEV_RIDGED_OCTAVES : c.octaves;
#else
    if (is_ridged) {                    // getRidgedMf(r * scale, ks, 9, 2.5, noiseOffset, noiseTilt)
        n = gnoise::ridged(nz.lane4, __fmul_rn(r.x, c.rscale), __fmul_rn(r.y, c.rscale), __fmul_rn(r.z, c.rscale),
                           DEV_RIDGED_OCTAVES, 2.5f, c.noise_offset, c.noise_tilt, c.rw[0], c.rwq);
        if (COUNT) cn.octaves += DEV_RIDGED_OCTAVES;
    } else {                            // getPerlinCloudNoise / octave_noise_3d
        n = gnoise::fbm(nz.lane4, r.x, r.y, r.z, c.nscale, c.octaves, c.pers, c.inv_maxamp);
        if (COUNT) cn.octaves += c.octaves;
    }
#endif

    float k;                            // emission coefficient (k >= 0 classes) or absorption exponent
    bool absorb = false;
    const float ival = rec[REC_IVAL];
    switch (cls) {
    case GAMER_CLASS_DISK: {            // galaxycomponents.cpp:41-59
        float p2 = fpow(fmaxf(fabsf(n), 0.01f), c.noise_tilt) + c.noise_offset;
        if (p2 < 0.0f) return;
        k = ival * p2 * rs;
        break;
    }
    case GAMER_CLASS_DUST: {            // galaxycomponents.cpp:64-84
        float p2 = fmaxf(n - c.noise_offset, 0.0f);
        p2 = (p2 > 0.0f) ? fpow(5.0f * p2, c.noise_tilt) : (c.noise_tilt > 0.0f ? 0.0f : (c.noise_tilt == 0.0f ? 1.0f : 10.0f));
        p2 = fminf(fmaxf(p2, -10.0f), 10.0f);
        k = -p2 * ival * 0.01f;
        absorb = true;
        break;
    }
    case GAMER_CLASS_DUST2: {           // galaxycomponents.cpp:85-114
        k = -fmaxf(n, 0.0f) * ival * 0.01f;
        absorb = true;
        break;
    }
    case GAMER_CLASS_DUST_POSITIVE: {   // galaxycomponents.cpp:116-136
        k = ival * fmaxf(n, 0.0f) * rs;
        break;
    }
    default: {                          // GAMER_CLASS_STARS, galaxycomponents.cpp:139-155
        float add = 0.0f;
        if (c.flags & DC_HAS_NOFF) {
            // (the winding is formed again here rather than kept across the noise call above)
            double w2 = 0.0;
            if ((c.flags & DC_HAS_ARM) && (c.flags & DC_HAS_WINDING)) w2 = __dmul_rn(rec_get_d(rec, REC_WD), c.winding_d);
            add = stars_helper_octaves(g, c.noise_offset, rec[0], rec[1], rec[2], nz.lane4, w2);
            if (COUNT) cn.octaves += 8;
        }
        float base = fabsf(n) + 1.0f + add;
        float val2;
        if (base >= 0.0f) val2 = fpow(base, c.noise_tilt);
        else if (c.flags & DC_TILT_INT) val2 = fpow(-base, c.noise_tilt);        // |pow(neg, int)|
        else { Ir = Ig = Ib = 0.0f; return; }   // pow -> NaN poisons I; Util::floor's fmax(0,NaN) then zeroes it
        k = ival * val2 * rs;
        break;
    }
    }
    if (absorb) { Ir *= fexp(k * c.spec[0]); Ig *= fexp(k * c.spec[1]); Ib *= fexp(k * c.spec[2]); }
    else { Ir = fmaf(k, c.spec[0], Ir); Ig = fmaf(k, c.spec[1], Ig); Ib = fmaf(k, c.spec[2], Ib); }
}

// Rasterizer::getIntensity (rasterizer.cpp:422-483) for one galaxy, entered by ALL 32 lanes of the warp (`active` is
// false for a lane outside the image or whose ray misses the ellipsoid: its ray ends before the first sample).
// Control arithmetic is exact (see march.cuh). The loop test `dot(p-origin, ll) < length + step` compares a float with
// a double; for a float d and double L, d < L  <=>  d < round_up_to_float(L), so the bound is kept as a float that is
// refreshed only when the step length changes.
template <bool COUNT, bool ORIENT_Y>
__device__ __forceinline__ void march(const DScene &sc, const DGalaxy &g, const Rec rec, bool active, f3 isp1, f3 isp2,
                                      f3 camrel, const NoiseCtx &nz, float &Ir, float &Ig, float &Ib, Counts &cn) {
    // The march has two kinds of samples with very different needs: off-slab samples (nine in ten in an exterior view)
    // take ~60 instructions and want the ray geometry in registers; an in-slab sample runs the components and their noise
    // octaves - thousands of instructions that need every register a thread has (80 at 24 warps per SM). ptxas does not split
    // live ranges: a value that is live across the in-slab work and does not fit is kept in local memory for its whole
    // life and reloaded in the per-sample loop. So the split is made by hand: the ray geometry lives in a small
    // per-thread record in shared memory (`rec`); the off-slab run loop loads it when it starts, the in-slab section
    // reads back the few values it needs, and nothing but the accumulated intensity is carried in registers between
    // the two.
    {
        const f3 origin = isp1;
        const f3 dseg = v_sub(isp1, isp2);
        const float length_f = v_length(dseg);      // the reference's `length` is this float, widened
        const f3 dir = v_normalized(dseg);
        rec[0] = origin.x; rec[1] = origin.y; rec[2] = origin.z;      // p starts at origin
        rec[3] = dir.x; rec[4] = dir.y; rec[5] = dir.z;
        rec[6] = origin.x; rec[7] = origin.y; rec[8] = origin.z;
        rec[REC_LEN] = length_f;
        rec[REC_MARCHED] = active ? 0.0f : INFINITY;            // distance marched so far (estimate, see the run loop)
        // the first test uses step = rayStep; a ray that is not marched has no bound at all
        rec[REC_THR] = active ? __double2float_ru(__dadd_rn((double)length_f, sc.ray_step)) : -INFINITY;
        // Every ray starts at the camera, so p - camera is parallel to dir and |p - camera| = dist0 - (distance marched)
        // up to rounding (bounded by sc.alive_margin like the loop test): it falls monotonically along the march. The
        // step clamp(|p - camera| * rayStep, minRayStep, 0.01) is therefore CERTAINLY 0.01 while marched < dist0 - d_hi -
        // margin and certainly minRayStep once marched > dist0 - d_lo + margin: samples inside those windows need
        // neither the distance nor the exact loop test (the fast sample of the run loop).
        const f3 oc = v_sub(origin, camrel);
        const float dist0 = sqrta(fmaf(oc.z, oc.z, fmaf(oc.y, oc.y, oc.x * oc.x)));
        const float sure_below = length_f - sc.alive_margin;
        rec[REC_HI_LIM] = (active && g.fast_ok) ? fminf(dist0 - sc.d_hi - sc.alive_margin, sure_below) : -INFINITY;
        rec[REC_LO_FROM] = (active && g.fast_ok) ? dist0 - sc.d_lo + sc.alive_margin : INFINITY;
    }
    if (COUNT && active) cn.rays++;
    const bool fast_ok = g.fast_ok != 0;
    // The 32 rays stay in lockstep, one sample per iteration: the run loop is left as soon as ANY of them reaches an
    // in-slab sample (a lane that ran ahead alone would execute its off-slab samples at a fraction of the warp's width),
    // and a lane whose ray has ended - its loop test stays false, p no longer moves - idles until the last one is done.
    const unsigned FULL = 0xffffffffu;
    for (;;) {
        bool need = false, any_alive;
        {   // ---- run of off-slab samples ---------------------------------------------------------------------
            f3 p = {rec[0], rec[1], rec[2]};
            const f3 dir = {rec[3], rec[4], rec[5]};
            const f3 camrel_ = v_sub({sc.cam[0], sc.cam[1], sc.cam[2]}, {g.pos[0], g.pos[1], g.pos[2]});   // = camrel, bit for bit
            const float length_f = rec[REC_LEN];
            float thr = rec[REC_THR];
            // The exact loop test costs ten instructions and decides only at the very end of a ray: d = -dot(p - origin,
            // dir) differs from the running sum of the steps by rounding alone (bounded by sc.alive_margin, scene_build.cpp),
            // so while that sum stays below length - margin the ray is certainly alive (the bound is length + step).
            float marched = rec[REC_MARCHED];
            const float sure_below = length_f - sc.alive_margin;
            float stepf = 0.f, weight = 0.f;
            // ll = normalized(isp2 - origin) of the reference is -dir bit for bit (negation commutes with every
            // rounding involved), so the loop test dot(p - origin, ll) < thr is evaluated as -dot(p - origin, dir) < thr.
            const f3 orient = {g.orient[0], g.orient[1], g.orient[2]};
            // the two clamped step states (rasterizer.cpp:449): step = 0.01 and step = minRayStep
            const float thr_hi = __double2float_ru(__dadd_rn((double)length_f, 0.01));
            const float thr_lo = __double2float_ru(__dadd_rn((double)length_f, sc.min_ray_step));
            const float hi_lim = rec[REC_HI_LIM], lo_from = rec[REC_LO_FROM];
            const float s_hi = __double2float_rn(0.01), s_lo = __double2float_rn(sc.min_ray_step);
            const float w_hi = (float)(0.01 * 200.0), w_lo = (float)(sc.min_ray_step * 200.0);
            // bulges of this galaxy in registers
            float bk0 = 0.f, br0 = 0.f, bs0r = 0.f, bs0g = 0.f, bs0b = 0.f, bk1 = 0.f, br1 = 0.f, bs1r = 0.f, bs1g = 0.f, bs1b = 0.f;
            if (g.n_bulge > 0) {
                const DComp &c = sc.comp[g.bulge_idx[0]];
                bk0 = c.strength * c.iscale * sc.ray_step_f; br0 = c.r0; bs0r = c.spec[0]; bs0g = c.spec[1]; bs0b = c.spec[2];
            }
            if (g.n_bulge > 1) {
                const DComp &c = sc.comp[g.bulge_idx[1]];
                bk1 = c.strength * c.iscale * sc.ray_step_f; br1 = c.r0; bs1r = c.spec[0]; bs1g = c.spec[1]; bs1b = c.spec[2];
            }
            // Fast samples. A lane may take them while marched < lim: the ray is then certainly alive and its step
            // certainly the clamped one of the lane's last general sample (windows of the ray record), so ds = dir * step
            // and the bulge coefficients kw = k * weight of that sample still hold. The warp takes a fast sample only when
            // EVERY lane that still has a ray can, and is off the slab: one vote, no distance, no loop test, no step logic.
            float lim = -INFINITY, kw0 = 0.f, kw1 = 0.f;
            f3 ds = {0.f, 0.f, 0.f};
            unsigned idle = 0u;                     // lanes whose ray has ended (or never began)
            for (;;) {
#ifndef K1_NO_FAST_RUN
                {   // two fast samples per trip: the bulge term is a serial chain of four MUFU latencies (sqrt, lg2, ex2,
                    // ex2); the chains of two consecutive samples are independent and overlap, and one vote serves both
                    const f3 p1 = v_sub(p, ds);
                    const float m1 = marched + stepf;
                    const float dott0 = ORIENT_Y ? p.y : v_dot(p, orient), dott1 = ORIENT_Y ? p1.y : v_dot(p1, orient);
#ifdef K1_DIAG_NOSLAB
                    const bool go = m1 < lim && dott0 == dott0 && dott1 == dott1;
#else
                    const bool go = m1 < lim && !(fabsf(dott0) <= g.slab_gate) && !(fabsf(dott1) <= g.slab_gate);
#endif
                    const unsigned b = __ballot_sync(FULL, go);
                    if (b != 0u && (b | idle) == FULL) {
                        if (go) {
                            const float q0 = sqrta(fmaf(p.z, p.z, fmaf(p.y, p.y, p.x * p.x)));
                            const float q1 = sqrta(fmaf(p1.z, p1.z, fmaf(p1.y, p1.y, p1.x * p1.x)));
                            const float i00 = bulge_term(kw0, br0, q0), i01 = bulge_term(kw0, br0, q1);
                            float i10 = 0.f, i11 = 0.f;
                            if (g.n_bulge > 1) { i10 = bulge_term(kw1, br1, q0); i11 = bulge_term(kw1, br1, q1); }
                            // accumulation in the reference's order: sample 0 (bulge 0, bulge 1), then sample 1
                            Ir = fmaf(i00, bs0r, Ir); Ig = fmaf(i00, bs0g, Ig); Ib = fmaf(i00, bs0b, Ib);
                            if (g.n_bulge > 1) { Ir = fmaf(i10, bs1r, Ir); Ig = fmaf(i10, bs1g, Ig); Ib = fmaf(i10, bs1b, Ib); }
                            Ir = fmaf(i01, bs0r, Ir); Ig = fmaf(i01, bs0g, Ig); Ib = fmaf(i01, bs0b, Ib);
                            if (g.n_bulge > 1) { Ir = fmaf(i11, bs1r, Ir); Ig = fmaf(i11, bs1g, Ig); Ib = fmaf(i11, bs1b, Ib); }
                            if (COUNT) { cn.evals += 2 * g.n_bulge; cn.samples += 2; cn.gates += 2 * (g.comp_count + g.noop_count); }
                            p = v_sub(p1, ds);
                            marched = m1 + stepf;
                        }
                        continue;
                    }
                }
#endif
                bool alive = marched < sure_below;
                if (!alive) {                       // the exact test, only near the ray's end: its operands come from the record
                    const f3 origin = {rec[6], rec[7], rec[8]};
                    alive = -v_dot(v_sub(p, origin), dir) < thr;
                }
                if (alive) {
                    // step = clamp(|p - camera| * rayStep, minRayStep, 0.01)
                    const f3 dc = v_sub(p, camrel_);
                    const float e2 = fmaf(dc.z, dc.z, fmaf(dc.y, dc.y, dc.x * dc.x));
                    if (e2 > sc.est2_hi) { stepf = s_hi; weight = w_hi; thr = thr_hi; lim = hi_lim; }
                    else if (e2 < sc.est2_lo) { stepf = s_lo; weight = w_lo; thr = thr_lo; lim = marched > lo_from ? sure_below : -INFINITY; }
                    else {
                        // the clamps as exact float thresholds on the length (scene_build.cpp), FP64 only in between
                        const float len_f = v_length(dc);
                        lim = -INFINITY;
                        if (len_f > sc.len_hi_thr) { stepf = s_hi; weight = w_hi; thr = thr_hi; }
                        else if (len_f <= sc.len_lo_thr) { stepf = s_lo; weight = w_lo; thr = thr_lo; }
                        else {
                            const double step = __dmul_rn((double)len_f, sc.ray_step);
                            stepf = __double2float_rn(step);
                            weight = (float)__dmul_rn(step, 200.0);
                            thr = __double2float_ru(__dadd_rn((double)length_f, step));
                        }
                    }
                    // getRadius (galaxycomponent.h:72-76); for orient = (0,1,0): (p.x*0 + p.y*1) + p.z*0 = p.y
                    const float dott = ORIENT_Y ? p.y : v_dot(p, orient);
                    need = !(fast_ok && !(fabsf(dott) <= g.slab_gate));           // an in-slab sample at p
#ifdef K1_DIAG_NOSLAB                   // timing experiment only (results are wrong): the march and the bulge alone
                    need = false;
#endif
                    if (!need) {
                        const float plen3 = sqrta(fmaf(p.z, p.z, fmaf(p.y, p.y, p.x * p.x)));
                        kw0 = bk0 * weight; kw1 = bk1 * weight;
                        bulge_add(kw0, br0, bs0r, bs0g, bs0b, plen3, Ir, Ig, Ib);
                        if (g.n_bulge > 1) bulge_add(kw1, br1, bs1r, bs1g, bs1b, plen3, Ir, Ig, Ib);
                        if (COUNT) { cn.evals += g.n_bulge; cn.samples++; cn.gates += g.comp_count + g.noop_count; }
                        ds = v_mul(dir, stepf);
                        p = v_sub(p, ds);
                        marched += stepf;
                        // I only grew: the floor of the general path is the identity here (scene_build.cpp, fast_ok)
                    }
                }
                const unsigned alive_mask = __ballot_sync(FULL, alive);
                idle = ~alive_mask;
                any_alive = alive_mask != 0u;
                if (__any_sync(FULL, need) || !any_alive) break;
            }
            rec[0] = p.x; rec[1] = p.y; rec[2] = p.z;
            rec[REC_THR] = thr; rec[REC_STEP] = stepf; rec[REC_WEIGHT] = weight; rec[REC_MARCHED] = marched;
        }
        if (!any_alive) break;
        if (need) {   // ---- one in-slab sample ------------------------------------------------------------------
            const f3 orient = {g.orient[0], g.orient[1], g.orient[2]};
            {
                const f3 sp = {rec[0], rec[1], rec[2]};
                const float dott = ORIENT_Y ? sp.y : v_dot(sp, orient);
                const f3 P = ORIENT_Y ? f3{sp.x, 0.0f, sp.z} : v_sub(sp, v_mul(orient, dott));   // p.y - p.y*1 = 0
                rec[REC_DOTT] = dott;
                rec[REC_PLEN] = sqrta(fmaf(P.z, P.z, fmaf(P.y, P.y, P.x * P.x)));
                rec[REC_WT] = __int_as_float(0x7fc00000);       // getWinding / getTheta: not formed yet (eval_component)
            }
#ifndef K1_NO_TWIRL_CACHE
            if (g.any_tshare) rec[REC_TG] = -1.0f;
#endif
            for (int k = 0; k < g.comp_count; k = __float_as_int(rec[REC_K]) + 1) {
                rec[REC_K] = __int_as_float(k);
                eval_component<COUNT, ORIENT_Y>(sc, g, sc.comp[g.comp_begin + k], rec, nz, Ir, Ig, Ib, cn);
            }
            if (COUNT) { cn.samples++; cn.gates += g.comp_count + g.noop_count; }
            const f3 dir = {rec[3], rec[4], rec[5]};
            const f3 p0 = {rec[0], rec[1], rec[2]};
            const float stepf = rec[REC_STEP];
            const f3 pn = v_sub(p0, v_mul(dir, stepf));
            rec[REC_MARCHED] = rec[REC_MARCHED] + stepf;
            rec[0] = pn.x; rec[1] = pn.y; rec[2] = pn.z;
            Ir = fmaxf(0.0f, Ir); Ig = fmaxf(0.0f, Ig); Ib = fmaxf(0.0f, Ib);   // Util::floor (util.h:62-64)
        }
        // all 32 lanes are here in every trip (the votes of the run loop name the full warp). Without the explicit
        // rendezvous the lanes that sat the in-slab sample out may go on through the run loop as a group of their own:
        // the full-mask votes make that correct, but each side then pays for the loop (ncu: 26 threads per instruction)
        __syncwarp();
    }
}

__device__ __forceinline__ int image_row(const DScene &sc, int L) {
    if (sc.band_rows > 0) return sc.y0 + (L / sc.band_rows) * sc.band_stride + (L % sc.band_rows);
    return sc.y0 + L;
}

// SINGLE: the frame holds one galaxy (the reference's CLI, the benchmarks, every shipped scene file): its index is a
// compile-time 0, so every galaxy field is a constant-bank operand of the instruction that uses it - no load, no
// register. The general kernel indexes the galaxy array at run time.
// MODE 2: SINGLE and the galaxy's orientation is exactly (0,1,0) with an identity rotmat (DGalaxy::orient_y) - every
// shipped galaxy file and the reference's CLI: dot(p, orient) is p.y and the twirl takes the short sequence of qtf.cuh.
template <bool COUNT, int MODE>
__global__ void __launch_bounds__(K1_THREADS, K1_MINBLOCKS)
k1_march(const __grid_constant__ DScene sc, float *__restrict__ out, unsigned long long *__restrict__ counters,
         const volatile int *abort_flag, unsigned int *progress, unsigned int n_tiles, unsigned int tiles_x) {
    constexpr bool SINGLE = MODE >= 1, ORIENT_Y = MODE == 2;
    gnoise::table_init();
    __syncthreads();

    const int lane = threadIdx.x & 31;
    NoiseCtx nz;
    nz.lane4 = gnoise::lane_const();
    Counts cn = {0, 0, 0, 0, 0};
    Rec rec;
    rec.base = reinterpret_cast<volatile float *>(gnoise::gn_dyn_smem + GN_DYN_SMEM) + threadIdx.x;

    // Persistent warps: the grid is one wave (K1_MINBLOCKS CTAs per SM, tables built once per CTA) and every warp draws
    // 8 x 4 pixel tiles from a ticket counter until none is left. Tile cost varies by two orders of magnitude (sky
    // versus the disk seen edge-on), so a fixed tile-to-CTA map left warps idle at a CTA's final barrier and SMs idle
    // at the end of the frame. progress[0] counts finished tiles (Rasterizer::getPercentDone), progress[1] is the ticket.
    for (;;) {
        unsigned int tile = 0;
        if (lane == 0) tile = atomicAdd(progress + 1, 1u);
        tile = __shfl_sync(0xffffffffu, tile, 0);
        if (tile >= n_tiles) break;
        if (*abort_flag) break;             // rasterizer.cpp:296-297; skipped pixels stay 0 (the frame is cleared before the pass)
        const int lx = (int)(tile % tiles_x) * 8 + (lane & 7);
        const int ly = (int)(tile / tiles_x) * 4 + (lane >> 3);
    // Every lane of the warp walks the same control flow (the march keeps the 32 rays in lockstep with warp votes);
    // a lane outside the image computes the ray of the nearest pixel inside it and marches nothing.
    const bool inimg = lx < sc.width && ly < sc.height;
    {
        const int cx = min(lx, sc.width - 1), cy = min(ly, sc.height - 1);
        // ---- GamerCamera::coord2ray (gamercamera.cpp:210-217) --------------------------------
        const int i = sc.x0 + cx, j = image_row(sc, cy);
        const double half = __dmul_rn((double)sc.full_size, 0.5);
        const float v0 = __double2float_rn(__dsub_rn(dmath::div_nc((double)i, half), 1.0));
        const float v1 = __double2float_rn(-__dsub_rn(dmath::div_nc((double)j, half), 1.0));
        f3 w;
        {
            const float *m = sc.inv_vp;
            w.x = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(v0, m[0]), __fmul_rn(v1, m[1])), m[2]), m[3]);
            w.y = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(v0, m[4]), __fmul_rn(v1, m[5])), m[6]), m[7]);
            w.z = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(v0, m[8]), __fmul_rn(v1, m[9])), m[10]), m[11]);
        }
        const f3 d = v_normalized(w);
        float *dst = out + 3 * ((size_t)cy * sc.width + cx);
        float Ir = 0.f, Ig = 0.f, Ib = 0.f;
        if (sc.flags & GAMER_FRAME_CONTINUE) { Ir = dst[0]; Ig = dst[1]; Ib = dst[2]; }   // unscaled I of the farther galaxies
        // ---- Rasterizer::renderPixel (rasterizer.cpp:379-416) --------------------------------
        for (int gi = 0; gi < (SINGLE ? 1 : sc.n_galaxies); gi++) {
            const DGalaxy &g = sc.gal[SINGLE ? 0 : gi];
            const f3 o = v_sub({sc.cam[0], sc.cam[1], sc.cam[2]}, {g.pos[0], g.pos[1], g.pos[2]});
            // Util::IntersectSphere (util.h:66-98)
            const f3 rD = {__fmul_rn(d.x, g.inv_axis2[0]), __fmul_rn(d.y, g.inv_axis2[1]), __fmul_rn(d.z, g.inv_axis2[2])};
            const f3 rO = {__fmul_rn(o.x, g.inv_axis2[0]), __fmul_rn(o.y, g.inv_axis2[1]), __fmul_rn(o.z, g.inv_axis2[2])};
            const double A = (double)v_dot(d, rD);
            const double B = __dmul_rn(2.0, (double)v_dot(d, rO));
            const double C = __dsub_rn((double)v_dot(o, rO), 1.0);
            const double S = __dsub_rn(__dmul_rn(B, B), __dmul_rn(__dmul_rn(4.0, A), C));
            bool hit = inimg && S > 0.0;
            const double sq = dmath::sqrt_nc(hit ? S : 1.0), twoA = __dmul_rn(2.0, A);
            const double t0 = dmath::div_nc(__dsub_rn(-B, sq), twoA);
            const double t1 = dmath::div_nc(__dadd_rn(-B, sq), twoA);
            f3 isp1 = v_add(o, v_mul(d, __double2float_rn(t0)));
            f3 isp2 = v_add(o, v_mul(d, __double2float_rn(t1)));
            if (t1 > 0.0) isp2 = o;
            if (t0 > 0.0 && t1 > 0.0) hit = false;
            march<COUNT, ORIENT_Y>(sc, g, rec, hit, isp1, isp2, o, nz, Ir, Ig, Ib, cn);
        }
        // rp.I * 0.01 / rayStep  (rasterizer.cpp:409): QVector3D*float then QVector3D/float
        if (!inimg) { }
        else if (sc.flags & GAMER_FRAME_PARTIAL) { dst[0] = Ir; dst[1] = Ig; dst[2] = Ib; }
        else {
            const float rrs = frcp_nc(sc.ray_step_f);
            dst[0] = fdiv_nc(__fmul_rn(Ir, sc.final_mul), sc.ray_step_f, rrs);
            dst[1] = fdiv_nc(__fmul_rn(Ig, sc.final_mul), sc.ray_step_f, rrs);
            dst[2] = fdiv_nc(__fmul_rn(Ib, sc.final_mul), sc.ray_step_f, rrs);
        }
    }
        if (lane == 0) atomicAdd(progress, 1u);
    }
    if (COUNT) {
        unsigned long long v[5] = {cn.rays, cn.samples, cn.gates, cn.evals, cn.octaves};
#pragma unroll
        for (int q = 0; q < 5; q++) {
            unsigned long long x = v[q];
            for (int off = 16; off > 0; off >>= 1) x += __shfl_down_sync(0xffffffffu, x, off);
            if (lane == 0 && x) atomicAdd(&counters[q], x);
        }
    }
}

// ---- K2: Buffer2D::PostProcess + ToColorBuffer (buffer2d.cpp:65-85,106-126) -------------------------
// HBM-bound: 12 B (+12 B stars) read, 8 B written per pixel. What is written is one BYTE per channel,
// (int)clamp(10 v), and the reference forms v through float -> double -> float -> double -> float: 13 conversions per
// pixel on the quarter-rate XU pipe, which was this kernel's limit (ncu r02: XU 76 % busy, HBM at half its rate).
// csrc/tone.cuh forms the same value in plain FP32 together with a proven bound on its distance from the reference's
// float, and decides the byte whenever no byte boundary lies within that bound; otherwise (about one pixel in 10^3 - 10^4)
// the pixel goes through tone_exact, the reference's own chain. The result is the reference's byte in every case.

// float(pow(double(x), gamma)) like Buffer2D::PostProcess; out of line: twelve inlined copies per thread spill
__device__ __noinline__ float pow_gamma(float x, double gamma) { return (float)pow((double)x, gamma); }
template <bool GAMMA>
__device__ __forceinline__ void tone_px(float x, float y, float z, float inv_exposure, double gamma, double saturation,
                                        const tone::ToneSat k, uint32_t &argb, uint32_t &shadow) {
    x = __fmul_rn(x, inv_exposure); y = __fmul_rn(y, inv_exposure); z = __fmul_rn(z, inv_exposure);
    if (GAMMA) { x = pow_gamma(x, gamma); y = pow_gamma(y, gamma); z = pow_gamma(z, gamma); }
    tone::tone_mapped(x, y, z, saturation, k, argb, shadow);
}

template <bool GAMMA>
__global__ void __launch_bounds__(256, 4)
k2_tonemap(const float *__restrict__ rgb, const float *__restrict__ stars, uint32_t *__restrict__ argb,
           uint32_t *__restrict__ shadow, size_t n, float inv_exposure, double gamma, double saturation,
           const tone::ToneSat ks) {
    // The scalar kernel: buffers that are not 16-byte aligned, frames of fewer than four pixels, and the n % 4 remainder
    // of k2_tonemap_v4. Persistent grid (4 CTAs per SM), grid-stride, FOUR independent pixels per thread and iteration
    // with all loads issued before the arithmetic: a single pixel per iteration is a chain of dependent load
    // latencies. The three 4-byte loads of a warp cover the same 384 contiguous bytes, so every fetched
    // sector is fully used.
    const size_t nth = (size_t)gridDim.x * blockDim.x;
    size_t k = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    for (; k + 3 * nth < n; k += 4 * nth) {
        float v[4][3];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const size_t q = 3 * (k + u * nth);
            v[u][0] = rgb[q]; v[u][1] = rgb[q + 1]; v[u][2] = rgb[q + 2];
        }
        if (stars) {
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const size_t q = 3 * (k + u * nth);
                v[u][0] = __fadd_rn(v[u][0], stars[q]); v[u][1] = __fadd_rn(v[u][1], stars[q + 1]); v[u][2] = __fadd_rn(v[u][2], stars[q + 2]);
            }
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            uint32_t o, h;
            tone_px<GAMMA>(v[u][0], v[u][1], v[u][2], inv_exposure, gamma, saturation, ks, o, h);
            if (argb)   argb[k + u * nth] = o;
            if (shadow) shadow[k + u * nth] = h;
        }
    }
    for (; k < n; k += nth) {
        float x = rgb[3 * k], y = rgb[3 * k + 1], z = rgb[3 * k + 2];
        if (stars) { x = __fadd_rn(x, stars[3 * k]); y = __fadd_rn(y, stars[3 * k + 1]); z = __fadd_rn(z, stars[3 * k + 2]); }
        uint32_t o, h;
        tone_px<GAMMA>(x, y, z, inv_exposure, gamma, saturation, ks, o, h);
        if (argb)   argb[k] = o;
        if (shadow) shadow[k] = h;
    }
}

// K2 for 16-byte aligned buffers (the normal case; launch_k2 checks): four CONSECUTIVE pixels per thread — three 16-byte
// loads (48 B of RGB), one 16-byte store per output image, a quarter of the memory instructions of the scalar kernel —
// and no grid-stride loop: one group per thread, 1536 resident threads per SM, the block scheduler does the balancing.
// The kernel is bound by instruction issue (ncu r02b: 70 warp instructions per pixel before this arrangement), so what
// is not arithmetic is kept out of it (and what remains above the copy-only kernel's 14.3 us is not issue either: 70 ->
// 54 instructions per pixel changed nothing, nor did 64..512 threads per block; a persistent variant fed by
// cp.async.bulk into per-warp or per-CTA shared-memory rings was 2-6 us SLOWER and is not kept): the four pixels are decided in FP32 first and ONE branch covers the rare group
// with an undecided pixel; the stars addend is a template parameter (as a run-time test it becomes 3 predicated loads
// and 12 predicated adds that issue for nothing); indices are 32-bit; the n % 4 remainder pixels go to the scalar kernel.
#ifndef K2_BLOCK
#define K2_BLOCK 256
#endif
#ifndef K2_MINB
#define K2_MINB 6
#endif
template <bool GAMMA>
__device__ __forceinline__ void expose_px(float &x, float &y, float &z, float inv_exposure, double gamma) {
    x = __fmul_rn(x, inv_exposure); y = __fmul_rn(y, inv_exposure); z = __fmul_rn(z, inv_exposure);
    if (GAMMA) { x = pow_gamma(x, gamma); y = pow_gamma(y, gamma); z = pow_gamma(z, gamma); }
}
// four consecutive pixels (48 B) of the frame, plus the star layer, exposed and gamma-mapped. (Loading the warp's 1536
// bytes with three fully coalesced loads and redistributing them through shared memory was measured: the copy-only
// kernel takes 14.27 us either way and the full kernel got slower, 16.35 -> 18.4 us.)
template <bool GAMMA, bool STARS>
__device__ __forceinline__ void k2_load_group(const float4 *__restrict__ r4, const float4 *__restrict__ s4, uint32_t tid,
                                              float inv_exposure, double gamma, float4 &a, float4 &b, float4 &c) {
    a = r4[3 * (size_t)tid]; b = r4[3 * (size_t)tid + 1]; c = r4[3 * (size_t)tid + 2];
    if (STARS) {
        const float4 p = s4[3 * (size_t)tid], q = s4[3 * (size_t)tid + 1], r = s4[3 * (size_t)tid + 2];
        a.x = __fadd_rn(a.x, p.x); a.y = __fadd_rn(a.y, p.y); a.z = __fadd_rn(a.z, p.z); a.w = __fadd_rn(a.w, p.w);
        b.x = __fadd_rn(b.x, q.x); b.y = __fadd_rn(b.y, q.y); b.z = __fadd_rn(b.z, q.z); b.w = __fadd_rn(b.w, q.w);
        c.x = __fadd_rn(c.x, r.x); c.y = __fadd_rn(c.y, r.y); c.z = __fadd_rn(c.z, r.z); c.w = __fadd_rn(c.w, r.w);
    }
    expose_px<GAMMA>(a.x, a.y, a.z, inv_exposure, gamma);
    expose_px<GAMMA>(a.w, b.x, b.y, inv_exposure, gamma);
    expose_px<GAMMA>(b.z, b.w, c.x, inv_exposure, gamma);
    expose_px<GAMMA>(c.y, c.z, c.w, inv_exposure, gamma);
}
template <bool GAMMA, bool STARS>
__device__ __noinline__ void k2_fix_group(const float4 *r4, const float4 *s4, uint4 *argb, uint4 *shadow, uint32_t tid,
                                          uint32_t undecided, float inv_exposure, double gamma, double saturation) {
    const float *rgb = reinterpret_cast<const float *>(r4 + 3 * tid), *st = reinterpret_cast<const float *>(s4 + 3 * tid);
    for (int u = 0; u < 4; u++) {
        if (!((undecided >> u) & 1u)) continue;
        float x = rgb[3 * u], y = rgb[3 * u + 1], z = rgb[3 * u + 2];
        if (STARS) { x = __fadd_rn(x, st[3 * u]); y = __fadd_rn(y, st[3 * u + 1]); z = __fadd_rn(z, st[3 * u + 2]); }
        expose_px<GAMMA>(x, y, z, inv_exposure, gamma);
        uint32_t o, h;
        tone::tone_mapped_exact(x, y, z, saturation, o, h);
        if (argb)   reinterpret_cast<uint32_t *>(argb + tid)[u] = o;
        if (shadow) reinterpret_cast<uint32_t *>(shadow + tid)[u] = h;
    }
}
template <bool GAMMA, bool STARS>
__global__ void __launch_bounds__(K2_BLOCK, K2_MINB)
k2_tonemap_v4(const float4 *__restrict__ r4, const float4 *__restrict__ s4, uint4 *__restrict__ argb,
              uint4 *__restrict__ shadow, uint32_t groups, float inv_exposure, double gamma, double saturation,
              const tone::ToneSat ks) {
    const uint32_t tid = blockIdx.x * K2_BLOCK + threadIdx.x;
    if (tid >= groups) return;
    float4 a, b, c;
    k2_load_group<GAMMA, STARS>(r4, s4, tid, inv_exposure, gamma, a, b, c);
    uint4 o, h;
#ifdef K2_DIAG_COPY      // timing experiment only (results are wrong): the memory pattern without the arithmetic
    o = make_uint4(__float_as_uint(a.x), __float_as_uint(a.w), __float_as_uint(b.z), __float_as_uint(c.y));
    h = make_uint4(__float_as_uint(a.y), __float_as_uint(b.x), __float_as_uint(b.w), __float_as_uint(c.z));
    const bool ok0 = true, ok1 = true, ok2 = true, ok3 = c.w != 12345.0f;
#else
    const bool ok0 = tone::tone_mapped_fast(a.x, a.y, a.z, ks, o.x, h.x);
    const bool ok1 = tone::tone_mapped_fast(a.w, b.x, b.y, ks, o.y, h.y);
    const bool ok2 = tone::tone_mapped_fast(b.z, b.w, c.x, ks, o.z, h.z);
    const bool ok3 = tone::tone_mapped_fast(c.y, c.z, c.w, ks, o.w, h.w);
#endif
    if (argb)   argb[tid] = o;
    if (shadow) shadow[tid] = h;
    // about one group in a thousand has an undecided pixel: its words are written again by the exact chain, out of line,
    // from pixels loaded again - nothing of the cold path is live in the path every thread takes
    const uint32_t undecided = (ok0 ? 0u : 1u) | (ok1 ? 0u : 2u) | (ok2 ? 0u : 4u) | (ok3 ? 0u : 8u);
    if (undecided) k2_fix_group<GAMMA, STARS>(r4, s4, argb, shadow, tid, undecided, inv_exposure, gamma, saturation);
}

// ---- K4: background star layer. Buffer2D::RenderGaussian (buffer2d.cpp:224-243) for every star --------
// One warp per star, lanes over the stamp. The reference combines stamps with a per-channel max, which
// commutes: an atomicMax on the bit pattern of the (non-negative) float gives the identical layer.
__global__ void __launch_bounds__(256)
k4_stars(const GamerStar *__restrict__ stars, int n, float *__restrict__ layer, int full_size, int x0, int y0,
         int width, int height, int band_rows, int band_stride) {
    const int star = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (star >= n) return;
    const GamerStar s = stars[star];
    const int w = s.w, half = w / 2, side = 2 * half;           // for (x = -w/2; x < w/2; x++)
    if (side <= 0) return;
    const double ss = (double)0.01f;
    for (int idx = lane; idx < side * side; idx += 32) {
        const int x = idx / side - half, y = idx % side - half;
        const int xx = s.x + x, yy = s.y + y;
        if (xx < 0 || xx >= full_size || yy < 0 || yy >= full_size) continue;
        // image pixel -> this context's local framebuffer (tile origin, interleaved bands)
        const int lx = xx - x0;
        int rel = yy - y0, ly;
        if (lx < 0 || lx >= width || rel < 0) continue;
        if (band_rows > 0) {
            const int r = rel % band_stride;
            if (r >= band_rows) continue;
            ly = (rel / band_stride) * band_rows + r;
        } else ly = rel;
        if (ly >= height) continue;
        const double dx = x / (double)w, dy = y / (double)w;
        const double d = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
        const float v = (float)exp(-d / ss);
        unsigned int *dst = reinterpret_cast<unsigned int *>(layer + 3 * ((size_t)ly * width + lx));
#pragma unroll
        for (int c = 0; c < 3; c++) {
            const float val = __fmul_rn(v, s.cs[c]);
            if (val > 0.0f) atomicMax(dst + c, __float_as_uint(val));
        }
    }
}

// ---- probes: the same device noise functions, one point per thread ----------------------------------
__global__ void __launch_bounds__(256)
kp_simplex(const float *__restrict__ xyz, int n, double frequency, int32_t *__restrict__ lat_out, float *__restrict__ val_out) {
    gnoise::table_init();
    __syncthreads();
    int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    uint32_t lane4 = gnoise::lane_const();
    gnoise::Skewed su, sv, sw;
    gnoise::skew(__dmul_rn((double)xyz[3 * q], frequency), __dmul_rn((double)xyz[3 * q + 1], frequency),
                 __dmul_rn((double)xyz[3 * q + 2], frequency), su, sv, sw);
    gnoise::Lattice L;
    float fu, fv, fw;
    uint32_t i7, j7, k7;
    gnoise::cell_of<true>(su, 0, i7, fu);
    gnoise::cell_of<true>(sv, 0, j7, fv);
    gnoise::cell_of<true>(sw, 0, k7, fw);
    // full 32-bit cells for the parity test: the plain cell, minus one where the fastfloor quirk fired
    L.i = su.hi - (int)((__funnelshift_l(su.lo, (uint32_t)su.hi, 7) - i7) >> 7);
    L.j = sv.hi - (int)((__funnelshift_l(sv.lo, (uint32_t)sv.hi, 7) - j7) >> 7);
    L.k = sw.hi - (int)((__funnelshift_l(sw.lo, (uint32_t)sw.hi, 7) - k7) >> 7);
    float v = 32.0f * gnoise::simplex_cell<true>(lane4, gnoise::lane16_of(lane4), i7, j7, k7, fu, fv, fw, &L);
    if (val_out) val_out[q] = v;
    if (lat_out) {
        int32_t *o = lat_out + 10 * (size_t)q;
        o[0] = L.i; o[1] = L.j; o[2] = L.k; o[3] = L.c1; o[4] = L.c2;
        o[5] = L.g0; o[6] = L.g1; o[7] = L.g2; o[8] = L.g3; o[9] = 0;
    }
}

__global__ void __launch_bounds__(256)
kp_fbm(const float *__restrict__ xyz, int n, int octaves, float pers, float inv_maxamp, double scale, float *__restrict__ out) {
    gnoise::table_init();
    __syncthreads();
    int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    out[q] = gnoise::fbm(gnoise::lane_const(), xyz[3 * q], xyz[3 * q + 1],
                         xyz[3 * q + 2], scale, octaves, pers, inv_maxamp);
}

struct RidgedW { float w[16]; };
__global__ void __launch_bounds__(256)
kp_ridged(const float *__restrict__ xyz, int n, int octaves, float lac, float offset, float gain, RidgedW rw,
          float *__restrict__ out) {
    gnoise::table_init();
    __syncthreads();
    int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    out[q] = gnoise::ridged(gnoise::lane_const(), xyz[3 * q], xyz[3 * q + 1],
                            xyz[3 * q + 2], octaves, lac, offset, gain, rw.w[0], rw.w[1] / rw.w[0]);
}

}  // namespace

// ---- launchers (host) -----------------------------------------------------------------------------
namespace gk {

}  // namespace gk
namespace {
// The noise code addresses its tables with immediates that assume the dynamic shared window begins at GN_DYN_BASE
// (noise.cuh). That offset is the driver's reservation and could differ under a debugger, a sanitizer or another driver:
// it is checked here, once per device, so that context creation fails with a message; the __trap() in table_init stays
// as an assertion only.
__device__ unsigned int g_dyn_smem_base_probe;
__global__ void k_dyn_smem_base() { g_dyn_smem_base_probe = (unsigned int)__cvta_generic_to_shared(gnoise::gn_dyn_smem); }
}  // namespace
namespace gk {

// kernels that use the noise table need > 48 KB of dynamic shared memory: opt in once per device
cudaError_t check_dyn_smem_base() {
    k_dyn_smem_base<<<1, 1, 1024>>>();
    cudaError_t e = cudaGetLastError();
    unsigned int base = 0;
    if (e == cudaSuccess) e = cudaMemcpyFromSymbol(&base, g_dyn_smem_base_probe, sizeof(base));
    if (e == cudaSuccess && base != GN_DYN_BASE) e = cudaErrorNotSupported;   // api.cu turns this into a message
    return e;
}

cudaError_t prepare_device() {
    cudaError_t e = cudaSuccess;
#define K1_ATTR(C, M) if (e == cudaSuccess) e = cudaFuncSetAttribute(k1_march<C, M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)K1_DYN_SMEM)
    K1_ATTR(true, 0); K1_ATTR(false, 0); K1_ATTR(true, 1); K1_ATTR(false, 1); K1_ATTR(true, 2); K1_ATTR(false, 2);
#undef K1_ATTR
    if (e == cudaSuccess) e = cudaFuncSetAttribute(kp_simplex, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GN_DYN_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(kp_fbm, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GN_DYN_SMEM);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(kp_ridged, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GN_DYN_SMEM);
    if (e == cudaSuccess) e = check_dyn_smem_base();
    return e;
}

cudaError_t launch_k1(const DScene &sc, float *out_dev, unsigned long long *counters_dev, const int *abort_dev,
                      unsigned int *progress_dev, bool counted, int sm_count, cudaStream_t st, unsigned int *n_tiles_out) {
    const unsigned int tiles_x = (sc.width + 7) / 8, tiles_y = (sc.height + 3) / 4, n_tiles = tiles_x * tiles_y;
    if (n_tiles_out) *n_tiles_out = n_tiles;
    const unsigned int warps_per_cta = K1_THREADS / 32;
    unsigned int grid = (n_tiles + warps_per_cta - 1) / warps_per_cta;
    const unsigned int wave = (unsigned int)sm_count * K1_MINBLOCKS;
    if (grid > wave) grid = wave;
    if (grid < 1) grid = 1;
    const int mode = sc.n_galaxies != 1 ? 0 : (sc.gal[0].orient_y ? 2 : 1);
#define K1_LAUNCH(C, M) k1_march<C, M><<<grid, K1_THREADS, K1_DYN_SMEM, st>>>(sc, out_dev, counters_dev, abort_dev, progress_dev, n_tiles, tiles_x)
    if (counted) { if (mode == 2) K1_LAUNCH(true, 2); else if (mode == 1) K1_LAUNCH(true, 1); else K1_LAUNCH(true, 0); }
    else         { if (mode == 2) K1_LAUNCH(false, 2); else if (mode == 1) K1_LAUNCH(false, 1); else K1_LAUNCH(false, 0); }
#undef K1_LAUNCH
    return cudaGetLastError();
}

cudaError_t launch_k2(const float *rgb, const float *stars, uint32_t *argb, uint32_t *shadow, size_t n, double exposure,
                      double gamma, double saturation, int sm_count, cudaStream_t st, int *n_launches) {
    if (n_launches) *n_launches = 1;
    size_t want = (n + 255) / 256;
    size_t cap = (size_t)sm_count * 4;          // 4 resident CTAs of 256 threads per SM (64 registers each), grid-stride
    int blocks = (int)(want < cap ? want : cap);
    if (blocks < 1) blocks = 1;
    const uintptr_t al = (uintptr_t)rgb | (uintptr_t)stars | (uintptr_t)argb | (uintptr_t)shadow;
    // the float constants of the launch are formed here, once: in the kernel they would be XU conversions per pixel
    const float inv_exposure = (float)(1.0 / exposure);
    const tone::ToneSat ks = tone::tone_sat(saturation);
    const bool g = gamma != 1.0;
    if ((al & 15u) == 0 && n >= 4 && (n >> 2) <= 0x7FFFFFFFu / 3u) {
        const uint32_t groups = (uint32_t)(n >> 2);          // one group of four consecutive pixels per thread
        const unsigned grid = (groups + K2_BLOCK - 1) / K2_BLOCK;
        const size_t done = (size_t)groups << 2, rest = n - done;          // the n % 4 remainder: the scalar kernel, one block
        cudaError_t e;
#define K2_LAUNCH(G, S) k2_tonemap_v4<G, S><<<grid, K2_BLOCK, 0, st>>>( \
            reinterpret_cast<const float4 *>(rgb), reinterpret_cast<const float4 *>(stars), reinterpret_cast<uint4 *>(argb), \
            reinterpret_cast<uint4 *>(shadow), groups, inv_exposure, gamma, saturation, ks)
        if (g) { if (stars) K2_LAUNCH(true, true); else K2_LAUNCH(true, false); }
        else   { if (stars) K2_LAUNCH(false, true); else K2_LAUNCH(false, false); }
#undef K2_LAUNCH
        e = cudaGetLastError();
        if (n_launches) *n_launches = rest ? 2 : 1;
        if (e != cudaSuccess || rest == 0) return e;
        rgb += 3 * done; if (stars) stars += 3 * done; if (argb) argb += done; if (shadow) shadow += done;
        if (g) k2_tonemap<true><<<1, 32, 0, st>>>(rgb, stars, argb, shadow, rest, inv_exposure, gamma, saturation, ks);
        else k2_tonemap<false><<<1, 32, 0, st>>>(rgb, stars, argb, shadow, rest, inv_exposure, gamma, saturation, ks);
        return cudaGetLastError();
    }
    if (g) k2_tonemap<true><<<blocks, 256, 0, st>>>(rgb, stars, argb, shadow, n, inv_exposure, gamma, saturation, ks);
    else k2_tonemap<false><<<blocks, 256, 0, st>>>(rgb, stars, argb, shadow, n, inv_exposure, gamma, saturation, ks);
    return cudaGetLastError();
}

cudaError_t launch_k4(const GamerStar *stars, int n, float *layer, const DScene &sc, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    k4_stars<<<(n + 7) / 8, 256, 0, st>>>(stars, n, layer, sc.full_size, sc.x0, sc.y0, sc.width, sc.height, sc.band_rows,
                                          sc.band_stride);
    return cudaGetLastError();
}

cudaError_t launch_probe_simplex(const float *xyz, int n, double frequency, int32_t *lat, float *val, cudaStream_t st) {
    kp_simplex<<<(n + 255) / 256, 256, GN_DYN_SMEM, st>>>(xyz, n, frequency, lat, val);
    return cudaGetLastError();
}
cudaError_t launch_probe_fbm(const float *xyz, int n, int octaves, float pers, float inv_maxamp, double scale, float *out,
                             cudaStream_t st) {
    kp_fbm<<<(n + 255) / 256, 256, GN_DYN_SMEM, st>>>(xyz, n, octaves, pers, inv_maxamp, scale, out);
    return cudaGetLastError();
}
cudaError_t launch_probe_ridged(const float *xyz, int n, int octaves, float lac, float offset, float gain, const float *rw,
                                float *out, cudaStream_t st) {
    RidgedW w;
    for (int i = 0; i < 16; i++) w.w[i] = (i < octaves) ? rw[i] : 0.f;
    kp_ridged<<<(n + 255) / 256, 256, GN_DYN_SMEM, st>>>(xyz, n, octaves, lac, offset, gain, w, out);
    return cudaGetLastError();
}

// ---- allocation-free fBm evaluation (gamer_b200_fbm_static) ---------------------------------------------------------
// kp_fbm with its staging memory (the points going in, the values coming out) in a file-scope device array instead of
// cudaMalloc'd buffers: for callers that may not allocate (CUDA graph capture, serving loops, harnesses that forbid it).
// Every GPU that loads this library holds the array (GK_STATIC_BYTES = 4 MB) until the process exits. In HOST code the
// array's name is only a stand-in for the symbol, not a device address: the address comes from cudaGetSymbolAddress,
// once per device, with that device current (fbm_static_setup below). All calls share the array and it keeps the
// last call's contents; nothing in it has to start at zero (the kernel writes every value it later reads back).
}  // namespace gk

#define GK_STATIC_MAX_POINTS 262144
#define GK_STATIC_BYTES (4u * GK_STATIC_MAX_POINTS * sizeof(float))
static __device__ __align__(256) float g_fbm_static_scratch[4 * GK_STATIC_MAX_POINTS];   // [0, 3 MAX): xyz; [3 MAX, 4 MAX): values

namespace gk {

int fbm_static_max_points() { return GK_STATIC_MAX_POINTS; }

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_static_mu;                    // one call at a time per process: the scratch is shared
float *g_static_ptr[kMaxDevices];          // device address of g_fbm_static_scratch, per device
bool g_static_ready[kMaxDevices];

// once per device, with that device current
cudaError_t fbm_static_setup(int dev) {
    if (g_static_ready[dev]) return cudaSuccess;
    cudaError_t e = prepare_device();      // kp_fbm needs > 48 KB of dynamic shared memory: an attribute, not an allocation
    if (e != cudaSuccess) return e;
    float *ptr = nullptr;
    e = cudaGetSymbolAddress((void **)&ptr, g_fbm_static_scratch);
    if (e != cudaSuccess) return e;
    if (!ptr) return cudaErrorInvalidSymbol;
    g_static_ptr[dev] = ptr;
    g_static_ready[dev] = true;
    return cudaSuccess;
}
}  // namespace

// Runs on the calling thread's current device, on the legacy default stream; blocking. No cudaMalloc / cudaFree /
// allocating library call. The caller (api.cu) has checked 0 < n <= GK_STATIC_MAX_POINTS.
cudaError_t fbm_static(const float *xyz_host, int n, int octaves, float pers, float inv_maxamp, double scale, float *out_host) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_static_mu);
    e = fbm_static_setup(dev);
    if (e != cudaSuccess) return e;
    float *xyz_dev = g_static_ptr[dev], *val_dev = g_static_ptr[dev] + 3 * (size_t)GK_STATIC_MAX_POINTS;
    e = cudaMemcpyAsync(xyz_dev, xyz_host, (size_t)n * 3 * sizeof(float), cudaMemcpyHostToDevice, cudaStreamLegacy);
    if (e != cudaSuccess) return e;
    kp_fbm<<<(n + 255) / 256, 256, GN_DYN_SMEM, cudaStreamLegacy>>>(xyz_dev, n, octaves, pers, inv_maxamp, scale, val_dev);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    e = cudaMemcpyAsync(out_host, val_dev, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, cudaStreamLegacy);
    if (e != cudaSuccess) return e;
    return cudaStreamSynchronize(cudaStreamLegacy);
}

}  // namespace gk
