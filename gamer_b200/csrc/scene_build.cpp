// scene_build.cpp — host: ABI scene (GamerFrame/GamerGalaxy/GamerComponent) -> DScene kernel parameter.
// The exact gate thresholds are found here by bisection over float bit patterns, evaluating the
// reference's double-precision expressions (galaxycomponent.h:61-81, galaxycomponent.cpp:58-65).
#include <cmath>
#include <cstdio>
#include <cstring>
#include <initializer_list>
#include <limits>
#include "scene_build.h"

namespace {

inline float bits_f(uint32_t b) { float f; std::memcpy(&f, &b, 4); return f; }

// largest non-negative float for which pred holds, pred monotone (true ... true false ... false).
// returns -1 if pred(0) is false, +inf if pred holds for every finite float.
template <class P> float largest_passing(P pred) {
    if (!pred(0.0f)) return -1.0f;
    uint32_t lo = 0, hi = 0x7f800000u;              // lo passes; hi = +inf
    if (pred(bits_f(hi - 1))) return std::numeric_limits<float>::infinity();
    hi = hi - 1;                                    // hi fails
    while (hi - lo > 1) {
        uint32_t mid = lo + (hi - lo) / 2;
        if (pred(bits_f(mid))) lo = mid; else hi = mid;
    }
    return bits_f(lo);
}

inline double ref_clamp(double v, double mi, double ma) { v = (v < ma) ? v : ma; v = (mi < v) ? v : mi; return v; }

// Qt 5.15 QVector3D::normalized in double, result float
void normalize3(const float in[3], float out[3]) {
    double len = double(in[0]) * in[0] + double(in[1]) * in[1] + double(in[2]) * in[2];
    if (std::fabs(len - 1.0) <= 1e-12) { out[0] = in[0]; out[1] = in[1]; out[2] = in[2]; return; }
    if (std::fabs(len) > 1e-12) {
        double s = std::sqrt(len);
        out[0] = float(in[0] / s); out[1] = float(in[1] / s); out[2] = float(in[2] / s);
        return;
    }
    out[0] = out[1] = out[2] = 0.f;
}

}  // namespace

// ---- getWinding as a certified table --------------------------------------------------------------------------------
// getWinding(rad) = atan(exp(-0.25 / (0.5 (rad + 0.05))) / windingB) * 2 * windingN (galaxycomponent.h:156-165) is, in
// s = 1 / (rad + 0.05), the function G(s) = atan(exp(-s / 2) / windingB) * 2 windingN: analytic, with its nearest
// singularities at distance pi from the real axis, so on intervals of width 0.25 a degree-9 Chebyshev interpolant is
// accurate to ~1e-17 of the local magnitude. The kernel evaluates one reciprocal and one Horner chain instead of two
// divisions, an exp and an atan in FP64 (a serial chain of ~90 operations, once per in-slab sample).
// What the reference consumes is the FLOAT twirl angle float(W * winding * 180): the kernel checks that the table's
// certified error bound cannot change that float and evaluates getWinding directly when it could (kernels.cu).
// The certification is done here, against libm, on the radii the kernel can meet: a table that misses WT_CERT_REL is dropped.
#define WT_CERT_REL 0x1p-46

namespace {
double winding_direct(const GamerGalaxy &g, float radius) {       // the reference's expression, in its operation order
    const double r = double(radius) + 0.05;
    return std::atan(std::exp(-0.25 / (0.5 * r)) / g.winding_b) * 2 * g.winding_n;
}
}  // namespace

double gk_winding_from_table(const DGalaxy &d, const double *coeffs, float radius) {
    const double s = 1.0 / (double(radius) + 0.05);
    const double u = (s - d.wt_s0) * (1.0 / WT_H);
    int i = int(u);
    if (i < 0) i = 0;
    if (i > d.wt_n - 1) i = d.wt_n - 1;
    const double x = std::fma(2.0, u - double(i), -1.0);
    const double *c = coeffs + d.wt_off + size_t(i) * WT_COEFFS;
    double v = c[WT_COEFFS - 1];
    for (int k = WT_COEFFS - 2; k >= 0; k--) v = std::fma(v, x, c[k]);
    return v;
}

void gk_build_winding_table(const GamerGalaxy &g, DGalaxy &d, std::vector<double> &coeffs) {
    d.wt_n = 0; d.wt_s0 = 0; d.wt_off = int32_t(coeffs.size());
    // radius = |P| / axis.x with P inside the ellipsoid: at most max(axis) / |axis.x| (a little more for rounding)
    const double amax = std::fmax(std::fabs(g.axis[0]), std::fmax(std::fabs(g.axis[1]), std::fabs(g.axis[2])));
    const double r_max = 1.02 * amax / std::fabs(g.axis[0]) + 0.01;
    const double s_hi = 20.0, s_lo = 1.0 / (r_max + 0.05);           // rad = 0 is s = 20
    const int n = int(std::ceil((s_hi - s_lo) / WT_H)) + 1;
    if (n > WT_MAX_INTERVALS || !(g.winding_b != 0)) return;
    const double s0 = s_hi - n * WT_H + 0.5 * WT_H;                  // the last interval ends half an interval beyond s = 20
    const size_t base = coeffs.size();
    coeffs.resize(base + size_t(n) * WT_COEFFS);
    auto G = [&](double s) { return std::atan(std::exp(-0.5 * s) / g.winding_b) * 2 * g.winding_n; };
    const double PI = 3.14159265358979323846;
    for (int i = 0; i < n; i++) {
        const double mid = s0 + (i + 0.5) * WT_H, half = 0.5 * WT_H;
        double fj[WT_COEFFS], ch[WT_COEFFS];
        for (int j = 0; j < WT_COEFFS; j++) fj[j] = G(mid + half * std::cos(PI * (j + 0.5) / WT_COEFFS));
        for (int k = 0; k < WT_COEFFS; k++) {                         // Chebyshev coefficients of the interpolant
            double a = 0;
            for (int j = 0; j < WT_COEFFS; j++) a += fj[j] * std::cos(PI * k * (j + 0.5) / WT_COEFFS);
            ch[k] = a * (k == 0 ? 1.0 : 2.0) / WT_COEFFS;
        }
        // Chebyshev -> monomial basis: T_0 = 1, T_1 = x, T_{k+1} = 2 x T_k - T_{k-1}
        double mono[WT_COEFFS] = {0}, Tkm1[WT_COEFFS] = {1}, Tk[WT_COEFFS] = {0, 1};
        for (int q = 0; q < WT_COEFFS; q++) mono[q] += ch[0] * Tkm1[q] + ch[1] * Tk[q];
        for (int k = 2; k < WT_COEFFS; k++) {
            double Tn[WT_COEFFS] = {0};
            for (int q = 0; q < WT_COEFFS; q++) {
                if (q > 0) Tn[q] += 2.0 * Tk[q - 1];
                Tn[q] -= Tkm1[q];
            }
            for (int q = 0; q < WT_COEFFS; q++) { mono[q] += ch[k] * Tn[q]; Tkm1[q] = Tk[q]; Tk[q] = Tn[q]; }
        }
        for (int q = 0; q < WT_COEFFS; q++) coeffs[base + size_t(i) * WT_COEFFS + q] = mono[q];
    }
    d.wt_s0 = s0; d.wt_n = n;
    // certification on float radii across the whole range (dense near both ends of every interval)
    double worst = 0;
    for (int i = 0; i < n && worst <= WT_CERT_REL; i++)
        for (int t = 0; t <= 16; t++) {
            const double s = s0 + (i + t / 16.0) * WT_H;
            const double r = 1.0 / s - 0.05;
            if (r < 0 || r > r_max) continue;
            const float radius = float(r);
            const double want = winding_direct(g, radius), got = gk_winding_from_table(d, coeffs.data(), radius);
            const double rel = std::fabs(got - want) / std::fmax(std::fabs(want), 1e-300);
            if (!(rel <= worst)) worst = rel;
        }
    if (!(worst <= WT_CERT_REL)) { coeffs.resize(base); d.wt_n = 0; }
}

int gk_build_scene(const GamerFrame *f, const GamerGalaxy *gal, const GamerComponent *comp, DScene *out, char *err,
                   size_t errlen, std::vector<double> *wtab_host) {
    auto fail = [&](const char *m) { if (err && errlen) std::snprintf(err, errlen, "%s", m); return GAMER_E_ARG; };
    if (!f || !out) return fail("null frame");
    if (f->width <= 0 || f->height <= 0 || f->full_size <= 0) return fail("frame size must be positive");
    if (f->n_galaxies < 0 || f->n_galaxies > DEV_MAX_GALAXIES) return fail("n_galaxies out of range (max 8 per frame)");
    if (f->n_components < 0) return fail("n_components negative");
    if ((f->n_galaxies && !gal) || (f->n_components && !comp)) return fail("null galaxy/component array");
    if (f->flags & ~(GAMER_FRAME_CONTINUE | GAMER_FRAME_PARTIAL | GAMER_FRAME_NO_RADIAL_GATE)) return fail("unknown frame flags");
    if (f->band_rows < 0 || (f->band_rows > 0 && f->band_stride < f->band_rows)) return fail("bad band_rows/band_stride");
    if (!(f->ray_step > 0.0) || !(f->min_ray_step > 0.0) || f->min_ray_step > 0.01) return fail("ray_step/min_ray_step out of range");

    std::memset(out, 0, sizeof(*out));
    DScene &s = *out;
    s.width = f->width; s.height = f->height; s.full_size = f->full_size;
    s.x0 = f->x0; s.y0 = f->y0; s.band_rows = f->band_rows; s.band_stride = f->band_stride;
    s.n_galaxies = f->n_galaxies;
    s.flags = f->flags;
    for (int i = 0; i < 16; i++) s.inv_vp[i] = f->inv_vp[i];
    for (int i = 0; i < 3; i++) s.cam[i] = f->cam[i];
    s.ray_step = f->ray_step;
    s.ray_step_f = float(f->ray_step);
    s.min_ray_step = f->min_ray_step;
    s.final_mul = float(0.01);
    {   // squared-distance bounds outside which clamp(|p-cam|*rayStep, min, 0.01) is certainly clamped
        double hi = 0.01 * (1.0 + 2e-5) / f->ray_step, lo = f->min_ray_step * (1.0 - 2e-5) / f->ray_step;
        s.est2_hi = float(hi * hi * (1.0 + 1e-6));
        s.est2_lo = float(lo * lo * (1.0 - 1e-6));
        s.d_hi = float(hi * (1.0 + 1e-5));
        s.d_lo = float(lo * (1.0 - 1e-5));
        // the same decision made exactly, on the float length the reference multiplies (rasterizer.cpp:449)
        const double rs = f->ray_step, mn = f->min_ray_step;
        auto step_of = [&](float L) { return ref_clamp(double(L) * rs, mn, 0.01); };
        const double step_hi = ref_clamp(0.01, mn, 0.01), step_lo = step_of(0.0f);
        s.len_hi_thr = largest_passing([&](float L) { return step_of(L) < step_hi; });
        s.len_lo_thr = largest_passing([&](float L) { return step_of(L) <= step_lo; });
        // Rounding bound for the cheap "certainly alive" test of the run loop. A march inside the unit-scale scene
        // (|p| < 4, length < 8) takes at most 8 / minRayStep samples; each adds at most 3 half-ulps of p's components
        // (2.4e-7 each at |p| < 4) to the position and one half-ulp of the running sum (< 8: 4.8e-7) to the sum.
        // 4e-6 / minRayStep covers both with a factor 2 to spare; 0.004 covers the normalisation of dir and the
        // first step. Scenes outside that scale, or without a lower step bound, always take the exact test.
        {
            bool unit_scale = mn > 0;
            for (int gi = 0; gi < f->n_galaxies; gi++)
                for (int i = 0; i < 3; i++)
                    if (!(std::fabs(gal[gi].axis[i]) <= 2.0f) || !(std::fabs(gal[gi].pos[i] - f->cam[i]) <= 2.0f)) unit_scale = false;
            s.alive_margin = unit_scale ? float(0.004 + 4e-6 / mn) : std::numeric_limits<float>::infinity();
        }
    }

    int nc = 0;
    for (int gi = 0; gi < f->n_galaxies; gi++) {
        const GamerGalaxy &g = gal[gi];
        DGalaxy &d = s.gal[gi];
        if (g.comp_begin < 0 || g.comp_count < 0 || g.comp_begin + g.comp_count > f->n_components)
            return fail("galaxy component range outside the component array");
        // The kernels use reciprocal / division / square-root sequences without the library's slow paths: their operands
        // must be finite, normal and non-zero. And the march must make progress: once ulp(p) exceeds the smallest step,
        // p - dir*step == p and the loop test never turns false (the reference would hang as well; a persistent kernel
        // that never ends wedges the context). |p| < 1024 keeps ulp(p) <= 6.1e-5 < min step for every legal step.
        for (int i = 0; i < 3; i++) {
            if (!std::isfinite(g.axis[i]) || !(std::fabs(g.axis[i]) >= 1e-6f) || !(std::fabs(g.axis[i]) <= 256.0f))
                return fail("galaxy axis must be finite, non-zero and at most 256");
            if (!std::isfinite(g.pos[i]) || !std::isfinite(g.orient[i]) || !std::isfinite(f->cam[i]))
                return fail("galaxy position / orientation and the camera must be finite");
            if (!(std::fabs(g.pos[i] - f->cam[i]) <= 512.0f)) return fail("galaxy farther than 512 from the camera");
        }
        if (!std::isfinite(g.winding_b) || g.winding_b == 0.0 || !std::isfinite(g.winding_n) || !std::isfinite(g.intensity_scale))
            return fail("windingB must be finite and non-zero, windingN and intensityScale finite");
        if (f->min_ray_step < 1e-4) return fail("min_ray_step below 1e-4");
        for (int i = 0; i < 3; i++) {
            d.pos[i] = g.pos[i];
            d.orient[i] = g.orient[i];
            d.inv_axis2[i] = float(1.0 / (g.axis[i] * g.axis[i]));   // util.h:68-70: float product, double divide
        }
        normalize3(g.orient, d.axis_n);
        for (int i = 0; i < 4; i++) { d.rot[i] = g.rotmat[i]; d.arm[i] = float(g.arm[i]); d.arm_d[i] = g.arm[i]; }
        d.rot_identity = (g.rotmat[0] == 1.f && g.rotmat[1] == 0.f && g.rotmat[2] == 0.f && g.rotmat[3] == 0.f);
        d.axis_x = g.axis[0];
        d.orient_y = d.rot_identity && g.orient[0] == 0.f && g.orient[1] == 1.f && g.orient[2] == 0.f &&
                     d.axis_n[0] == 0.f && d.axis_n[1] == 1.f && d.axis_n[2] == 0.f;
        d.inv_winding_b_d = 1.0 / g.winding_b;
        d.inv_axis_x = 1.0f / g.axis[0];
        d.inv_winding_b = float(1.0 / g.winding_b);
        d.winding_2n = float(2.0 * g.winding_n);
        d.winding_b = g.winding_b;
        d.winding_n = g.winding_n;
        d.no_arms = (g.no_arms == 1) ? 1 : (g.no_arms == 2) ? 2 : (g.no_arms == 3) ? 3 : 4;
        d.comp_begin = nc;
        d.slab_gate = -1.0f;
        d.any_arm = 0;
        d.n_bulge = 0;
        for (int k = 0; k < g.comp_count; k++) {
            const GamerComponent &c = comp[g.comp_begin + k];
            if (!(c.active == 1)) continue;                                   // rasterizer.cpp:458
            if (c.class_id < GAMER_CLASS_BULGE || c.class_id > GAMER_CLASS_STARS_SMALL) return fail("unknown class_id");
            if (c.class_id == GAMER_CLASS_STARS_SMALL) { d.noop_count++; continue; }   // no-op, see gamer_b200.h
            for (double v : {c.strength, c.arm, c.z0, c.r0, c.inner, c.delta, c.winding, c.scale, c.noise_offset, c.noise_tilt, c.ks})
                if (!std::isfinite(v)) return fail("component parameters must be finite");
            if (c.class_id != GAMER_CLASS_BULGE && !(std::fabs(c.z0) >= 1e-9)) return fail("component z0 must be non-zero");
            if (std::fabs(c.scale) >= 65536.0) return fail("noise scale too large (|scale| must stay below 65536)");
            if (nc >= DEV_MAX_COMPONENTS) return fail("too many active components (max 64 per frame)");
            DComp &q = s.comp[nc++];
            q.cls = c.class_id;
            q.flags = 0;
            if (c.arm != 0) { q.flags |= DC_HAS_ARM; d.any_arm = 1; }
            if (c.winding != 0) q.flags |= DC_HAS_WINDING;
            if (c.noise_offset != 0) q.flags |= DC_HAS_NOFF;
            const double inner = 1.0f * c.inner;
            if (inner == 0.0) q.flags |= DC_INNER_ZERO;
            if (c.noise_tilt == std::floor(c.noise_tilt) && std::fabs(c.noise_tilt) < 1e9) {
                q.flags |= DC_TILT_INT;
                if (std::fmod(std::fabs(c.noise_tilt), 2.0) == 1.0) q.flags |= DC_TILT_ODD;
            }
            q.iscale = float(g.intensity_scale);
            q.strength = (c.class_id == GAMER_CLASS_BULGE) ? float(c.strength) : float(c.strength * g.intensity_scale);
            q.arm15 = float(c.arm * 15);
            q.inv_z0 = float(1.0 / std::fabs(c.z0));
            q.inv_inner = float(1.0 / inner);
            q.delta = float(c.delta);
            q.delta_d = c.delta;
            q.winding = float(c.winding);
            q.winding_d = c.winding;
            q.noise_offset = float(c.noise_offset);
            q.noise_tilt = float(c.noise_tilt);
            q.rscale = float(c.scale);
            q.r0 = float(c.r0);
            q.hz2 = float(2.0 * 1.4426950408889634 / std::fabs(c.z0));
            q.strength4 = 4.0f * q.strength;
            q.inner_k = float(1.0 / (double(g.axis[0]) * inner));
            q.bk = q.strength * q.iscale * float(f->ray_step);
            for (int i = 0; i < 3; i++) q.spec[i] = c.spec[i];

            // slab gate (galaxycomponent.h:61-64): h = |height/z0|; h > 2 -> 0
            const double z0 = c.z0;
            q.h_gate = largest_passing([&](float a) { double h = std::fabs(double(a) / z0); return !(h > 2.0) && h == h; });
            if (c.class_id == GAMER_CLASS_BULGE) {
                if (d.n_bulge < 2) d.bulge_idx[d.n_bulge] = nc - 1;
                d.n_bulge++;
            } else if (q.h_gate > d.slab_gate) d.slab_gate = q.h_gate;
            // radial gate (galaxycomponent.h:72-81, galaxycomponent.cpp:58-65)
            const float ax = g.axis[0];
            const double r0h = c.r0 * 0.5f;
            const double radial_floor = (f->flags & GAMER_FRAME_NO_RADIAL_GATE) ? 0.0 : 0.001;
            auto radial = [&](float plen) {
                float radius = plen / ax;
                double r = std::exp(-double(radius) / r0h);
                double in = ref_clamp(r - 0.01f, 0, 1);
                if (in > 0.1) in = 0.1;
                return in > radial_floor;   // 0.001 as written; 0 when the test is compiled away (gamer_b200.h)
            };
            if (c.r0 > 0 && ax > 0) q.plen_thr = largest_passing(radial);
            else q.plen_thr = radial(1.0f) ? std::numeric_limits<float>::infinity() : -1.0f;
            if (std::isinf(q.plen_thr)) { q.plen_lo = q.plen_hi = q.plen_thr; }
            else if (q.plen_thr < 0) { q.plen_lo = q.plen_hi = -1.0f; }
            else { q.plen_lo = q.plen_thr * (1.0f - 2e-6f); q.plen_hi = q.plen_thr * (1.0f + 2e-6f); }
            q.k_radial = float(-1.4426950408889634 / (r0h * double(ax)));

            // noise set-up
            int oct = 0;
            q.nscale = 0.0;
            if (c.class_id == GAMER_CLASS_DISK) { oct = 10; q.nscale = c.scale * 0.1f; }      // galaxycomponent.h:96
            if (c.class_id == GAMER_CLASS_DUST) { oct = 9;  q.nscale = c.scale * 0.1f; }
            if (c.class_id == GAMER_CLASS_STARS) { oct = 10; q.nscale = 0.01f * c.scale * 100; } // galaxycomponents.cpp:141
            q.pers = float(c.ks);
            q.octaves = oct;
            double amp = 1, maxamp = 0;
            for (int o = 0; o < oct; o++) { maxamp += amp; amp *= c.ks; }
            q.inv_maxamp = oct ? float(1.0 / maxamp) : 0.f;
            if (std::fabs(q.nscale) >= 65536.0) return fail("noise scale too large (|scale| must stay below 65536)");
            double freq = c.ks;                                                        // noise.cpp:119-124
            const double w = -0.05f;
            for (int o = 0; o < DEV_RIDGED_OCTAVES; o++) { q.rw[o] = float(std::pow(freq, w)); freq *= 2.5; }
            q.rwq = float(std::pow(2.5, w));
        }
        d.comp_count = nc - d.comp_begin;
        d.wt_n = 0;
        if (wtab_host && d.any_arm) gk_build_winding_table(g, d, *wtab_host);
        // twirl sharing: every class but bulge and stars evaluates its noise at twirl(p, winding) with
        // winding = getWinding(r) * cp.winding (0 without arm or winding): components with the same factor get the same
        // point, so the kernel keeps the last twirled point of a sample and the group it belongs to
        d.any_tshare = 0;
        for (int a = 0; a < d.comp_count; a++) {
            DComp &qa = s.comp[d.comp_begin + a];
            qa.tgroup = -1;
            if (qa.cls == GAMER_CLASS_BULGE || qa.cls >= GAMER_CLASS_STARS) continue;
            const bool wa = (qa.flags & DC_HAS_ARM) && (qa.flags & DC_HAS_WINDING);
            int first = -1, members = 0;
            for (int b = 0; b < d.comp_count; b++) {
                const DComp &qb = s.comp[d.comp_begin + b];
                if (qb.cls == GAMER_CLASS_BULGE || qb.cls >= GAMER_CLASS_STARS) continue;
                const bool wb = (qb.flags & DC_HAS_ARM) && (qb.flags & DC_HAS_WINDING);
                if (wa != wb || (wa && qa.winding_d != qb.winding_d)) continue;
                if (first < 0) first = b;
                members++;
            }
            if (members >= 2) { qa.tgroup = first; d.any_tshare = 1; }
        }
        d.fast_ok = d.n_bulge <= 2;
        // the bulge-only path also skips Util::floor (util.h:62-64): a bulge adds max(k v, 0) * spectrum, which
        // cannot take a non-negative I below zero unless a spectrum component is negative (or NaN)
        for (int b = 0; b < d.n_bulge && b < 2; b++)
            for (int ch = 0; ch < 3; ch++)
                if (!(s.comp[d.bulge_idx[b]].spec[ch] >= 0.0f)) d.fast_ok = 0;
    }
    return GAMER_OK;
}
