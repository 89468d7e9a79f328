// scene_dev.h — the scene as the kernels see it: one POD passed as a __grid_constant__ kernel
// parameter (constant bank, warp-uniform reads), built on the host from the ABI structs.
//
// What is precomputed here, in double exactly as the reference would evaluate it, so that the
// kernel's discontinuous decisions agree with the CPU renderer bit for bit:
//   * h_gate     — getHeightModulation's `h > 2` cut (galaxycomponent.h:61-64) as a threshold on the
//                  float |dott| the kernel computes with the same unfused float ops as QVector3D;
//   * plen_thr   — calculateIntensity's `intensity > 0.001` cut (galaxycomponent.cpp:62-65) as a
//                  threshold on the float |P| (getRadius, galaxycomponent.h:72-76);
// everything else is the component's parameters narrowed to float once.
#pragma once
#include <stdint.h>

#define DEV_MAX_GALAXIES   8
#define DEV_MAX_COMPONENTS 64
#define DEV_RIDGED_OCTAVES 9
// getWinding as a table (scene_build.cpp, winding_table): per galaxy up to WT_MAX_INTERVALS intervals of width WT_H in
// s = 1 / (radius + 0.05), a degree-9 polynomial (WT_COEFFS doubles, monomials in x in [-1, 1]) on each
#define WT_MAX_INTERVALS 96
#define WT_COEFFS 10
#define WT_H 0.25
#define WT_DOUBLES_PER_GALAXY (WT_MAX_INTERVALS * WT_COEFFS)

// flags
#define DC_HAS_ARM      1    // cp.arm != 0                (galaxycomponent.cpp:69)
#define DC_HAS_WINDING  2    // cp.winding != 0            (galaxycomponent.cpp:72)
#define DC_HAS_NOFF     4    // cp.noiseOffset != 0        (galaxycomponents.cpp:144)
#define DC_INNER_ZERO   8    // smoothstep(0, 0, r) -> 1   (util.cpp:114-120)
#define DC_TILT_INT     16   // noiseTilt is an integer (pow of a negative base is defined)
#define DC_TILT_ODD     32

struct DComp {
    int32_t cls;
    int32_t flags;
    float   strength;        // cp.strength (bulge) or cp.strength * gi.intensityScale (others)
    float   iscale;          // gi.intensityScale
    float   arm15;           // cp.arm * 15
    float   h_gate;          // pass the slab gate iff |dott| <= h_gate  (exact)
    float   inv_z0;
    float   plen_thr;        // pass the radial gate iff |P| <= plen_thr (exact)
    float   plen_lo, plen_hi;// outside [lo,hi] the fast float |P| decides; inside, the exact one
    float   k_radial;        // exp(-radius/(r0*0.5)) = exp2(|P| * k_radial)
    float   inv_inner;       // 1 / cp.inner
    float   delta;
    float   winding;
    float   pers;            // cp.ks used as persistence
    float   inv_maxamp;      // 1 / sum_k pers^k
    float   noise_offset;
    float   noise_tilt;
    float   rscale;          // float(cp.scale): getRidgedMf input scale (QVector3D * double)
    float   r0;
    float   spec[3];
    float   rw[DEV_RIDGED_OCTAVES];   // pow(ks * 2.5^o, -0.05f)  (noise.cpp:119); the kernel uses rw[0] and the ratio below
    float   rwq;                      // pow(2.5, -0.05f): rw[o] = rw[0] * rwq^o
    int32_t octaves;         // fBm octave count of the class (10 disk/stars, 9 dust; 0 otherwise)
    int32_t tgroup;          // twirl sharing: first component of this galaxy with the same twirl angle, -1 = nobody shares it
    double  nscale;          // fBm start frequency, as a double the way the reference forms it
    double  winding_d;       // cp.winding as the double the reference multiplies getWinding() with
    // value-arithmetic constants of the gate code (kernels.cu, eval_component)
    float   hz2;             // 2 log2(e) / |z0|: sech^2(h) = 4 E / (1 + E)^2 with E = exp(-2h) = exp2(-|dott| * hz2)
    float   strength4;       // 4 * strength (the 4 of sech^2)
    float   inner_k;         // 1 / (axis.x * inner): smoothstep argument = |P| * inner_k
    float   bk;              // bulge: strength * iscale * float(rayStep), the factor of `weight` in its emission
    double  delta_d;         // cp.delta as the double the reference adds in getTheta (negative arm exponents only)
};

struct DGalaxy {
    float   pos[3];
    float   orient[3];       // as given (getRadius uses it un-normalised)
    float   axis_n[3];       // normalised orientation (twirl axis, QQuaternion::fromAxisAndAngle)
    float   rot[4];          // rotmat w,x,y,z
    float   inv_axis2[3];    // float(1.0/(a*a)) per axis  (util.h:68-70)
    float   inv_axis_x;      // 1 / axis.x
    float   axis_x;
    float   inv_winding_b;
    float   winding_2n;      // 2 * windingN
    float   arm[4];
    int32_t no_arms;         // 1,2,3 or 4 (anything else means 4: galaxycomponent.h:120-137)
    int32_t rot_identity;
    int32_t comp_begin, comp_count;   // ACTIVE components only, file order
    int32_t n_bulge;                  // active bulges; the first two are also listed for the off-slab fast path
    int32_t bulge_idx[2];
    int32_t fast_ok;                  // n_bulge <= 2: off-slab samples may take the bulge-only path
    float   slab_gate;                // max h_gate over the non-bulge components (-1 if none)
    int32_t noop_count;               // active "stars small" components: visited by the reference, no-op here
    int32_t any_arm;                  // some active component has arm != 0 (getWinding / getTheta are then needed)
    int32_t any_tshare;               // some components share a twirl angle (DComp::tgroup >= 0)
    double  winding_b, winding_n;     // getWinding is evaluated in double (its result feeds the twirl angle)
    double  inv_winding_b_d;          // 1 / windingB (double)
    double  arm_d[4];                 // arm1..arm4 as the reference's doubles (negative arm exponents only)
    double  wt_s0;                    // winding table: s of the first interval's left end
    int32_t wt_n;                     // intervals (0: no table for this galaxy, getWinding is evaluated directly)
    int32_t wt_off;                   // offset of this galaxy's coefficients in DScene::wtab (doubles)
    int32_t orient_y;                 // orientation and twirl axis are exactly (0,1,0) and rotmat is the identity:
                                      // dot(p, orient) = p.y, P = (p.x, 0, p.z), twirl_y (qtf.cuh) — every shipped galaxy
    int32_t _padg;
};

struct DScene {
    int32_t width, height, full_size, x0, y0, band_rows, band_stride;
    int32_t n_galaxies;
    int32_t flags;           // GAMER_FRAME_*
    int32_t _padf;
    float   inv_vp[16];
    float   cam[3];
    float   ray_step_f;      // float(rayStep)
    double  ray_step;
    double  min_ray_step;
    float   final_mul;       // float(0.01)
    float   est2_hi, est2_lo; // |p-cam|^2 above / below which the step is certainly clamped to 0.01 / minRayStep
    float   len_hi_thr;      // exact: the step is the upper clamp iff |p-cam| (float, as QVector3D::length) > len_hi_thr
    float   len_lo_thr;      // exact: ... the lower clamp iff |p-cam| <= len_lo_thr (tested second)
    float   alive_margin;    // bound on |(-dot(p - origin, dir)) - sum of steps| over a whole march (+inf: always test exactly)
    float   d_hi, d_lo;      // |p - cam| above d_hi: the step is certainly 0.01; below d_lo: certainly minRayStep (sqrt of est2_*)
    const double *wtab;      // device memory: the winding tables of this frame's galaxies (api.cu uploads them), or null
    DGalaxy gal[DEV_MAX_GALAXIES];
    DComp   comp[DEV_MAX_COMPONENTS];
};
