// noise.cuh — 3-D simplex noise, fBm and ridged multifractal for sm_100a.
//
// Replaces source/noise/simplexnoise.cpp:243-333 (raw_noise_3d), noise.cpp:162-180
// (Noise::octave_noise_3d) and noise.cpp:81-128 (Noise::getRidgedMf).
//
// Design (not a translation of the reference's double-precision table walk):
//  * Lattice selection is integer work on a 32.32 fixed-point image of the SKEWED coordinate
//    u = x + (x+y+z)/3.  u is formed once per fBm call in FP64 with exactly the reference's
//    operations, so floor() agrees with the CPU renderer bit for bit; because every octave doubles
//    the frequency and scaling by 2 commutes with IEEE rounding, octave k's skewed coordinate is
//    exactly 2^k * u — a 64-bit left shift.  Integer cell = high word, in-cell offset = low word.
//  * In-cell offsets are the fractional parts (fu,fv,fw) in [0,1); the unskewed corner vectors
//    follow from x0 = fu - (fu+fv+fw)/6, all FP32 on small numbers, so FP32 loses nothing even at
//    coordinates ~1e5 where a float product would have no fractional bits left.
//  * The permutation table lives in shared memory replicated once per bank (257 entries x 32
//    lanes x 4 B = 32 KB): lane l only ever reads bank l, so the data-dependent hash walk is
//    bank-conflict free.  An entry is perm[e] << 7, i.e. the value is already the byte offset of the
//    next lookup: one IADD + one LOP3 + one LDS per level.  The last level reads a second 32 KB table
//    whose rows hold the gradient vector itself (one conflict-free 128-bit load per corner).
//  * The reference's fastfloor quirk (simplexnoise.h:130: exact non-positive integers floor one
//    too low) is reproduced on a rare path.
#pragma once
#include <stdint.h>

#ifndef FBM_UNROLL
#define FBM_UNROLL 1
#endif
#define GN_PRAGMA(x) _Pragma(#x)
#define GN_UNROLL(n) GN_PRAGMA(unroll n)

namespace gnoise {

// Ken Perlin's permutation, as in simplexnoise.h:160-185 (first 256 entries; the second copy is &255).
__constant__ uint8_t c_perm[256] = {
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

// Gradient code of g = perm % 12 for grad3 (simplexnoise.h:142-146): dot(grad3[g], (x,y,z)) = ±a ± b with
//   g<4: (a,b)=(x,y)   4<=g<8: (x,z)   g>=8: (y,z);  sign(a) = g&1, sign(b) = (g>>1)&1.
// ±a ± b = sa (a + (sa sb) b): bit 31 = sign(a) XOR sign(b) (flips b with one LOP3), bit 19 = sign(a) (applied to the
// sum through a predicate), bit 17 = "a is y", bit 18 = "b is y"; bits 20..23 = g itself (probe only). The low
// 7 bits of an entry stay ZERO: an index operand may then carry 7 garbage bits below the <<7 position (the
// fraction bits a funnel shift brings along) without a carry into the index.
__host__ __device__ inline uint32_t grad_code(uint32_t g) {
    uint32_t code = (g << 20);
    const uint32_t sa = g & 1u, sb = (g >> 1) & 1u;
    if (sa ^ sb) code |= 0x80000000u;
    if (sa) code |= 0x80000u;
    if (g >= 8) code |= 0x20000u;
    if (g < 4) code |= 0x40000u;
    return code;
}

// 257 entries (entry 256 repeats entry 0, like perm[256] of the reference's doubled table): the "+1" of a
// corner offset is then a +128-byte displacement AFTER the index has been wrapped, i.e. an immediate of the
// load instead of index arithmetic.
#define GN_TABLE_ENTRIES 257
#define GN_TABLE_WORDS (GN_TABLE_ENTRIES * 32)
#define GN_ADDR_MASK 0x7F80u
// Two tables in dynamic shared memory, both with a 128-byte row per entry so that one masked index serves both:
//   hash table      row e = 32 x u32, one per lane (lane l reads bank l only: conflict free), value perm[e] << 7
//                   (| the gradient code above, used by the probe) — the first two levels of the hash walk;
//   gradient table  row e = 8 x float4 (gx, gy, gz, g as bits) of grad3[perm[e] % 12], one per lane of a
//                   quarter warp (a 128-bit shared load is served one quarter warp at a time, and its 8 lanes
//                   read 8 different 16-byte slots: conflict free) — the last level: ONE 128-bit load returns the
//                   gradient itself, so the dot product is three FMA-pipe instructions instead of bit tests,
//                   selects and sign logic on the ALU pipe (the kernel is issue/ALU bound, the LSU is not).
// Static shared memory starts at window offset 0x400 (the first KB is reserved); kernels that use the noise code
// take GN_DYN_SMEM bytes of dynamic shared memory and no static shared memory, so the tables sit at fixed
// window addresses and the base is an immediate of the load. table_init() verifies the base and traps otherwise.
#define GN_DYN_BASE 0x400u
#define GN_HASH_ADDR GN_DYN_BASE
#define GN_GRAD_ADDR (GN_DYN_BASE + GN_TABLE_WORDS * 4u)
#define GN_DYN_SMEM (2u * GN_TABLE_WORDS * 4u)
// the same two addresses as plain decimal literals for the inline PTX (checked below)
#define GN_HASH_IMM 1024
#define GN_GRAD_IMM 33920
#define GN_STR2(x) #x
#define GN_STR(x) GN_STR2(x)
static_assert(GN_HASH_IMM == GN_HASH_ADDR && GN_GRAD_IMM == GN_GRAD_ADDR, "PTX immediates out of step with the table layout");

extern __shared__ __align__(16) unsigned char gn_dyn_smem[];

// grad3 of simplexnoise.h:142-146
__host__ __device__ inline void grad3_of(uint32_t g, float &gx, float &gy, float &gz) {
    const float sa = (g & 1u) ? -1.0f : 1.0f, sb = (g & 2u) ? -1.0f : 1.0f;
    if (g < 4) { gx = sa; gy = sb; gz = 0.0f; }
    else if (g < 8) { gx = sa; gy = 0.0f; gz = sb; }
    else { gx = 0.0f; gy = sa; gz = sb; }
}

// Fill both tables; call from all threads of the block, then __syncthreads().
__device__ __forceinline__ void table_init() {
    if ((uint32_t)__cvta_generic_to_shared(gn_dyn_smem) != GN_DYN_BASE) __trap();
    uint32_t *tbl = reinterpret_cast<uint32_t *>(gn_dyn_smem);
    uint32_t *grd = tbl + GN_TABLE_WORDS;
    for (uint32_t w = threadIdx.x + threadIdx.y * blockDim.x; w < GN_TABLE_WORDS; w += blockDim.x * blockDim.y) {
        const uint32_t p = c_perm[(w >> 5) & 255u], g = p % 12u;
        tbl[w] = (p << 7) | grad_code(g);
        float gx, gy, gz;
        grad3_of(g, gx, gy, gz);
        const uint32_t c = w & 3u;
        grd[w] = c == 0 ? __float_as_uint(gx) : c == 1 ? __float_as_uint(gy) : c == 2 ? __float_as_uint(gz) : g;
    }
}
// per-lane constant of the lookup address: the lane's bank in a hash-table row
__device__ __forceinline__ uint32_t lane_const() { return (threadIdx.x & 31u) << 2; }

// one level of the hash walk: row offset of entry ((off7 >> 7) & 255) plus this lane's slot in the row ...
__device__ __forceinline__ uint32_t addr_of(uint32_t lane4, uint32_t off7) { return (off7 & GN_ADDR_MASK) | lane4; }
// ... and the hash entry at that address, or the one after it
__device__ __forceinline__ uint32_t ld0(uint32_t a) {
    uint32_t v; asm("ld.shared.u32 %0, [%1+" GN_STR(GN_HASH_IMM) "];" : "=r"(v) : "r"(a)); return v;
}
__device__ __forceinline__ uint32_t ld1(uint32_t a) {
    uint32_t v; asm("ld.shared.u32 %0, [%1+" GN_STR(GN_HASH_IMM) "+128];" : "=r"(v) : "r"(a)); return v;
}
// ... or the gradient at that address / the one after it (address built with the lane's 16-byte slot)
struct Grad { float x, y, z; uint32_t g; };
__device__ __forceinline__ Grad gd0(uint32_t a) {
    Grad v; asm("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4+" GN_STR(GN_GRAD_IMM) "];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=r"(v.g) : "r"(a)); return v;
}
__device__ __forceinline__ Grad gd1(uint32_t a) {
    Grad v; asm("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4+" GN_STR(GN_GRAD_IMM) "+128];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=r"(v.g) : "r"(a)); return v;
}
__device__ __forceinline__ float grad_dot(const Grad &e, float x, float y, float z) {
    return fmaf(e.z, z, fmaf(e.y, y, e.x * x));
}

struct Lattice {   // filled only by the probe path
    int i, j, k, c1, c2, g0, g1, g2, g3;
};

// Skewed coordinate in 32.32 fixed point; d keeps the double for the rare exact-integer test.
struct Skewed {
    uint32_t lo;
    int32_t hi;
    double d;
};
__device__ __forceinline__ Skewed to_fixed(double u) {
    Skewed s;
    long long f = __double2ll_rd(u * 4294967296.0);
    s.lo = (uint32_t)f;
    s.hi = (int32_t)(f >> 32);
    s.d = u;
    return s;
}

// integer cell and in-cell fraction of octave k (skewed coordinate 2^k * u); the fraction is returned
// BIASED BY +1 (simplex_cell removes the bias inside an FMA it needs anyway). CHECK: also handle the
// reference's fastfloor quirk; callers hoist the (almost never true) need for it out of their loops.
template <bool CHECK>
__device__ __forceinline__ void cell_of(const Skewed &s, int k, uint32_t &cell7, float &frac) {
    uint32_t lo = s.lo << k;
    // (cell << 7) | 7 fraction bits: the upper word of the 64-bit value shifted by k + 7. The 7 low bits are
    // harmless (table entries have zeros there and the address mask drops them).
    cell7 = __funnelshift_l(s.lo, (uint32_t)s.hi, k + 7);
    frac = __uint_as_float((lo >> 9) | 0x3F800000u);   // 1 + fraction (top 23 bits): no int->float conversion
    if (CHECK && lo == 0u) {   // rare: looks like an exact integer in fixed point
        double uk = s.d * (double)(1u << k);
        if (uk <= 0.0 && uk == rint(uk)) {   // simplexnoise.h:130 fastfloor: (int)x - 1 for exact x <= 0
            cell7 -= 128u;
            frac = 2.0f;
        }
    }
}

// One raw simplex value from cells (ci,cj,ck) and fractions (fu,fv,fw).  Returns 32*(n0+n1+n2+n3).
template <bool PROBE>
__device__ __forceinline__ float simplex_cell(uint32_t lane4, uint32_t I7, uint32_t J7, uint32_t K7,
                                              float fu, float fv, float fw, Lattice *lat) {
    const float G3 = 1.0f / 6.0f;
    float t = fmaf(fu + fv + fw, G3, 0.5f);      // fu,fv,fw carry a +1 bias: (sum-3)/6 + 1
    float x0 = fu - t, y0 = fv - t, z0 = fw - t;

    // the six orderings of simplexnoise.cpp:267-276 as three comparisons
    bool a = x0 >= y0, b = y0 >= z0, c = x0 >= z0;
    bool i1 = a && c, j1 = !a && b, k1 = !b && !c;
    bool i2 = a || c, j2 = !a || b, k2 = !(b && c);

    const float G3m = G3 - 1.0f, G6 = 2.0f * G3, G6m = 2.0f * G3 - 1.0f;
    float x1 = x0 + (i1 ? G3m : G3), y1 = y0 + (j1 ? G3m : G3), z1 = z0 + (k1 ? G3m : G3);
    float x2 = x0 + (i2 ? G6m : G6), y2 = y0 + (j2 ? G6m : G6), z2 = z0 + (k2 ? G6m : G6);
    float x3 = x0 - 0.5f, y3 = y0 - 0.5f, z3 = z0 - 0.5f;

    // hash walk: perm[ii+oi+perm[jj+oj+perm[kk+ok]]], offsets pre-scaled by 128. An address is the wrapped index
    // of the corner WITHOUT its +1 offsets; the offsets pick the entry at the address or the one after it.
    const uint32_t aK = addr_of(lane4, K7);
    const uint32_t A0 = ld0(aK), A1 = ld1(aK);
    const uint32_t aJ0 = addr_of(lane4, A0 + J7), aJ1 = addr_of(lane4, A1 + J7);
    const uint32_t aB1 = k1 ? aJ1 : aJ0, aB2 = k2 ? aJ1 : aJ0;
    const uint32_t B0 = ld0(aJ0);
    const uint32_t B1 = j1 ? ld1(aB1) : ld0(aB1);
    const uint32_t B2 = j2 ? ld1(aB2) : ld0(aB2);
    const uint32_t B3 = ld1(aJ1);
    const uint32_t lane16 = (lane4 << 2) & 0x70u;      // (lane % 8) * 16: this lane's slot in a gradient row
    const uint32_t aC1 = addr_of(lane16, B1 + I7), aC2 = addr_of(lane16, B2 + I7);
    const Grad C0 = gd0(addr_of(lane16, B0 + I7));
    const Grad C1 = i1 ? gd1(aC1) : gd0(aC1);
    const Grad C2 = i2 ? gd1(aC2) : gd0(aC2);
    const Grad C3 = gd1(addr_of(lane16, B3 + I7));

    float t0 = fmaxf(0.6f - x0 * x0 - y0 * y0 - z0 * z0, 0.0f);
    float t1 = fmaxf(0.6f - x1 * x1 - y1 * y1 - z1 * z1, 0.0f);
    float t2 = fmaxf(0.6f - x2 * x2 - y2 * y2 - z2 * z2, 0.0f);
    float t3 = fmaxf(0.6f - x3 * x3 - y3 * y3 - z3 * z3, 0.0f);
    t0 *= t0; t1 *= t1; t2 *= t2; t3 *= t3;
    float n = t0 * t0 * grad_dot(C0, x0, y0, z0);
    n = fmaf(t1 * t1, grad_dot(C1, x1, y1, z1), n);
    n = fmaf(t2 * t2, grad_dot(C2, x2, y2, z2), n);
    n = fmaf(t3 * t3, grad_dot(C3, x3, y3, z3), n);

    if (PROBE) {
        lat->c1 = (int)i1 + 2 * (int)j1 + 4 * (int)k1;
        lat->c2 = (int)i2 + 2 * (int)j2 + 4 * (int)k2;
        lat->g0 = (int)C0.g; lat->g1 = (int)C1.g; lat->g2 = (int)C2.g; lat->g3 = (int)C3.g;
    }
    return n;        // 1/32 of the raw value: callers fold the factor into their own scale
}

// skewed coordinates of the point (x,y,z) (doubles), reference operation order (simplexnoise.cpp:247-251)
__device__ __forceinline__ void skew(double x, double y, double z, Skewed &su, Skewed &sv, Skewed &sw) {
    const double F3 = 1.0 / 3.0;
    double s = __dmul_rn(__dadd_rn(__dadd_rn(x, y), z), F3);
    su = to_fixed(__dadd_rn(x, s));
    sv = to_fixed(__dadd_rn(y, s));
    sw = to_fixed(__dadd_rn(z, s));
}

// one octave: cells/fractions of (2^k u, 2^k v, 2^k w) and the raw simplex value. QUIRK selects the
// (almost never needed) variant that checks for the reference's fastfloor behaviour at exact integers.
template <bool QUIRK>
__device__ __forceinline__ float octave(uint32_t lane4, const Skewed &su, const Skewed &sv, const Skewed &sw, int k) {
    uint32_t ci, cj, ck;
    float fu, fv, fw;
    cell_of<QUIRK>(su, k, ci, fu);
    cell_of<QUIRK>(sv, k, cj, fv);
    cell_of<QUIRK>(sw, k, ck, fw);
    return simplex_cell<false>(lane4, ci, cj, ck, fu, fv, fw, nullptr);
}

// can an octave k <= kmax of this point hit an exact integer in fixed point? (needs the low 32-kmax bits zero)
__device__ __forceinline__ bool maybe_integer(const Skewed &su, const Skewed &sv, const Skewed &sw, int kmax) {
    const uint32_t m = 0xFFFFFFFFu >> kmax;
    return ((su.lo & m) == 0u) | ((sv.lo & m) == 0u) | ((sw.lo & m) == 0u);
}

template <bool QUIRK>
__device__ __forceinline__ float fbm_sum(uint32_t lane4, const Skewed &su, const Skewed &sv, const Skewed &sw,
                                         int octaves, float pers) {
    float total = 0.0f, amp = 1.0f;
GN_UNROLL(FBM_UNROLL)
    for (int k = 0; k < octaves; k++) {
        total = fmaf(octave<QUIRK>(lane4, su, sv, sw, k), amp, total);
        amp *= pers;
    }
    return total;
}
// out of line: points on the exact-integer planes of the skewed lattice (fastfloor quirk, simplexnoise.h:130)
__device__ __noinline__ float fbm_quirk(uint32_t lane4, double x, double y, double z, int octaves, float pers) {
    Skewed su, sv, sw;
    skew(x, y, z, su, sv, sw);
    return fbm_sum<true>(lane4, su, sv, sw, octaves, pers);
}
__device__ __noinline__ float octave_quirk(uint32_t lane4, float vx, float vy, float vz) {
    Skewed su, sv, sw;
    skew((double)vx, (double)vy, (double)vz, su, sv, sw);
    return octave<true>(lane4, su, sv, sw, 0);
}

// Noise::octave_noise_3d (noise.cpp:162-180): sum_k raw(x f_k) a_k / sum a_k, f_k = scale 2^k, a_k = pers^k.
// (px,py,pz) are the float coordinates the reference widens to double; scale is the double frequency.
__device__ __forceinline__ float fbm(uint32_t lane4, float px, float py, float pz, double scale,
                                     int octaves, float pers, float inv_maxamp) {
    const double x = __dmul_rn((double)px, scale), y = __dmul_rn((double)py, scale), z = __dmul_rn((double)pz, scale);
    Skewed su, sv, sw;
    skew(x, y, z, su, sv, sw);
    float total;
    if (maybe_integer(su, sv, sw, octaves - 1)) total = fbm_quirk(lane4, x, y, z, octaves, pers);
    else total = fbm_sum<false>(lane4, su, sv, sw, octaves, pers);
    return total * (32.0f * inv_maxamp);
}

// Noise::getRidgedMf (noise.cpp:81-128). v is the float QVector3D (already r*scale); it is scaled by
// float(lacunarity) per octave exactly like `vt = vt * lacunarity`. rw[o] = pow(frequency_o, -0.05f).
__device__ __forceinline__ float ridged(uint32_t lane4, float vx, float vy, float vz, int octaves,
                                        float lac, float offset, float gain, const float *rw) {
    float value = 0.0f, weight = 1.0f;
#pragma unroll 1
    for (int o = 0; o < octaves; o++) {
        Skewed su, sv, sw;
        skew((double)vx, (double)vy, (double)vz, su, sv, sw);
        float signal;
        if (maybe_integer(su, sv, sw, 0)) signal = octave_quirk(lane4, vx, vy, vz);
        else signal = octave<false>(lane4, su, sv, sw, 0);
        signal = offset - fabsf(32.0f * signal);
        signal *= signal;
        signal *= weight;
        weight = fminf(fmaxf(signal * gain, 0.0f), 1.0f);
        value = fmaf(signal, rw[o], value);
        vx = __fmul_rn(vx, lac); vy = __fmul_rn(vy, lac); vz = __fmul_rn(vz, lac);
    }
    return fmaf(value, 1.25f, -1.0f);
}

}  // namespace gnoise
