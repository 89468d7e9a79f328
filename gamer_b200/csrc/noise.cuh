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
//  * The permutation table lives in shared memory replicated once per bank (512 entries x 32
//    lanes x 4 B = 64 KB): lane l only ever reads bank l, so the data-dependent hash walk is
//    bank-conflict free.  An entry is perm[e & 255] << 7, i.e. the value is already the byte offset of
//    the next lookup: one IADD + one LDS per level.  The last level reads a second 64 KB table
//    whose rows hold the gradient vector itself (one conflict-free 128-bit load per corner).
//  * The reference's fastfloor quirk (simplexnoise.h:130: exact non-positive integers floor one
//    too low) is reproduced on a rare path.
#pragma once
#include <stdint.h>

#ifndef FBM_UNROLL
#define FBM_UNROLL 1
#endif
#ifndef GN_NOISE_LINKAGE
#define GN_NOISE_LINKAGE __noinline__
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

// 512 entries (entry e holds perm[e & 255], the reference's own doubled table, simplexnoise.h:160-185): an index
// "cell + perm[...]" (< 511) then needs no wrap, so one level of the hash walk is ONE add - the masked cell offset
// with the lane's slot already OR-ed in is prepared once per coordinate and octave - and a corner's "+1" is a
// +128-byte immediate of the load.
#define GN_TABLE_ENTRIES 512
#define GN_TABLE_WORDS (GN_TABLE_ENTRIES * 32)
#define GN_ADDR_MASK 0x7F80u
// Two tables in dynamic shared memory, both with a 128-byte row per entry:
//   hash table      row e = 32 x u32, one per lane (lane l reads bank l only: conflict free), value perm[e & 255] << 7,
//                   i.e. already the row offset of the next level - the first two levels of the hash walk;
//   gradient table  row e = 8 x float4 (gx, gy, gz, g as bits) of grad3[perm[e & 255] % 12], one per lane of a
//                   quarter warp (a 128-bit shared load is served one quarter warp at a time, and its 8 lanes
//                   read 8 different 16-byte slots: conflict free) - the last level: ONE 128-bit load returns the
//                   gradient itself, so the dot product is three FMA-pipe instructions instead of bit tests,
//                   selects and sign logic on the ALU pipe (the kernel is issue/ALU bound, the LSU is not).
// 2 x 64 KB: the ray-march kernel runs one persistent CTA of 24 warps per SM, so the tables exist once per SM.
// Static shared memory starts at window offset 0x400 (the first KB is reserved); kernels that use the noise code
// take GN_DYN_SMEM bytes of dynamic shared memory and no static shared memory, so the tables sit at fixed
// window addresses and the base is an immediate of the load. table_init() verifies the base and traps otherwise.
#define GN_DYN_BASE 0x400u
#define GN_HASH_ADDR GN_DYN_BASE
#define GN_GRAD_ADDR (GN_DYN_BASE + GN_TABLE_WORDS * 4u)
#define GN_DYN_SMEM (2u * GN_TABLE_WORDS * 4u)
// the same two addresses as plain decimal literals for the inline PTX (checked below)
#define GN_HASH_IMM 1024
#define GN_GRAD_IMM 66560
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
        tbl[w] = p << 7;
        float gx, gy, gz;
        grad3_of(g, gx, gy, gz);
        const uint32_t c = w & 3u;
        grd[w] = c == 0 ? __float_as_uint(gx) : c == 1 ? __float_as_uint(gy) : c == 2 ? __float_as_uint(gz) : g;
    }
}
// per-lane constant of the lookup address: the lane's bank in a hash-table row
__device__ __forceinline__ uint32_t lane_const() { return (threadIdx.x & 31u) << 2; }
// (lane % 8) * 16: the lane's 16-byte slot in a gradient-table row
__device__ __forceinline__ uint32_t lane16_of(uint32_t lane4) { return (lane4 << 2) & 0x70u; }

// a gradient-table entry: grad3[perm % 12] and the index itself (read by the probes only)
struct Grad { float x, y, z; uint32_t g; };
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

// packed FP32 pairs (FFMA2 / FMUL2 of sm_100a): one issue slot for two lanes' worth of work. ptxas allocates the two
// halves in adjacent registers and folds negations and immediates, so packing and unpacking cost nothing.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float a, float b) { f32x2 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void upk2(f32x2 v, float &a, float &b) { asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }

// One raw simplex value from cells (ci,cj,ck) and fractions (fu,fv,fw).  Returns (n0+n1+n2+n3), i.e. 1/32 of the
// reference's value. The kernel is bound by instruction issue and, inside this function, by the half-rate ALU pipe
// (compare / select / logic / min-max: 2 cycles per warp instruction), so the code is arranged to keep work OFF it:
//   * the six orderings of simplexnoise.cpp:267-276 are eight compare instructions whose boolean combination rides
//     on the compare itself; the corner offsets are then PREDICATED ADDS with immediates (FMA pipe) — no selects;
//   * the hash walk loads all four (j,k) combinations unconditionally and picks corner 1 / corner 2 by three
//     predicated 128-bit gradient loads each (exactly one of them executes) — no address selects;
//   * max(t, 0) is the saturation modifier of the last FMA of t (t <= 0.6 < 1, so the upper clamp never acts);
//   * corners are processed as pairs (0,3) and (1,2) with packed FFMA2 / FMUL2.
#ifdef GN_DIAG_LDS32   // timing experiment only (results are wrong): a quarter of the shared-memory wavefronts per gradient
#define GN_LDG(r0, r1, r2, r3, addr) "ld.shared.b32 " r0 ", [" addr "];\n\tmov.b32 " r1 ", " r0 ";\n\tmov.b32 " r2 ", " r0 ";\n\tmov.b32 " r3 ", " r0 ";\n\t"
#else
#define GN_LDG(r0, r1, r2, r3, addr) "ld.shared.v4.b32 {" r0 "," r1 "," r2 "," r3 "}, [" addr "];\n\t"
#endif
template <bool PROBE>
__device__ __forceinline__ float simplex_cell(uint32_t lane4, uint32_t lane16, uint32_t I7, uint32_t J7, uint32_t K7,
                                              float fu, float fv, float fw, Lattice *lat) {
    // fu,fv,fw carry a +1 bias: t = (sum-3)/6 + 1. Multiply and add with immediates (an FMA would need 1/6 in a
    // register, i.e. one more instruction on the ALU pipe or one more live register)
    float t = __fadd_rn(__fmul_rn(fu + fv + fw, 1.0f / 6.0f), 0.5f);
    float x0 = fu - t, y0 = fv - t, z0 = fw - t;
    float x1, y1, z1, x2, y2, z2;
    Grad C0, C1, C2, C3;
    // hash walk: perm[ii+oi+perm[jj+oj+perm[kk+ok]]], offsets pre-scaled by 128. An address is the wrapped index
    // of the corner WITHOUT its +1 offsets; the offsets pick the entry at the address or the one after it.
    asm("{\n\t"
        ".reg .pred pb, pc, i1, i2, j1, j2, k1, k2;\n\t"
        ".reg .b32 aK, A0, A1, aJ0, aJ1, B00, B01, B10, B11, c00, c01, c10, c11;\n\t"
        "setp.ge.f32 pc, %0, %2;\n\t"
        "setp.ge.f32 pb, %1, %2;\n\t"
        "setp.ge.and.f32 i1, %0, %1, pc;\n\t"
        "setp.ge.or.f32  i2, %0, %1, pc;\n\t"
        "setp.lt.and.f32 j1, %0, %1, pb;\n\t"
        "setp.lt.or.f32  j2, %0, %1, pb;\n\t"
        "setp.lt.and.f32 k1, %1, %2, !pc;\n\t"
        "setp.lt.or.f32  k2, %1, %2, !pc;\n\t"
        // corner offsets (simplexnoise.cpp:283-291): +G3 (-1 along the largest axis), +2*G3 (-1 along the two largest)
        "add.f32 %3, %0, 0f3E2AAAAB;\n\t"
        "add.f32 %4, %1, 0f3E2AAAAB;\n\t"
        "add.f32 %5, %2, 0f3E2AAAAB;\n\t"
        "@i1 add.f32 %3, %0, 0fBF555555;\n\t"
        "@j1 add.f32 %4, %1, 0fBF555555;\n\t"
        "@k1 add.f32 %5, %2, 0fBF555555;\n\t"
        "add.f32 %6, %0, 0f3EAAAAAB;\n\t"
        "add.f32 %7, %1, 0f3EAAAAAB;\n\t"
        "add.f32 %8, %2, 0f3EAAAAAB;\n\t"
        "@i2 add.f32 %6, %0, 0fBF2AAAAA;\n\t"
        "@j2 add.f32 %7, %1, 0fBF2AAAAA;\n\t"
        "@k2 add.f32 %8, %2, 0fBF2AAAAA;\n\t"
        // level k
        "lop3.b32 aK, %27, 0x7F80, %25, 0xEA;\n\t"
        "ld.shared.u32 A0, [aK+" GN_STR(GN_HASH_IMM) "];\n\t"
        "ld.shared.u32 A1, [aK+" GN_STR(GN_HASH_IMM) "+128];\n\t"
        // level j: all four (j, k) offset combinations; %28 = masked j cell offset | lane slot
        "add.u32 aJ0, A0, %28;\n\t"
        "add.u32 aJ1, A1, %28;\n\t"
        "ld.shared.u32 B00, [aJ0+" GN_STR(GN_HASH_IMM) "];\n\t"
        "ld.shared.u32 B10, [aJ0+" GN_STR(GN_HASH_IMM) "+128];\n\t"
        "ld.shared.u32 B01, [aJ1+" GN_STR(GN_HASH_IMM) "];\n\t"
        "ld.shared.u32 B11, [aJ1+" GN_STR(GN_HASH_IMM) "+128];\n\t"
        // level i: gradient rows; %29 = masked i cell offset | 16-byte lane slot
        "add.u32 c00, B00, %29;\n\t"
        "add.u32 c10, B10, %29;\n\t"
        "add.u32 c01, B01, %29;\n\t"
        "add.u32 c11, B11, %29;\n\t"
        GN_LDG("%9", "%10", "%11", "%12", "c00+" GN_STR(GN_GRAD_IMM))
        GN_LDG("%21", "%22", "%23", "%24", "c11+" GN_STR(GN_GRAD_IMM) "+128")
        "@i1 " GN_LDG("%13", "%14", "%15", "%16", "c00+" GN_STR(GN_GRAD_IMM) "+128")
        "@j1 " GN_LDG("%13", "%14", "%15", "%16", "c10+" GN_STR(GN_GRAD_IMM))
        "@k1 " GN_LDG("%13", "%14", "%15", "%16", "c01+" GN_STR(GN_GRAD_IMM))
        "@!i2 " GN_LDG("%17", "%18", "%19", "%20", "c11+" GN_STR(GN_GRAD_IMM))
        "@!j2 " GN_LDG("%17", "%18", "%19", "%20", "c01+" GN_STR(GN_GRAD_IMM) "+128")
        "@!k2 " GN_LDG("%17", "%18", "%19", "%20", "c10+" GN_STR(GN_GRAD_IMM) "+128")
        "}"
        : "+f"(x0), "+f"(y0), "+f"(z0), "=f"(x1), "=f"(y1), "=f"(z1), "=f"(x2), "=f"(y2), "=f"(z2),
          "=f"(C0.x), "=f"(C0.y), "=f"(C0.z), "=r"(C0.g), "=f"(C1.x), "=f"(C1.y), "=f"(C1.z), "=r"(C1.g),
          "=f"(C2.x), "=f"(C2.y), "=f"(C2.z), "=r"(C2.g), "=f"(C3.x), "=f"(C3.y), "=f"(C3.z), "=r"(C3.g)
        : "r"(lane4), "r"(lane16), "r"(K7), "r"((J7 & GN_ADDR_MASK) | lane4), "r"((I7 & GN_ADDR_MASK) | lane16));
    const float x3 = x0 - 0.5f, y3 = y0 - 0.5f, z3 = z0 - 0.5f;

    // t_c = max(0.6 - |x_c|^2, 0) for the corner pairs (0,3) and (1,2), then t^4
    const f32x2 c06 = pk2(0.6f, 0.6f);
    f32x2 ta = fma2(pk2(-x0, -x3), pk2(x0, x3), c06), tb = fma2(pk2(-x1, -x2), pk2(x1, x2), c06);
    ta = fma2(pk2(-y0, -y3), pk2(y0, y3), ta);
    tb = fma2(pk2(-y1, -y2), pk2(y1, y2), tb);
    float t0, t1, t2, t3;
    upk2(ta, t0, t3);
    upk2(tb, t1, t2);
    t0 = __saturatef(fmaf(-z0, z0, t0)); t3 = __saturatef(fmaf(-z3, z3, t3));
    t1 = __saturatef(fmaf(-z1, z1, t1)); t2 = __saturatef(fmaf(-z2, z2, t2));
    ta = pk2(t0, t3); tb = pk2(t1, t2);
    ta = mul2(ta, ta); tb = mul2(tb, tb);
    ta = mul2(ta, ta); tb = mul2(tb, tb);
    f32x2 nn = mul2(ta, pk2(grad_dot(C0, x0, y0, z0), grad_dot(C3, x3, y3, z3)));
    nn = fma2(tb, pk2(grad_dot(C1, x1, y1, z1), grad_dot(C2, x2, y2, z2)), nn);
    float na, nb;
    upk2(nn, na, nb);

    if (PROBE) {
        const bool a = x0 >= y0, b = y0 >= z0, c = x0 >= z0;
        lat->c1 = (int)(a && c) + 2 * (int)(!a && b) + 4 * (int)(!b && !c);
        lat->c2 = (int)(a || c) + 2 * (int)(!a || b) + 4 * (int)(!(b && c));
        lat->g0 = (int)C0.g; lat->g1 = (int)C1.g; lat->g2 = (int)C2.g; lat->g3 = (int)C3.g;
    }
    return na + nb;        // 1/32 of the raw value: callers fold the factor into their own scale
}

// skewed coordinates of the point (x,y,z) (doubles), reference operation order (simplexnoise.cpp:247-251)
__device__ __forceinline__ void skew(double x, double y, double z, Skewed &su, Skewed &sv, Skewed &sw) {
    const double F3 = 1.0 / 3.0;
    double s = __dmul_rn(__dadd_rn(__dadd_rn(x, y), z), F3);
    su = to_fixed(__dadd_rn(x, s));
    sv = to_fixed(__dadd_rn(y, s));
    sw = to_fixed(__dadd_rn(z, s));
}

// The same three fixed-point images from a FLOAT point (the ridged multifractal, whose point is a float QVector3D),
// without the three double -> int64 conversions: x * 2^32 is exact in float and widens exactly; the reference's
// roundings s = RN(RN(RN(x + y) + z) / 3), u = RN(x + s) commute with the power-of-two scaling, so the doubles below are
// exactly 2^32 times the reference's; and floor(u 2^32) is read off the significand after a round-DOWN add of
// 1.5 * 2^52 (low word = the fraction, the next 20 bits = the cell modulo 2^20; |u 2^32| < 2^51). `d` is not set.
__device__ __forceinline__ void skew_float(float x, float y, float z, Skewed &su, Skewed &sv, Skewed &sw) {
    const double X = (double)__fmul_rn(x, 4294967296.0f), Y = (double)__fmul_rn(y, 4294967296.0f),
                 Z = (double)__fmul_rn(z, 4294967296.0f);
    const double s = __dmul_rn(__dadd_rn(__dadd_rn(X, Y), Z), 1.0 / 3.0);
    const double M = 6755399441055744.0;
    const double tu = __dadd_rd(__dadd_rn(X, s), M), tv = __dadd_rd(__dadd_rn(Y, s), M), tw = __dadd_rd(__dadd_rn(Z, s), M);
    su.lo = (uint32_t)__double2loint(tu); su.hi = __double2hiint(tu);
    sv.lo = (uint32_t)__double2loint(tv); sv.hi = __double2hiint(tv);
    sw.lo = (uint32_t)__double2loint(tw); sw.hi = __double2hiint(tw);
}

// one octave: cells/fractions of (2^k u, 2^k v, 2^k w) and the raw simplex value. QUIRK selects the
// (almost never needed) variant that checks for the reference's fastfloor behaviour at exact integers.
template <bool QUIRK>
__device__ __forceinline__ float octave(uint32_t lane4, uint32_t lane16, const Skewed &su, const Skewed &sv, const Skewed &sw, int k) {
    uint32_t ci, cj, ck;
    float fu, fv, fw;
    cell_of<QUIRK>(su, k, ci, fu);
    cell_of<QUIRK>(sv, k, cj, fv);
    cell_of<QUIRK>(sw, k, ck, fw);
    return simplex_cell<false>(lane4, lane16, ci, cj, ck, fu, fv, fw, nullptr);
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
    const uint32_t lane16 = lane16_of(lane4);
GN_UNROLL(FBM_UNROLL)
    for (int k = 0; k < octaves; k++) {
        total = fmaf(octave<QUIRK>(lane4, lane16, su, sv, sw, k), amp, total);
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
    return octave<true>(lane4, lane16_of(lane4), su, sv, sw, 0);
}

// Noise::octave_noise_3d (noise.cpp:162-180): sum_k raw(x f_k) a_k / sum a_k, f_k = scale 2^k, a_k = pers^k.
// (px,py,pz) are the float coordinates the reference widens to double; scale is the double frequency.
// NOT inlined, on purpose: the call is where the register allocator splits live ranges. The octave loop wants ~45
// registers and the march around it ~60 of the 80 a thread has (24 warps per SM); inlined, ptxas keeps the march state
// in local memory for its whole life and reloads it in the per-sample loop. Out of line, what does not fit is
// saved and restored around the call - once per noise evaluation (> 1000 instructions), never per sample.
__device__ GN_NOISE_LINKAGE float fbm(uint32_t lane4, float px, float py, float pz, double scale,
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
// The spectral weights pow(frequency_o, -0.05f), frequency_o = ks * lac^o, are rw0 * (lac^-0.05)^o: a running product
// (value arithmetic; a negative ks gives NaN either way, like the reference's pow).
__device__ GN_NOISE_LINKAGE float ridged(uint32_t lane4, float vx, float vy, float vz, int octaves,
                                        float lac, float offset, float gain, float rw0, float rwq) {
    float value = 0.0f, weight = 1.0f, rw = rw0;
    const uint32_t lane16 = lane16_of(lane4);
#pragma unroll 1
    for (int o = 0; o < octaves; o++) {
        Skewed su, sv, sw;
        skew_float(vx, vy, vz, su, sv, sw);
        float signal;
        if (maybe_integer(su, sv, sw, 0)) signal = octave_quirk(lane4, vx, vy, vz);
        else signal = octave<false>(lane4, lane16, su, sv, sw, 0);
        signal = offset - fabsf(32.0f * signal);
        signal *= signal;
        signal *= weight;
        weight = fminf(fmaxf(signal * gain, 0.0f), 1.0f);
        value = fmaf(signal, rw, value);
        rw *= rwq;
        vx = __fmul_rn(vx, lac); vy = __fmul_rn(vy, lac); vz = __fmul_rn(vz, lac);
    }
    return fmaf(value, 1.25f, -1.0f);
}

}  // namespace gnoise
