// tone.cuh — the per-pixel arithmetic of K2 (Buffer2D::PostProcess + ToColorBuffer, buffer2d.cpp:65-85,106-126),
// after exposure and gamma: saturation about the mean, x10, clamp, truncate to a byte, pack two ARGB32 words.
//
// What is written is one BYTE per channel, (int)clamp(10 v), and the reference forms v through
// float -> double -> float -> double -> float: 13 conversions per pixel on the quarter-rate XU pipe, which was this
// kernel's limit (ncu r02: XU 76 % busy, HBM at half its rate). tone_decide forms the same v in plain FP32 together
// with a proven bound on its distance from the reference's float, and decides the byte whenever no integer lies within
// that bound of 10 v; otherwise (about one pixel in 10^4) the pixel goes through tone_exact, the reference's own chain.
// The result is the reference's byte in every case.
//
// Compiles for the host as well (built with -ffp-contract=off): tests/test_cpu.py checks, on values placed at the
// integer boundaries, that a decided byte always equals the exact chain's.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDA_ARCH__)
#define TN_FN __device__ __forceinline__
#define TN_SLOW __device__ __noinline__
#define TN_FMUL(a, b) __fmul_rn((a), (b))
#define TN_FADD(a, b) __fadd_rn((a), (b))
#define TN_DMUL(a, b) __dmul_rn((a), (b))
#define TN_DSUB(a, b) __dsub_rn((a), (b))
#define TN_DFMA(a, b, c) __fma_rn((a), (b), (c))
#else
#if defined(__CUDACC__)
#define TN_FN __host__ __device__ inline
#define TN_SLOW __host__ __device__ inline
#else
#define TN_FN inline
#define TN_SLOW inline
#endif
#define TN_FMUL(a, b) ((a) * (b))
#define TN_FADD(a, b) ((a) + (b))
#define TN_DMUL(a, b) ((a) * (b))
#define TN_DSUB(a, b) ((a) - (b))
#define TN_DFMA(a, b, c) fma((a), (b), (c))
#endif

namespace tone {

// (int) of the clamped value: Util::clamp is std::min(255, v) then std::max(0, v), NaN becomes 255. On the device
// v + 2^23 rounded DOWN leaves floor(v) in the low significand bits (an FADD instead of a conversion).
TN_FN uint32_t byte_of10(float v) {
    v = (v < 255.0f) ? v : 255.0f; v = (0.0f < v) ? v : 0.0f;
#if defined(__CUDA_ARCH__)
    return __float_as_uint(__fadd_rd(v, 8388608.0f)) & 0xFFu;
#else
    return (uint32_t)(int)v;
#endif
}
// the reference's chain on the exposed (and gamma-mapped) pixel
// returns the three bytes as 0x00BBGGRR (by value: an out-of-line call with reference parameters goes through the stack)
TN_SLOW uint32_t tone_exact(float x, float y, float z, double saturation) {
    // center = double(float sum) / 3.0 by the IEEE division itself (this path is rare and out of line): an infinite
    // sum must stay infinite (inf / 3), which a divide-free q + (s - 3 q) / 3 does not give
    const double center = (double)TN_FADD(TN_FADD(x, y), z) / 3.0;
    const float tx = (float)TN_DSUB(center, (double)x), ty = (float)TN_DSUB(center, (double)y), tz = (float)TN_DSUB(center, (double)z);
    x = (float)TN_DSUB(center, TN_DMUL(saturation, (double)tx));
    y = (float)TN_DSUB(center, TN_DMUL(saturation, (double)ty));
    z = (float)TN_DSUB(center, TN_DMUL(saturation, (double)tz));
    return byte_of10(TN_FMUL(x, 10.0f)) | (byte_of10(TN_FMUL(y, 10.0f)) << 8) | (byte_of10(TN_FMUL(z, 10.0f)) << 16);
}

// The pixel in FP32. With s the (identical) float sum, C = RN64(s/3), T = RN32(C - x), X = RN32(C - S T),
// X10 = RN32(10 X) the reference's values and c = RN32(s RN32(1/3)), c10 = RN32(10 c), h = RN32(c10 - 1/2),
// t = RN32(c - x), U = RN32(h - S10 t) (one FMA, S10 = RN32(10 S)) ours, U standing for X10 - 1/2:
//   |c - C| <= 1.6 2^-24 |c|,   |t - T| <= 1.01 |c - C| + 1.01 2^-23 |t|,   |c10 - 10 c| <= 10 2^-24 |c|,
//   |h - (c10 - 1/2)| <= 2^-24 (10 |c| + 1/2),   |S10 t - 10 S T| <= 10 |S| 2^-23 (0.81 |c| + 1.51 |t|),
//   |U + 1/2 - X10| <= 10 2^-23 (1.8 |c| + 0.81 |S| |c| + 1.51 |S| |t| + 0.15 |U|) + 9e-8.
// E = 2.5e-6 (|c| (1.25 + |S|) + |S| sum|t| + 0.2 sum|U|) + 2e-7, one bound for the three channels, exceeds that by a
// factor >= 1.39 in every term (2.5e-6 = 10 x 2^-23 x 2.1). It is built from SUMS so that an infinity or NaN anywhere
// in the pixel makes E non-finite and the comparison below false.
// The byte changes only where X10 crosses an integer in 1..255, i.e. where X10 - 1/2 crosses k + 1/2, k = 0..254.
// u = U clamped to [0, 255] has those crossings at the same places and none outside; with n = RNE(u), the nearest
// integer, the distance of u to the nearest crossing is 1/2 - |u - n|, so the pixel is decided when the LARGEST |u - n|
// of its channels is below 1/2 - E. The byte is then n itself: floor(X10) = RNE(X10 - 1/2) unless X10 is an integer,
// and an integer X10 is a crossing, never decided. n comes out of the add that rounds: -u - 1.5 2^23 has the bits
// 0xCB4000nn, n in the low byte and a top byte with its sign bit set, which pack_px turns into the alpha byte 0xFF.
struct ToneSat { float s10, a, a1; };      // 10 S as a float, |S|, 1.25 + |S|
TN_FN ToneSat tone_sat(double saturation) {
    ToneSat k; k.s10 = (float)(10.0 * saturation); k.a = fabsf((float)saturation); k.a1 = 1.25f + k.a; return k;
}
TN_FN uint32_t tone_bits(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}
// one channel: the word holding the byte, and |u - n|
TN_FN uint32_t tone_channel(float U, float &dist) {
    const float u = fminf(fmaxf(U, 0.0f), 255.0f);
    const float m = TN_FADD(-u, -12582912.0f);              // -(1.5 2^23 + n)
    const float n = TN_FADD(-m, -12582912.0f);
    dist = fabsf(TN_FADD(u, -n));
    return tone_bits(m);
}
// ARGB32 (r in the low byte: Buffer2D::ToColorBuffer's first image) and the shadow image (b in the low byte) from the
// three channel words: four byte permutes, the alpha byte by sign replication of a top byte
TN_FN void pack_px(uint32_t mr, uint32_t mg, uint32_t mb, uint32_t &argb, uint32_t &shadow) {
#if defined(__CUDA_ARCH__)
    uint32_t rg, bg;
    asm("prmt.b32 %0, %1, %2, 0x0040;" : "=r"(rg) : "r"(mr), "r"(mg));          // [r, g, ., .]
    asm("prmt.b32 %0, %1, %2, 0x0040;" : "=r"(bg) : "r"(mb), "r"(mg));          // [b, g, ., .]
    asm("prmt.b32 %0, %1, %2, 0xF410;" : "=r"(argb) : "r"(rg), "r"(mb));        // [r, g, b, sign(mb.3)]
    asm("prmt.b32 %0, %1, %2, 0xF410;" : "=r"(shadow) : "r"(bg), "r"(mr));      // [b, g, r, sign(mr.3)]
#else
    const uint32_t r = mr & 0xFFu, g = mg & 0xFFu, b = mb & 0xFFu;
    argb   = 0xFF000000u | (b << 16) | (g << 8) | r;
    shadow = 0xFF000000u | (r << 16) | (g << 8) | b;
#endif
}
TN_FN bool tone_decide(float x, float y, float z, const ToneSat k, uint32_t &mr, uint32_t &mg, uint32_t &mb) {
    const float c = TN_FMUL(TN_FADD(TN_FADD(x, y), z), 1.0f / 3.0f), h = TN_FADD(TN_FMUL(c, 10.0f), -0.5f);
    const float tx = TN_FADD(c, -x), ty = TN_FADD(c, -y), tz = TN_FADD(c, -z);
    const float Ux = fmaf(-k.s10, tx, h), Uy = fmaf(-k.s10, ty, h), Uz = fmaf(-k.s10, tz, h);
    const float st = TN_FADD(TN_FADD(fabsf(tx), fabsf(ty)), fabsf(tz)), su = TN_FADD(TN_FADD(fabsf(Ux), fabsf(Uy)), fabsf(Uz));
    const float room = fmaf(-2.5e-6f, fmaf(k.a, st, fmaf(0.2f, su, TN_FMUL(fabsf(c), k.a1))), 0.5f - 2e-7f);      // 1/2 - E
    float dx, dy, dz;
    mr = tone_channel(Ux, dx); mg = tone_channel(Uy, dy); mb = tone_channel(Uz, dz);
    return fmaxf(fmaxf(dx, dy), dz) < room;
}

// the two words of a pixel, decided in FP32 (the words are valid only if true is returned) / by the exact chain
TN_FN bool tone_mapped_fast(float x, float y, float z, const ToneSat k, uint32_t &argb, uint32_t &shadow) {
    uint32_t mr, mg, mb;
    const bool ok = tone_decide(x, y, z, k, mr, mg, mb);
    pack_px(mr, mg, mb, argb, shadow);
    return ok;
}
TN_FN void tone_mapped_exact(float x, float y, float z, double saturation, uint32_t &argb, uint32_t &shadow) {
    const uint32_t p = tone_exact(x, y, z, saturation);          // 0x00BBGGRR
    argb = 0xFF000000u | p;
    shadow = 0xFF000000u | ((p & 0xFFu) << 16) | (p & 0xFF00u) | (p >> 16);
}
// decided = the pixel was decided in FP32 (reported for the tests; the words are right either way)
TN_FN bool tone_mapped(float x, float y, float z, double saturation, const ToneSat k, uint32_t &argb, uint32_t &shadow) {
    const bool ok = tone_mapped_fast(x, y, z, k, argb, shadow);
    if (!ok) tone_mapped_exact(x, y, z, saturation, argb, shadow);
    return ok;
}

}  // namespace tone
