// Micro-benchmark: issue cost of packed FP32 (FFMA2/FADD2/FMUL2) against scalar FP32 on sm_100a, alone and
// mixed with ALU-pipe work. Prints warp-instructions per cycle per SM sub-partition for each variant.
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
#define ITER 4096
__device__ __forceinline__ void ffma2(float &a0, float &a1, float b0, float b1, float c0, float c1) {
    unsigned long long A, B, C;
    asm("mov.b64 %0, {%1,%2};" : "=l"(A) : "f"(a0), "f"(a1));
    asm("mov.b64 %0, {%1,%2};" : "=l"(B) : "f"(b0), "f"(b1));
    asm("mov.b64 %0, {%1,%2};" : "=l"(C) : "f"(c0), "f"(c1));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(A) : "l"(A), "l"(B), "l"(C));
    asm("mov.b64 {%0,%1}, %2;" : "=f"(a0), "=f"(a1) : "l"(A));
}
template <int MODE>
__global__ void __launch_bounds__(256) kern(float *out, float s, uint32_t m) {
    float a[8];
    uint32_t u[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { a[i] = threadIdx.x * 0.001f + i; u[i] = threadIdx.x * 77u + i; }
    const float b = s, c = s * 0.5f;
#pragma unroll 1
    for (int it = 0; it < ITER; it++) {
        if (MODE == 0 || MODE == 2 || MODE == 4) {
#pragma unroll
            for (int i = 0; i < 8; i++) a[i] = fmaf(a[i], b, c);
        }
        if (MODE == 1 || MODE == 3 || MODE == 5) {
#pragma unroll
            for (int i = 0; i < 8; i += 2) ffma2(a[i], a[i + 1], b, b, c, c);
        }
        if (MODE == 2 || MODE == 3) {
#pragma unroll
            for (int i = 0; i < 8; i++) u[i] = (u[i] & m) ^ (u[(i + 1) & 7] >> 1);   // LOP3 + SHF
        }
        if (MODE == 4 || MODE == 5) {
#pragma unroll
            for (int i = 0; i < 4; i++) u[i] = (u[i] & m) ^ (u[(i + 1) & 7] >> 1);
        }
        if (MODE == 6) {
#pragma unroll
            for (int i = 0; i < 8; i++) u[i] = (u[i] & m) ^ (u[(i + 1) & 7] >> 1);
        }
        if (MODE == 7) {   // FSEL / FMNMX / FSETP style
#pragma unroll
            for (int i = 0; i < 8; i++) a[i] = fmaxf(a[i], a[(i + 1) & 7] > b ? c : s);
        }
    }
    float r = 0; uint32_t q = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) { r += a[i]; q ^= u[i]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = r + __uint_as_float(q & 0xff);
}
template <int MODE> void run(const char *name, int per_iter, float *out) {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    int blocks = sms * 8;
    kern<MODE><<<blocks, 256>>>(out, 1.0001f, 0x7fffffffu);
    cudaEventRecord(e0);
    kern<MODE><<<blocks, 256>>>(out, 1.0001f, 0x7fffffffu);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double warps = (double)blocks * 8, inst = warps * ITER * per_iter;
    double cyc = ms * 1e-3 * clk * 1e3;
    printf("%-34s %8.3f ms  %.3f warp-inst/clk/SMSP (nominal %d inst/iter, clock %d kHz)\n", name, ms, inst / cyc / (sms * 4), per_iter, clk);
}
int main() {
    float *out; cudaMalloc(&out, 148 * 8 * 256 * 4 * 4);
    run<0>("8 FFMA", 8, out);
    run<1>("4 FFMA2", 4, out);
    run<6>("8 LOP3 + 8 SHF", 16, out);
    run<2>("8 FFMA + 16 ALU", 24, out);
    run<3>("4 FFMA2 + 16 ALU", 20, out);
    run<4>("8 FFMA + 8 ALU", 16, out);
    run<5>("4 FFMA2 + 8 ALU", 12, out);
    run<7>("8 x (FSETP+FSEL+FMNMX)", 24, out);
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
