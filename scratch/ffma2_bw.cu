// micro-benchmark: issue rate of FFMA vs FFMA2 / FADD.RM vs FADD2.RM on sm_100a (per SM sub-partition)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 addrm2(u64 a, u64 b) { u64 d; asm volatile("add.rm.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ float fma1(float a, float b, float c) { float d; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }
__device__ __forceinline__ float addrm1(float a, float b) { float d; asm volatile("add.rm.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b)); return d; }

template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, int iters, float s) {
  float a[8];
  u64 A[8];
  for (int i = 0; i < 8; i++) { a[i] = threadIdx.x * 0.001f + i; A[i] = pk(a[i], a[i] + 0.5f); }
  const u64 S = pk(s, s), C = pk(0.25f, 0.5f);
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int u = 0; u < 4; u++) {
#pragma unroll
      for (int i = 0; i < 8; i++) {
        if (MODE == 0) a[i] = fma1(a[i], s, 0.25f);
        if (MODE == 1) A[i] = fma2(A[i], S, C);
        if (MODE == 2) a[i] = addrm1(a[i], s);
        if (MODE == 3) A[i] = addrm2(A[i], S);
        if (MODE == 4) { a[i] = fma1(a[i], s, 0.25f); A[i] = fma2(A[i], S, C); }  // mixed 1:1
      }
    }
  }
  float r = 0;
  for (int i = 0; i < 8; i++) r += a[i] + __uint_as_float((uint32_t)A[i]) + __uint_as_float((uint32_t)(A[i] >> 32));
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int MODE>
void run(const char* name, int per_iter) {
  float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
  const int iters = 20000;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int blocks_per_sm = 1; blocks_per_sm <= 4; blocks_per_sm *= 2) {
    k<MODE><<<148 * blocks_per_sm, 256>>>(d, 100, 1.0001f);
    cudaEventRecord(e0);
    k<MODE><<<148 * blocks_per_sm, 256>>>(d, iters, 1.0001f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const double warp_instr = (double)iters * 32 * per_iter * 8 * blocks_per_sm;  // per SM (8 warps per block)
    // cycles at the nominal max clock (upper bound on cycles actually run)
    const double cycles = ms * 1e-3 * clk * 1e3;
    printf("%-10s warps/SMSP=%d  %.3f ms  warp-instr/clk/SMSP = %.3f (at %d MHz)\n", name, 2 * blocks_per_sm, ms, warp_instr / 4 / cycles, clk / 1000);
  }
  cudaFree(d);
}
int main() {
  run<0>("FFMA", 1); run<1>("FFMA2", 1); run<2>("FADD.RM", 1); run<3>("FADD2.RM", 1); run<4>("FFMA+FFMA2", 2);
  return 0;
}
