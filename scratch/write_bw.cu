// micro-benchmark: achievable DRAM write / read / copy bandwidth with different store flavours
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
template <int MODE>
__global__ void wr(uint4* p, size_t n, uint32_t v) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, st = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += st) {
    uint4 x = make_uint4(v + (uint32_t)i, v, v, v);
    if (MODE == 0) p[i] = x;
    else if (MODE == 1) asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p + i), "r"(x.x), "r"(x.y), "r"(x.z), "r"(x.w) : "memory");
    else if (MODE == 2) asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p + i), "r"(x.x), "r"(x.y), "r"(x.z), "r"(x.w) : "memory");
    else if (MODE == 3) asm volatile("st.global.wt.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p + i), "r"(x.x), "r"(x.y), "r"(x.z), "r"(x.w) : "memory");
  }
}
__global__ void rd(const uint4* p, size_t n, uint32_t* out) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, st = (size_t)gridDim.x * blockDim.x;
  uint32_t acc = 0;
  for (; i < n; i += st) { uint4 x; asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(x.x), "=r"(x.y), "=r"(x.z), "=r"(x.w) : "l"(p + i)); acc ^= x.x ^ x.y ^ x.z ^ x.w; }
  if (acc == 0x12345) *out = acc;
}
// mixed: read r bytes for every w bytes written (like apply: 1.5625 read / 4 written)
__global__ void mix(const uint4* src, uint4* dst, size_t n_dst, int ratio_num, int ratio_den) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, st = (size_t)gridDim.x * blockDim.x;
  for (; i < n_dst; i += st) {
    uint4 x = make_uint4((uint32_t)i, 1, 2, 3);
    if ((i % ratio_den) < (size_t)ratio_num) { uint4 y; asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(y.x), "=r"(y.y), "=r"(y.z), "=r"(y.w) : "l"(src + i)); x.x ^= y.x; x.y ^= y.y; }
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(dst + i), "r"(x.x), "r"(x.y), "r"(x.z), "r"(x.w) : "memory");
  }
}
template <class F> float timeit(F f) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int i = 0; i < 2; i++) f();
  cudaEventRecord(a); for (int i = 0; i < 5; i++) f(); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b); return ms / 5;
}
int main() {
  const size_t bytes = (size_t)8 << 30, n = bytes / 16;
  uint4 *p, *q; uint32_t* o;
  cudaMalloc(&p, bytes); cudaMalloc(&q, bytes); cudaMalloc(&o, 4);
  cudaMemset(p, 1, bytes); cudaMemset(q, 1, bytes);
  for (int grid : {148 * 8, 148 * 16, 148 * 64}) {
    printf("grid %d\n", grid);
    printf("  st default      %.0f GB/s\n", bytes / 1e6 / timeit([&] { wr<0><<<grid, 256>>>(p, n, 1); }));
    printf("  st no_allocate  %.0f GB/s\n", bytes / 1e6 / timeit([&] { wr<1><<<grid, 256>>>(p, n, 1); }));
    printf("  st .cs          %.0f GB/s\n", bytes / 1e6 / timeit([&] { wr<2><<<grid, 256>>>(p, n, 1); }));
    printf("  st .wt          %.0f GB/s\n", bytes / 1e6 / timeit([&] { wr<3><<<grid, 256>>>(p, n, 1); }));
    printf("  ld              %.0f GB/s\n", bytes / 1e6 / timeit([&] { rd<<<grid, 256>>>(p, n, o); }));
    printf("  copy 1:1 (r+w)  %.0f GB/s\n", 2.0 * bytes / 1e6 / timeit([&] { mix<<<grid, 256>>>(p, q, n, 1, 1); }));
    printf("  mix 0.39:1 r:w  %.0f GB/s (write part %.0f)\n", 1.39 * bytes / 1e6 / timeit([&] { mix<<<grid, 256>>>(p, q, n, 39, 100); }), bytes / 1e6 / timeit([&] { mix<<<grid, 256>>>(p, q, n, 39, 100); }));
  }
  printf("  cudaMemset      %.0f GB/s\n", bytes / 1e6 / timeit([&] { cudaMemsetAsync(p, 0, bytes); }));
  printf("  cudaMemcpy D2D  %.0f GB/s (r+w)\n", 2.0 * bytes / 1e6 / timeit([&] { cudaMemcpyAsync(q, p, bytes, cudaMemcpyDeviceToDevice); }));
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
}
