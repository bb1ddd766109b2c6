// tonemap.cu — A3: UltraHdr::toneMap (reference lib/src/ultrahdr.cpp:517-558) on sm_100a.
//
// Integer only, bit exact: 8-bit sample = top 8 bits of the 16-bit P010 word
// ((v>>6)>>2 & 0xff == v>>8), UV de-interleaved into planar U and V, the stride tails
// [width, luma_stride) and [width/2, chroma_stride) zero-filled like the reference's memsets.
// HBM-bound: 3 B/px read, 1.5 B/px written. One thread moves 8 pixels x 4 rows (luma + chroma).
#include "uhdr_io.cuh"
#include "uhdr_kernels.cuh"

namespace uhdr {
namespace {

// high bytes of the two u16 halves of a and b -> 4 packed bytes (a.lo, a.hi, b.lo, b.hi)
__device__ __forceinline__ uint32_t hi_bytes(uint32_t a, uint32_t b) { return __byte_perm(a, b, 0x7531); }

// one luma group: 8 destination bytes of row `row` starting at x0 = 8 * grp (may reach into the stride tail)
template <bool VEC>
__device__ __forceinline__ void luma_group(const Planes& pl, const Geometry& g, int row, int grp) {
  const int x0 = grp * 8;
  const uint16_t* src = (const uint16_t*)pl.hdr_y + (size_t)row * g.hdr_luma_stride + x0;
  uint8_t* dst = (uint8_t*)pl.out + (size_t)row * g.dst_luma_stride + x0;
  if (VEC && x0 + 8 <= g.width) {
    const uint4 v = ldg16(src);
    stg8(dst, make_uint2(hi_bytes(v.x, v.y), hi_bytes(v.z, v.w)));
  } else {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      const int x = x0 + i;
      if (x < g.width) dst[i] = (uint8_t)(src[i] >> 8);
      else if (x < g.dst_luma_stride) dst[i] = 0;  // ultrahdr.cpp:533-535
    }
  }
}

// one chroma group: 4 U + 4 V destination bytes of chroma row `row` starting at x0 = 4 * grp
template <bool VEC>
__device__ __forceinline__ void chroma_group(const Planes& pl, const Geometry& g, int row, int grp) {
  const int x0 = grp * 4, cw = g.width / 2;
  const uint16_t* src = (const uint16_t*)pl.hdr_uv + (size_t)row * g.hdr_chroma_stride + 2 * x0;
  uint8_t* du = (uint8_t*)pl.out_u + (size_t)row * g.dst_chroma_stride + x0;
  uint8_t* dv = (uint8_t*)pl.out_v + (size_t)row * g.dst_chroma_stride + x0;
  if (VEC && x0 + 4 <= cw) {
    const uint4 v = ldg16(src);  // u0 v0 | u1 v1 | u2 v2 | u3 v3
    const uint32_t a = hi_bytes(v.x, v.y), b = hi_bytes(v.z, v.w);  // a = u0 v0 u1 v1, b = u2 v2 u3 v3
    stg4(du, __byte_perm(a, b, 0x6420));
    stg4(dv, __byte_perm(a, b, 0x7531));
  } else {
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const int x = x0 + i;
      if (x < cw) {
        du[i] = (uint8_t)(src[2 * i] >> 8);
        dv[i] = (uint8_t)(src[2 * i + 1] >> 8);
      } else if (x < g.dst_chroma_stride) {  // ultrahdr.cpp:551-554
        du[i] = 0;
        dv[i] = 0;
      }
    }
  }
}

// One thread = one unit of 8 pixels x 4 rows: four 16-byte luma loads and two 16-byte chroma loads
// (96 bytes) are in flight before the first store, which is what it takes to cover HBM latency at
// 2048 threads per SM. Consecutive lanes take consecutive units of a row, so every warp-level
// request is one contiguous 512-byte (load) / 256-byte or 128-byte (store) segment.
template <bool VEC>
__global__ void __launch_bounds__(256) tonemap_kernel(ToneMapParams p) {
  const Planes pl = p.batch ? p.batch[blockIdx.z] : p.single;
  const Geometry& g = p.g;
  const uint32_t gprL = (g.dst_luma_stride + 7) >> 3;    // luma groups per row (8 dest bytes each)
  const uint32_t gprC = (g.dst_chroma_stride + 3) >> 2;  // chroma groups per row (4 U + 4 V bytes)
  const uint32_t upr = max(gprL, gprC);                  // units per unit-row
  const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t uy = idx / upr, ux = idx - uy * upr;
  const int r0 = 4 * (int)uy, c0 = 2 * (int)uy;
  if (r0 >= g.height) return;
  if (VEC && 8 * (int)ux + 8 <= g.width && r0 + 4 <= g.height) {
    const uint16_t* sy = (const uint16_t*)pl.hdr_y + (size_t)r0 * g.hdr_luma_stride + 8 * ux;
    const uint16_t* sc = (const uint16_t*)pl.hdr_uv + (size_t)c0 * g.hdr_chroma_stride + 8 * ux;
    uint4 y[4], c[2];
#pragma unroll
    for (int r = 0; r < 4; r++) y[r] = ldg16(sy + (size_t)r * g.hdr_luma_stride);
#pragma unroll
    for (int r = 0; r < 2; r++) c[r] = ldg16(sc + (size_t)r * g.hdr_chroma_stride);
    uint8_t* dy = (uint8_t*)pl.out + (size_t)r0 * g.dst_luma_stride + 8 * ux;
#pragma unroll
    for (int r = 0; r < 4; r++)
      stg8(dy + (size_t)r * g.dst_luma_stride, make_uint2(hi_bytes(y[r].x, y[r].y), hi_bytes(y[r].z, y[r].w)));
    uint8_t* du = (uint8_t*)pl.out_u + (size_t)c0 * g.dst_chroma_stride + 4 * ux;
    uint8_t* dv = (uint8_t*)pl.out_v + (size_t)c0 * g.dst_chroma_stride + 4 * ux;
#pragma unroll
    for (int r = 0; r < 2; r++) {
      const uint32_t a = hi_bytes(c[r].x, c[r].y), b = hi_bytes(c[r].z, c[r].w);
      stg4(du + (size_t)r * g.dst_chroma_stride, __byte_perm(a, b, 0x6420));
      stg4(dv + (size_t)r * g.dst_chroma_stride, __byte_perm(a, b, 0x7531));
    }
    return;
  }
  // edges: right-most groups, stride tails, a last unit-row of two rows
  for (int r = r0; r < min(r0 + 4, g.height); r++)
    if (ux < gprL) luma_group<VEC>(pl, g, r, (int)ux);
  for (int r = c0; r < min(c0 + 2, g.height / 2); r++)
    if (ux < gprC) chroma_group<VEC>(pl, g, r, (int)ux);
}

bool aligned(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) == 0; }
}  // namespace

// The vector path needs 16-byte aligned source rows and 8/4-byte aligned destination rows; the
// caller (capi.cu) passes vec_ok computed over every image of the batch via planes_vec_ok().
bool tonemap_vec_ok(const Geometry& g, const Planes& pl) {
  return aligned(pl.hdr_y, 16) && aligned(pl.hdr_uv, 16) && aligned(pl.out, 8) && aligned(pl.out_u, 4) &&
         aligned(pl.out_v, 4) && g.hdr_luma_stride % 8 == 0 && g.hdr_chroma_stride % 8 == 0 &&
         g.dst_luma_stride % 8 == 0 && g.dst_chroma_stride % 4 == 0;
}

cudaError_t launch_tonemap_impl(const ToneMapParams& p, int n, bool vec, cudaStream_t s) {
  const Geometry& g = p.g;
  const long long gprL = (g.dst_luma_stride + 7) / 8, gprC = (g.dst_chroma_stride + 3) / 4;
  const long long units = (gprL > gprC ? gprL : gprC) * ((g.height + 3) / 4);
  if (units == 0 || n == 0) return cudaSuccess;
  dim3 grid((unsigned)((units + 255) / 256), 1, (unsigned)n);
  if (vec) tonemap_kernel<true><<<grid, 256, 0, s>>>(p);
  else tonemap_kernel<false><<<grid, 256, 0, s>>>(p);
  return cudaGetLastError();
}

}  // namespace uhdr
