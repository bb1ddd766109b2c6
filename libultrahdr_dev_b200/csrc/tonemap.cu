// tonemap.cu — A3: UltraHdr::toneMap (reference lib/src/ultrahdr.cpp:517-558) on sm_100a.
//
// Integer only, bit exact: 8-bit sample = top 8 bits of the 16-bit P010 word
// ((v>>6)>>2 & 0xff == v>>8), UV de-interleaved into planar U and V, the stride tails
// [width, luma_stride) and [width/2, chroma_stride) zero-filled like the reference's memsets.
// HBM-bound: 3 B/px read, 1.5 B/px written. One thread moves one 16-byte group (8 samples) with a
// 128-bit load; luma and chroma groups share one flat index space so every warp is full.
#include "uhdr_kernels.cuh"

namespace uhdr {
namespace {

__device__ __forceinline__ uint4 ld_stream16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
// high bytes of the two u16 halves of a and b -> 4 packed bytes (a.lo, a.hi, b.lo, b.hi)
__device__ __forceinline__ uint32_t hi_bytes(uint32_t a, uint32_t b) { return __byte_perm(a, b, 0x7531); }

template <bool VEC>
__global__ void __launch_bounds__(256) tonemap_kernel(ToneMapParams p) {
  const Planes pl = p.batch ? p.batch[blockIdx.z] : p.single;
  const Geometry& g = p.g;
  const int gprL = (g.dst_luma_stride + 7) >> 3;    // luma groups per row (8 dest bytes each)
  const int gprC = (g.dst_chroma_stride + 3) >> 2;  // chroma groups per row (4 U + 4 V bytes)
  const long long nL = (long long)g.height * gprL;
  const long long nC = (long long)(g.height / 2) * gprC;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < nL) {
    const int row = (int)(idx / gprL), grp = (int)(idx - (long long)row * gprL);
    const int x0 = grp * 8;
    const uint16_t* src = (const uint16_t*)pl.hdr_y + (size_t)row * g.hdr_luma_stride + x0;
    uint8_t* dst = (uint8_t*)pl.out + (size_t)row * g.dst_luma_stride + x0;
    if (VEC && x0 + 8 <= g.width) {
      uint4 v = ld_stream16(src);
      uint2 o = make_uint2(hi_bytes(v.x, v.y), hi_bytes(v.z, v.w));
      *reinterpret_cast<uint2*>(dst) = o;
    } else {
#pragma unroll
      for (int i = 0; i < 8; i++) {
        int x = x0 + i;
        if (x < g.width) dst[i] = (uint8_t)(src[i] >> 8);
        else if (x < g.dst_luma_stride) dst[i] = 0;  // ultrahdr.cpp:533-535
      }
    }
  } else if (idx < nL + nC) {
    const long long c = idx - nL;
    const int row = (int)(c / gprC), grp = (int)(c - (long long)row * gprC);
    const int x0 = grp * 4, cw = g.width / 2;
    const uint16_t* src = (const uint16_t*)pl.hdr_uv + (size_t)row * g.hdr_chroma_stride + 2 * x0;
    uint8_t* du = (uint8_t*)pl.out_u + (size_t)row * g.dst_chroma_stride + x0;
    uint8_t* dv = (uint8_t*)pl.out_v + (size_t)row * g.dst_chroma_stride + x0;
    if (VEC && x0 + 4 <= cw) {
      uint4 v = ld_stream16(src);  // u0 v0 | u1 v1 | u2 v2 | u3 v3
      uint32_t a = hi_bytes(v.x, v.y), b = hi_bytes(v.z, v.w);  // a = u0 v0 u1 v1, b = u2 v2 u3 v3
      *reinterpret_cast<uint32_t*>(du) = __byte_perm(a, b, 0x6420);
      *reinterpret_cast<uint32_t*>(dv) = __byte_perm(a, b, 0x7531);
    } else {
#pragma unroll
      for (int i = 0; i < 4; i++) {
        int x = x0 + i;
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
  long long groups = (long long)g.height * ((g.dst_luma_stride + 7) / 8) +
                     (long long)(g.height / 2) * ((g.dst_chroma_stride + 3) / 4);
  if (groups == 0 || n == 0) return cudaSuccess;
  dim3 grid((unsigned)((groups + 255) / 256), 1, (unsigned)n);
  if (vec) tonemap_kernel<true><<<grid, 256, 0, s>>>(p);
  else tonemap_kernel<false><<<grid, 256, 0, s>>>(p);
  return cudaGetLastError();
}

}  // namespace uhdr
