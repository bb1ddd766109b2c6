// generate.cu — A1: UltraHdr::generateGainMap per-map-pixel loop (reference
// lib/src/ultrahdr.cpp:314-336) on sm_100a.
//
// One thread = one gain-map pixel = one 4x4 block of the SDR (YUV420 u8) and HDR (P010) images.
// HBM-bound (4.5625 B/px: 3 B P010 + 1.5 B 420 read, 1/16 B written). A warp covers 128 image
// columns, so every row request is one fully coalesced 256 B (P010 Y/UV), 128 B (420 Y) or 64 B
// (420 U, V) segment; the 14 independent loads per thread are issued before any use.
// The 4x4 box average reproduces the reference's float accumulation order exactly
// (gainmapmath.cpp:605-615: dy-major, dx-minor, one division at the end), so the only deviation
// from the reference is the fp32 transcendental chain (~1e-6 relative, see uhdr_math.cuh).
#include "uhdr_kernels.cuh"
#include "uhdr_math.cuh"

namespace uhdr {
namespace {

__device__ __forceinline__ uint2 ldg8(const void* p) {
  uint2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ uint32_t ldg4(const void* p) {
  uint32_t r;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ uint32_t ldg2(const void* p) {
  uint16_t r;
  asm volatile("ld.global.nc.L1::no_allocate.u16 %0, [%1];" : "=h"(r) : "l"(p));
  return r;
}

// getYuv420Pixel scaling, gainmapmath.cpp:579-580
__device__ __forceinline__ float sdr_luma(uint32_t v) { return (float)v * (1 / 255.0f); }
__device__ __forceinline__ float sdr_chroma(uint32_t v) { return (float)((int)v - 128) * (1 / 255.0f); }
// getP010Pixel scaling, gainmapmath.cpp:593-600 (no limited-range clamp)
__device__ __forceinline__ float hdr_luma(uint32_t v16) { return (float)((int)(v16 >> 6) - 64) * (1 / 876.0f); }
__device__ __forceinline__ float hdr_chroma(uint32_t v16) { return (float)((int)(v16 >> 6) - 64) * (1 / 896.0f) - 0.5f; }

template <bool VEC>
__global__ void __launch_bounds__(256) generate_kernel(GenerateParams p) {
  const Geometry& g = p.g;
  const int mx = blockIdx.x * 32 + threadIdx.x;
  const int my = blockIdx.y * 8 + threadIdx.y;
  if (mx >= g.map_width || my >= g.map_height) return;
  const Planes pl = p.batch ? p.batch[blockIdx.z] : p.single;

  // ---- gather the raw samples of the 4x4 block (all loads first) ----
  uint32_t sy[4][4], su[2][2], sv[2][2];  // [row][col]
  uint32_t hy[4][4], hu[2][2], hv[2][2];
  const int x0 = mx * 4, y0 = my * 4;
  {
    const uint8_t* y8 = (const uint8_t*)pl.sdr_y + (size_t)y0 * g.sdr_luma_stride + x0;
    const uint8_t* u8 = (const uint8_t*)pl.sdr_u + (size_t)(y0 >> 1) * g.sdr_chroma_stride + (x0 >> 1);
    const uint8_t* v8 = (const uint8_t*)pl.sdr_v + (size_t)(y0 >> 1) * g.sdr_chroma_stride + (x0 >> 1);
    const uint16_t* y16 = (const uint16_t*)pl.hdr_y + (size_t)y0 * g.hdr_luma_stride + x0;
    const uint16_t* uv16 = (const uint16_t*)pl.hdr_uv + (size_t)(y0 >> 1) * g.hdr_chroma_stride + x0;
    if (VEC) {
      uint2 a[4], c[2];
      uint32_t b[4], d[2], e[2];
#pragma unroll
      for (int r = 0; r < 4; r++) a[r] = ldg8(y16 + (size_t)r * g.hdr_luma_stride);
#pragma unroll
      for (int r = 0; r < 2; r++) c[r] = ldg8(uv16 + (size_t)r * g.hdr_chroma_stride);
#pragma unroll
      for (int r = 0; r < 4; r++) b[r] = ldg4(y8 + (size_t)r * g.sdr_luma_stride);
#pragma unroll
      for (int r = 0; r < 2; r++) {
        d[r] = ldg2(u8 + (size_t)r * g.sdr_chroma_stride);
        e[r] = ldg2(v8 + (size_t)r * g.sdr_chroma_stride);
      }
#pragma unroll
      for (int r = 0; r < 4; r++) {
        hy[r][0] = a[r].x & 0xffff; hy[r][1] = a[r].x >> 16; hy[r][2] = a[r].y & 0xffff; hy[r][3] = a[r].y >> 16;
        sy[r][0] = b[r] & 0xff; sy[r][1] = (b[r] >> 8) & 0xff; sy[r][2] = (b[r] >> 16) & 0xff; sy[r][3] = b[r] >> 24;
      }
#pragma unroll
      for (int r = 0; r < 2; r++) {
        hu[r][0] = c[r].x & 0xffff; hv[r][0] = c[r].x >> 16; hu[r][1] = c[r].y & 0xffff; hv[r][1] = c[r].y >> 16;
        su[r][0] = d[r] & 0xff; su[r][1] = d[r] >> 8; sv[r][0] = e[r] & 0xff; sv[r][1] = e[r] >> 8;
      }
    } else {
#pragma unroll
      for (int r = 0; r < 4; r++)
#pragma unroll
        for (int c = 0; c < 4; c++) {
          hy[r][c] = y16[(size_t)r * g.hdr_luma_stride + c];
          sy[r][c] = y8[(size_t)r * g.sdr_luma_stride + c];
        }
#pragma unroll
      for (int r = 0; r < 2; r++)
#pragma unroll
        for (int c = 0; c < 2; c++) {
          hu[r][c] = uv16[(size_t)r * g.hdr_chroma_stride + 2 * c];
          hv[r][c] = uv16[(size_t)r * g.hdr_chroma_stride + 2 * c + 1];
          su[r][c] = u8[(size_t)r * g.sdr_chroma_stride + c];
          sv[r][c] = v8[(size_t)r * g.sdr_chroma_stride + c];
        }
    }
  }

  // ---- 4x4 box averages in the reference's accumulation order (samplePixels) ----
  Color3 s = {0.f, 0.f, 0.f}, h = {0.f, 0.f, 0.f};
#pragma unroll
  for (int r = 0; r < 4; r++)
#pragma unroll
    for (int c = 0; c < 4; c++) {
      s.r += sdr_luma(sy[r][c]); s.g += sdr_chroma(su[r >> 1][c >> 1]); s.b += sdr_chroma(sv[r >> 1][c >> 1]);
      h.r += hdr_luma(hy[r][c]); h.g += hdr_chroma(hu[r >> 1][c >> 1]); h.b += hdr_chroma(hv[r >> 1][c >> 1]);
    }
  s.r *= 0.0625f; s.g *= 0.0625f; s.b *= 0.0625f;  // "/16.0f" is exact either way
  h.r *= 0.0625f; h.g *= 0.0625f; h.b *= 0.0625f;

  // ---- SDR: YUV->RGB, sRGB^-1, luminance * 203 (ultrahdr.cpp:316-324) ----
  const LumaCoefs lk = p.luma;
  Color3 sg = yuv_to_rgb(p.sdr_k, s);
  Color3 sl = {srgb_inv_oetf(sg.r), srgb_inv_oetf(sg.g), srgb_inv_oetf(sg.b)};
  const float sdr_nits = luminance(lk, sl) * 203.0f;

  // ---- HDR: YUV->RGB, inverse OETF, gamut conversion into the SDR gamut, luminance (:326-330) ----
  Color3 hg = yuv_to_rgb(p.hdr_k, h);
  Color3 hl;
  if (p.hdr_tf == 1) hl = {hlg_inv_oetf(hg.r), hlg_inv_oetf(hg.g), hlg_inv_oetf(hg.b)};
  else if (p.hdr_tf == 2) hl = {pq_inv_oetf(hg.r), pq_inv_oetf(hg.g), pq_inv_oetf(hg.b)};
  else hl = hg;
  if (p.use_gamut) hl = mat3_apply(p.gamut, hl);
  const float hdr_nits = luminance(lk, hl) * p.hdr_white_nits;

  // ---- encodeGain (gainmapmath.cpp:529-541), truncating ----
  float gain = 1.0f;
  if (sdr_nits > 0.0f) gain = __fdiv_rn(hdr_nits, sdr_nits);
  float v = __fdiv_rn(mufu_lg2(gain) - p.log2_min, p.log2_max - p.log2_min) * 255.0f;
  uint32_t code = __float2uint_rz(v);
  // A clamped gain is the common case (saturated map bytes); there the reference's double-precision
  // quotient against the float-rounded log2 bounds decides between e.g. 254 and 255, so those two
  // codes come from the host, computed exactly as the reference does.
  if (gain <= p.min_boost) code = p.code_at_min;
  if (gain >= p.max_boost) code = p.code_at_max;
  ((uint8_t*)pl.out)[(size_t)my * g.map_width + mx] = (uint8_t)code;
}

bool aligned(const void* q, size_t a) { return (reinterpret_cast<uintptr_t>(q) & (a - 1)) == 0; }
}  // namespace

bool generate_vec_ok(const Geometry& g, const Planes& pl) {
  return aligned(pl.hdr_y, 8) && aligned(pl.hdr_uv, 8) && aligned(pl.sdr_y, 4) && aligned(pl.sdr_u, 2) &&
         aligned(pl.sdr_v, 2) && g.hdr_luma_stride % 4 == 0 && g.hdr_chroma_stride % 4 == 0 &&
         g.sdr_luma_stride % 4 == 0 && g.sdr_chroma_stride % 2 == 0;
}

cudaError_t launch_generate_impl(const GenerateParams& p, int n, bool vec, cudaStream_t s) {
  if (p.g.map_width == 0 || p.g.map_height == 0 || n == 0) return cudaSuccess;
  dim3 block(32, 8, 1);
  dim3 grid((p.g.map_width + 31) / 32, (p.g.map_height + 7) / 8, n);
  if (vec) generate_kernel<true><<<grid, block, 0, s>>>(p);
  else generate_kernel<false><<<grid, block, 0, s>>>(p);
  return cudaGetLastError();
}

}  // namespace uhdr
