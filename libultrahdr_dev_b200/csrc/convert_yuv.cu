// convert_yuv.cu — N1: JpegR::convertYuv (reference lib/src/jpegr.cpp:1132-1206) +
// transformYuv420 (lib/src/gainmapmath.cpp:483-520) on sm_100a: in-place re-encode of a YUV420
// image between the BT.709 / BT.601 / BT.2100 YUV matrices, per 2x2 quad: each of the 4 pixels is
// transformed with the quad's (U,V), luma is written back rounded (+0.5, clip), chroma gets the
// average of the 4 transformed (U,V).
//
// One thread = 4 px x 4 rows (4-aligned planes; packed fp32x2 quad-pair transform) or two horizontally adjacent
// quads (4 px x 2 rows, scalar fallback): 32-bit luma loads/stores, 16-bit chroma loads/stores. In place, so plain (coherent) loads; every byte is read and written by the
// same thread only. HBM-bound: 1.5 B/px read + 1.5 B/px written.
#include "uhdr_convert.cuh"
#include "uhdr_kernels.cuh"
#include "uhdr_math.cuh"

namespace uhdr {
namespace {

template <bool VEC>
__global__ void __launch_bounds__(256) convert_yuv_kernel(ConvertYuvParams p) {
  const Geometry& g = p.g;
  const int x0 = (blockIdx.x * 32 + threadIdx.x) * 4;
  const int cw = g.width / 2, ch = g.height / 2;
  const Planes pl = p.batch ? p.batch[blockIdx.z] : p.single;
  if (VEC) {
    // thread = 4 px x 4 rows = two quad pairs on packed fp32x2 lanes (uhdr_convert.cuh), same roundings; all eight
    // loads are issued before the first transform
    const int y0 = (blockIdx.y * 8 + threadIdx.y) * 4;
    if ((x0 >> 1) >= cw || (y0 >> 1) >= ch) return;
    uint8_t* yp = (uint8_t*)pl.sdr_y + (size_t)y0 * g.sdr_luma_stride + x0;
    uint8_t* up = (uint8_t*)pl.sdr_u + (size_t)(y0 >> 1) * g.sdr_chroma_stride + (x0 >> 1);
    uint8_t* vp = (uint8_t*)pl.sdr_v + (size_t)(y0 >> 1) * g.sdr_chroma_stride + (x0 >> 1);
    const bool two = (y0 >> 1) + 1 < ch;
    uint32_t a[4] = {0, 0, 0, 0}, c[2] = {0, 0}, d[2] = {0, 0};
#pragma unroll
    for (int r = 0; r < 4; r++)
      if (r < 2 || two) a[r] = *reinterpret_cast<uint32_t*>(yp + (size_t)r * g.sdr_luma_stride);
#pragma unroll
    for (int r = 0; r < 2; r++)
      if (r < 1 || two) {
        c[r] = *reinterpret_cast<uint16_t*>(up + (size_t)r * g.sdr_chroma_stride);
        d[r] = *reinterpret_cast<uint16_t*>(vp + (size_t)r * g.sdr_chroma_stride);
      }
#pragma unroll
    for (int r = 0; r < 2; r++) {
      if (r == 1 && !two) break;
      transform_quad_pair(p.m, a[2 * r], a[2 * r + 1], c[r], d[r]);
      *reinterpret_cast<uint32_t*>(yp + (size_t)(2 * r) * g.sdr_luma_stride) = a[2 * r];
      *reinterpret_cast<uint32_t*>(yp + (size_t)(2 * r + 1) * g.sdr_luma_stride) = a[2 * r + 1];
      *reinterpret_cast<uint16_t*>(up + (size_t)r * g.sdr_chroma_stride) = (uint16_t)c[r];
      *reinterpret_cast<uint16_t*>(vp + (size_t)r * g.sdr_chroma_stride) = (uint16_t)d[r];
    }
    return;
  }
  const int y0 = (blockIdx.y * 8 + threadIdx.y) * 2;
  if ((x0 >> 1) >= cw || (y0 >> 1) >= ch) return;
  uint8_t* yp = (uint8_t*)pl.sdr_y + (size_t)y0 * g.sdr_luma_stride + x0;
  uint8_t* up = (uint8_t*)pl.sdr_u + (size_t)(y0 >> 1) * g.sdr_chroma_stride + (x0 >> 1);
  uint8_t* vp = (uint8_t*)pl.sdr_v + (size_t)(y0 >> 1) * g.sdr_chroma_stride + (x0 >> 1);
  const int nq = VEC ? 2 : min(2, cw - (x0 >> 1));
  Quad q[2];
  if (VEC) {
    uint32_t a = *reinterpret_cast<uint32_t*>(yp), b = *reinterpret_cast<uint32_t*>(yp + g.sdr_luma_stride);
    uint32_t c = *reinterpret_cast<uint16_t*>(up), d = *reinterpret_cast<uint16_t*>(vp);
#pragma unroll
    for (int k = 0; k < 2; k++) {
      q[k].y[0][0] = (a >> (16 * k)) & 0xff; q[k].y[0][1] = (a >> (16 * k + 8)) & 0xff;
      q[k].y[1][0] = (b >> (16 * k)) & 0xff; q[k].y[1][1] = (b >> (16 * k + 8)) & 0xff;
      q[k].u = (c >> (8 * k)) & 0xff; q[k].v = (d >> (8 * k)) & 0xff;
    }
  } else {
    for (int k = 0; k < nq; k++) {
      for (int r = 0; r < 2; r++)
        for (int c = 0; c < 2; c++) q[k].y[r][c] = yp[(size_t)r * g.sdr_luma_stride + 2 * k + c];
      q[k].u = up[k]; q[k].v = vp[k];
    }
  }
#pragma unroll
  for (int k = 0; k < 2; k++) if (k < nq) transform_quad(p.m, q[k]);
  if (VEC) {
    uint32_t a = q[0].y[0][0] | (q[0].y[0][1] << 8) | (q[1].y[0][0] << 16) | (q[1].y[0][1] << 24);
    uint32_t b = q[0].y[1][0] | (q[0].y[1][1] << 8) | (q[1].y[1][0] << 16) | (q[1].y[1][1] << 24);
    *reinterpret_cast<uint32_t*>(yp) = a;
    *reinterpret_cast<uint32_t*>(yp + g.sdr_luma_stride) = b;
    *reinterpret_cast<uint16_t*>(up) = (uint16_t)(q[0].u | (q[1].u << 8));
    *reinterpret_cast<uint16_t*>(vp) = (uint16_t)(q[0].v | (q[1].v << 8));
  } else {
    for (int k = 0; k < nq; k++) {
      for (int r = 0; r < 2; r++)
        for (int c = 0; c < 2; c++) yp[(size_t)r * g.sdr_luma_stride + 2 * k + c] = (uint8_t)q[k].y[r][c];
      up[k] = (uint8_t)q[k].u; vp[k] = (uint8_t)q[k].v;
    }
  }
}

bool aligned(const void* q, size_t a) { return (reinterpret_cast<uintptr_t>(q) & (a - 1)) == 0; }
}  // namespace

bool convert_yuv_vec_ok(const Geometry& g, const Planes& pl) {
  return g.width % 4 == 0 && aligned(pl.sdr_y, 4) && aligned(pl.sdr_u, 2) && aligned(pl.sdr_v, 2) &&
         g.sdr_luma_stride % 4 == 0 && g.sdr_chroma_stride % 2 == 0;
}

cudaError_t launch_convert_yuv_impl(const ConvertYuvParams& p, int n, bool vec, cudaStream_t s) {
  if (p.g.width < 2 || p.g.height < 2 || n == 0) return cudaSuccess;
  dim3 block(32, 8, 1);
  if (vec) {
    dim3 grid((p.g.width + 127) / 128, (p.g.height + 31) / 32, n);  // thread = 4 x 4 pixels
    convert_yuv_kernel<true><<<grid, block, 0, s>>>(p);
  } else {
    dim3 grid((p.g.width + 127) / 128, (p.g.height + 15) / 16, n);  // thread = 4 x 2 pixels
    convert_yuv_kernel<false><<<grid, block, 0, s>>>(p);
  }
  return cudaGetLastError();
}

}  // namespace uhdr
