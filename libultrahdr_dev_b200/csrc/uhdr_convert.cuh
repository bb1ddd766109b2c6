// uhdr_convert.cuh — transformYuv420 (reference lib/src/gainmapmath.cpp:483-520) on one 2x2 quad,
// shared by convert_yuv.cu (row N1) and the fused API-0 kernel in generate.cu (row N2).
//
// Bit exact against the reference's x86-64 build (files are compiled with -fmad=false, so a * b + c
// keeps its two roundings). All six YUV->YUV matrices (gainmapmath.cpp:447-481) have the first
// column (1, 0, 0): 1 * y is exact and 0 * y adds +0, so the chroma outputs do not depend on the
// pixel's luma and are evaluated once per quad; the reference's running sum of four equal values
// and the division by 4 are kept as written (3 * nu is not always representable).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace uhdr {

__device__ __forceinline__ uint32_t clip_u8(float v) {  // static_cast<uint8_t>(CLIP3(v, 0, 255))
  v = v < 0.0f ? 0.0f : (v > 255.0f ? 255.0f : v);
  return __float2uint_rz(v);
}

struct Quad { uint32_t y[2][2]; uint32_t u, v; };

// m[0] == 1, m[3] == 0, m[6] == 0 (checked on the host, capi.cu)
__device__ __forceinline__ void transform_quad(const float* m, Quad& q) {
  const float uf = (float)((int)q.u - 128) * (1 / 255.0f), vf = (float)((int)q.v - 128) * (1 / 255.0f);
  const float yu = m[1] * uf, yv = m[2] * vf;
  const float nu = m[4] * uf + m[5] * vf, nv = m[7] * uf + m[8] * vf;
#pragma unroll
  for (int r = 0; r < 2; r++)
#pragma unroll
    for (int c = 0; c < 2; c++) {
      const float yf = (float)q.y[r][c] * (1 / 255.0f);
      const float ny = (yf + yu) + yv;  // m[0]*y + m[1]*u + m[2]*v, left to right
      q.y[r][c] = clip_u8(ny * 255.0f + 0.5f);
    }
  const float su = ((nu + nu) + nu) + nu, sv = ((nv + nv) + nv) + nv;  // (yuv1 + yuv2 + yuv3 + yuv4), gainmapmath.cpp:499-500
  q.u = clip_u8(su / 4.0f * 255.0f + 128.0f + 0.5f);
  q.v = clip_u8(sv / 4.0f * 255.0f + 128.0f + 0.5f);
}

}  // namespace uhdr
