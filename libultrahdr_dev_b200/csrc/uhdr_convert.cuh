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

// ---- packed fp32x2 version for 4-aligned blocks (fused API-0 kernel, generate.cu) -------------------------------
// Same arithmetic, same roundings: the int -> float conversions are I2F.U8 with a byte selector (the XU pipe is idle in
// these kernels; byte-indexed shared-memory tables cost 3 issue slots per sample and were 8-13 % slower), the * (1 / 255),
// sums and products run on FADD2 / FMUL2 lanes (IEEE per lane), and
// CLIP3 + truncation is one saturating float -> u8 conversion (F2IP.U8.F32.TRUNC). ptxas 12.9 contracts a
// mul.rn.f32x2 feeding an add.rn.f32x2 into FFMA2 even under -fmad=false, so every "product + term" below adds
// through fma(product, 1, term), which is the same two-rounding result.
typedef unsigned long long cv_f32x2;
__device__ __forceinline__ cv_f32x2 cv_pk2(float lo, float hi) {
  cv_f32x2 r;
  asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void cv_unpk2(cv_f32x2 v, float& lo, float& hi) { asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ cv_f32x2 cv_add2(cv_f32x2 a, cv_f32x2 b) {
  cv_f32x2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ cv_f32x2 cv_mul2(cv_f32x2 a, cv_f32x2 b) {
  cv_f32x2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ cv_f32x2 cv_sum2(cv_f32x2 prod, cv_f32x2 term) {  // rn(prod + term), not contractible
  cv_f32x2 r;
  const cv_f32x2 one = cv_pk2(1.0f, 1.0f);
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(prod), "l"(one), "l"(term));
  return r;
}
__device__ __forceinline__ uint32_t cv_sat_u8(float v) {
  uint32_t r;
  asm("cvt.rzi.u8.f32 %0, %1;" : "=r"(r) : "f"(v));
  return r;
}
// two pixel rows x four columns = two horizontally adjacent quads, in place on the packed words:
// y0w / y1w = the 4 luma bytes of each row, uw / vw = U0 U1 / V0 V1 in the low 16 bits.
__device__ __forceinline__ void transform_quad_pair(const float* m, uint32_t& y0w, uint32_t& y1w,
                                                    uint32_t& uw, uint32_t& vw) {
  const cv_f32x2 k255 = cv_pk2(255.0f, 255.0f), half = cv_pk2(0.5f, 0.5f), k128 = cv_pk2(128.0f, 128.0f), quarter = cv_pk2(0.25f, 0.25f);
  const cv_f32x2 m47 = cv_pk2(m[4], m[7]), m58 = cv_pk2(m[5], m[8]), k255inv = cv_pk2(1 / 255.0f, 1 / 255.0f);
  uint32_t yo[2][4], uo[2], vo[2];
#pragma unroll
  for (int qc = 0; qc < 2; qc++) {
    // (float)(c - 128) * (1 / 255.0f): the int -> float conversion of c - 128 equals (float)c - 128 exactly
    float uf, vf, yu, yv;
    const cv_f32x2 uv = cv_mul2(cv_add2(cv_pk2((float)((uw >> (8 * qc)) & 0xffu), (float)((vw >> (8 * qc)) & 0xffu)), cv_pk2(-128.0f, -128.0f)), k255inv);
    cv_unpk2(uv, uf, vf);
    cv_unpk2(cv_mul2(uv, cv_pk2(m[1], m[2])), yu, yv);
    const cv_f32x2 YU = cv_pk2(yu, yu), YV = cv_pk2(yv, yv);
    // lanes = (nu, nv) = (m4*u + m5*v, m7*u + m8*v); then the reference's running sum of four equal values, / 4
    const cv_f32x2 n = cv_sum2(cv_mul2(m58, cv_pk2(vf, vf)), cv_mul2(m47, cv_pk2(uf, uf)));
    const cv_f32x2 s4 = cv_add2(cv_add2(cv_add2(n, n), n), n);
    const cv_f32x2 c = cv_add2(cv_sum2(cv_mul2(cv_mul2(s4, quarter), k255), k128), half);  // su / 4 * 255 + 128 + 0.5
    float cu, cvv;
    cv_unpk2(c, cu, cvv);
    uo[qc] = cv_sat_u8(cu);
    vo[qc] = cv_sat_u8(cvv);
#pragma unroll
    for (int r = 0; r < 2; r++) {
      const uint32_t w = r ? y1w : y0w;
      // (float)y * (1 / 255.0f) as transform_quad computes it: the byte -> float conversions select their byte themselves
      // (I2F.U8 .Bn on the otherwise idle XU pipe), the products share one FMUL2 per pixel pair
      const cv_f32x2 yf = cv_mul2(cv_pk2((float)((w >> (16 * qc)) & 0xffu), (float)((w >> (16 * qc + 8)) & 0xffu)), k255inv);
      const cv_f32x2 ny = cv_add2(cv_add2(yf, YU), YV);  // m[0]*y + m[1]*u + m[2]*v, left to right
      float a0, a1;
      cv_unpk2(cv_sum2(cv_mul2(ny, k255), half), a0, a1);
      yo[r][2 * qc] = cv_sat_u8(a0);
      yo[r][2 * qc + 1] = cv_sat_u8(a1);
    }
  }
  y0w = __byte_perm(__byte_perm(yo[0][0], yo[0][1], 0x0040), __byte_perm(yo[0][2], yo[0][3], 0x0040), 0x5410);
  y1w = __byte_perm(__byte_perm(yo[1][0], yo[1][1], 0x0040), __byte_perm(yo[1][2], yo[1][3], 0x0040), 0x5410);
  uw = __byte_perm(uo[0], uo[1], 0x0040);
  vw = __byte_perm(vo[0], vo[1], 0x0040);
}

}  // namespace uhdr
