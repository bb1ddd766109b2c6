// uhdr_math.cuh — device-side pixel math of the gain-map hot path (sm_100a).
//
// Device restatement of the reference's L1 library (lib/src/gainmapmath.cpp,
// lib/include/ultrahdr/gainmapmath.h); each function cites the lines it follows. The reference
// evaluates its transcendentals in double (glibc) and narrows to float; here they are fp32
// built from MUFU lg2/ex2/rsq plus FMA-pipe polynomials, formulated so that the ill-conditioned
// spots (PQ's x^0.0127 and (.)^78.84) stay accurate to ~1e-6 relative. All files are compiled
// with -fmad=false so that plain a*b+c keeps the reference's two roundings; fmaf() is written
// explicitly where contraction is wanted.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace uhdr {

struct Color3 { float r, g, b; };  // (r,g,b) or (y,u,v): reference `Color`, gainmapmath.h:38-51

// ---- MUFU wrappers (one SASS MUFU each; .ftz: all arguments here are normal or zero) ----
__device__ __forceinline__ float mufu_lg2(float x) { float r; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float mufu_ex2(float x) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float mufu_rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float mufu_sqrt(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

__device__ __forceinline__ float clamp01(float v) { return __saturatef(v); }  // gainmapmath.cpp:115-118

// ---- luminance / YUV matrices: gainmapmath.cpp:121-140,177-196,208-248 ----
// gamut index 0 = BT.709, 1 = P3 (BT.601 YUV matrix), 2 = BT.2100
struct YuvCoefs { float cr, gcb, gcr, cb; };
__host__ __device__ inline YuvCoefs yuv_coefs(int m) {
  const float yr[3] = {0.2126f, 0.299f, 0.2627f}, yg[3] = {0.7152f, 0.587f, 0.6780f},
              yb[3] = {0.0722f, 0.114f, 0.0593f};
  const float cb[3] = {1.8556f, 1.772f, 1.8814f}, cr[3] = {1.5748f, 1.402f, 1.4746f};
  YuvCoefs k;
  k.cr = cr[m];
  k.cb = cb[m];
  k.gcb = yb[m] * cb[m] / yg[m];  // float ops, like the reference's static initialisers
  k.gcr = yr[m] * cr[m] / yg[m];
  return k;
}
struct LumaCoefs { float r, g, b; };
__host__ __device__ inline LumaCoefs luma_coefs(int gamut) {
  const float t[3][3] = {{0.2126f, 0.7152f, 0.0722f}, {0.20949f, 0.72160f, 0.06891f}, {0.2627f, 0.6780f, 0.0593f}};
  return {t[gamut][0], t[gamut][1], t[gamut][2]};
}

// gainmapmath.cpp:142-146 / 198-202 / 250-254: separate mul and add roundings (-fmad=false)
__device__ __forceinline__ Color3 yuv_to_rgb(const YuvCoefs& k, Color3 e) {
  Color3 o;
  o.r = clamp01(e.r + k.cr * e.b);
  o.g = clamp01(e.r - k.gcb * e.g - k.gcr * e.b);
  o.b = clamp01(e.r + k.cb * e.g);
  return o;
}
__device__ __forceinline__ float luminance(const LumaCoefs& k, Color3 e) {
  return k.r * e.r + k.g * e.g + k.b * e.b;
}
// Contracted variants for the generate kernel: its result is a truncated 8-bit code of a log2 whose
// inputs already carry the MUFU error (2^-22), so the last-ulp difference of an FMA is immaterial
// and each contraction saves an issue slot in an issue-bound kernel.
__device__ __forceinline__ Color3 yuv_to_rgb_fma(const YuvCoefs& k, Color3 e) {
  Color3 o;
  o.r = clamp01(fmaf(k.cr, e.b, e.r));
  o.g = clamp01(fmaf(-k.gcr, e.b, fmaf(-k.gcb, e.g, e.r)));
  o.b = clamp01(fmaf(k.cb, e.g, e.r));
  return o;
}
__device__ __forceinline__ float luminance_fma(const LumaCoefs& k, Color3 e) {
  return fmaf(k.b, e.b, fmaf(k.g, e.g, k.r * e.r));
}

// ---- sRGB inverse OETF, gainmapmath.cpp:149-155 ----
// The two divisions by constants are multiplications by the double-rounded reciprocals (1 ulp,
// far below the MUFU error): a correctly rounded division costs ~10 instructions each.
__device__ __forceinline__ float srgb_inv_oetf(float x) {
  float t = fmaf(x, (float)(1.0 / 1.055), (float)(0.055 / 1.055));
  float p = mufu_ex2(2.4f * mufu_lg2(t));
  return x <= 0.04045f ? x * (float)(1.0 / 12.92) : p;
}

// ---- HLG, gainmapmath.cpp:257-286 ----
#define UHDR_HLG_A 0.17883277f
#define UHDR_HLG_B 0.28466892f
#define UHDR_HLG_C 0.55991073f
__device__ __forceinline__ float hlg_oetf(float e) {
  float lo = mufu_sqrt(3.0f * e);
  // a*ln(12e-b)+c = (a*ln2)*lg2(12e-b)+c
  float hi = fmaf(UHDR_HLG_A * 0.69314718056f, mufu_lg2(fmaxf(12.0f * e - UHDR_HLG_B, 1e-30f)), UHDR_HLG_C);
  return e <= 1.0f / 12.0f ? lo : hi;
}
__device__ __forceinline__ float hlg_inv_oetf(float x) {
  float lo = x * x * (1.0f / 3.0f);
  float hi = (mufu_ex2((x - UHDR_HLG_C) * (1.44269504089f / UHDR_HLG_A)) + UHDR_HLG_B) * (1.0f / 12.0f);
  return x <= 0.5f ? lo : hi;
}

// ---- PQ, gainmapmath.cpp:305-338 ----
// expm1 on |u| <= 0.12 (Taylor to u^7: next term 2e-12)
__device__ __forceinline__ float expm1_small(float u) {
  float p = 1.0f / 5040.0f;
  p = fmaf(p, u, 1.0f / 720.0f);
  p = fmaf(p, u, 1.0f / 120.0f);
  p = fmaf(p, u, 1.0f / 24.0f);
  p = fmaf(p, u, 1.0f / 6.0f);
  p = fmaf(p, u, 0.5f);
  p = fmaf(p, u, 1.0f);
  return p * u;
}
// pqInvOetf: ((128p-107)/(2413-2392p))^6.2774, p = x^0.0126833. With em = p-1 computed by expm1
// the cancellation in 128p-107 = 21+128em and 2413-2392p = 21-2392em is harmless.
__device__ __forceinline__ float pq_inv_oetf(float x) {
  float u = (0.0126833f * 0.69314718056f) * mufu_lg2(fmaxf(x, 1e-4f));
  float em = expm1_small(u);
  float num = fmaf(128.0f, em, 21.0f), den = fmaf(-2392.0f, em, 21.0f);
  float ratio = fmaxf(num, 0.0f) * mufu_rcp(den);
  float r = mufu_ex2(6.2773946361f * mufu_lg2(ratio));
  return x <= 0.0001f ? 0.0f : r;
}
// pqOetf: ((c1+c2 p)/(1+c3 p))^m2, p = e^m1. Since c2-c3 = 1-c1 = k, the base is
// 1 - k(1-p)/(1+c3 p) =: 1-d, d in [0,0.164]; ln(1-d) = -2 atanh(d/(2-d)) by odd series.
__device__ __forceinline__ float pq_oetf(float e) {
  const float m1 = 2610.0f / 16384.0f, m2 = 2523.0f / 4096.0f * 128.0f;
  const float c3 = 2392.0f / 4096.0f * 32.0f, k = 1.0f - 3424.0f / 4096.0f;
  float p = mufu_ex2(m1 * mufu_lg2(e));
  float d = k * (1.0f - p) * mufu_rcp(fmaf(c3, p, 1.0f));
  float s = d * mufu_rcp(2.0f - d);  // in [0, 0.0894]
  float s2 = s * s;
  float q = 1.0f / 11.0f;
  q = fmaf(q, s2, 1.0f / 9.0f);
  q = fmaf(q, s2, 1.0f / 7.0f);
  q = fmaf(q, s2, 1.0f / 5.0f);
  q = fmaf(q, s2, 1.0f / 3.0f);
  q = fmaf(q, s2, 1.0f);
  float ln1md = -2.0f * s * q;
  float r = mufu_ex2((m2 * 1.44269504089f) * ln1md);
  return e <= 0.0f ? 0.0f : r;
}

// ---- gamut 3x3, gainmapmath.cpp:359-393; index [sdr_gamut][hdr_gamut], identity when equal ----
struct Mat3 { float m[9]; };
__host__ __device__ inline Mat3 gamut_matrix(int to, int from) {
  const float t[3][3][9] = {
      {{1, 0, 0, 0, 1, 0, 0, 0, 1},
       {1.22482f, -0.22490f, -0.00007f, -0.04196f, 1.04199f, 0.00001f, -0.01961f, -0.07865f, 1.09831f},
       {1.66045f, -0.58764f, -0.07286f, -0.12445f, 1.13282f, -0.00837f, -0.01811f, -0.10057f, 1.11878f}},
      {{0.82254f, 0.17755f, 0.00006f, 0.03312f, 0.96684f, -0.00001f, 0.01706f, 0.07240f, 0.91049f},
       {1, 0, 0, 0, 1, 0, 0, 0, 1},
       {1.34369f, -0.28223f, -0.06135f, -0.06533f, 1.07580f, -0.01051f, 0.00283f, -0.01957f, 1.01679f}},
      {{0.62740f, 0.32930f, 0.04332f, 0.06904f, 0.91958f, 0.01138f, 0.01636f, 0.08799f, 0.89555f},
       {0.75378f, 0.19862f, 0.04754f, 0.04576f, 0.94177f, 0.01250f, -0.00121f, 0.01757f, 0.98359f},
       {1, 0, 0, 0, 1, 0, 0, 0, 1}}};
  Mat3 r;
  for (int i = 0; i < 9; i++) r.m[i] = t[to][from][i];
  return r;
}
__device__ __forceinline__ Color3 mat3_apply_fma(const Mat3& k, Color3 e) {
  return {fmaf(k.m[2], e.b, fmaf(k.m[1], e.g, k.m[0] * e.r)), fmaf(k.m[5], e.b, fmaf(k.m[4], e.g, k.m[3] * e.r)),
          fmaf(k.m[8], e.b, fmaf(k.m[7], e.g, k.m[6] * e.r))};
}
__device__ __forceinline__ Color3 mat3_apply(const Mat3& k, Color3 e) {
  return {k.m[0] * e.r + k.m[1] * e.g + k.m[2] * e.b, k.m[3] * e.r + k.m[4] * e.g + k.m[5] * e.b,
          k.m[6] * e.r + k.m[7] * e.g + k.m[8] * e.b};
}

// ---- encodeGain, gainmapmath.cpp:524-541 (truncating). The reference evaluates log2 in double against float-rounded
// log2 bounds; a gain clamped to min / max (the common case: saturated map bytes) takes its code from the host, which
// computes it exactly as the reference does (capi.cu: gain_code_of).
struct GainCode {
  float min_boost, max_boost, log2_min;
  float log2_max;                // LUT numeric mode only (encode_gain_exact)
  float code_scale;              // 255 / (log2_max - log2_min), the float difference like the reference, in double
  int code_at_min, code_at_max;  // encodeGain of a gain clamped to min / max
};
__device__ __forceinline__ uint32_t encode_gain(float sdr_nits, float hdr_nits, const GainCode& k) {
  float gain = 1.0f;
  if (sdr_nits > 0.0f) gain = hdr_nits * mufu_rcp(sdr_nits);
  const float v = (mufu_lg2(gain) - k.log2_min) * k.code_scale;
  uint32_t code = __float2uint_rz(v);
  if (gain <= k.min_boost) code = (uint32_t)k.code_at_min;
  if (gain >= k.max_boost) code = (uint32_t)k.code_at_max;
  return code;
}

// LUT numeric mode: the reference's own chain, float quotient, clamp, then double log2 / quotient / * 255 and a
// truncating cast (gainmapmath.cpp:529-541). libdevice's double log2 is within 1 ulp of glibc's, so the truncated code
// can differ only when the double product sits within ~1e-16 of an integer.
__device__ __forceinline__ uint32_t encode_gain_exact(float sdr_nits, float hdr_nits, const GainCode& k) {
  float gain = 1.0f;
  if (sdr_nits > 0.0f) gain = __fdiv_rn(hdr_nits, sdr_nits);
  if (gain < k.min_boost) gain = k.min_boost;
  if (gain > k.max_boost) gain = k.max_boost;
  const double v = (log2((double)gain) - (double)k.log2_min) / (double)(k.log2_max - k.log2_min) * 255.0;
  return (uint32_t)(uint8_t)(int)v;
}
// nearest-entry LUT index, `static_cast<uint32_t>(x * (N - 1) + 0.5)` then CLIP3 (gainmapmath.cpp:162-167): the float
// product, + 0.5 in double (exact for a product below 2^23), truncation. x >= 0 here (clamped colours, IDW gains).
__device__ __forceinline__ uint32_t lut_index(float x, int n) {
  const float pos = x * (float)(n - 1);
  const float fl = truncf(pos);
  uint32_t idx = __float2uint_rz(fl) + ((pos - fl) >= 0.5f ? 1u : 0u);
  return idx > (uint32_t)(n - 1) ? (uint32_t)(n - 1) : idx;
}

// ---- packers ----
// floatToHalf, gainmapmath.h:136-147: bit-exact integer routine (+0x1000 rounding, saturate to
// 0x7FFF, flush e<=101 to zero).
__device__ __forceinline__ uint32_t float_to_half_ref(float f) {
  const uint32_t b = __float_as_uint(f) + 0x00001000u;
  const int e = (int)((b & 0x7F800000u) >> 23);
  const uint32_t m = b & 0x007FFFFFu;
  uint32_t r = (b & 0x80000000u) >> 16;
  if (e > 112) r |= (((uint32_t)(e - 112) << 10) & 0x7C00u) | (m >> 13);
  if (e < 113 && e > 101) r |= (((0x007FF000u + m) >> (125 - e)) + 1u) >> 1;
  if (e > 143) r |= 0x7FFFu;
  return r;
}
// colorToRgba1010102 channel, gainmapmath.cpp:722-727: truncation, then & 0x3ff
__device__ __forceinline__ uint32_t to_10bit(float v) { return 0x3ffu & (uint32_t)__float2uint_rz(v * 1023.0f); }

}  // namespace uhdr
