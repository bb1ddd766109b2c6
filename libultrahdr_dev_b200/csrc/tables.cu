// tables.cu — host-side construction (double precision) of the shared-memory transfer-function
// tables used by apply_fast.cu. Each table is piecewise linear on log-spaced segments (a float's
// exponent + top mantissa bits pick the segment); per segment the line interpolates the target
// at the two Chebyshev nodes, which halves the maximum error of an end-point chord.
//
// Targets follow the reference's formulas evaluated in double, as the reference itself does under
// g++/glibc (DESIGN.md "numerics"): srgbInvOetf gainmapmath.cpp:149-155, hlgOetf :259-267,
// pqOetf :309-314. Measured against the oracle the tables give >= 99.9 % exact 10-bit codes and
// never more than 1 LSB (tests/test_gpu_parity.py).
#include <cmath>
#include <cstdint>
#include <cstring>

#include "uhdr_kernels.cuh"

namespace uhdr {
namespace {

double hlg_oetf_d(double e) {
  const double a = 0.17883277, b = 0.28466892, c = 0.55991073;
  if (e <= 1.0 / 12.0) return sqrt(3.0 * (e < 0 ? 0 : e));
  return a * log(12.0 * e - b) + c;
}
double pq_oetf_d(double e) {
  if (e <= 0.0) return 0.0;
  const double m1 = 2610.0 / 16384.0, m2 = 2523.0 / 4096.0 * 128.0;
  const double c1 = 3424.0 / 4096.0, c2 = 2413.0 / 4096.0 * 32.0, c3 = 2392.0 / 4096.0 * 32.0;
  const double p = pow(e, m1);
  return pow((c1 + c2 * p) / (1.0 + c3 * p), m2);
}

float bits_to_float(uint32_t b) {
  float f;
  memcpy(&f, &b, 4);
  return f;
}

template <class F>
void fill(float2* tab, int lo_exp, int hi_exp, int bits, F fn, bool nonneg = false) {
  const int n = (hi_exp - lo_exp) << bits;
  const uint32_t step = 1u << (23 - bits);
  const double k = sqrt(2.0) / 4.0;
  for (int i = 0; i < n; i++) {
    const uint32_t b0 = ((uint32_t)(lo_exp + 127) << 23) + (uint32_t)i * step;
    const double x0 = bits_to_float(b0), x1 = bits_to_float(b0 + step), h = x1 - x0;
    const double n0 = x0 + h * (0.5 - k), n1 = x0 + h * (0.5 + k);
    const double f0 = fn(n0), f1 = fn(n1);
    const double slope = (f1 - f0) / (n1 - n0);
    tab[i].x = (float)(f0 - slope * n0);
    tab[i].y = (float)slope;
    if (nonneg) {  // lift the line so that it is >= 0 at both ends of its segment (first segments only)
      const double lo = fmin((double)tab[i].x + (double)tab[i].y * x0, (double)tab[i].x + (double)tab[i].y * x1);
      if (lo < 0) {
        float a = (float)((double)tab[i].x - lo);
        while ((double)a + (double)tab[i].y * x0 < 0 || (double)a + (double)tab[i].y * x1 < 0) a = nextafterf(a, 1e30f);
        tab[i].x = a;
      }
    }
  }
}

double srgb_inv_d(double c) { return c <= 0.04045 ? c / 12.92 : pow((c + 0.055) / 1.055, 2.4); }

// S table on uniform segments of c in [0,1]. Segments wholly inside the linear part of
// srgbInvOetf hold the exact line (a = 0, so c = 0 gives exactly +0: floatToHalf,
// gainmapmath.h:136-147, would turn a negative zero-ish value into 0x8000).
void fill_s(float2* tab, int segs, int entries) {
  const double k = sqrt(2.0) / 4.0, h = 1.0 / segs;
  for (int i = 0; i < segs; i++) {
    const double x0 = i * h, x1 = (i + 1) * h;
    if (x1 <= 0.04045) {
      tab[i].x = 0.0f;
      tab[i].y = (float)(1.0 / 12.92);
      continue;
    }
    const double n0 = x0 + h * (0.5 - k), n1 = x0 + h * (0.5 + k);
    const double f0 = srgb_inv_d(n0), f1 = srgb_inv_d(n1);
    const double slope = (f1 - f0) / (n1 - n0);
    tab[i].x = (float)(f0 - slope * n0);
    tab[i].y = (float)slope;
  }
  for (int i = segs; i < entries; i++) tab[i] = tab[segs - 1];  // c == 1.0 lands on index `segs`
}

}  // namespace

void build_apply_tables(float2* s, float2* s_hi, float2* o_hlg, float2* o_pq, float2* p_hlg, float2* p_pq, float2* l_pow, float* e255) {
  fill_s(s, kSSegs, kSEntries);
  fill_s(s_hi, kSSegsHi, kSEntriesHi);
  const double e0 = ldexp(1.0, kOLoExp);
  fill(o_hlg, kOLoExp, kOHiExp, kOBits, [e0](double x) { return 1023.0 * hlg_oetf_d(x - e0); }, true);
  fill(o_pq, kOLoExp, kOHiExp, kOBits, [e0](double x) { return 1023.0 * pq_oetf_d(x - e0); }, true);
  // q-domain tables: argument q~ = q + q0, value 1023 * OETF(q^2.4)
  const double q0 = ldexp(1.0, kPLoExp);
  auto q24 = [q0](double x) { return x > q0 ? pow(x - q0, 2.4) : 0.0; };
  fill(p_hlg, kPLoExp, kPHiExp, kPBits, [q24](double x) { return 1023.0 * hlg_oetf_d(q24(x)); }, true);
  fill(p_pq, kPLoExp, kPHiExp, kPBits, [q24](double x) { return 1023.0 * pq_oetf_d(q24(x)); }, true);
  fill(l_pow, kLLoExp, kLHiExp, kLBits, [](double x) { return pow(x, 2.4); });
  for (int i = 0; i < 256; i++) e255[i] = (float)i / 255.0f;  // gainmapmath.cpp:700-703
}

// ---------------------------------------------------------------- LUT numeric mode
// The nearest-entry tables of the reference's -DUSE_*_LUT=1 build (gainmapmath.cpp:21-64): entry i holds the analytic
// function at (float)i / (float)(N - 1), evaluated the way g++ / glibc evaluate the reference's expressions: float operands,
// the transcendental in double (pow / exp / log take the double overload because one operand is a double literal or the
// float is promoted), narrowed to float on return. This file is host code compiled by the same g++ / glibc.
namespace {
float ref_srgb_inv(float x) {  // gainmapmath.cpp:149-155
  if (x <= 0.04045f) return x / 12.92f;
  return (float)pow((double)((x + 0.055f) / 1.055f), 2.4);
}
const float kHlgA = 0.17883277f, kHlgB = 0.28466892f, kHlgC = 0.55991073f;  // gainmapmath.cpp:257
float ref_hlg_oetf(float e) {  // :259-265
  if (e <= 1.0f / 12.0f) return sqrtf(3.0f * e);
  return (float)((double)kHlgA * log((double)(12.0f * e - kHlgB)) + (double)kHlgC);
}
float ref_hlg_inv(float x) {  // :280-286
  if (x <= 0.5f) return (float)(pow((double)x, 2.0) / (double)3.0f);
  return (float)((exp((double)((x - kHlgC) / kHlgA)) + (double)kHlgB) / (double)12.0f);
}
const float kPqM1 = 2610.0f / 16384.0f, kPqM2 = 2523.0f / 4096.0f * 128.0f;  // :305-307
const float kPqC1 = 3424.0f / 4096.0f, kPqC2 = 2413.0f / 4096.0f * 32.0f, kPqC3 = 2392.0f / 4096.0f * 32.0f;
float ref_pq_oetf(float e) {  // :309-312
  if (e <= 0.0f) return 0.0f;
  const double p = pow((double)e, (double)kPqM1);
  return (float)pow(((double)kPqC1 + (double)kPqC2 * p) / (1 + (double)kPqC3 * p), (double)kPqM2);
}
float ref_pq_inv(float x) {  // :327-338
  if (x <= 0.0001f) return 0.0f;
  const double p = pow((double)x, (double)0.0126833f);
  return (float)pow(((double)128.0f * p - (double)107.0f) / ((double)2413.0f - (double)2392.0f * p), (double)6.2773946361f);
}
void fill_lut(float* t, int n, float (*fn)(float)) {
  for (int i = 0; i < n; i++) t[i] = fn((float)i / (float)(n - 1));
}
}  // namespace

void build_lut_tables(float* srgb_inv, float* hlg_inv, float* pq_inv, float* hlg_oetf, float* pq_oetf) {
  fill_lut(srgb_inv, kLutSrgbInv, ref_srgb_inv);
  fill_lut(hlg_inv, kLutHlgInv, ref_hlg_inv);
  fill_lut(pq_inv, kLutPqInv, ref_pq_inv);
  fill_lut(hlg_oetf, kLutHlgOetf, ref_hlg_oetf);
  fill_lut(pq_oetf, kLutPqOetf, ref_pq_oetf);
}

// GainLUT(metadata, displayBoost), gainmapmath.h:161-169: log2 of the float boosts in double, the blend narrowed to
// float, `logBoost * boostFactor` in float, exp2 in double narrowed to float.
void build_gain_lut(float* tab, float min_boost, float max_boost, float display_boost) {
  const float boost_factor = display_boost > 0 ? display_boost / max_boost : 1.0f;
  for (int i = 0; i < kLutGain; i++) {
    const float v = (float)i / (float)(kLutGain - 1);
    const float log_boost = (float)(log2((double)min_boost) * (double)(1.0f - v) + log2((double)max_boost) * (double)v);
    tab[i] = (float)exp2((double)(log_boost * boost_factor));
  }
}

}  // namespace uhdr
