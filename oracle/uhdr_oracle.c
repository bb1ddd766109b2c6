/*
 * uhdr_oracle.c — TEST INFRASTRUCTURE ONLY. CPU restatement, in plain C, of the reference's
 * gain-map hot path. Nothing in the product library may include, link or call this file; only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg use it, as the checker.
 *
 * Parity status: PINNED. tests/test_oracle_vs_ref.py compares every function below, bit for bit,
 * with the unmodified reference compiled by oracle/Makefile (oracle/_ref/libuhdr_ref.so) and with
 * the committed golden fixtures generated from it (tests/golden/), and tests/test_oracle_kat.py
 * checks the known-answer values of the reference's tests/gainmapmath_test.cpp.
 *
 * Numerical character that is restated on purpose (see DESIGN.md "oracle numerics"):
 * the reference calls unqualified pow/exp/exp2/log/log2 from namespace ultrahdr, which under
 * g++/glibc bind to the DOUBLE overloads; mixed float/double expressions are therefore evaluated
 * in double and narrowed to float only where the reference stores into a float. Every such spot
 * is marked "double" below. Plain float expressions stay float (x86-64 SSE, no FMA contraction:
 * this file is built with -ffp-contract=off -march=x86-64).
 *
 * Citations are file:line in the reference tree (lib/src/gainmapmath.cpp = "gmm.cpp",
 * lib/include/ultrahdr/gainmapmath.h = "gmm.h", lib/src/ultrahdr.cpp = "uhdr.cpp").
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "uhdr_b200.h"

typedef struct { float a, b, c; } tri; /* (r,g,b) or (y,u,v) — reference `Color`, gmm.h:38-51 */

static const float SDR_WHITE_NITS = 203.0f; /* gmm.h:34 */
static const float HLG_MAX_NITS = 1000.0f;  /* gmm.h:35 */
static const float PQ_MAX_NITS = 10000.0f;  /* gmm.h:36 */

/* ------------------------------------------------------------------ helpers */
static float clamp01(float v) { return v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v); } /* gmm.cpp:115-118 */

/* ------------------------------------------------------------------ luminance, gmm.cpp:121-123,177-179,208-210 */
static const float LUMA_COEF[3][3] = {{0.2126f, 0.7152f, 0.0722f},
                                      {0.20949f, 0.72160f, 0.06891f},
                                      {0.2627f, 0.6780f, 0.0593f}};
float orc_luminance(int gamut, float r, float g, float b) {
  const float* k = LUMA_COEF[gamut < 0 || gamut > 2 ? 2 : gamut];
  return k[0] * r + k[1] * g + k[2] * b;
}

/* ------------------------------------------------------------------ YUV<->RGB, gmm.cpp:129-146,184-202,215-254
 * gamut 0: BT.709 matrix, 1: BT.601 matrix (Display-P3), 2: BT.2100 matrix.
 * Luma weights used for Y' (601 differs from the P3 luminance weights, gmm.cpp:184). */
static const float YW[3][3] = {{0.2126f, 0.7152f, 0.0722f}, {0.299f, 0.587f, 0.114f}, {0.2627f, 0.6780f, 0.0593f}};
static const float CB[3] = {1.8556f, 1.772f, 1.8814f};
static const float CR[3] = {1.5748f, 1.402f, 1.4746f};

static void g_coefs(int m, float* gcb, float* gcr) {
  /* float products/quotients, as the reference's `static const float` initialisers
   * (gmm.cpp:139-140, 195-196, 247-248): B*Cb/G and R*Cr/G */
  *gcb = YW[m][2] * CB[m] / YW[m][1];
  *gcr = YW[m][0] * CR[m] / YW[m][1];
}

static tri yuv_to_rgb(int m, tri e) {
  float gcb, gcr;
  g_coefs(m, &gcb, &gcr);
  tri o;
  o.a = clamp01(e.a + CR[m] * e.c);
  o.b = clamp01(e.a - gcb * e.b - gcr * e.c);
  o.c = clamp01(e.a + CB[m] * e.b);
  return o;
}
static tri rgb_to_yuv(int m, tri e) { /* gmm.cpp:131-134,187-190,217-220 */
  float y = YW[m][0] * e.a + YW[m][1] * e.b + YW[m][2] * e.c;
  tri o = {y, (e.c - y) / CB[m], (e.a - y) / CR[m]};
  return o;
}
void orc_yuvToRgb(int gamut, const float* yuv, float* rgb) {
  tri o = yuv_to_rgb(gamut, (tri){yuv[0], yuv[1], yuv[2]});
  rgb[0] = o.a; rgb[1] = o.b; rgb[2] = o.c;
}
void orc_rgbToYuv(int gamut, const float* rgb, float* yuv) {
  tri o = rgb_to_yuv(gamut, (tri){rgb[0], rgb[1], rgb[2]});
  yuv[0] = o.a; yuv[1] = o.b; yuv[2] = o.c;
}

/* ------------------------------------------------------------------ transfer functions */
/* gmm.cpp:149-155. pow is the double overload (exponent literal 2.4 is a double). */
float orc_srgbInvOetf(float x) {
  if (x <= 0.04045f) return x / 12.92f;
  return (float)pow((double)((x + 0.055f) / 1.055f), 2.4);
}

static const float HLG_A = 0.17883277f, HLG_B = 0.28466892f, HLG_C = (float)0.55991073; /* gmm.cpp:257 */

/* gmm.cpp:259-265. sqrt(float) narrows exactly to sqrtf; the log branch is evaluated in double:
 * float * double + float. */
float orc_hlgOetf(float e) {
  if (e <= 1.0f / 12.0f) return sqrtf(3.0f * e);
  return (float)((double)HLG_A * log((double)(12.0f * e - HLG_B)) + (double)HLG_C);
}
/* gmm.cpp:280-286. double: pow(x,2)/3 and (exp(.)+b)/12 */
float orc_hlgInvOetf(float x) {
  if (x <= 0.5f) return (float)(pow((double)x, 2.0) / (double)3.0f);
  return (float)((exp((double)((x - HLG_C) / HLG_A)) + (double)HLG_B) / (double)12.0f);
}

/* gmm.cpp:305-307 */
static const float PQ_M1 = 2610.0f / 16384.0f, PQ_M2 = 2523.0f / 4096.0f * 128.0f;
static const float PQ_C1 = 3424.0f / 4096.0f, PQ_C2 = 2413.0f / 4096.0f * 32.0f,
                   PQ_C3 = 2392.0f / 4096.0f * 32.0f;
/* gmm.cpp:309-312, all double after the first pow */
float orc_pqOetf(float e) {
  if (e <= 0.0f) return 0.0f;
  double p = pow((double)e, (double)PQ_M1);
  return (float)pow(((double)PQ_C1 + (double)PQ_C2 * p) / (1 + (double)PQ_C3 * p), (double)PQ_M2);
}
/* gmm.cpp:327-338 */
static const float PQI_A = 128.0f, PQI_B = 107.0f, PQI_C = 2413.0f, PQI_D = 2392.0f,
                   PQI_E = 6.2773946361f, PQI_F = 0.0126833f;
float orc_pqInvOetf(float x) {
  if (x <= 0.0001f) return 0.0f;
  double p = pow((double)x, (double)PQI_F);
  return (float)pow(((double)PQI_A * p - (double)PQI_B) / ((double)PQI_C - (double)PQI_D * p),
                    (double)PQI_E);
}

/* ---- nearest-entry LUT variants (the -DUSE_*_LUT=1 build), gmm.cpp:21-64,162-167,269-275,
 * 292-298,316-322,344-350; sizes gmm.h:268-269,329-330,342-343,355-356,368-369 ---- */
#define N_SRGB_INV 1024
#define N_HLG_OETF 65536
#define N_HLG_INV 4096
#define N_PQ_OETF 65536
#define N_PQ_INV 4096
static float *T_SRGB_INV, *T_HLG_OETF, *T_HLG_INV, *T_PQ_OETF, *T_PQ_INV;
static float* build_table(int n, float (*fn)(float)) {
  float* t = (float*)malloc(sizeof(float) * (size_t)n);
  for (int i = 0; i < n; i++) t[i] = fn((float)i / (float)(n - 1));
  return t;
}
static float lut_lookup(float** tab, int n, float (*fn)(float), float x) {
  if (!*tab) *tab = build_table(n, fn); /* tests are single-threaded at first touch */
  /* float * (N-1 as float) + 0.5 (double), truncating conversion; then CLIP3 on the unsigned value */
  double pos = (double)(x * (float)(n - 1)) + 0.5;
  uint32_t idx = (uint32_t)(int64_t)pos;
  if (idx > (uint32_t)(n - 1)) idx = (uint32_t)(n - 1);
  return (*tab)[idx];
}
float orc_srgbInvOetfLUT(float x) { return lut_lookup(&T_SRGB_INV, N_SRGB_INV, orc_srgbInvOetf, x); }
float orc_hlgOetfLUT(float x) { return lut_lookup(&T_HLG_OETF, N_HLG_OETF, orc_hlgOetf, x); }
float orc_hlgInvOetfLUT(float x) { return lut_lookup(&T_HLG_INV, N_HLG_INV, orc_hlgInvOetf, x); }
float orc_pqOetfLUT(float x) { return lut_lookup(&T_PQ_OETF, N_PQ_OETF, orc_pqOetf, x); }
float orc_pqInvOetfLUT(float x) { return lut_lookup(&T_PQ_INV, N_PQ_INV, orc_pqInvOetf, x); }

/* ------------------------------------------------------------------ gamut 3x3s, gmm.cpp:359-393
 * index [to][from]; HDR is converted INTO the SDR gamut (selector gmm.cpp:397-440). */
static const float GAMUT[3][3][9] = {
    /* to 709 */
    {{1, 0, 0, 0, 1, 0, 0, 0, 1},
     {1.22482f, -0.22490f, -0.00007f, -0.04196f, 1.04199f, 0.00001f, -0.01961f, -0.07865f, 1.09831f},
     {1.66045f, -0.58764f, -0.07286f, -0.12445f, 1.13282f, -0.00837f, -0.01811f, -0.10057f, 1.11878f}},
    /* to P3 */
    {{0.82254f, 0.17755f, 0.00006f, 0.03312f, 0.96684f, -0.00001f, 0.01706f, 0.07240f, 0.91049f},
     {1, 0, 0, 0, 1, 0, 0, 0, 1},
     {1.34369f, -0.28223f, -0.06135f, -0.06533f, 1.07580f, -0.01051f, 0.00283f, -0.01957f, 1.01679f}},
    /* to 2100 */
    {{0.62740f, 0.32930f, 0.04332f, 0.06904f, 0.91958f, 0.01138f, 0.01636f, 0.08799f, 0.89555f},
     {0.75378f, 0.19862f, 0.04754f, 0.04576f, 0.94177f, 0.01250f, -0.00121f, 0.01757f, 0.98359f},
     {1, 0, 0, 0, 1, 0, 0, 0, 1}}};
static tri mat3(const float* m, tri e) {
  tri o = {m[0] * e.a + m[1] * e.b + m[2] * e.c, m[3] * e.a + m[4] * e.b + m[5] * e.c,
           m[6] * e.a + m[7] * e.b + m[8] * e.c};
  return o;
}
int orc_hdrConversion(int sdr_gamut, int hdr_gamut, const float* in, float* out) {
  if (sdr_gamut < 0 || sdr_gamut > 2 || hdr_gamut < 0 || hdr_gamut > 2) return -1;
  tri e = {in[0], in[1], in[2]};
  tri o = (sdr_gamut == hdr_gamut) ? e : mat3(GAMUT[sdr_gamut][hdr_gamut], e); /* identity returns e itself, gmm.h:391 */
  out[0] = o.a; out[1] = o.b; out[2] = o.c;
  return 0;
}

/* ------------------------------------------------------------------ YUV->YUV 3x3s, gmm.cpp:447-481
 * 0: 709->601, 1: 709->2100, 2: 601->709, 3: 601->2100, 4: 2100->709, 5: 2100->601 */
static const float YUV2YUV[6][9] = {
    {1.0f, 0.101579f, 0.196076f, 0.0f, 0.989854f, -0.110653f, 0.0f, -0.072453f, 0.983398f},
    {1.0f, -0.016969f, 0.096312f, 0.0f, 0.995306f, -0.051192f, 0.0f, 0.011507f, 1.002637f},
    {1.0f, -0.118188f, -0.212685f, 0.0f, 1.018640f, 0.114618f, 0.0f, 0.075049f, 1.025327f},
    {1.0f, -0.128245f, -0.115879f, 0.0f, 1.010016f, 0.061592f, 0.0f, 0.086969f, 1.029350f},
    {1.0f, 0.018149f, -0.095132f, 0.0f, 1.004123f, 0.051267f, 0.0f, -0.011524f, 0.996782f},
    {1.0f, 0.117887f, 0.105521f, 0.0f, 0.995211f, -0.059549f, 0.0f, -0.084085f, 0.976518f}};
void orc_yuvToYuv(int idx, const float* in, float* out) {
  tri o = mat3(YUV2YUV[idx], (tri){in[0], in[1], in[2]});
  out[0] = o.a; out[1] = o.b; out[2] = o.c;
}

/* ------------------------------------------------------------------ pixel accessors */
static tri yuv420_px(const uhdrb200_image* im, size_t x, size_t y) { /* gmm.cpp:562-581 */
  const uint8_t* luma = (const uint8_t*)im->data;
  const uint8_t* chroma = (const uint8_t*)im->chroma_data;
  size_t cs = im->chroma_stride;
  size_t v_off = cs * (im->height / 2);
  size_t ci = x / 2 + (y / 2) * cs;
  int yy = luma[x + y * im->luma_stride], uu = chroma[ci], vv = chroma[v_off + ci];
  tri o = {(float)yy * (1 / 255.0f), (float)(uu - 128) * (1 / 255.0f), (float)(vv - 128) * (1 / 255.0f)};
  return o;
}
static tri p010_px(const uhdrb200_image* im, size_t x, size_t y) { /* gmm.cpp:583-601, no range clamp */
  const uint16_t* luma = (const uint16_t*)im->data;
  const uint16_t* chroma = (const uint16_t*)im->chroma_data;
  size_t ls = im->luma_stride == 0 ? im->width : im->luma_stride;
  size_t ui = (y >> 1) * im->chroma_stride + (x & ~(size_t)1);
  int yy = luma[y * ls + x] >> 6, uu = chroma[ui] >> 6, vv = chroma[ui + 1] >> 6;
  tri o = {(float)(yy - 64) * (1 / 876.0f), (float)(uu - 64) * (1 / 896.0f) - 0.5f,
           (float)(vv - 64) * (1 / 896.0f) - 0.5f};
  return o;
}
/* gmm.cpp:605-615: sequential float accumulation, dy-major, then one division */
static tri box_sample(const uhdrb200_image* im, size_t s, size_t x, size_t y, int p010) {
  tri acc = {0.0f, 0.0f, 0.0f};
  for (size_t dy = 0; dy < s; ++dy)
    for (size_t dx = 0; dx < s; ++dx) {
      tri p = p010 ? p010_px(im, x * s + dx, y * s + dy) : yuv420_px(im, x * s + dx, y * s + dy);
      acc.a += p.a; acc.b += p.b; acc.c += p.c;
    }
  float n = (float)(s * s);
  acc.a /= n; acc.b /= n; acc.c /= n;
  return acc;
}
void orc_getYuv420Pixel(const uhdrb200_image* im, size_t x, size_t y, float* o) {
  tri t = yuv420_px(im, x, y); o[0] = t.a; o[1] = t.b; o[2] = t.c;
}
void orc_getP010Pixel(const uhdrb200_image* im, size_t x, size_t y, float* o) {
  tri t = p010_px(im, x, y); o[0] = t.a; o[1] = t.b; o[2] = t.c;
}
void orc_sampleYuv420(const uhdrb200_image* im, size_t s, size_t x, size_t y, float* o) {
  tri t = box_sample(im, s, x, y, 0); o[0] = t.a; o[1] = t.b; o[2] = t.c;
}
void orc_sampleP010(const uhdrb200_image* im, size_t s, size_t x, size_t y, float* o) {
  tri t = box_sample(im, s, x, y, 1); o[0] = t.a; o[1] = t.b; o[2] = t.c;
}

/* ------------------------------------------------------------------ gain encode / apply */
/* gmm.cpp:529-541; the quotient chain is double, the final cast truncates */
static uint8_t encode_gain(float y_sdr, float y_hdr, float mn, float mx, float l2mn, float l2mx) {
  float gain = 1.0f;
  if (y_sdr > 0.0f) gain = y_hdr / y_sdr;
  if (gain < mn) gain = mn;
  if (gain > mx) gain = mx;
  double v = (log2((double)gain) - (double)l2mn) / (double)(l2mx - l2mn) * (double)255.0f;
  return (uint8_t)v;
}
int orc_encodeGain(float y_sdr, float y_hdr, float mn, float mx) { /* gmm.cpp:524-527 */
  return encode_gain(y_sdr, y_hdr, mn, mx, (float)log2((double)mn), (float)log2((double)mx));
}
/* gmm.cpp:543-555 */
void orc_applyGain(const float* rgb, float gain, float mn, float mx, float display_boost,
                   int use_display_boost, float* out) {
  float log_boost = (float)(log2((double)mn) * (double)(1.0f - gain) + log2((double)mx) * (double)gain);
  float f = use_display_boost ? (float)exp2((double)(log_boost * display_boost / mx))
                              : (float)exp2((double)log_boost);
  out[0] = rgb[0] * f; out[1] = rgb[1] * f; out[2] = rgb[2] * f;
}
/* GainLUT gmm.h:151-182 + applyGainLUT gmm.cpp:557-560 */
static void gain_lut_build(float* tab, float mn, float mx, float display_boost) {
  float boost_factor = display_boost > 0 ? display_boost / mx : 1.0f;
  for (int i = 0; i < 1024; i++) {
    float v = (float)i / 1023.0f;
    float log_boost = (float)(log2((double)mn) * (double)(1.0f - v) + log2((double)mx) * (double)v);
    tab[i] = (float)exp2((double)(log_boost * boost_factor));
  }
}
static float gain_lut_get(const float* tab, float gain) {
  uint32_t idx = (uint32_t)(int64_t)((double)(gain * 1023.0f) + 0.5);
  if (idx > 1023u) idx = 1023u;
  return tab[idx];
}
void orc_applyGainLUT(const float* rgb, float gain, float mn, float mx, float display_boost, float* out) {
  float tab[1024];
  gain_lut_build(tab, mn, mx, display_boost);
  float f = gain_lut_get(tab, gain);
  out[0] = rgb[0] * f; out[1] = rgb[1] * f; out[2] = rgb[2] * f;
}

/* ------------------------------------------------------------------ Shepard IDW, gmm.h:184-228, gmm.cpp:69-110 */
static float dist(float x1, float x2, float y1, float y2) {
  return sqrtf(((y2 - y1) * (y2 - y1)) + (x2 - x1) * (x2 - x1));
}
static void idw_fill(float* w, int s, int inc_r, int inc_b) {
  for (int y = 0; y < s; y++)
    for (int x = 0; x < s; x++) {
      float px = ((float)x) / s, py = ((float)y) / s;
      int cx = (int)floorf(px), cy = (int)floorf(py);
      int nx = cx + inc_r, ny = cy + inc_b;
      float* o = w + (y * s + x) * 4;
      float d1 = dist(px, (float)cx, py, (float)cy);
      if (d1 == 0) { o[0] = 1.f; o[1] = 0.f; o[2] = 0.f; o[3] = 0.f; continue; }
      float w1 = 1.f / d1;
      float w2 = 1.f / dist(px, (float)cx, py, (float)ny);
      float w3 = 1.f / dist(px, (float)nx, py, (float)cy);
      float w4 = 1.f / dist(px, (float)nx, py, (float)ny);
      float tot = w1 + w2 + w3 + w4;
      o[0] = w1 / tot; o[1] = w2 / tot; o[2] = w3 / tot; o[3] = w4 / tot;
    }
}
void orc_idwTable(int scale, int table, float* out) {
  static const int inc[4][2] = {{1, 1}, {0, 1}, {1, 0}, {0, 0}}; /* mWeights, NR, NB, C: gmm.h:191-194 */
  idw_fill(out, scale, inc[table][0], inc[table][1]);
}
/* gmm.cpp:686-720; the map is addressed with map->width as its stride */
static float sample_map_idw(const uhdrb200_image* map, size_t s, size_t x, size_t y, float* const tabs[4]) {
  const uint8_t* m = (const uint8_t*)map->data;
  size_t xl = x / s, xu = xl + 1, yl = y / s, yu = yl + 1;
  if (xl > map->width - 1) xl = map->width - 1;
  if (xu > map->width - 1) xu = map->width - 1;
  if (yl > map->height - 1) yl = map->height - 1;
  if (yu > map->height - 1) yu = map->height - 1;
  float e1 = (float)m[xl + yl * map->width] / 255.0f, e2 = (float)m[xl + yu * map->width] / 255.0f;
  float e3 = (float)m[xu + yl * map->width] / 255.0f, e4 = (float)m[xu + yu * map->width] / 255.0f;
  int ox = (int)(x % s), oy = (int)(y % s);
  const float* w = tabs[0];
  if (xl == xu && yl == yu) w = tabs[3];
  else if (xl == xu) w = tabs[1];
  else if (yl == yu) w = tabs[2];
  w += oy * s * 4 + ox * 4;
  return e1 * w[0] + e2 * w[1] + e3 * w[2] + e4 * w[3];
}
float orc_sampleMapIdw(const uhdrb200_image* map, size_t s, size_t x, size_t y) {
  float* tabs[4];
  for (int t = 0; t < 4; t++) { tabs[t] = (float*)malloc(sizeof(float) * s * s * 4); orc_idwTable((int)s, t, tabs[t]); }
  float r = sample_map_idw(map, s, x, y, tabs);
  for (int t = 0; t < 4; t++) free(tabs[t]);
  return r;
}
/* gmm.cpp:639-684 — float-scale variant; dead code on the path (uhdr.cpp:440 is always false for a
 * size_t scale) but kept so the reference's SampleMap KAT (IDW == float Shepard) can be restated. */
static size_t clampz(size_t v, size_t lo, size_t hi) { return v < lo ? lo : (hi < v ? hi : v); }
static float pyth(float dx, float dy) { return (float)sqrt(pow((double)dx, 2.0) + pow((double)dy, 2.0)); }
float orc_sampleMapFloat(const uhdrb200_image* map, float s, size_t x, size_t y) {
  const uint8_t* m = (const uint8_t*)map->data;
  float xm = (float)x / s, ym = (float)y / s;
  size_t xl = (size_t)floorf(xm), xu = xl + 1, yl = (size_t)floorf(ym), yu = yl + 1;
  xl = clampz(xl, 0, map->width - 1); xu = clampz(xu, 0, map->width - 1);
  yl = clampz(yl, 0, map->height - 1); yu = clampz(yu, 0, map->height - 1);
  float e1 = (float)m[xl + yl * map->width] / 255.0f, d1 = pyth(xm - (float)xl, ym - (float)yl);
  if (d1 == 0.0f) return e1;
  float e2 = (float)m[xl + yu * map->width] / 255.0f, d2 = pyth(xm - (float)xl, ym - (float)yu);
  if (d2 == 0.0f) return e2;
  float e3 = (float)m[xu + yl * map->width] / 255.0f, d3 = pyth(xm - (float)xu, ym - (float)yl);
  if (d3 == 0.0f) return e3;
  float e4 = (float)m[xu + yu * map->width] / 255.0f, d4 = pyth(xm - (float)xu, ym - (float)yu);
  if (d4 == 0.0f) return e2; /* sic: the reference returns e2 here, gmm.cpp:674 */
  float w1 = 1.0f / d1, w2 = 1.0f / d2, w3 = 1.0f / d3, w4 = 1.0f / d4, tot = w1 + w2 + w3 + w4;
  return e1 * (w1 / tot) + e2 * (w2 / tot) + e3 * (w3 / tot) + e4 * (w4 / tot);
}

/* ------------------------------------------------------------------ packers */
uint16_t orc_floatToHalf(float f) { /* gmm.h:136-147: +0x1000 rounding, saturate 0x7FFF, flush e<=101 */
  uint32_t bits; memcpy(&bits, &f, 4);
  const uint32_t b = bits + 0x00001000u;
  const int32_t e = (int32_t)((b & 0x7F800000u) >> 23);
  const uint32_t m = b & 0x007FFFFFu;
  uint32_t r = (b & 0x80000000u) >> 16;
  if (e > 112) r |= (((uint32_t)(e - 112) << 10) & 0x7C00u) | (m >> 13);
  if (e < 113 && e > 101) r |= (((0x007FF000u + m) >> (125 - e)) + 1) >> 1;
  if (e > 143) r |= 0x7FFFu;
  return (uint16_t)r;
}
uint32_t orc_colorToRgba1010102(float r, float g, float b) { /* gmm.cpp:722-727, truncating */
  return (0x3ffu & (uint32_t)(r * 1023.0f)) | ((0x3ffu & (uint32_t)(g * 1023.0f)) << 10) |
         ((0x3ffu & (uint32_t)(b * 1023.0f)) << 20) | (0x3u << 30);
}
uint64_t orc_colorToRgbaF16(float r, float g, float b) { /* gmm.cpp:729-732 */
  return (uint64_t)orc_floatToHalf(r) | ((uint64_t)orc_floatToHalf(g) << 16) |
         ((uint64_t)orc_floatToHalf(b) << 32) | ((uint64_t)orc_floatToHalf(1.0f) << 48);
}

/* ================================================================== the three path entry points */
static int use_lut_mode = 0; /* 0 = as built (analytic); 1 = the -DUSE_*_LUT=1 variant */
void orc_set_lut_mode(int on) { use_lut_mode = on != 0; }
int orc_is_lut_variant(void) { return use_lut_mode; }

static tri map3(float (*fn)(float), tri e) { tri o = {fn(e.a), fn(e.b), fn(e.c)}; return o; }

/* uhdr.cpp:185-358. The map is written into map_buf (the reference new[]s it, uhdr.cpp:209). */
int orc_generate_gainmap(const uhdrb200_image* sdr, const uhdrb200_image* hdr, int hdr_tf,
                         uhdrb200_metadata* md, uhdrb200_image* dest, int sdr_is_601,
                         void* map_buf, size_t map_buf_bytes) {
  if (!sdr || !hdr || !md || !dest || !sdr->data || !sdr->chroma_data || !hdr->data || !hdr->chroma_data)
    return UHDRB200_ERROR_BAD_PTR; /* :189-194 */
  if (sdr->width != hdr->width || sdr->height != hdr->height) return UHDRB200_ERROR_RESOLUTION_MISMATCH; /* :195-198 */
  if (sdr->colorGamut == UHDRB200_COLORGAMUT_UNSPECIFIED || hdr->colorGamut == UHDRB200_COLORGAMUT_UNSPECIFIED)
    return UHDRB200_ERROR_INVALID_COLORGAMUT; /* :199-202 */
  size_t mw = sdr->width / 4, mh = sdr->height / 4; /* kMapDimensionScaleFactor, ultrahdr.h:213 */
  float (*inv)(float) = NULL;
  float white;
  switch (hdr_tf) { /* :222-248 */
    case UHDRB200_TF_LINEAR: white = HLG_MAX_NITS; break;
    case UHDRB200_TF_HLG: inv = use_lut_mode ? orc_hlgInvOetfLUT : orc_hlgInvOetf; white = HLG_MAX_NITS; break;
    case UHDRB200_TF_PQ: inv = use_lut_mode ? orc_pqInvOetfLUT : orc_pqInvOetf; white = PQ_MAX_NITS; break;
    default: return UHDRB200_ERROR_INVALID_TRANS_FUNC;
  }
  if (!map_buf || map_buf_bytes < mw * mh) return -20004; /* ERROR_ULTRAHDR_BUFFER_TOO_SMALL (harness only) */
  dest->data = map_buf; dest->width = mw; dest->height = mh; /* :209-216 */
  dest->colorGamut = UHDRB200_COLORGAMUT_UNSPECIFIED; dest->luma_stride = mw;
  dest->chroma_data = NULL; dest->chroma_stride = 0; dest->pixelFormat = UHDRB200_PIX_FMT_MONOCHROME;
  memset(md->version, 0, sizeof md->version); strcpy(md->version, "1.0"); /* :250-257 */
  md->maxContentBoost = white / SDR_WHITE_NITS; md->minContentBoost = 1.0f; md->gamma = 1.0f;
  md->offsetSdr = 0.0f; md->offsetHdr = 0.0f; md->hdrCapacityMin = 1.0f; md->hdrCapacityMax = md->maxContentBoost;
  float l2mn = (float)log2((double)md->minContentBoost), l2mx = (float)log2((double)md->maxContentBoost); /* :259-260 */
  int sg = sdr->colorGamut, hg = hdr->colorGamut;
  if (sg < 0 || sg > 2 || hg < 0 || hg > 2) return UHDRB200_ERROR_INVALID_COLORGAMUT; /* :280-282,299-301 */
  int sdr_matrix = sdr_is_601 ? 1 : sg; /* :284-286 */
  float (*sinv)(float) = use_lut_mode ? orc_srgbInvOetfLUT : orc_srgbInvOetf;
  uint8_t* out = (uint8_t*)map_buf;
  for (size_t y = 0; y < mh; ++y)
    for (size_t x = 0; x < mw; ++x) { /* :314-336 */
      tri s_rgb = map3(sinv, yuv_to_rgb(sdr_matrix, box_sample(sdr, 4, x, y, 0)));
      float s_nits = orc_luminance(sg, s_rgb.a, s_rgb.b, s_rgb.c) * SDR_WHITE_NITS;
      tri h_rgb = yuv_to_rgb(hg, box_sample(hdr, 4, x, y, 1));
      if (inv) h_rgb = map3(inv, h_rgb);
      if (sg != hg) h_rgb = mat3(GAMUT[sg][hg], h_rgb);
      float h_nits = orc_luminance(sg, h_rgb.a, h_rgb.b, h_rgb.c) * white; /* luminance by SDR gamut, :265-283 */
      out[x + y * mw] = encode_gain(s_nits, h_nits, md->minContentBoost, md->maxContentBoost, l2mn, l2mx);
    }
  return UHDRB200_NO_ERROR;
}

/* uhdr.cpp:360-515 */
int orc_apply_gainmap(const uhdrb200_image* sdr, const uhdrb200_image* map, const uhdrb200_metadata* md,
                      int fmt, float max_display_boost, uhdrb200_image* dest) {
  if (!sdr || !map || !md || !dest || !sdr->data || !sdr->chroma_data || !map->data) return UHDRB200_ERROR_BAD_PTR;
  if (strncmp(md->version, "1.0", sizeof md->version) != 0) return UHDRB200_ERROR_BAD_METADATA; /* :369 */
  if (md->gamma != 1.0f) return UHDRB200_ERROR_BAD_METADATA;                                  /* :373 */
  if (md->offsetSdr != 0.0f || md->offsetHdr != 0.0f) return UHDRB200_ERROR_BAD_METADATA;     /* :377 */
  if (md->hdrCapacityMin != md->minContentBoost || md->hdrCapacityMax != md->maxContentBoost)
    return UHDRB200_ERROR_BAD_METADATA; /* :381 */
  if (sdr->width % map->width != 0 || sdr->height % map->height != 0) return UHDRB200_ERROR_UNSUPPORTED_MAP_SCALE_FACTOR;
  if (sdr->width * map->height != sdr->height * map->width) return UHDRB200_ERROR_UNSUPPORTED_MAP_SCALE_FACTOR;
  size_t s = sdr->width / map->width, w = sdr->width, h = sdr->height;
  dest->width = w; dest->height = h; dest->colorGamut = sdr->colorGamut; /* :411-413 */
  float* tabs[4];
  for (int t = 0; t < 4; t++) { tabs[t] = (float*)malloc(sizeof(float) * s * s * 4); orc_idwTable((int)s, t, tabs[t]); }
  float db = max_display_boost < md->maxContentBoost ? max_display_boost : md->maxContentBoost; /* std::min, :415 */
  float lut[1024];
  if (use_lut_mode) gain_lut_build(lut, md->minContentBoost, md->maxContentBoost, db);
  float (*sinv)(float) = use_lut_mode ? orc_srgbInvOetfLUT : orc_srgbInvOetf;
  float (*hlg)(float) = use_lut_mode ? orc_hlgOetfLUT : orc_hlgOetf;
  float (*pq)(float) = use_lut_mode ? orc_pqOetfLUT : orc_pqOetf;
  for (size_t y = 0; y < h; ++y)
    for (size_t x = 0; x < w; ++x) { /* :427-496 */
      tri lin = map3(sinv, yuv_to_rgb(1, yuv420_px(sdr, x, y))); /* always the 601 matrix, :431 */
      float g = sample_map_idw(map, s, x, y, tabs);
      float o[3], in[3] = {lin.a, lin.b, lin.c};
      if (use_lut_mode) { float f = gain_lut_get(lut, g); o[0] = in[0] * f; o[1] = in[1] * f; o[2] = in[2] * f; }
      else orc_applyGain(in, g, md->minContentBoost, md->maxContentBoost, db, 1, o);
      tri e = {o[0] / db, o[1] / db, o[2] / db}; /* :451 */
      size_t i = x + y * w;
      switch (fmt) {
        case UHDRB200_OUTPUT_HDR_LINEAR: ((uint64_t*)dest->data)[i] = orc_colorToRgbaF16(e.a, e.b, e.c); break;
        case UHDRB200_OUTPUT_HDR_LINEAR_RGB_10BIT: { /* :460-468 */
          uint16_t* p = (uint16_t*)dest->data;
          p[i] = (uint16_t)(0x3ff & (uint32_t)(e.a * 1023.0f));
          p[w * h + i] = (uint16_t)(0x3ff & (uint32_t)(e.b * 1023.0f));
          p[w * h * 2 + i] = (uint16_t)(0x3ff & (uint32_t)(e.c * 1023.0f));
          break;
        }
        case UHDRB200_OUTPUT_HDR_HLG: { tri q = map3(hlg, e); ((uint32_t*)dest->data)[i] = orc_colorToRgba1010102(q.a, q.b, q.c); break; }
        case UHDRB200_OUTPUT_HDR_PQ: { tri q = map3(pq, e); ((uint32_t*)dest->data)[i] = orc_colorToRgba1010102(q.a, q.b, q.c); break; }
        default: break; /* :491-493: silently writes nothing */
      }
    }
  for (int t = 0; t < 4; t++) free(tabs[t]);
  return UHDRB200_NO_ERROR;
}

/* uhdr.cpp:517-558 */
int orc_tonemap(const uhdrb200_image* src, uhdrb200_image* dest) {
  if (!src || !dest) return UHDRB200_ERROR_BAD_PTR;
  if (src->width != dest->width || src->height != dest->height) return UHDRB200_ERROR_RESOLUTION_MISMATCH;
  const uint16_t* sy = (const uint16_t*)src->data;
  uint8_t* dy = (uint8_t*)dest->data;
  for (size_t y = 0; y < src->height; ++y) {
    const uint16_t* s = sy + y * src->luma_stride;
    uint8_t* d = dy + y * dest->luma_stride;
    for (size_t x = 0; x < src->width; ++x) d[x] = (uint8_t)(((s[x] >> 6) >> 2) & 0xff);
    if (dest->width != dest->luma_stride) memset(d + dest->width, 0, dest->luma_stride - dest->width);
  }
  const uint16_t* suv = (const uint16_t*)src->chroma_data;
  uint8_t* du = (uint8_t*)dest->chroma_data;
  uint8_t* dv = du + dest->chroma_stride * dest->height / 2;
  for (size_t y = 0; y < src->height / 2; ++y) {
    const uint16_t* s = suv + y * src->chroma_stride;
    uint8_t* u = du + y * dest->chroma_stride;
    uint8_t* v = dv + y * dest->chroma_stride;
    for (size_t x = 0; x < src->width / 2; ++x) {
      u[x] = (uint8_t)(((s[x << 1] >> 6) >> 2) & 0xff);
      v[x] = (uint8_t)(((s[(x << 1) + 1] >> 6) >> 2) & 0xff);
    }
    if (dest->width / 2 != dest->chroma_stride) {
      memset(u + dest->width / 2, 0, dest->chroma_stride - dest->width / 2);
      memset(v + dest->width / 2, 0, dest->chroma_stride - dest->width / 2);
    }
  }
  dest->colorGamut = src->colorGamut;
  return UHDRB200_NO_ERROR;
}

/* N1: JpegR::convertYuv jpegr.cpp:1132-1206 + transformYuv420 gmm.cpp:483-520 */
static float clip255(float v) { return v < 0 ? 0 : (v > 255 ? 255 : v); } /* CLIP3, gmm.h:25 */
int orc_convert_yuv(uhdrb200_image* im, int src_enc, int dst_enc) {
  if (!im) return UHDRB200_ERROR_BAD_PTR;
  if (src_enc == UHDRB200_COLORGAMUT_UNSPECIFIED || dst_enc == UHDRB200_COLORGAMUT_UNSPECIFIED)
    return UHDRB200_ERROR_INVALID_COLORGAMUT;
  if (src_enc < 0 || src_enc > 2 || dst_enc < 0 || dst_enc > 2) return UHDRB200_ERROR_INVALID_COLORGAMUT;
  if (src_enc == dst_enc) return UHDRB200_NO_ERROR;
  static const int sel[3][3] = {{-1, 0, 1}, {2, -1, 3}, {4, 5, -1}};
  const float* m = YUV2YUV[sel[src_enc][dst_enc]];
  uint8_t* luma = (uint8_t*)im->data;
  uint8_t* chroma = (uint8_t*)im->chroma_data;
  size_t ls = im->luma_stride, cs = im->chroma_stride;
  size_t v_plane = cs * im->height / 2; /* gmm.cpp:507 */
  for (size_t cy = 0; cy < im->height / 2; ++cy)
    for (size_t cx = 0; cx < im->width / 2; ++cx) {
      tri p[4];
      for (int k = 0; k < 4; k++) p[k] = mat3(m, yuv420_px(im, cx * 2 + (k & 1), cy * 2 + (k >> 1)));
      float nu = (((p[0].b + p[1].b) + p[2].b) + p[3].b) / 4.0f;
      float nv = (((p[0].c + p[1].c) + p[2].c) + p[3].c) / 4.0f;
      for (int k = 0; k < 4; k++)
        luma[(cx * 2 + (k & 1)) + (cy * 2 + (k >> 1)) * ls] = (uint8_t)clip255(p[k].a * 255.0f + 0.5f);
      chroma[cx + cy * cs] = (uint8_t)clip255(nu * 255.0f + 128.0f + 0.5f);
      chroma[v_plane + cx + cy * cs] = (uint8_t)clip255(nv * 255.0f + 128.0f + 0.5f);
    }
  return UHDRB200_NO_ERROR;
}

/* N3: SDR output pack, examples/ultrahdr_app.cpp:682-734 (convertYuv420ToRGBImage): full-range
 * BT.601 YUV420 -> RGBA8888, nearest chroma sample (j/2, i/2), true divisions by 255.0f, the
 * 3-term dot products summed left to right in float, *255 + 0.5, CLIP3 to [0,255], truncation,
 * alpha 255. Matrix literals are the reference's double expressions narrowed to float (:34-35). */
int orc_yuv420_to_rgba8888(const uhdrb200_image* im, uhdrb200_image* dest) {
  if (!im || !dest || !im->data || !im->chroma_data || !dest->data) return UHDRB200_ERROR_BAD_PTR;
  static const float c[9] = {1, 0, 1.402, 1, (-0.202008 / 0.587), (-0.419198 / 0.587), 1.0, 1.772, 0.0};
  const uint8_t* y = (const uint8_t*)im->data;
  const uint8_t* u = (const uint8_t*)im->chroma_data;
  const uint8_t* v = u + im->chroma_stride * (im->height / 2);
  uint32_t* out = (uint32_t*)dest->data;
  const size_t os = dest->luma_stride >= im->width ? dest->luma_stride : im->width;
  for (size_t i = 0; i < im->height; i++)
    for (size_t j = 0; j < im->width; j++) {
      float y0 = (float)y[im->luma_stride * i + j];
      float u0 = (float)(u[im->chroma_stride * (i / 2) + (j / 2)] - 128);
      float v0 = (float)(v[im->chroma_stride * (i / 2) + (j / 2)] - 128);
      y0 /= 255.0f;
      u0 /= 255.0f;
      v0 /= 255.0f;
      float r = c[0] * y0 + c[1] * u0 + c[2] * v0;
      float g = c[3] * y0 + c[4] * u0 + c[5] * v0;
      float b = c[6] * y0 + c[7] * u0 + c[8] * v0;
      r = clip255(r * 255.0f + 0.5f);
      g = clip255(g * 255.0f + 0.5f);
      b = clip255(b * 255.0f + 0.5f);
      out[i * os + j] = (uint32_t)(int32_t)r | ((uint32_t)(int32_t)g << 8) | ((uint32_t)(int32_t)b << 16) | (255u << 24);
    }
  dest->width = im->width;
  dest->height = im->height;
  dest->colorGamut = im->colorGamut;
  return UHDRB200_NO_ERROR;
}

/* N3: copy-with-restride of a YUV420 image, jpegr.cpp:315-359 / ultrahdr_api.cpp:437-479: rows are
 * copied `width` (`width/2`) bytes at a time and the destination stride tails are zero-filled. */
int orc_copy_yuv420(const uhdrb200_image* src, uhdrb200_image* dst) {
  if (!src || !dst || !src->data || !src->chroma_data || !dst->data || !dst->chroma_data) return UHDRB200_ERROR_BAD_PTR;
  if (dst->luma_stride < src->width || dst->chroma_stride < src->width / 2) return UHDRB200_ERROR_INVALID_STRIDE;
  const size_t w = src->width, h = src->height;
  for (size_t i = 0; i < h; i++) {
    uint8_t* d = (uint8_t*)dst->data + i * dst->luma_stride;
    memcpy(d, (const uint8_t*)src->data + i * src->luma_stride, w);
    memset(d + w, 0, dst->luma_stride - w);
  }
  for (int pl = 0; pl < 2; pl++) {
    const uint8_t* s0 = (const uint8_t*)src->chroma_data + (size_t)pl * src->chroma_stride * (h / 2);
    uint8_t* d0 = (uint8_t*)dst->chroma_data + (size_t)pl * dst->chroma_stride * (h / 2);
    for (size_t i = 0; i < h / 2; i++) {
      memcpy(d0 + i * dst->chroma_stride, s0 + i * src->chroma_stride, w / 2);
      memset(d0 + i * dst->chroma_stride + w / 2, 0, dst->chroma_stride - w / 2);
    }
  }
  dst->width = src->width;
  dst->height = src->height;
  dst->colorGamut = src->colorGamut;
  return UHDRB200_NO_ERROR;
}

/* ------------------------------------------------------------------------------------------------------------------
 * Row N4: dequantise + 8x8 inverse DCT + range-limit of JPEG coefficient planes, what libjpeg-turbo does inside
 * jpeg_read_raw_data for the reference's JpegDecoderHelper (lib/src/jpegdecoderhelper.cpp:291-307: raw_data_out,
 * dct_method = JDCT_ISLOW). libjpeg-turbo (tag 3.0.1, fetched at configure time, CMakeLists.txt:254-256) is NOT in
 * /root/reference: this restates the published IJG algorithm (jidctint.c, jpeg_idct_islow: CONST_BITS 13, PASS1_BITS 2;
 * jdmaster.c prepare_range_limit_table). Pinned empirically: tests/test_jpeg_idct.py compares it bit for bit with
 * libjpeg-turbo's own luma output (through Pillow) on the reference's JPEG fixtures. */
static void islow_1d(const int32_t* in, int stride_in, int32_t* out, int stride_out, int shift) {
  int32_t z1, z2, z3, z4, z5, tmp0, tmp1, tmp2, tmp3, tmp10, tmp11, tmp12, tmp13;
  z2 = in[2 * stride_in]; z3 = in[6 * stride_in];
  z1 = (z2 + z3) * 4433;            /* FIX_0_541196100 */
  tmp2 = z1 + z3 * (-15137);        /* FIX_1_847759065 */
  tmp3 = z1 + z2 * 6270;            /* FIX_0_765366865 */
  z2 = in[0]; z3 = in[4 * stride_in];
  tmp0 = (int32_t)((uint32_t)(z2 + z3) << 13);
  tmp1 = (int32_t)((uint32_t)(z2 - z3) << 13);
  tmp10 = tmp0 + tmp3; tmp13 = tmp0 - tmp3; tmp11 = tmp1 + tmp2; tmp12 = tmp1 - tmp2;
  tmp0 = in[7 * stride_in]; tmp1 = in[5 * stride_in]; tmp2 = in[3 * stride_in]; tmp3 = in[1 * stride_in];
  z1 = tmp0 + tmp3; z2 = tmp1 + tmp2; z3 = tmp0 + tmp2; z4 = tmp1 + tmp3;
  z5 = (z3 + z4) * 9633;            /* FIX_1_175875602 */
  tmp0 *= 2446; tmp1 *= 16819; tmp2 *= 25172; tmp3 *= 12299; /* 0.298631336 2.053119869 3.072711026 1.501321110 */
  z1 *= -7373; z2 *= -20995; z3 *= -16069; z4 *= -3196;      /* 0.899976223 2.562915447 1.961570560 0.390180644 */
  z3 += z5; z4 += z5;
  tmp0 += z1 + z3; tmp1 += z2 + z4; tmp2 += z2 + z3; tmp3 += z1 + z4;
  const int32_t rnd = (int32_t)1 << (shift - 1);
  out[0 * stride_out] = (tmp10 + tmp3 + rnd) >> shift; out[7 * stride_out] = (tmp10 - tmp3 + rnd) >> shift;
  out[1 * stride_out] = (tmp11 + tmp2 + rnd) >> shift; out[6 * stride_out] = (tmp11 - tmp2 + rnd) >> shift;
  out[2 * stride_out] = (tmp12 + tmp1 + rnd) >> shift; out[5 * stride_out] = (tmp12 - tmp1 + rnd) >> shift;
  out[3 * stride_out] = (tmp13 + tmp0 + rnd) >> shift; out[4 * stride_out] = (tmp13 - tmp0 + rnd) >> shift;
}
static uint8_t jpeg_range_limit(int32_t x) { /* table index x & RANGE_MASK (1023), centred on 128 */
  int i = (int)(x & 1023);
  return (uint8_t)(i < 128 ? i + 128 : (i < 512 ? 255 : (i < 896 ? 0 : i - 896)));
}
/* one block: 64 coefficients (natural order) x 64 quantisation steps -> 8x8 samples */
void orc_jpeg_idct_islow_block(const int16_t* coef, const uint16_t* quant, uint8_t* out8x8) {
  int32_t ws[64], in[64], row[8];
  for (int i = 0; i < 64; i++) in[i] = (int32_t)coef[i] * (int32_t)quant[i];
  for (int c = 0; c < 8; c++) islow_1d(in + c, 8, ws + c, 8, 13 - 2);
  for (int r = 0; r < 8; r++) {
    islow_1d(ws + r * 8, 1, row, 1, 13 + 2 + 3);
    for (int c = 0; c < 8; c++) out8x8[r * 8 + c] = jpeg_range_limit(row[c]);
  }
}
/* a whole component: blocks in raster order; only the visible width x height samples are written */
int orc_jpeg_idct_plane(const int16_t* coeffs, const uint16_t* quant, size_t blocks_w, size_t blocks_h, uint8_t* plane,
                        size_t width, size_t height, size_t stride) {
  if (!coeffs || !quant || !plane) return UHDRB200_ERROR_BAD_PTR;
  if (width > blocks_w * 8 || height > blocks_h * 8) return UHDRB200_ERROR_RESOLUTION_MISMATCH;
  uint8_t blk[64];
  for (size_t by = 0; by < blocks_h; by++)
    for (size_t bx = 0; bx < blocks_w; bx++) {
      orc_jpeg_idct_islow_block(coeffs + (by * blocks_w + bx) * 64, quant, blk);
      for (int r = 0; r < 8; r++)
        for (int c = 0; c < 8; c++)
          if (by * 8 + r < height && bx * 8 + c < width) plane[(by * 8 + r) * stride + bx * 8 + c] = blk[r * 8 + c];
    }
  return UHDRB200_NO_ERROR;
}
