// apply_q_common.cuh — device helpers shared by the applyGainMap kernels (apply_fast.cu, apply_pipe.cu, kat.cu):
// packed fp32x2 arithmetic, the table evaluators (p_eval / l_eval), the table staging (TMA bulk copy) and the q-domain
// evaluation of one quad-row of a 4x4 block (quad_row_eval). See apply_fast.cu for the formulation.
#pragma once
#include <type_traits>

#include "uhdr_io.cuh"
#include "uhdr_kernels.cuh"
#include "uhdr_math.cuh"

namespace uhdr {
namespace {

#ifndef UHDR_APPLY_MUFU_CHANNELS
#define UHDR_APPLY_MUFU_CHANNELS 2  // colour channels whose srgbInvOetf runs on the XU pipe (lg2/ex2) instead of the smem table
#endif
#ifndef UHDR_APPLY_MUFU_CHANNELS_LINEAR
#define UHDR_APPLY_MUFU_CHANNELS_LINEAR 2  // same, for the linear-output kernels (RGBA F16, planar 10-bit)
#endif
#ifndef UHDR_APPLY_MIN_BLOCKS
#define UHDR_APPLY_MIN_BLOCKS 3
#endif
constexpr int kThreads = 256;
#ifndef UHDR_APPLYQ_THREADS
#define UHDR_APPLYQ_THREADS 256
#endif
constexpr int kQThreads = UHDR_APPLYQ_THREADS;  // block size of apply_q_kernel (its tile is kQThreads / 32 warps x CELLS cell rows tall)
constexpr int kTileCellsX = 32, kTileCellsY = 8;
#ifndef UHDR_APPLYQ_UNROLL
#define UHDR_APPLYQ_UNROLL 1  // unroll factor of apply_q_kernel's cell walk
#endif
constexpr int kQUnroll = UHDR_APPLYQ_UNROLL;

enum { kPack1010102 = 0, kPackF16 = 1, kPackPlanar10 = 2 };

// log-spaced table: value = a + b * x on the segment selected by the float's bits; `tab` is pre-biased by -base
template <int SHIFT>
__device__ __forceinline__ float seg_eval(const float2* __restrict__ tab, float x) {
  const float2 ab = tab[__float_as_uint(x) >> SHIFT];
  return fmaf(ab.y, x, ab.x);
}

// uniform table over [0,1] with SEGS segments: floor(c * SEGS) lands in the low mantissa bits of
// c * SEGS + 2^23 rounded toward -inf (one FFMA.RM); `biased` = shared address of the table minus
// 8 * 0x4B000000 (mod 2^32), so the entry address is one shift-add (LEA) away.
template <int SEGS>
__device__ __forceinline__ float uni_eval(uint32_t biased, float c) {
  float m;
  asm("fma.rm.f32 %0, %1, %2, 0f4B000000;" : "=f"(m) : "f"(c), "f"((float)SEGS));
  float a, b;
  asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(a), "=f"(b) : "r"(biased + (__float_as_uint(m) << 3)));
  return fmaf(b, c, a);
}

// floor(v) for 0 <= v < 2^22 as an integer in the low mantissa bits: one FADD with round-to-minus-
// infinity on the FMA pipe instead of an F2I on the (quarter-rate) XU pipe.
__device__ __forceinline__ uint32_t floor_bits(float v) {
  float r;
  asm("add.rm.f32 %0, %1, 0f4B000000;" : "=f"(r) : "f"(v));  // + 2^23
  return __float_as_uint(r);
}

// srgbInvOetf (gainmapmath.cpp:149-155) on the XU pipe: 2 MUFU, ~4e-7 relative
__device__ __forceinline__ float srgb_inv_mufu(float c) {
  const float pw = mufu_ex2(2.4f * mufu_lg2(fmaf(c, (float)(1.0 / 1.055), (float)(0.055 / 1.055))));
  return c <= 0.04045f ? c * (float)(1.0 / 12.92) : pw;
}

// same, on a static shared array: the table's link-time offset and -BASE fold into the LDS immediate
template <int SHIFT, int BASE, int N>
__device__ __forceinline__ float seg_eval_static(const float2 (&tab)[N], float x) {
  const float2 ab = tab[(int)(__float_as_uint(x) >> SHIFT) - BASE];
  return fmaf(ab.y, x, ab.x);
}

// P table whose shared-memory ADDRESS is the float's bit-field. 16 octaves of segments form a ring of
// kPRingBytes (128 B << kPBits: 16 KB at 7 bits per octave); the table (13 octaves) sits at a shared address congruent to
// kPRing modulo the ring size, so entry address = ((bits >> (20 - kPBits)) & (kPRingBytes - 8)) | hi with
// `hi` = the ring-aligned part of the table address: SHF + LOP3, no base-register add.
constexpr uint32_t kPRingBytes = 128u << kPBits;
constexpr uint32_t kPRing = 8u * (kPBase & ((16 << kPBits) - 1));  // offset of the first entry inside its ring
static_assert(kPRing + kPEntries * 8 <= kPRingBytes, "P table must not wrap inside its ring");
__device__ __forceinline__ float p_eval(uint32_t hi, float x) {
  const uint32_t sh = __float_as_uint(x) >> (20 - kPBits);
  const uint32_t addr = (sh & (kPRingBytes - 8u)) | hi;
  float a, b;
  asm("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(a), "=f"(b) : "r"(addr));
  return fmaf(b, x, a);
}

// Dynamic shared memory the q kernel needs when its dynamic region starts at shared address dyn0
// (= reserved shared memory per block + the kernel's static shared memory): header, padding up to
// the ring position, table. The kernel recomputes the same layout and traps if it does not fit.
__host__ __device__ constexpr uint32_t q_table_addr(uint32_t dyn0) {
  return ((dyn0 + 16u - kPRing + kPRingBytes - 1u) & ~(kPRingBytes - 1u)) + kPRing;
}

__device__ __forceinline__ float add_rm(float a, float b) {
  float r;
  asm("add.rm.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
}

// ---- packed fp32x2 arithmetic (sm_100a FFMA2 / FADD2): one issue slot for two lanes of work ----
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float lo, float hi) {
  f32x2 r;
  asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpk2(f32x2 v, float& lo, float& hi) { asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
  f32x2 d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ f32x2 add_rm2(f32x2 a, f32x2 b) {  // both lanes rounded towards -inf
  f32x2 d;
  asm("add.rm.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}

__device__ __forceinline__ float ex2_fast(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}


// ---------------------------------------------------------------------------------------------
// q-domain kernel for the RGBA1010102 outputs (HLG / PQ).
//
// Above its linear toe srgbInvOetf(c) = Phi^2.4 with Phi = (c + 0.055) / 1.055 (gainmapmath.cpp:153),
// and Phi is LINEAR in the 8-bit luma: Phi = y * K + o, K = 1 / (255 * 1.055), o per channel and
// per 2x2 chroma quad. With f' = f^(1/2.4) (the exponent's 1/2.4 is folded into the constant-bank
// IDW weights) the reference chain  code = 1023 * OETF(f * srgbInvOetf(c))  is a function of
// q = Phi * f' alone: code = P(q), P(q) = 1023 * OETF(q^2.4), ONE shared-memory table look-up per
// channel instead of srgbInvOetf + OETF (2 look-ups or 1 look-up + 2 MUFU). Per channel:
// FFMA.SAT (Phi), FFMA (q + q0), SHF + LOP3 (segment), LDS.64, FFMA (value), FADD.RM (floor).
// The linear toe (c <= 0.04045, 8-bit code <= 10) has Phi = (c / 12.92)^(1/2.4): a quad-row of the
// cell that contains such a channel (tested conservatively on min(y) and min(o)) takes a branch
// that evaluates that power with MUFU lg2/ex2; everything downstream is shared.
// (Phi * f')^2.4 for the linear outputs: two MUFU; q == 0 gives lg2 = -inf and ex2(-inf) = +0
__device__ __forceinline__ float pow24(float q) { return mufu_ex2(2.4f * mufu_lg2(q)); }

// {lo, hi} floats -> packed half2 with truncation (F2FP.PACK_AB.RZ): on the normal-half range this is exactly the
// reference's floatToHalf (gainmapmath.h:136-147) once 0x1000 has been added to the float's bits
__device__ __forceinline__ uint32_t pack_half2_rz(float lo, float hi) {
  uint32_t d;
  asm("cvt.rz.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}

// L table (q^2.4, kLBits per octave): tL = shared address of the table - 8 * kLBase; SHF + LEA + LDS.64 + FFMA
__device__ __forceinline__ float l_eval(uint32_t tL, float q) {
  const uint32_t addr = ((__float_as_uint(q) >> (23 - kLBits)) << 3) + tL;
  float a, b;
  asm("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(a), "=f"(b) : "r"(addr));
  return fmaf(b, q, a);
}

#ifndef UHDR_APPLYQ_F32X2
#define UHDR_APPLYQ_F32X2 1  // 0: scalar FFMA / FADD everywhere (the pre-FFMA2 kernel)
#endif

#ifndef UHDR_L_MUFU_CHANNELS
#define UHDR_L_MUFU_CHANNELS 1  // channels of the linear-output kernels evaluated as ex2(2.4 * lg2(q)) instead of the L table
#endif
constexpr int kLMufu = UHDR_L_MUFU_CHANNELS;

#if UHDR_APPLYQ_F32X2
// q_rows for RGBA1010102 above the toe, without the `& 0x3ff` wrap, on packed fp32x2 arithmetic
// srgbInvOetf's linear toe in the Phi domain: for c <= 0.04045 (Phi <= kPhiThr) the factor is (c / 12.92)^(1/2.4) with
// c = 1.055 * Phi - 0.055 (clamped at 0: the reference clamps c, gainmapmath.cpp:142-146); 2 MUFU + 4 FMA/ALU-pipe
// operations per channel. At the threshold both branches agree to 2e-6 relative, so deciding on Phi instead of on the
// reference's own c changes no code.
__device__ __forceinline__ float toe_fix(float phi) {
  constexpr float kPhiThr = (float)((0.04045 + 0.055) / 1.055);
  constexpr float kToeA = (float)(1.0 / 2.4), kToeB = (float)(-3.6915341649192007 / 2.4);  // lg2(12.92) = 3.69153...
  const float c = __saturatef(fmaf(phi, 1.055f, -0.055f));
  const float t = mufu_ex2(fmaf(mufu_lg2(c), kToeA, kToeB));  // c == 0: lg2 = -inf, ex2(-inf) = +0
  return phi <= kPhiThr ? t : phi;
}

// kSat: lanes with `satcell` (interior cell, four map samples 255) take the host-computed exact code vs8[.] (+ 0.5, so
// that the shared floor returns it) for every channel clamped to Phi == 1; see ApplyFastParams::vsat.
template <bool kToe = false, bool kSat = false>
__device__ __forceinline__ void q_rows_paired(const uint32_t hiP, const float (&oR)[2], const float (&oG)[2], const float (&oB)[2],
                                              const float (&yf)[2][4], const float* __restrict__ f8, uint8_t* orow, const size_t ostride,
                                              const bool satcell = false, const float* __restrict__ vs8 = nullptr) {
  constexpr float KP = (float)(1.0 / (255.0 * 1.055));
  constexpr float q0 = 1.0f / (float)(1 << -kPLoExp);
  // the two pixels of a quad's row as one f32x2 lane pair: q, the three floors and the field move take one
  // FFMA2 / FADD2 per pair instead of one FFMA / FADD per pixel (same arithmetic, half the issue slots)
  const f32x2 q0q0 = pk2(q0, q0), two23 = pk2(8388608.0f, 8388608.0f), k1024 = pk2(1024.0f, 1024.0f);
  const f32x2 gOff = pk2(8388608.0f - 8589934592.0f, 8388608.0f - 8589934592.0f);
  const f32x2 bOff = pk2(8388608.0f + 1872.0f, 8388608.0f + 1872.0f);
#pragma unroll
  for (int rr = 0; rr < 2; rr++) {
    uint32_t w[4];
#pragma unroll
    for (int c = 0; c < 2; c++) {
      const float y0 = yf[rr][2 * c], y1 = yf[rr][2 * c + 1];
      const f32x2 ff = pk2(f8[rr * 4 + 2 * c], f8[rr * 4 + 2 * c + 1]);
      float pR0 = __saturatef(fmaf(y0, KP, oR[c])), pR1 = __saturatef(fmaf(y1, KP, oR[c]));
      float pG0 = __saturatef(fmaf(y0, KP, oG[c])), pG1 = __saturatef(fmaf(y1, KP, oG[c]));
      float pB0 = __saturatef(fmaf(y0, KP, oB[c])), pB1 = __saturatef(fmaf(y1, KP, oB[c]));
      if (kToe) {  // a quad-row some lane of the warp has a toe channel in: every channel takes the exact select
        pR0 = toe_fix(pR0); pR1 = toe_fix(pR1);
        pG0 = toe_fix(pG0); pG1 = toe_fix(pG1);
        pB0 = toe_fix(pB0); pB1 = toe_fix(pB1);
      }
      const f32x2 qR = fma2(pk2(pR0, pR1), ff, q0q0);
      const f32x2 qG = fma2(pk2(pG0, pG1), ff, q0q0);
      const f32x2 qB = fma2(pk2(pB0, pB1), ff, q0q0);
      float a0, a1;
      unpk2(qR, a0, a1);
      float r0 = p_eval(hiP, a0), r1 = p_eval(hiP, a1);
      unpk2(qG, a0, a1);
      float g0 = p_eval(hiP, a0), g1 = p_eval(hiP, a1);
      unpk2(qB, a0, a1);
      float b0v = p_eval(hiP, a0), b1v = p_eval(hiP, a1);
      if (kSat) {
        const float s0 = vs8[rr * 4 + 2 * c], s1 = vs8[rr * 4 + 2 * c + 1];
        r0 = satcell && pR0 >= 1.0f ? s0 : r0; r1 = satcell && pR1 >= 1.0f ? s1 : r1;
        g0 = satcell && pG0 >= 1.0f ? s0 : g0; g1 = satcell && pG1 >= 1.0f ? s1 : g1;
        b0v = satcell && pB0 >= 1.0f ? s0 : b0v; b1v = satcell && pB1 >= 1.0f ? s1 : b1v;
      }
      const f32x2 vR = pk2(r0, r1), vG = pk2(g0, g1), vB = pk2(b0v, b1v);
      // floor + pack as in the scalar path below, two pixels per operation
      const f32x2 t = add_rm2(vR, fma2(add_rm2(vG, two23), k1024, gOff));
      const f32x2 Bm = add_rm2(vB, bOff);
      float t0, t1, b0, b1;
      unpk2(t, t0, t1);
      unpk2(Bm, b0, b1);
      w[2 * c] = (__float_as_uint(b0) << 20) + __float_as_uint(t0);
      w[2 * c + 1] = (__float_as_uint(b1) << 20) + __float_as_uint(t1);
    }
    stg16(orow + (size_t)rr * ostride, make_uint4(w[0], w[1], w[2], w[3]));
  }
}

// q_rows for the linear outputs above the toe (L table): q = Phi * f' and the 10-bit scale + floor on packed fp32x2
template <int PACK, bool NORM, bool MASK>
__device__ __forceinline__ void q_rows_paired_lin(const uint32_t tL, const float (&oR)[2], const float (&oG)[2], const float (&oB)[2],
                                                  const float (&yf)[2][4], const float* __restrict__ f8, uint8_t* orow, const size_t ostride,
                                                  const size_t plane_bytes) {
  static_assert(PACK != kPackF16 || NORM, "the F16 L-table kernel relies on the host's normal-half guarantee");
  constexpr float KP = (float)(1.0 / (255.0 * 1.055));
#pragma unroll
  for (int rr = 0; rr < 2; rr++) {
    uint32_t w[4][PACK == kPackF16 ? 2 : 3];  // F16: two words per pixel; planar: [pixel pair][channel] u16 pairs
#pragma unroll
    for (int c = 0; c < 2; c++) {
      const float y0 = yf[rr][2 * c], y1 = yf[rr][2 * c + 1];
      const f32x2 ff = pk2(f8[rr * 4 + 2 * c], f8[rr * 4 + 2 * c + 1]);
      float e[3][2];
      {
        // q^2.4: table look-up (LSU) for the first channels, two MUFU for the last kLMufu ones, so that the shared-memory
        // data pipe and the XU pipe share the work
        float a0, a1;
        unpk2(mul2(pk2(__saturatef(fmaf(y0, KP, oR[c])), __saturatef(fmaf(y1, KP, oR[c]))), ff), a0, a1);
        e[0][0] = kLMufu >= 3 ? pow24(a0) : l_eval(tL, a0);
        e[0][1] = kLMufu >= 3 ? pow24(a1) : l_eval(tL, a1);
        unpk2(mul2(pk2(__saturatef(fmaf(y0, KP, oG[c])), __saturatef(fmaf(y1, KP, oG[c]))), ff), a0, a1);
        e[1][0] = kLMufu >= 2 ? pow24(a0) : l_eval(tL, a0);
        e[1][1] = kLMufu >= 2 ? pow24(a1) : l_eval(tL, a1);
        unpk2(mul2(pk2(__saturatef(fmaf(y0, KP, oB[c])), __saturatef(fmaf(y1, KP, oB[c]))), ff), a0, a1);
        e[2][0] = kLMufu >= 1 ? pow24(a0) : l_eval(tL, a0);
        e[2][1] = kLMufu >= 1 ? pow24(a1) : l_eval(tL, a1);
      }
      if (PACK == kPackF16) {
#pragma unroll
        for (int h = 0; h < 2; h++) {
          const uint32_t bR = __float_as_uint(e[0][h]) + 0x1000u, bG = __float_as_uint(e[1][h]) + 0x1000u, bB = __float_as_uint(e[2][h]) + 0x1000u;
          // floatToHalf (gainmapmath.h:136-147) on the normal-half range; the host only selects the L table when
          // every value of a toe-free quad-row is a normal half (lin_table_ok)
          w[2 * c + h][0] = pack_half2_rz(__uint_as_float(bR), __uint_as_float(bG));
          w[2 * c + h][1] = pack_half2_rz(__uint_as_float(bB), 1.0f);
        }
      } else {
        // 0x3ff & uint(v * 1023) (ultrahdr.cpp:460-468): v >= 0, so truncation = floor = FADD.RM
        const f32x2 k1023 = pk2(1023.0f, 1023.0f), two23 = pk2(8388608.0f, 8388608.0f);
#pragma unroll
        for (int ch = 0; ch < 3; ch++) {
          float m0, m1;
          unpk2(add_rm2(mul2(pk2(e[ch][0], e[ch][1]), k1023), two23), m0, m1);
          // the integers sit in the low mantissa bits: one PRMT takes both low halves (dropping the 0x4B00 of the magic
          // number) into the u16 pair; `& 0x3ff` only when a code can exceed 1023 (displayBoost < maxContentBoost)
          const uint32_t pr = __byte_perm(__float_as_uint(m0), __float_as_uint(m1), 0x5410);
          w[c][ch] = MASK ? (pr & 0x03ff03ffu) : pr;
        }
      }
    }
    uint8_t* o = orow + (size_t)rr * ostride;
    if (PACK == kPackF16) {  // colorToRgbaF16, gainmapmath.cpp:729-732
      stg32(o, make_uint4(w[0][0], w[0][1], w[1][0], w[1][1]), make_uint4(w[2][0], w[2][1], w[3][0], w[3][1]));
    } else {  // three planar u16 planes, ultrahdr.cpp:460-468
#pragma unroll
      for (int ch = 0; ch < 3; ch++) stg8(o + ch * plane_bytes, make_uint2(w[0][ch], w[1][ch]));
    }
  }
}
#endif

// two pixel rows (one row of 2x2 chroma quads) of a cell: Phi -> q -> P(q) -> floor -> pack -> 128-bit store per row
// (RGBA1010102), or Phi -> q -> q^2.4 -> half / 10-bit pack (linear outputs)
template <int PACK, bool MASK, bool kToe, bool LTAB, bool NORM>
__device__ __forceinline__ void q_rows(const uint32_t hiP, const uint32_t uq, const uint32_t vq,
                                       const float (&oR)[2], const float (&oG)[2], const float (&oB)[2], const float (&yf)[2][4],
                                       const float* __restrict__ f8, uint8_t* orow, const size_t ostride, const size_t plane_bytes,
                                       const bool satcell = false, const float* __restrict__ vs8 = nullptr, const float* __restrict__ rs8 = nullptr) {
  constexpr float KP = (float)(1.0 / (255.0 * 1.055));
  constexpr float q0 = 1.0f / (float)(1 << -kPLoExp);  // 2^kPLoExp as an FFMA immediate
        float rv[2] = {0.f, 0.f}, gq[2] = {0.f, 0.f}, bu[2] = {0.f, 0.f};
        if (kToe) {  // chroma terms in the reference's own order for the c <= 0.04045 test (ultrahdr.cpp:430-431)
#pragma unroll
          for (int c = 0; c < 2; c++) {
            const float u = (float)((int)((uq >> (8 * c)) & 0xff) - 128) * (1 / 255.0f);
            const float v = (float)((int)((vq >> (8 * c)) & 0xff) - 128) * (1 / 255.0f);
            rv[c] = 1.402f * v;
            gq[c] = -((0.114f * 1.772f / 0.587f) * u + (0.299f * 1.402f / 0.587f) * v);
            bu[c] = 1.772f * u;
          }
        }
#pragma unroll
        for (int rr = 0; rr < 2; rr++) {
          uint32_t w[4][PACK == kPack1010102 ? 1 : (PACK == kPackF16 ? 2 : 3)];
#pragma unroll
          for (int i = 0; i < 4; i++) {
            const int c = i >> 1;
            const float y = yf[rr][i], fi = f8[rr * 4 + i];
            float phR = __saturatef(fmaf(y, KP, oR[c]));
            float phG = __saturatef(fmaf(y, KP, oG[c]));
            float phB = __saturatef(fmaf(y, KP, oB[c]));
            float cR = 0.f, cG = 0.f, cB = 0.f;
            bool tR = false, tG = false, tB = false;
            if (kToe) {
              cR = __saturatef(fmaf(y, 1 / 255.0f, rv[c]));
              cG = __saturatef(fmaf(y, 1 / 255.0f, gq[c]));
              cB = __saturatef(fmaf(y, 1 / 255.0f, bu[c]));
              tR = cR <= 0.04045f;
              tG = cG <= 0.04045f;
              tB = cB <= 0.04045f;
              if (PACK == kPack1010102) {
                // (c / 12.92)^(1/2.4) = 2^(lg2(c) / 2.4 - lg2(12.92) / 2.4); c == 0 gives lg2 = -inf and ex2(-inf) = +0
                constexpr float kToeA = (float)(1.0 / 2.4), kToeB = (float)(-3.6915341649192007 / 2.4);  // lg2(12.92) = 3.69153...
                if (tR) phR = mufu_ex2(fmaf(mufu_lg2(cR), kToeA, kToeB));
                if (tG) phG = mufu_ex2(fmaf(mufu_lg2(cG), kToeA, kToeB));
                if (tB) phB = mufu_ex2(fmaf(mufu_lg2(cB), kToeA, kToeB));
              }
            }
            if (PACK == kPack1010102) {
              // the P table is >= 0 everywhere (tables.cu), so floor == the reference's truncation
              float vR = p_eval(hiP, fmaf(phR, fi, q0));
              float vG = p_eval(hiP, fmaf(phG, fi, q0));
              float vB = p_eval(hiP, fmaf(phB, fi, q0));
              if (kToe && satcell) {  // the special path also serves saturated cells (ApplyFastParams::vsat)
                const float vs = vs8[rr * 4 + i];
                vR = phR >= 1.0f ? vs : vR;
                vG = phG >= 1.0f ? vs : vG;
                vB = phB >= 1.0f ? vs : vB;
              }
              if (MASK) {  // codes above 1023 wrap like the reference's `& 0x3ff` (gainmapmath.cpp:722-727)
                w[i][0] = (floor_bits(vR) & 0x3ffu) | ((floor_bits(vG) & 0x3ffu) << 10) | ((floor_bits(vB) & 0x3ffu) << 20) | 0xC0000000u;
              } else {
                // floor + pack in 5 operations: G floored, moved up 10 bits inside the float (exact FFMA), R floored
                // on top of it by one FADD.RM; B floored with a bias whose << 20 also supplies the alpha bits and
                // cancels the 0x4B000000 of the R|G word.
                const float Gm = add_rm(vG, 8388608.0f);
                const float cGs = fmaf(Gm, 1024.0f, 8388608.0f - 8589934592.0f);  // 2^23 + 1024 * G
                const float t = add_rm(vR, cGs);
                const float Bm = add_rm(vB, 8388608.0f + 1872.0f);  // (0x750 << 20) == 0xC0000000 - 0x4B000000
                w[i][0] = (__float_as_uint(Bm) << 20) + __float_as_uint(t);
              }
            } else {
              // linear light: f * srgbInvOetf(c) = (Phi * f')^2.4 above the toe, f * c / 12.92 on it
              float eR, eG, eB;
              if (LTAB) {  // hiP carries the L table's biased shared address
                // channels on the toe have Phi down to 0: keep their (discarded) look-up inside the table
                constexpr float kQLo = 1.0f / (float)(1 << -kLLoExp);
                const float qR = kToe ? fmaxf(phR * fi, kQLo) : phR * fi;
                const float qG = kToe ? fmaxf(phG * fi, kQLo) : phG * fi;
                const float qB = kToe ? fmaxf(phB * fi, kQLo) : phB * fi;
                eR = l_eval(hiP, qR); eG = l_eval(hiP, qG); eB = l_eval(hiP, qB);
              } else {
                eR = pow24(phR * fi); eG = pow24(phG * fi); eB = pow24(phB * fi);
              }
              if (kToe && satcell) {  // ApplyFastParams::rsat
                const float rs = rs8[rr * 4 + i];
                eR = phR >= 1.0f ? rs : eR;
                eG = phG >= 1.0f ? rs : eG;
                eB = phB >= 1.0f ? rs : eB;
              }
              if (kToe) {
                const float f24 = pow24(fi);
                if (tR) eR = cR * (float)(1.0 / 12.92) * f24;
                if (tG) eG = cG * (float)(1.0 / 12.92) * f24;
                if (tB) eB = cB * (float)(1.0 / 12.92) * f24;
              }
              if (PACK == kPackF16) {
                // floatToHalf (gainmapmath.h:136-147), see apply_fast_kernel: shift + re-bias on the normal-half range
                const uint32_t bR = __float_as_uint(eR) + 0x1000u, bG = __float_as_uint(eG) + 0x1000u, bB = __float_as_uint(eB) + 0x1000u;
                if (NORM && !kToe) {  // host guarantee: every value of a toe-free quad-row is a normal half
                  w[i][0] = pack_half2_rz(__uint_as_float(bR), __uint_as_float(bG));
                  w[i][1] = pack_half2_rz(__uint_as_float(bB), 1.0f);
                } else {
                  uint32_t hR, hG, hB;
                  if (min(bR, min(bG, bB)) >= 0x38800000u) {
                    hR = (bR >> 13) - 0x1C000u;
                    hG = (bG >> 13) - 0x1C000u;
                    hB = (bB >> 13) - 0x1C000u;
                  } else {
                    hR = bR >= 0x38800000u ? (bR >> 13) - 0x1C000u : 0u;
                    hG = bG >= 0x38800000u ? (bG >> 13) - 0x1C000u : 0u;
                    hB = bB >= 0x38800000u ? (bB >> 13) - 0x1C000u : 0u;
                    if (bR - 0x33000000u < 0x05800000u) hR = float_to_half_ref(eR);
                    if (bG - 0x33000000u < 0x05800000u) hG = float_to_half_ref(eG);
                    if (bB - 0x33000000u < 0x05800000u) hB = float_to_half_ref(eB);
                  }
                  w[i][0] = hR | (hG << 16);
                  w[i][1] = hB | (0x3C00u << 16);
                }
              } else {
                // 0x3ff & uint(v * 1023) (ultrahdr.cpp:460-468): v >= 0, so truncation = floor = FADD.RM
                w[i][0] = floor_bits(eR * 1023.0f) & 0x3ffu;
                w[i][1] = floor_bits(eG * 1023.0f) & 0x3ffu;
                w[i][2] = floor_bits(eB * 1023.0f) & 0x3ffu;
              }
            }
          }
          uint8_t* o = orow + (size_t)rr * ostride;
          if (PACK == kPack1010102) {
            stg16(o, make_uint4(w[0][0], w[1][0], w[2][0], w[3][0]));
          } else if (PACK == kPackF16) {  // colorToRgbaF16, gainmapmath.cpp:729-732
            stg32(o, make_uint4(w[0][0], w[0][1], w[1][0], w[1][1]), make_uint4(w[2][0], w[2][1], w[3][0], w[3][1]));
          } else {  // three planar u16 planes, ultrahdr.cpp:460-468
#pragma unroll
            for (int ch = 0; ch < 3; ch++)
              stg8(o + ch * plane_bytes, make_uint2(w[0][ch] | (w[1][ch] << 16), w[2][ch] | (w[3][ch] << 16)));
          }
        }
}

#ifndef UHDR_APPLYQ_MIN_BLOCKS
#define UHDR_APPLYQ_MIN_BLOCKS 3  // 80 registers: room for the software prefetch without spills
#endif

#ifndef UHDR_APPLYQ_L_MIN_BLOCKS
#define UHDR_APPLYQ_L_MIN_BLOCKS 3  // L-table kernels: 3 resident blocks measured faster than 4 (523 vs 509 Gpix/s planar)
#endif

#ifndef UHDR_APPLYQ_PREFETCH
#define UHDR_APPLYQ_PREFETCH 1  // loads of cell k + 1 issued before cell k is evaluated (needs the 80-register budget)
#endif

// Stages the block's table (P for RGBA1010102, L for the linear outputs with LTAB) with one TMA bulk copy and returns
// the biased shared address the look-ups use (0 when the instantiation uses no table).
template <int PACK, bool LTAB>
__device__ __forceinline__ uint32_t stage_q_table(const ApplyFastParams& p, const int tid) {
  uint32_t hiP = 0;
  if (LTAB) {
    // dynamic shared memory: [mbarrier 16 B][L table]
    extern __shared__ __align__(16) uint8_t smem_dyn[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem_dyn);
    if (tid == 0) mbar_init(bar, 1);
    __syncthreads();
    if (tid == 0) {
      constexpr uint32_t lb = kLEntries * sizeof(float2);
      mbar_expect_tx(bar, lb);
      bulk_g2s(smem_dyn + 16, p.tab.l_pow, lb, bar);
    }
    mbar_wait(bar, 0);
    hiP = smem_u32(smem_dyn + 16) - 8u * (uint32_t)kLBase;
  } else if (PACK == kPack1010102) {
    // dynamic shared memory: [mbarrier 16 B] ... [P table at shared address == kPRing (mod the ring size)]
    extern __shared__ __align__(16) uint8_t smem_dyn[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem_dyn);
    const uint32_t dyn0 = smem_u32(smem_dyn);
    const uint32_t pAddr = q_table_addr(dyn0);
    {
      uint32_t dyn_size;
      asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(dyn_size));
      if (pAddr - dyn0 + kPEntries * (uint32_t)sizeof(float2) > dyn_size) __trap();
    }
    hiP = pAddr - kPRing;
    if (tid == 0) mbar_init(bar, 1);
    __syncthreads();
    if (tid == 0) {
      constexpr uint32_t pb = kPEntries * sizeof(float2);
      mbar_expect_tx(bar, pb);
      bulk_g2s(smem_dyn + (pAddr - dyn0), p.output_format == 3 ? p.tab.p_hlg : p.tab.p_pq, pb, bar);
    }
    mbar_wait(bar, 0);
  }
  return hiP;
}

// One quad-row of a 4x4 block (2 pixel rows x 4 columns = two 2x2 chroma quads) given its 8 gain factors f' = f^(1/2.4):
// Phi-domain chroma offsets, int -> float luma, the conservative toe test and the look-up / pack / store of both rows.
// yw0 / yw1 = the 4 luma bytes of the two rows, uq / vq = the two chroma bytes, orow = output address of the first row.
// VOTE: `vote_mask` names the lanes executing this call together; when any of them has a toe channel the whole
// group takes the toe-capable path (one pass through one path instead of both paths one after the other).
template <int PACK, bool MASK, bool LTAB, bool NORM, bool VOTE = false>
__device__ __forceinline__ void quad_row_eval(const uint32_t hiP, const uint32_t yw0, const uint32_t yw1, const uint32_t uq, const uint32_t vq,
                                              const float (&f)[8], uint8_t* orow, const size_t ostride, const size_t plane_bytes,
                                              const unsigned vote_mask = 0xffffffffu, const bool satcell = false,
                                              const float* __restrict__ vs8 = nullptr, const float* __restrict__ rs8 = nullptr) {
  // Phi-domain constants (double on the host side of the compiler): K = 1/(255*1.055), chroma gains of the
  // BT.601 matrix (gainmapmath.cpp:177-196) times K, offsets = 0.055/1.055 - 128 * gain
  constexpr double dK = 1.0 / (255.0 * 1.055);
  constexpr double dCr = 1.402, dCb = 1.772;
  constexpr double dGcb = 0.114 * 1.772 / 0.587, dGcr = 0.299 * 1.402 / 0.587;
  constexpr float KP = (float)dK;
  constexpr float kR = (float)(dCr * dK), kB = (float)(dCb * dK), kGu = (float)(-dGcb * dK), kGv = (float)(-dGcr * dK);
  constexpr float oR0 = (float)(0.055 / 1.055 - 128.0 * dCr * dK), oB0 = (float)(0.055 / 1.055 - 128.0 * dCb * dK);
  constexpr float oG0 = (float)(0.055 / 1.055 + 128.0 * (dGcb + dGcr) * dK);
  constexpr float kPhiToe = 0.0905f;  // Phi(0.04045) = 0.090474; anything at or below 0.0905 takes the exact branch
        // Phi-domain chroma offsets of the two 2x2 quads of this quad-row
        float oR[2], oG[2], oB[2], yf[2][4];
#if UHDR_APPLYQ_F32X2
        {  // both quads per FFMA2
          const f32x2 UB = pk2((float)(uq & 0xff), (float)((uq >> 8) & 0xff));
          const f32x2 VB = pk2((float)(vq & 0xff), (float)((vq >> 8) & 0xff));
          unpk2(fma2(VB, pk2(kR, kR), pk2(oR0, oR0)), oR[0], oR[1]);
          unpk2(fma2(UB, pk2(kB, kB), pk2(oB0, oB0)), oB[0], oB[1]);
          unpk2(fma2(VB, pk2(kGv, kGv), fma2(UB, pk2(kGu, kGu), pk2(oG0, oG0))), oG[0], oG[1]);
        }
#else
#pragma unroll
        for (int c = 0; c < 2; c++) {
          const float ub = (float)((uq >> (8 * c)) & 0xff), vb = (float)((vq >> (8 * c)) & 0xff);
          oR[c] = fmaf(vb, kR, oR0);
          oB[c] = fmaf(ub, kB, oB0);
          oG[c] = fmaf(vb, kGv, fmaf(ub, kGu, oG0));
        }
#endif
#pragma unroll
        for (int rr = 0; rr < 2; rr++)
#pragma unroll
          for (int i = 0; i < 4; i++) yf[rr][i] = (float)(((rr ? yw1 : yw0) >> (8 * i)) & 0xff);
        const float ym0 = fminf(fminf(yf[0][0], yf[0][1]), fminf(yf[1][0], yf[1][1]));
        const float ym1 = fminf(fminf(yf[0][2], yf[0][3]), fminf(yf[1][2], yf[1][3]));
        const float om0 = fminf(oR[0], fminf(oG[0], oB[0])), om1 = fminf(oR[1], fminf(oG[1], oB[1]));
#if UHDR_APPLYQ_F32X2
        float ph0, ph1;
        unpk2(fma2(pk2(ym0, ym1), pk2(KP, KP), pk2(om0, om1)), ph0, ph1);
        const bool toe = fminf(ph0, ph1) <= kPhiToe;
#else
        const bool toe = fminf(fmaf(ym0, KP, om0), fmaf(ym1, KP, om1)) <= kPhiToe;
#endif
#ifndef UHDR_SAT_VARIANT
#define UHDR_SAT_VARIANT 0  // 1: a third instantiation of the packed path for saturated cells without toe channels (more code, spills)
#endif
        if constexpr (PACK == kPack1010102 && !MASK && UHDR_APPLYQ_F32X2) {
          // the packed path, and one special path that serves both the toe channels and the exact codes of saturated cells
          const bool t = VOTE ? __any_sync(vote_mask, toe) : toe;
          const bool sc = VOTE ? __any_sync(vote_mask, satcell) : satcell;
          if (UHDR_SAT_VARIANT ? t : (t || sc)) {
            if constexpr (VOTE) q_rows_paired<true, true>(hiP, oR, oG, oB, yf, f, orow, ostride, satcell, vs8);
            else q_rows<PACK, MASK, true, LTAB, NORM>(hiP, uq, vq, oR, oG, oB, yf, f, orow, ostride, plane_bytes, satcell, vs8, rs8);
          } else if (UHDR_SAT_VARIANT && sc) {
            q_rows_paired<false, true>(hiP, oR, oG, oB, yf, f, orow, ostride, satcell, vs8);
          } else {
            q_rows_paired<false, false>(hiP, oR, oG, oB, yf, f, orow, ostride);
          }
          return;
        }
        const bool special = VOTE ? __any_sync(vote_mask, toe || satcell) : (toe || satcell);
        if (special) q_rows<PACK, MASK, true, LTAB, NORM>(hiP, uq, vq, oR, oG, oB, yf, f, orow, ostride, plane_bytes, satcell, vs8, rs8);
#if UHDR_APPLYQ_F32X2
        else if constexpr (LTAB) q_rows_paired_lin<PACK, NORM, MASK>(hiP, oR, oG, oB, yf, f, orow, ostride, plane_bytes);
#endif
        else q_rows<PACK, MASK, false, LTAB, NORM>(hiP, uq, vq, oR, oG, oB, yf, f, orow, ostride, plane_bytes);
}

}  // namespace
}  // namespace uhdr
