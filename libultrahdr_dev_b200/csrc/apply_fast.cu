// apply_fast.cu — A2: UltraHdr::applyGainMap (reference lib/src/ultrahdr.cpp:427-496) on table-driven kernels for
// 4-aligned planes, restructured so that it is bound by instruction issue / HBM instead of the SFU. Three kernels:
//
//   apply_q_kernel   map scale factor 4 (kMapDimensionScaleFactor: what every libultrahdr encode writes). The kernel
//                    the benchmark measures. One thread = one gain-map cell = 4x4 pixels, a lane walks CELLS cells
//                    downwards with the next cell's loads in flight; ONE look-up per channel in the q domain (below).
//   apply_qs_kernel  any integer scale factor (specialised for 1 and 2): thread = 4x4 pixel block, general Shepard-IDW
//                    sampling, the same q-domain evaluation (stage_q_table, quad_row_eval are shared).
//   apply_fast_kernel  the previous two-table design (srgbInvOetf table + OETF table), kept selectable for A/B runs and
//                    for linear outputs whose values can leave the normal-half range.
//
// Common structure:
//   * the four surrounding map samples, the 2x2 chroma samples and the plane pointers are fetched once per 16 pixels;
//     luma comes in as four 32-bit loads (a warp reads 128 contiguous bytes per row) and the packed output leaves as
//     128/256-bit stores (a warp writes 512/1024 contiguous bytes per row).
//   * the Shepard-IDW weights of the interior table (gainmapmath.h:184-228) sit in the kernel parameter block, i.e. in
//     the constant bank: with the 4x4 loop fully unrolled every weight is a constant-bank operand of an FFMA/FFMA2.
//     Cells on the right/bottom map border (tables NR/NB/C, gainmapmath.cpp:710-716) read their weights from global memory.
//   * transfer functions (3 pow per pixel + sqrt/log or 9 pow in the reference) are piecewise-linear tables in shared
//     memory on log-spaced segments: the float's exponent/mantissa bits are the index, one 64-bit LDS + one FFMA per
//     evaluation. The tables (built in double on the host, tables.cu) are staged with one TMA bulk copy per block and
//     the blocks are persistent (grid = SMs x resident blocks, grid-stride over tiles).
//   * the gain factor 2^(logBoost * displayBoost / maxContentBoost) / displayBoost (gainmapmath.cpp:550-555,
//     ultrahdr.cpp:451) is three FFMA (FFMA2 on pixel pairs) + one MUFU.EX2 per pixel.
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

// PACK selects the output packing; MASK (RGBA1010102 only) keeps the reference's `& 0x3ff` wrap for
// codes above 1023, which can only occur when displayBoost < maxContentBoost.
// The RGBA1010102 kernel holds its tables in STATIC shared memory (S 12 KB + O 28 KB + 1 KB) so
// that a lookup address is `mask(bits >> n) + link-time constant` (no base-register add); the
// linear-output kernels use a 4x finer S table (48 KB, dynamic shared memory) and no O table.
template <int PACK, bool MASK>
__global__ void __launch_bounds__(kThreads, UHDR_APPLY_MIN_BLOCKS) apply_fast_kernel(const __grid_constant__ ApplyFastParams p) {
  constexpr bool kHasO = PACK == kPack1010102;
  constexpr int kSSeg = kHasO ? kSSegs : kSSegsHi;
  constexpr int kSN = kHasO ? kSEntries : kSEntriesHi;
  extern __shared__ __align__(16) uint8_t smem_dyn[];
  __shared__ __align__(16) float2 sS_static[kHasO ? kSEntries : 1];
  __shared__ __align__(16) float2 sO_static[kHasO ? kOEntries : 1];
  __shared__ __align__(16) float sE[256];
  __shared__ __align__(8) uint64_t bar;
  float2* sS = kHasO ? sS_static : reinterpret_cast<float2*>(smem_dyn);
  float2* sO = sO_static;
  const int tid = threadIdx.x;

  // ---- stage the tables: one elected thread issues the TMA bulk copies, everyone waits on the mbarrier
  if (tid == 0) mbar_init(&bar, 1);
  __syncthreads();
  if (tid == 0) {
    constexpr uint32_t sb = kSN * sizeof(float2), ob = kOEntries * sizeof(float2), eb = 256 * sizeof(float);
    mbar_expect_tx(&bar, sb + eb + (kHasO ? ob : 0));
    bulk_g2s(sS, kHasO ? p.tab.s : p.tab.s_hi, sb, &bar);
    if (kHasO) bulk_g2s(sO, p.output_format == 3 ? p.tab.o_hlg : p.tab.o_pq, ob, &bar);
    bulk_g2s(sE, p.tab.e255, eb, &bar);
  }
  mbar_wait(&bar, 0);
  // p.s_bias = 8 * 0x4B000000 (mod 2^32) comes from the parameter block so that it stays a run-time value and the
  // entry address is one LEA: (bits << 3) + tS
  const uint32_t tS = smem_u32(sS) - p.s_bias;
  const float2* tO = sO - kOBase;  // index = bits >> (23 - kOBits)

  const Geometry& g = p.g;
  const int mw = g.map_width, mh = g.map_height;
  const int tiles_per_image = p.tiles_x * p.tiles_y;
  const float k601_cr = 1.402f, k601_cb = 1.772f;
  const float k601_gcb = 0.114f * 1.772f / 0.587f, k601_gcr = 0.299f * 1.402f / 0.587f;  // gainmapmath.cpp:195-196
  const size_t npix = (size_t)g.width * g.height;

  // tile -> (image z, tile row ty, tile column tx), advanced incrementally: no division in the loop
  const uint32_t tiles_x = p.tiles_x, tiles_y = p.tiles_y;
  uint32_t tx, ty, z, dtx, dty, dz;
  {
    const uint32_t b0 = blockIdx.x, st = gridDim.x;
    z = b0 / tiles_per_image;
    uint32_t rem = b0 - z * tiles_per_image;
    ty = rem / tiles_x;
    tx = rem - ty * tiles_x;
    dz = st / tiles_per_image;
    rem = st - dz * tiles_per_image;
    dty = rem / tiles_x;
    dtx = rem - dty * tiles_x;
  }
  const uint32_t ls = g.sdr_luma_stride, cs = g.sdr_chroma_stride;

  for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
    const uint32_t cx = tx * kTileCellsX + (tid & 31), cy = ty * kTileCellsY + (tid >> 5);
    const uint32_t zi = z;
    tx += dtx;
    if (tx >= tiles_x) { tx -= tiles_x; ty++; }
    ty += dty;
    if (ty >= tiles_y) { ty -= tiles_y; z++; }
    z += dz;
    if (cx >= (uint32_t)mw || cy >= (uint32_t)mh) continue;
    const uint8_t *sdr_y, *sdr_u, *sdr_v, *map;
    uint8_t* outp;
    if (p.batch) {
      const Planes* pl = p.batch + zi;
      sdr_y = (const uint8_t*)__ldg((const unsigned long long*)&pl->sdr_y);
      sdr_u = (const uint8_t*)__ldg((const unsigned long long*)&pl->sdr_u);
      sdr_v = (const uint8_t*)__ldg((const unsigned long long*)&pl->sdr_v);
      map = (const uint8_t*)__ldg((const unsigned long long*)&pl->map_in);
      outp = (uint8_t*)__ldg((const unsigned long long*)&pl->out);
    } else {
      sdr_y = (const uint8_t*)p.single.sdr_y;
      sdr_u = (const uint8_t*)p.single.sdr_u;
      sdr_v = (const uint8_t*)p.single.sdr_v;
      map = (const uint8_t*)p.single.map_in;
      outp = (uint8_t*)p.single.out;
    }
    // 32-bit offsets inside a plane (planes are < 4 GB: width, height <= 2^15 after the capi checks)
    const uint8_t* yp = sdr_y + (4u * cy * ls + 4u * cx);
    const uint32_t coff = 2u * cy * cs + 2u * cx;
    const uint8_t* up = sdr_u + coff;
    const uint8_t* vp = sdr_v + coff;

    // ---- all global loads of the cell first
    uint32_t yw[4], uw[2], vw[2];
#pragma unroll
    for (int r = 0; r < 4; r++) yw[r] = ldg4(yp + r * ls);
#pragma unroll
    for (int r = 0; r < 2; r++) {
      uw[r] = ldg2(up + r * cs);
      vw[r] = ldg2(vp + r * cs);
    }
    const uint32_t cxr = min(cx + 1, (uint32_t)mw - 1), cyb = min(cy + 1, (uint32_t)mh - 1);  // gainmapmath.cpp:690-697
    const uint32_t b1 = __ldg(map + (cx + cy * mw)), b2 = __ldg(map + (cx + cyb * mw));
    const uint32_t b3 = __ldg(map + (cxr + cy * mw)), b4 = __ldg(map + (cxr + cyb * mw));
    const float e1 = sE[b1], e2 = sE[b2], e3 = sE[b3], e4 = sE[b4];  // e1=(xl,yl) e2=(xl,yu) e3=(xu,yl) e4=(xu,yu)

    // ---- gain factor of the 16 pixels: f = 2^(gain * A + B); p.w holds the IDW weights times A
    float f[16];
    if (cx + 1 < (uint32_t)mw && cy + 1 < (uint32_t)mh) {
#pragma unroll
      for (int i = 0; i < 16; i++) {
        float ex = fmaf(e1, p.w[i][0], p.B);
        ex = fmaf(e2, p.w[i][1], ex);
        ex = fmaf(e3, p.w[i][2], ex);
        ex = fmaf(e4, p.w[i][3], ex);
        f[i] = ex2_fast(ex);
      }
    } else {  // map border: tables NR (1), NB (2), C (3)
      const int t = (cx + 1 < (uint32_t)mw) ? 2 : ((cy + 1 < (uint32_t)mh) ? 1 : 3);
      const float4* wt = reinterpret_cast<const float4*>(p.idw) + t * 16;
#pragma unroll
      for (int i = 0; i < 16; i++) {
        const float4 w = __ldg(wt + i);
        float gain = e1 * w.x;
        gain = fmaf(e2, w.y, gain);
        gain = fmaf(e3, w.z, gain);
        gain = fmaf(e4, w.w, gain);
        f[i] = ex2_fast(fmaf(gain, p.A, p.B));
      }
    }

    // ---- chroma terms of the four 2x2 quads: R = y + cr*v, G = y - gcb*u - gcr*v, B = y + cb*u (always BT.601, ultrahdr.cpp:430-431)
    float rv[2][2], gq[2][2], bu[2][2];
#pragma unroll
    for (int r = 0; r < 2; r++)
#pragma unroll
      for (int c = 0; c < 2; c++) {
        const float u = (float)((int)((uw[r] >> (8 * c)) & 0xff) - 128) * (1 / 255.0f);
        const float v = (float)((int)((vw[r] >> (8 * c)) & 0xff) - 128) * (1 / 255.0f);
        rv[r][c] = k601_cr * v;
        gq[r][c] = -(k601_gcb * u + k601_gcr * v);
        bu[r][c] = k601_cb * u;
      }

#pragma unroll
    for (int r = 0; r < 4; r++) {
      uint32_t px[4][3];  // 10-bit codes or half bits
#pragma unroll
      for (int i = 0; i < 4; i++) {
        const float yf = (float)((yw[r] >> (8 * i)) & 0xff);
        const float fi = f[r * 4 + i];
        const float cR = __saturatef(fmaf(yf, 1 / 255.0f, rv[r >> 1][i >> 1]));
        const float cG = __saturatef(fmaf(yf, 1 / 255.0f, gq[r >> 1][i >> 1]));
        const float cB = __saturatef(fmaf(yf, 1 / 255.0f, bu[r >> 1][i >> 1]));
        constexpr int kX = kHasO ? UHDR_APPLY_MUFU_CHANNELS : UHDR_APPLY_MUFU_CHANNELS_LINEAR;
        const float sR = kX >= 2 ? srgb_inv_mufu(cR) : uni_eval<kSSeg>(tS, cR);
        const float sG = kX >= 3 ? srgb_inv_mufu(cG) : uni_eval<kSSeg>(tS, cG);
        const float sB = kX >= 1 ? srgb_inv_mufu(cB) : uni_eval<kSSeg>(tS, cB);
        if (PACK == kPack1010102) {
          // the O table is >= 0 everywhere (tables.cu), so floor == the reference's truncation
          px[i][0] = floor_bits(seg_eval<23 - kOBits>(tO, fmaf(sR, fi, p.e0)));
          px[i][1] = floor_bits(seg_eval<23 - kOBits>(tO, fmaf(sG, fi, p.e0)));
          px[i][2] = floor_bits(seg_eval<23 - kOBits>(tO, fmaf(sB, fi, p.e0)));
          if (MASK) { px[i][0] &= 0x3ffu; px[i][1] &= 0x3ffu; px[i][2] &= 0x3ffu; }
        } else if (PACK == kPackF16) {
          // floatToHalf (gainmapmath.h:136-147). Values here are in [0, 4): for the normal-half range
          // (exponent field of bits + 0x1000 >= 113) the routine reduces to a shift and a re-bias;
          // darker values (half subnormals, flush to zero) take the full bit-exact routine.
          const float eR = sR * fi, eG = sG * fi, eB = sB * fi;
          const uint32_t bR = __float_as_uint(eR) + 0x1000u, bG = __float_as_uint(eG) + 0x1000u, bB = __float_as_uint(eB) + 0x1000u;
          if (min(bR, min(bG, bB)) >= 0x38800000u) {
            px[i][0] = (bR >> 13) - 0x1C000u;
            px[i][1] = (bG >> 13) - 0x1C000u;
            px[i][2] = (bB >> 13) - 0x1C000u;
          } else {  // per channel: normal / flushed to zero (exponent field <= 101) / half subnormal (rare)
            px[i][0] = bR >= 0x38800000u ? (bR >> 13) - 0x1C000u : 0u;
            px[i][1] = bG >= 0x38800000u ? (bG >> 13) - 0x1C000u : 0u;
            px[i][2] = bB >= 0x38800000u ? (bB >> 13) - 0x1C000u : 0u;
            if (bR - 0x33000000u < 0x05800000u) px[i][0] = float_to_half_ref(eR);
            if (bG - 0x33000000u < 0x05800000u) px[i][1] = float_to_half_ref(eG);
            if (bB - 0x33000000u < 0x05800000u) px[i][2] = float_to_half_ref(eB);
          }
        } else {
          px[i][0] = to_10bit(sR * fi);
          px[i][1] = to_10bit(sG * fi);
          px[i][2] = to_10bit(sB * fi);
        }
      }
      const size_t pix = (size_t)((4u * cy + r) * (uint32_t)g.width + 4u * cx);
      if (PACK == kPack1010102) {  // colorToRgba1010102, gainmapmath.cpp:722-727
        uint32_t w[4];
        // px = 0x4B000000 + code (or the masked code); codes <= 1023, so the fields do not overlap and
        // "| alpha" and the removal of the float bias fold into one additive constant
        constexpr uint32_t kBias = MASK ? 0u : (0x4B000000u + (0x4B000000u << 10) + (0x4B000000u << 20));
        constexpr uint32_t kAdd = 0xC0000000u - kBias;
#pragma unroll
        for (int i = 0; i < 4; i++) w[i] = px[i][0] + (px[i][1] << 10) + (px[i][2] << 20) + kAdd;
        stg16(outp + pix * 4, make_uint4(w[0], w[1], w[2], w[3]));
      } else if (PACK == kPackF16) {  // colorToRgbaF16, gainmapmath.cpp:729-732
        uint32_t w[8];
#pragma unroll
        for (int i = 0; i < 4; i++) {
          w[2 * i] = px[i][0] | (px[i][1] << 16);
          w[2 * i + 1] = px[i][2] | (0x3C00u << 16);
        }
        stg16(outp + pix * 8, make_uint4(w[0], w[1], w[2], w[3]));
        stg16(outp + pix * 8 + 16, make_uint4(w[4], w[5], w[6], w[7]));
      } else {  // three planar u16 planes, ultrahdr.cpp:460-468
#pragma unroll
        for (int ch = 0; ch < 3; ch++)
          stg8(outp + (ch * npix + pix) * 2, make_uint2(px[0][ch] | (px[1][ch] << 16), px[2][ch] | (px[3][ch] << 16)));
      }
    }
  }
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
__device__ __forceinline__ void q_rows_paired(const uint32_t hiP, const float (&oR)[2], const float (&oG)[2], const float (&oB)[2],
                                              const float (&yf)[2][4], const float* __restrict__ f8, uint8_t* orow, const size_t ostride) {
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
      const f32x2 qR = fma2(pk2(__saturatef(fmaf(y0, KP, oR[c])), __saturatef(fmaf(y1, KP, oR[c]))), ff, q0q0);
      const f32x2 qG = fma2(pk2(__saturatef(fmaf(y0, KP, oG[c])), __saturatef(fmaf(y1, KP, oG[c]))), ff, q0q0);
      const f32x2 qB = fma2(pk2(__saturatef(fmaf(y0, KP, oB[c])), __saturatef(fmaf(y1, KP, oB[c]))), ff, q0q0);
      float a0, a1;
      unpk2(qR, a0, a1);
      const f32x2 vR = pk2(p_eval(hiP, a0), p_eval(hiP, a1));
      unpk2(qG, a0, a1);
      const f32x2 vG = pk2(p_eval(hiP, a0), p_eval(hiP, a1));
      unpk2(qB, a0, a1);
      const f32x2 vB = pk2(p_eval(hiP, a0), p_eval(hiP, a1));
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
                                       const float* __restrict__ f8, uint8_t* orow, const size_t ostride, const size_t plane_bytes) {
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
              const float vR = p_eval(hiP, fmaf(phR, fi, q0));
              const float vG = p_eval(hiP, fmaf(phG, fi, q0));
              const float vB = p_eval(hiP, fmaf(phB, fi, q0));
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
template <int PACK, bool MASK, bool LTAB, bool NORM>
__device__ __forceinline__ void quad_row_eval(const uint32_t hiP, const uint32_t yw0, const uint32_t yw1, const uint32_t uq, const uint32_t vq,
                                              const float (&f)[8], uint8_t* orow, const size_t ostride, const size_t plane_bytes) {
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
        if (toe) q_rows<PACK, MASK, true, LTAB, NORM>(hiP, uq, vq, oR, oG, oB, yf, f, orow, ostride, plane_bytes);
#if UHDR_APPLYQ_F32X2
        else if constexpr (PACK == kPack1010102 && !MASK) q_rows_paired(hiP, oR, oG, oB, yf, f, orow, ostride);
        else if constexpr (LTAB) q_rows_paired_lin<PACK, NORM, MASK>(hiP, oR, oG, oB, yf, f, orow, ostride, plane_bytes);
#endif
        else q_rows<PACK, MASK, false, LTAB, NORM>(hiP, uq, vq, oR, oG, oB, yf, f, orow, ostride, plane_bytes);
}

// Tile = 32 x 8 * CELLS map cells (128 x 128 pixels for CELLS = 4). A lane owns one column of
// cells; a warp walks CELLS cell rows downwards, so the lower pair of map samples of one cell is
// the upper pair of the next (2 map loads per cell) and every plane pointer advances by a constant.
// CELLS = p.cells is chosen per launch (pick_cells): 8 for large batches (fewer tile prologues: 740 vs 728 Gpix/s at 4,
// 711 at 16), and for a launch of only a few hundred tiles the height that fills the persistent grid in the fewest
// rounds (a single 4K image: 5 cells -> 420 tiles in one round instead of 510 in two).
template <int PACK, bool MASK, bool BATCH, bool LTAB, bool NORM>
__global__ void __launch_bounds__(kQThreads, LTAB ? UHDR_APPLYQ_L_MIN_BLOCKS : UHDR_APPLYQ_MIN_BLOCKS) apply_q_kernel(const __grid_constant__ ApplyFastParams p) {
  const uint32_t kQCells = (uint32_t)p.cells, kQTileRows = (kQThreads / 32) * kQCells;  // cell rows per tile: warps x CELLS
  const int tid = threadIdx.x;
  const uint32_t hiP = stage_q_table<PACK, LTAB>(p, tid);

  const Geometry& g = p.g;
  const uint32_t mw = g.map_width, mh = g.map_height;
  const uint32_t tiles_per_image = p.tiles_x * p.tiles_y;

  // tile -> (image z, tile row ty, tile column tx), advanced incrementally: no division in the loop
  const uint32_t tiles_x = p.tiles_x, tiles_y = p.tiles_y;
  uint32_t tx, ty, z, dtx, dty, dz;
  {
    const uint32_t b0 = blockIdx.x, st = gridDim.x;
    z = b0 / tiles_per_image;
    uint32_t rem = b0 - z * tiles_per_image;
    ty = rem / tiles_x;
    tx = rem - ty * tiles_x;
    dz = st / tiles_per_image;
    rem = st - dz * tiles_per_image;
    dty = rem / tiles_x;
    dtx = rem - dty * tiles_x;
  }
  const uint32_t ls = g.sdr_luma_stride, cs = g.sdr_chroma_stride;
  constexpr int kBpp = PACK == kPack1010102 ? 4 : (PACK == kPackF16 ? 8 : 2);  // bytes per pixel (per plane for planar)
  const size_t ostride = (size_t)g.width * kBpp;
  const size_t plane_bytes = (size_t)g.width * g.height * 2;

  for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
    const uint32_t cx = tx * kTileCellsX + (tid & 31);
    uint32_t cy = ty * kQTileRows + (tid >> 5) * kQCells;
    const uint32_t zi = z;
    tx += dtx;
    if (tx >= tiles_x) { tx -= tiles_x; ty++; }
    ty += dty;
    if (ty >= tiles_y) { ty -= tiles_y; z++; }
    z += dz;
    if (cx >= mw || cy >= mh) continue;
    const uint8_t *sdr_y, *sdr_u, *sdr_v, *map;
    uint8_t* outp;
    if (BATCH) {
      const Planes* pl = p.batch + zi;
      sdr_y = (const uint8_t*)__ldg((const unsigned long long*)&pl->sdr_y);
      sdr_u = (const uint8_t*)__ldg((const unsigned long long*)&pl->sdr_u);
      sdr_v = (const uint8_t*)__ldg((const unsigned long long*)&pl->sdr_v);
      map = (const uint8_t*)__ldg((const unsigned long long*)&pl->map_in);
      outp = (uint8_t*)__ldg((const unsigned long long*)&pl->out);
    } else {
      sdr_y = (const uint8_t*)p.single.sdr_y;
      sdr_u = (const uint8_t*)p.single.sdr_u;
      sdr_v = (const uint8_t*)p.single.sdr_v;
      map = (const uint8_t*)p.single.map_in;
      outp = (uint8_t*)p.single.out;
    }
    // 32-bit offsets inside a plane (planes are < 2 GB after the capi checks)
    const uint8_t* yp = sdr_y + (4u * cy * ls + 4u * cx);
    const uint32_t coff = 2u * cy * cs + 2u * cx;
    const uint8_t* up = sdr_u + coff;
    const uint8_t* vp = sdr_v + coff;
    const uint8_t* mp = map + (cy * mw + cx);
    uint8_t* orow = outp + (size_t)(4u * cy * (uint32_t)g.width + 4u * cx) * kBpp;
    const uint32_t dx = cx + 1 < mw ? 1u : 0u;  // xu = min(xl + 1, mw - 1), gainmapmath.cpp:690-697
    // map samples as 0..255 floats: the 1/255 of gainmapmath.cpp:700-703 is folded into the weights
    float e1 = (float)__ldg(mp), e3 = (float)__ldg(mp + dx);  // e1=(xl,yl) e3=(xu,yl)
    const uint32_t kmax = min(kQCells, mh - cy);

#if UHDR_APPLYQ_PREFETCH
    // software pipeline: the 10 loads of cell k + 1 are in flight while cell k is evaluated
    uint32_t nyw[4], nuw[2], nvw[2], ne2, ne4;
    bool nlast;
    auto fetch = [&](uint32_t cyy) {
#pragma unroll
      for (int r = 0; r < 4; r++) nyw[r] = ldg4(yp + r * ls);
#pragma unroll
      for (int r = 0; r < 2; r++) {
        nuw[r] = ldg2(up + r * cs);
        nvw[r] = ldg2(vp + r * cs);
      }
      nlast = cyy + 1 >= mh;  // yu = min(yl + 1, mh - 1)
      const uint8_t* mp2 = nlast ? mp : mp + mw;
      ne2 = __ldg(mp2);
      ne4 = __ldg(mp2 + dx);
      yp += 4u * ls;
      up += 2u * cs;
      vp += 2u * cs;
      mp = mp2;
    };
    fetch(cy);
#endif
#pragma unroll 1
    for (uint32_t k = 0; k < kmax; k++, cy++) {
      uint32_t yw[4], uw[2], vw[2];
#if UHDR_APPLYQ_PREFETCH
#pragma unroll
      for (int r = 0; r < 4; r++) yw[r] = nyw[r];
#pragma unroll
      for (int r = 0; r < 2; r++) {
        uw[r] = nuw[r];
        vw[r] = nvw[r];
      }
      const bool last = nlast;
      const float e2 = (float)ne2, e4 = (float)ne4;  // e2=(xl,yu) e4=(xu,yu)
      if (k + 1 < kmax) fetch(cy + 1);
#else
#pragma unroll
      for (int r = 0; r < 4; r++) yw[r] = ldg4(yp + r * ls);
#pragma unroll
      for (int r = 0; r < 2; r++) {
        uw[r] = ldg2(up + r * cs);
        vw[r] = ldg2(vp + r * cs);
      }
      const bool last = cy + 1 >= mh;  // yu = min(yl + 1, mh - 1)
      const uint8_t* mp2 = last ? mp : mp + mw;
      const float e2 = (float)__ldg(mp2), e4 = (float)__ldg(mp2 + dx);  // e2=(xl,yu) e4=(xu,yu)
      yp += 4u * ls;
      up += 2u * cs;
      vp += 2u * cs;
      mp = mp2;
#endif

      const float d2 = e2 - e1, d3 = e3 - e1, d4 = e4 - e1, exb = fmaf(e1, p.Aq, p.Bq);
      const bool interior = dx && !last;

#pragma unroll
      for (int qr = 0; qr < 2; qr++) {
        // ---- f' = f^(1/2.4) of the 8 pixels of this quad-row
        float f[8];
        if (interior) {
          // the four weights sum to 1: sum_k e_k w_k = e1 + sum_{k>1} (e_k - e1) w_k, three FFMAs per pixel
#if UHDR_APPLYQ_F32X2
          const f32x2 D2 = pk2(d2, d2), D3 = pk2(d3, d3), D4 = pk2(d4, d4), EXB = pk2(exb, exb);
#pragma unroll
          for (int j = 0; j < 4; j++) {  // ... issued as FFMA2 on horizontally adjacent pixel pairs
            const int i = qr * 8 + 2 * j;
            f32x2 ex = fma2(D2, *reinterpret_cast<const f32x2*>(&p.wq[0][i]), EXB);
            ex = fma2(D3, *reinterpret_cast<const f32x2*>(&p.wq[1][i]), ex);
            ex = fma2(D4, *reinterpret_cast<const f32x2*>(&p.wq[2][i]), ex);
            float a, b;
            unpk2(ex, a, b);
            f[2 * j] = ex2_fast(a);
            f[2 * j + 1] = ex2_fast(b);
          }
#else
#pragma unroll
          for (int i = 0; i < 8; i++) {
            float ex = fmaf(d2, p.wq[0][qr * 8 + i], exb);
            ex = fmaf(d3, p.wq[1][qr * 8 + i], ex);
            ex = fmaf(d4, p.wq[2][qr * 8 + i], ex);
            f[i] = ex2_fast(ex);
          }
#endif
        } else {  // map border: tables NR (1), NB (2), C (3), gainmapmath.cpp:710-716
          const int t = dx ? 2 : (last ? 3 : 1);
          const float4* wt = reinterpret_cast<const float4*>(p.idw) + t * 16 + qr * 8;
#pragma unroll
          for (int i = 0; i < 8; i++) {
            const float4 w = __ldg(wt + i);
            float gain = e1 * w.x;
            gain = fmaf(e2, w.y, gain);
            gain = fmaf(e3, w.z, gain);
            gain = fmaf(e4, w.w, gain);
            f[i] = ex2_fast(fmaf(gain, p.Aq, p.Bq));
          }
        }
        quad_row_eval<PACK, MASK, LTAB, NORM>(hiP, yw[2 * qr], yw[2 * qr + 1], uw[qr], vw[qr], f, orow + (size_t)(2 * qr) * ostride, ostride, plane_bytes);
      }
      e1 = e2;
      e3 = e4;
      orow += 4 * ostride;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Any integer map scale factor (sampleMap with a size_t scale, gainmapmath.cpp:686-720) on the same q-domain tables:
// thread = one 4x4 PIXEL block; its 16 gain factors come from the general Shepard-IDW sampling (map samples and the
// weight row of tables mWeights / NR / NB / C fetched per map cell the block touches), everything downstream
// (quad_row_eval) is shared with the scale-4 kernel. x / scale uses a host-computed reciprocal (exact for x < 2^16).
// SCALE = 1 / 2 are compile-time specialisations (no interpolation at scale 1; one 3x3 map patch per block and the four
// interior weight rows from the constant bank at scale 2, same operation order as the general code, so the results are
// identical); SCALE = 0 is the general run-time scale.
template <int PACK, bool MASK, bool BATCH, bool LTAB, bool NORM, int SCALE>
__global__ void __launch_bounds__(kThreads, 3) apply_qs_kernel(const __grid_constant__ ApplyFastParams p) {
  const int tid = threadIdx.x;
  const uint32_t hiP = stage_q_table<PACK, LTAB>(p, tid);
  const Geometry& g = p.g;
  const uint32_t mw = g.map_width, mh = g.map_height, sc = (uint32_t)g.map_scale, magic = p.scale_magic;
  const uint32_t bw = (uint32_t)g.width >> 2, bh = (uint32_t)g.height >> 2;  // 4x4 pixel blocks
  const uint32_t tiles_x = p.tiles_x, tiles_per_image = p.tiles_x * p.tiles_y;  // tiles of 32 x 8 blocks
  const uint32_t ls = g.sdr_luma_stride, cs = g.sdr_chroma_stride;
  constexpr int kBpp = PACK == kPack1010102 ? 4 : (PACK == kPackF16 ? 8 : 2);  // bytes per pixel (per plane for planar)
  const size_t ostride = (size_t)g.width * kBpp;
  const size_t plane_bytes = (size_t)g.width * g.height * 2;
  const float4* idw = reinterpret_cast<const float4*>(p.idw);  // [table][oy][ox] rows of 4 weights
  const uint32_t ss = sc * sc;
  auto div_s = [&](uint32_t x) { return magic ? __umulhi(x, magic) : x; };  // magic == 0 <=> scale 1

  // tile -> (image z, tile row ty, tile column tx), advanced incrementally: no division in the loop
  const uint32_t tiles_y = p.tiles_y;
  uint32_t tx, ty, zn, dtx, dty, dz;
  {
    const uint32_t b0 = blockIdx.x, st = gridDim.x;
    zn = b0 / tiles_per_image;
    uint32_t rem = b0 - zn * tiles_per_image;
    ty = rem / tiles_x;
    tx = rem - ty * tiles_x;
    dz = st / tiles_per_image;
    rem = st - dz * tiles_per_image;
    dty = rem / tiles_x;
    dtx = rem - dty * tiles_x;
  }
  for (uint32_t tile = blockIdx.x; tile < (uint32_t)p.n_tiles; tile += gridDim.x) {
    const uint32_t bx = tx * 32 + (tid & 31), by = ty * 8 + (tid >> 5), z = zn;
    tx += dtx;
    if (tx >= tiles_x) { tx -= tiles_x; ty++; }
    ty += dty;
    if (ty >= tiles_y) { ty -= tiles_y; zn++; }
    zn += dz;
    if (bx >= bw || by >= bh) continue;
    const uint8_t *sdr_y, *sdr_u, *sdr_v, *map;
    uint8_t* outp;
    if (BATCH) {
      const Planes* pl = p.batch + z;
      sdr_y = (const uint8_t*)__ldg((const unsigned long long*)&pl->sdr_y);
      sdr_u = (const uint8_t*)__ldg((const unsigned long long*)&pl->sdr_u);
      sdr_v = (const uint8_t*)__ldg((const unsigned long long*)&pl->sdr_v);
      map = (const uint8_t*)__ldg((const unsigned long long*)&pl->map_in);
      outp = (uint8_t*)__ldg((const unsigned long long*)&pl->out);
    } else {
      sdr_y = (const uint8_t*)p.single.sdr_y;
      sdr_u = (const uint8_t*)p.single.sdr_u;
      sdr_v = (const uint8_t*)p.single.sdr_v;
      map = (const uint8_t*)p.single.map_in;
      outp = (uint8_t*)p.single.out;
    }
    const uint8_t* yp = sdr_y + (4u * by * ls + 4u * bx);
    const uint32_t coff = 2u * by * cs + 2u * bx;
    uint8_t* orow = outp + (size_t)(4u * by * (uint32_t)g.width + 4u * bx) * kBpp;
    uint32_t yw[4], uw[2], vw[2];
#pragma unroll
    for (int r = 0; r < 4; r++) yw[r] = ldg4(yp + r * ls);
#pragma unroll
    for (int r = 0; r < 2; r++) {
      uw[r] = ldg2(sdr_u + coff + r * cs);
      vw[r] = ldg2(sdr_v + coff + r * cs);
    }
    const uint32_t x0 = 4u * bx, xl0 = div_s(x0), ox0 = x0 - xl0 * sc;
    // scale 2: the 3x3 map samples the block's four cells use (as 0..255 floats), interior blocks only
    float P2[3][3];
    bool patch = false;
    if (SCALE == 2) {
      patch = 2u * bx + 2u < mw && 2u * by + 2u < mh;  // no cell of the block on the right / bottom map border
      if (patch) {
#pragma unroll
        for (int a = 0; a < 3; a++) {
          const uint8_t* mr = map + (2u * by + a) * mw + 2u * bx;
          const uint32_t two = ldg2(mr);
          P2[a][0] = (float)(two & 0xff);
          P2[a][1] = (float)(two >> 8);
          P2[a][2] = (float)__ldg(mr + 2);
        }
      }
    }
#pragma unroll
    for (int qr = 0; qr < 2; qr++) {
      float f[8];
      if (SCALE == 1) {  // offsets (0,0): weights [1,0,0,0] in all four tables, the gain is the pixel's own map sample
#pragma unroll
        for (int rr = 0; rr < 2; rr++) {
          const uint32_t mwrd = ldg4(map + (4u * by + 2u * qr + rr) * mw + 4u * bx);
#pragma unroll
          for (int i = 0; i < 4; i++) f[rr * 4 + i] = ex2_fast(fmaf((float)((mwrd >> (8 * i)) & 0xff), p.Aq, p.Bq));
        }
      } else if (SCALE == 2 && patch) {  // cell row qr of the block; pixel (rr, i): cell column i >> 1, offsets (rr, i & 1)
#pragma unroll
        for (int rr = 0; rr < 2; rr++)
#pragma unroll
          for (int i = 0; i < 4; i++) {
            const int cxi = i >> 1, ox = i & 1;
            const float e1 = P2[qr][cxi], e2 = P2[qr + 1][cxi], e3 = P2[qr][cxi + 1], e4 = P2[qr + 1][cxi + 1];
            float gain = e1;  // offsets (0,0): e1 * 1 + e2 * 0 + e3 * 0 + e4 * 0
            if (rr | ox) {
              const float* w = p.w2[rr * 2 + ox];
              gain = e1 * w[0];
              gain = fmaf(e2, w[1], gain);
              gain = fmaf(e3, w[2], gain);
              gain = fmaf(e4, w[3], gain);
            }
            f[rr * 4 + i] = ex2_fast(fmaf(gain, p.Aq, p.Bq));
          }
      } else {
#pragma unroll
        for (int rr = 0; rr < 2; rr++) {
          const uint32_t y = 4u * by + 2u * qr + rr, yl = div_s(y), oy = y - yl * sc;
          const bool ylast = yl + 1 >= mh;  // y_upper = min(y_lower + 1, height - 1)
          const uint8_t* m0 = map + yl * mw;
          const uint8_t* m1 = ylast ? m0 : m0 + mw;
          const float4* wrow = idw + oy * sc;
          uint32_t xl = xl0, ox = ox0, t = 0;
          float e1 = 0.f, e2 = 0.f, e3 = 0.f, e4 = 0.f;
#pragma unroll
          for (int i = 0; i < 4; i++) {
            if (i == 0 || ox == 0) {  // first pixel of the row segment or of a new map cell
              const bool xlast = xl + 1 >= mw;
              const uint32_t xu = xlast ? xl : xl + 1;
              e1 = (float)__ldg(m0 + xl);
              e2 = (float)__ldg(m1 + xl);
              e3 = (float)__ldg(m0 + xu);
              e4 = (float)__ldg(m1 + xu);
              t = xlast ? (ylast ? 3u : 1u) : (ylast ? 2u : 0u);  // C / NR / NB / interior, gainmapmath.cpp:710-716
            }
            const float4 w = __ldg(wrow + t * ss + ox);
            float gain = e1 * w.x;
            gain = fmaf(e2, w.y, gain);
            gain = fmaf(e3, w.z, gain);
            gain = fmaf(e4, w.w, gain);
            f[rr * 4 + i] = ex2_fast(fmaf(gain, p.Aq, p.Bq));
            if (++ox == sc) {
              ox = 0;
              xl++;
            }
          }
        }
      }
      quad_row_eval<PACK, MASK, LTAB, NORM>(hiP, yw[2 * qr], yw[2 * qr + 1], uw[qr], vw[qr], f, orow + (size_t)(2 * qr) * ostride, ostride, plane_bytes);
    }
  }
}

bool aligned(const void* q, size_t a) { return (reinterpret_cast<uintptr_t>(q) & (a - 1)) == 0; }

template <int PACK, bool MASK>
cudaError_t launch(const ApplyFastParams& p, cudaStream_t s) {
  const size_t smem = PACK == kPack1010102 ? 0 : kSEntriesHi * sizeof(float2);
  static int sms_of[64], per_sm_of[64];  // per device, filled on first use (benign race: same values)
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  dev &= 63;
  if (per_sm_of[dev] == 0) {
    int n_sm = 0, n = 0;
    if ((e = cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
    if (smem && (e = cudaFuncSetAttribute(apply_fast_kernel<PACK, MASK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, apply_fast_kernel<PACK, MASK>, kThreads, smem)) != cudaSuccess) return e;
    sms_of[dev] = n_sm;
    per_sm_of[dev] = n > 0 ? n : 1;
  }
  const int sms = sms_of[dev], per_sm = per_sm_of[dev];
  const int grid = p.n_tiles < sms * per_sm ? p.n_tiles : sms * per_sm;
  apply_fast_kernel<PACK, MASK><<<grid, kThreads, smem, s>>>(p);
  return cudaGetLastError();
}

// Cells per lane and tile: the persistent grid runs ceil(tiles / cap) rounds of `cells` cell-times each. Large launches
// take 8 (measured best); small ones the height with the cheapest rounds x (cells + per-tile overhead), ties to the taller.
int pick_cells(int tiles_x, int map_height, int n_img, int cap) {
  const int warps = kQThreads / 32;
  if ((long long)tiles_x * ((map_height + warps * 8 - 1) / (warps * 8)) * n_img >= 8LL * cap) return 8;
  int best = 1;
  double best_cost = 1e30;
  for (int c = 1; c <= 8; c++) {
    const long long tiles = (long long)tiles_x * ((map_height + warps * c - 1) / (warps * c)) * n_img;
    const long long rounds = (tiles + cap - 1) / cap;
    const double cost = (double)rounds * (c + 0.25);
    if (cost <= best_cost) {
      best_cost = cost;
      best = c;
    }
  }
  return best;
}

template <int PACK, bool MASK, bool BATCH, bool LTAB, bool NORM>
cudaError_t launch_q(const ApplyFastParams& p0, cudaStream_t s) {
  static int sms_of[64], per_sm_of[64], smem_of[64];
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  dev &= 63;
  if (per_sm_of[dev] == 0) {
    int n_sm = 0, n = 0, reserved = 0;
    cudaFuncAttributes fa;
    if ((e = cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
    if ((e = cudaDeviceGetAttribute(&reserved, cudaDevAttrReservedSharedMemoryPerBlock, dev)) != cudaSuccess) return e;
    if ((e = cudaFuncGetAttributes(&fa, apply_q_kernel<PACK, MASK, BATCH, LTAB, NORM>)) != cudaSuccess) return e;
    const uint32_t dyn0 = (uint32_t)reserved + (uint32_t)fa.sharedSizeBytes;
    const int smem = LTAB ? 16 + (int)(kLEntries * sizeof(float2))
                          : (PACK == kPack1010102 ? (int)(q_table_addr(dyn0) - dyn0 + kPEntries * sizeof(float2)) + 64 : 0);  // + slack for the base alignment
    if ((e = cudaFuncSetAttribute(apply_q_kernel<PACK, MASK, BATCH, LTAB, NORM>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)) != cudaSuccess) return e;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, apply_q_kernel<PACK, MASK, BATCH, LTAB, NORM>, kQThreads, smem)) != cudaSuccess) return e;
    sms_of[dev] = n_sm;
    smem_of[dev] = smem;
    per_sm_of[dev] = n > 0 ? n : 1;
  }
  ApplyFastParams p = p0;
  const int n_img = p.n_tiles / (p.tiles_x * p.tiles_y);
  const int cap = sms_of[dev] * per_sm_of[dev];
  p.cells = pick_cells(p.tiles_x, p.g.map_height, n_img, cap);
  const int kQTileRows = (kQThreads / 32) * p.cells;
  p.tiles_y = (p.g.map_height + kQTileRows - 1) / kQTileRows;  // this kernel's tiles are kQTileRows cell rows tall
  p.n_tiles = p.tiles_x * p.tiles_y * n_img;
  const int grid = p.n_tiles < cap ? p.n_tiles : cap;
  apply_q_kernel<PACK, MASK, BATCH, LTAB, NORM><<<grid, kQThreads, (size_t)smem_of[dev], s>>>(p);
  return cudaGetLastError();
}

template <int PACK, bool MASK>
cudaError_t launch_q(const ApplyFastParams& p, cudaStream_t s) {
  constexpr bool kLin = PACK != kPack1010102;
  if (kLin && p.lin_table_ok) {  // lin_table_ok also implies normal halves (capi.cu)
    constexpr bool kNorm = PACK == kPackF16;
    return p.batch ? launch_q<PACK, MASK, true, kLin, kNorm>(p, s) : launch_q<PACK, MASK, false, kLin, kNorm>(p, s);
  }
  return p.batch ? launch_q<PACK, MASK, true, false, false>(p, s) : launch_q<PACK, MASK, false, false, false>(p, s);
}

template <int PACK, bool MASK, bool BATCH, bool LTAB, bool NORM, int SCALE>
cudaError_t launch_qs(const ApplyFastParams& p0, cudaStream_t s) {
  static int sms_of[64], per_sm_of[64], smem_of[64];
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  dev &= 63;
  if (per_sm_of[dev] == 0) {
    int n_sm = 0, n = 0, reserved = 0;
    cudaFuncAttributes fa;
    if ((e = cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
    if ((e = cudaDeviceGetAttribute(&reserved, cudaDevAttrReservedSharedMemoryPerBlock, dev)) != cudaSuccess) return e;
    if ((e = cudaFuncGetAttributes(&fa, apply_qs_kernel<PACK, MASK, BATCH, LTAB, NORM, SCALE>)) != cudaSuccess) return e;
    const uint32_t dyn0 = (uint32_t)reserved + (uint32_t)fa.sharedSizeBytes;
    const int smem = LTAB ? 16 + (int)(kLEntries * sizeof(float2))
                          : (PACK == kPack1010102 ? (int)(q_table_addr(dyn0) - dyn0 + kPEntries * sizeof(float2)) + 64 : 0);
    if ((e = cudaFuncSetAttribute(apply_qs_kernel<PACK, MASK, BATCH, LTAB, NORM, SCALE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)) != cudaSuccess) return e;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, apply_qs_kernel<PACK, MASK, BATCH, LTAB, NORM, SCALE>, kThreads, smem)) != cudaSuccess) return e;
    sms_of[dev] = n_sm;
    smem_of[dev] = smem;
    per_sm_of[dev] = n > 0 ? n : 1;
  }
  const int cap = sms_of[dev] * per_sm_of[dev];
  const int grid = p0.n_tiles < cap ? p0.n_tiles : cap;
  apply_qs_kernel<PACK, MASK, BATCH, LTAB, NORM, SCALE><<<grid, kThreads, (size_t)smem_of[dev], s>>>(p0);
  return cudaGetLastError();
}

template <int PACK, bool MASK, int SCALE>
cudaError_t launch_qs(const ApplyFastParams& p, cudaStream_t s) {
  constexpr bool kLin = PACK != kPack1010102;
  if (kLin && p.lin_table_ok) {
    constexpr bool kNorm = PACK == kPackF16;
    return p.batch ? launch_qs<PACK, MASK, true, kLin, kNorm, SCALE>(p, s) : launch_qs<PACK, MASK, false, kLin, kNorm, SCALE>(p, s);
  }
  return p.batch ? launch_qs<PACK, MASK, true, false, false, SCALE>(p, s) : launch_qs<PACK, MASK, false, false, false, SCALE>(p, s);
}

template <int PACK, bool MASK>
cudaError_t launch_qs(const ApplyFastParams& p, cudaStream_t s) {
  switch (p.scale_special) {  // 1 / 2 when the map planes are aligned for the specialised loads (capi.cu)
    case 1: return launch_qs<PACK, MASK, 1>(p, s);
    case 2: return launch_qs<PACK, MASK, 2>(p, s);
    default: return launch_qs<PACK, MASK, 0>(p, s);
  }
}

}  // namespace

bool apply_fast_ok(const Geometry& g, const Planes& pl) {
  return g.map_scale == 4 && g.width % 4 == 0 && g.height % 4 == 0 && g.map_width * 4 == g.width &&
         g.map_height * 4 == g.height && aligned(pl.sdr_y, 4) && aligned(pl.sdr_u, 2) && aligned(pl.sdr_v, 2) &&
         aligned(pl.out, 16) && g.sdr_luma_stride % 4 == 0 && g.sdr_chroma_stride % 2 == 0 &&
         ((size_t)g.width * g.height * 2) % 8 == 0 && (uint64_t)g.height * (uint64_t)g.sdr_luma_stride < (1ull << 31) &&
         (uint64_t)g.height * (uint64_t)g.width < (1ull << 31);
}

// the any-scale q-domain kernel: 4-aligned planes, map = image / scale exactly, coordinates below 2^16 (reciprocal division)
bool apply_qs_ok(const Geometry& g, const Planes& pl) {
  return g.map_scale >= 1 && g.map_scale < 65536 && g.width % 4 == 0 && g.height % 4 == 0 && g.width < 65536 && g.height < 65536 &&
         g.map_width * g.map_scale == g.width && g.map_height * g.map_scale == g.height && aligned(pl.sdr_y, 4) &&
         aligned(pl.sdr_u, 2) && aligned(pl.sdr_v, 2) && aligned(pl.out, 16) && g.sdr_luma_stride % 4 == 0 &&
         g.sdr_chroma_stride % 2 == 0 && ((size_t)g.width * g.height * 2) % 8 == 0 &&
         (uint64_t)g.height * (uint64_t)g.sdr_luma_stride < (1ull << 31) && (uint64_t)g.height * (uint64_t)g.width < (1ull << 31);
}

cudaError_t launch_apply_fast_impl(ApplyFastParams& p, int n, cudaStream_t s) {
  if (p.g.width == 0 || p.g.height == 0 || n == 0) return cudaSuccess;
  if (p.any_scale) {
    p.tiles_x = (p.g.width / 4 + 31) / 32;
    p.tiles_y = (p.g.height / 4 + 7) / 8;
    p.n_tiles = p.tiles_x * p.tiles_y * n;
    // x / scale == umulhi(x, ceil(2^32 / scale)) for x < 2^16 <= 2^32 / scale; scale 1 is the identity (magic 0)
    p.scale_magic = p.g.map_scale == 1 ? 0u : (uint32_t)((0x100000000ull + (uint64_t)p.g.map_scale - 1) / (uint64_t)p.g.map_scale);
    switch (p.output_format) {
      case 1: return launch_qs<kPackF16, false>(p, s);
      case 4: return p.need_mask ? launch_qs<kPackPlanar10, true>(p, s) : launch_qs<kPackPlanar10, false>(p, s);
      default: return p.need_mask ? launch_qs<kPack1010102, true>(p, s) : launch_qs<kPack1010102, false>(p, s);
    }
  }
  p.tiles_x = (p.g.map_width + kTileCellsX - 1) / kTileCellsX;
  p.tiles_y = (p.g.map_height + kTileCellsY - 1) / kTileCellsY;
  p.n_tiles = p.tiles_x * p.tiles_y * n;
  switch (p.output_format) {
    case 1: return p.two_table ? launch<kPackF16, false>(p, s) : launch_q<kPackF16, false>(p, s);
    case 4:
      if (p.two_table) return launch<kPackPlanar10, false>(p, s);
      return p.need_mask ? launch_q<kPackPlanar10, true>(p, s) : launch_q<kPackPlanar10, false>(p, s);
    default:
      if (!p.two_table) return p.need_mask ? launch_q<kPack1010102, true>(p, s) : launch_q<kPack1010102, false>(p, s);
      return p.need_mask ? launch<kPack1010102, true>(p, s) : launch<kPack1010102, false>(p, s);
  }
}

}  // namespace uhdr
