// apply_fast.cu — A2 fast path: UltraHdr::applyGainMap (reference lib/src/ultrahdr.cpp:427-496) for
// the layout every decode produces (map scale factor 4 = kMapDimensionScaleFactor, 4-aligned
// planes), restructured so that it is bound by instruction issue / HBM instead of the SFU:
//
//   * one thread = one gain-map cell = 4x4 pixels: the four surrounding map samples, the 2x2
//     chroma samples and the cell's plane pointers are fetched once per 16 pixels; luma comes in
//     as four 32-bit loads (a warp reads 128 contiguous bytes per row) and the packed output
//     leaves as 128-bit stores (a warp writes 512 contiguous bytes per row).
//   * the Shepard-IDW weights of the interior table (gainmapmath.h:184-228) sit in the kernel
//     parameter block, i.e. in the constant bank: with the 4x4 loop fully unrolled every weight
//     is an immediate constant-bank operand of an FFMA. Cells on the right/bottom map border
//     (tables NR/NB/C, gainmapmath.cpp:710-716) read their weights from global memory.
//   * srgbInvOetf (3 pow per pixel in the reference) and hlgOetf/pqOetf (sqrt/log or 9 pow per
//     pixel) are piecewise-linear tables in shared memory on log-spaced segments: the float's
//     exponent/mantissa bits are the index, one 64-bit LDS + one FFMA per evaluation. The tables
//     (built in double on the host, tables.cu) are staged with one TMA bulk copy per block and the
//     blocks are persistent (grid = SMs x resident blocks, grid-stride over 128x32-pixel tiles).
//   * the gain factor 2^(logBoost * displayBoost / maxContentBoost) / displayBoost
//     (gainmapmath.cpp:550-555, ultrahdr.cpp:451) is one FFMA + one MUFU.EX2 per pixel.
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

}  // namespace

bool apply_fast_ok(const Geometry& g, const Planes& pl) {
  return g.map_scale == 4 && g.width % 4 == 0 && g.height % 4 == 0 && g.map_width * 4 == g.width &&
         g.map_height * 4 == g.height && aligned(pl.sdr_y, 4) && aligned(pl.sdr_u, 2) && aligned(pl.sdr_v, 2) &&
         aligned(pl.out, 16) && g.sdr_luma_stride % 4 == 0 && g.sdr_chroma_stride % 2 == 0 &&
         ((size_t)g.width * g.height * 2) % 8 == 0 && (uint64_t)g.height * (uint64_t)g.sdr_luma_stride < (1ull << 31) &&
         (uint64_t)g.height * (uint64_t)g.width < (1ull << 31);
}

cudaError_t launch_apply_fast_impl(ApplyFastParams& p, int n, cudaStream_t s) {
  if (p.g.width == 0 || p.g.height == 0 || n == 0) return cudaSuccess;
  p.tiles_x = (p.g.map_width + kTileCellsX - 1) / kTileCellsX;
  p.tiles_y = (p.g.map_height + kTileCellsY - 1) / kTileCellsY;
  p.n_tiles = p.tiles_x * p.tiles_y * n;
  switch (p.output_format) {
    case 1: return launch<kPackF16, false>(p, s);
    case 4: return launch<kPackPlanar10, false>(p, s);
    default: return p.need_mask ? launch<kPack1010102, true>(p, s) : launch<kPack1010102, false>(p, s);
  }
}

}  // namespace uhdr
