// apply_fast.cu — A2: UltraHdr::applyGainMap (reference lib/src/ultrahdr.cpp:427-496) on table-driven kernels for
// 4-aligned planes, restructured so that it is bound by instruction issue / HBM instead of the SFU. Three kernels:
//
//   apply_q_kernel   map scale factor 4 (kMapDimensionScaleFactor: what every libultrahdr encode writes). The kernel
//                    the benchmark measures. One thread = one gain-map cell = 4x4 pixels, a lane walks CELLS cells
//                    downwards with the next cell's loads in flight; ONE look-up per channel in the q domain (below).
//   apply_qs_kernel  any integer scale factor (specialised for 1, 2 and the multiples of 4): thread = 4x4 pixel block, general Shepard-IDW
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
#include <cstdlib>
#include <mutex>

#include "apply_q_common.cuh"

namespace uhdr {
namespace {

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
  pdl_trigger();
  const uint32_t hiP = stage_q_table<PACK, LTAB>(p, tid);  // constants only: may overlap the previous kernel's tail
  pdl_wait();

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
#ifndef UHDR_APPLYQ_VOTE
#define UHDR_APPLYQ_VOTE 1  // a quad-row with a toe channel in ANY lane sends the whole warp down the packed toe-select path
#endif
    // the lanes of a warp share cy and kmax (a warp is one row of 32 cells), so those inside the map walk in lock step
    const unsigned amask = __ballot_sync(0xffffffffu, cx < mw && cy < mh);
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
#pragma unroll kQUnroll
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
      const bool satcell = e1 == 255.0f && d2 == 0.0f && d3 == 0.0f && d4 == 0.0f;  // ApplyFastParams::vsat
      const int sat_t = interior ? 0 : (dx ? 2 : (last ? 3 : 1));                   // which IDW table the cell uses

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
        quad_row_eval<PACK, MASK, LTAB, NORM, UHDR_APPLYQ_VOTE != 0>(hiP, yw[2 * qr], yw[2 * qr + 1], uw[qr], vw[qr], f, orow + (size_t)(2 * qr) * ostride,
                                                                     ostride, plane_bytes, amask, satcell, p.vsat[sat_t] + qr * 8, p.rsat[sat_t] + qr * 8);
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
#ifndef UHDR_APPLYQS_MIN_BLOCKS
#define UHDR_APPLYQS_MIN_BLOCKS 4  // 64 registers (0-8 B of spills): scale 1 0.59 -> 0.65, scale 2 0.49 -> 0.53 of the HBM peak against 3 blocks x 80 registers
#endif
__global__ void __launch_bounds__(kThreads, UHDR_APPLYQS_MIN_BLOCKS) apply_qs_kernel(const __grid_constant__ ApplyFastParams p) {
  const int tid = threadIdx.x;
  pdl_trigger();
  const uint32_t hiP = stage_q_table<PACK, LTAB>(p, tid);  // constants only: may overlap the previous kernel's tail
  pdl_wait();
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
    float E4[4] = {0.f, 0.f, 0.f, 0.f};
    const float4* w4 = idw;
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
      } else if (SCALE == 4) {  // scale factor a multiple of 4 (8, 16, ...; 4 with unaligned map): the 4x4 block lies inside ONE
                                // map cell, so its four map samples and its weight table are loaded once (at qr == 0)
        if (qr == 0) {
          const uint32_t yl = div_s(4u * by);
          const bool xlast = xl0 + 1 >= mw, ylast = yl + 1 >= mh;
          const uint8_t* m0 = map + yl * mw;
          const uint8_t* m1 = ylast ? m0 : m0 + mw;
          const uint32_t xu = xlast ? xl0 : xl0 + 1;
          E4[0] = (float)__ldg(m0 + xl0);
          E4[1] = (float)__ldg(m1 + xl0);
          E4[2] = (float)__ldg(m0 + xu);
          E4[3] = (float)__ldg(m1 + xu);
          const uint32_t t = xlast ? (ylast ? 3u : 1u) : (ylast ? 2u : 0u);  // C / NR / NB / interior, gainmapmath.cpp:710-716
          w4 = idw + t * ss + (4u * by - yl * sc) * sc + ox0;  // weight row of the block's first pixel row
        }
#pragma unroll
        for (int rr = 0; rr < 2; rr++)
#pragma unroll
          for (int i = 0; i < 4; i++) {
            const float4 w = __ldg(w4 + (2 * qr + rr) * sc + i);
            float gain = E4[0] * w.x;
            gain = fmaf(E4[1], w.y, gain);
            gain = fmaf(E4[2], w.z, gain);
            gain = fmaf(E4[3], w.w, gain);
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
  static int sms_of[64], per_sm_of[64];  // per device, filled on first use under the lock (host threads launch concurrently)
  static std::mutex mu;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  dev &= 63;
  std::unique_lock<std::mutex> lk(mu);
  if (per_sm_of[dev] == 0) {
    int n_sm = 0, n = 0;
    if ((e = cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
    if (smem && (e = cudaFuncSetAttribute(apply_fast_kernel<PACK, MASK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, apply_fast_kernel<PACK, MASK>, kThreads, smem)) != cudaSuccess) return e;
    sms_of[dev] = n_sm;
    per_sm_of[dev] = n > 0 ? n : 1;
  }
  const int sms = sms_of[dev], per_sm = per_sm_of[dev];
  lk.unlock();
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
  static int sms_of[64], per_sm_of[64], smem_of[64];  // per device, filled on first use under the lock
  static std::mutex mu;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  dev &= 63;
  std::unique_lock<std::mutex> lk(mu);
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
  const int cap = sms_of[dev] * per_sm_of[dev], smem_bytes = smem_of[dev];
  lk.unlock();
  p.cells = pick_cells(p.tiles_x, p.g.map_height, n_img, cap);
  {
    static const int forced = [] {  // experiments: UHDRB200_Q_CELLS=1..16 overrides the tile height
      const char* e = getenv("UHDRB200_Q_CELLS");
      return e ? atoi(e) : 0;
    }();
    if (forced >= 1 && forced <= 16) p.cells = forced;
  }
  const int kQTileRows = (kQThreads / 32) * p.cells;
  p.tiles_y = (p.g.map_height + kQTileRows - 1) / kQTileRows;  // this kernel's tiles are kQTileRows cell rows tall
  p.n_tiles = p.tiles_x * p.tiles_y * n_img;
  const int grid = p.n_tiles < cap ? p.n_tiles : cap;
  return launch_pdl(apply_q_kernel<PACK, MASK, BATCH, LTAB, NORM>, dim3(grid), dim3(kQThreads), (size_t)smem_bytes, s, p.pdl != 0, p);
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
  static int sms_of[64], per_sm_of[64], smem_of[64];  // per device, filled on first use under the lock
  static std::mutex mu;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  dev &= 63;
  std::unique_lock<std::mutex> lk(mu);
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
  const int cap = sms_of[dev] * per_sm_of[dev], smem_bytes = smem_of[dev];
  lk.unlock();
  const int grid = p0.n_tiles < cap ? p0.n_tiles : cap;
  return launch_pdl(apply_qs_kernel<PACK, MASK, BATCH, LTAB, NORM, SCALE>, dim3(grid), dim3(kThreads), (size_t)smem_bytes, s, p0.pdl != 0, p0);
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
    case 4: return launch_qs<PACK, MASK, 4>(p, s);  // any multiple of 4
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
  if (p.pipe && !p.two_table) return launch_apply_pipe_impl(p, n, s);
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
