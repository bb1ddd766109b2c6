// apply_pipe.cu — A2: UltraHdr::applyGainMap (reference lib/src/ultrahdr.cpp:427-496), map scale factor 4, as a
// TMA-fed producer/consumer pipeline per WARP. Same arithmetic as apply_q_kernel (apply_fast.cu; the q-domain
// evaluation lives in apply_q_common.cuh), different data movement:
//
//   * a warp owns a strip of 32 map cells (128 pixels) and walks `cells` cell rows down it; strips are handed out
//     round-robin over all resident warps of a persistent grid (unit u -> warp u mod #warps), and a warp's sequence
//     of steps runs on across its units, so the pipeline below never drains;
//   * per step (one cell row = 4 pixel rows) ten lanes of the warp each issue ONE 1-D bulk copy (cp.async.bulk,
//     global -> shared, completion on an mbarrier): 4 luma rows of 128 B, 2 + 2 chroma rows of 64 B, the lower map
//     row (32 samples + 16 of halo) and, for the first step of a strip, the upper map row. kStages steps are in
//     flight per warp; the consumer side of the same warp waits on the stage's mbarrier, reads its samples with
//     conflict-free LDS and refills the stage when it is done with it. No block-level synchronisation after start-up;
//   * compared with apply_q_kernel this removes the per-lane global address arithmetic (10 loads per cell, each with
//     a 64-bit address) and the 11 software-prefetch registers, and hides the load latency kStages steps deep
//     instead of one.
//
// Layout requirements (apply_pipe_ok): map width a multiple of 16 cells (image width % 64 == 0), luma / chroma strides
// and all plane pointers 16-byte aligned (the bulk copy's granularity). Everything else takes apply_q_kernel.
#include <cstring>
#include <mutex>

#include "apply_q_common.cuh"

namespace uhdr {
namespace {

#ifndef UHDR_PIPE_STAGES
#define UHDR_PIPE_STAGES 4  // power of two
#endif
#ifndef UHDR_PIPE_MIN_BLOCKS
#define UHDR_PIPE_MIN_BLOCKS 3
#endif
#ifndef UHDR_PIPE_L_MIN_BLOCKS
#define UHDR_PIPE_L_MIN_BLOCKS 3
#endif
constexpr int kPipeThreads = 256, kPipeWarps = kPipeThreads / 32, kStages = UHDR_PIPE_STAGES;
static_assert((kStages & (kStages - 1)) == 0, "stage count must be a power of two");
// one stage: 4 luma rows x 128 B | 2 U rows x 64 B | 2 V rows x 64 B | lower map row (48 B, padded) | upper map row
constexpr uint32_t kOffU = 512, kOffV = 640, kOffMapLo = 768, kOffMapUp = 832, kStageBytes = 896;

__device__ __forceinline__ uint32_t lds_u32(uint32_t a) {
  uint32_t r;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(r) : "r"(a));
  return r;
}
__device__ __forceinline__ uint32_t lds_u16(uint32_t a) {
  uint16_t r;
  asm volatile("ld.shared.u16 %0, [%1];" : "=h"(r) : "r"(a));
  return r;
}
__device__ __forceinline__ uint32_t lds_u8(uint32_t a) {
  uint32_t r;
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(r) : "r"(a));
  return r;
}
// mbarrier / bulk-copy helpers on shared ADDRESSES (uhdr_io.cuh has the pointer forms)
__device__ __forceinline__ void mbar_init_a(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx_a(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait_a(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s_a(uint32_t dst, const void* gsrc, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(gsrc),
               "r"(bytes), "r"(bar)
               : "memory");
}

// Shared-memory layout behind the table of stage_q_table (absolute shared addresses; dyn0 = start of the dynamic
// region): [table ...][kPipeWarps * kStages mbarriers][128-byte aligned: kPipeWarps * kStages stages]
template <int PACK, bool LTAB>
__host__ __device__ constexpr uint32_t pipe_bars_addr(uint32_t dyn0) {
  const uint32_t tab_end = LTAB ? dyn0 + 16u + (uint32_t)(kLEntries * sizeof(float2))
                                : (PACK == kPack1010102 ? q_table_addr(dyn0) + (uint32_t)(kPEntries * sizeof(float2)) : dyn0 + 16u);
  return (tab_end + 15u) & ~15u;
}
template <int PACK, bool LTAB>
__host__ __device__ constexpr uint32_t pipe_stages_addr(uint32_t dyn0) {
  return (pipe_bars_addr<PACK, LTAB>(dyn0) + (uint32_t)(kPipeWarps * kStages * 8) + 127u) & ~127u;
}
template <int PACK, bool LTAB>
__host__ __device__ constexpr uint32_t pipe_smem_bytes(uint32_t dyn0) {
  return pipe_stages_addr<PACK, LTAB>(dyn0) + (uint32_t)(kPipeWarps * kStages) * kStageBytes - dyn0;
}

// unit -> (image z, row block ty, strip tx), advanced incrementally by the grid's warp count (no division per unit)
struct Walk {
  uint32_t u, tx, ty, z;
};

template <int PACK, bool MASK, bool BATCH, bool LTAB, bool NORM>
__global__ void __launch_bounds__(kPipeThreads, LTAB ? UHDR_PIPE_L_MIN_BLOCKS : UHDR_PIPE_MIN_BLOCKS) apply_pipe_kernel(const __grid_constant__ ApplyFastParams p) {
  // programmatic dependent launch: the next kernel of the stream may start its prologue now; its own griddepcontrol.wait
  // still orders its loads and stores behind this whole grid
  pdl_trigger();
  extern __shared__ __align__(16) uint8_t smem_dyn[];
  const int tid = threadIdx.x;
  const uint32_t lane = tid & 31, wid = tid >> 5;
  const uint32_t dyn0 = smem_u32(smem_dyn);
  {
    uint32_t dyn_size;
    asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(dyn_size));
    if (pipe_smem_bytes<PACK, LTAB>(dyn0) > dyn_size) __trap();
  }
  const uint32_t bars = pipe_bars_addr<PACK, LTAB>(dyn0) + wid * (uint32_t)(kStages * 8);
  const uint32_t stages0 = pipe_stages_addr<PACK, LTAB>(dyn0) + wid * (uint32_t)kStages * kStageBytes;
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < kStages; s++) mbar_init_a(bars + 8 * s, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  const uint32_t hiP = stage_q_table<PACK, LTAB>(p, tid);
  // everything above touches only constants; the planes below may be the previous kernel's output
  pdl_wait();

  const Geometry& g = p.g;
  const uint32_t mw = g.map_width, mh = g.map_height, cells = (uint32_t)p.cells;
  const uint32_t tiles_x = p.tiles_x, rows_y = p.tiles_y, n_units = (uint32_t)p.n_tiles;
  const uint32_t ls = g.sdr_luma_stride, cs = g.sdr_chroma_stride;
  constexpr int kBpp = PACK == kPack1010102 ? 4 : (PACK == kPackF16 ? 8 : 2);  // bytes per pixel (per plane for planar)
  const size_t ostride = (size_t)g.width * kBpp;
  const size_t plane_bytes = (size_t)g.width * g.height * 2;

  // ---- unit walkers: the consumer's (cw) and the producer's (pw), kStages steps ahead, over the same sequence
  uint32_t dtx, dty, dz;
  Walk cw;
  {
    const uint32_t upi = tiles_x * rows_y, st = gridDim.x * (uint32_t)kPipeWarps;
    cw.u = blockIdx.x * (uint32_t)kPipeWarps + wid;
    cw.z = cw.u / upi;
    uint32_t rem = cw.u - cw.z * upi;
    cw.ty = rem / tiles_x;
    cw.tx = rem - cw.ty * tiles_x;
    dz = st / upi;
    rem = st - dz * upi;
    dty = rem / tiles_x;
    dtx = rem - dty * tiles_x;
  }
  Walk pw = cw;
  auto advance = [&](Walk& w) {
    w.u += gridDim.x * (uint32_t)kPipeWarps;
    w.tx += dtx;
    if (w.tx >= tiles_x) { w.tx -= tiles_x; w.ty++; }
    w.ty += dty;
    if (w.ty >= rows_y) { w.ty -= rows_y; w.z++; }
    w.z += dz;
  };

  // ---- producer roles: lanes 0-3 luma rows, 4-5 U rows, 6-7 V rows, 8 lower map row, 9 upper map row (first step only)
  const uint32_t role = lane < 4 ? 0u : (lane < 6 ? 1u : (lane < 8 ? 2u : (lane == 8 ? 3u : (lane == 9 ? 4u : 5u))));
  const uint32_t r_in = role == 0 ? lane : (role == 1 ? lane - 4 : (role == 2 ? lane - 6 : 0u));
  const uint32_t row_mul = role == 0 ? 4u : (role <= 2 ? 2u : 1u);
  const uint32_t row_stride = role == 0 ? ls : (role <= 2 ? cs : mw);  // bytes per row of my plane
  const uint32_t cell_bytes = role == 0 ? 4u : (role <= 2 ? 2u : 1u);  // bytes per map cell in one row of my plane
  const uint32_t dst_off = role == 0 ? r_in * 128u : (role == 1 ? kOffU + r_in * 64u : (role == 2 ? kOffV + r_in * 64u : (role == 3 ? kOffMapLo : kOffMapUp)));
  const int field = role == 0 ? 0 : (role == 1 ? 1 : (role == 2 ? 2 : 5));  // index of my plane pointer inside Planes
  const uint8_t* pbase = nullptr;  // my plane + the strip's x offset
  uint32_t psize = 0, ptotal = 0, pmsz = 0, pk = 0, pkmax = 0, pcy = 0;
  bool pvalid = pw.u < n_units;
  auto p_begin_unit = [&]() {
    pcy = pw.ty * cells;
    pkmax = min(cells, mh - pcy);
    pk = 0;
    const uint32_t cx0 = 32u * pw.tx;
    const uint32_t wc = min(32u, mw - cx0);   // 16 or 32 cells (mw % 16 == 0)
    pmsz = wc + (cx0 + 32u < mw ? 16u : 0u);   // + the right-hand halo of the map row
    ptotal = 24u * wc + pmsz;                  // 4 x 4wc luma + 4 x 2wc chroma + lower map row
    if (role < 5) {
      const unsigned long long* pl = reinterpret_cast<const unsigned long long*>(BATCH ? (p.batch + pw.z) : &p.single);
      const uint8_t* plane = reinterpret_cast<const uint8_t*>(BATCH ? __ldg(pl + field) : pl[field]);
      pbase = plane + (size_t)cx0 * cell_bytes;
      psize = role >= 3 ? pmsz : wc * cell_bytes;
    }
  };
  // issues the copies of the producer's current step into stage n % kStages and advances the producer
  auto produce = [&](uint32_t n) {
    const uint32_t s = n & (uint32_t)(kStages - 1);
    const uint32_t bar = bars + 8u * s;
    if (lane == 0) mbar_expect_tx_a(bar, ptotal + (pk == 0 ? pmsz : 0u));
    __syncwarp();
    if (role < 4 || (role == 4 && pk == 0)) {
      uint32_t row = row_mul * pcy + r_in;
      if (role == 3) row = min(pcy + 1u, mh - 1u);  // yu = min(yl + 1, mh - 1), gainmapmath.cpp:690-697
      bulk_g2s_a(stages0 + s * kStageBytes + dst_off, pbase + (size_t)row * row_stride, psize, bar);
    }
    pk++;
    pcy++;
    if (pk == pkmax) {
      advance(pw);
      pvalid = pw.u < n_units;
      if (pvalid) p_begin_unit();
    }
  };
  if (pvalid) p_begin_unit();
#pragma unroll 1
  for (uint32_t i = 0; i < (uint32_t)kStages && pvalid; i++) produce(i);

  uint32_t n = 0;  // consumer step counter: stage n % kStages, phase (n / kStages) & 1
#pragma unroll 1
  while (cw.u < n_units) {
    uint32_t cy = cw.ty * cells;
    const uint32_t kmax = min(cells, mh - cy);
    const uint32_t cx = 32u * cw.tx + lane;
    const bool active = cx < mw;
    const unsigned amask = __ballot_sync(0xffffffffu, active);
    const uint32_t dx = cx + 1 < mw ? 1u : 0u;  // xu = min(xl + 1, mw - 1), gainmapmath.cpp:690-697
    uint8_t* orow;
    {
      uint8_t* outp = BATCH ? (uint8_t*)__ldg((const unsigned long long*)&(p.batch + cw.z)->out) : (uint8_t*)p.single.out;
      orow = outp + (size_t)(4u * cy * (uint32_t)g.width + 4u * cx) * kBpp;
    }
    float e1 = 0.f, e3 = 0.f;
#pragma unroll 1
    for (uint32_t k = 0; k < kmax; k++, cy++, n++) {
      const uint32_t s = n & (uint32_t)(kStages - 1);
      mbar_wait_a(bars + 8u * s, (n / (uint32_t)kStages) & 1u);
      const uint32_t st = stages0 + s * kStageBytes;
      if (active) {
        uint32_t yw[4], uw[2], vw[2];
#pragma unroll
        for (int r = 0; r < 4; r++) yw[r] = lds_u32(st + r * 128 + lane * 4);
#pragma unroll
        for (int r = 0; r < 2; r++) {
          uw[r] = lds_u16(st + kOffU + r * 64 + lane * 2);
          vw[r] = lds_u16(st + kOffV + r * 64 + lane * 2);
        }
        // map samples as 0..255 floats: the 1/255 of gainmapmath.cpp:700-703 is folded into the weights
        const float e2 = (float)lds_u8(st + kOffMapLo + lane), e4 = (float)lds_u8(st + kOffMapLo + lane + dx);  // (xl,yu) (xu,yu)
        if (k == 0) {
          e1 = (float)lds_u8(st + kOffMapUp + lane);       // (xl,yl)
          e3 = (float)lds_u8(st + kOffMapUp + lane + dx);  // (xu,yl)
        }
        const bool last = cy + 1 >= mh;
        const float d2 = e2 - e1, d3 = e3 - e1, d4 = e4 - e1, exb = fmaf(e1, p.Aq, p.Bq);
        const bool interior = dx && !last;
      const bool satcell = e1 == 255.0f && d2 == 0.0f && d3 == 0.0f && d4 == 0.0f;  // ApplyFastParams::vsat
      const int sat_t = interior ? 0 : (dx ? 2 : (last ? 3 : 1));                   // which IDW table the cell uses
#pragma unroll
        for (int qr = 0; qr < 2; qr++) {
          // ---- f' = f^(1/2.4) of the 8 pixels of this quad-row (see apply_q_kernel)
          float f[8];
          if (interior) {
            const f32x2 D2 = pk2(d2, d2), D3 = pk2(d3, d3), D4 = pk2(d4, d4), EXB = pk2(exb, exb);
#pragma unroll
            for (int j = 0; j < 4; j++) {
              const int i = qr * 8 + 2 * j;
              f32x2 ex = fma2(D2, *reinterpret_cast<const f32x2*>(&p.wq[0][i]), EXB);
              ex = fma2(D3, *reinterpret_cast<const f32x2*>(&p.wq[1][i]), ex);
              ex = fma2(D4, *reinterpret_cast<const f32x2*>(&p.wq[2][i]), ex);
              float a, b;
              unpk2(ex, a, b);
              f[2 * j] = ex2_fast(a);
              f[2 * j + 1] = ex2_fast(b);
            }
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
          quad_row_eval<PACK, MASK, LTAB, NORM, true>(hiP, yw[2 * qr], yw[2 * qr + 1], uw[qr], vw[qr], f, orow + (size_t)(2 * qr) * ostride, ostride,
                                                      plane_bytes, amask, satcell, p.vsat[sat_t] + qr * 8, p.rsat[sat_t] + qr * 8);
        }
        e1 = e2;
        e3 = e4;
      }
      orow += 4 * ostride;
      // every lane's samples of this stage are in registers and consumed: refill it with step n + kStages
      __syncwarp();
      if (pvalid) produce(n + (uint32_t)kStages);
    }
    advance(cw);
  }
}

struct PipeLaunchInfo {
  int sms = 0, per_sm = 0, smem = 0;
  cudaError_t err = cudaSuccess;
};

// per device and kernel instantiation, filled once under a lock (several host threads may launch concurrently)
template <int PACK, bool MASK, bool BATCH, bool LTAB, bool NORM>
const PipeLaunchInfo& pipe_info(int dev) {
  static std::mutex mu;
  static PipeLaunchInfo info[64];
  static bool have[64];
  std::lock_guard<std::mutex> lk(mu);
  PipeLaunchInfo& li = info[dev & 63];
  if (have[dev & 63]) return li;
  auto kern = apply_pipe_kernel<PACK, MASK, BATCH, LTAB, NORM>;
  int reserved = 0, n = 0;
  cudaFuncAttributes fa;
  cudaError_t e;
  li.err = cudaSuccess;
  if ((e = cudaDeviceGetAttribute(&li.sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess ||
      (e = cudaDeviceGetAttribute(&reserved, cudaDevAttrReservedSharedMemoryPerBlock, dev)) != cudaSuccess ||
      (e = cudaFuncGetAttributes(&fa, kern)) != cudaSuccess) {
    li.err = e;
    return li;
  }
  const uint32_t dyn0 = (uint32_t)reserved + (uint32_t)fa.sharedSizeBytes;
  li.smem = (int)pipe_smem_bytes<PACK, LTAB>(dyn0) + 128;  // + slack for the alignment of the dynamic region's base
  if ((e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, li.smem)) != cudaSuccess ||
      (e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, kPipeThreads, li.smem)) != cudaSuccess) {
    li.err = e;
    return li;
  }
  li.per_sm = n > 0 ? n : 1;
  have[dev & 63] = true;
  return li;
}

// Cell rows per unit: 8 for large launches (fewer upper-map-row reloads); for a launch with fewer units than a few per
// resident warp, the height that spreads the work over the warps in the fewest rounds (ties to the taller unit).
int pipe_pick_cells(int tiles_x, int map_height, int n_img, int warps) {
  if ((long long)tiles_x * ((map_height + 7) / 8) * n_img >= 8LL * warps) return 8;
  int best = 1;
  double best_cost = 1e30;
  for (int c = 1; c <= 8; c++) {
    const long long units = (long long)tiles_x * ((map_height + c - 1) / c) * n_img;
    const long long rounds = (units + warps - 1) / warps;
    const double cost = (double)rounds * (c + 0.15);
    if (cost <= best_cost) {
      best_cost = cost;
      best = c;
    }
  }
  return best;
}

template <int PACK, bool MASK, bool BATCH, bool LTAB, bool NORM>
cudaError_t launch_pipe(const ApplyFastParams& p0, int n_img, cudaStream_t s) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  const PipeLaunchInfo& li = pipe_info<PACK, MASK, BATCH, LTAB, NORM>(dev);
  if (li.err != cudaSuccess) return li.err;
  ApplyFastParams p = p0;
  const int warps = li.sms * li.per_sm * kPipeWarps;
  p.tiles_x = (p.g.map_width + 31) / 32;
  p.cells = pipe_pick_cells(p.tiles_x, p.g.map_height, n_img, warps);
  p.tiles_y = (p.g.map_height + p.cells - 1) / p.cells;
  p.n_tiles = p.tiles_x * p.tiles_y * n_img;
  const int blocks_needed = (p.n_tiles + kPipeWarps - 1) / kPipeWarps;
  const int grid = blocks_needed < li.sms * li.per_sm ? blocks_needed : li.sms * li.per_sm;
  // programmatic dependent launch: this kernel's start-up (table staging, barrier set-up) may overlap the tail of the
  // previous kernel in the stream; its griddepcontrol.wait orders everything else
  return launch_pdl(apply_pipe_kernel<PACK, MASK, BATCH, LTAB, NORM>, dim3((unsigned)grid), dim3(kPipeThreads), (size_t)li.smem, s, p0.pdl != 0, p);
}

template <int PACK, bool MASK>
cudaError_t launch_pipe(const ApplyFastParams& p, int n_img, cudaStream_t s) {
  constexpr bool kLin = PACK != kPack1010102;
  if (kLin && p.lin_table_ok) {  // lin_table_ok also implies normal halves (capi.cu)
    constexpr bool kNorm = PACK == kPackF16;
    return p.batch ? launch_pipe<PACK, MASK, true, kLin, kNorm>(p, n_img, s) : launch_pipe<PACK, MASK, false, kLin, kNorm>(p, n_img, s);
  }
  return p.batch ? launch_pipe<PACK, MASK, true, false, false>(p, n_img, s) : launch_pipe<PACK, MASK, false, false, false>(p, n_img, s);
}

}  // namespace

// the bulk copies move 16-byte granules: map width a multiple of 16 cells, 16-byte aligned planes and strides
bool apply_pipe_ok(const Geometry& g, const Planes& pl) {
  auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  return apply_fast_ok(g, pl) && g.map_width % 16 == 0 && g.sdr_luma_stride % 16 == 0 && g.sdr_chroma_stride % 16 == 0 && al16(pl.sdr_y) &&
         al16(pl.sdr_u) && al16(pl.sdr_v) && al16(pl.map_in) && al16(pl.out);
}

cudaError_t launch_apply_pipe_impl(const ApplyFastParams& p, int n, cudaStream_t s) {
  switch (p.output_format) {
    case 1: return launch_pipe<kPackF16, false>(p, n, s);
    case 4: return p.need_mask ? launch_pipe<kPackPlanar10, true>(p, n, s) : launch_pipe<kPackPlanar10, false>(p, n, s);
    default: return p.need_mask ? launch_pipe<kPack1010102, true>(p, n, s) : launch_pipe<kPack1010102, false>(p, n, s);
  }
}

}  // namespace uhdr
