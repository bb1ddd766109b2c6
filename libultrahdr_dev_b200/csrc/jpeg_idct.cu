// jpeg_idct.cu — row N4 (first slice): the coefficient-domain back half of the reference's JPEG decode,
// JpegDecoderHelper::decompressImage with raw_data_out and dct_method = JDCT_ISLOW (lib/src/jpegdecoderhelper.cpp:
// 291-307, 380-435): dequantise + 8x8 inverse DCT + range-limit, written straight into the planar Y | U | V (or grey)
// buffer UltraHdr::applyGainMap consumes. Entropy decoding stays on the host (libjpeg's jpeg_read_coefficients() hands
// out exactly the arrays this kernel reads), so a decoded primary image and its gain map never exist in host memory.
//
// The arithmetic is the IJG "slow-but-accurate" integer IDCT (jidctint.c, jpeg_idct_islow: CONST_BITS = 13,
// PASS1_BITS = 2, the Loeffler-Ligtenberg-Moschytz factorisation with 12 multiplies per 1-D pass). libjpeg-turbo's
// source is NOT in /root/reference (fetched at configure time, CMakeLists.txt:254-256, tag 3.0.1): the algorithm is
// restated from its published description; parity is anchored on libjpeg-turbo's own output for the reference's JPEG
// fixtures, obtained through Pillow in tests/test_jpeg_idct.py (luma planes, bit for bit).
//
// Mapping: one thread per 8x8 block, a warp per 32 consecutive blocks, a persistent grid with one two-deep cp.async
// pipeline per warp (see jpeg_idct_kernel). HBM traffic: 128 B of coefficients in, 64 B of samples out per block = 3 B per
// sample.
#include <cstring>
#include <mutex>

#include "uhdr_io.cuh"
#include "uhdr_kernels.cuh"

namespace uhdr {
namespace {

constexpr int kConstBits = 13, kPass1Bits = 2;
constexpr int FIX_0_298631336 = 2446, FIX_0_390180644 = 3196, FIX_0_541196100 = 4433, FIX_0_765366865 = 6270;
constexpr int FIX_0_899976223 = 7373, FIX_1_175875602 = 9633, FIX_1_501321110 = 12299, FIX_1_847759065 = 15137;
constexpr int FIX_1_961570560 = 16069, FIX_2_053119869 = 16819, FIX_2_562915447 = 20995, FIX_3_072711026 = 25172;

// One 1-D pass on eight values (even part, odd part, butterflies); SHIFT = CONST_BITS - PASS1_BITS for the column pass,
// CONST_BITS + PASS1_BITS + 3 for the row pass. 32-bit two's-complement arithmetic, as libjpeg for 8-bit samples.
// PRE > 0 runs the same pass on 2^PRE times the values: every term is an input times a constant or an input shifted left,
// so scaling the constants and the shift is exact modulo 2^32, and the final arithmetic shift by SHIFT + PRE then returns
// bits SHIFT .. 31 - PRE of the unscaled sum SIGN-EXTENDED: with PRE = 4 and SHIFT = 18 that is the ten-bit wrap of
// libjpeg's range-limit index (sext10 below) for free.
template <int SHIFT, int PRE = 0>
__device__ __forceinline__ void idct_1d(const int (&in)[8], int (&out)[8]) {
  constexpr int K = 1 << PRE;
  int z2 = in[2], z3 = in[6];
  int z1 = (z2 + z3) * (FIX_0_541196100 * K);
  int tmp2 = z1 - z3 * (FIX_1_847759065 * K);
  int tmp3 = z1 + z2 * (FIX_0_765366865 * K);
  z2 = in[0];
  z3 = in[4];
  int tmp0 = (int)((unsigned)(z2 + z3) << (kConstBits + PRE));
  int tmp1 = (int)((unsigned)(z2 - z3) << (kConstBits + PRE));
  const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
  tmp0 = in[7];
  tmp1 = in[5];
  tmp2 = in[3];
  tmp3 = in[1];
  z1 = tmp0 + tmp3;
  z2 = tmp1 + tmp2;
  z3 = tmp0 + tmp2;
  int z4 = tmp1 + tmp3;
  const int z5 = (z3 + z4) * (FIX_1_175875602 * K);
  tmp0 *= FIX_0_298631336 * K;
  tmp1 *= FIX_2_053119869 * K;
  tmp2 *= FIX_3_072711026 * K;
  tmp3 *= FIX_1_501321110 * K;
  z1 *= -FIX_0_899976223 * K;
  z2 *= -FIX_2_562915447 * K;
  z3 *= -FIX_1_961570560 * K;
  z4 *= -FIX_0_390180644 * K;
  z3 += z5;
  z4 += z5;
  tmp0 += z1 + z3;
  tmp1 += z2 + z4;
  tmp2 += z2 + z3;
  tmp3 += z1 + z4;
  constexpr int kRound = 1 << (SHIFT - 1 + PRE);
  out[0] = (tmp10 + tmp3 + kRound) >> (SHIFT + PRE);
  out[7] = (tmp10 - tmp3 + kRound) >> (SHIFT + PRE);
  out[1] = (tmp11 + tmp2 + kRound) >> (SHIFT + PRE);
  out[6] = (tmp11 - tmp2 + kRound) >> (SHIFT + PRE);
  out[2] = (tmp12 + tmp1 + kRound) >> (SHIFT + PRE);
  out[5] = (tmp12 - tmp1 + kRound) >> (SHIFT + PRE);
  out[3] = (tmp13 + tmp0 + kRound) >> (SHIFT + PRE);
  out[4] = (tmp13 - tmp0 + kRound) >> (SHIFT + PRE);
}

// libjpeg's range_limit table (jdmaster.c prepare_range_limit_table) indexed with `x & RANGE_MASK` (1023) is x + 128
// clamped to 0..255 for -512 <= x < 512 plus a documented wrap-around for corrupt data beyond that; both together are
// clamp(s + 128, 0, 255) with s = the low ten bits of x sign-extended, which idct_1d<18, 4> returns directly.
// clamp(s + 128, 0, 255) = clamp(s, -128, 127) + 128, and + 128 on a byte is a flip of its top bit: two saturating signed
// packs (cvt.pack d, x, y, z = sat(y) | sat(x) << 8 | z << 16) and one XOR for four samples a | b << 8 | c << 16 | d << 24.
__device__ __forceinline__ uint32_t pack4_level_shift(int a, int b, int c, int d) {
  uint32_t hi, r;
  asm("cvt.pack.sat.s8.s32.b32 %0, %1, %2, 0;" : "=r"(hi) : "r"(d), "r"(c));
  asm("cvt.pack.sat.s8.s32.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(b), "r"(a), "r"(hi));
  return r ^ 0x80808080u;
}

// ---- one thread per 8x8 block -------------------------------------------------------------------------------------------
// The whole block lives in 64 registers: both passes run without a shared-memory hand-over, a shuffle or a barrier between
// them (1130 instructions per block against 1750 for eight lanes per block with a shared-memory transpose, the first
// version of this kernel: 0.44 -> 0.58 of the HBM peak; the kernel is bound by instruction issue, not by HBM). A warp's 32 blocks are 4 KB of consecutive coefficients ("tile"): the warp copies them with eight fully
// coalesced 16-byte asynchronous copies per lane (cp.async, no registers in between) into a padded shared-memory tile (144
// bytes per block: the eight lanes of a quarter-warp then read eight different 16-byte bank groups) and each lane reads
// its own block back with eight 16-byte loads. The grid is persistent and every WARP runs its own two-deep pipeline: the
// copies of its next tile are in flight while it transforms the current one, so no block waits for HBM after the first.
// Blocks are numbered in raster order over a component and the components' tiles are numbered one after the other, so a
// warp may straddle two block rows (each lane derives its own position) but never two components.
#ifndef UHDR_IDCT_BLOCK_MIN_BLOCKS
#define UHDR_IDCT_BLOCK_MIN_BLOCKS 4
#endif
constexpr int kIdctBThreads = 128, kTileStride = 144;  // bytes per block in the shared-memory tile (128 + 16 of padding)
constexpr int kTileBytes = 32 * kTileStride;            // one warp's tile

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// two signed 16-bit halves of a times two unsigned bytes of b: lo = a.h0 * b.b0 + a.h1 * b.b1, hi = a.h0 * b.b2 + a.h1 * b.b3
__device__ __forceinline__ int dp2a_lo(uint32_t a, int b) {
  int d;
  asm("dp2a.lo.s32.u32 %0, %1, %2, 0;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}
__device__ __forceinline__ int dp2a_hi(uint32_t a, int b) {
  int d;
  asm("dp2a.hi.s32.u32 %0, %1, %2, 0;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}

struct IdctTile {
  int comp;
  uint32_t first, count;  // first block of the tile inside its component, blocks in it (1..32)
};
__device__ __forceinline__ IdctTile locate_tile(const JpegIdctParams& p, const uint32_t (&tiles)[4], uint32_t t) {
  IdctTile r;
  r.comp = 0;
#pragma unroll
  for (int k = 0; k < 3; k++) {
    if (r.comp == k && t >= tiles[k]) {
      t -= tiles[k];
      r.comp = k + 1;
    }
  }
  const JpegIdctComponent& c = p.comp[r.comp];
  const uint32_t nblk = (uint32_t)c.blocks_w * (uint32_t)c.blocks_h;
  r.first = t * 32;
  r.count = nblk - r.first < 32 ? nblk - r.first : 32;
  return r;
}

template <bool ALIGNED>  // every component's coefficient array is 16-byte aligned (the host looks at the pointers)
__global__ void __launch_bounds__(kIdctBThreads, UHDR_IDCT_BLOCK_MIN_BLOCKS) jpeg_idct_kernel(const __grid_constant__ JpegIdctParams p) {
  __shared__ __align__(16) unsigned char tile[2][kIdctBThreads / 32][kTileBytes];
  __shared__ __align__(16) int qs[4][64], qp[4][32];  // steps in natural order; pairs packed for dp2a
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  uint32_t tiles[4], total = 0;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    tiles[k] = k < p.n_components ? ((uint32_t)p.comp[k].blocks_w * (uint32_t)p.comp[k].blocks_h + 31) / 32 : 0;
    total += tiles[k];
  }
  const uint32_t n_warps = gridDim.x * (kIdctBThreads / 32), t0 = blockIdx.x * (kIdctBThreads / 32) + warp;
  auto issue = [&](uint32_t t, int buf) {  // this warp's copies of tile t; always one commit so that the groups count evenly
    if (t < total) {
      const IdctTile it = locate_tile(p, tiles, t);
      const unsigned char* src = reinterpret_cast<const unsigned char*>(p.comp[it.comp].coeffs + (size_t)it.first * 64);
      unsigned char* wt = tile[buf][warp];
#pragma unroll
      for (int j = 0; j < 8; j++) {
        const uint32_t g = j * 32 + lane;  // 16-byte chunk g of the tile: block g / 8, coefficient row g % 8
        if (g >= it.count * 8) continue;
        unsigned char* dst = wt + (g >> 3) * kTileStride + (g & 7) * 16;
        if (ALIGNED) {
          cp_async16(dst, src + (size_t)g * 16);
        } else {  // a DEVICE-pointer caller's array that is only 2-byte aligned: plain loads, ordered by the __syncwarp below
          const uint16_t* s2 = reinterpret_cast<const uint16_t*>(src + (size_t)g * 16);
          uint32_t v[4];
#pragma unroll
          for (int k = 0; k < 4; k++) v[k] = ldg2(s2 + 2 * k) | ldg2(s2 + 2 * k + 1) << 16;
          *reinterpret_cast<uint4*>(dst) = make_uint4(v[0], v[1], v[2], v[3]);
        }
      }
    }
    cp_async_commit();
  };
  issue(t0, 0);
  bool big = false;  // a 16-bit quantisation table (steps above 255): plain multiplies instead of dp2a
  for (uint32_t i = tid; i < (uint32_t)p.n_components * 64; i += kIdctBThreads) {
    const uint32_t v = __ldg(p.comp[i >> 6].quant + (i & 63));
    qs[i >> 6][i & 63] = (int)v;
    big |= v > 255;
  }
  for (uint32_t i = tid; i < (uint32_t)p.n_components * 32; i += kIdctBThreads) {
    const uint16_t* q2 = p.comp[i >> 5].quant + 2 * (i & 31);
    qp[i >> 5][i & 31] = (int)(((uint32_t)__ldg(q2) & 0xff) | ((uint32_t)__ldg(q2 + 1) & 0xff) << 24);  // bytes 0 and 3: see dp2a_lo / dp2a_hi
  }
  const bool wide_q = __syncthreads_or(big);
  int buf = 0;
  for (uint32_t t = t0; t < total; t += n_warps, buf ^= 1) {
    issue(t + n_warps, buf ^ 1);
    cp_async_wait<1>();
    __syncwarp();
    const IdctTile it = locate_tile(p, tiles, t);
    const JpegIdctComponent& c = p.comp[it.comp];
    const unsigned char* wt = tile[buf][warp] + lane * kTileStride;
    int w[64];  // lanes beyond it.count transform stale tile bytes and store nothing
    if (!wide_q) {  // DEQUANTIZE(coef, quantval): sign-extending unpack and multiply in one instruction per coefficient
      const int* q = qp[it.comp];
#pragma unroll
      for (int r = 0; r < 8; r++) {
        const uint4 u = *reinterpret_cast<const uint4*>(wt + r * 16);
        const int4 q4 = *reinterpret_cast<const int4*>(q + 4 * r);
        w[8 * r + 0] = dp2a_lo(u.x, q4.x);
        w[8 * r + 1] = dp2a_hi(u.x, q4.x);
        w[8 * r + 2] = dp2a_lo(u.y, q4.y);
        w[8 * r + 3] = dp2a_hi(u.y, q4.y);
        w[8 * r + 4] = dp2a_lo(u.z, q4.z);
        w[8 * r + 5] = dp2a_hi(u.z, q4.z);
        w[8 * r + 6] = dp2a_lo(u.w, q4.w);
        w[8 * r + 7] = dp2a_hi(u.w, q4.w);
      }
    } else {
      const int* q = qs[it.comp];
#pragma unroll
      for (int r = 0; r < 8; r++) {
        const uint4 u = *reinterpret_cast<const uint4*>(wt + r * 16);
        const int4 qa = *reinterpret_cast<const int4*>(q + 8 * r), qb = *reinterpret_cast<const int4*>(q + 8 * r + 4);
        w[8 * r + 0] = (int)(int16_t)(u.x & 0xffff) * qa.x;
        w[8 * r + 1] = ((int)u.x >> 16) * qa.y;
        w[8 * r + 2] = (int)(int16_t)(u.y & 0xffff) * qa.z;
        w[8 * r + 3] = ((int)u.y >> 16) * qa.w;
        w[8 * r + 4] = (int)(int16_t)(u.z & 0xffff) * qb.x;
        w[8 * r + 5] = ((int)u.z >> 16) * qb.y;
        w[8 * r + 6] = (int)(int16_t)(u.w & 0xffff) * qb.z;
        w[8 * r + 7] = ((int)u.w >> 16) * qb.w;
      }
    }
    __syncwarp();  // the tile is free: the next iteration's copies overwrite it
#pragma unroll
    for (int col = 0; col < 8; col++) {  // pass 1: columns
      int v[8], o[8];
#pragma unroll
      for (int k = 0; k < 8; k++) v[k] = w[8 * k + col];
      idct_1d<kConstBits - kPass1Bits>(v, o);
#pragma unroll
      for (int k = 0; k < 8; k++) w[8 * k + col] = o[k];
    }
    const uint32_t bw = (uint32_t)c.blocks_w, i = it.first + lane, by = i / bw, bx = i - by * bw;
    const size_t x = (size_t)bx * 8;
    const bool live = lane < it.count && x < c.width;
    uint8_t* dst = c.plane + (size_t)by * 8 * c.stride + x;
    const bool vec = x + 8 <= c.width && ((reinterpret_cast<uintptr_t>(dst) | c.stride) & 7) == 0;
    const uint32_t rows = !live ? 0 : ((size_t)by * 8 + 8 <= c.height ? 8 : (uint32_t)(c.height - (size_t)by * 8 < 8 ? c.height - (size_t)by * 8 : 0));
#pragma unroll
    for (int r = 0; r < 8; r++) {  // pass 2: rows, stored as they are finished
      int v[8], o[8];
#pragma unroll
      for (int k = 0; k < 8; k++) v[k] = w[8 * r + k];
      idct_1d<kConstBits + kPass1Bits + 3, 4>(v, o);  // o = sext10 of libjpeg's descaled sum
      const uint32_t lo = pack4_level_shift(o[0], o[1], o[2], o[3]), hi = pack4_level_shift(o[4], o[5], o[6], o[7]);
      if ((uint32_t)r < rows) {
        if (vec) {
          stg8(dst, make_uint2(lo, hi));
        } else {  // ragged right edge or an unaligned plane: byte stores
          const uint32_t n = c.width - x < 8 ? (uint32_t)(c.width - x) : 8;
          for (uint32_t k = 0; k < n; k++) dst[k] = (uint8_t)((k < 4 ? lo >> (8 * k) : hi >> (8 * (k - 4))) & 0xff);
        }
      }
      dst += c.stride;
    }
  }
  cp_async_wait<0>();
}

}  // namespace

cudaError_t launch_jpeg_idct_impl(const JpegIdctParams& p, cudaStream_t s) {
  if (p.n_components <= 0) return cudaSuccess;
  size_t bw = 0, bh = 0;
  for (int i = 0; i < p.n_components; i++) {
    bw = p.comp[i].blocks_w > bw ? p.comp[i].blocks_w : bw;
    bh = p.comp[i].blocks_h > bh ? p.comp[i].blocks_h : bh;
  }
  if (bw == 0 || bh == 0) return cudaSuccess;
  size_t tiles = 0;
  bool aligned = true;
  for (int i = 0; i < p.n_components; i++) {
    aligned = aligned && (reinterpret_cast<uintptr_t>(p.comp[i].coeffs) & 15) == 0;
    const size_t nblk = p.comp[i].blocks_w * p.comp[i].blocks_h;
    if (nblk > 0x7fffffffu) return cudaErrorInvalidValue;  // jpeg_check_layout bounds the codec's images, this the raw entry point
    tiles += (nblk + 31) / 32;
  }
  static int sms_of[64], per_sm_of[64];  // persistent grid: resident blocks per device, asked once
  static std::mutex mu;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  dev &= 63;
  std::unique_lock<std::mutex> lk(mu);
  if (per_sm_of[dev] == 0) {
    int n_sm = 0, n = 0;
    if ((e = cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(jpeg_idct_kernel<true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared)) != cudaSuccess ||
        (e = cudaFuncSetAttribute(jpeg_idct_kernel<false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared)) != cudaSuccess)
      return e;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, jpeg_idct_kernel<true>, kIdctBThreads, 0)) != cudaSuccess) return e;
    sms_of[dev] = n_sm;
    per_sm_of[dev] = n > 0 ? n : 1;
  }
  const size_t resident = (size_t)sms_of[dev] * per_sm_of[dev];
  lk.unlock();
  const size_t want = (tiles + kIdctBThreads / 32 - 1) / (kIdctBThreads / 32);
  const unsigned grid = (unsigned)(want < resident ? want : resident);  // any grid is correct: every warp strides over the tiles
  if (aligned) jpeg_idct_kernel<true><<<grid, kIdctBThreads, 0, s>>>(p);
  else jpeg_idct_kernel<false><<<grid, kIdctBThreads, 0, s>>>(p);
  return cudaGetLastError();
}

}  // namespace uhdr
