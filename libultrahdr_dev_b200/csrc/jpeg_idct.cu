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
// Mapping: 8 lanes per 8x8 block, 4 blocks per warp, 8 warps per CTA = 32 consecutive blocks of one block row. Lane r of
// a block reads coefficient column r, dequantises, runs the column pass, passes the block through a padded shared-memory
// workspace once, runs the row pass on row r and stores 8 output bytes (one 8-byte store; a block row is contiguous in
// the plane). HBM traffic: 128 B of coefficients in, 64 B of samples out per block = 3 B per sample.
#include <cstring>

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
template <int SHIFT>
__device__ __forceinline__ void idct_1d(const int (&in)[8], int (&out)[8]) {
  int z2 = in[2], z3 = in[6];
  int z1 = (z2 + z3) * FIX_0_541196100;
  int tmp2 = z1 - z3 * FIX_1_847759065;
  int tmp3 = z1 + z2 * FIX_0_765366865;
  z2 = in[0];
  z3 = in[4];
  int tmp0 = (int)((unsigned)(z2 + z3) << kConstBits);
  int tmp1 = (int)((unsigned)(z2 - z3) << kConstBits);
  const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
  tmp0 = in[7];
  tmp1 = in[5];
  tmp2 = in[3];
  tmp3 = in[1];
  z1 = tmp0 + tmp3;
  z2 = tmp1 + tmp2;
  z3 = tmp0 + tmp2;
  int z4 = tmp1 + tmp3;
  const int z5 = (z3 + z4) * FIX_1_175875602;
  tmp0 *= FIX_0_298631336;
  tmp1 *= FIX_2_053119869;
  tmp2 *= FIX_3_072711026;
  tmp3 *= FIX_1_501321110;
  z1 *= -FIX_0_899976223;
  z2 *= -FIX_2_562915447;
  z3 *= -FIX_1_961570560;
  z4 *= -FIX_0_390180644;
  z3 += z5;
  z4 += z5;
  tmp0 += z1 + z3;
  tmp1 += z2 + z4;
  tmp2 += z2 + z3;
  tmp3 += z1 + z4;
  constexpr int kRound = 1 << (SHIFT - 1);
  out[0] = (tmp10 + tmp3 + kRound) >> SHIFT;
  out[7] = (tmp10 - tmp3 + kRound) >> SHIFT;
  out[1] = (tmp11 + tmp2 + kRound) >> SHIFT;
  out[6] = (tmp11 - tmp2 + kRound) >> SHIFT;
  out[2] = (tmp12 + tmp1 + kRound) >> SHIFT;
  out[5] = (tmp12 - tmp1 + kRound) >> SHIFT;
  out[3] = (tmp13 + tmp0 + kRound) >> SHIFT;
  out[4] = (tmp13 - tmp0 + kRound) >> SHIFT;
}

// libjpeg's range_limit table (jdmaster.c prepare_range_limit_table) indexed with `x & RANGE_MASK` (1023) is x + 128
// clamped to 0..255 for -512 <= x < 512 plus a documented wrap-around for corrupt data beyond that; both together are
// clamp(sext10(x) + 128, 0, 255). Two samples per saturating pack.
__device__ __forceinline__ int sext10(int x) {
  int r;
  asm("bfe.s32 %0, %1, 0, 10;" : "=r"(r) : "r"(x));
  return r;
}
__device__ __forceinline__ uint32_t pack4_u8_sat(int a, int b, int c, int d) {  // bytes: a | b << 8 | c << 16 | d << 24
  uint32_t hi, r;  // cvt.pack d, x, y, z = sat(y) | sat(x) << 8 | z << 16
  asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, 0;" : "=r"(hi) : "r"(d), "r"(c));
  asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(b), "r"(a), "r"(hi));
  return r;
}

constexpr int kIdctThreads = 256, kBlocksPerCta = kIdctThreads / 8;
// shared-memory workspace per WARP: the four 8x8 blocks side by side as one 8 x 32 matrix with rows padded to 33 words:
// a column-pass write (row k, lanes = 32 consecutive columns) and a row-pass read (lane (block b, row r) reads word
// r * 33 + 8 b + k: bank r + 8 b + k, all 32 different) are both free of bank conflicts
constexpr int kWsRow = 33, kWsWarp = 8 * kWsRow;

// CTA = a strip of 32 block columns walked kRowsPerCta block rows down (grid.x strips, grid.y row groups, grid.z
// components); 8 lanes per block. Lane r reads coefficient COLUMN r (eight 2-byte loads; the eight lanes of a block cover
// each 16-byte row together), dequantises with the column's eight quantisation steps held in registers, runs the column
// pass, hands the block over through shared memory once, runs the row pass on row r and stores its eight samples.
constexpr int kRowsPerCta = 8;  // block rows a CTA walks down its strip of 32 block columns

__global__ void __launch_bounds__(kIdctThreads) jpeg_idct_kernel(const __grid_constant__ JpegIdctParams p) {
  __shared__ int ws[(kIdctThreads / 32) * kWsWarp];
  const JpegIdctComponent& c = p.comp[blockIdx.z];
  const uint32_t tid = threadIdx.x, r = tid & 7, lb = tid >> 3;
  const uint32_t bw = (uint32_t)c.blocks_w, bh = (uint32_t)c.blocks_h, bx = blockIdx.x * kBlocksPerCta + lb;
  const uint32_t by0 = blockIdx.y * kRowsPerCta;
  if (by0 >= bh || bx >= bw) return;  // the components of one launch differ in size; a warp's four blocks may straddle the right edge
  const uint32_t by1 = by0 + kRowsPerCta < bh ? by0 + kRowsPerCta : bh;
  const unsigned lanes = __activemask();  // the lanes of this warp that own a block column (uniform over the walk)
  int* w = ws + (tid >> 5) * kWsWarp + (lb & 3) * 8;  // this block's eight columns inside its warp's matrix
  int q[8];
#pragma unroll
  for (int k = 0; k < 8; k++) q[k] = (int)__ldg(c.quant + 8 * k + r);  // the quantisation steps of column r
  const size_t row_step = (size_t)bw * 64;
  const int16_t* cp = c.coeffs + ((size_t)by0 * bw + bx) * 64 + r;
  int nv[8];
#pragma unroll
  for (int k = 0; k < 8; k++) nv[k] = (int)__ldg(cp + 8 * k);
  const size_t x = (size_t)bx * 8;
  const bool wide = x + 8 <= c.width;
  for (uint32_t by = by0; by < by1; by++) {
    int v[8], o[8];
#pragma unroll
    for (int k = 0; k < 8; k++) v[k] = nv[k] * q[k];  // DEQUANTIZE(coef, quantval)
    cp += row_step;
    if (by + 1 < by1) {  // the next block's coefficients are in flight while this one is transformed
#pragma unroll
      for (int k = 0; k < 8; k++) nv[k] = (int)__ldg(cp + 8 * k);
    }
    idct_1d<kConstBits - kPass1Bits>(v, o);  // pass 1: column r
#pragma unroll
    for (int k = 0; k < 8; k++) w[k * kWsRow + r] = o[k];
    __syncwarp(lanes);
#pragma unroll
    for (int k = 0; k < 8; k++) v[k] = w[r * kWsRow + k];
    __syncwarp(lanes);  // the workspace is rewritten by the next block row
    idct_1d<kConstBits + kPass1Bits + 3>(v, o);  // pass 2: row r
    const size_t y = (size_t)by * 8 + r;
    if (y >= c.height || x >= c.width) continue;
    const uint32_t lo = pack4_u8_sat(sext10(o[0]) + 128, sext10(o[1]) + 128, sext10(o[2]) + 128, sext10(o[3]) + 128);
    const uint32_t hi = pack4_u8_sat(sext10(o[4]) + 128, sext10(o[5]) + 128, sext10(o[6]) + 128, sext10(o[7]) + 128);
    uint8_t* dst = c.plane + y * c.stride + x;
    if (wide && ((reinterpret_cast<uintptr_t>(dst) & 7) == 0)) {
      stg8(dst, make_uint2(lo, hi));
    } else {  // ragged right edge or an unaligned plane: byte stores
      const size_t n = c.width - x < 8 ? c.width - x : 8;
      for (size_t k = 0; k < n; k++) dst[k] = (uint8_t)((k < 4 ? lo >> (8 * k) : hi >> (8 * (k - 4))) & 0xff);
    }
  }
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
  if (bh > 65535u * kRowsPerCta) return cudaErrorInvalidValue;  // far beyond any JPEG (65 535 x 65 535 samples)
  const dim3 grid((unsigned)((bw + kBlocksPerCta - 1) / kBlocksPerCta), (unsigned)((bh + kRowsPerCta - 1) / kRowsPerCta), (unsigned)p.n_components);
  jpeg_idct_kernel<<<grid, kIdctThreads, 0, s>>>(p);
  return cudaGetLastError();
}

}  // namespace uhdr
