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
// Mapping: 8 lanes per 8x8 block, 4 blocks per warp, 8 warps per CTA. Lane r of a block loads coefficient row r as one
// 16-byte vector and dequantises it into a padded shared-memory workspace, runs the column pass on column r, then the
// row pass on row r, and stores 8 output bytes (one 8-byte store; a block row is contiguous in the plane). HBM traffic:
// 128 B of coefficients in, 64 B of samples out per block = 3 B per sample.
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

// libjpeg's range_limit table (jdmaster.c prepare_range_limit_table) indexed with `x & RANGE_MASK` (1023): x + 128
// clamped to 0..255 for -512 <= x < 512, and its documented wrap-around for corrupt data beyond that
__device__ __forceinline__ uint32_t range_limit(int x) {
  const int i = x & 1023;
  return (uint32_t)(i < 128 ? i + 128 : (i < 512 ? 255 : (i < 896 ? 0 : i - 896)));
}

constexpr int kIdctThreads = 256, kBlocksPerCta = kIdctThreads / 8;
constexpr int kWsStride = 8 * 9 + 1;  // words per block in shared memory: rows padded to 9, blocks staggered by 1 (bank spread)

__global__ void __launch_bounds__(kIdctThreads) jpeg_idct_kernel(const __grid_constant__ JpegIdctParams p) {
  __shared__ int ws[kBlocksPerCta * kWsStride];
  __shared__ int sq[64];
  const JpegIdctComponent& c = p.comp[blockIdx.y];
  const int tid = threadIdx.x, r = tid & 7, lb = tid >> 3;
  if (tid < 64) sq[tid] = (int)c.quant[tid];
  __syncthreads();
  const size_t n_blocks = c.blocks_w * c.blocks_h;
  int* w = ws + lb * kWsStride;
  for (size_t b0 = (size_t)blockIdx.x * kBlocksPerCta; b0 < n_blocks; b0 += (size_t)gridDim.x * kBlocksPerCta) {
    const size_t b = b0 + lb;
    const bool live = b < n_blocks;
    int v[8], o[8];
    if (live) {  // coefficient row r, dequantised (DEQUANTIZE(coef, quantval) = coef * quantval)
      const uint4 q = ldg16(c.coeffs + b * 64 + r * 8);
      const uint32_t qw[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
      for (int i = 0; i < 4; i++) {
        w[r * 9 + 2 * i] = (int)(short)(qw[i] & 0xffff) * sq[r * 8 + 2 * i];
        w[r * 9 + 2 * i + 1] = (int)(short)(qw[i] >> 16) * sq[r * 8 + 2 * i + 1];
      }
    }
    __syncwarp();
    if (live) {  // pass 1: column r
#pragma unroll
      for (int k = 0; k < 8; k++) v[k] = w[k * 9 + r];
      idct_1d<kConstBits - kPass1Bits>(v, o);
    }
    __syncwarp();
    if (live) {
#pragma unroll
      for (int k = 0; k < 8; k++) w[k * 9 + r] = o[k];
    }
    __syncwarp();
    if (live) {  // pass 2: row r, range-limit, store the visible part of the row
#pragma unroll
      for (int k = 0; k < 8; k++) v[k] = w[r * 9 + k];
      idct_1d<kConstBits + kPass1Bits + 3>(v, o);
      const size_t by = b / c.blocks_w, bx = b - by * c.blocks_w;
      const size_t y = by * 8 + r, x = bx * 8;
      if (y < c.height && x < c.width) {
        uint32_t lo = 0, hi = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
          lo |= range_limit(o[k]) << (8 * k);
          hi |= range_limit(o[k + 4]) << (8 * k);
        }
        uint8_t* dst = c.plane + y * c.stride + x;
        if (x + 8 <= c.width && ((reinterpret_cast<uintptr_t>(dst) & 7) == 0)) {
          stg8(dst, make_uint2(lo, hi));
        } else {  // ragged right edge or an unaligned plane: byte stores
          const size_t n = c.width - x < 8 ? c.width - x : 8;
          for (size_t k = 0; k < n; k++) dst[k] = (uint8_t)((k < 4 ? lo >> (8 * k) : hi >> (8 * (k - 4))) & 0xff);
        }
      }
    }
    __syncwarp();
  }
}

}  // namespace

cudaError_t launch_jpeg_idct_impl(const JpegIdctParams& p, cudaStream_t s) {
  if (p.n_components <= 0) return cudaSuccess;
  size_t most = 0;
  for (int i = 0; i < p.n_components; i++) {
    const size_t n = p.comp[i].blocks_w * p.comp[i].blocks_h;
    most = n > most ? n : most;
  }
  if (most == 0) return cudaSuccess;
  size_t ctas = (most + kBlocksPerCta - 1) / kBlocksPerCta;
  if (ctas > 148 * 16) ctas = 148 * 16;  // grid-stride beyond 16 resident-CTA waves per SM
  jpeg_idct_kernel<<<dim3((unsigned)ctas, (unsigned)p.n_components), kIdctThreads, 0, s>>>(p);
  return cudaGetLastError();
}

}  // namespace uhdr
