// jpeg_enc.cu — row N4 (third slice): the reference's JPEG ENCODE of the SDR image and of the gain map,
// JpegEncoderHelper::compressImage / encode (lib/src/jpegencoderhelper.cpp:86-232; call sites jpegr.cpp:234-239 SDR image,
// :1037-1050 gain map), on the device: raw 4:2:0 planes (or one grey plane) in device memory -> forward DCT, quantisation
// and Huffman coding in CUDA -> a baseline JFIF file. The planes the fused API-0 kernel and generateGainMap leave in device
// memory never cross PCIe; only the compressed bytes come down.
//
// What the reference asks libjpeg for (setJpegCompressStruct, :113-130): jpeg_set_defaults + jpeg_set_quality(quality,
// TRUE), raw_data_in, JDCT_ISLOW, sampling 2x2 / 1x1 / 1x1 (or one 1x1 component), the typical Huffman tables, no restart
// intervals; rows and columns beyond the image inside the last MCUs are ZERO samples (compressYuv's padding, :147-205).
// libjpeg-turbo's source is NOT in /root/reference: the codec is restated from ITU-T T.81 (Annex A.3.3 FDCT, Annex F.1.2
// Huffman encoding, Annex K tables, Annex B markers) and from the published IJG algorithm descriptions (jfdctint.c "slow
// but accurate" LL&M FDCT with CONST_BITS 13 / PASS1_BITS 2; jcdctmgr.c quantisation by (|x| + q/2) / q on the 8x scaled
// DCT output; jcparam.c quality scaling; jcmarker.c marker order). Parity is anchored on libjpeg-turbo's own OUTPUT:
// tests/test_gpu_jpeg_encode.py requires files BYTE-IDENTICAL to the ones libjpeg-turbo (through Pillow) writes from the
// same samples.
//
// Kernels: fdct_quant_kernel (8 lanes per 8x8 block, like jpeg_idct_kernel), huff_size_kernel (bits per block in scan
// order), an exclusive scan, huff_emit_kernel (every block writes its bits at its offset; words shared with a neighbour
// through atomicOr). Byte stuffing (FF -> FF 00) and the markers are written by the host while it copies the segment out.
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include <emmintrin.h>

#include "uhdr_io.cuh"
#include "uhdr_kernels.cuh"

namespace uhdr {
namespace {

// ITU-T T.81 Annex K: Tables K.1 / K.2 (quantisation, zig-zag order as a DQT segment carries them) and K.3-K.6 (the typical
// Huffman tables: code-length counts and symbol values), which jpeg_set_defaults installs (jpegencoderhelper.cpp:124).
const uint8_t kStdQuantLuma[64] = {
    16, 11, 12, 14, 12, 10, 16, 14, 13, 14, 18, 17, 16, 19, 24, 40,
    26, 24, 22, 22, 24, 49, 35, 37, 29, 40, 58, 51, 61, 60, 57, 51,
    56, 55, 64, 72, 92, 78, 64, 68, 87, 69, 55, 56, 80, 109, 81, 87,
    95, 98, 103, 104, 103, 62, 77, 113, 121, 112, 100, 120, 92, 101, 103, 99,
};
const uint8_t kStdQuantChroma[64] = {
    17, 18, 18, 24, 21, 24, 47, 26, 26, 47, 99, 66, 56, 66, 99, 99,
    99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99,
    99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99,
    99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99,
};
const uint8_t kStdDcLumaBits[16] = {
    0, 1, 5, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0,
};
const uint8_t kStdDcLumaVals[12] = {
    0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11,
};
const uint8_t kStdDcChromaBits[16] = {
    0, 3, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0,
};
const uint8_t kStdDcChromaVals[12] = {
    0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11,
};
const uint8_t kStdAcLumaBits[16] = {
    0, 2, 1, 3, 3, 2, 4, 3, 5, 5, 4, 4, 0, 0, 1, 125,
};
const uint8_t kStdAcLumaVals[162] = {
    1, 2, 3, 0, 4, 17, 5, 18, 33, 49, 65, 6, 19, 81, 97, 7, 34, 113,
    20, 50, 129, 145, 161, 8, 35, 66, 177, 193, 21, 82, 209, 240, 36, 51, 98, 114,
    130, 9, 10, 22, 23, 24, 25, 26, 37, 38, 39, 40, 41, 42, 52, 53, 54, 55,
    56, 57, 58, 67, 68, 69, 70, 71, 72, 73, 74, 83, 84, 85, 86, 87, 88, 89,
    90, 99, 100, 101, 102, 103, 104, 105, 106, 115, 116, 117, 118, 119, 120, 121, 122, 131,
    132, 133, 134, 135, 136, 137, 138, 146, 147, 148, 149, 150, 151, 152, 153, 154, 162, 163,
    164, 165, 166, 167, 168, 169, 170, 178, 179, 180, 181, 182, 183, 184, 185, 186, 194, 195,
    196, 197, 198, 199, 200, 201, 202, 210, 211, 212, 213, 214, 215, 216, 217, 218, 225, 226,
    227, 228, 229, 230, 231, 232, 233, 234, 241, 242, 243, 244, 245, 246, 247, 248, 249, 250,
};
const uint8_t kStdAcChromaBits[16] = {
    0, 2, 1, 2, 4, 4, 3, 4, 7, 5, 4, 4, 0, 1, 2, 119,
};
const uint8_t kStdAcChromaVals[162] = {
    0, 1, 2, 3, 17, 4, 5, 33, 49, 6, 18, 65, 81, 7, 97, 113, 19, 34,
    50, 129, 8, 20, 66, 145, 161, 177, 193, 9, 35, 51, 82, 240, 21, 98, 114, 209,
    10, 22, 36, 52, 225, 37, 241, 23, 24, 25, 26, 38, 39, 40, 41, 42, 53, 54,
    55, 56, 57, 58, 67, 68, 69, 70, 71, 72, 73, 74, 83, 84, 85, 86, 87, 88,
    89, 90, 99, 100, 101, 102, 103, 104, 105, 106, 115, 116, 117, 118, 119, 120, 121, 122,
    130, 131, 132, 133, 134, 135, 136, 137, 138, 146, 147, 148, 149, 150, 151, 152, 153, 154,
    162, 163, 164, 165, 166, 167, 168, 169, 170, 178, 179, 180, 181, 182, 183, 184, 185, 186,
    194, 195, 196, 197, 198, 199, 200, 201, 202, 210, 211, 212, 213, 214, 215, 216, 217, 218,
    226, 227, 228, 229, 230, 231, 232, 233, 234, 242, 243, 244, 245, 246, 247, 248, 249, 250,
};

const uint8_t kEncZigZag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

// jpeg_quality_scaling + jpeg_add_quant_table with force_baseline (jcparam.c): quality 1..100 -> percentage scale,
// entry = clamp((std * scale + 50) / 100, 1, 255); tables in ZIG-ZAG order like the DQT segment.
void scaled_quant(const uint8_t* std_zz, int quality, uint8_t* out_zz) {
  quality = quality < 1 ? 1 : (quality > 100 ? 100 : quality);
  const int scale = quality < 50 ? 5000 / quality : 200 - quality * 2;
  for (int i = 0; i < 64; i++) {
    long v = ((long)std_zz[i] * scale + 50L) / 100L;
    out_zz[i] = (uint8_t)(v <= 0 ? 1 : (v > 255 ? 255 : v));
  }
}

// Annex C: code and length of every symbol of a table (EHUFCO / EHUFSI)
void build_enc_table(const uint8_t* bits, const uint8_t* vals, JpegEncHuff* t) {
  memset(t, 0, sizeof *t);
  int code = 0, k = 0;
  for (int l = 1; l <= 16; l++) {
    for (int i = 0; i < bits[l - 1]; i++, k++, code++) {
      t->code[vals[k]] = (uint16_t)code;
      t->len[vals[k]] = (uint8_t)l;
    }
    code <<= 1;
  }
}

struct Out {
  uint8_t* p;
  size_t cap, n;
  bool ok;
  void put(const void* src, size_t len) {
    if (n + len > cap) { ok = false; return; }
    memcpy(p + n, src, len);
    n += len;
  }
  void u8(int v) { uint8_t b = (uint8_t)v; put(&b, 1); }
  void u16(int v) { u8(v >> 8); u8(v & 0xff); }
};

void put_dht(Out& o, int tc, int th, const uint8_t* bits, const uint8_t* vals) {
  int n = 0;
  for (int i = 0; i < 16; i++) n += bits[i];
  o.u16(0xFFC4);
  o.u16(2 + 1 + 16 + n);
  o.u8((tc << 4) | th);
  o.put(bits, 16);
  o.put(vals, (size_t)n);
}

}  // namespace

void jpeg_enc_tables(int quality, bool grey, JpegEncTables* t) {
  memset(t, 0, sizeof *t);
  uint8_t qz[2][64];
  scaled_quant(kStdQuantLuma, quality, qz[0]);
  scaled_quant(kStdQuantChroma, quality, qz[1]);
  for (int k = 0; k < 64; k++) {
    t->quant_zz[0][k] = qz[0][k];
    t->quant_zz[1][k] = qz[1][k];
    t->quant[0][kEncZigZag[k]] = qz[0][k];
    t->quant[1][kEncZigZag[k]] = qz[1][k];
  }
  build_enc_table(kStdDcLumaBits, kStdDcLumaVals, &t->dc[0]);
  build_enc_table(kStdAcLumaBits, kStdAcLumaVals, &t->ac[0]);
  build_enc_table(kStdDcChromaBits, kStdDcChromaVals, &t->dc[1]);
  build_enc_table(kStdAcChromaBits, kStdAcChromaVals, &t->ac[1]);
  (void)grey;
}

// SOI, JFIF APP0, [APP2: the caller's ICC payload], DQT per table, SOF0, DHT per table, SOS (jcmarker.c: write_file_header,
// jpeg_write_marker, write_frame_header, write_scan_header). Returns the bytes written, 0 when `cap` is too small.
size_t jpeg_write_headers(uint8_t* out, size_t cap, int width, int height, bool grey, const JpegEncTables& t, const void* icc, size_t icc_bytes) {
  Out o{out, cap, 0, true};
  o.u16(0xFFD8);
  static const uint8_t jfif[14] = {'J', 'F', 'I', 'F', 0, 1, 1, 0, 0, 1, 0, 1, 0, 0};  // version 1.01, no units, 1:1, no thumbnail
  o.u16(0xFFE0);
  o.u16(16);
  o.put(jfif, 14);
  if (icc && icc_bytes) {  // jpeg_write_marker(&cinfo, JPEG_APP0 + 2, iccBuffer, iccSize), jpegencoderhelper.cpp:100-102
    if (icc_bytes > 65533) return 0;
    o.u16(0xFFE2);
    o.u16((int)icc_bytes + 2);
    o.put(icc, icc_bytes);
  }
  for (int q = 0; q < (grey ? 1 : 2); q++) {
    o.u16(0xFFDB);
    o.u16(67);
    o.u8(q);
    o.put(t.quant_zz[q], 64);
  }
  o.u16(0xFFC0);
  o.u16(8 + 3 * (grey ? 1 : 3));
  o.u8(8);
  o.u16(height);
  o.u16(width);
  o.u8(grey ? 1 : 3);
  if (grey) {
    o.u8(1); o.u8(0x11); o.u8(0);
  } else {
    o.u8(1); o.u8(0x22); o.u8(0);
    o.u8(2); o.u8(0x11); o.u8(1);
    o.u8(3); o.u8(0x11); o.u8(1);
  }
  put_dht(o, 0, 0, kStdDcLumaBits, kStdDcLumaVals);
  put_dht(o, 1, 0, kStdAcLumaBits, kStdAcLumaVals);
  if (!grey) {
    put_dht(o, 0, 1, kStdDcChromaBits, kStdDcChromaVals);
    put_dht(o, 1, 1, kStdAcChromaBits, kStdAcChromaVals);
  }
  o.u16(0xFFDA);
  o.u16(6 + 2 * (grey ? 1 : 3));
  o.u8(grey ? 1 : 3);
  o.u8(1); o.u8(0x00);
  if (!grey) {
    o.u8(2); o.u8(0x11);
    o.u8(3); o.u8(0x11);
  }
  o.u8(0); o.u8(63); o.u8(0);  // Ss, Se, Ah / Al
  return o.ok ? o.n : 0;
}

// copies the entropy-coded bytes inserting the stuffed zero after every FF (F.1.2.3); returns bytes written or 0
size_t jpeg_stuff(const uint8_t* src, size_t n, uint8_t* dst, size_t cap) {
  size_t o = 0, i = 0;
  const __m128i ff = _mm_set1_epi8((char)0xFF);
  while (i < n) {
    while (i + 16 <= n && o + 17 <= cap) {  // 16 bytes at a time while no FF is among them
      const __m128i v = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + i));
      const int mask = _mm_movemask_epi8(_mm_cmpeq_epi8(v, ff));
      _mm_storeu_si128(reinterpret_cast<__m128i*>(dst + o), v);
      if (mask) {  // keep the bytes in front of the first FF; the FF itself is handled below
        const int k = __builtin_ctz((unsigned)mask);
        o += (size_t)k;
        i += (size_t)k;
        break;
      }
      o += 16;
      i += 16;
    }
    if (i >= n) break;
    if (o + 2 > cap) return 0;
    const uint8_t b = src[i++];
    dst[o++] = b;
    if (b == 0xFF) dst[o++] = 0x00;
  }
  return o;
}

// ================================================================ device
namespace {

constexpr int kEncConstBits = 13, kEncPass1Bits = 2;
constexpr int EFIX_0_298631336 = 2446, EFIX_0_390180644 = 3196, EFIX_0_541196100 = 4433, EFIX_0_765366865 = 6270;
constexpr int EFIX_0_899976223 = 7373, EFIX_1_175875602 = 9633, EFIX_1_501321110 = 12299, EFIX_1_847759065 = 15137;
constexpr int EFIX_1_961570560 = 16069, EFIX_2_053119869 = 16819, EFIX_2_562915447 = 20995, EFIX_3_072711026 = 25172;

__device__ __forceinline__ int descale(int x, int n) { return (x + (1 << (n - 1))) >> n; }

// One 1-D pass of the LL&M forward DCT on eight values. FIRST: the row pass (results scaled up by 2^PASS1_BITS), else the
// column pass (that scaling removed again; the overall factor of 8 stays and is divided out by the quantisation).
template <bool FIRST>
__device__ __forceinline__ void fdct_1d(const int (&d)[8], int (&o)[8]) {
  const int tmp0 = d[0] + d[7], tmp7 = d[0] - d[7], tmp1 = d[1] + d[6], tmp6 = d[1] - d[6];
  const int tmp2 = d[2] + d[5], tmp5 = d[2] - d[5], tmp3 = d[3] + d[4], tmp4 = d[3] - d[4];
  const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
  constexpr int kShift = FIRST ? kEncConstBits - kEncPass1Bits : kEncConstBits + kEncPass1Bits;
  if (FIRST) {
    o[0] = (int)((unsigned)(tmp10 + tmp11) << kEncPass1Bits);
    o[4] = (int)((unsigned)(tmp10 - tmp11) << kEncPass1Bits);
  } else {
    o[0] = descale(tmp10 + tmp11, kEncPass1Bits);
    o[4] = descale(tmp10 - tmp11, kEncPass1Bits);
  }
  int z1 = (tmp12 + tmp13) * EFIX_0_541196100;
  o[2] = descale(z1 + tmp13 * EFIX_0_765366865, kShift);
  o[6] = descale(z1 + tmp12 * (-EFIX_1_847759065), kShift);
  z1 = tmp4 + tmp7;
  int z2 = tmp5 + tmp6, z3 = tmp4 + tmp6, z4 = tmp5 + tmp7;
  const int z5 = (z3 + z4) * EFIX_1_175875602;
  const int t4 = tmp4 * EFIX_0_298631336, t5 = tmp5 * EFIX_2_053119869, t6 = tmp6 * EFIX_3_072711026, t7 = tmp7 * EFIX_1_501321110;
  z1 *= -EFIX_0_899976223;
  z2 *= -EFIX_2_562915447;
  z3 *= -EFIX_1_961570560;
  z4 *= -EFIX_0_390180644;
  z3 += z5;
  z4 += z5;
  o[7] = descale(t4 + z1 + z3, kShift);
  o[5] = descale(t5 + z2 + z4, kShift);
  o[3] = descale(t6 + z2 + z3, kShift);
  o[1] = descale(t7 + z1 + z4, kShift);
}

// zig-zag position -> natural index (jpeg_natural_order), folded at compile time in the unrolled store loop
__device__ constexpr int zig_natural(int pos) {
  constexpr int t[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                         41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                         30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
  return t[pos];
}

constexpr int kFdctThreads = 128, kFdctTileStride = 144;  // bytes per block record in the shared-memory tile (128 + 16 of padding)

// One THREAD per 8x8 block, blocks in raster order over the component (grid.x chunks of 128 blocks, grid.z components): the
// block lives in 64 registers, both passes run without a hand-over. Eight 8-byte loads fetch the samples (adjacent lanes
// = adjacent blocks: 256 contiguous bytes per warp and row; samples beyond the image are ZERO, jpegencoderhelper.cpp:
// 147-205, then centred like every sample). Quantisation (jcdctmgr.c: (|x| + q/2) / q with q = 8 * step, sign restored)
// divides by multiplication: M = floor(2^32 / q) + 1 gives umulhi(n, M) == n / q for every n < 2^32 / q (q <= 2040, n <
// 2^15 here). The 64 coefficients leave in zig-zag order: each lane writes its 128-byte record into a padded tile and the
// warp stores its 32 consecutive records with coalesced 16-byte stores.
__global__ void __launch_bounds__(kFdctThreads, 4) fdct_quant_kernel(const __grid_constant__ JpegEncParams p) {
  __shared__ __align__(16) unsigned char tile[kFdctThreads * kFdctTileStride];
  __shared__ __align__(8) uint2 qm[64];  // per natural index: q / 2 and the reciprocal M
  const JpegEncComponent& c = p.comp[blockIdx.z];
  const uint32_t nblk = c.blocks_w * c.blocks_h, tid = threadIdx.x, lane = tid & 31;
  if (blockIdx.x * kFdctThreads >= nblk) return;  // the components of one launch differ in size
  if (tid < 64) {
    const uint32_t q = (uint32_t)c.quant[tid] << 3;
    qm[tid] = make_uint2(q >> 1, (uint32_t)(0x100000000ull / q) + 1);
  }
  __syncthreads();
  const uint32_t first = blockIdx.x * kFdctThreads + (tid & ~31u);  // the warp's first block
  if (first >= nblk) return;
  const uint32_t in_warp = nblk - first < 32 ? nblk - first : 32;
  const uint32_t i = first + lane < nblk ? first + lane : nblk - 1;  // idle lanes of the last warp repeat its last block
  const uint32_t by = i / c.blocks_w, bx = i - by * c.blocks_w;
  const size_t x0 = (size_t)bx * 8;
  int w[64];
  {
    const uint8_t* src = c.plane + (size_t)by * 8 * c.stride + x0;
    const bool vec = x0 + 8 <= c.width && ((reinterpret_cast<uintptr_t>(src) | c.stride) & 7) == 0;
    uint2 raw[8];
#pragma unroll
    for (int r = 0; r < 8; r++) {
      raw[r] = make_uint2(0, 0);  // zero samples
      const size_t y = (size_t)by * 8 + r;
      if (y < c.height && x0 < c.width) {
        if (vec) {
          raw[r] = ldg8(src + r * c.stride);
        } else {
          for (int k = 0; k < 8; k++)
            if (x0 + k < c.width) (k < 4 ? raw[r].x : raw[r].y) |= (uint32_t)src[r * c.stride + k] << (8 * (k & 3));
        }
      }
    }
#pragma unroll
    for (int r = 0; r < 8; r++) {  // pass 1: rows
      int v[8], o[8];
      const uint32_t lo = raw[r].x ^ 0x80808080u, hi = raw[r].y ^ 0x80808080u;  // sample - 128 as a signed byte
#pragma unroll
      for (int k = 0; k < 4; k++) {
        v[k] = (int)(int8_t)(lo >> (8 * k));
        v[4 + k] = (int)(int8_t)(hi >> (8 * k));
      }
      fdct_1d<true>(v, o);
#pragma unroll
      for (int k = 0; k < 8; k++) w[8 * r + k] = o[k];
    }
  }
#pragma unroll
  for (int col = 0; col < 8; col++) {  // pass 2: columns, quantised as they are finished
    int v[8], o[8];
#pragma unroll
    for (int k = 0; k < 8; k++) v[k] = w[8 * k + col];
    fdct_1d<false>(v, o);
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const uint2 t = qm[8 * k + col];
      const int x = o[k];
      const int m = (int)__umulhi((uint32_t)(x < 0 ? -x : x) + t.x, t.y);
      w[8 * k + col] = x < 0 ? -m : m;
    }
  }
  unsigned char* wt = tile + (tid & ~31u) * kFdctTileStride;
#pragma unroll
  for (int j = 0; j < 8; j++) {  // the record in zig-zag order
    uint32_t u[4];
#pragma unroll
    for (int k = 0; k < 4; k++)
      u[k] = ((uint32_t)w[zig_natural(8 * j + 2 * k)] & 0xffffu) | ((uint32_t)w[zig_natural(8 * j + 2 * k + 1)] << 16);
    *reinterpret_cast<uint4*>(wt + lane * kFdctTileStride + j * 16) = make_uint4(u[0], u[1], u[2], u[3]);
  }
  __syncwarp();
  unsigned char* dst = reinterpret_cast<unsigned char*>(c.coeffs + (size_t)first * 64);
#pragma unroll
  for (int j = 0; j < 8; j++) {
    const uint32_t g = j * 32 + lane;  // 16-byte chunk g of the warp's records: block g / 8, chunk g % 8
    if (g < in_warp * 8) stg16(dst + (size_t)g * 16, *reinterpret_cast<const uint4*>(wt + (g >> 3) * kFdctTileStride + (g & 7) * 16));
  }
}

// ---- Huffman coding (F.1.2): one thread per block, blocks in scan order
__device__ __forceinline__ uint32_t nbits_of(uint32_t a) { return 32u - (uint32_t)__clz((int)a); }  // SSSS of |v| (0 for 0)

struct EncBlockRef {
  const int16_t* zz;  // the block's 64 coefficients in zig-zag order
  int pred;           // DC prediction: the previous block of the same component in scan order (F.1.2.1.1)
  uint32_t tbl;       // 0 luma tables, 1 chroma tables
};
__device__ __forceinline__ const int16_t* enc_block_addr(const JpegScan& sc, uint32_t comp, uint32_t mcu, uint32_t i) {
  const uint32_t h = sc.n_comp == 1 ? 1u : sc.comp_h[comp], v = sc.n_comp == 1 ? 1u : sc.comp_v[comp];
  const uint32_t my = mcu / sc.mcus_x, mx = mcu - my * sc.mcus_x;
  const uint32_t bx = mx * h + i % h, by = my * v + i / h;
  return sc.coeffs[comp] + ((size_t)by * sc.comp_blocks_w[comp] + bx) * 64;
}
__device__ __forceinline__ EncBlockRef enc_block(const JpegScan& sc, uint32_t b) {
  const uint32_t B = (uint32_t)sc.blocks_per_mcu, mcu = b / B, cim = b - mcu * B;
  const uint32_t comp = sc.blk_comp[cim];
  const uint32_t h = sc.n_comp == 1 ? 1u : sc.comp_h[comp], v = sc.n_comp == 1 ? 1u : sc.comp_v[comp], nb = h * v;
  const uint32_t i = sc.blk_dy[cim] * h + sc.blk_dx[cim];
  EncBlockRef r;
  r.zz = enc_block_addr(sc, comp, mcu, i);
  r.tbl = comp == 0 ? 0u : 1u;
  if (i > 0) r.pred = enc_block_addr(sc, comp, mcu, i - 1)[0];
  else if (mcu > 0) r.pred = enc_block_addr(sc, comp, mcu - 1, nb - 1)[0];
  else r.pred = 0;
  return r;
}

// walks one block and hands every (code, length) pair, extra bits appended, to `put` in stream order
template <class Put>
__device__ __forceinline__ void encode_block(const JpegEncTables* __restrict__ t, const EncBlockRef& blk, Put put) {
  uint4 raw[8];
#pragma unroll
  for (int i = 0; i < 8; i++) raw[i] = __ldg(reinterpret_cast<const uint4*>(blk.zz) + i);
  const JpegEncHuff& dc = t->dc[blk.tbl];
  const JpegEncHuff& ac = t->ac[blk.tbl];
  auto coef = [&](int k) -> int {  // static k after unrolling: registers only
    const uint4& q = raw[k >> 3];
    const uint32_t wv = ((k >> 1) & 3) == 0 ? q.x : (((k >> 1) & 3) == 1 ? q.y : (((k >> 1) & 3) == 2 ? q.z : q.w));
    return (int)(int16_t)((k & 1) ? (wv >> 16) : (wv & 0xffffu));
  };
  {  // DC difference, F.1.2.1: category SSSS, then the low SSSS bits of diff (diff - 1 when negative)
    const int diff = coef(0) - blk.pred;
    const uint32_t a = (uint32_t)(diff < 0 ? -diff : diff), s = nbits_of(a);
    const uint32_t extra = (uint32_t)(diff < 0 ? diff - 1 : diff) & ((1u << s) - 1u);
    put(((uint32_t)__ldg(&dc.code[s]) << s) | extra, (uint32_t)__ldg(&dc.len[s]) + s);
  }
  uint32_t run = 0;
#pragma unroll
  for (int k = 1; k < 64; k++) {  // AC coefficients, F.1.2.2: run / size symbols, ZRL for runs above 15, EOB
    const int v = coef(k);
    if (v == 0) {
      run++;
    } else {
      while (run > 15) {
        put((uint32_t)__ldg(&ac.code[0xF0]), (uint32_t)__ldg(&ac.len[0xF0]));
        run -= 16;
      }
      const uint32_t a = (uint32_t)(v < 0 ? -v : v), s = nbits_of(a);
      const uint32_t extra = (uint32_t)(v < 0 ? v - 1 : v) & ((1u << s) - 1u), sym = (run << 4) | s;
      put(((uint32_t)__ldg(&ac.code[sym]) << s) | extra, (uint32_t)__ldg(&ac.len[sym]) + s);
      run = 0;
    }
  }
  if (run > 0) put((uint32_t)__ldg(&ac.code[0]), (uint32_t)__ldg(&ac.len[0]));
}

__global__ void __launch_bounds__(128) huff_size_kernel(const __grid_constant__ JpegScan sc, const JpegEncTables* __restrict__ t,
                                                        uint32_t* __restrict__ bits) {
  const uint32_t b = blockIdx.x * 128 + threadIdx.x;
  if (b >= sc.total_blocks) return;
  uint32_t n = 0;
  encode_block(t, enc_block(sc, b), [&](uint32_t, uint32_t len) { n += len; });
  bits[b] = n;
}

// every block writes its bits at its offset; a 32-bit word shared with a neighbouring block is completed by atomicOr on
// the zero-filled buffer, words inside a block's own range are plain stores. Words hold the stream bytes in order.
__global__ void __launch_bounds__(128) huff_emit_kernel(const __grid_constant__ JpegScan sc, const JpegEncTables* __restrict__ t,
                                                        const uint32_t* __restrict__ offset, uint32_t* __restrict__ out) {
  const uint32_t b = blockIdx.x * 128 + threadIdx.x;
  if (b >= sc.total_blocks) return;
  const uint32_t off = offset[b];
  uint32_t* wp = out + (off >> 5);
  uint64_t acc = 0;
  uint32_t nacc = off & 31u;  // the first word's leading bits belong to the previous block: zeros here, ORed in there
  bool first = true;
  encode_block(t, enc_block(sc, b), [&](uint32_t code, uint32_t len) {
    acc = (acc << len) | code;
    nacc += len;
    if (nacc >= 32) {
      nacc -= 32;
      const uint32_t word = __byte_perm((uint32_t)(acc >> nacc), 0, 0x0123);
      if (first) atomicOr(wp, word);
      else *wp = word;
      first = false;
      wp++;
    }
  });
  if (nacc) atomicOr(wp, __byte_perm((uint32_t)(acc << (32 - nacc)), 0, 0x0123));
}

// ---- exclusive scan of the per-block bit counts (three phases, 2048 elements per CTA)
constexpr int kScanChunk = 2048, kScanThreads = 256, kScanPer = kScanChunk / kScanThreads;
__device__ __forceinline__ uint32_t cta_inclusive(uint32_t v, uint32_t* warp_sums) {
  const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t o = __shfl_up_sync(0xffffffffu, v, d);
    if (lane >= (uint32_t)d) v += o;
  }
  if (lane == 31) warp_sums[w] = v;
  __syncthreads();
  uint32_t base = 0;
  for (uint32_t i = 0; i < w; i++) base += warp_sums[i];
  __syncthreads();
  return v + base;
}
__global__ void __launch_bounds__(kScanThreads) scan_partial_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ partial, uint32_t n) {
  __shared__ uint32_t wsum[kScanThreads / 32];
  const uint32_t j0 = blockIdx.x * kScanChunk + threadIdx.x * kScanPer;
  uint32_t s = 0;
  for (uint32_t i = 0; i < (uint32_t)kScanPer; i++)
    if (j0 + i < n) s += in[j0 + i];
  const uint32_t incl = cta_inclusive(s, wsum);
  if (threadIdx.x == kScanThreads - 1) partial[blockIdx.x] = incl;
}
__global__ void __launch_bounds__(1024) scan_offsets_kernel(uint32_t* __restrict__ partial, uint32_t chunks, uint32_t* __restrict__ total) {
  __shared__ uint32_t wsum[32];
  __shared__ uint32_t chunk_total;
  uint32_t carry = 0;
  for (uint32_t base = 0; base < chunks; base += 1024) {
    const uint32_t i = base + threadIdx.x, v = i < chunks ? partial[i] : 0;
    const uint32_t incl = cta_inclusive(v, wsum);
    if (threadIdx.x == 1023) chunk_total = incl;
    if (i < chunks) partial[i] = carry + incl - v;
    __syncthreads();
    carry += chunk_total;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = carry;
}
__global__ void __launch_bounds__(kScanThreads) scan_apply_kernel(const uint32_t* __restrict__ in, const uint32_t* __restrict__ partial,
                                                                  uint32_t* __restrict__ out, uint32_t n) {
  __shared__ uint32_t wsum[kScanThreads / 32];
  const uint32_t j0 = blockIdx.x * kScanChunk + threadIdx.x * kScanPer;
  uint32_t d[kScanPer], s = 0;
  for (uint32_t i = 0; i < (uint32_t)kScanPer; i++) {
    d[i] = j0 + i < n ? in[j0 + i] : 0;
    s += d[i];
  }
  uint32_t run = cta_inclusive(s, wsum) - s + partial[blockIdx.x];
  for (uint32_t i = 0; i < (uint32_t)kScanPer; i++) {
    if (j0 + i < n) out[j0 + i] = run;
    run += d[i];
  }
}

}  // namespace

cudaError_t launch_jpeg_fdct_impl(const JpegEncParams& p, cudaStream_t s) {
  uint64_t most = 0;
  for (int i = 0; i < p.n_components; i++) {
    const uint64_t nblk = (uint64_t)p.comp[i].blocks_w * p.comp[i].blocks_h;
    if (nblk > 0x7fffffffu || (reinterpret_cast<uintptr_t>(p.comp[i].coeffs) & 15) != 0) return cudaErrorInvalidValue;  // 16-byte stores
    most = nblk > most ? nblk : most;
  }
  if (most == 0 || p.n_components <= 0) return cudaSuccess;
  fdct_quant_kernel<<<dim3((unsigned)((most + kFdctThreads - 1) / kFdctThreads), 1, (unsigned)p.n_components), kFdctThreads, 0, s>>>(p);
  return cudaGetLastError();
}

size_t jpeg_enc_workspace_bytes(uint32_t total_blocks) {
  const size_t n = ((size_t)total_blocks + 63) & ~(size_t)63;
  return n * 8 + ((n + kScanChunk - 1) / kScanChunk + 64) * 4 + 64;
}

// bits per block -> offsets -> total (left in `total_bits_dev`, one u32) -> the bit stream in `out_words` (zero-filled by
// the caller, capacity >= ceil(total / 32) + 1 words; the caller reads the total back before it sizes the copy)
cudaError_t jpeg_huff_encode(const JpegScan& sc, const JpegEncTables* tables_dev, void* workspace, uint32_t* total_bits_dev, cudaStream_t s,
                             uint32_t** offsets_out) {
  const uint32_t n = sc.total_blocks;
  if (n == 0) return cudaSuccess;
  const size_t np = ((size_t)n + 63) & ~(size_t)63;
  uint32_t* bits = (uint32_t*)workspace;
  uint32_t* offs = bits + np;
  uint32_t* partial = offs + np;
  const uint32_t chunks = (n + kScanChunk - 1) / kScanChunk;
  huff_size_kernel<<<(n + 127) / 128, 128, 0, s>>>(sc, tables_dev, bits);
  scan_partial_kernel<<<chunks, kScanThreads, 0, s>>>(bits, partial, n);
  scan_offsets_kernel<<<1, 1024, 0, s>>>(partial, chunks, total_bits_dev);
  scan_apply_kernel<<<chunks, kScanThreads, 0, s>>>(bits, partial, offs, n);
  if (offsets_out) *offsets_out = offs;
  return cudaGetLastError();
}
cudaError_t jpeg_huff_emit(const JpegScan& sc, const JpegEncTables* tables_dev, const uint32_t* offsets, uint32_t* out_words, cudaStream_t s) {
  if (sc.total_blocks == 0) return cudaSuccess;
  huff_emit_kernel<<<(sc.total_blocks + 127) / 128, 128, 0, s>>>(sc, tables_dev, offsets, out_words);
  return cudaGetLastError();
}

}  // namespace uhdr
