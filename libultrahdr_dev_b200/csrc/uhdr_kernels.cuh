// uhdr_kernels.cuh — launch interface between the C ABI (capi.cu) and the kernels.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

#include "uhdr_math.cuh"

namespace uhdr {

// Device plane pointers of ONE image (all device memory).
struct Planes {
  const void* sdr_y;   // u8  luma of the YUV420 image
  const void* sdr_u;   // u8  U plane
  const void* sdr_v;   // u8  V plane  (= chroma_data + chroma_stride*(height/2), gainmapmath.cpp:568)
  const void* hdr_y;   // u16 luma of the P010 image
  const void* hdr_uv;  // u16 interleaved UV
  const void* map_in;  // u8  gain map (apply)
  void* out;           // map (generate) / packed pixels (apply) / dest luma (toneMap)
  void* out_u;         // toneMap: dest U plane
  void* out_v;         // toneMap: dest V plane
};

// Geometry shared by every image of a launch. Strides in ELEMENTS, as in the reference structs.
struct Geometry {
  int width, height;
  int sdr_luma_stride, sdr_chroma_stride;
  int hdr_luma_stride, hdr_chroma_stride;
  int dst_luma_stride, dst_chroma_stride;  // toneMap destination
  int map_width, map_height, map_scale;
};

// ---- LUT numeric mode (lut tables of the reference's -DUSE_*_LUT=1 build, gainmapmath.cpp:21-64) ----
// Nearest-entry tables over [0, 1]: index = uint32(x * (N - 1) + 0.5) clipped to N - 1 (gainmapmath.cpp:162-167, 269-275,
// 292-298, 316-322, 344-350); sizes gainmapmath.h:268-269, 329-330, 342-343, 355-356, 368-369. Built on the host with the
// reference's own float / double arithmetic (tables.cu), one device copy per GPU (capi.cu: get_lut_tables).
constexpr int kLutSrgbInv = 1024, kLutHlgInv = 4096, kLutPqInv = 4096, kLutHlgOetf = 65536, kLutPqOetf = 65536;
constexpr int kLutGain = 1024;  // GainLUT, gainmapmath.h:150-183 (one table per metadata / display boost)
struct LutTables {
  const float* srgb_inv;
  const float* hlg_inv;
  const float* pq_inv;
  const float* hlg_oetf;
  const float* pq_oetf;
};
void build_lut_tables(float* srgb_inv, float* hlg_inv, float* pq_inv, float* hlg_oetf, float* pq_oetf);
void build_gain_lut(float* tab, float min_boost, float max_boost, float display_boost);

struct GenerateParams {
  Geometry g;
  int lut;          // 1: LUT numeric mode (nearest-entry tables, the reference's float order, exact encodeGain)
  LutTables lt;
  YuvCoefs sdr_k;   // YUV matrix of the SDR image (BT.601 when sdr_is_601, ultrahdr.cpp:284-286)
  YuvCoefs hdr_k;   // YUV matrix of the HDR image (ultrahdr.cpp:288-302)
  LumaCoefs luma;   // luminance weights of the SDR gamut, used for BOTH images (ultrahdr.cpp:265-283)
  Mat3 gamut;       // HDR gamut -> SDR gamut (getHdrConversionFn, gainmapmath.cpp:397-440)
  int use_gamut;    // 0 when both gamuts are equal (identityConversion)
  int hdr_tf;       // 0 linear, 1 HLG, 2 PQ
  float hdr_white_nits;
  GainCode gc;      // encodeGain constants (uhdr_math.cuh), filled by the host
  int pdl;          // 1: launch with programmatic stream serialization (uhdr_io.cuh: launch_pdl)
  int convert;                   // fused API-0 only: 1 = the SDR planes are written re-encoded with `conv` (row N1, jpegr.cpp:223-225)
  float conv[9];                 // YUV->YUV matrix, gainmapmath.cpp:447-481
  Planes single;
  const Planes* batch;  // device array, indexed by blockIdx.z; nullptr => `single`
};

struct ApplyParams {
  Geometry g;
  int lut;                // 1: LUT numeric mode: srgbInvOetfLUT, applyGainLUT, hlg/pqOetfLUT (ultrahdr.cpp:433-485)
  LutTables lt;
  const float* gain_lut;  // device, kLutGain floats
  int output_format;      // UHDRB200_OUTPUT_*
  float log2_min, log2_max;
  float max_boost;        // maxContentBoost (gainmapmath.cpp:553: logBoost * displayBoost / maxContentBoost)
  float inv_display_boost;
  float display_boost;
  const float* idw;       // device: 4 tables x scale^2 x 4 floats (mWeights, NR, NB, C)
  const float* e255;      // device: 256 floats, i / 255.0f correctly rounded (ApplyTables::e255)
  Planes single;
  const Planes* batch;
};

// ---- table-driven fast path of applyGainMap (apply_fast.cu) ----
// Piecewise-linear tables (value = a + b * x per segment), built on the host in double
// (tables.cu) and staged into shared memory by a TMA bulk copy.
//   S table: srgbInvOetf (gainmapmath.cpp:149-155) on UNIFORM segments of c in [0,1]:
//            index = floor(c * kSSegs) comes out of one FFMA with round-to-minus-infinity.
//   O table: x = e + 2^kOLoExp in [2^kOLoExp, 2^kOHiExp) -> 1023 * hlgOetf(e) or 1023 * pqOetf(e)
//            (gainmapmath.cpp:259-267, 309-314; the 1023 of colorToRgba1010102, :722-727, is folded in) on
//            LOG-spaced segments: the float's exponent and top kOBits mantissa bits are the index.
constexpr int kSSegs = 1536;      // RGBA1010102 kernel: 12 KB, next to the 28 KB O table in static shared memory
constexpr int kSSegsHi = 6144;    // linear outputs (RGBA F16, planar 10-bit) need relative accuracy: 48 KB, no O table
constexpr int kSEntries = kSSegs + 2;      // + entry for c == 1.0, padded to a 16-byte multiple
constexpr int kSEntriesHi = kSSegsHi + 2;
constexpr int kOLoExp = -26, kOHiExp = 2, kOBits = 7;
constexpr int kOEntries = (kOHiExp - kOLoExp) << kOBits;      // 3584 float2
constexpr int kOBase = (kOLoExp + 127) << kOBits;
//   P table (q-domain kernel, apply_fast.cu): srgbInvOetf above its linear toe is Phi(c)^2.4 with
//            Phi = (c + 0.055) / 1.055 LINEAR in the 8-bit luma, so with f' = f^(1/2.4) the whole chain
//            OETF(f * srgbInvOetf(c)) is a function of q = Phi * f' alone:
//            P(q + 2^kPLoExp) = 1023 * OETF(q^2.4), log-spaced segments like the O table.
#ifndef UHDR_P_BITS
#define UHDR_P_BITS 7  // 7 bits per octave: 99.93 % of the 10-bit codes exact (8 bits: 99.98 %), 5 % faster (fewer bank conflicts)
#endif
constexpr int kPLoExp = -12, kPHiExp = 1, kPBits = UHDR_P_BITS;
constexpr int kPEntries = (kPHiExp - kPLoExp) << kPBits;      // 1664 float2 = 13 KB
constexpr int kPBase = (kPLoExp + 127) << kPBits;
//   L table (q-domain kernel, linear outputs RGBA F16 / planar 10-bit): L(q) = q^2.4 = f * srgbInvOetf(c) above the
//            toe, 9 bits per octave (relative error 8e-7) over q in [2^-6, 2).
#ifndef UHDR_L_BITS
#define UHDR_L_BITS 9  // 99.96 % of the halves bit-exact, same as 10 bits; 28 KB instead of 56 KB and 14 % faster. 8 bits: 99.89 %, +26 %
#endif
constexpr int kLLoExp = -6, kLHiExp = 1, kLBits = UHDR_L_BITS;
constexpr int kLEntries = (kLHiExp - kLLoExp) << kLBits;      // 3584 float2 = 28 KB
constexpr int kLBase = (kLLoExp + 127) << kLBits;

struct ApplyTables {       // device pointers, one set per device
  const float2* s;         // kSEntries
  const float2* s_hi;      // kSEntriesHi
  const float2* o_hlg;     // kOEntries
  const float2* o_pq;      // kOEntries
  const float2* p_hlg;     // kPEntries
  const float2* p_pq;      // kPEntries
  const float2* l_pow;     // kLEntries
  const float* e255;       // 256 floats: i / 255.0f, correctly rounded (gainmapmath.cpp:700-703)
};

struct ApplyFastParams {
  Geometry g;              // map_scale == 4, width % 4 == 0, height % 4 == 0
  int output_format;
  float A, B;              // gain factor f = 2^(gain * A + B) = 2^(logBoost * displayBoost / maxContentBoost) / displayBoost
  float e0;                // 2^kOLoExp
  uint32_t s_bias;         // 0x4B000000 << 3 (see uni_eval in apply_fast.cu)
  int need_mask;           // 1 when 10-bit codes can exceed 1023 (displayBoost < maxContentBoost): `& 0x3ff` wraps like the reference
  float Aq, Bq, q0;        // q-domain kernel: f' = 2^(gain255 * Aq + Bq) = f^(1/2.4), gain255 = 255 * gain; q0 = 2^kPLoExp
  int two_table;           // 1: the S+O table kernels (apply_fast_kernel) instead of the q-domain kernels
  int lin_table_ok;        // 1: q = Phi * f' stays inside the L table for every non-toe channel (host check)
  int cells;               // apply_q_kernel: map cells a lane walks per tile (set by launch_q / pick_cells)
  int any_scale;           // 1: apply_qs_kernel (any integer map scale factor, thread = 4x4 pixels) instead of the scale-4 kernels
  uint32_t scale_magic;    // ceil(2^32 / map_scale), 0 for scale 1 (set by launch_apply_fast_impl)
  int scale_special;       // 1 / 2: compile-time specialisation of apply_qs_kernel (map planes aligned to 4 / 2 bytes); 4: any
                           // scale factor that is a multiple of 4 (a 4x4 block lies inside one map cell); else 0
  int pipe;                // 1: apply_pipe_kernel (TMA-fed per-warp pipeline, apply_pipe.cu) instead of apply_q_kernel
  int pdl;                 // 1: launch with programmatic stream serialization (the kernels call griddepcontrol.*)
  // A channel clamped to c = 1 (ultrahdr.cpp:430-431 -> clampPixelFloat) inside a cell whose four map samples are all 255
  // (blown highlight under a saturated map: gain = the weight sum, 1 or 1 - 1..2 ulp) lands exactly on a code boundary of
  // the reference's float chain, finer than the q domain resolves; the host evaluates that chain exactly (capi.cu:
  // fill_sat_cell) per IDW table (0 interior, 1 NR, 2 NB, 3 C, gainmapmath.cpp:710-716) and pixel position [oy * 4 + ox]:
  float vsat[4][16];       // RGBA1010102: the reference's 10-bit code (before `& 0x3ff`) + 0.5
  float rsat[4][16];       // linear outputs: the reference's value f * srgbInvOetf(1) / displayBoost
  float w2[4][4];          // ShepardsIDW::mWeights of scale 2, [oy * 2 + ox][k]
  // Aq x interior IDW weights of corners 2..4 (they include the 1/255 of the map samples), [corner - 2][oy * 4 + ox]:
  // horizontally adjacent pixels are one 8-byte pair for FFMA2
  alignas(8) float wq[3][16];
  float w[16][4];          // A x interior IDW weights [oy * 4 + ox][k] (ShepardsIDW::mWeights, gainmapmath.h:184-228)
  const float* idw;        // all four tables (device), used by the cells on the right/bottom map border
  ApplyTables tab;
  int tiles_x, tiles_y, n_tiles;  // tiles of 32 x 8 map cells (128 x 32 pixels)
  Planes single;
  const Planes* batch;
};

struct ToneMapParams {
  Geometry g;
  Planes single;
  const Planes* batch;
};

struct ConvertYuvParams {
  Geometry g;       // in-place on the sdr_* planes (luma/chroma strides = sdr_*)
  float m[9];       // YUV->YUV matrix, gainmapmath.cpp:447-481
  Planes single;
  const Planes* batch;
};

struct SdrPackParams {   // row N3: yuv420 -> RGBA8888 (out = packed pixels, dst_luma_stride in pixels) and copy_yuv420
  Geometry g;            // (out/out_u/out_v = destination planes, dst_luma_stride / dst_chroma_stride in bytes)
  Planes single;
  const Planes* batch;
};

// row N4 (jpeg_idct.cu): one colour component of a JPEG in the coefficient domain -> its u8 plane
struct JpegIdctComponent {
  const int16_t* coeffs;      // blocks_h x blocks_w blocks of 64 quantised coefficients, natural (row-major) order
  const uint16_t* quant;      // 64 quantisation steps, natural order
  size_t blocks_w, blocks_h;  // geometry of the coefficient array (whole MCUs, as jpeg_read_coefficients stores it)
  uint8_t* plane;             // output samples
  size_t width, height, stride;  // visible samples and the plane's row stride in bytes
};
struct JpegIdctParams {
  int n_components;
  JpegIdctComponent comp[4];
};
cudaError_t launch_jpeg_idct_impl(const JpegIdctParams& p, cudaStream_t s);

// row N4 (jpeg_huff.cu): baseline JPEG entropy decoding on the device. The host parses the markers (jpeg_parse), removes the
// byte stuffing and uploads the entropy-coded segment; the device decodes it with one thread per kJpegSubseqBits-bit
// subsequence (self-synchronising parallel Huffman decoding) into the coefficient arrays jpeg_idct_kernel reads.
#ifndef UHDR_JPEG_SUBSEQ_BITS
#define UHDR_JPEG_SUBSEQ_BITS 1024
#endif
constexpr int kJpegSubseqBits = UHDR_JPEG_SUBSEQ_BITS;   // bits of entropy-coded data one thread owns (a multiple of 32)
constexpr int kJpegLookBits = 10;       // first-level Huffman look-up: codes of up to 10 bits in one shared-memory load
struct JpegHuffTable {                  // one DC or AC table (ITU-T T.81 Annex C / F.2.2.3), built by jpeg_parse
  // first level, indexed by the next kJpegLookBits bits: code length | size << 5 | run << 9 | (length + size) << 13 |
  // step << 18 for a code of up to kJpegLookBits bits (size = SSSS extra bits, run = RRRR, 0 in a DC table; step = what the
  // symbol adds to the coefficient index: run + 1, 16 for ZRL, 64 for EOB, 1 for DC); 0 = a longer code
  uint32_t fast[1 << kJpegLookBits];
  int32_t maxcode[18];                  // largest code of each length 1..16 (-1: none); [17] = sentinel
  int32_t valoff[18];                   // index into vals of the first code of each length minus that code
  uint8_t vals[256];
};
struct JpegDeviceTables {               // one upload per decode
  JpegHuffTable dc[2], ac[2];           // baseline: table ids 0 and 1
  uint16_t quant[4][64];                // natural (row-major) order
};
struct JpegScan {                       // passed by value to the kernels
  int n_comp, blocks_per_mcu;
  uint32_t mcus_x, mcus_y, total_blocks;
  uint8_t blk_comp[10], blk_dx[10], blk_dy[10];  // block i of an MCU: component, block offset inside the MCU
  uint8_t comp_h[4], comp_v[4], comp_dc[4], comp_ac[4];
  uint32_t dc_mask, ac_mask;            // bit i: DC / AC table id (0 or 1) of block i of an MCU
  uint32_t comp_blocks_w[4], comp_blocks_h[4];   // coefficient array geometry (whole MCUs)
  int16_t* coeffs[4];                   // device, zero-filled before the write pass
  const uint32_t* data;                 // device: the entropy-coded segment without byte stuffing, zero padded
  uint64_t total_bits;
  uint32_t n_subseq;
  // restart intervals (T.81 B.2.4.4, F.2.1.3.1): every interval starts at a subsequence boundary of `data` (the host pads),
  // in a known state (block 0 of MCU interval * restart_interval, DC predictions zero)
  uint32_t restart_interval;            // MCUs per interval, 0 = none (one interval)
  uint32_t n_intervals;
  const uint32_t* interval_of;          // device, per subsequence: its interval (nullptr without restart intervals)
  const uint32_t* interval_first;       // device, per interval: its first subsequence
  const JpegDeviceTables* tables;       // device
};
struct JpegHeader {                     // host-side result of jpeg_parse
  int width, height, n_comp;
  int comp_id[4], comp_h[4], comp_v[4], comp_tq[4], comp_td[4], comp_ta[4];
  int hmax, vmax, restart_interval;
  size_t scan_begin, scan_end;          // byte range of the entropy-coded segment inside the file
  JpegDeviceTables tables;
};
// status: 0 ok, 1 not a JPEG / corrupt header, 2 a JPEG this decoder does not handle (progressive, arithmetic, 12-bit,
// several scans); `why` receives a short reason
int jpeg_parse(const uint8_t* data, size_t len, JpegHeader* h, const char** why);
// copies the entropy-coded segment that starts at `begin` into dst without its byte stuffing (FF 00 -> FF) up to the
// first marker that is not RSTn, which it reports (0: the data ran out); returns the number of bytes. At every RSTn (at
// most max_intervals - 1 of them) the output is zero padded to a multiple of kJpegSubseqBits / 8 bytes and the byte offset
// of the next interval is appended to interval_start (which starts as {0}). dst capacity: end - begin + max_intervals *
// kJpegSubseqBits / 8.
size_t jpeg_destuff(const uint8_t* data, size_t begin, size_t end, uint8_t* dst, int* marker, uint32_t max_intervals,
                    uint32_t* interval_start, uint32_t* n_intervals);
// workspace per subsequence: 2 x u32 exit state (ping-pong), u32 block count, u32 scanned count, 2 x u8 dirty flags
size_t jpeg_huff_workspace_bytes(uint32_t n_subseq);
// the whole entropy decode on stream s; synchronises the stream while iterating to the fixed point. `iters` (optional)
// receives the number of synchronisation passes.
cudaError_t jpeg_huff_decode(const JpegScan& sc, void* workspace, cudaStream_t s, int* iters);

// *_vec_ok: true when the planes/strides of an image allow the 128-bit vector path (the caller
// ANDs it over a batch). launch_*_impl return the cudaError_t of the launch; `n` = images (grid.z).
bool generate_vec_ok(const Geometry& g, const Planes& pl);
bool apply_vec_ok(const Geometry& g, const Planes& pl);
bool tonemap_vec_ok(const Geometry& g, const Planes& pl);
bool convert_yuv_vec_ok(const Geometry& g, const Planes& pl);
cudaError_t launch_generate_impl(const GenerateParams& p, int n, bool vec, cudaStream_t s);
bool api0_fused_ok(const Geometry& g, const Planes& pl);
cudaError_t launch_api0_fused_impl(const GenerateParams& p, int n, cudaStream_t s);
cudaError_t launch_apply_impl(const ApplyParams& p, int n, bool vec, cudaStream_t s);
// fast path: scale 4, vector-aligned planes, gain factor bounded so that e + e0 < 2^kOHiExp
bool apply_fast_ok(const Geometry& g, const Planes& pl);
bool apply_qs_ok(const Geometry& g, const Planes& pl);
cudaError_t launch_apply_fast_impl(ApplyFastParams& p, int n, cudaStream_t s);
// scale 4 on the TMA-fed pipeline: additionally 16-byte aligned planes / strides and a map width % 16 == 0
bool apply_pipe_ok(const Geometry& g, const Planes& pl);
cudaError_t launch_apply_pipe_impl(const ApplyFastParams& p, int n, cudaStream_t s);
// host: fills the double-precision-built tables (tables.cu)
void build_apply_tables(float2* s, float2* s_hi, float2* o_hlg, float2* o_pq, float2* p_hlg, float2* p_pq, float2* l_pow, float* e255);
cudaError_t launch_tonemap_impl(const ToneMapParams& p, int n, bool vec, cudaStream_t s);
// row N4 (jpeg_enc.cu): baseline JPEG encode of raw 4:2:0 / grey planes (JpegEncoderHelper::compressImage)
struct JpegEncHuff {       // EHUFCO / EHUFSI of T.81 Annex C for one table
  uint16_t code[256];
  uint8_t len[256];
};
struct JpegEncTables {     // one upload per encode
  uint16_t quant[2][64];   // natural order (the FDCT kernel), [0] luma, [1] chroma
  uint8_t quant_zz[2][64]; // zig-zag order (the DQT segments)
  JpegEncHuff dc[2], ac[2];
};
struct JpegEncComponent {
  const uint8_t* plane;    // device
  size_t width, height, stride;  // visible samples; beyond them the encoder sees zero samples
  uint32_t blocks_w, blocks_h;   // whole MCUs
  const uint16_t* quant;   // device, natural order
  int16_t* coeffs;         // device: blocks_h x blocks_w records of 64 quantised coefficients in ZIG-ZAG order
};
struct JpegEncParams {
  int n_components;
  JpegEncComponent comp[3];
};
void jpeg_enc_tables(int quality, bool grey, JpegEncTables* t);
size_t jpeg_write_headers(uint8_t* out, size_t cap, int width, int height, bool grey, const JpegEncTables& t, const void* icc, size_t icc_bytes);
size_t jpeg_stuff(const uint8_t* src, size_t n, uint8_t* dst, size_t cap);
cudaError_t launch_jpeg_fdct_impl(const JpegEncParams& p, cudaStream_t s);
size_t jpeg_enc_workspace_bytes(uint32_t total_blocks);
cudaError_t jpeg_huff_encode(const JpegScan& sc, const JpegEncTables* tables_dev, void* workspace, uint32_t* total_bits_dev, cudaStream_t s,
                             uint32_t** offsets_out);
cudaError_t jpeg_huff_emit(const JpegScan& sc, const JpegEncTables* tables_dev, const uint32_t* offsets, uint32_t* out_words, cudaStream_t s);

// test hook (include/uhdr_b200_test.h): out[i] = fn(in[i]) with the device functions of uhdr_math.cuh / apply_q_common.cuh
cudaError_t launch_kat_impl(int fn, const float* in, float* out, size_t n, const ApplyTables& tab, const GainCode& gc, cudaStream_t s);
bool sdr_pack_vec_ok(const Geometry& g, const Planes& pl);
cudaError_t launch_sdr_pack_impl(const SdrPackParams& p, int n, bool vec, cudaStream_t s);
cudaError_t launch_copy_yuv420_impl(const SdrPackParams& p, int n, cudaStream_t s);
cudaError_t launch_convert_yuv_impl(const ConvertYuvParams& p, int n, bool vec, cudaStream_t s);

}  // namespace uhdr
