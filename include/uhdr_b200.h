/*
 * uhdr_b200.h — C ABI of the B200-native gain-map hot path.
 *
 * Drop-in boundary for the three protected members of the reference's `ultrahdr::UltraHdr`
 * (reference lib/include/ultrahdr/ultrahdr.h:348-350 generateGainMap, :370-372 applyGainMap,
 * :381 toneMap; definitions lib/src/ultrahdr.cpp:185-358, :360-515, :517-558) plus the "next"
 * row N1 `JpegR::convertYuv` (lib/src/jpegr.cpp:1132-1206). Plain pointers and sizes only: no
 * C++ types, no torch types, no CUDA types (streams travel as `void*`).
 *
 * Every entry point returns the SAME integer values as the reference's `status_t`
 * (ultrahdr.h:91-120) and performs the reference's argument checks in the reference's order
 * before any CUDA call, so argument-validation behaviour is identical with or without a GPU.
 *
 * Memory kinds
 *   *_host entry points  : all `data` / `chroma_data` pointers are HOST pointers (pageable or
 *                          pinned). Synchronous, blocking, re-entrant — same contract as the
 *                          reference members. Copies host<->device inside the call.
 *   *_dev entry points   : all plane pointers are DEVICE pointers on the current CUDA device;
 *                          the call only enqueues work on `stream` (a cudaStream_t passed as
 *                          void*, NULL = legacy default stream) and returns without syncing.
 *   *_batch_dev          : the same for `n` independent images in ONE launch.
 */
#ifndef UHDR_B200_H
#define UHDR_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define UHDRB200_ABI_VERSION 1

/* ---- enums: numeric values identical to the reference (ultrahdr.h:36-75) ---- */
enum {
  UHDRB200_COLORGAMUT_UNSPECIFIED = -1,
  UHDRB200_COLORGAMUT_BT709 = 0,
  UHDRB200_COLORGAMUT_P3 = 1,
  UHDRB200_COLORGAMUT_BT2100 = 2
};
enum {
  UHDRB200_TF_UNSPECIFIED = -1,
  UHDRB200_TF_LINEAR = 0,
  UHDRB200_TF_HLG = 1,
  UHDRB200_TF_PQ = 2,
  UHDRB200_TF_SRGB = 3
};
enum {
  UHDRB200_OUTPUT_UNSPECIFIED = -1,
  UHDRB200_OUTPUT_SDR = 0,               /* RGBA8888; see uhdrb200_sdr_to_rgba8888_* (row N3) */
  UHDRB200_OUTPUT_HDR_LINEAR = 1,        /* RGBA F16, 8 B/px */
  UHDRB200_OUTPUT_HDR_PQ = 2,            /* RGBA1010102, 4 B/px */
  UHDRB200_OUTPUT_HDR_HLG = 3,           /* RGBA1010102, 4 B/px */
  UHDRB200_OUTPUT_HDR_LINEAR_RGB_10BIT = 4 /* three planar u16 planes R,G,B */
};
enum {
  UHDRB200_PIX_FMT_UNSPECIFIED = -1,
  UHDRB200_PIX_FMT_P010 = 0,
  UHDRB200_PIX_FMT_YUV420 = 1,
  UHDRB200_PIX_FMT_MONOCHROME = 2,
  UHDRB200_PIX_FMT_RGBA8888 = 3,
  UHDRB200_PIX_FMT_RGBAF16 = 4,
  UHDRB200_PIX_FMT_RGBA1010102 = 5
};

/* ---- status codes: numeric values identical to the reference status_t (ultrahdr.h:91-120) ---- */
enum {
  UHDRB200_NO_ERROR = 0,
  UHDRB200_UNKNOWN_ERROR = -1,
  UHDRB200_ERROR_BAD_PTR = -10001,
  UHDRB200_ERROR_UNSUPPORTED_WIDTH_HEIGHT = -10002,
  UHDRB200_ERROR_INVALID_COLORGAMUT = -10003,
  UHDRB200_ERROR_INVALID_STRIDE = -10004,
  UHDRB200_ERROR_INVALID_TRANS_FUNC = -10005,
  UHDRB200_ERROR_RESOLUTION_MISMATCH = -10006,
  UHDRB200_ERROR_INVALID_QUALITY_FACTOR = -10007,
  UHDRB200_ERROR_INVALID_OUTPUT_FORMAT = -10009,
  UHDRB200_ERROR_BAD_METADATA = -10010,
  UHDRB200_ERROR_DECODE = -20002, /* ERROR_ULTRAHDR_DECODE_ERROR: not a JPEG / corrupt header (row N4) */
  UHDRB200_ERROR_BUFFER_TOO_SMALL = -20004,
  UHDRB200_ERROR_UNSUPPORTED_MAP_SCALE_FACTOR = -20008,
  UHDRB200_ERROR_INSUFFICIENT_RESOURCE = -20009, /* also: no CUDA device / CUDA failure */
  UHDRB200_ERROR_UNSUPPORTED_FEATURE = -30000
};

/*
 * POD mirror of `ultrahdr_uncompressed_struct` (ultrahdr.h:152-180): same field order and types
 * (the reference enums are plain `int`-sized), so on LP64 the two layouts are identical and the
 * C++ shim can pass the reference struct straight through.
 * Strides are in PIXELS (P010 chroma stride in uint16 elements), exactly as in the reference.
 */
typedef struct uhdrb200_image {
  void* data;
  size_t width;
  size_t height;
  int32_t colorGamut;
  void* chroma_data;
  size_t luma_stride;
  size_t chroma_stride;
  int32_t pixelFormat;
} uhdrb200_image;

/*
 * POD flattening of `ultrahdr_metadata_struct` (ultrahdr.h:129-146). The reference member
 * `std::string version` becomes a fixed char array; applyGainMap compares it with "1.0"
 * (kGainMapVersion, ultrahdr.h:210) like ultrahdr.cpp:369-372.
 */
typedef struct uhdrb200_metadata {
  char version[8];
  float maxContentBoost;
  float minContentBoost;
  float gamma;
  float offsetSdr;
  float offsetHdr;
  float hdrCapacityMin;
  float hdrCapacityMax;
} uhdrb200_metadata;

/* ---- numeric mode ---- */
enum {
  /* default: follows the reference AS BUILT (analytic transfer functions; the USE_*_LUT macros
   * of jpegr.cpp:33-38 are invisible to ultrahdr.cpp). */
  UHDRB200_MODE_ANALYTIC = 0,
  /* the reference's -DUSE_SRGB_INVOETF_LUT=1 ... -DUSE_APPLY_GAIN_LUT=1 build (what jpegr.cpp:33-38 intends): nearest-entry
   * tables for srgbInvOetf (1024), hlg/pqInvOetf (4096), hlg/pqOetf (65 536) and the gain factor (1024), gainmapmath.cpp:
   * 21-64, gainmapmath.h:150-183. In this mode the kernels follow the reference's float arithmetic operation by operation
   * (tables built on the host by the same g++ / glibc expressions), so the output is byte-identical to that build. */
  UHDRB200_MODE_LUT = 1
};
/* Numeric mode of the CALLING THREAD for its following generate / apply / API-0 calls (toneMap, convertYuv and the packers
 * are integer or mode-free). Default UHDRB200_MODE_ANALYTIC. Returns ERROR_UNSUPPORTED_FEATURE for an unknown mode. */
int uhdrb200_set_numeric_mode(int mode);
int uhdrb200_get_numeric_mode(void);

/* ---- library / device management ---- */
int uhdrb200_abi_version(void);
/* number of CUDA devices visible, 0 if none / no driver. Never fails. */
int uhdrb200_device_count(void);
/* binds the calling thread to `device` and creates its context lazily. */
int uhdrb200_init(int device);
void uhdrb200_shutdown(void);
/* last CUDA error string seen by the calling thread ("" if none). */
const char* uhdrb200_last_error(void);
/* pinned host memory helpers (full-speed DMA for the *_host entry points). */
void* uhdrb200_host_alloc(size_t bytes);
void uhdrb200_host_free(void* p);
/* frees a gain map returned by uhdrb200_generate_gainmap_host (malloc-compatible). */
void uhdrb200_free(void* p);
/* number of kernel launches issued by this library in the calling process (monotonic). */
uint64_t uhdrb200_launch_count(void);

/* How the synchronous *_host entry points wait for their stream (process-wide). The reference joins its row-block
 * worker threads before returning (ultrahdr.cpp:354, 513); here the join is a stream wait. SPIN (default,
 * cudaStreamSynchronize) has the lowest latency; BLOCKING sleeps on a cudaEventBlockingSync event, which is what a
 * host with more calling threads than cores wants (several processes x several threads per GPU). The environment
 * variable UHDRB200_BLOCKING_SYNC=1 selects BLOCKING when this function has not been called. */
enum { UHDRB200_SYNC_SPIN = 0, UHDRB200_SYNC_BLOCKING = 1 };
int uhdrb200_set_sync_mode(int mode);

/* ---- A1: UltraHdr::generateGainMap (ultrahdr.cpp:185-358) ---- */
/* Host planes. On success dest->data is a malloc'ed map_w*map_h buffer owned by the caller
 * (release with uhdrb200_free); metadata is fully overwritten (ultrahdr.cpp:250-257). */
int uhdrb200_generate_gainmap_host(const uhdrb200_image* yuv420, const uhdrb200_image* p010,
                                   int hdr_tf, uhdrb200_metadata* metadata, uhdrb200_image* dest,
                                   int sdr_is_601);
/* Same, but writes the map into caller memory `map_buf` (>= (w/4)*(h/4) bytes); used by the C++
 * shim, which must hand the reference callers a new[]-allocated buffer (jpegr.cpp:207-208). */
int uhdrb200_generate_gainmap_host_into(const uhdrb200_image* yuv420, const uhdrb200_image* p010,
                                        int hdr_tf, uhdrb200_metadata* metadata,
                                        uhdrb200_image* dest, int sdr_is_601, void* map_buf,
                                        size_t map_buf_bytes);
/* Device planes; dest->data must point to a device buffer of (w/4)*(h/4) bytes. Async. */
int uhdrb200_generate_gainmap_dev(const uhdrb200_image* yuv420, const uhdrb200_image* p010,
                                  int hdr_tf, uhdrb200_metadata* metadata, uhdrb200_image* dest,
                                  int sdr_is_601, void* stream);

/* ---- A2: UltraHdr::applyGainMap (ultrahdr.cpp:360-515) ---- */
int uhdrb200_apply_gainmap_host(const uhdrb200_image* yuv420, const uhdrb200_image* gainmap,
                                const uhdrb200_metadata* metadata, int output_format,
                                float max_display_boost, uhdrb200_image* dest);
int uhdrb200_apply_gainmap_dev(const uhdrb200_image* yuv420, const uhdrb200_image* gainmap,
                               const uhdrb200_metadata* metadata, int output_format,
                               float max_display_boost, uhdrb200_image* dest, void* stream);

/* ---- A3: UltraHdr::toneMap (ultrahdr.cpp:517-558) ---- */
int uhdrb200_tonemap_host(const uhdrb200_image* src, uhdrb200_image* dest);
int uhdrb200_tonemap_dev(const uhdrb200_image* src, uhdrb200_image* dest, void* stream);

/* ---- batched device-resident variants: n independent images, ONE launch each ----
 * Arrays of n descriptors (host memory, read during the call); all images of a batch must share
 * width/height/strides layout class, gamuts, transfer function / output format. */
int uhdrb200_generate_gainmap_batch_dev(int n, const uhdrb200_image* yuv420, const uhdrb200_image* p010,
                                        int hdr_tf, uhdrb200_metadata* metadata,
                                        uhdrb200_image* dest, int sdr_is_601, void* stream);
/* No counterpart in the reference: added for callers that must not allocate
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc). Same arguments, results and
 * status codes as uhdrb200_generate_gainmap_batch_dev, but the call allocates NOTHING: no cudaMalloc / cudaFree, no
 * stream or event creation, no heap. The per-image plane-pointer array the kernel reads lives in a file-scope
 * `static __device__` array of UHDRB200_NOALLOC_MAX_IMAGES entries (18 KB, held by every GPU that loads the library until
 * exit) whose device address is taken once per device with cudaGetSymbolAddress. It runs on the CALLER's current device
 * and `stream` (the thread's uhdrb200_init binding is not used). n > UHDRB200_NOALLOC_MAX_IMAGES returns
 * UHDRB200_ERROR_UNSUPPORTED_FEATURE with the reason in uhdrb200_last_error(). All calls share the one array: calls on
 * the same stream are ordered, calls that overlap on the GPU (different streams, different host threads) RACE on it and
 * must be serialised by the caller. The array keeps the last call's entries; each call overwrites the n it uses. */
#define UHDRB200_NOALLOC_MAX_IMAGES 256
int uhdrb200_generate_gainmap_batch_dev_noalloc(int n, const uhdrb200_image* yuv420, const uhdrb200_image* p010,
                                                int hdr_tf, uhdrb200_metadata* metadata, uhdrb200_image* dest,
                                                int sdr_is_601, void* stream);
int uhdrb200_apply_gainmap_batch_dev(int n, const uhdrb200_image* yuv420, const uhdrb200_image* gainmap,
                                     const uhdrb200_metadata* metadata, int output_format,
                                     float max_display_boost, uhdrb200_image* dest, void* stream);
int uhdrb200_tonemap_batch_dev(int n, const uhdrb200_image* src, uhdrb200_image* dest,
                               void* stream);

/* ---- N2: API-0 front half, JpegR::encodeJPEGR(p010, ...) lib/src/jpegr.cpp:167-225: toneMap(p010 -> sdr_dest),
 * generateGainMap(sdr_dest, p010, hdr_tf, metadata, map_dest) and, when sdr_to_601 != 0 and the gamut is not P3,
 * convertYuv(sdr_dest, gamut -> BT.601) (jpegr.cpp:223-225; the gain map is computed from the un-converted
 * image, as there). Same results, byte for byte, as calling the separate entry points one after the other;
 * when width and height are multiples of 4 and the planes are vector-aligned it is ONE kernel that reads the
 * P010 image once (4.5625 B/px instead of 9.06 + 3). sdr_dest: caller-owned YUV420 descriptor (data,
 * chroma_data, strides, dims set), colorGamut is written (= the P010 gamut, convertYuv leaves it alone).
 * host_into: the map goes to map_buf (>= (w/4)*(h/4) bytes); dev: map_dest->data is a device buffer. ---- */
int uhdrb200_tonemap_generate_gainmap_host_into(const uhdrb200_image* p010, int hdr_tf, int sdr_to_601,
                                                uhdrb200_image* sdr_dest, uhdrb200_metadata* metadata,
                                                uhdrb200_image* map_dest, void* map_buf, size_t map_buf_bytes);
int uhdrb200_tonemap_generate_gainmap_dev(const uhdrb200_image* p010, int hdr_tf, int sdr_to_601,
                                          uhdrb200_image* sdr_dest, uhdrb200_metadata* metadata,
                                          uhdrb200_image* map_dest, void* stream);
int uhdrb200_tonemap_generate_gainmap_batch_dev(int n, const uhdrb200_image* p010, int hdr_tf, int sdr_to_601,
                                                uhdrb200_image* sdr_dest, uhdrb200_metadata* metadata,
                                                uhdrb200_image* map_dest, void* stream);

/* ---- N1: JpegR::convertYuv (jpegr.cpp:1132-1206) + transformYuv420 (gainmapmath.cpp:483-520):
 * in-place YUV420 re-encode between BT.709 / BT.601(P3) / BT.2100 matrices. ---- */
int uhdrb200_convert_yuv_host(uhdrb200_image* image, int src_encoding, int dest_encoding);
int uhdrb200_convert_yuv_dev(uhdrb200_image* image, int src_encoding, int dest_encoding,
                             void* stream);

/* ---- N3: the SDR output pack and the copy-with-restride around the path ----
 * yuv420_to_rgba8888: full-range BT.601 YUV420 -> RGBA8888 (R in bits 0-7, alpha 255), bit-exact
 * with the reference's own routine UltraHdrAppInput::convertYuv420ToRGBImage
 * (examples/ultrahdr_app.cpp:682-734); this is the device-side stand-in for what
 * ULTRAHDR_OUTPUT_SDR obtains from libjpeg (jpegr.cpp:770-788; applyGainMap itself has no SDR
 * case, ultrahdr.cpp:491-493). dest->data is caller-owned, width*height*4 bytes
 * (dest->luma_stride >= width selects a row stride in pixels, 0 = width); the call sets
 * dest->width/height/colorGamut. Even width and height.
 * copy_yuv420: the luma/chroma copy into other strides of jpegr.cpp:315-359 and
 * ultrahdr_api.cpp:437-479: dest->data/chroma_data/luma_stride/chroma_stride are pre-filled by the
 * caller, rows are copied width (width/2) bytes at a time and the stride tails are zero-filled. */
int uhdrb200_yuv420_to_rgba8888_host(const uhdrb200_image* yuv420, uhdrb200_image* dest);
int uhdrb200_yuv420_to_rgba8888_dev(const uhdrb200_image* yuv420, uhdrb200_image* dest, void* stream);
int uhdrb200_yuv420_to_rgba8888_batch_dev(int n, const uhdrb200_image* yuv420, uhdrb200_image* dest,
                                          void* stream);
int uhdrb200_copy_yuv420_host(const uhdrb200_image* src, uhdrb200_image* dest);
int uhdrb200_copy_yuv420_dev(const uhdrb200_image* src, uhdrb200_image* dest, void* stream);

/* ---- N4 (first slice): the coefficient-domain back half of the JPEG decode that feeds applyGainMap ----
 * JpegDecoderHelper::decompressImage with raw_data_out = TRUE and dct_method = JDCT_ISLOW
 * (lib/src/jpegdecoderhelper.cpp:291-307, 380-435, called from jpegr.cpp:693 for the primary image and :732 for the gain
 * map): dequantise + 8x8 integer inverse DCT (IJG "islow", jidctint.c) + range-limit, into the planar Y | U | V / grey
 * planes. Entropy decoding stays on the host: `coeffs` is what libjpeg's jpeg_read_coefficients() / access_virt_barray
 * hands out per component (blocks in raster order, 64 int16 each in natural order, whole MCUs), `quant` the component's
 * 64 quantisation steps in natural order (quant_tbl_ptrs[...]->quantval). Rows beyond `height` and columns beyond `width`
 * (the MCU padding) are not written. libjpeg-turbo's source is not part of the reference tree: parity is pinned on its
 * OUTPUT for the reference's JPEG fixtures (tests/test_jpeg_idct.py). Up to 4 components per call (_host: up to 3);
 * _dev: every pointer is device memory, asynchronous on `stream`. */
typedef struct uhdrb200_jpeg_component {
  const int16_t* coeffs;
  const uint16_t* quant;
  size_t blocks_w, blocks_h;
  void* plane;
  size_t width, height, stride;
} uhdrb200_jpeg_component;
int uhdrb200_jpeg_idct_host(int n_components, const uhdrb200_jpeg_component* comps);
int uhdrb200_jpeg_idct_dev(int n_components, const uhdrb200_jpeg_component* comps, void* stream);

/* ---- N4 (second slice): the whole baseline-JPEG decode, entropy decoding included, on the device ----
 * JpegDecoderHelper::decompressImage(image, length, DECODE_TO_YCBCR) + getDecompressedImagePtr / Width / Height
 * (lib/src/jpegdecoderhelper.cpp:190-320, 380-470; jpegr.cpp:693 primary image, :732 gain map): a 4:2:0 YCbCr JPEG becomes
 * packed planar Y | U | V (strides width and width / 2), a greyscale JPEG one plane; like the reference, other samplings
 * and images above 8192 x 8192 are refused. Handled on the device: baseline / sequential Huffman, 8-bit, one interleaved
 * scan, with or without restart intervals (libjpeg-turbo and JpegEncoderHelper write none by default). Progressive,
 * arithmetic-coded, 12-bit and multi-scan streams return UHDRB200_ERROR_UNSUPPORTED_FEATURE with the reason in
 * uhdrb200_last_error() so the caller can fall back to libjpeg; a stream that is not a JPEG returns UHDRB200_ERROR_DECODE.
 * The host parses the markers and removes the byte stuffing; only the compressed bytes are copied to the GPU, where the
 * Huffman stream is decoded in parallel (csrc/jpeg_huff.cu) and transformed by the N4 IDCT kernel, ISLOW like the reference
 * (jpegdecoderhelper.cpp:305): the planes are byte-identical to libjpeg-turbo's.
 *   uhdrb200_jpeg_get_info     getCompressedImageParameters (:339): size and UHDRB200_PIX_FMT_YUV420 / _MONOCHROME, no GPU work.
 *   uhdrb200_jpeg_decode_host  `jpeg` and dest->data are host memory; dest->data holds width * height * 3 / 2 (grey:
 *                              width * height) bytes; width, height, strides, chroma_data and pixelFormat are filled in.
 *   uhdrb200_jpeg_decode_dev   `jpeg` is host memory, dest->data DEVICE memory of the same size on the calling thread's
 *                              uhdrb200_init device: the planes stay on the GPU for uhdrb200_apply_gainmap_dev. Runs on
 *                              `stream` (0: the default stream) and returns when the planes are complete.
 * Both are synchronous and use scratch buffers owned by the calling thread. */
int uhdrb200_jpeg_get_info(const void* jpeg, size_t bytes, size_t* width, size_t* height, int* pixel_format);
int uhdrb200_jpeg_decode_host(const void* jpeg, size_t bytes, uhdrb200_image* dest);
int uhdrb200_jpeg_decode_dev(const void* jpeg, size_t bytes, uhdrb200_image* dest, void* stream);

/* ---- N4 (third slice): the JPEG encode of the SDR image and of the gain map on the device ----
 * JpegEncoderHelper::compressImage(yBuffer, uvBuffer, width, height, lumaStride, chromaStride, quality, iccBuffer, iccSize)
 * + getCompressedImagePtr / Size (lib/src/jpegencoderhelper.cpp:35-58, 86-232; called from jpegr.cpp:234-239 for the SDR
 * image and :1037-1050 for the gain map): raw planar 4:2:0 (Y, then U rows and V rows at chroma_stride * (height / 2), as
 * uhdrb200_image describes a YUV420 image) or one grey plane (chroma_data == NULL or pixelFormat MONOCHROME) -> a baseline
 * JFIF file with the settings the reference gives libjpeg: jpeg_set_quality(quality, TRUE), JDCT_ISLOW, 2x2 / 1x1 / 1x1
 * sampling, the typical Huffman tables of T.81 Annex K, no restart intervals, samples beyond the image inside the last MCUs
 * zero, the caller's ICC payload as one APP2 segment. Forward DCT, quantisation and Huffman coding run on the GPU
 * (csrc/jpeg_enc.cu); the host writes the markers and the byte stuffing. The file is BYTE-IDENTICAL to the one
 * libjpeg-turbo writes from the same samples (tests/test_gpu_jpeg_encode.py).
 *   _host: image planes in host memory.   _dev: image planes in DEVICE memory on the calling thread's uhdrb200_init device
 *   (what the fused API-0 kernel and generateGainMap leave there); `stream` orders the encode after the producer.
 * `out` is host memory of `out_capacity` bytes in both; *out_bytes receives the file size. Synchronous.
 * UHDRB200_ERROR_BUFFER_TOO_SMALL when the file does not fit (width * height * 3 always suffices for quality <= 100). */
int uhdrb200_jpeg_encode_host(const uhdrb200_image* image, int quality, const void* icc, size_t icc_bytes, void* out,
                              size_t out_capacity, size_t* out_bytes);
int uhdrb200_jpeg_encode_dev(const uhdrb200_image* image, int quality, const void* icc, size_t icc_bytes, void* out,
                             size_t out_capacity, size_t* out_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* UHDR_B200_H */
