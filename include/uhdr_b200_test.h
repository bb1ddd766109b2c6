/*
 * uhdr_b200_test.h — TEST HOOKS of libuhdr_b200.so. Not part of the drop-in boundary (include/uhdr_b200.h): nothing a
 * libultrahdr caller needs is declared here, and the shim never calls these. They exist so that tests/ and bench.py can
 * (a) force a particular applyGainMap kernel family for A/B measurements and cross-checks, and (b) run the device-side
 * pixel math (csrc/uhdr_math.cuh, the shared-memory table evaluators) over an array of arguments, which is how the
 * known-answer tests of the reference's tests/gainmapmath_test.cpp are applied to the device functions one by one.
 */
#ifndef UHDR_B200_TEST_H
#define UHDR_B200_TEST_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* applyGainMap kernel selection FOR THE CALLING THREAD ONLY (thread-local; other threads keep AUTO):
 * AUTO      what a caller gets: the q-domain kernels (scale 4, scale 1 / 2, any integer scale) when the layout allows
 *           them, else the analytic kernel;
 * GENERIC   the analytic kernel (any layout, MUFU-evaluated transfer functions);
 * TWO_TABLE scale-4 inputs through the srgbInvOetf-table + OETF-table kernel (previous design, kept as a cross-check);
 * Q_PIPE    scale-4 inputs through the TMA-fed per-warp pipeline (apply_pipe.cu) where the layout allows 16-byte bulk
 *           copies: same arithmetic as AUTO's kernel, measured slower on coherent images, kept for A/B runs. */
enum { UHDRB200_APPLY_AUTO = 0, UHDRB200_APPLY_GENERIC = 1, UHDRB200_APPLY_TWO_TABLE = 2, UHDRB200_APPLY_Q_PIPE = 3 };
int uhdrb200_test_set_apply_mode(int mode);

/* Evaluates device function `fn` over n floats (device pointers, asynchronous on `stream`): out[i] = fn(in[i]).
 * Integer-valued functions return their result as a float (exact below 2^24). For the two-argument functions
 * `in` holds n pairs. Returns a status_t value. */
enum {
  UHDRB200_FN_SRGB_INV_OETF = 0, /* gainmapmath.cpp:149-155 */
  UHDRB200_FN_HLG_OETF = 1,      /* :259-267 */
  UHDRB200_FN_HLG_INV_OETF = 2,  /* :280-290 */
  UHDRB200_FN_PQ_OETF = 3,       /* :309-314 */
  UHDRB200_FN_PQ_INV_OETF = 4,   /* :330-342 */
  UHDRB200_FN_FLOAT_TO_HALF = 5, /* gainmapmath.h:136-147 */
  UHDRB200_FN_TO_10BIT = 6,      /* one channel of colorToRgba1010102, gainmapmath.cpp:722-727 */
  UHDRB200_FN_P_HLG = 7,         /* shared-memory table: 1023 * hlgOetf(q^2.4), argument q */
  UHDRB200_FN_P_PQ = 8,          /* shared-memory table: 1023 * pqOetf(q^2.4), argument q */
  UHDRB200_FN_L_POW = 9,         /* shared-memory table: q^2.4 = f * srgbInvOetf(c) above the toe, argument q in [2^-6, 2) */
  UHDRB200_FN_TOE_FIX = 10,      /* Phi -> srgbInvOetf(c)^(1/2.4) with c = 1.055 Phi - 0.055 (toe select, apply_q_common.cuh) */
  UHDRB200_FN_ENCODE_GAIN = 11,  /* pairs (y_sdr, y_hdr) -> encodeGain byte for HLG metadata (min 1, max 1000/203), :524-541 */
  UHDRB200_FN_COUNT = 12
};
int uhdrb200_test_eval(int fn, const float* in_dev, float* out_dev, size_t n, void* stream);

/* Host-only parts of the JPEG encoder (csrc/jpeg_enc.cu), callable without a GPU: everything uhdrb200_jpeg_encode_* writes in
 * front of the entropy-coded segment (SOI, JFIF APP0, optional ICC APP2, DQT, SOF0, DHT, SOS) for the given quality and
 * size, and the byte stuffing (FF -> FF 00) of an entropy-coded segment. Both return the bytes written, 0 when `cap` is
 * too small. */
size_t uhdrb200_test_jpeg_headers(int quality, int grey, int width, int height, const void* icc, size_t icc_bytes, void* out, size_t cap);
size_t uhdrb200_test_jpeg_stuff(const void* src, size_t n, void* out, size_t cap);

/* Host-only front end of the JPEG decoder (csrc/jpeg_huff.cu), callable without a GPU: the removal of the byte stuffing
 * from an entropy-coded segment src[0, n) as uhdrb200_jpeg_decode_* does it before the upload. FF 00 -> FF, fill bytes
 * (FF FF) dropped, every RSTn marker replaced by zero padding up to the next multiple of `uhdrb200_test_jpeg_subseq_bytes()`
 * with the new interval's offset appended to interval_start (at most max_intervals entries, the first is 0), any other
 * marker ends the segment and is returned in *marker (0 when the data ran out first). `out` needs n + max_intervals *
 * subseq_bytes + 64 bytes. Returns the bytes written. */
size_t uhdrb200_test_jpeg_destuff(const void* src, size_t n, void* out, int* marker, unsigned max_intervals, unsigned* interval_start,
                                  unsigned* n_intervals);
size_t uhdrb200_test_jpeg_subseq_bytes(void);

#ifdef __cplusplus
}
#endif
#endif /* UHDR_B200_TEST_H */
