// jpeg_huff.cu — row N4 (second slice): the entropy-decoding front half of the reference's JPEG decode,
// JpegDecoderHelper::decompressImage (lib/src/jpegdecoderhelper.cpp:190-320: jpeg_read_header, the 4:2:0 / greyscale
// checks, jpeg_start_decompress + jpeg_read_raw_data), on the device: a baseline (SOF0, Huffman, 8-bit, one interleaved
// scan) JPEG's entropy-coded segment is decoded into the per-component coefficient arrays that jpeg_idct.cu turns into
// planes. Only the compressed bytes cross PCIe (about a tenth of the decoded planes).
//
// libjpeg-turbo's source is NOT in /root/reference (fetched at configure time, CMakeLists.txt:254-256): the format is
// restated from ITU-T T.81 (Annex B markers, Annex C code tables, Annex F.2.2 decoding procedures); parity is anchored on
// libjpeg-turbo's own output for the reference's JPEG fixtures and for generated images (tests/test_gpu_jpeg_decode.py).
//
// Parallel decoding (a Huffman stream has no random access): the segment is cut into subsequences of kJpegSubseqBits
// bits, one thread each. Huffman codes self-synchronise: a decoder started at a wrong bit position falls into step with
// the true code boundaries after a few symbols with high probability. So
//   1. every thread decodes its subsequence from a GUESSED state (block start at its first bit) and records the state in
//      which it leaves (bit overshoot into the next subsequence, block-in-MCU, coefficient index);
//   2. repeatedly, every thread whose predecessor's exit state changed re-decodes its subsequence from that exit state;
//      thread 0's entry state is exact, so exactness spreads forward, and through self-synchronisation the fixed point is
//      reached after a handful of passes (each pass re-decodes only what changed); the passes also count the blocks
//      each subsequence completes;
//   3. an exclusive scan of the block counts gives every subsequence the index of the block it starts in;
//   4. a last pass decodes once more from the exact entry states and writes the coefficients (DC as differences);
//   5. a scan per component in decoding order turns the DC differences into DC values (T.81 F.2.2.1).
// (After Weissenberger and Schmidt, "Massively Parallel Huffman Decoding on GPUs", ICPP 2018, and their JPEG follow-up.)
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include <emmintrin.h>

#include "uhdr_kernels.cuh"

namespace uhdr {

// ================================================================ host: markers, tables, byte stuffing
namespace {

const uint8_t kZigZag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                             41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                             30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

// T.81 Annex C: canonical codes in order of increasing length; F.2.2.3: MAXCODE / VALPTR for the bit-serial decode
// length | size << 5 | run << 9 | (length + size) << 13 | step << 18: `step` is what the symbol adds to the coefficient index
// (run + 1; 16 for ZRL; 64 for EOB, which ends the block; 1 for a DC symbol)
uint32_t fast_entry(int len, int sym, bool ac) {
  const int s = sym & 15, r = ac ? sym >> 4 : 0;
  const int step = !ac ? 1 : (s ? r + 1 : (r == 15 ? 16 : 64));
  return (uint32_t)len | ((uint32_t)s << 5) | ((uint32_t)r << 9) | ((uint32_t)(len + s) << 13) | ((uint32_t)step << 18);
}
bool build_huff(const uint8_t* bits, const uint8_t* vals, int nvals, bool ac, JpegHuffTable* t) {
  memset(t, 0, sizeof *t);
  memcpy(t->vals, vals, (size_t)nvals);
  int code = 0, k = 0;
  for (int l = 1; l <= 16; l++) {
    t->valoff[l] = k - code;
    if (bits[l - 1]) {
      if (code + bits[l - 1] > (1 << l)) return false;  // over-subscribed
      for (int i = 0; i < bits[l - 1]; i++, k++, code++) {
        if (l <= kJpegLookBits) {
          const int lo = code << (kJpegLookBits - l), n = 1 << (kJpegLookBits - l);
          for (int j = 0; j < n; j++) t->fast[lo + j] = fast_entry(l, vals[k], ac);
        }
      }
      t->maxcode[l] = code - 1;
    } else {
      t->maxcode[l] = -1;
    }
    code <<= 1;
  }
  t->maxcode[17] = 0x7fffffff;
  return k == nvals;
}

int be16(const uint8_t* p) { return (p[0] << 8) | p[1]; }

}  // namespace

int jpeg_parse(const uint8_t* d, size_t len, JpegHeader* h, const char** why) {
  static const char* dummy;
  if (!why) why = &dummy;
  memset(h, 0, sizeof *h);
  *why = "not a JPEG";
  if (len < 4 || d[0] != 0xFF || d[1] != 0xD8) return 1;
  size_t pos = 2;
  bool have_sof = false, have_q[4] = {false, false, false, false}, have_dc[4] = {false, false, false, false},
       have_ac[4] = {false, false, false, false};
  for (;;) {
    *why = "truncated before the scan";
    if (pos + 4 > len) return 1;
    if (d[pos] != 0xFF) { *why = "marker expected"; return 1; }
    const int m = d[pos + 1];
    if (m == 0xFF) { pos++; continue; }  // fill byte
    pos += 2;
    if (m == 0xD8 || m == 0x01 || (m >= 0xD0 && m <= 0xD7)) continue;
    if (m == 0xD9) { *why = "no scan before EOI"; return 1; }
    const size_t ln = (size_t)be16(d + pos);
    if (ln < 2 || pos + ln > len) { *why = "bad segment length"; return 1; }
    const uint8_t* seg = d + pos + 2;
    const size_t n = ln - 2;
    switch (m) {
      case 0xDB:  // DQT, B.2.4.1
        for (size_t i = 0; i < n;) {
          const int pq = seg[i] >> 4, tq = seg[i] & 15;
          i++;
          if (tq > 3 || i + (pq ? 128 : 64) > n) { *why = "bad DQT"; return 1; }
          for (int k = 0; k < 64; k++) h->tables.quant[tq][kZigZag[k]] = (uint16_t)(pq ? be16(seg + i + 2 * k) : seg[i + k]);
          i += pq ? 128 : 64;
          have_q[tq] = true;
        }
        break;
      case 0xC0:  // SOF0 baseline; SOF1 (extended sequential Huffman) shares the syntax when 8-bit
      case 0xC1:
        if (n < 6 || have_sof) { *why = "bad SOF"; return 1; }
        if (seg[0] != 8) { *why = "sample precision other than 8 bits"; return 2; }
        h->height = be16(seg + 1);
        h->width = be16(seg + 3);
        h->n_comp = seg[5];
        if (h->width == 0 || h->height == 0) { *why = "zero image dimension (DNL is not handled)"; return 2; }
        if ((h->n_comp != 1 && h->n_comp != 3) || n < 6 + 3 * (size_t)h->n_comp) { *why = "component count other than 1 or 3"; return 2; }
        for (int c = 0; c < h->n_comp; c++) {
          h->comp_id[c] = seg[6 + 3 * c];
          h->comp_h[c] = seg[7 + 3 * c] >> 4;
          h->comp_v[c] = seg[7 + 3 * c] & 15;
          h->comp_tq[c] = seg[8 + 3 * c];
          if (h->comp_h[c] < 1 || h->comp_h[c] > 2 || h->comp_v[c] < 1 || h->comp_v[c] > 2 || h->comp_tq[c] > 3) {
            *why = "sampling factor above 2";
            return 2;
          }
          h->hmax = h->comp_h[c] > h->hmax ? h->comp_h[c] : h->hmax;
          h->vmax = h->comp_v[c] > h->vmax ? h->comp_v[c] : h->vmax;
        }
        have_sof = true;
        break;
      case 0xC2: case 0xC3: case 0xC5: case 0xC6: case 0xC7: case 0xC9: case 0xCA: case 0xCB: case 0xCD: case 0xCE: case 0xCF:
        *why = "progressive, lossless or arithmetic-coded JPEG";
        return 2;
      case 0xC4:  // DHT, B.2.4.2
        for (size_t i = 0; i < n;) {
          if (i + 17 > n) { *why = "bad DHT"; return 1; }
          const int tc = seg[i] >> 4, th = seg[i] & 15;
          int cnt = 0;
          for (int k = 0; k < 16; k++) cnt += seg[i + 1 + k];
          if (tc > 1 || cnt > 256 || i + 17 + (size_t)cnt > n) { *why = "bad DHT"; return 1; }
          if (th > 1) { *why = "Huffman table id above 1 (not baseline)"; return 2; }
          if (!build_huff(seg + i + 1, seg + i + 17, cnt, tc != 0, tc ? &h->tables.ac[th] : &h->tables.dc[th])) { *why = "invalid Huffman table"; return 1; }
          (tc ? have_ac : have_dc)[th] = true;
          i += 17 + (size_t)cnt;
        }
        break;
      case 0xDD:  // DRI
        if (n < 2) { *why = "bad DRI"; return 1; }
        h->restart_interval = be16(seg);
        break;
      case 0xDA: {  // SOS, B.2.3
        if (!have_sof) { *why = "SOS before SOF"; return 1; }
        if (n < 1 || n < 1 + 2 * (size_t)seg[0] + 3) { *why = "bad SOS"; return 1; }
        if (seg[0] != h->n_comp) { *why = "non-interleaved scans (one scan per component)"; return 2; }
        for (int k = 0; k < h->n_comp; k++) {
          int c = -1;
          for (int j = 0; j < h->n_comp; j++)
            if (h->comp_id[j] == seg[1 + 2 * k]) c = j;
          if (c != k) { *why = "scan components out of frame order"; return 2; }
          h->comp_td[c] = seg[2 + 2 * k] >> 4;
          h->comp_ta[c] = seg[2 + 2 * k] & 15;
          if (h->comp_td[c] > 1 || h->comp_ta[c] > 1 || !have_dc[h->comp_td[c]] || !have_ac[h->comp_ta[c]] || !have_q[h->comp_tq[c]]) {
            *why = "scan refers to a table that was not defined";
            return 1;
          }
        }
        h->scan_begin = pos + ln;
        h->scan_end = len;  // jpeg_destuff finds the marker that ends the segment while it copies
        *why = "";
        return 0;
      }
      default: break;  // APPn, COM, ...
    }
    pos += ln;
  }
}

size_t jpeg_destuff(const uint8_t* d, size_t begin, size_t end, uint8_t* dst, int* marker, uint32_t max_intervals,
                    uint32_t* interval_start, uint32_t* n_intervals) {
  constexpr size_t kAlign = kJpegSubseqBits / 8;
  size_t o = 0, i = begin;
  *marker = 0;
  *n_intervals = 1;
  interval_start[0] = 0;
  const __m128i ff = _mm_set1_epi8((char)0xFF);
  while (i < end) {
    // 16 bytes at a time while no FF is among them (an FF comes along every ~256 bytes of entropy-coded data; memchr +
    // memcpy per run cost ~50 ns each, three times the copy itself)
    while (i + 16 <= end) {
      const __m128i v = _mm_loadu_si128(reinterpret_cast<const __m128i*>(d + i));
      const int mask = _mm_movemask_epi8(_mm_cmpeq_epi8(v, ff));
      _mm_storeu_si128(reinterpret_cast<__m128i*>(dst + o), v);  // dst has the slack (capacity covers end - begin + padding)
      if (mask) {
        const int k = __builtin_ctz((unsigned)mask);  // bytes before the first FF
        o += (size_t)k;
        i += (size_t)k;
        break;
      }
      o += 16;
      i += 16;
    }
    while (i < end && d[i] != 0xFF && i + 16 > end) dst[o++] = d[i++];  // the last few bytes
    if (i >= end) break;
    if (d[i] != 0xFF) continue;
    const int m = i + 1 < end ? d[i + 1] : 0xD9;  // a lone FF at the very end: treat as EOI
    if (m == 0x00) {         // stuffed byte: FF 00 -> FF
      dst[o++] = 0xFF;
      i += 2;
    } else if (m == 0xFF) {  // fill byte before a marker
      i++;
    } else if (m >= 0xD0 && m <= 0xD7 && *n_intervals < max_intervals) {  // RSTn: the next interval starts byte aligned
      const size_t padded = (o + kAlign - 1) / kAlign * kAlign;
      memset(dst + o, 0, padded - o);
      o = padded;
      interval_start[(*n_intervals)++] = (uint32_t)o;
      i += 2;
    } else {                 // EOI (or any other marker): the entropy-coded segment ends here
      *marker = m;
      break;
    }
  }
  return o;
}

// ================================================================ device: the decoder
namespace {

#ifndef UHDR_JPEG_HYP
#define UHDR_JPEG_HYP 1  // phase hypotheses for interleaved scans (below); 0: guessed start + fixed-point rounds only
#endif
constexpr int kHuffThreads = 128;  // subsequences per CTA
#ifndef UHDR_JPEG_WARM
#define UHDR_JPEG_WARM 32
#endif
constexpr int kHuffWarm = UHDR_JPEG_WARM;  // subsequences a CTA decodes in front of its own in the first round (a multiple of 32)
__constant__ uint8_t c_zigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                    41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                    30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};  // T.81 Figure A.6

// A decoder state at a subsequence boundary, packed into 32 bits: how many bits the last symbol reached into the next
// subsequence (a symbol spans at most 16 + 15 bits), the block inside the MCU and the coefficient index inside the block.
__device__ __forceinline__ uint32_t pack_state(uint32_t over, uint32_t c, uint32_t k) { return (over << 16) | (c << 8) | k; }

struct HuffSmem {
  JpegHuffTable dc[2], ac[2];
};

__device__ __forceinline__ void load_tables(HuffSmem* sm, const JpegDeviceTables* t) {
  const uint32_t* src = reinterpret_cast<const uint32_t*>(t);  // dc[2], ac[2] are the first members
  uint32_t* dst = reinterpret_cast<uint32_t*>(sm);
  for (uint32_t i = threadIdx.x; i < sizeof(HuffSmem) / 4; i += blockDim.x) dst[i] = __ldg(src + i);
  __syncthreads();
}

// The stream as a 64-bit window kept in registers (MSB first): `buf` holds at least 32 valid bits before every symbol
// (a symbol spans at most 16 + 15), the word after the window is already in flight (`next`) when the window is topped up,
// so no load sits on the symbol-to-symbol dependency chain. The segment is zero padded: reads past its end are safe.
struct BitWindow {
  const uint32_t* __restrict__ ptr;  // the word after `next_raw`
  uint32_t hi, lo;    // the window, MSB first: hi = the next 32 bits, lo = the bits after them (`avail` - 32 valid)
  uint32_t next_raw;  // the following word as loaded (little endian): swapped only when it enters the window, so that the
                      // load issued at one top-up is not waited for before the next
  int avail;
  __device__ __forceinline__ static uint32_t be(uint32_t v) { return __byte_perm(v, 0, 0x0123); }
  __device__ __forceinline__ void open(const uint32_t* d, uint64_t p) {
    const uint32_t sh = (uint32_t)p & 31u;
    ptr = d + (p >> 5);
    const uint32_t w0 = be(__ldg(ptr)), w1 = be(__ldg(ptr + 1));
    next_raw = __ldg(ptr + 2);
    ptr += 3;
    hi = __funnelshift_l(w1, w0, sh);
    lo = w1 << sh;
    avail = 64 - (int)sh;
  }
  __device__ __forceinline__ uint32_t peek() const { return hi; }
  __device__ __forceinline__ void skip(uint32_t n) {  // n <= 31
    hi = __funnelshift_l(lo, hi, n);
    lo <<= n;
    avail -= (int)n;
    if (avail <= 32) {  // lo holds avail - 32 <= 0 valid bits: hi is short of 32 - avail bits at its low end
      const uint32_t next = be(next_raw);
      next_raw = __ldg(ptr++);
      const uint32_t sh = (uint32_t)(32 - avail);
      hi |= sh ? next >> (32 - sh) : 0u;
      lo = sh ? next << sh : next;
      avail += 32;
    }
  }
};

// codes of kJpegLookBits + 1 .. 16 bits: T.81 F.2.2.3 (Figure F.16), one length at a time; returns a first-level entry
__device__ __noinline__ uint32_t huff_long_code(const HuffSmem* sm, uint32_t table, uint32_t bits) {  // table: 0, 1 = DC, 2, 3 = AC
  const JpegHuffTable& t = table < 2 ? sm->dc[table] : sm->ac[table - 2];
  const bool ac = table >= 2;
  for (uint32_t l = kJpegLookBits + 1; l <= 16; l++) {
    const int code = (int)(bits >> (32 - l));
    if (code <= t.maxcode[l]) {
      const uint32_t sym = t.vals[(t.valoff[l] + code) & 255], s = sym & 15u, r = ac ? sym >> 4 : 0u;
      const uint32_t step = !ac ? 1u : (s ? r + 1u : (r == 15u ? 16u : 64u));
      return l | (s << 5) | (r << 9) | ((l + s) << 13) | (step << 18);
    }
  }
  return 16u | (16u << 13) | ((ac ? 64u : 1u) << 18);  // no code (only while decoding from a guessed position): 16 bits, EOB / zero difference
}

// EXTEND, T.81 F.2.2.1 (Figure F.12)
__device__ __forceinline__ int extend(uint32_t v, uint32_t s) { return v < (1u << (s - 1)) ? (int)v - (int)(1u << s) + 1 : (int)v; }

// position of the block in progress, advanced block by block without divisions
struct BlockCursor {
  uint32_t mx, my, blk;
  __device__ __forceinline__ void open(const JpegScan& sc, uint32_t b) {
    blk = b;
    const uint32_t mcu = b / (uint32_t)sc.blocks_per_mcu;
    my = mcu / sc.mcus_x;
    mx = mcu - my * sc.mcus_x;
  }
  __device__ __forceinline__ int16_t* addr(const JpegScan& sc, uint32_t c) const {
    const uint32_t comp = sc.blk_comp[c];
    const uint32_t bx = mx * sc.comp_h[comp] + sc.blk_dx[c], by = my * sc.comp_v[comp] + sc.blk_dy[c];
    return sc.coeffs[comp] + ((size_t)by * sc.comp_blocks_w[comp] + bx) * 64;
  }
  __device__ __forceinline__ void next(const JpegScan& sc, uint32_t c_after) {  // c_after == 0: the MCU is complete
    blk++;
    if (c_after == 0 && ++mx == sc.mcus_x) {
      mx = 0;
      my++;
    }
  }
};

// Decodes from bit p in state (c, k) until the first symbol that starts at or after p_end (or, when writing, until block
// blk_end - 1, the last one of the image or of the restart interval, is complete). blk = scan-order index of the block in progress. Returns the exit state; `done`
// counts the blocks completed. The symbol-to-symbol dependency chain is: table entry (one shared-memory load indexed by
// the window), its `length + size` field, the window shift; everything else hangs off it.
template <bool WRITE>
__device__ __forceinline__ uint32_t decode_span(const JpegScan& sc, const HuffSmem* sm, uint64_t p, const uint64_t p_end, uint32_t c, uint32_t k,
                                                uint32_t blk, const uint32_t blk_end, uint32_t& done) {
  const uint32_t B = (uint32_t)sc.blocks_per_mcu, dc_sel = sc.dc_mask, ac_sel = sc.ac_mask;
  done = 0;
  if (p >= p_end) return pack_state(0, c, k);
  BlockCursor cur;
  int16_t* out = nullptr;
  if (WRITE) {
    if (blk >= blk_end) return pack_state(0, c, k);
    cur.open(sc, blk);
    out = cur.addr(sc, c);
  }
  // shared-memory byte addresses of the four first-level tables (HuffSmem = dc[0], dc[1], ac[0], ac[1])
  const uint32_t sm0 = (uint32_t)__cvta_generic_to_shared(sm);
  constexpr uint32_t kTab = (uint32_t)sizeof(JpegHuffTable);
  uint32_t tdc = sm0 + ((dc_sel >> c) & 1u) * kTab, tac = sm0 + (2u + ((ac_sel >> c) & 1u)) * kTab;
  uint32_t n = 0;
  BitWindow bw;
  bw.open(sc.data, p);
  int left = (int)(p_end - p);  // bits before the end of the span (<= kJpegSubseqBits); <= 0: the next symbol belongs to the successor
  while (left > 0) {
    const uint32_t bits = bw.peek();
    const bool ac = k != 0;
    const uint32_t t = ac ? tac : tdc;
    uint32_t e;
    asm("ld.shared.u32 %0, [%1];" : "=r"(e) : "r"(t + ((bits >> (32 - kJpegLookBits)) << 2)));  // JpegHuffTable::fast comes first
    if (e == 0) e = huff_long_code(sm, (t - sm0) / kTab, bits);
    const uint32_t used = (e >> 13) & 31u;
    if (WRITE) {
      const uint32_t s = (e >> 5) & 15u, kpos = k + ((e >> 9) & 15u);  // zig-zag index of this coefficient (a DC entry has run 0)
      if (s && kpos < 64) out[c_zigzag[kpos]] = (int16_t)extend((bits << (e & 31u)) >> (32 - s), s);
    }
    k += e >> 18;  // run + 1, or 16 for ZRL, or 64 for EOB (F.2.2.2): the block is complete at k >= 64
    bw.skip(used);
    left -= (int)used;
    if (k >= 64) {  // block complete (k > 64 only for corrupt data or a guessed state)
      k = 0;
      n++;
      if (++c == B) c = 0;
      tdc = sm0 + ((dc_sel >> c) & 1u) * kTab;
      tac = sm0 + (2u + ((ac_sel >> c) & 1u)) * kTab;
      if (WRITE) {
        cur.next(sc, c);
        if (cur.blk >= blk_end) break;  // the image's (or the restart interval's) last block is complete
        out = cur.addr(sc, c);
      }
    }
  }
  done = n;
  return pack_state((uint32_t)(left < 0 ? -left : 0), c, k);
}

// a subsequence whose entry state is known without decoding: the first one of the scan or of a restart interval
__device__ __forceinline__ bool hard_start(const JpegScan& sc, uint32_t t) {
  return t == 0 || (sc.interval_of != nullptr && __ldg(sc.interval_first + __ldg(sc.interval_of + t)) == t);
}

__device__ __forceinline__ void span_of(const JpegScan& sc, uint32_t t, uint64_t& b, uint64_t& e) {
  b = (uint64_t)t * kJpegSubseqBits;
  e = b + kJpegSubseqBits;
  if (e > sc.total_bits) e = sc.total_bits;
  if (b > sc.total_bits) b = sc.total_bits;
}

// Passes 1 and 2. A CTA owns kHuffThreads consecutive subsequences and runs the fixed-point iteration on them in shared
// memory: in every step a thread whose left neighbour's exit state changed in the previous step decodes its subsequence
// again from that state. FIRST: all threads start from the guessed state (thread 0 of CTA 0 from the exact one). Later
// rounds: a CTA whose predecessor's LAST exit state changed in the previous round re-enters with it; all others have
// nothing to do. A round in which no CTA's last exit state changes is the global fixed point (the host checks
// n_changed). Steps cost one subsequence's decode latency each, not a kernel launch.
// In the FIRST round a CTA additionally decodes the kHuffWarm subsequences BEFORE its own 128 (lanes 0 .. kHuffWarm - 1,
// results not stored): the chain that starts at the guessed state of the first warm-up lane has fallen into step by the
// time it reaches the CTA's own subsequences (for every stream whose synchronisation distance is below kHuffWarm
// subsequences), so the later rounds find the CTA's exit states already exact and only verify them.
// MODE 2 (after the phase-hypothesis passes below): every subsequence has an exit state already; all of them are decoded
// once more from their left neighbour's state (the proof, and the repair where a hypothesis chain was broken).
constexpr int kRoundLater = 0, kRoundFirst = 1, kRoundVerify = 2;
template <int MODE>
__global__ void __launch_bounds__(kHuffThreads + kHuffWarm) huff_round_kernel(const __grid_constant__ JpegScan sc, uint32_t* __restrict__ exit_state,
                                                                              uint32_t* __restrict__ count, const uint8_t* __restrict__ cta_dirty_in,
                                                                              uint8_t* __restrict__ cta_dirty_out, uint32_t* __restrict__ n_changed,
                                                                              uint32_t* __restrict__ max_steps) {
  constexpr bool FIRST = MODE == kRoundFirst;
  __shared__ HuffSmem sm;
  __shared__ uint32_t exit_s[kHuffThreads + kHuffWarm];
  __shared__ uint8_t dirty_s[2][kHuffThreads + kHuffWarm];
  constexpr uint32_t kLead = FIRST ? kHuffWarm : 0;  // lanes in front of the CTA's own subsequences
  const uint32_t i = threadIdx.x;
  const int64_t tt = (int64_t)blockIdx.x * kHuffThreads + (int64_t)i - (int64_t)kLead;
  const uint32_t t = tt < 0 ? 0u : (uint32_t)tt;
  if (MODE == kRoundLater && (blockIdx.x == 0 || !cta_dirty_in[blockIdx.x - 1])) {  // CTA-uniform
    if (i == 0) cta_dirty_out[blockIdx.x] = 0;
    return;
  }
  load_tables(&sm, sc.tables);
  const bool live = tt >= 0 && t < sc.n_subseq, owned = i >= kLead;
  const bool hard = live && hard_start(sc, t);  // entry state exact from the start: never decoded again
  uint64_t b = 0, e = 0;
  if (live) span_of(sc, t, b, e);
  uint32_t cur = 0, cnt = 0;
  bool ch = false;
  if (FIRST) {
    if (live) cur = decode_span<false>(sc, &sm, b, e, 0, 0, 0, 0, cnt);
    ch = live;
  } else if (live) {
    cur = exit_state[t];
    cnt = count[t];
  }
  const uint32_t entry_exit = cur;  // this subsequence's exit state before the round
  if (!FIRST && i == 0 && !hard) {  // re-enter with the predecessor CTA's last exit state
    const uint32_t st = exit_state[t - 1];
    const uint32_t now = decode_span<false>(sc, &sm, b + (st >> 16), e, (st >> 8) & 0xFFu, st & 0xFFu, 0, 0, cnt);
    ch = now != cur;
    cur = now;
  }
  if (MODE == kRoundVerify) ch = live;  // every lane decodes again from its neighbour's state in the first step
  exit_s[i] = cur;
  dirty_s[0][i] = ch ? 1 : 0;
  int ph = 0;
  uint32_t steps = 0;
  for (;;) {
    steps++;
    __syncthreads();
    const bool work = live && i > 0 && !hard && dirty_s[ph][i - 1];
    const uint32_t st = work ? exit_s[i - 1] : 0;
    __syncthreads();  // every thread has read its neighbour before any thread publishes a new state
    ch = false;
    if (work) {
      const uint32_t now = decode_span<false>(sc, &sm, b + (st >> 16), e, (st >> 8) & 0xFFu, st & 0xFFu, 0, 0, cnt);
      ch = now != cur;
      cur = now;
      exit_s[i] = now;
    }
    dirty_s[ph ^ 1][i] = ch ? 1 : 0;
    if (!__syncthreads_or(ch ? 1 : 0)) break;
    ph ^= 1;
  }
  if (live && owned) {
    exit_state[t] = cur;
    count[t] = cnt;
  }
  const uint32_t last = min((blockIdx.x + 1) * kHuffThreads, sc.n_subseq) - 1;  // the CTA's last live subsequence
  if (i == 0 && max_steps) atomicMax(max_steps, steps);  // UHDRB200_JPEG_TIMING: the longest in-CTA chain
  if (live && owned && t == last) {
    const bool moved = FIRST || cur != entry_exit;
    cta_dirty_out[blockIdx.x] = moved ? 1 : 0;
    if (moved) atomicAdd(n_changed, 1u);
  }
}

// ---- phase hypotheses (interleaved scans: blocks_per_mcu = H > 1) ---------------------------------------------------------
// Why the chains above are long: a decoder started at a subsequence border falls into step with the bit positions and the
// coefficient index within ~100 bits, but which BLOCK OF THE MCU it is in (luma or chroma tables) only by luck; started
// with the right block it is in step almost at once (scripts/huff_chain_sim.py: mean 0.09 subsequences against 1.5, longest
// 2.7 against 10 in a stream of 1300). So every subsequence is decoded once per possible block-in-MCU (pass A: exit state
// X[j][h], blocks N[j][h]), then once more from each of its left neighbour's H exit states (pass B: the EXACT exit state
// Y[j][h] and block count M[j][h] of subsequence j given that the neighbour's true hypothesis is h, and the hypothesis of
// j that left in the same state, Lnk[j][h]). The true hypotheses are then a composition of H -> H maps along the stream,
// i.e. a scan (pass C), and one verification round (huff_round_kernel<kRoundVerify>) proves the result and repairs the few
// places where no hypothesis of j left in the state Y[j][h] (a broken link: the chain continues with hypothesis 0).
constexpr int kHypReal = 8;     // X, N: one slot per guessed block of the MCU (H <= 8)
constexpr int kHypStride = 16;  // Y, M, Lnk: slots 0 .. 7 = "the neighbour's true hypothesis is h", 8 + h = the virtual hypothesis below

__global__ void __launch_bounds__(kHuffThreads) huff_hyp_kernel(const __grid_constant__ JpegScan sc, uint32_t* __restrict__ X, uint32_t* __restrict__ N) {
  __shared__ HuffSmem sm;
  load_tables(&sm, sc.tables);
  const uint32_t H = (uint32_t)sc.blocks_per_mcu, g = blockIdx.x * kHuffThreads + threadIdx.x, j = g / H, h = g - j * H;
  if (j >= sc.n_subseq) return;
  uint64_t b, e;
  span_of(sc, j, b, e);
  uint32_t cnt = 0;
  X[j * kHypReal + h] = decode_span<false>(sc, &sm, b, e, h, 0, 0, 0, cnt);
  N[j * kHypReal + h] = cnt;
}

// the hypothesis of subsequence j that leaves in state y; failing that (none has fallen into step inside j) one that leaves
// in the same block of the MCU, which falls into step early in j + 1; failing that 0. A wrong choice costs steps of the
// verification round, never exactness.
__device__ __forceinline__ uint32_t link_of(const uint32_t* __restrict__ X, uint32_t j, uint32_t H, uint32_t y, bool& exact) {
  uint32_t l = 0;
  exact = false;
  for (uint32_t k = H; k-- > 0;) {
    const uint32_t x = X[j * kHypReal + k];
    if (x == y) {
      l = k;
      exact = true;
    } else if (!exact && ((x ^ y) & 0xFF00u) == 0) {
      l = k;
    }
  }
  return l;
}

// Thread (j, h): subsequence j decoded from the exit state of hypothesis h of j - 1. When no hypothesis of j leaves in the
// resulting (exact) state y, the same thread goes on through j + 1 from y and looks for ITS exit state among the hypotheses
// of j + 1: the pair is recorded as the virtual hypothesis 8 + h of j (slots 8 + h of j + 1 hold its exit state, block
// count and link), which repairs every chain that falls into step within two subsequences.
__global__ void __launch_bounds__(kHuffThreads) huff_link_kernel(const __grid_constant__ JpegScan sc, const uint32_t* __restrict__ X,
                                                                 const uint32_t* __restrict__ N, uint32_t* __restrict__ Y, uint32_t* __restrict__ M,
                                                                 uint8_t* __restrict__ Lnk) {
  __shared__ HuffSmem sm;
  load_tables(&sm, sc.tables);
  const uint32_t H = (uint32_t)sc.blocks_per_mcu, g = blockIdx.x * kHuffThreads + threadIdx.x, j = g / H, h = g - j * H;
  if (j >= sc.n_subseq) return;
  uint32_t y, m, l = 0;
  if (hard_start(sc, j)) {  // the entry state is known: block 0, coefficient 0 at the border itself = hypothesis 0
    y = X[j * kHypReal];
    m = N[j * kHypReal];
  } else {
    uint64_t b, e;
    span_of(sc, j, b, e);
    const uint32_t st = X[(j - 1) * kHypReal + h];
    y = decode_span<false>(sc, &sm, b + (st >> 16), e, (st >> 8) & 0xFFu, st & 0xFFu, 0, 0, m);
    bool exact;
    l = link_of(X, j, H, y, exact);
    if (!exact && j + 1 < sc.n_subseq && !hard_start(sc, j + 1)) {
      span_of(sc, j + 1, b, e);
      uint32_t m2;
      const uint32_t y2 = decode_span<false>(sc, &sm, b + (y >> 16), e, (y >> 8) & 0xFFu, y & 0xFFu, 0, 0, m2);
      bool exact2;
      const uint32_t l2 = link_of(X, j + 1, H, y2, exact2);
      const size_t v = (size_t)(j + 1) * kHypStride + 8 + h;
      Y[v] = y2;
      M[v] = m2;
      Lnk[v] = (uint8_t)l2;
      l = 8 + h;
    }
  }
  Y[(size_t)j * kHypStride + h] = y;
  M[(size_t)j * kHypStride + h] = m;
  Lnk[(size_t)j * kHypStride + h] = (uint8_t)l;
}

// pass C: the true hypothesis of every subsequence = Lnk[j][true hypothesis of j - 1], a scan of 16 -> 16 maps (four bits
// per entry in one 64-bit word; slots nobody wrote hold arbitrary links, which only ever feed other unused slots); one CTA,
// a chunk of consecutive subsequences per thread
__device__ __forceinline__ uint32_t map_get(uint64_t f, uint32_t h) { return (uint32_t)(f >> (4 * h)) & 15u; }
__device__ __forceinline__ uint64_t map_then(uint64_t f, uint64_t g) {  // h -> g[f[h]]
  uint64_t r = 0;
#pragma unroll
  for (uint32_t h = 0; h < (uint32_t)kHypStride; h++) r |= (uint64_t)map_get(g, map_get(f, h)) << (4 * h);
  return r;
}
__global__ void __launch_bounds__(1024) huff_select_kernel(const uint32_t* __restrict__ Y, const uint32_t* __restrict__ M, const uint8_t* __restrict__ Lnk,
                                                           uint32_t* __restrict__ exit_state, uint32_t* __restrict__ count, uint32_t n) {
  __shared__ uint64_t part[1024];
  constexpr uint64_t kIdentity = 0xFEDCBA9876543210ull;
  const uint32_t per = (n + 1023) / 1024, lo = threadIdx.x * per, hi = min(lo + per, n);
  uint64_t f = kIdentity;
  for (uint32_t j = lo; j < hi; j++) {
    uint64_t mj = 0;
#pragma unroll
    for (uint32_t h = 0; h < (uint32_t)kHypStride; h++) mj |= (uint64_t)(Lnk[(size_t)j * kHypStride + h] & 15u) << (4 * h);
    f = map_then(f, mj);
  }
  part[threadIdx.x] = f;
  __syncthreads();
  for (uint32_t d = 1; d < 1024; d <<= 1) {  // Hillis-Steele inclusive scan under composition (earlier chunks first)
    const uint64_t v = threadIdx.x >= d ? part[threadIdx.x - d] : kIdentity;
    __syncthreads();
    part[threadIdx.x] = map_then(v, part[threadIdx.x]);
    __syncthreads();
  }
  uint32_t h = threadIdx.x ? map_get(part[threadIdx.x - 1], 0) : 0u;  // subsequence 0 is a hard start: whatever enters, 0 leaves
  for (uint32_t j = lo; j < hi; j++) {
    exit_state[j] = Y[(size_t)j * kHypStride + h];
    count[j] = M[(size_t)j * kHypStride + h];
    h = Lnk[(size_t)j * kHypStride + h] & 15u;
  }
}

// single-CTA exclusive scan of the block counts (n_subseq is a few ten thousand: one pass of 1024 threads, serial chunks)
__global__ void __launch_bounds__(1024) huff_scan_kernel(const uint32_t* __restrict__ count, uint32_t* __restrict__ first_block, uint32_t n) {
  __shared__ uint32_t part[1024];
  const uint32_t per = (n + 1023) / 1024, lo = threadIdx.x * per, hi = min(lo + per, n);
  uint32_t s = 0;
  for (uint32_t i = lo; i < hi; i++) s += count[i];
  part[threadIdx.x] = s;
  __syncthreads();
  for (uint32_t d = 1; d < 1024; d <<= 1) {  // Hillis-Steele inclusive scan
    const uint32_t v = threadIdx.x >= d ? part[threadIdx.x - d] : 0;
    __syncthreads();
    part[threadIdx.x] += v;
    __syncthreads();
  }
  uint32_t run = threadIdx.x ? part[threadIdx.x - 1] : 0;
  for (uint32_t i = lo; i < hi; i++) {
    first_block[i] = run;
    run += count[i];
  }
}

// pass 4: decode from the exact entry states and write the coefficients (DC as differences). With restart intervals the
// block a subsequence starts in is counted from its interval's first block, which is known (interval * restart_interval MCUs).
__global__ void __launch_bounds__(kHuffThreads) huff_write_kernel(const __grid_constant__ JpegScan sc, const uint32_t* __restrict__ exit_state,
                                                                  const uint32_t* __restrict__ first_block) {
  __shared__ HuffSmem sm;
  load_tables(&sm, sc.tables);
  const uint32_t t = blockIdx.x * kHuffThreads + threadIdx.x;
  if (t >= sc.n_subseq) return;
  uint32_t blk = first_block[t], blk_end = sc.total_blocks;
  bool hard = t == 0;
  if (sc.interval_of != nullptr) {
    const uint32_t iv = sc.interval_of[t], t0 = sc.interval_first[iv], per = sc.restart_interval * (uint32_t)sc.blocks_per_mcu;
    const uint64_t base = (uint64_t)iv * per;
    if (base >= sc.total_blocks) return;  // more intervals in the stream than the image has room for
    blk = (uint32_t)base + (blk - first_block[t0]);
    blk_end = (uint32_t)min((uint64_t)sc.total_blocks, base + per);
    hard = t == t0;
  }
  const uint32_t st = hard ? 0 : exit_state[t - 1];
  uint64_t b, e;
  span_of(sc, t, b, e);
  uint32_t n;
  decode_span<true>(sc, &sm, b + (st >> 16), e, (st >> 8) & 0xFFu, st & 0xFFu, blk, blk_end, n);
}

// pass 5: DC prediction (F.2.2.1: DIFF + PRED) as a scan over one component's blocks in DECODING order: block j of the
// component lies in MCU j / (h v) at offset (j % (h v)) inside it. grid.y = component; three phases through `partial`.
constexpr int kDcChunk = 2048, kDcThreads = 256, kDcPer = kDcChunk / kDcThreads;
__device__ __forceinline__ int16_t* dc_addr(const JpegScan& sc, uint32_t comp, uint32_t j) {
  const uint32_t h = sc.comp_h[comp], v = sc.comp_v[comp], nb = sc.n_comp == 1 ? 1 : h * v;
  const uint32_t mcu = j / nb, i = j - mcu * nb;
  const uint32_t my = mcu / sc.mcus_x, mx = mcu - my * sc.mcus_x;
  const uint32_t hh = sc.n_comp == 1 ? 1 : h, vv = sc.n_comp == 1 ? 1 : v;
  const uint32_t bx = mx * hh + i % hh, by = my * vv + i / hh;
  return sc.coeffs[comp] + ((size_t)by * sc.comp_blocks_w[comp] + bx) * 64;
}
__device__ __forceinline__ uint32_t comp_blocks(const JpegScan& sc, uint32_t comp) {
  return sc.n_comp == 1 ? sc.total_blocks : sc.mcus_x * sc.mcus_y * sc.comp_h[comp] * sc.comp_v[comp];
}
__device__ __forceinline__ int block_sum(int v, int* warp_sums) {  // CTA-wide inclusive scan of one value per thread
  const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int o = __shfl_up_sync(0xffffffffu, v, d);
    if (lane >= (uint32_t)d) v += o;
  }
  if (lane == 31) warp_sums[w] = v;
  __syncthreads();
  int base = 0;
  for (uint32_t i = 0; i < w; i++) base += warp_sums[i];
  __syncthreads();
  return v + base;
}
__global__ void __launch_bounds__(kDcThreads) dc_partial_kernel(const __grid_constant__ JpegScan sc, int* __restrict__ partial, uint32_t chunks) {
  __shared__ int ws[kDcThreads / 32];
  const uint32_t comp = blockIdx.y, n = comp_blocks(sc, comp), j0 = blockIdx.x * kDcChunk + threadIdx.x * kDcPer;
  int s = 0;
  for (uint32_t i = 0; i < (uint32_t)kDcPer; i++)
    if (j0 + i < n) s += dc_addr(sc, comp, j0 + i)[0];
  const int incl = block_sum(s, ws);
  if (threadIdx.x == kDcThreads - 1) partial[comp * chunks + blockIdx.x] = incl;
}
__global__ void __launch_bounds__(1024) dc_offsets_kernel(int* __restrict__ partial, uint32_t chunks) {  // grid.x = component
  __shared__ int ws[32];
  int* p = partial + blockIdx.x * chunks;
  int carry = 0;
  for (uint32_t base = 0; base < chunks; base += 1024) {  // exclusive scan in place, 1024 chunk sums at a time
    const uint32_t i = base + threadIdx.x;
    const int v = i < chunks ? p[i] : 0;
    const int incl = block_sum(v, ws);
    __shared__ int total;
    if (threadIdx.x == 1023) total = incl;
    if (i < chunks) p[i] = carry + incl - v;
    __syncthreads();
    carry += total;
    __syncthreads();
  }
}
__global__ void __launch_bounds__(kDcThreads) dc_apply_kernel(const __grid_constant__ JpegScan sc, const int* __restrict__ partial, uint32_t chunks) {
  __shared__ int ws[kDcThreads / 32];
  const uint32_t comp = blockIdx.y, n = comp_blocks(sc, comp), j0 = blockIdx.x * kDcChunk + threadIdx.x * kDcPer;
  int d[kDcPer], s = 0;
  for (uint32_t i = 0; i < (uint32_t)kDcPer; i++) {
    d[i] = j0 + i < n ? (int)dc_addr(sc, comp, j0 + i)[0] : 0;
    s += d[i];
  }
  int run = block_sum(s, ws) - s + partial[comp * chunks + blockIdx.x];
  for (uint32_t i = 0; i < (uint32_t)kDcPer; i++) {
    run += d[i];
    if (j0 + i < n) dc_addr(sc, comp, j0 + i)[0] = (int16_t)run;
  }
}

// DC prediction with restart intervals: the prediction restarts at every interval, i.e. every restart_interval * h * v
// elements of a component's decoding-order sequence. One warp per (component, interval): a warp-wide scan, 32 elements a step.
__global__ void __launch_bounds__(kDcThreads) dc_interval_kernel(const __grid_constant__ JpegScan sc) {
  const uint32_t comp = blockIdx.y, n = comp_blocks(sc, comp), lane = threadIdx.x & 31;
  const uint32_t seg = blockIdx.x * (kDcThreads / 32) + (threadIdx.x >> 5);
  const uint32_t nb = sc.n_comp == 1 ? 1u : (uint32_t)sc.comp_h[comp] * sc.comp_v[comp], len = sc.restart_interval * nb;
  const uint64_t j0 = (uint64_t)seg * len;
  if (j0 >= n) return;  // warp-uniform
  const uint32_t j1 = (uint32_t)min((uint64_t)n, j0 + len);
  int carry = 0;
  for (uint32_t j = (uint32_t)j0; j < j1; j += 32) {
    const bool in = j + lane < j1;
    int16_t* p = in ? dc_addr(sc, comp, j + lane) : nullptr;
    int v = in ? (int)p[0] : 0;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int o = __shfl_up_sync(0xffffffffu, v, d);
      if (lane >= (uint32_t)d) v += o;
    }
    v += carry;
    if (in) p[0] = (int16_t)v;
    carry = __shfl_sync(0xffffffffu, v, 31);
  }
}

}  // namespace

size_t jpeg_huff_workspace_bytes(uint32_t n_subseq) {
  const size_t n = ((size_t)n_subseq + 63) & ~(size_t)63;
  const size_t dc_chunks = 3 * (((size_t)8192 * 8192 / 64 + kDcChunk - 1) / kDcChunk);  // largest image jpeg_parse admits: 65535^2 is clamped by the caller
  return n * (3 * 4 + 2) + 256 + dc_chunks * 4 + 64 + n * (kHypReal * 8 + kHypStride * 9);  // + X, N | Y, M, Lnk of the phase hypotheses
}

cudaError_t jpeg_huff_decode(const JpegScan& sc, void* workspace, cudaStream_t s, int* iters) {
  cudaError_t e;
  const uint32_t n = sc.n_subseq;
  if (n == 0 || sc.total_blocks == 0) return cudaSuccess;
  const size_t np = ((size_t)n + 63) & ~(size_t)63;
  uint32_t* exit_a = (uint32_t*)workspace;
  uint32_t* count = exit_a + np;
  uint32_t* first = count + np;
  uint8_t* dirty_a = (uint8_t*)(first + np);  // one flag per CTA, ping-pong
  uint8_t* dirty_b = dirty_a + np;
  uint32_t* n_changed = (uint32_t*)(dirty_b + np);  // 64 counters, one per round
  int* partial = (int*)(n_changed + 64);
  if ((e = cudaMemsetAsync(n_changed, 0, 256, s)) != cudaSuccess) return e;
  const unsigned grid = (n + kHuffThreads - 1) / kHuffThreads;
  static const bool timing = [] { const char* v = getenv("UHDRB200_JPEG_TIMING"); return v && v[0] == '1'; }();
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  if (timing) {
    for (auto& x : ev) cudaEventCreate(&x);
    cudaEventRecord(ev[0], s);
  }
  uint32_t* step_stat = timing ? n_changed + 62 : nullptr;  // slots 62, 63: longest chain of the first / the later rounds
  const size_t dc_chunks = 3 * (((size_t)8192 * 8192 / 64 + kDcChunk - 1) / kDcChunk);  // as in jpeg_huff_workspace_bytes
  uint32_t* hyp_x = (uint32_t*)(((uintptr_t)(partial + dc_chunks) + 63) & ~(uintptr_t)63);
  uint32_t* hyp_n = hyp_x + np * kHypReal;
  uint32_t* hyp_y = hyp_n + np * kHypReal;
  uint32_t* hyp_m = hyp_y + np * kHypStride;
  uint8_t* hyp_l = (uint8_t*)(hyp_m + np * kHypStride);
  const uint32_t H = (uint32_t)sc.blocks_per_mcu;
  // Where the hypotheses pay (4K, one host thread, same box): 0.49 -> 0.47 ms at quality 75, 1.32 -> 1.17 at 90, 1.47 -> 1.43 at
  // 95, 1.36 -> 1.16 with a restart interval per MCU row. Where they do not: streams with few end-of-block codes (quality
  // 100: 200 bits per block, 3.5 -> 3.9 ms), whose chains wait for the COEFFICIENT INDEX to fall into step, which no guess of
  // the block helps, and short restart intervals, which bound every chain anyway (16 MCUs: 0.99 -> 1.02 ms).
  const uint64_t bits_per_block = sc.total_bits / sc.total_blocks;
  const bool short_intervals = sc.restart_interval != 0 && (uint64_t)n < 16ull * sc.n_intervals;
  const bool hyp = UHDR_JPEG_HYP && H >= 2 && H <= (uint32_t)kHypReal && bits_per_block < 128 && !short_intervals;
  int pass = 0;
  uint8_t* din = dirty_a;
  uint8_t* dout = dirty_b;
  bool settled = false;
  if (hyp) {
    const unsigned hgrid = (unsigned)(((uint64_t)n * H + kHuffThreads - 1) / kHuffThreads);
    huff_hyp_kernel<<<hgrid, kHuffThreads, 0, s>>>(sc, hyp_x, hyp_n);
    huff_link_kernel<<<hgrid, kHuffThreads, 0, s>>>(sc, hyp_x, hyp_n, hyp_y, hyp_m, hyp_l);
    huff_select_kernel<<<1, 1024, 0, s>>>(hyp_y, hyp_m, hyp_l, exit_a, count, n);
    huff_round_kernel<kRoundVerify><<<grid, kHuffThreads, 0, s>>>(sc, exit_a, count, dirty_b, dirty_a, n_changed, step_stat);
    if (timing) cudaEventRecord(ev[1], s);
    uint32_t moved = 0;
    if ((e = cudaMemcpyAsync(&moved, n_changed, 4, cudaMemcpyDeviceToHost, s)) != cudaSuccess) return e;
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return e;
    settled = moved == 0;  // no CTA's last exit state moved: every entry state is exact
    pass = 3;              // launches so far beyond the first (for the caller's launch count)
  } else {
    huff_round_kernel<kRoundFirst><<<grid, kHuffThreads + kHuffWarm, 0, s>>>(sc, exit_a, count, dirty_b, dirty_a, n_changed, step_stat);
    if (timing) cudaEventRecord(ev[1], s);
  }
  // Rounds across CTAs, two per host round trip: after the first, every CTA has re-entered with its predecessor's exact
  // exit state (a CTA's LAST exit state is almost always right already, 128 subsequences after a guessed start); the
  // second finds nothing to do and proves the fixed point. A round that changes nothing costs a launch.
  int rounds = 0;
  while (!settled) {
    uint32_t last_changed = 0;
    for (int k = 0; k < 2; k++) {
      pass++;
      rounds++;
      if (pass % 64 == 0) {  // the 64 counters come round again; the previous pair has been read
        if ((e = cudaMemsetAsync(n_changed, 0, 256, s)) != cudaSuccess) return e;
      }
      huff_round_kernel<kRoundLater><<<grid, kHuffThreads, 0, s>>>(sc, exit_a, count, din, dout, n_changed + pass % 64, timing ? step_stat + 1 : nullptr);
      uint8_t* td = din; din = dout; dout = td;
    }
    if ((e = cudaMemcpyAsync(&last_changed, n_changed + pass % 64, 4, cudaMemcpyDeviceToHost, s)) != cudaSuccess) return e;
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return e;
    if (last_changed == 0) break;           // nothing moved in the last round: every entry state is exact
    if ((uint32_t)rounds > grid + 4) break;   // cannot happen: exactness advances by at least one CTA per round
  }
  if (iters) *iters = pass;
  if (timing) cudaEventRecord(ev[2], s);
  huff_scan_kernel<<<1, 1024, 0, s>>>(count, first, n);
  huff_write_kernel<<<grid, kHuffThreads, 0, s>>>(sc, exit_a, first);
  if (timing) {
    cudaEventRecord(ev[3], s);
    cudaEventSynchronize(ev[3]);
    float a = 0, b = 0, c = 0;
    cudaEventElapsedTime(&a, ev[0], ev[1]);
    cudaEventElapsedTime(&b, ev[1], ev[2]);
    cudaEventElapsedTime(&c, ev[2], ev[3]);
    uint32_t st2[2] = {0, 0};
    cudaMemcpy(st2, n_changed + 62, 8, cudaMemcpyDeviceToHost);
    fprintf(stderr, "[uhdrb200 jpeg huffman] %u subsequences: first round %.0f us (longest chain %u steps), %d more rounds %.0f us (%u steps), scan + write pass %.0f us\n",
            n, a * 1e3, st2[0], pass, b * 1e3, st2[1], c * 1e3);
    for (auto& x : ev) cudaEventDestroy(x);
  }
  uint32_t max_blocks = 0;
  for (int c = 0; c < sc.n_comp; c++) {
    const uint32_t nb = sc.n_comp == 1 ? sc.total_blocks : sc.mcus_x * sc.mcus_y * sc.comp_h[c] * sc.comp_v[c];
    max_blocks = nb > max_blocks ? nb : max_blocks;
  }
  if (sc.restart_interval) {
    const uint32_t segs = (sc.mcus_x * sc.mcus_y + sc.restart_interval - 1) / sc.restart_interval;
    dc_interval_kernel<<<dim3((segs + kDcThreads / 32 - 1) / (kDcThreads / 32), sc.n_comp), kDcThreads, 0, s>>>(sc);
    return cudaGetLastError();
  }
  const uint32_t chunks = (max_blocks + kDcChunk - 1) / kDcChunk;
  dc_partial_kernel<<<dim3(chunks, sc.n_comp), kDcThreads, 0, s>>>(sc, partial, chunks);
  dc_offsets_kernel<<<sc.n_comp, 1024, 0, s>>>(partial, chunks);
  dc_apply_kernel<<<dim3(chunks, sc.n_comp), kDcThreads, 0, s>>>(sc, partial, chunks);
  return cudaGetLastError();
}

}  // namespace uhdr
