// generate.cu — A1: UltraHdr::generateGainMap per-map-pixel loop (reference
// lib/src/ultrahdr.cpp:314-336) on sm_100a.
//
// One thread = one gain-map pixel = one 4x4 block of the SDR (YUV420 u8) and HDR (P010) images.
// HBM-bound (4.5625 B/px: 3 B P010 + 1.5 B 420 read, 1/16 B written). A warp covers 128 image
// columns, so every row request is one fully coalesced 256 B (P010 Y/UV), 128 B (420 Y) or 64 B
// (420 U, V) segment; the 14 independent loads per thread are issued before any use.
// The 4x4 box average (gainmapmath.cpp:605-615) is an exact integer sum (dp4a / dp2a) converted
// once; with the fp32 transcendental chain (~1e-6 relative, uhdr_math.cuh) that is the only
// deviation from the reference: map bytes +-1 LSB, >= 99.9 % exact.
#include "uhdr_convert.cuh"
#include "uhdr_io.cuh"
#include "uhdr_kernels.cuh"
#include "uhdr_math.cuh"

namespace uhdr {
namespace {

// FUSED = API-0 (lib/src/jpegr.cpp:167-225): the SDR image is not an input but toneMap(HDR)
// (ultrahdr.cpp:517-558: 8-bit sample = top 8 bits of the P010 word). The thread derives its
// 4x4 SDR block from the P010 block it already holds, WRITES it to the SDR planes (with the
// reference's zero-filled stride tails) and feeds it to the same gain computation: P010 is read
// once and nothing is re-read (4.5625 B/px instead of 4.5 + 4.5625). Requires width, height % 4 == 0.
// CONVERT (with FUSED): the SDR block is stored re-encoded by convertYuv (row N1, jpegr.cpp:223-225,
// transformYuv420 gainmapmath.cpp:483-520) while the gain computation keeps the un-converted samples.
// LUT = the LUT numeric mode (graded against the reference's -DUSE_*_LUT=1 build): after the same gather the thread
// follows the reference's float arithmetic operation by operation (sequential box sum, unfused YUV matrix, nearest-entry
// tables, double log2 in encodeGain) instead of the fp32 fast chain.
// Occupancy bounds (blocks of 256 threads per SM), measured on B200 (scripts/api0_convert_variants.sh): the fused kernels
// are bound by instruction issue and latency, not by HBM, and gain from more resident warps as long as nothing spills:
// with convertYuv 54 registers / 4 blocks 0.60-0.67 of the HBM peak, 48 registers / 5 blocks 0.75, 40 registers / 6
// blocks (52 B of spills) 0.72.
#ifndef UHDR_API0_CONVERT_MIN_BLOCKS
#define UHDR_API0_CONVERT_MIN_BLOCKS 5
#endif
#ifndef UHDR_API0_MIN_BLOCKS
#define UHDR_API0_MIN_BLOCKS 5  // without convertYuv: 66 registers / 3 blocks 0.72, 4 blocks 0.81, 48 registers / 5 blocks 0.86 (same script)
#endif
#ifndef UHDR_GENERATE_MIN_BLOCKS
#define UHDR_GENERATE_MIN_BLOCKS 1
#endif
template <bool VEC, bool FUSED, bool CONVERT = false, bool LUT = false>
__global__ void __launch_bounds__(256, LUT ? 1 : (CONVERT ? UHDR_API0_CONVERT_MIN_BLOCKS : (FUSED ? UHDR_API0_MIN_BLOCKS : UHDR_GENERATE_MIN_BLOCKS)))
    generate_kernel(GenerateParams p) {
  pdl_trigger();
  const Geometry& g = p.g;
  const int mx = blockIdx.x * 32 + threadIdx.x;
  const int my = blockIdx.y * 8 + threadIdx.y;
  if (mx >= g.map_width || my >= g.map_height) return;
  pdl_wait();  // planes (and the descriptor array) may come from the previous kernel of the stream
  const Planes pl = p.batch ? p.batch[blockIdx.z] : p.single;

  // ---- gather the raw samples of the 4x4 block as packed words (all loads first) ----
  uint32_t sy[4];      // 4 x u8 luma per row
  uint32_t su[2], sv[2];  // 2 x u8 chroma per chroma row (low 16 bits)
  uint2 hy[4];         // 4 x u16 luma per row
  uint2 huv[2];        // U0 V0 | U1 V1 per chroma row
  const int x0 = mx * 4, y0 = my * 4;
  {
    const uint8_t* y8 = (const uint8_t*)pl.sdr_y + (size_t)y0 * g.sdr_luma_stride + x0;
    const uint8_t* u8 = (const uint8_t*)pl.sdr_u + (size_t)(y0 >> 1) * g.sdr_chroma_stride + (x0 >> 1);
    const uint8_t* v8 = (const uint8_t*)pl.sdr_v + (size_t)(y0 >> 1) * g.sdr_chroma_stride + (x0 >> 1);
    const uint16_t* y16 = (const uint16_t*)pl.hdr_y + (size_t)y0 * g.hdr_luma_stride + x0;
    const uint16_t* uv16 = (const uint16_t*)pl.hdr_uv + (size_t)(y0 >> 1) * g.hdr_chroma_stride + x0;
    if (FUSED) {
#pragma unroll
      for (int r = 0; r < 4; r++) hy[r] = ldg8(y16 + (size_t)r * g.hdr_luma_stride);
#pragma unroll
      for (int r = 0; r < 2; r++) huv[r] = ldg8(uv16 + (size_t)r * g.hdr_chroma_stride);
      uint8_t* oy = const_cast<uint8_t*>(y8);
      uint8_t* ou = const_cast<uint8_t*>(u8);
      uint8_t* ov = const_cast<uint8_t*>(v8);
#pragma unroll
      for (int r = 0; r < 4; r++) sy[r] = __byte_perm(hy[r].x, hy[r].y, 0x7531);  // high byte of each u16 == (v >> 6) >> 2, ultrahdr.cpp:530-531
#pragma unroll
      for (int r = 0; r < 2; r++) {
        su[r] = __byte_perm(huv[r].x, huv[r].y, 0x7751) & 0xFFFFu;  // U0 U1
        sv[r] = __byte_perm(huv[r].x, huv[r].y, 0x7773) & 0xFFFFu;  // V0 V1
      }
      uint32_t oy_w[4], ou_w[2], ov_w[2];  // what is stored: the tone-mapped block, re-encoded when CONVERT
#pragma unroll
      for (int r = 0; r < 4; r++) oy_w[r] = sy[r];
#pragma unroll
      for (int r = 0; r < 2; r++) { ou_w[r] = su[r]; ov_w[r] = sv[r]; }
      if (CONVERT) {
#pragma unroll
        for (int qr = 0; qr < 2; qr++)
          transform_quad_pair(p.conv, oy_w[2 * qr], oy_w[2 * qr + 1], ou_w[qr], ov_w[qr]);
      }
#pragma unroll
      for (int r = 0; r < 4; r++) stg4(oy + (size_t)r * g.sdr_luma_stride, oy_w[r]);
#pragma unroll
      for (int r = 0; r < 2; r++) {
        stg2(ou + (size_t)r * g.sdr_chroma_stride, ou_w[r]);
        stg2(ov + (size_t)r * g.sdr_chroma_stride, ov_w[r]);
      }
      if (mx == g.map_width - 1) {  // stride tails, ultrahdr.cpp:533-535, 551-554
        for (int x = g.width; x < g.sdr_luma_stride; x++)
#pragma unroll
          for (int r = 0; r < 4; r++) oy[(size_t)r * g.sdr_luma_stride + (x - x0)] = 0;
        for (int x = g.width / 2; x < g.sdr_chroma_stride; x++)
#pragma unroll
          for (int r = 0; r < 2; r++) {
            ou[(size_t)r * g.sdr_chroma_stride + (x - (x0 >> 1))] = 0;
            ov[(size_t)r * g.sdr_chroma_stride + (x - (x0 >> 1))] = 0;
          }
      }
    } else if (VEC) {
#pragma unroll
      for (int r = 0; r < 4; r++) hy[r] = ldg8(y16 + (size_t)r * g.hdr_luma_stride);
#pragma unroll
      for (int r = 0; r < 2; r++) huv[r] = ldg8(uv16 + (size_t)r * g.hdr_chroma_stride);
#pragma unroll
      for (int r = 0; r < 4; r++) sy[r] = ldg4(y8 + (size_t)r * g.sdr_luma_stride);
#pragma unroll
      for (int r = 0; r < 2; r++) {
        su[r] = ldg2(u8 + (size_t)r * g.sdr_chroma_stride);
        sv[r] = ldg2(v8 + (size_t)r * g.sdr_chroma_stride);
      }
    } else {
#pragma unroll
      for (int r = 0; r < 4; r++) {
        const uint16_t* h = y16 + (size_t)r * g.hdr_luma_stride;
        const uint8_t* q = y8 + (size_t)r * g.sdr_luma_stride;
        hy[r] = make_uint2((uint32_t)h[0] | ((uint32_t)h[1] << 16), (uint32_t)h[2] | ((uint32_t)h[3] << 16));
        sy[r] = (uint32_t)q[0] | ((uint32_t)q[1] << 8) | ((uint32_t)q[2] << 16) | ((uint32_t)q[3] << 24);
      }
#pragma unroll
      for (int r = 0; r < 2; r++) {
        const uint16_t* h = uv16 + (size_t)r * g.hdr_chroma_stride;
        huv[r] = make_uint2((uint32_t)h[0] | ((uint32_t)h[1] << 16), (uint32_t)h[2] | ((uint32_t)h[3] << 16));
        su[r] = (uint32_t)u8[(size_t)r * g.sdr_chroma_stride] | ((uint32_t)u8[(size_t)r * g.sdr_chroma_stride + 1] << 8);
        sv[r] = (uint32_t)v8[(size_t)r * g.sdr_chroma_stride] | ((uint32_t)v8[(size_t)r * g.sdr_chroma_stride + 1] << 8);
      }
    }
  }

  // ---- 4x4 box averages (samplePixels, gainmapmath.cpp:605-615) ----
  // The reference sums 16 float-converted samples sequentially and divides by 16. Here the
  // samples are summed EXACTLY as integers (byte / 16-bit dot products with 1) and converted
  // once: the result differs from the reference's float accumulation by a few 1e-7 relative,
  // far inside the +-1 LSB contract (measured exact-match fraction in tests/ and DESIGN.md),
  // and it removes ~300 unpack/convert/add instructions per map pixel from an issue-bound kernel.
  // Each chroma sample covers 4 of the 16 pixels, so chroma averages are sums of 4 samples / 4.
  if (LUT) {
    Color3 s = {0.0f, 0.0f, 0.0f}, h = {0.0f, 0.0f, 0.0f};
#pragma unroll
    for (int r = 0; r < 4; r++)  // samplePixels: dy-major, sequential float sums of the per-pixel conversions
#pragma unroll
      for (int i = 0; i < 4; i++) {
        const int y8 = (int)((sy[r] >> (8 * i)) & 0xFFu);  // getYuv420Pixel, gainmapmath.cpp:562-581
        const int u8 = (int)((su[r >> 1] >> (8 * (i >> 1))) & 0xFFu), v8 = (int)((sv[r >> 1] >> (8 * (i >> 1))) & 0xFFu);
        s.r += (float)y8 * (1 / 255.0f);
        s.g += (float)(u8 - 128) * (1 / 255.0f);
        s.b += (float)(v8 - 128) * (1 / 255.0f);
        const uint32_t yw = i < 2 ? hy[r].x : hy[r].y, cw = i < 2 ? huv[r >> 1].x : huv[r >> 1].y;
        const int y10 = (int)(((yw >> (16 * (i & 1))) & 0xFFFFu) >> 6);  // getP010Pixel, :583-601
        const int u10 = (int)((cw & 0xFFFFu) >> 6), v10 = (int)(cw >> 22);
        h.r += (float)(y10 - 64) * (1 / 876.0f);
        h.g += (float)(u10 - 64) * (1 / 896.0f) - 0.5f;
        h.b += (float)(v10 - 64) * (1 / 896.0f) - 0.5f;
      }
    s = {s.r / 16.0f, s.g / 16.0f, s.b / 16.0f};
    h = {h.r / 16.0f, h.g / 16.0f, h.b / 16.0f};
    const Color3 sg = yuv_to_rgb(p.sdr_k, s);  // ultrahdr.cpp:316-324 with srgbInvOetfLUT
    const Color3 sl = {__ldg(p.lt.srgb_inv + lut_index(sg.r, kLutSrgbInv)), __ldg(p.lt.srgb_inv + lut_index(sg.g, kLutSrgbInv)),
                       __ldg(p.lt.srgb_inv + lut_index(sg.b, kLutSrgbInv))};
    const float sdr_nits = luminance(p.luma, sl) * 203.0f;
    const Color3 hg = yuv_to_rgb(p.hdr_k, h);  // :326-330 with hlgInvOetfLUT / pqInvOetfLUT
    Color3 hl = hg;
    if (p.hdr_tf == 1 || p.hdr_tf == 2) {
      const float* t = p.hdr_tf == 1 ? p.lt.hlg_inv : p.lt.pq_inv;  // both tables have 4096 entries
      hl = {__ldg(t + lut_index(hg.r, kLutHlgInv)), __ldg(t + lut_index(hg.g, kLutHlgInv)), __ldg(t + lut_index(hg.b, kLutHlgInv))};
    }
    if (p.use_gamut) hl = mat3_apply(p.gamut, hl);
    const float hdr_nits = luminance(p.luma, hl) * p.hdr_white_nits;
    ((uint8_t*)pl.out)[(size_t)my * g.map_width + mx] = (uint8_t)encode_gain_exact(sdr_nits, hdr_nits, p.gc);
    return;
  }
  uint32_t ys = 0, us, vs, yh = 0, uh = 0, vh = 0;
#pragma unroll
  for (int r = 0; r < 4; r++) {
    ys = __dp4a(sy[r], 0x01010101u, ys);
    yh = __dp2a_lo(hy[r].x & 0xFFC0FFC0u, 0x0101u, yh);  // 10 MSBs only (getP010Pixel: >> 6)
    yh = __dp2a_lo(hy[r].y & 0xFFC0FFC0u, 0x0101u, yh);
  }
  us = __dp4a(su[0] | (su[1] << 16), 0x01010101u, 0u);
  vs = __dp4a(sv[0] | (sv[1] << 16), 0x01010101u, 0u);
#pragma unroll
  for (int r = 0; r < 2; r++) {
    const uint32_t a = huv[r].x & 0xFFC0FFC0u, b = huv[r].y & 0xFFC0FFC0u;
    uh += (a & 0xFFFFu) + (b & 0xFFFFu);
    vh += (a >> 16) + (b >> 16);
  }
  Color3 s, h;
  s.r = (float)ys * (1.0f / (255.0f * 16.0f));                    // y / 255
  s.g = (float)((int)us - 4 * 128) * (1.0f / (255.0f * 4.0f));        // (u - 128) / 255
  s.b = (float)((int)vs - 4 * 128) * (1.0f / (255.0f * 4.0f));
  h.r = (float)((int)(yh >> 6) - 16 * 64) * (1.0f / (876.0f * 16.0f));  // (y - 64) / 876, no range clamp
  h.g = fmaf((float)((int)(uh >> 6) - 4 * 64), 1.0f / (896.0f * 4.0f), -0.5f);  // (u - 64) / 896 - 0.5
  h.b = fmaf((float)((int)(vh >> 6) - 4 * 64), 1.0f / (896.0f * 4.0f), -0.5f);

  // ---- SDR: YUV->RGB, sRGB^-1, luminance * 203 (ultrahdr.cpp:316-324) ----
  const LumaCoefs lk = p.luma;
  Color3 sg = yuv_to_rgb_fma(p.sdr_k, s);
  Color3 sl = {srgb_inv_oetf(sg.r), srgb_inv_oetf(sg.g), srgb_inv_oetf(sg.b)};
  const float sdr_nits = luminance_fma(lk, sl) * 203.0f;

  // ---- HDR: YUV->RGB, inverse OETF, gamut conversion into the SDR gamut, luminance (:326-330) ----
  Color3 hg = yuv_to_rgb_fma(p.hdr_k, h);
  Color3 hl;
  if (p.hdr_tf == 1) hl = {hlg_inv_oetf(hg.r), hlg_inv_oetf(hg.g), hlg_inv_oetf(hg.b)};
  else if (p.hdr_tf == 2) hl = {pq_inv_oetf(hg.r), pq_inv_oetf(hg.g), pq_inv_oetf(hg.b)};
  else hl = hg;
  if (p.use_gamut) hl = mat3_apply_fma(p.gamut, hl);
  const float hdr_nits = luminance_fma(lk, hl) * p.hdr_white_nits;

  // ---- encodeGain (gainmapmath.cpp:529-541), truncating; clamped gains take the host-computed codes ----
  const uint32_t code = encode_gain(sdr_nits, hdr_nits, p.gc);
  ((uint8_t*)pl.out)[(size_t)my * g.map_width + mx] = (uint8_t)code;
}

bool aligned(const void* q, size_t a) { return (reinterpret_cast<uintptr_t>(q) & (a - 1)) == 0; }
}  // namespace

bool generate_vec_ok(const Geometry& g, const Planes& pl) {
  return aligned(pl.hdr_y, 8) && aligned(pl.hdr_uv, 8) && aligned(pl.sdr_y, 4) && aligned(pl.sdr_u, 2) &&
         aligned(pl.sdr_v, 2) && g.hdr_luma_stride % 4 == 0 && g.hdr_chroma_stride % 4 == 0 &&
         g.sdr_luma_stride % 4 == 0 && g.sdr_chroma_stride % 2 == 0;
}

cudaError_t launch_generate_impl(const GenerateParams& p, int n, bool vec, cudaStream_t s) {
  if (p.g.map_width == 0 || p.g.map_height == 0 || n == 0) return cudaSuccess;
  dim3 block(32, 8, 1);
  dim3 grid((p.g.map_width + 31) / 32, (p.g.map_height + 7) / 8, n);
  if (p.lut) {
    if (vec) return launch_pdl(generate_kernel<true, false, false, true>, grid, block, 0, s, p.pdl != 0, p);
    return launch_pdl(generate_kernel<false, false, false, true>, grid, block, 0, s, p.pdl != 0, p);
  }
  if (vec) return launch_pdl(generate_kernel<true, false, false>, grid, block, 0, s, p.pdl != 0, p);
  return launch_pdl(generate_kernel<false, false, false>, grid, block, 0, s, p.pdl != 0, p);
}

// API-0 fused toneMap + generateGainMap: the sdr_* planes of `Planes` are OUTPUTS here.
bool api0_fused_ok(const Geometry& g, const Planes& pl) {
  return g.width % 4 == 0 && g.height % 4 == 0 && g.width > 0 && g.height > 0 && generate_vec_ok(g, pl);
}
cudaError_t launch_api0_fused_impl(const GenerateParams& p, int n, cudaStream_t s) {
  if (p.g.map_width == 0 || p.g.map_height == 0 || n == 0) return cudaSuccess;
  dim3 block(32, 8, 1);
  dim3 grid((p.g.map_width + 31) / 32, (p.g.map_height + 7) / 8, n);
  if (p.lut) {
    if (p.convert) return launch_pdl(generate_kernel<true, true, true, true>, grid, block, 0, s, p.pdl != 0, p);
    return launch_pdl(generate_kernel<true, true, false, true>, grid, block, 0, s, p.pdl != 0, p);
  }
  if (p.convert) return launch_pdl(generate_kernel<true, true, true>, grid, block, 0, s, p.pdl != 0, p);
  return launch_pdl(generate_kernel<true, true, false>, grid, block, 0, s, p.pdl != 0, p);
}

}  // namespace uhdr
