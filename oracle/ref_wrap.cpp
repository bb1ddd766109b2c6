// ref_wrap.cpp — TEST INFRASTRUCTURE ONLY (never linked into the product library).
//
// C wrapper around the UNMODIFIED reference hot path, compiled by oracle/Makefile from the
// reference sources where they lie under /root/reference (gainmapmath.cpp and lines 1-558 of
// ultrahdr.cpp); nothing of the reference is copied into this repository. The protected members
// are reached the way the reference's own test does it (tests/jpegr_test.cpp:2203-2208): a
// subclass with using-declarations.
//
// Exposes (a) the three path entry points with the POD structs of include/uhdr_b200.h and (b) the
// L1 pixel-math functions one by one, so tests can pin oracle/uhdr_oracle.c against the real code.
#include <cstring>
#include <string>

#include "ultrahdr/gainmapmath.h"
#include "ultrahdr/ultrahdr.h"

#include "uhdr_b200.h"

namespace {
struct RefPath : public ultrahdr::UltraHdr {
  using ultrahdr::UltraHdr::applyGainMap;
  using ultrahdr::UltraHdr::generateGainMap;
  using ultrahdr::UltraHdr::toneMap;
};

ultrahdr::ultrahdr_uncompressed_struct toRef(const uhdrb200_image* im) {
  ultrahdr::ultrahdr_uncompressed_struct r;
  r.data = im->data;
  r.width = im->width;
  r.height = im->height;
  r.colorGamut = static_cast<ultrahdr::ultrahdr_color_gamut>(im->colorGamut);
  r.chroma_data = im->chroma_data;
  r.luma_stride = im->luma_stride;
  r.chroma_stride = im->chroma_stride;
  r.pixelFormat = static_cast<ultrahdr::ultrahdr_pixel_format>(im->pixelFormat);
  return r;
}
void fromRef(const ultrahdr::ultrahdr_uncompressed_struct& r, uhdrb200_image* im) {
  im->data = r.data;
  im->width = r.width;
  im->height = r.height;
  im->colorGamut = r.colorGamut;
  im->chroma_data = r.chroma_data;
  im->luma_stride = r.luma_stride;
  im->chroma_stride = r.chroma_stride;
  im->pixelFormat = r.pixelFormat;
}
ultrahdr::ultrahdr_metadata_struct toRef(const uhdrb200_metadata* m) {
  ultrahdr::ultrahdr_metadata_struct r;
  r.version = std::string(m->version, strnlen(m->version, sizeof(m->version)));
  r.maxContentBoost = m->maxContentBoost;
  r.minContentBoost = m->minContentBoost;
  r.gamma = m->gamma;
  r.offsetSdr = m->offsetSdr;
  r.offsetHdr = m->offsetHdr;
  r.hdrCapacityMin = m->hdrCapacityMin;
  r.hdrCapacityMax = m->hdrCapacityMax;
  return r;
}
void fromRef(const ultrahdr::ultrahdr_metadata_struct& r, uhdrb200_metadata* m) {
  memset(m->version, 0, sizeof(m->version));
  strncpy(m->version, r.version.c_str(), sizeof(m->version) - 1);
  m->maxContentBoost = r.maxContentBoost;
  m->minContentBoost = r.minContentBoost;
  m->gamma = r.gamma;
  m->offsetSdr = r.offsetSdr;
  m->offsetHdr = r.offsetHdr;
  m->hdrCapacityMin = r.hdrCapacityMin;
  m->hdrCapacityMax = r.hdrCapacityMax;
}
ultrahdr::Color mk(float a, float b, float c) { return {{{a, b, c}}}; }
}  // namespace

extern "C" {

// 1 when built with the six -DUSE_*_LUT=1 macros (the "LUT variant"), else 0 (as-built).
int ref_is_lut_variant(void) {
#if USE_SRGB_INVOETF_LUT
  return 1;
#else
  return 0;
#endif
}

// generateGainMap: the map is copied into map_buf (size checked) and the reference's new[] buffer
// is released here, so callers never own reference memory.
int ref_generate_gainmap(const uhdrb200_image* yuv420, const uhdrb200_image* p010, int hdr_tf,
                         uhdrb200_metadata* metadata, uhdrb200_image* dest, int sdr_is_601,
                         void* map_buf, size_t map_buf_bytes) {
  ultrahdr::ultrahdr_uncompressed_struct sdr, hdr, out;
  ultrahdr::ultrahdr_metadata_struct md;
  if (yuv420) sdr = toRef(yuv420);
  if (p010) hdr = toRef(p010);
  out.data = nullptr;
  int st = RefPath::generateGainMap(yuv420 ? &sdr : nullptr, p010 ? &hdr : nullptr,
                                    static_cast<ultrahdr::ultrahdr_transfer_function>(hdr_tf),
                                    metadata ? &md : nullptr, dest ? &out : nullptr,
                                    sdr_is_601 != 0);
  if (st != ultrahdr::ULTRAHDR_NO_ERROR) return st;
  size_t n = out.width * out.height;
  if (map_buf == nullptr || map_buf_bytes < n) {
    delete[] reinterpret_cast<uint8_t*>(out.data);
    return ultrahdr::ERROR_ULTRAHDR_BUFFER_TOO_SMALL;
  }
  memcpy(map_buf, out.data, n);
  delete[] reinterpret_cast<uint8_t*>(out.data);
  out.data = map_buf;
  fromRef(out, dest);
  fromRef(md, metadata);
  return st;
}

int ref_apply_gainmap(const uhdrb200_image* yuv420, const uhdrb200_image* gainmap,
                      const uhdrb200_metadata* metadata, int output_format,
                      float max_display_boost, uhdrb200_image* dest) {
  ultrahdr::ultrahdr_uncompressed_struct sdr, map, out;
  ultrahdr::ultrahdr_metadata_struct md;
  if (yuv420) sdr = toRef(yuv420);
  if (gainmap) map = toRef(gainmap);
  if (metadata) md = toRef(metadata);
  if (dest) out = toRef(dest);
  RefPath r;
  int st = r.applyGainMap(yuv420 ? &sdr : nullptr, gainmap ? &map : nullptr,
                          metadata ? &md : nullptr,
                          static_cast<ultrahdr::ultrahdr_output_format>(output_format),
                          max_display_boost, dest ? &out : nullptr);
  if (dest) fromRef(out, dest);
  return st;
}

int ref_tonemap(const uhdrb200_image* src, uhdrb200_image* dest) {
  ultrahdr::ultrahdr_uncompressed_struct s, d;
  if (src) s = toRef(src);
  if (dest) d = toRef(dest);
  RefPath r;
  int st = r.toneMap(src ? &s : nullptr, dest ? &d : nullptr);
  if (dest) fromRef(d, dest);
  return st;
}

// JpegR::convertYuv lives in jpegr.cpp (needs libjpeg headers, not buildable here); its body is a
// selector + a double loop over transformYuv420 (jpegr.cpp:1197-1203). The selector and loop are
// restated here; the per-quad arithmetic is the reference's own transformYuv420.
int ref_convert_yuv(uhdrb200_image* image, int src_encoding, int dest_encoding) {
  using namespace ultrahdr;
  if (image == nullptr) return ERROR_ULTRAHDR_BAD_PTR;
  if (src_encoding == ULTRAHDR_COLORGAMUT_UNSPECIFIED ||
      dest_encoding == ULTRAHDR_COLORGAMUT_UNSPECIFIED)
    return ERROR_ULTRAHDR_INVALID_COLORGAMUT;
  ColorTransformFn table[3][3] = {{nullptr, yuv709To601, yuv709To2100},
                                  {yuv601To709, nullptr, yuv601To2100},
                                  {yuv2100To709, yuv2100To601, nullptr}};
  if (src_encoding < 0 || src_encoding > 2 || dest_encoding < 0 || dest_encoding > 2)
    return ERROR_ULTRAHDR_INVALID_COLORGAMUT;
  if (src_encoding == dest_encoding) return ULTRAHDR_NO_ERROR;
  ColorTransformFn fn = table[src_encoding][dest_encoding];
  ultrahdr_uncompressed_struct im = toRef(image);
  for (size_t y = 0; y < im.height / 2; ++y)
    for (size_t x = 0; x < im.width / 2; ++x) transformYuv420(&im, x, y, fn);
  return ULTRAHDR_NO_ERROR;
}

// ---- L1 functions, one by one (for pinning oracle/uhdr_oracle.c) ----
float ref_srgbInvOetf(float x) { return ultrahdr::srgbInvOetf(x); }
float ref_srgbInvOetfLUT(float x) { return ultrahdr::srgbInvOetfLUT(x); }
float ref_hlgOetf(float x) { return ultrahdr::hlgOetf(x); }
float ref_hlgOetfLUT(float x) { return ultrahdr::hlgOetfLUT(x); }
float ref_hlgInvOetf(float x) { return ultrahdr::hlgInvOetf(x); }
float ref_hlgInvOetfLUT(float x) { return ultrahdr::hlgInvOetfLUT(x); }
float ref_pqOetf(float x) { return ultrahdr::pqOetf(x); }
float ref_pqOetfLUT(float x) { return ultrahdr::pqOetfLUT(x); }
float ref_pqInvOetf(float x) { return ultrahdr::pqInvOetf(x); }
float ref_pqInvOetfLUT(float x) { return ultrahdr::pqInvOetfLUT(x); }
float ref_luminance(int gamut, float r, float g, float b) {
  ultrahdr::Color c = mk(r, g, b);
  switch (gamut) {
    case 0: return ultrahdr::srgbLuminance(c);
    case 1: return ultrahdr::p3Luminance(c);
    default: return ultrahdr::bt2100Luminance(c);
  }
}
void ref_yuvToRgb(int gamut, const float* yuv, float* rgb) {
  ultrahdr::Color c = mk(yuv[0], yuv[1], yuv[2]), o;
  switch (gamut) {
    case 0: o = ultrahdr::srgbYuvToRgb(c); break;
    case 1: o = ultrahdr::p3YuvToRgb(c); break;
    default: o = ultrahdr::bt2100YuvToRgb(c); break;
  }
  rgb[0] = o.r; rgb[1] = o.g; rgb[2] = o.b;
}
void ref_rgbToYuv(int gamut, const float* rgb, float* yuv) {
  ultrahdr::Color c = mk(rgb[0], rgb[1], rgb[2]), o;
  switch (gamut) {
    case 0: o = ultrahdr::srgbRgbToYuv(c); break;
    case 1: o = ultrahdr::p3RgbToYuv(c); break;
    default: o = ultrahdr::bt2100RgbToYuv(c); break;
  }
  yuv[0] = o.y; yuv[1] = o.u; yuv[2] = o.v;
}
// returns 0 and writes out, or -1 when the reference selector returns nullptr
int ref_hdrConversion(int sdr_gamut, int hdr_gamut, const float* in, float* out) {
  ultrahdr::ColorTransformFn fn =
      ultrahdr::getHdrConversionFn(static_cast<ultrahdr::ultrahdr_color_gamut>(sdr_gamut),
                                   static_cast<ultrahdr::ultrahdr_color_gamut>(hdr_gamut));
  if (fn == nullptr) return -1;
  ultrahdr::Color o = fn(mk(in[0], in[1], in[2]));
  out[0] = o.r; out[1] = o.g; out[2] = o.b;
  return 0;
}
// idx: 0 709->601, 1 709->2100, 2 601->709, 3 601->2100, 4 2100->709, 5 2100->601
void ref_yuvToYuv(int idx, const float* in, float* out) {
  ultrahdr::ColorTransformFn fns[6] = {ultrahdr::yuv709To601,  ultrahdr::yuv709To2100,
                                       ultrahdr::yuv601To709,  ultrahdr::yuv601To2100,
                                       ultrahdr::yuv2100To709, ultrahdr::yuv2100To601};
  ultrahdr::Color o = fns[idx](mk(in[0], in[1], in[2]));
  out[0] = o.y; out[1] = o.u; out[2] = o.v;
}
int ref_encodeGain(float y_sdr, float y_hdr, float min_boost, float max_boost) {
  ultrahdr::ultrahdr_metadata_struct md;
  md.minContentBoost = min_boost;
  md.maxContentBoost = max_boost;
  return ultrahdr::encodeGain(y_sdr, y_hdr, &md);
}
void ref_applyGain(const float* rgb, float gain, float min_boost, float max_boost,
                   float display_boost, int use_display_boost, float* out) {
  ultrahdr::ultrahdr_metadata_struct md;
  md.minContentBoost = min_boost;
  md.maxContentBoost = max_boost;
  ultrahdr::Color o = use_display_boost
                          ? ultrahdr::applyGain(mk(rgb[0], rgb[1], rgb[2]), gain, &md, display_boost)
                          : ultrahdr::applyGain(mk(rgb[0], rgb[1], rgb[2]), gain, &md);
  out[0] = o.r; out[1] = o.g; out[2] = o.b;
}
void ref_applyGainLUT(const float* rgb, float gain, float min_boost, float max_boost,
                      float display_boost, float* out) {
  ultrahdr::ultrahdr_metadata_struct md;
  md.minContentBoost = min_boost;
  md.maxContentBoost = max_boost;
  ultrahdr::GainLUT lut(&md, display_boost);
  ultrahdr::Color o = ultrahdr::applyGainLUT(mk(rgb[0], rgb[1], rgb[2]), gain, lut);
  out[0] = o.r; out[1] = o.g; out[2] = o.b;
}
void ref_getYuv420Pixel(const uhdrb200_image* im, size_t x, size_t y, float* out) {
  ultrahdr::ultrahdr_uncompressed_struct r = toRef(im);
  ultrahdr::Color o = ultrahdr::getYuv420Pixel(&r, x, y);
  out[0] = o.y; out[1] = o.u; out[2] = o.v;
}
void ref_getP010Pixel(const uhdrb200_image* im, size_t x, size_t y, float* out) {
  ultrahdr::ultrahdr_uncompressed_struct r = toRef(im);
  ultrahdr::Color o = ultrahdr::getP010Pixel(&r, x, y);
  out[0] = o.y; out[1] = o.u; out[2] = o.v;
}
void ref_sampleYuv420(const uhdrb200_image* im, size_t scale, size_t x, size_t y, float* out) {
  ultrahdr::ultrahdr_uncompressed_struct r = toRef(im);
  ultrahdr::Color o = ultrahdr::sampleYuv420(&r, scale, x, y);
  out[0] = o.y; out[1] = o.u; out[2] = o.v;
}
void ref_sampleP010(const uhdrb200_image* im, size_t scale, size_t x, size_t y, float* out) {
  ultrahdr::ultrahdr_uncompressed_struct r = toRef(im);
  ultrahdr::Color o = ultrahdr::sampleP010(&r, scale, x, y);
  out[0] = o.y; out[1] = o.u; out[2] = o.v;
}
float ref_sampleMapIdw(const uhdrb200_image* map, size_t scale, size_t x, size_t y) {
  ultrahdr::ultrahdr_uncompressed_struct r = toRef(map);
  ultrahdr::ShepardsIDW idw(static_cast<int>(scale));
  return ultrahdr::sampleMap(&r, scale, x, y, idw);
}
float ref_sampleMapFloat(const uhdrb200_image* map, float scale, size_t x, size_t y) {
  ultrahdr::ultrahdr_uncompressed_struct r = toRef(map);
  return ultrahdr::sampleMap(&r, scale, x, y);
}
// table: 0 mWeights, 1 NR, 2 NB, 3 C; out has scale*scale*4 floats
void ref_idwTable(int scale, int table, float* out) {
  ultrahdr::ShepardsIDW idw(scale);
  float* src = table == 0 ? idw.mWeights : table == 1 ? idw.mWeightsNR : table == 2 ? idw.mWeightsNB : idw.mWeightsC;
  memcpy(out, src, sizeof(float) * scale * scale * 4);
}
uint32_t ref_colorToRgba1010102(float r, float g, float b) {
  return ultrahdr::colorToRgba1010102(mk(r, g, b));
}
uint64_t ref_colorToRgbaF16(float r, float g, float b) {
  return ultrahdr::colorToRgbaF16(mk(r, g, b));
}
uint16_t ref_floatToHalf(float f) { return ultrahdr::floatToHalf(f); }

}  // extern "C"
