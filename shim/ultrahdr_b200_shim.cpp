// ultrahdr_b200_shim.cpp — link-time replacement for the three hot-path members of the reference's
// `ultrahdr::UltraHdr` (declared lib/include/ultrahdr/ultrahdr.h:348-350, 370-372, 381; defined in
// lib/src/ultrahdr.cpp:185-358, 360-515, 517-558). Same signatures, same status codes, same
// ownership rules; the bodies forward to the C ABI of include/uhdr_b200.h, i.e. to the sm_100a
// kernels in libuhdr_b200.so. Compile this file INSTEAD of those three definitions (see
// INTEGRATION.md): JpegR, HeifR, UltraHdr::convert and ultrahdr_api.cpp link unchanged on top.
//
// Build: g++ -std=c++17 -I<reference>/lib/include -I<reference> -I<this repo>/include -c ...
// The reference headers are NOT part of this repository; they come from the reference tree.
#include <cstddef>
#include <cstring>
#include <memory>
#include <new>
#include <string>

#include "ultrahdr/ultrahdr.h"

#include "uhdr_b200.h"

namespace ultrahdr {
namespace {

// uhdrb200_image is a field-for-field mirror of ultrahdr_uncompressed_struct (ultrahdr.h:152-180)
static_assert(sizeof(ultrahdr_uncompressed_struct) == sizeof(uhdrb200_image), "descriptor size");
static_assert(offsetof(ultrahdr_uncompressed_struct, data) == offsetof(uhdrb200_image, data), "data");
static_assert(offsetof(ultrahdr_uncompressed_struct, width) == offsetof(uhdrb200_image, width), "width");
static_assert(offsetof(ultrahdr_uncompressed_struct, height) == offsetof(uhdrb200_image, height), "height");
static_assert(offsetof(ultrahdr_uncompressed_struct, colorGamut) == offsetof(uhdrb200_image, colorGamut), "colorGamut");
static_assert(offsetof(ultrahdr_uncompressed_struct, chroma_data) == offsetof(uhdrb200_image, chroma_data), "chroma_data");
static_assert(offsetof(ultrahdr_uncompressed_struct, luma_stride) == offsetof(uhdrb200_image, luma_stride), "luma_stride");
static_assert(offsetof(ultrahdr_uncompressed_struct, chroma_stride) == offsetof(uhdrb200_image, chroma_stride), "chroma_stride");
static_assert(offsetof(ultrahdr_uncompressed_struct, pixelFormat) == offsetof(uhdrb200_image, pixelFormat), "pixelFormat");
static_assert(sizeof(ultrahdr_color_gamut) == sizeof(int32_t) && sizeof(ultrahdr_pixel_format) == sizeof(int32_t), "enum size");

inline uhdrb200_image* pod(uhdr_uncompressed_ptr p) { return reinterpret_cast<uhdrb200_image*>(p); }

void to_pod(const ultrahdr_metadata_struct& m, uhdrb200_metadata* o) {
  memset(o, 0, sizeof *o);
  strncpy(o->version, m.version.c_str(), sizeof(o->version) - 1);
  if (m.version.size() >= sizeof(o->version)) o->version[0] = '\0';  // cannot equal "1.0": BAD_METADATA, like :369-372
  o->maxContentBoost = m.maxContentBoost;
  o->minContentBoost = m.minContentBoost;
  o->gamma = m.gamma;
  o->offsetSdr = m.offsetSdr;
  o->offsetHdr = m.offsetHdr;
  o->hdrCapacityMin = m.hdrCapacityMin;
  o->hdrCapacityMax = m.hdrCapacityMax;
}

}  // namespace

// ultrahdr.cpp:185-358
status_t UltraHdr::generateGainMap(uhdr_uncompressed_ptr yuv420_image_ptr, uhdr_uncompressed_ptr p010_image_ptr,
                                   ultrahdr_transfer_function hdr_tf, ultrahdr_metadata_ptr metadata,
                                   uhdr_uncompressed_ptr dest, bool sdr_is_601) {
  if (yuv420_image_ptr == nullptr || p010_image_ptr == nullptr || metadata == nullptr || dest == nullptr ||
      yuv420_image_ptr->data == nullptr || yuv420_image_ptr->chroma_data == nullptr ||
      p010_image_ptr->data == nullptr || p010_image_ptr->chroma_data == nullptr) {
    return ERROR_ULTRAHDR_BAD_PTR;
  }
  // the callers adopt dest->data with unique_ptr<uint8_t[]> (jpegr.cpp:207-208): it must come from new[]
  const size_t map_bytes = (yuv420_image_ptr->width / kMapDimensionScaleFactor) * (yuv420_image_ptr->height / kMapDimensionScaleFactor);
  std::unique_ptr<uint8_t[]> map(new (std::nothrow) uint8_t[map_bytes ? map_bytes : 1]);
  if (!map) return ERROR_ULTRAHDR_INSUFFICIENT_RESOURCE;
  uhdrb200_metadata md;
  memset(&md, 0, sizeof md);
  uhdrb200_image out;
  memset(&out, 0, sizeof out);
  const int st = uhdrb200_generate_gainmap_host_into(pod(yuv420_image_ptr), pod(p010_image_ptr), static_cast<int>(hdr_tf), &md,
                                                     &out, sdr_is_601 ? 1 : 0, map.get(), map_bytes);
  if (st != UHDRB200_NO_ERROR) return static_cast<status_t>(st);  // dest and metadata untouched, map freed (:217-218)
  metadata->version = md.version;  // kGainMapVersion, :250
  metadata->maxContentBoost = md.maxContentBoost;
  metadata->minContentBoost = md.minContentBoost;
  metadata->gamma = md.gamma;
  metadata->offsetSdr = md.offsetSdr;
  metadata->offsetHdr = md.offsetHdr;
  metadata->hdrCapacityMin = md.hdrCapacityMin;
  metadata->hdrCapacityMax = md.hdrCapacityMax;
  *pod(dest) = out;            // width, height, strides, gamut UNSPECIFIED, MONOCHROME (:209-216)
  dest->data = map.release();  // ownership passes to the caller (:356)
  return ULTRAHDR_NO_ERROR;
}

// ultrahdr.cpp:360-515
status_t UltraHdr::applyGainMap(uhdr_uncompressed_ptr yuv420_image_ptr, uhdr_uncompressed_ptr gainmap_image_ptr,
                                ultrahdr_metadata_ptr metadata, ultrahdr_output_format output_format,
                                float max_display_boost, uhdr_uncompressed_ptr dest) {
  if (yuv420_image_ptr == nullptr || gainmap_image_ptr == nullptr || metadata == nullptr || dest == nullptr ||
      yuv420_image_ptr->data == nullptr || yuv420_image_ptr->chroma_data == nullptr || gainmap_image_ptr->data == nullptr) {
    return ERROR_ULTRAHDR_BAD_PTR;
  }
  uhdrb200_metadata md;
  to_pod(*metadata, &md);
  return static_cast<status_t>(uhdrb200_apply_gainmap_host(pod(yuv420_image_ptr), pod(gainmap_image_ptr), &md,
                                                           static_cast<int>(output_format), max_display_boost, pod(dest)));
}

// ultrahdr.cpp:517-558
status_t UltraHdr::toneMap(uhdr_uncompressed_ptr src, uhdr_uncompressed_ptr dest) {
  if (src == nullptr || dest == nullptr) return ERROR_ULTRAHDR_BAD_PTR;
  return static_cast<status_t>(uhdrb200_tonemap_host(pod(src), pod(dest)));
}

}  // namespace ultrahdr
