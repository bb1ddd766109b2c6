// shim_wrap.cpp — TEST INFRASTRUCTURE ONLY. The same C wrapper surface as ref_wrap.cpp
// (`shim_generate_gainmap`, `shim_apply_gainmap`, `shim_tonemap`), but linked against
// shim/ultrahdr_b200_shim.cpp + libuhdr_b200.so instead of the reference's ultrahdr.cpp: the
// reference's class declaration and calling convention on top, the CUDA kernels underneath.
// Proves the drop-in boundary compiles and links against the unmodified reference headers.
#include <cstring>
#include <memory>
#include <string>

#include "ultrahdr/ultrahdr.h"

#include "uhdr_b200.h"

namespace {
struct ShimPath : public ultrahdr::UltraHdr {
  using ultrahdr::UltraHdr::applyGainMap;
  using ultrahdr::UltraHdr::generateGainMap;
  using ultrahdr::UltraHdr::toneMap;
};
ultrahdr::ultrahdr_uncompressed_struct toRef(const uhdrb200_image* im) {
  ultrahdr::ultrahdr_uncompressed_struct r;
  memcpy(&r, im, sizeof r);
  return r;
}
}  // namespace

extern "C" {

int shim_generate_gainmap(const uhdrb200_image* yuv420, const uhdrb200_image* p010, int hdr_tf, uhdrb200_metadata* metadata,
                          uhdrb200_image* dest, int sdr_is_601, void* map_buf, size_t map_buf_bytes) {
  ultrahdr::ultrahdr_uncompressed_struct sdr, hdr, out;
  ultrahdr::ultrahdr_metadata_struct md;
  if (yuv420) sdr = toRef(yuv420);
  if (p010) hdr = toRef(p010);
  out.data = nullptr;
  int st = ShimPath::generateGainMap(yuv420 ? &sdr : nullptr, p010 ? &hdr : nullptr,
                                     static_cast<ultrahdr::ultrahdr_transfer_function>(hdr_tf), metadata ? &md : nullptr,
                                     dest ? &out : nullptr, sdr_is_601 != 0);
  if (st != ultrahdr::ULTRAHDR_NO_ERROR) return st;
  size_t n = out.width * out.height;
  std::unique_ptr<uint8_t[]> owned(reinterpret_cast<uint8_t*>(out.data));  // exactly what jpegr.cpp:207-208 does
  if (map_buf == nullptr || map_buf_bytes < n) return ultrahdr::ERROR_ULTRAHDR_BUFFER_TOO_SMALL;
  memcpy(map_buf, out.data, n);
  out.data = map_buf;
  memcpy(dest, &out, sizeof out);
  memset(metadata->version, 0, sizeof metadata->version);
  strncpy(metadata->version, md.version.c_str(), sizeof(metadata->version) - 1);
  metadata->maxContentBoost = md.maxContentBoost;
  metadata->minContentBoost = md.minContentBoost;
  metadata->gamma = md.gamma;
  metadata->offsetSdr = md.offsetSdr;
  metadata->offsetHdr = md.offsetHdr;
  metadata->hdrCapacityMin = md.hdrCapacityMin;
  metadata->hdrCapacityMax = md.hdrCapacityMax;
  return st;
}

int shim_apply_gainmap(const uhdrb200_image* yuv420, const uhdrb200_image* gainmap, const uhdrb200_metadata* metadata,
                       int output_format, float max_display_boost, uhdrb200_image* dest) {
  ultrahdr::ultrahdr_uncompressed_struct sdr, map, out;
  ultrahdr::ultrahdr_metadata_struct md;
  if (yuv420) sdr = toRef(yuv420);
  if (gainmap) map = toRef(gainmap);
  if (dest) out = toRef(dest);
  if (metadata) {
    md.version = std::string(metadata->version, strnlen(metadata->version, sizeof(metadata->version)));
    md.maxContentBoost = metadata->maxContentBoost;
    md.minContentBoost = metadata->minContentBoost;
    md.gamma = metadata->gamma;
    md.offsetSdr = metadata->offsetSdr;
    md.offsetHdr = metadata->offsetHdr;
    md.hdrCapacityMin = metadata->hdrCapacityMin;
    md.hdrCapacityMax = metadata->hdrCapacityMax;
  }
  ShimPath r;
  int st = r.applyGainMap(yuv420 ? &sdr : nullptr, gainmap ? &map : nullptr, metadata ? &md : nullptr,
                          static_cast<ultrahdr::ultrahdr_output_format>(output_format), max_display_boost, dest ? &out : nullptr);
  if (dest) memcpy(dest, &out, sizeof out);
  return st;
}

int shim_tonemap(const uhdrb200_image* src, uhdrb200_image* dest) {
  ultrahdr::ultrahdr_uncompressed_struct s, d;
  if (src) s = toRef(src);
  if (dest) d = toRef(dest);
  ShimPath r;
  int st = r.toneMap(src ? &s : nullptr, dest ? &d : nullptr);
  if (dest) memcpy(dest, &d, sizeof d);
  return st;
}

}  // extern "C"
