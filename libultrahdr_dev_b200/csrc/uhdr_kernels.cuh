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

struct GenerateParams {
  Geometry g;
  YuvCoefs sdr_k;   // YUV matrix of the SDR image (BT.601 when sdr_is_601, ultrahdr.cpp:284-286)
  YuvCoefs hdr_k;   // YUV matrix of the HDR image (ultrahdr.cpp:288-302)
  LumaCoefs luma;   // luminance weights of the SDR gamut, used for BOTH images (ultrahdr.cpp:265-283)
  Mat3 gamut;       // HDR gamut -> SDR gamut (getHdrConversionFn, gainmapmath.cpp:397-440)
  int use_gamut;    // 0 when both gamuts are equal (identityConversion)
  int hdr_tf;       // 0 linear, 1 HLG, 2 PQ
  float hdr_white_nits;
  float min_boost, max_boost, log2_min, log2_max;
  int code_at_min, code_at_max;  // encodeGain of a gain clamped to min/max, evaluated on the host in double
  Planes single;
  const Planes* batch;  // device array, indexed by blockIdx.z; nullptr => `single`
};

struct ApplyParams {
  Geometry g;
  int output_format;      // UHDRB200_OUTPUT_*
  float log2_min, log2_max;
  float max_boost;        // maxContentBoost (gainmapmath.cpp:553: logBoost * displayBoost / maxContentBoost)
  float inv_display_boost;
  float display_boost;
  const float* idw;       // device: 4 tables x scale^2 x 4 floats (mWeights, NR, NB, C)
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

// *_vec_ok: true when the planes/strides of an image allow the 128-bit vector path (the caller
// ANDs it over a batch). launch_*_impl return the cudaError_t of the launch; `n` = images (grid.z).
bool generate_vec_ok(const Geometry& g, const Planes& pl);
bool apply_vec_ok(const Geometry& g, const Planes& pl);
bool tonemap_vec_ok(const Geometry& g, const Planes& pl);
bool convert_yuv_vec_ok(const Geometry& g, const Planes& pl);
cudaError_t launch_generate_impl(const GenerateParams& p, int n, bool vec, cudaStream_t s);
cudaError_t launch_apply_impl(const ApplyParams& p, int n, bool vec, cudaStream_t s);
cudaError_t launch_tonemap_impl(const ToneMapParams& p, int n, bool vec, cudaStream_t s);
cudaError_t launch_convert_yuv_impl(const ConvertYuvParams& p, int n, bool vec, cudaStream_t s);

}  // namespace uhdr
