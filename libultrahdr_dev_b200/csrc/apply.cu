// apply.cu — A2: UltraHdr::applyGainMap per-pixel loop (reference lib/src/ultrahdr.cpp:427-496)
// on sm_100a: BT.601 YUV->RGB, sRGB^-1, Shepard-IDW gain sampling, applyGain with display boost,
// HLG/PQ OETF and packing to RGBA F16 / RGBA1010102 / planar 10-bit.
//
// Generic kernel (any integer map scale factor): one thread = 4 pixels x 2 rows, so the two chroma
// samples it loads are shared by all 8 pixels, luma comes in as one 32-bit load per row and the
// packed output leaves as 128-bit stores. The four IDW weight tables (gainmapmath.h:184-228) are
// staged in shared memory.
#include "uhdr_kernels.cuh"
#include "uhdr_math.cuh"

namespace uhdr {
namespace {

constexpr int kMaxSmemScale = 16;  // 4 tables * 16*16*4 floats = 16 KB

__device__ __forceinline__ uint32_t ldg4(const void* p) {
  uint32_t r;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ uint32_t ldg2(const void* p) {
  uint16_t r;
  asm volatile("ld.global.nc.L1::no_allocate.u16 %0, [%1];" : "=h"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ void stg16(void* p, uint4 v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void stg8(void* p, uint2 v) {
  asm volatile("st.global.L1::no_allocate.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}

// sampleMap with the IDW tables, gainmapmath.cpp:686-720 (the map is addressed with its width), split in two so that
// neighbouring pixels of one map cell share the four samples: `MapCell` holds e1..e4 = sample / 255.0f (the correctly
// rounded quotients come from a 256-entry shared-memory table, built on the host as (float)i / 255.0f) and the weight
// table the cell selects; `cell_gain` is the left-to-right float sum of gainmapmath.cpp:718-719.
struct MapCell {
  float e1, e2, e3, e4;
  int t;  // 0 mWeights, 1 mWeightsNR, 2 mWeightsNB, 3 mWeightsC
};
__device__ __forceinline__ MapCell load_cell(const uint8_t* __restrict__ map, int mw, int mh, int xl, int yl, const float* e255) {
  const int xu = min(xl + 1, mw - 1), yu = min(yl + 1, mh - 1);
  xl = min(xl, mw - 1);
  yl = min(yl, mh - 1);
  MapCell c;
  c.e1 = e255[__ldg(map + xl + yl * mw)];
  c.e2 = e255[__ldg(map + xl + yu * mw)];
  c.e3 = e255[__ldg(map + xu + yl * mw)];
  c.e4 = e255[__ldg(map + xu + yu * mw)];
  c.t = 0;
  if (xl == xu && yl == yu) c.t = 3;
  else if (xl == xu) c.t = 1;
  else if (yl == yu) c.t = 2;
  return c;
}
__device__ __forceinline__ float cell_gain(const MapCell& c, int s, int ox, int oy, const float* __restrict__ tabs) {
  const float4 w = *reinterpret_cast<const float4*>(tabs + ((c.t * s + oy) * s + ox) * 4);
  return c.e1 * w.x + c.e2 * w.y + c.e3 * w.z + c.e4 * w.w;
}

// one pixel: ultrahdr.cpp:429-451. LUT (numeric mode LUT, the reference's -DUSE_*_LUT=1 build): srgbInvOetfLUT (:433-434),
// applyGainLUT (:446-447, gainmapmath.cpp:557-560) from the two 1024-entry tables in shared memory, then the correctly
// rounded `/ display_boost` (:451).
template <bool LUT>
__device__ __forceinline__ Color3 recover(const ApplyParams& p, const YuvCoefs& k601, uint32_t Y, uint32_t U,
                                          uint32_t V, float gain, const float* s_srgb, const float* s_gain) {
  Color3 yuv = {(float)Y * (1 / 255.0f), (float)((int)U - 128) * (1 / 255.0f), (float)((int)V - 128) * (1 / 255.0f)};
  Color3 rgb = yuv_to_rgb(k601, yuv);
  if (LUT) {
    const float f = s_gain[lut_index(gain, kLutGain)];
    const Color3 lin = {s_srgb[lut_index(rgb.r, kLutSrgbInv)], s_srgb[lut_index(rgb.g, kLutSrgbInv)], s_srgb[lut_index(rgb.b, kLutSrgbInv)]};
    return {__fdiv_rn(lin.r * f, p.display_boost), __fdiv_rn(lin.g * f, p.display_boost), __fdiv_rn(lin.b * f, p.display_boost)};
  }
  Color3 lin = {srgb_inv_oetf(rgb.r), srgb_inv_oetf(rgb.g), srgb_inv_oetf(rgb.b)};
  // applyGain(e, gain, metadata, displayBoost), gainmapmath.cpp:550-555
  float log_boost = p.log2_min * (1.0f - gain) + p.log2_max * gain;
  float f = mufu_ex2(__fdiv_rn(log_boost * p.display_boost, p.max_boost)) * p.inv_display_boost;  // "/ display_boost", :451
  return {lin.r * f, lin.g * f, lin.b * f};
}

__device__ __forceinline__ uint32_t pack1010102(Color3 e) {
  return to_10bit(e.r) | (to_10bit(e.g) << 10) | (to_10bit(e.b) << 20) | 0xC0000000u;
}

template <bool VEC, bool LUT>
__global__ void __launch_bounds__(256) apply_kernel(ApplyParams p) {
  extern __shared__ float s_idw[];  // [IDW tables when they fit][i / 255.0f][LUT: srgbInvOetf table | gain table]
  const Geometry& g = p.g;
  const int s = g.map_scale;
  const float* tabs = p.idw;
  const int n_idw = s <= kMaxSmemScale ? 16 * s * s : 0;
  const float* s_e255 = s_idw + n_idw;
  const float* s_srgb = s_e255 + 256;
  const float* s_gain = s_srgb + kLutSrgbInv;
  const int tid = threadIdx.y * 32 + threadIdx.x;
  if (n_idw) {
    for (int i = tid; i < n_idw; i += 256) s_idw[i] = p.idw[i];
    tabs = s_idw;
  }
  s_idw[n_idw + tid] = p.e255[tid];
  if (LUT) {
    for (int i = tid; i < kLutSrgbInv; i += 256) s_idw[n_idw + 256 + i] = p.lt.srgb_inv[i];
    for (int i = tid; i < kLutGain; i += 256) s_idw[n_idw + 256 + kLutSrgbInv + i] = p.gain_lut[i];
  }
  __syncthreads();
  const int x0 = (blockIdx.x * 32 + threadIdx.x) * 4;
  const int y0 = (blockIdx.y * 8 + threadIdx.y) * 2;
  if (x0 >= g.width || y0 >= g.height) return;
  const Planes pl = p.batch ? p.batch[blockIdx.z] : p.single;
  const YuvCoefs k601 = {1.402f, 0.114f * 1.772f / 0.587f, 0.299f * 1.402f / 0.587f, 1.772f};  // always BT.601, ultrahdr.cpp:430-431
  const uint8_t* map = (const uint8_t*)pl.map_in;
  const size_t npix = (size_t)g.width * g.height;

  uint32_t Y[2][4], U[2], V[2];
  const uint8_t* yp = (const uint8_t*)pl.sdr_y + (size_t)y0 * g.sdr_luma_stride + x0;
  const uint8_t* up = (const uint8_t*)pl.sdr_u + (size_t)(y0 >> 1) * g.sdr_chroma_stride + (x0 >> 1);
  const uint8_t* vp = (const uint8_t*)pl.sdr_v + (size_t)(y0 >> 1) * g.sdr_chroma_stride + (x0 >> 1);
  const int nx = VEC ? 4 : min(4, g.width - x0), ny = VEC ? 2 : min(2, g.height - y0);
  if (VEC) {
    uint32_t a = ldg4(yp), b = ldg4(yp + g.sdr_luma_stride), c = ldg2(up), d = ldg2(vp);
#pragma unroll
    for (int i = 0; i < 4; i++) { Y[0][i] = (a >> (8 * i)) & 0xff; Y[1][i] = (b >> (8 * i)) & 0xff; }
    U[0] = c & 0xff; U[1] = c >> 8; V[0] = d & 0xff; V[1] = d >> 8;
  } else {
#pragma unroll
    for (int r = 0; r < 2; r++)
#pragma unroll
      for (int i = 0; i < 4; i++) Y[r][i] = (r < ny && i < nx) ? yp[(size_t)r * g.sdr_luma_stride + i] : 0;
#pragma unroll
    for (int i = 0; i < 2; i++) { U[i] = (2 * i < nx) ? up[i] : 128; V[i] = (2 * i < nx) ? vp[i] : 128; }
  }

  // map cell of the thread's first pixel: the only two integer divisions; the other seven pixels step from it
  const int xl0 = x0 / s, yl0 = y0 / s;
  const int ox0 = x0 - xl0 * s;
  int yl = yl0, oy = y0 - yl0 * s;
#pragma unroll
  for (int r = 0; r < 2; r++) {
    if (!VEC && r >= ny) break;
    const int y = y0 + r;
    Color3 e[4];
    int xl = xl0, ox = ox0;
    MapCell cell = load_cell(map, g.map_width, g.map_height, xl, yl, s_e255);
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const float gain = cell_gain(cell, s, ox, oy, tabs);
      e[i] = recover<LUT>(p, k601, Y[r][i], U[i >> 1], V[i >> 1], gain, s_srgb, s_gain);
      if (i < 3 && ++ox == s) {  // the next pixel starts a new cell (columns past the right edge are never stored)
        ox = 0;
        cell = load_cell(map, g.map_width, g.map_height, ++xl, yl, s_e255);
      }
    }
    if (++oy == s) {
      oy = 0;
      ++yl;
    }
    const size_t pix = (size_t)y * g.width + x0;
    switch (p.output_format) {
      case 1: {  // HDR_LINEAR -> RGBA F16, ultrahdr.cpp:455-459
        uint32_t w[8];
#pragma unroll
        for (int i = 0; i < 4; i++) {
          w[2 * i] = float_to_half_ref(e[i].r) | (float_to_half_ref(e[i].g) << 16);
          w[2 * i + 1] = float_to_half_ref(e[i].b) | (0x3C00u << 16);
        }
        uint64_t* o = (uint64_t*)pl.out + pix;
        if (VEC) {
          stg16(o, make_uint4(w[0], w[1], w[2], w[3]));
          stg16(o + 2, make_uint4(w[4], w[5], w[6], w[7]));
        } else {
          for (int i = 0; i < nx; i++) o[i] = (uint64_t)w[2 * i] | ((uint64_t)w[2 * i + 1] << 32);
        }
        break;
      }
      case 4: {  // HDR_LINEAR_RGB_10BIT -> three planar u16 planes, ultrahdr.cpp:460-468
        uint16_t* o = (uint16_t*)pl.out + pix;
        uint32_t c[3][4];
#pragma unroll
        for (int i = 0; i < 4; i++) { c[0][i] = to_10bit(e[i].r); c[1][i] = to_10bit(e[i].g); c[2][i] = to_10bit(e[i].b); }
#pragma unroll
        for (int ch = 0; ch < 3; ch++) {
          if (VEC) stg8(o + ch * npix, make_uint2(c[ch][0] | (c[ch][1] << 16), c[ch][2] | (c[ch][3] << 16)));
          else for (int i = 0; i < nx; i++) o[ch * npix + i] = (uint16_t)c[ch][i];
        }
        break;
      }
      case 3:    // HDR_HLG -> RGBA1010102, ultrahdr.cpp:469-479
      case 2: {  // HDR_PQ  -> RGBA1010102, ultrahdr.cpp:480-490
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
          Color3 q;
          if (LUT) {  // hlgOetfLUT / pqOetfLUT, 65 536 entries each (ultrahdr.cpp:470-485): L1 / L2 resident
            const float* t = p.output_format == 3 ? p.lt.hlg_oetf : p.lt.pq_oetf;
            q = {__ldg(t + lut_index(e[i].r, kLutHlgOetf)), __ldg(t + lut_index(e[i].g, kLutHlgOetf)), __ldg(t + lut_index(e[i].b, kLutHlgOetf))};
          } else {
            q = p.output_format == 3 ? Color3{hlg_oetf(e[i].r), hlg_oetf(e[i].g), hlg_oetf(e[i].b)}
                                     : Color3{pq_oetf(e[i].r), pq_oetf(e[i].g), pq_oetf(e[i].b)};
          }
          w[i] = pack1010102(q);
        }
        uint32_t* o = (uint32_t*)pl.out + pix;
        if (VEC) stg16(o, make_uint4(w[0], w[1], w[2], w[3]));
        else for (int i = 0; i < nx; i++) o[i] = w[i];
        break;
      }
      default: break;  // ultrahdr.cpp:491-493: unknown format writes nothing
    }
  }
}

bool aligned(const void* q, size_t a) { return (reinterpret_cast<uintptr_t>(q) & (a - 1)) == 0; }
}  // namespace

bool apply_vec_ok(const Geometry& g, const Planes& pl) {
  return g.width % 4 == 0 && g.height % 2 == 0 && aligned(pl.sdr_y, 4) && aligned(pl.sdr_u, 2) &&
         aligned(pl.sdr_v, 2) && aligned(pl.out, 16) && g.sdr_luma_stride % 4 == 0 &&
         g.sdr_chroma_stride % 2 == 0 && ((size_t)g.width * g.height * 2) % 8 == 0;
}

cudaError_t launch_apply_impl(const ApplyParams& p, int n, bool vec, cudaStream_t s) {
  if (p.g.width == 0 || p.g.height == 0 || n == 0) return cudaSuccess;
  dim3 block(32, 8, 1);
  dim3 grid((p.g.width + 127) / 128, (p.g.height + 15) / 16, n);
  size_t smem = (p.g.map_scale <= kMaxSmemScale ? sizeof(float) * 16 * p.g.map_scale * p.g.map_scale : 0) + sizeof(float) * 256;
  if (p.lut) {
    smem += sizeof(float) * (kLutSrgbInv + kLutGain);
    if (vec) apply_kernel<true, true><<<grid, block, smem, s>>>(p);
    else apply_kernel<false, true><<<grid, block, smem, s>>>(p);
    return cudaGetLastError();
  }
  if (vec) apply_kernel<true, false><<<grid, block, smem, s>>>(p);
  else apply_kernel<false, false><<<grid, block, smem, s>>>(p);
  return cudaGetLastError();
}

}  // namespace uhdr
