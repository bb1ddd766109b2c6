// sdr_pack.cu — row N3: the SDR output pack and the copy-with-restride of a YUV420 image.
//
// yuv420_to_rgba8888: full-range BT.601 YUV420 -> RGBA8888 with the arithmetic of the reference's
// own routine (examples/ultrahdr_app.cpp:682-734, matrix literals :34-35): nearest chroma sample,
// true divisions by 255.0f, dot products summed left to right in float, * 255 + 0.5, clip to
// [0, 255], truncation, alpha 255. BIT EXACT: this file is compiled with -fmad=false, so every
// `a * b + c` below keeps the two roundings of the x86-64 reference build; the divisions come from a
// 256-entry table of correctly rounded i / 255.0f (division is odd-symmetric, so chroma uses
// |u - 128| and a sign). HBM-bound: 1.5 B/px read, 4 B/px written.
//
// copy_yuv420: jpegr.cpp:315-359 / ultrahdr_api.cpp:437-479 — copy into other strides, zero-filling
// [width, luma_stride) and [width/2, chroma_stride) like the reference's memsets.
#include "uhdr_io.cuh"
#include "uhdr_kernels.cuh"

namespace uhdr {
namespace {

__device__ __forceinline__ float add_rm(float a, float b) {
  float r;
  asm("add.rm.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
}

struct Chroma {  // per 2x2 quad: the chroma terms of the three dot products
  float rv, gu, gv, bu;
};
__device__ __forceinline__ Chroma chroma_terms(const float* sE, uint32_t ub, uint32_t vb) {
  const int ui = (int)ub - 128, vi = (int)vb - 128;
  const float u0 = ui < 0 ? -sE[-ui] : sE[ui];  // float(u - 128) / 255.0f
  const float v0 = vi < 0 ? -sE[-vi] : sE[vi];
  Chroma c;
  c.rv = (float)1.402 * v0;
  c.gu = (float)(-0.202008 / 0.587) * u0;
  c.gv = (float)(-0.419198 / 0.587) * v0;
  c.bu = (float)1.772 * u0;
  return c;
}
// one pixel: (c0*y0 + c1*u0) + c2*v0 with c0 = 1 and the zero coefficients contributing +-0
__device__ __forceinline__ uint32_t pack_pixel(float y0, const Chroma& c) {
  float r = y0 + c.rv;
  float g = (y0 + c.gu) + c.gv;
  float b = y0 + c.bu;
  r = fminf(fmaxf(r * 255.0f + 0.5f, 0.0f), 255.0f);
  g = fminf(fmaxf(g * 255.0f + 0.5f, 0.0f), 255.0f);
  b = fminf(fmaxf(b * 255.0f + 0.5f, 0.0f), 255.0f);
  // truncation of a value in [0, 255] = floor: FADD.RM with 2^23 leaves the integer in the low mantissa bits;
  // the B magic also carries 0xB400 so that its << 16 supplies alpha and cancels the 0x4B000000 of the R word
  const uint32_t rm = __float_as_uint(add_rm(r, 8388608.0f));
  const uint32_t gm = __float_as_uint(add_rm(g, 8388608.0f));
  const uint32_t bm = __float_as_uint(add_rm(b, 8388608.0f + 46080.0f));
  return rm + (gm << 8) + (bm << 16);
}

template <bool VEC>
__global__ void __launch_bounds__(256) yuv420_to_rgba8888_kernel(SdrPackParams p) {
  __shared__ float sE[256];
  sE[threadIdx.x] = __fdiv_rn((float)threadIdx.x, 255.0f);
  __syncthreads();
  const Planes pl = p.batch ? p.batch[blockIdx.z] : p.single;
  const Geometry& g = p.g;
  const uint8_t* py = (const uint8_t*)pl.sdr_y;
  const uint8_t* pu = (const uint8_t*)pl.sdr_u;
  const uint8_t* pv = (const uint8_t*)pl.sdr_v;
  uint32_t* out = (uint32_t*)pl.out;
  const int cx = blockIdx.x * 32 + (threadIdx.x & 31), cy = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (VEC) {  // thread = 4 x 4 pixels; a warp reads 128 contiguous luma bytes and writes 512 contiguous bytes per row
    if (4 * cx >= g.width || 4 * cy >= g.height) return;
    uint32_t yw[4], uw[2], vw[2];
#pragma unroll
    for (int r = 0; r < 4; r++) yw[r] = ldg4(py + (size_t)(4 * cy + r) * g.sdr_luma_stride + 4 * cx);
#pragma unroll
    for (int r = 0; r < 2; r++) {
      uw[r] = ldg2(pu + (size_t)(2 * cy + r) * g.sdr_chroma_stride + 2 * cx);
      vw[r] = ldg2(pv + (size_t)(2 * cy + r) * g.sdr_chroma_stride + 2 * cx);
    }
#pragma unroll
    for (int qr = 0; qr < 2; qr++) {
      Chroma c[2];
#pragma unroll
      for (int q = 0; q < 2; q++) c[q] = chroma_terms(sE, (uw[qr] >> (8 * q)) & 0xff, (vw[qr] >> (8 * q)) & 0xff);
#pragma unroll
      for (int rr = 0; rr < 2; rr++) {
        const int r = 2 * qr + rr;
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; i++) w[i] = pack_pixel(sE[(yw[r] >> (8 * i)) & 0xff], c[i >> 1]);
        stg16(out + (size_t)(4 * cy + r) * g.dst_luma_stride + 4 * cx, make_uint4(w[0], w[1], w[2], w[3]));
      }
    }
  } else {  // thread = one 2x2 quad, scalar accesses, any even width/height and any alignment
    const int qx = cx, qy = cy;
    if (2 * qx >= g.width || 2 * qy >= g.height) return;
    const Chroma c = chroma_terms(sE, pu[(size_t)qy * g.sdr_chroma_stride + qx], pv[(size_t)qy * g.sdr_chroma_stride + qx]);
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int x = 2 * qx + (k & 1), y = 2 * qy + (k >> 1);
      if (x < g.width && y < g.height) out[(size_t)y * g.dst_luma_stride + x] = pack_pixel(sE[py[(size_t)y * g.sdr_luma_stride + x]], c);
    }
  }
}

// flat byte copy of the three planes: index space = dst rows x dst stride, grid.y = plane
__global__ void __launch_bounds__(256) copy_yuv420_kernel(SdrPackParams p) {
  const Planes pl = p.batch ? p.batch[blockIdx.z] : p.single;
  const Geometry& g = p.g;
  const int plane = blockIdx.y;
  const int w = plane ? g.width / 2 : g.width, h = plane ? g.height / 2 : g.height;
  const int ss = plane ? g.sdr_chroma_stride : g.sdr_luma_stride, ds = plane ? g.dst_chroma_stride : g.dst_luma_stride;
  const uint8_t* src = (const uint8_t*)(plane == 0 ? pl.sdr_y : plane == 1 ? pl.sdr_u : pl.sdr_v);
  uint8_t* dst = (uint8_t*)(plane == 0 ? pl.out : plane == 1 ? pl.out_u : pl.out_v);
  const size_t total = (size_t)h * ds;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int row = (int)(i / ds), x = (int)(i - (size_t)row * ds);
    dst[i] = x < w ? src[(size_t)row * ss + x] : 0;
  }
}

bool aligned(const void* q, size_t a) { return (reinterpret_cast<uintptr_t>(q) & (a - 1)) == 0; }
}  // namespace

bool sdr_pack_vec_ok(const Geometry& g, const Planes& pl) {
  return g.width % 4 == 0 && g.height % 4 == 0 && aligned(pl.sdr_y, 4) && aligned(pl.sdr_u, 2) && aligned(pl.sdr_v, 2) &&
         aligned(pl.out, 16) && g.sdr_luma_stride % 4 == 0 && g.sdr_chroma_stride % 2 == 0 && g.dst_luma_stride % 4 == 0;
}

cudaError_t launch_sdr_pack_impl(const SdrPackParams& p, int n, bool vec, cudaStream_t s) {
  if (p.g.width == 0 || p.g.height == 0 || n == 0) return cudaSuccess;
  if (vec) {
    dim3 grid((p.g.width / 4 + 31) / 32, (p.g.height / 4 + 7) / 8, n);
    yuv420_to_rgba8888_kernel<true><<<grid, 256, 0, s>>>(p);
  } else {
    dim3 grid(((p.g.width + 1) / 2 + 31) / 32, ((p.g.height + 1) / 2 + 7) / 8, n);
    yuv420_to_rgba8888_kernel<false><<<grid, 256, 0, s>>>(p);
  }
  return cudaGetLastError();
}

cudaError_t launch_copy_yuv420_impl(const SdrPackParams& p, int n, cudaStream_t s) {
  if (p.g.width == 0 || p.g.height == 0 || n == 0) return cudaSuccess;
  const size_t total = (size_t)p.g.height * p.g.dst_luma_stride;
  unsigned bx = (unsigned)((total + 255) / 256);
  if (bx > 148u * 16u) bx = 148u * 16u;
  dim3 grid(bx, 3, n);
  copy_yuv420_kernel<<<grid, 256, 0, s>>>(p);
  return cudaGetLastError();
}

}  // namespace uhdr
