// sdr_pack.cu — row N3: the SDR output pack and the copy-with-restride of a YUV420 image.
//
// yuv420_to_rgba8888: full-range BT.601 YUV420 -> RGBA8888 with the arithmetic of the reference's
// own routine (examples/ultrahdr_app.cpp:682-734, matrix literals :34-35): nearest chroma sample,
// true divisions by 255.0f, dot products summed left to right in float, * 255 + 0.5, clip to
// [0, 255], truncation, alpha 255. BIT EXACT: this file is compiled with -fmad=false, so every
// `a * b + c` below keeps the two roundings of the x86-64 reference build; the divisions come from a
// shared-memory tables of correctly rounded i / 255.0f. The vector kernel (4-aligned planes) keeps five
// byte-indexed tables (y / 255 and the four chroma products coeff * ((b - 128) / 255)), does the sums,
// the * 255 and the + 0.5 on packed fp32x2 (FADD2 / FMUL2 / FFMA2 by 1: IEEE per lane, two roundings kept) and clips +
// truncates with one saturating F2IP.U8.F32.TRUNC per channel. HBM-bound: 1.5 B/px read, 4 B/px written.
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

// ---- vector kernel: thread = kPackCells vertically adjacent 4x4 cells, packed fp32x2 arithmetic ----
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float lo, float hi) {
  f32x2 r;
  asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpk2(f32x2 v, float& lo, float& hi) { asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {  // two IEEE round-to-nearest additions
  f32x2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
// rn(rn(a * b) + c): ptxas 12.9 contracts mul.rn.f32x2 feeding add.rn.f32x2 into one FFMA2 even under -fmad=false
// (it does not for the scalar forms), so the addition is written as fma(m, 1, c), which is the same sum exactly
__device__ __forceinline__ f32x2 mul_add2(f32x2 a, f32x2 b, f32x2 c, f32x2 one) {
  f32x2 r;
  const f32x2 m = mul2(a, b);
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(m), "l"(one), "l"(c));
  return r;
}
// uint8_t(clip(v, 0, 255)) (ultrahdr_app.cpp:725-731): the float -> u8 conversion saturates and truncates
__device__ __forceinline__ uint32_t clip_u8(float v) {
  uint32_t r;
  asm("cvt.rzi.u8.f32 %0, %1;" : "=r"(r) : "f"(v));
  return r;
}
__device__ __forceinline__ float lds_at(uint32_t addr) {
  float v;
  asm("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
  return v;
}

#ifndef UHDR_PACK_CELLS
#define UHDR_PACK_CELLS 4
#endif
constexpr int kPackCells = UHDR_PACK_CELLS;

__global__ void __launch_bounds__(256) yuv420_to_rgba8888_vec_kernel(SdrPackParams p) {
  // byte-indexed tables: [0] y / 255, [1] 1.402 * c, [2] gu * c, [3] gv * c, [4] 1.772 * c with c = (b - 128) / 255
  __shared__ float sT[5][256];
  {
    const int t = threadIdx.x;
    const float c = __fdiv_rn((float)(t - 128), 255.0f);  // correctly rounded division is odd-symmetric
    sT[0][t] = __fdiv_rn((float)t, 255.0f);
    sT[1][t] = (float)1.402 * c;
    sT[2][t] = (float)(-0.202008 / 0.587) * c;
    sT[3][t] = (float)(-0.419198 / 0.587) * c;
    sT[4][t] = (float)1.772 * c;
  }
  __syncthreads();
  const Planes pl = p.batch ? p.batch[blockIdx.z] : p.single;
  const Geometry& g = p.g;
  const int cx = blockIdx.x * 32 + (threadIdx.x & 31);
  if (4 * cx >= g.width) return;
  const uint32_t tY = (uint32_t)__cvta_generic_to_shared(&sT[0][0]);
  const f32x2 k255 = pk2(255.0f, 255.0f), half = pk2(0.5f, 0.5f), one = pk2(1.0f, 1.0f);
  const uint8_t* py = (const uint8_t*)pl.sdr_y;
  const uint8_t* pu = (const uint8_t*)pl.sdr_u;
  const uint8_t* pv = (const uint8_t*)pl.sdr_v;
  uint32_t* out = (uint32_t*)pl.out;
  // software pipeline: the 8 loads of cell k + 1 are in flight while cell k is converted and stored
  const int cy0 = blockIdx.y * kPackCells * 8 + (threadIdx.x >> 5);  // a block covers 32 pixel rows per step
  if (4 * cy0 >= g.height) return;
  uint32_t nyw[4], nuw[2], nvw[2];
  auto fetch = [&](int cy) {
#pragma unroll
    for (int r = 0; r < 4; r++) nyw[r] = ldg4(py + (size_t)(4 * cy + r) * g.sdr_luma_stride + 4 * cx);
#pragma unroll
    for (int r = 0; r < 2; r++) {
      nuw[r] = ldg2(pu + (size_t)(2 * cy + r) * g.sdr_chroma_stride + 2 * cx);
      nvw[r] = ldg2(pv + (size_t)(2 * cy + r) * g.sdr_chroma_stride + 2 * cx);
    }
  };
  fetch(cy0);
#pragma unroll 1
  for (int k = 0; k < kPackCells; k++) {
    const int cy = cy0 + 8 * k;
    if (4 * cy >= g.height) return;
    uint32_t yw[4], uw[2], vw[2];
#pragma unroll
    for (int r = 0; r < 4; r++) yw[r] = nyw[r];
#pragma unroll
    for (int r = 0; r < 2; r++) {
      uw[r] = nuw[r];
      vw[r] = nvw[r];
    }
    if (k + 1 < kPackCells && 4 * (cy + 8) < g.height) fetch(cy + 8);
#pragma unroll
    for (int qr = 0; qr < 2; qr++) {
      f32x2 RV[2], GU[2], GV[2], BU[2];
#pragma unroll
      for (int q = 0; q < 2; q++) {  // table offsets = byte * 4
        const uint32_t u4 = q ? (uw[qr] >> 6) & 0x3fcu : (uw[qr] << 2) & 0x3fcu;
        const uint32_t v4 = q ? (vw[qr] >> 6) & 0x3fcu : (vw[qr] << 2) & 0x3fcu;
        const float rv = lds_at(tY + 1024u + v4), gu = lds_at(tY + 2048u + u4), gv = lds_at(tY + 3072u + v4), bu = lds_at(tY + 4096u + u4);
        RV[q] = pk2(rv, rv);
        GU[q] = pk2(gu, gu);
        GV[q] = pk2(gv, gv);
        BU[q] = pk2(bu, bu);
      }
#pragma unroll
      for (int rr = 0; rr < 2; rr++) {
        const int r = 2 * qr + rr;
        uint32_t w[4];
#pragma unroll
        for (int q = 0; q < 2; q++) {
          const uint32_t b0 = q ? (yw[r] >> 14) & 0x3fcu : (yw[r] << 2) & 0x3fcu;
          const uint32_t b1 = q ? (yw[r] >> 22) & 0x3fcu : (yw[r] >> 6) & 0x3fcu;
          const f32x2 y2 = pk2(lds_at(tY + b0), lds_at(tY + b1));
          // (c0*y0 + c1*u0) + c2*v0 with c0 = 1 and the zero coefficients contributing +-0 (ultrahdr_app.cpp:712-722)
          const f32x2 R = mul_add2(add2(y2, RV[q]), k255, half, one);
          const f32x2 G = mul_add2(add2(add2(y2, GU[q]), GV[q]), k255, half, one);
          const f32x2 B = mul_add2(add2(y2, BU[q]), k255, half, one);
          float r0, r1, g0, g1, c0, c1;
          unpk2(R, r0, r1);
          unpk2(G, g0, g1);
          unpk2(B, c0, c1);
          w[2 * q] = __byte_perm(__byte_perm(clip_u8(r0), clip_u8(g0), 0x0040), __byte_perm(clip_u8(c0), 0xffu, 0x0040), 0x5410);
          w[2 * q + 1] = __byte_perm(__byte_perm(clip_u8(r1), clip_u8(g1), 0x0040), __byte_perm(clip_u8(c1), 0xffu, 0x0040), 0x5410);
        }
        stg16(out + (size_t)(4 * cy + r) * g.dst_luma_stride + 4 * cx, make_uint4(w[0], w[1], w[2], w[3]));
      }
    }
  }
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
    dim3 grid((p.g.width / 4 + 31) / 32, (p.g.height / 4 + 8 * kPackCells - 1) / (8 * kPackCells), n);
    yuv420_to_rgba8888_vec_kernel<<<grid, 256, 0, s>>>(p);
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
