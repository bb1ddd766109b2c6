// kat.cu — TEST HOOK (include/uhdr_b200_test.h): evaluates the device-side pixel math one function at a time over an
// array, so that the known-answer values of the reference's tests/gainmapmath_test.cpp (:467-473 sRGB, :752-806 HLG / PQ,
// :980-1038 encodeGain, :1229-1262 packers) and dense grids against the oracle reach csrc/uhdr_math.cuh and the
// shared-memory table evaluators of apply_q_common.cuh directly, not only through whole-kernel parity.
#include <cstring>

#include "apply_q_common.cuh"
#include "../../include/uhdr_b200_test.h"

namespace uhdr {
namespace {

__global__ void kat_math_kernel(int fn, const float* __restrict__ in, float* __restrict__ out, size_t n, GainCode gc) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    float r = 0.f;
    switch (fn) {
      case UHDRB200_FN_SRGB_INV_OETF: r = srgb_inv_oetf(in[i]); break;
      case UHDRB200_FN_HLG_OETF: r = hlg_oetf(in[i]); break;
      case UHDRB200_FN_HLG_INV_OETF: r = hlg_inv_oetf(in[i]); break;
      case UHDRB200_FN_PQ_OETF: r = pq_oetf(in[i]); break;
      case UHDRB200_FN_PQ_INV_OETF: r = pq_inv_oetf(in[i]); break;
      case UHDRB200_FN_FLOAT_TO_HALF: r = (float)float_to_half_ref(in[i]); break;
      case UHDRB200_FN_TO_10BIT: r = (float)to_10bit(in[i]); break;
      case UHDRB200_FN_TOE_FIX: r = toe_fix(in[i]); break;
      case UHDRB200_FN_ENCODE_GAIN: r = (float)encode_gain(in[2 * i], in[2 * i + 1], gc); break;
      default: break;
    }
    out[i] = r;
  }
}

// table evaluators exactly as the apply kernels use them: table staged by stage_q_table, looked up by p_eval / l_eval
template <bool LTAB>
__global__ void __launch_bounds__(256) kat_table_kernel(const __grid_constant__ ApplyFastParams p, const float* __restrict__ in,
                                                       float* __restrict__ out, size_t n) {
  const uint32_t hi = stage_q_table<LTAB ? kPackF16 : kPack1010102, LTAB>(p, threadIdx.x);
  constexpr float q0 = 1.0f / (float)(1 << -kPLoExp);
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    out[i] = LTAB ? l_eval(hi, in[i]) : p_eval(hi, in[i] + q0);
}

}  // namespace

cudaError_t launch_kat_impl(int fn, const float* in, float* out, size_t n, const ApplyTables& tab, const GainCode& gc, cudaStream_t s) {
  if (n == 0) return cudaSuccess;
  const int grid = (int)((n + 255) / 256 < 1184 ? (n + 255) / 256 : 1184);
  if (fn != UHDRB200_FN_P_HLG && fn != UHDRB200_FN_P_PQ && fn != UHDRB200_FN_L_POW) {
    kat_math_kernel<<<grid, 256, 0, s>>>(fn, in, out, n, gc);
    return cudaGetLastError();
  }
  ApplyFastParams p;
  memset(&p, 0, sizeof p);
  p.tab = tab;
  p.output_format = fn == UHDRB200_FN_P_HLG ? 3 : 2;
  int dev = 0, reserved = 0;
  cudaError_t e;
  if ((e = cudaGetDevice(&dev)) != cudaSuccess) return e;
  if ((e = cudaDeviceGetAttribute(&reserved, cudaDevAttrReservedSharedMemoryPerBlock, dev)) != cudaSuccess) return e;
  cudaFuncAttributes fa;
  if (fn == UHDRB200_FN_L_POW) {
    const int smem = 16 + (int)(kLEntries * sizeof(float2));
    if ((e = cudaFuncSetAttribute(kat_table_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)) != cudaSuccess) return e;
    kat_table_kernel<true><<<grid, 256, smem, s>>>(p, in, out, n);
  } else {
    if ((e = cudaFuncGetAttributes(&fa, kat_table_kernel<false>)) != cudaSuccess) return e;
    const uint32_t dyn0 = (uint32_t)reserved + (uint32_t)fa.sharedSizeBytes;
    const int smem = (int)(q_table_addr(dyn0) - dyn0 + kPEntries * sizeof(float2)) + 64;
    if ((e = cudaFuncSetAttribute(kat_table_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)) != cudaSuccess) return e;
    kat_table_kernel<false><<<grid, 256, smem, s>>>(p, in, out, n);
  }
  return cudaGetLastError();
}

}  // namespace uhdr
