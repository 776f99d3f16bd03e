// A6 / A9 glue: y = act(x * scale + shift) per channel in one pass, written as fp32 and / or bf16.
// Used for the eval-mode BatchNorm + GELU that follows the pooled segment-max (SerializedPooling.forward
// point_transformer_v3m1_base.py:433-437) so the (N, C) pooled features are read once and written once per dtype
// instead of four elementwise passes.
#include "common.cuh"

namespace ptv3 {

__device__ __forceinline__ float gelu_exact(float x) { return 0.5f * x * (1.f + erff(x * 0.70710678118654752f)); }

// 4 channels per thread (16-byte fp32 / 8-byte bf16 accesses); c % 4 == 0
template <typename TX>
__global__ void __launch_bounds__(256) affine_act_kernel(const TX* __restrict__ x, const float* __restrict__ scale,
                                                         const float* __restrict__ shift, int act, int64_t n, int c,
                                                         float* __restrict__ out32, __nv_bfloat16* __restrict__ out16) {
  const int64_t total = n * (int64_t)(c >> 2);
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int cq = (int)(t % (c >> 2)) << 2;
    const int64_t e = t << 2;
    float v[4];
    if constexpr (sizeof(TX) == 4) {
      const float4 f = *reinterpret_cast<const float4*>(x + e);
      v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w;
    } else {
      const uint2 u = *reinterpret_cast<const uint2*>(x + e);
      const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&u.x);
      const __nv_bfloat162 b = *reinterpret_cast<const __nv_bfloat162*>(&u.y);
      v[0] = __low2float(a); v[1] = __high2float(a); v[2] = __low2float(b); v[3] = __high2float(b);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float y = v[j];
      if (scale) y *= scale[cq + j];
      if (shift) y += shift[cq + j];
      if (act == 1) y = gelu_exact(y);
      else if (act == 2) y = fmaxf(y, 0.f);
      v[j] = y;
    }
    if (out32) *reinterpret_cast<float4*>(out32 + e) = make_float4(v[0], v[1], v[2], v[3]);
    if (out16) {
      const __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
      uint2 u;
      u.x = *reinterpret_cast<const uint32_t*>(&a);
      u.y = *reinterpret_cast<const uint32_t*>(&b);
      *reinterpret_cast<uint2*>(out16 + e) = u;
    }
  }
}

// exact GELU of a bf16 matrix, bf16 out (may be in place): the hidden activation of the MLP at C >= 512 between the two
// library GEMMs (MLP.forward :237-247).  16-byte accesses, two independent vectors per thread and iteration, and the
// 15-instruction erf of common.cuh (|error| 1.5e-7, invisible after the bf16 rounding) instead of erff(): with erff the
// kernel was instruction-bound at 0.45 of the HBM rate.
__device__ __forceinline__ uint4 gelu8(uint4 u) {
  uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float lo = gelu_erf(__uint_as_float(w[j] << 16)), hi = gelu_erf(__uint_as_float(w[j] & 0xffff0000u));
    const __nv_bfloat162 r = __floats2bfloat162_rn(lo, hi);
    w[j] = *reinterpret_cast<const uint32_t*>(&r);
  }
  return make_uint4(w[0], w[1], w[2], w[3]);
}
__global__ void __launch_bounds__(256) gelu_bf16_kernel(const uint4* __restrict__ x, uint4* __restrict__ y, int64_t n8) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; t + stride < n8; t += 2 * stride) {
    const uint4 a = x[t], b = x[t + stride];
    y[t] = gelu8(a);
    y[t + stride] = gelu8(b);
  }
  if (t < n8) y[t] = gelu8(x[t]);
}

}  // namespace ptv3

using namespace ptv3;

extern "C" int ptv3_affine_act(const void* x, int dtype_x, const float* scale, const float* shift, int act, int64_t n,
                               int c, float* out32, void* out16, void* stream) {
  if (n < 0 || c < 4 || (c & 3) || act < 0 || act > 2) return PTV3_EINVAL;
  if (n == 0) return PTV3_OK;
  if (!x || (!out32 && !out16)) return PTV3_EINVAL;
  const int64_t total = n * (int64_t)(c >> 2);
  const unsigned grid = (unsigned)grid_for(total, 256);
  cudaStream_t st = as_stream(stream);
  if (dtype_x == PTV3_BF16 && act == 1 && !scale && !shift && !out32 && (c & 7) == 0 &&
      (reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(out16) & 15) == 0) {
    const int64_t n8 = n * (int64_t)(c >> 3);
    gelu_bf16_kernel<<<(unsigned)grid_for((n8 + 1) / 2, 256), 256, 0, st>>>((const uint4*)x, (uint4*)out16, n8);
    return check_launch();
  }
  if (dtype_x == PTV3_F32)
    affine_act_kernel<float><<<grid, 256, 0, st>>>((const float*)x, scale, shift, act, n, c, out32,
                                                   (__nv_bfloat16*)out16);
  else if (dtype_x == PTV3_BF16)
    affine_act_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, scale, shift, act, n, c, out32,
                                                           (__nv_bfloat16*)out16);
  else
    return PTV3_EINVAL;
  return check_launch();
}
