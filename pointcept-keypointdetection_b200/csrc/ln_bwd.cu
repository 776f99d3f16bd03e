// LayerNorm backward for the training step (Block.forward :318-338 has three LayerNorms per block; torch's generic
// kernels spend more time on them than our attention backward).  Same row layout as ln_residual_vec_kernel
// (segment.cu): a row of C = L * 8 * VPT channels is spread over L lanes, 8 contiguous channels per lane and pass.
//   xhat = (x - mean) * rstd                 (mean / rstd recomputed from x: nothing is saved by the forward)
//   dx   = rstd * (dy*gamma - mean_c(dy*gamma) - xhat * mean_c(dy*gamma*xhat))
//   dgamma += sum_rows dy * xhat,  dbeta += sum_rows dy    (per-thread partials -> shared atomics -> global atomics)
#include "common.cuh"

namespace ptv3 {

namespace {
template <typename T>
__device__ __forceinline__ void ld8v(const T* p, float (&v)[8]);
template <>
__device__ __forceinline__ void ld8v<float>(const float* p, float (&v)[8]) {
  float4 a = reinterpret_cast<const float4*>(p)[0], b = reinterpret_cast<const float4*>(p)[1];
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
template <>
__device__ __forceinline__ void ld8v<__nv_bfloat16>(const __nv_bfloat16* p, float (&v)[8]) {
  uint4 u = *reinterpret_cast<const uint4*>(p);
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    v[2 * i] = __uint_as_float(w[i] << 16);
    v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}
}  // namespace

template <typename TX, typename TDY, int L, int VPT>
__global__ void __launch_bounds__(256) ln_bwd_vec_kernel(const TX* __restrict__ x, const TDY* __restrict__ dy,
                                                         const float* __restrict__ gamma, float eps, int64_t n,
                                                         float* __restrict__ dx, float* __restrict__ dgamma,
                                                         float* __restrict__ dbeta) {
  constexpr int C = L * 8 * VPT;
  constexpr int ROWS_PER_WARP = 32 / L;
  __shared__ float red[2 * C];
  for (int e = threadIdx.x; e < 2 * C; e += 256) red[e] = 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31, sub = lane % L;
  const int64_t warp_global = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t warps_total = (int64_t)gridDim.x * (blockDim.x >> 5);
  const float inv_c = 1.f / (float)C;
  float gam[VPT][8], pg[VPT][8], pb[VPT][8];
#pragma unroll
  for (int p = 0; p < VPT; ++p) {
    ld8v<float>(gamma + (p * L + sub) * 8, gam[p]);
#pragma unroll
    for (int i = 0; i < 8; ++i) pg[p][i] = pb[p][i] = 0.f;
  }
  // the loop bound is warp-uniform: lanes of one warp own different rows, and every lane must reach the shuffles
  // (and the __syncthreads() after the loop) together -- rows past n are processed as zeros and not stored
  for (int64_t base = warp_global * ROWS_PER_WARP; base < n; base += warps_total * ROWS_PER_WARP) {
    const int64_t row = base + lane / L;
    const bool valid = row < n;
    float v[VPT][8], g[VPT][8];
    float s = 0.f;
#pragma unroll
    for (int p = 0; p < VPT; ++p) {
      if (valid) {
        ld8v<TX>(x + row * C + (p * L + sub) * 8, v[p]);
        ld8v<TDY>(dy + row * C + (p * L + sub) * 8, g[p]);
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) v[p][i] = g[p][i] = 0.f;
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) s += v[p][i];
    }
#pragma unroll
    for (int d = L / 2; d > 0; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
    const float mean = s * inv_c;
    float q = 0.f;
#pragma unroll
    for (int p = 0; p < VPT; ++p)
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        v[p][i] -= mean;
        q += v[p][i] * v[p][i];
      }
#pragma unroll
    for (int d = L / 2; d > 0; d >>= 1) q += __shfl_xor_sync(0xffffffffu, q, d);
    const float rstd = rsqrtf(q * inv_c + eps);
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int p = 0; p < VPT; ++p)
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        v[p][i] *= rstd;                       // xhat
        pg[p][i] = fmaf(g[p][i], v[p][i], pg[p][i]);
        pb[p][i] += g[p][i];
        g[p][i] *= gam[p][i];                  // dy * gamma
        s1 += g[p][i];
        s2 = fmaf(g[p][i], v[p][i], s2);
      }
#pragma unroll
    for (int d = L / 2; d > 0; d >>= 1) {
      s1 += __shfl_xor_sync(0xffffffffu, s1, d);
      s2 += __shfl_xor_sync(0xffffffffu, s2, d);
    }
    s1 *= inv_c;
    s2 *= inv_c;
#pragma unroll
    for (int p = 0; p < VPT; ++p) {
      float o[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] = rstd * (g[p][i] - s1 - v[p][i] * s2);
      if (valid) {
        float4* dst = reinterpret_cast<float4*>(dx + row * C + (p * L + sub) * 8);
        dst[0] = make_float4(o[0], o[1], o[2], o[3]);
        dst[1] = make_float4(o[4], o[5], o[6], o[7]);
      }
    }
  }
#pragma unroll
  for (int p = 0; p < VPT; ++p)
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      atomicAdd(&red[(p * L + sub) * 8 + i], pg[p][i]);
      atomicAdd(&red[C + (p * L + sub) * 8 + i], pb[p][i]);
    }
  __syncthreads();
  for (int e = threadIdx.x; e < C; e += 256) {
    atomicAdd(dgamma + e, red[e]);
    atomicAdd(dbeta + e, red[C + e]);
  }
}

}  // namespace ptv3

using namespace ptv3;

extern "C" int ptv3_layer_norm_bwd(const void* x, int dtype_x, const void* dy, int dtype_dy, const float* gamma, float eps,
                                   int64_t n, int c, float* dx, float* dgamma, float* dbeta, void* stream) {
  if (n < 0 || (c != 32 && c != 64 && c != 128 && c != 256 && c != 512)) return PTV3_EUNSUPPORTED;
  if (n == 0) return PTV3_OK;
  if (!x || !dy || !gamma || !dx || !dgamma || !dbeta) return PTV3_EINVAL;
  cudaStream_t st = as_stream(stream);
  const int rows_per_cta = 8 * (32 / (c >= 256 ? 32 : c / 8));
  int64_t g = cdiv(n, rows_per_cta);
  const int cap = kNumSMs * 8;
  if (g > cap) g = cap;
#define LAUNCHVEC(TX, TDY, L, VPT) \
  ln_bwd_vec_kernel<TX, TDY, L, VPT><<<(unsigned)g, 256, 0, st>>>((const TX*)x, (const TDY*)dy, gamma, eps, n, dx, dgamma, dbeta)
#define LAUNCHC(TX, TDY)                          \
  do {                                            \
    if (c == 32) LAUNCHVEC(TX, TDY, 4, 1);        \
    else if (c == 64) LAUNCHVEC(TX, TDY, 8, 1);   \
    else if (c == 128) LAUNCHVEC(TX, TDY, 16, 1); \
    else if (c == 256) LAUNCHVEC(TX, TDY, 32, 1); \
    else LAUNCHVEC(TX, TDY, 32, 2);               \
  } while (0)
  if (dtype_x == PTV3_F32 && dtype_dy == PTV3_F32) LAUNCHC(float, float);
  else if (dtype_x == PTV3_F32 && dtype_dy == PTV3_BF16) LAUNCHC(float, __nv_bfloat16);
  else if (dtype_x == PTV3_BF16 && dtype_dy == PTV3_F32) LAUNCHC(__nv_bfloat16, float);
  else if (dtype_x == PTV3_BF16 && dtype_dy == PTV3_BF16) LAUNCHC(__nv_bfloat16, __nv_bfloat16);
  else return PTV3_EINVAL;
#undef LAUNCHC
#undef LAUNCHVEC
  return check_launch();
}
