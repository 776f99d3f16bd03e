// Column sums of an (n, c) matrix -> fp32 (c): the bias gradient of every Linear in the training step (torch's generic
// reduce kernel spent 6.3 ms per step on these).  The matrix is walked as a flat array of 16-byte vectors with a
// grid stride that is a multiple of the vectors per row, so a thread always meets the same 8 (bf16) / 4 (fp32)
// columns and keeps their partial sums in registers; then shared atomics, then one global atomic per column and CTA.
#include "common.cuh"

namespace ptv3 {

template <typename T, int E>   // E = elements per 16-byte vector
__global__ void __launch_bounds__(256) colsum_kernel(const T* __restrict__ x, int64_t n, int c, float* __restrict__ out) {
  extern __shared__ float red[];   // [c]
  for (int e = threadIdx.x; e < c; e += blockDim.x) red[e] = 0.f;
  __syncthreads();
  const int g = c / E;                                      // vectors per row; blockDim.x is a multiple of g
  const int64_t total = n * (int64_t)g;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  float acc[E];
#pragma unroll
  for (int i = 0; i < E; ++i) acc[i] = 0.f;
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < total; v += stride) {
    const uint4 u = reinterpret_cast<const uint4*>(x)[v];
    if constexpr (E == 8) {
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        acc[2 * i] += __uint_as_float(w[i] << 16);
        acc[2 * i + 1] += __uint_as_float(w[i] & 0xffff0000u);
      }
    } else {
      acc[0] += __uint_as_float(u.x);
      acc[1] += __uint_as_float(u.y);
      acc[2] += __uint_as_float(u.z);
      acc[3] += __uint_as_float(u.w);
    }
  }
  const int col = (int)(threadIdx.x % g) * E;
#pragma unroll
  for (int i = 0; i < E; ++i) atomicAdd(&red[col + i], acc[i]);
  __syncthreads();
  for (int e = threadIdx.x; e < c; e += blockDim.x) atomicAdd(out + e, red[e]);
}

}  // namespace ptv3

using namespace ptv3;

extern "C" int ptv3_column_sum(const void* x, int dtype, int64_t n, int c, float* out, void* stream) {
  const int E = dtype == PTV3_BF16 ? 8 : 4;
  if (n < 0 || c < 1 || (dtype != PTV3_BF16 && dtype != PTV3_F32)) return PTV3_EINVAL;
  if (c % E != 0 || c / E > 256) return PTV3_EUNSUPPORTED;
  if (!out) return PTV3_EINVAL;
  cudaStream_t st = as_stream(stream);
  cudaMemsetAsync(out, 0, (size_t)c * 4, st);
  if (n == 0) return PTV3_OK;
  if (!x) return PTV3_EINVAL;
  const int g = c / E;
  const int threads = (256 / g) * g;
  int64_t blocks = cdiv(n * (int64_t)g, (int64_t)threads * 8);
  if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
  if (blocks < 1) blocks = 1;
  if (dtype == PTV3_BF16)
    colsum_kernel<__nv_bfloat16, 8><<<(unsigned)blocks, threads, (size_t)c * 4, st>>>((const __nv_bfloat16*)x, n, c, out);
  else
    colsum_kernel<float, 4><<<(unsigned)blocks, threads, (size_t)c * 4, st>>>((const float*)x, n, c, out);
  return check_launch();
}
