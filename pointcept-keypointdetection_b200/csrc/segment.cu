// A6/A7 floating point: CSR segment reductions (torch_scatter.segment_csr
// semantics), unpool gather-add with fused GELU, fused residual LayerNorm.
// HBM-bound: 16-byte vector loads along the channel dimension, one pass.
#include "common.cuh"

namespace ptv3 {

enum { RED_SUM = 0, RED_MEAN = 1, RED_MAX = 2, RED_MIN = 3 };

__device__ __forceinline__ float red_init(int reduce) {
  return reduce == RED_MAX ? -INFINITY : reduce == RED_MIN ? INFINITY : 0.f;
}
__device__ __forceinline__ float red_op(int reduce, float a, float b) {
  return reduce == RED_MAX ? fmaxf(a, b) : reduce == RED_MIN ? fminf(a, b) : a + b;
}

// 4 channels per thread
template <typename T>
__device__ __forceinline__ void load4(const T* p, float v[4]);
template <>
__device__ __forceinline__ void load4<float>(const float* p, float v[4]) {
  float4 t = *reinterpret_cast<const float4*>(p);
  v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
template <>
__device__ __forceinline__ void load4<__nv_bfloat16>(const __nv_bfloat16* p, float v[4]) {
  uint2 t = *reinterpret_cast<const uint2*>(p);
  v[0] = __uint_as_float(t.x << 16); v[1] = __uint_as_float(t.x & 0xffff0000u);
  v[2] = __uint_as_float(t.y << 16); v[3] = __uint_as_float(t.y & 0xffff0000u);
}
template <typename T>
__device__ __forceinline__ void store4(T* p, const float v[4]);
template <>
__device__ __forceinline__ void store4<float>(float* p, const float v[4]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
}
template <>
__device__ __forceinline__ void store4<__nv_bfloat16>(__nv_bfloat16* p, const float v[4]) {
  __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
  uint2 t;
  t.x = *reinterpret_cast<uint32_t*>(&a);
  t.y = *reinterpret_cast<uint32_t*>(&b);
  *reinterpret_cast<uint2*>(p) = t;
}

// short segments: one thread per (segment, 4-channel group); threads of a warp
// cover consecutive channel groups (coalesced rows).  c % 4 == 0.
template <typename T>
__global__ void __launch_bounds__(256) seg_short_vec4_kernel(const T* __restrict__ src,
                                                             const int64_t* __restrict__ gather,
                                                             const int64_t* __restrict__ indptr, int64_t m, int c4,
                                                             int reduce, T* __restrict__ out) {
  int64_t total = m * c4;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    int64_t k = t / c4;
    int g = (int)(t - k * c4);
    int64_t a = indptr[k], b = indptr[k + 1];
    float acc[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[i] = red_init(reduce);
    for (int64_t j = a; j < b; ++j) {
      int64_t row = gather ? gather[j] : j;
      float v[4];
      load4<T>(src + row * (int64_t)c4 * 4 + g * 4, v);
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[i] = red_op(reduce, acc[i], v[i]);
    }
    if (b <= a) {
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[i] = 0.f;
    } else if (reduce == RED_MEAN) {
      float inv = 1.f / (float)(b - a);
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[i] *= inv;
    }
    store4<T>(out + k * (int64_t)c4 * 4 + g * 4, acc);
  }
}

// short segments, any channel count (c = 3 for coordinates): one thread per (segment, channel).
template <typename T>
__global__ void __launch_bounds__(256) seg_short_scalar_kernel(const T* __restrict__ src,
                                                               const int64_t* __restrict__ gather,
                                                               const int64_t* __restrict__ indptr, int64_t m, int c,
                                                               int reduce, T* __restrict__ out) {
  const int64_t total = m * c;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t k = t / c;
    const int ch = (int)(t - k * c);
    const int64_t a = indptr[k], b = indptr[k + 1];
    float acc = red_init(reduce);
    for (int64_t j = a; j < b; ++j) {
      const int64_t row = gather ? gather[j] : j;
      acc = red_op(reduce, acc, Elem<T>::load(src + row * c + ch));
    }
    if (b <= a) acc = 0.f;
    else if (reduce == RED_MEAN) acc *= 1.f / (float)(b - a);
    Elem<T>::store(out + t, acc);
  }
}

// generic (any c, any segment length): one CTA per segment; warp w strides the
// rows, lanes stride the channels; cross-warp combine through shared memory.
template <typename T>
__global__ void __launch_bounds__(256) seg_block_kernel(const T* __restrict__ src, const int64_t* __restrict__ gather,
                                                        const int64_t* __restrict__ indptr, int64_t m, int c,
                                                        int reduce, T* __restrict__ out) {
  __shared__ float part[8][33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int64_t k = blockIdx.x; k < m; k += gridDim.x) {
    int64_t a = indptr[k], b = indptr[k + 1];
    for (int c0 = 0; c0 < c; c0 += 32) {
      int ch = c0 + lane;
      float acc = red_init(reduce);
      if (ch < c)
        for (int64_t j = a + warp; j < b; j += 8) {
          int64_t row = gather ? gather[j] : j;
          acc = red_op(reduce, acc, Elem<T>::load(src + row * c + ch));
        }
      part[warp][lane] = acc;
      __syncthreads();
      if (warp == 0 && ch < c) {
        float r = part[0][lane];
#pragma unroll
        for (int w = 1; w < 8; ++w) r = red_op(reduce, r, part[w][lane]);
        if (b <= a) r = 0.f;
        else if (reduce == RED_MEAN) r /= (float)(b - a);
        Elem<T>::store(out + k * c + ch, r);
      }
      __syncthreads();
    }
  }
}

__device__ __forceinline__ float gelu_libm(float x) { return 0.5f * x * (1.f + erff(x * 0.70710678118654752440f)); }

template <typename T, bool GELU>
__global__ void __launch_bounds__(256) unpool_add_kernel(const T* __restrict__ skip, const T* __restrict__ up,
                                                         const int64_t* __restrict__ inverse, int64_t n, int c4,
                                                         float* __restrict__ out, T* __restrict__ skip_act,
                                                         bool second_is_sum) {
  int64_t total = n * c4;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    int64_t i = t / c4;
    int g = (int)(t - i * c4);
    float a[4], b[4], o[4];
    load4<T>(skip + t * 4, a);
    load4<T>(up + (inverse[i] * c4 + g) * 4, b);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      if (GELU) {
        a[q] = gelu_libm(a[q]);
        b[q] = gelu_libm(b[q]);
      }
      o[q] = a[q] + b[q];
    }
    store4<float>(out + t * 4, o);
    if (skip_act) store4<T>(skip_act + t * 4, second_is_sum ? o : a);
  }
}

// One warp per row, MAXV channels per lane (c <= 32*MAXV).
//   t = gamma ? LayerNorm(x; gamma, beta) : x ;  t *= row_scale[row] (optional) ;  r = residual ? residual + t : t
//   res_out = r (fp32, optional) ;  out2 = LayerNorm(r; gamma2, beta2) (optional)
template <typename TX, typename TO, int MAXV>
__global__ void __launch_bounds__(256) ln_residual_kernel(const TX* __restrict__ x, const float* __restrict__ gamma,
                                                          const float* __restrict__ beta, float eps,
                                                          const float* __restrict__ residual,
                                                          float* __restrict__ res_out, const float* __restrict__ gamma2,
                                                          const float* __restrict__ beta2, TO* __restrict__ out2,
                                                          int64_t n, int c, const float* __restrict__ row_scale) {
  const int lane = threadIdx.x & 31;
  int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t stride = (int64_t)gridDim.x * (blockDim.x >> 5);
  const float inv_c = 1.f / (float)c;
  for (; row < n; row += stride) {
    float v[MAXV];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      int ch = i * 32 + lane;
      v[i] = ch < c ? Elem<TX>::load(x + row * c + ch) : 0.f;
      s += v[i];
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
    float mean = s * inv_c, q = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      int ch = i * 32 + lane;
      float dlt = ch < c ? v[i] - mean : 0.f;
      q += dlt * dlt;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) q += __shfl_xor_sync(0xffffffffu, q, d);
    float rstd = rsqrtf(q * inv_c + eps);
    if (gamma == nullptr) { mean = 0.f; rstd = 1.f; }
    const float rs = row_scale ? row_scale[row] : 1.f;     // DropPath: 0 or 1 / keep for the whole row of the branch
    float s2 = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      int ch = i * 32 + lane;
      if (ch < c) {
        float y = (v[i] - mean) * rstd;
        if (gamma) y = y * gamma[ch] + beta[ch];
        y *= rs;
        if (residual) y += residual[row * c + ch];
        v[i] = y;
        if (res_out) res_out[row * c + ch] = y;
        s2 += y;
      }
    }
    if (out2 && gamma2 == nullptr) {
#pragma unroll
      for (int i = 0; i < MAXV; ++i) {
        int ch = i * 32 + lane;
        if (ch < c) Elem<TO>::store(out2 + row * c + ch, v[i]);
      }
    } else if (out2) {
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) s2 += __shfl_xor_sync(0xffffffffu, s2, d);
      float mean2 = s2 * inv_c, q2 = 0.f;
#pragma unroll
      for (int i = 0; i < MAXV; ++i) {
        int ch = i * 32 + lane;
        float dlt = ch < c ? v[i] - mean2 : 0.f;
        q2 += dlt * dlt;
      }
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) q2 += __shfl_xor_sync(0xffffffffu, q2, d);
      float rstd2 = rsqrtf(q2 * inv_c + eps);
#pragma unroll
      for (int i = 0; i < MAXV; ++i) {
        int ch = i * 32 + lane;
        if (ch < c) Elem<TO>::store(out2 + row * c + ch, (v[i] - mean2) * rstd2 * gamma2[ch] + beta2[ch]);
      }
    }
  }
}


// ---- vectorised variant: every thread owns 8 contiguous channels (16 B of bf16 / 32 B of fp32), a row is spread
// over L = c/8 lanes (VPT = 2: 16 channels per thread for c = 512) and reduced with xor-shuffles inside the warp.
template <typename T>
__device__ __forceinline__ void ld8(const T* p, float (&v)[8]);
template <>
__device__ __forceinline__ void ld8<float>(const float* p, float (&v)[8]) {
  float4 a = reinterpret_cast<const float4*>(p)[0], b = reinterpret_cast<const float4*>(p)[1];
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
template <>
__device__ __forceinline__ void ld8<__nv_bfloat16>(const __nv_bfloat16* p, float (&v)[8]) {
  uint4 u = *reinterpret_cast<const uint4*>(p);
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    v[2 * i] = __uint_as_float(w[i] << 16);
    v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}
template <typename T>
__device__ __forceinline__ void st8(T* p, const float (&v)[8]);
template <>
__device__ __forceinline__ void st8<float>(float* p, const float (&v)[8]) {
  reinterpret_cast<float4*>(p)[0] = make_float4(v[0], v[1], v[2], v[3]);
  reinterpret_cast<float4*>(p)[1] = make_float4(v[4], v[5], v[6], v[7]);
}
template <>
__device__ __forceinline__ void st8<__nv_bfloat16>(__nv_bfloat16* p, const float (&v)[8]) {
  uint32_t w[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    w[i] = *reinterpret_cast<uint32_t*>(&h);
  }
  *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
}

template <typename TX, typename TO, int L, int VPT>
__global__ void __launch_bounds__(256) ln_residual_vec_kernel(const TX* __restrict__ x, const float* __restrict__ gamma,
                                                              const float* __restrict__ beta, float eps,
                                                              const float* __restrict__ residual,
                                                              float* __restrict__ res_out,
                                                              const float* __restrict__ gamma2,
                                                              const float* __restrict__ beta2, TO* __restrict__ out2,
                                                              int64_t n, const float* __restrict__ row_scale) {
  constexpr int C = L * 8 * VPT;
  constexpr int ROWS_PER_WARP = 32 / L;
  const int lane = threadIdx.x & 31, sub = lane % L;
  const int64_t warp_global = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t warps_total = (int64_t)gridDim.x * (blockDim.x >> 5);
  const float inv_c = 1.f / (float)C;
  for (int64_t row = warp_global * ROWS_PER_WARP + lane / L; row < n; row += warps_total * ROWS_PER_WARP) {
    float v[VPT][8];
    float s = 0.f;
#pragma unroll
    for (int p = 0; p < VPT; ++p) {
      ld8<TX>(x + row * C + (p * L + sub) * 8, v[p]);
#pragma unroll
      for (int i = 0; i < 8; ++i) s += v[p][i];
    }
    if (gamma) {
#pragma unroll
      for (int d = L / 2; d > 0; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
      const float mean = s * inv_c;
      float q = 0.f;
#pragma unroll
      for (int p = 0; p < VPT; ++p)
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          float d0 = v[p][i] - mean;
          q += d0 * d0;
        }
#pragma unroll
      for (int d = L / 2; d > 0; d >>= 1) q += __shfl_xor_sync(0xffffffffu, q, d);
      const float rstd = rsqrtf(q * inv_c + eps);
#pragma unroll
      for (int p = 0; p < VPT; ++p) {
        float g[8], b[8];
        ld8<float>(gamma + (p * L + sub) * 8, g);
        ld8<float>(beta + (p * L + sub) * 8, b);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[p][i] = (v[p][i] - mean) * rstd * g[i] + b[i];
      }
    }
    if (row_scale) {
      const float rs = row_scale[row];
#pragma unroll
      for (int p = 0; p < VPT; ++p)
#pragma unroll
        for (int i = 0; i < 8; ++i) v[p][i] *= rs;
    }
    float s2 = 0.f;
#pragma unroll
    for (int p = 0; p < VPT; ++p) {
      if (residual) {
        float r[8];
        ld8<float>(residual + row * C + (p * L + sub) * 8, r);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[p][i] += r[i];
      }
      if (res_out) st8<float>(res_out + row * C + (p * L + sub) * 8, v[p]);
#pragma unroll
      for (int i = 0; i < 8; ++i) s2 += v[p][i];
    }
    if (out2 == nullptr) continue;
    if (gamma2) {
#pragma unroll
      for (int d = L / 2; d > 0; d >>= 1) s2 += __shfl_xor_sync(0xffffffffu, s2, d);
      const float mean2 = s2 * inv_c;
      float q2 = 0.f;
#pragma unroll
      for (int p = 0; p < VPT; ++p)
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          float d0 = v[p][i] - mean2;
          q2 += d0 * d0;
        }
#pragma unroll
      for (int d = L / 2; d > 0; d >>= 1) q2 += __shfl_xor_sync(0xffffffffu, q2, d);
      const float rstd2 = rsqrtf(q2 * inv_c + eps);
#pragma unroll
      for (int p = 0; p < VPT; ++p) {
        float g[8], b[8];
        ld8<float>(gamma2 + (p * L + sub) * 8, g);
        ld8<float>(beta2 + (p * L + sub) * 8, b);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[p][i] = (v[p][i] - mean2) * rstd2 * g[i] + b[i];
      }
    }
#pragma unroll
    for (int p = 0; p < VPT; ++p) st8<TO>(out2 + row * C + (p * L + sub) * 8, v[p]);
  }
}

template <typename T>
static int seg_dispatch(const T* src, const int64_t* gather, const int64_t* indptr, int64_t m, int64_t c, int reduce,
                        T* out, int64_t hint_rows_per_seg, cudaStream_t st) {
  if ((c % 4) == 0 && hint_rows_per_seg <= 64) {
    int c4 = (int)(c / 4);
    seg_short_vec4_kernel<T><<<grid_for(m * c4, 256), 256, 0, st>>>(src, gather, indptr, m, c4, reduce, out);
  } else if (hint_rows_per_seg <= 64) {
    seg_short_scalar_kernel<T><<<grid_for(m * c, 256), 256, 0, st>>>(src, gather, indptr, m, (int)c, reduce, out);
  } else {
    int g = (int)(m < (int64_t)kNumSMs * 16 ? m : (int64_t)kNumSMs * 16);
    seg_block_kernel<T><<<g, 256, 0, st>>>(src, gather, indptr, m, (int)c, reduce, out);
  }
  return check_launch();
}

}  // namespace ptv3

using namespace ptv3;

extern "C" {

// `m < 0` selects the long-segment kernel explicitly (|m| segments): used by scatter_mean over whole scenes.
int ptv3_segment_csr(const void* src, int dtype, const int64_t* gather, const int64_t* indptr, int64_t m, int64_t c,
                     int reduce, void* out, void* stream) {
  int64_t hint = 1;
  if (m < 0) {
    m = -m;
    hint = 1 << 20;
  }
  if (c < 1 || reduce < 0 || reduce > 3 || !indptr) return PTV3_EINVAL;
  if (m == 0) return PTV3_OK;
  if (!src || !out) return PTV3_EINVAL;
  if (dtype == PTV3_F32)
    return seg_dispatch<float>((const float*)src, gather, indptr, m, c, reduce, (float*)out, hint, as_stream(stream));
  if (dtype == PTV3_BF16)
    return seg_dispatch<__nv_bfloat16>((const __nv_bfloat16*)src, gather, indptr, m, c, reduce, (__nv_bfloat16*)out,
                                       hint, as_stream(stream));
  return PTV3_EINVAL;
}

int ptv3_unpool_add(const void* skip, const void* up, int dtype, const int64_t* inverse, int64_t n, int64_t c, int gelu,
                    void* out, void* skip_act, void* stream) {
  if (n < 0 || c < 4 || (c % 4) != 0 || !inverse) return PTV3_EINVAL;
  if (n == 0) return PTV3_OK;
  int c4 = (int)(c / 4);
  int g = grid_for(n * c4, 256);
  cudaStream_t st = as_stream(stream);
  const bool sum2 = (gelu & 2) != 0;     // bit 1: the second output holds the sum (PT-v3m2) instead of act(skip) (PT-v3m1)
  gelu &= 1;
  if (dtype == PTV3_F32) {
    if (gelu)
      unpool_add_kernel<float, true><<<g, 256, 0, st>>>((const float*)skip, (const float*)up, inverse, n, c4,
                                                        (float*)out, (float*)skip_act, sum2);
    else
      unpool_add_kernel<float, false><<<g, 256, 0, st>>>((const float*)skip, (const float*)up, inverse, n, c4,
                                                         (float*)out, (float*)skip_act, sum2);
  } else if (dtype == PTV3_BF16) {
    if (gelu)
      unpool_add_kernel<__nv_bfloat16, true><<<g, 256, 0, st>>>((const __nv_bfloat16*)skip, (const __nv_bfloat16*)up,
                                                                inverse, n, c4, (float*)out,
                                                                (__nv_bfloat16*)skip_act, sum2);
    else
      unpool_add_kernel<__nv_bfloat16, false><<<g, 256, 0, st>>>((const __nv_bfloat16*)skip, (const __nv_bfloat16*)up,
                                                                 inverse, n, c4, (float*)out,
                                                                 (__nv_bfloat16*)skip_act, sum2);
  } else {
    return PTV3_EINVAL;
  }
  return check_launch();
}

int ptv3_ln_residual(const void* x, int dtype_x, const float* gamma, const float* beta, float eps,
                     const float* residual, float* res_out, const float* gamma2, const float* beta2, void* out2,
                     int dtype_out2, int64_t n, int c, void* stream) {
  return ptv3_ln_residual_rs(x, dtype_x, gamma, beta, eps, nullptr, residual, res_out, gamma2, beta2, out2, dtype_out2,
                             n, c, stream);
}

int ptv3_ln_residual_rs(const void* x, int dtype_x, const float* gamma, const float* beta, float eps,
                        const float* row_scale, const float* residual, float* res_out, const float* gamma2,
                        const float* beta2, void* out2, int dtype_out2, int64_t n, int c, void* stream) {
  if (n < 0 || c < 1 || c > 1024 || (gamma && !beta)) return PTV3_EINVAL;
  if (gamma2 && !beta2) return PTV3_EINVAL;
  if (n == 0) return PTV3_OK;
  int rows_per_cta = 8;
  int g = (int)cdiv(n, rows_per_cta);
  int cap = kNumSMs * 8 * 4;
  if (g > cap) g = cap;
  cudaStream_t st = as_stream(stream);
  if (c == 32 || c == 64 || c == 128 || c == 256 || c == 512 || c == 768) {
    // rows are spread over c/8 lanes; enough CTAs for ~8 resident warps of loads per SM sub-partition
    const int l = c >= 512 ? 32 : c / 8;
    const int64_t rows_per_cta_v = 8 * (32 / l);
    int gv = (int)cdiv(n, rows_per_cta_v);
    const int capv = kNumSMs * 8 * 8;
    if (gv > capv) gv = capv;
#define LAUNCHVEC(TX, TO, L, VPT)                                                                                  \
  ln_residual_vec_kernel<TX, TO, L, VPT><<<gv, 256, 0, st>>>((const TX*)x, gamma, beta, eps, residual, res_out, \
                                                            gamma2, beta2, (TO*)out2, n, row_scale)
#define LAUNCHC(TX, TO)                          \
  do {                                           \
    if (c == 32) LAUNCHVEC(TX, TO, 4, 1);        \
    else if (c == 64) LAUNCHVEC(TX, TO, 8, 1);   \
    else if (c == 128) LAUNCHVEC(TX, TO, 16, 1); \
    else if (c == 256) LAUNCHVEC(TX, TO, 32, 1); \
    else if (c == 512) LAUNCHVEC(TX, TO, 32, 2); \
    else LAUNCHVEC(TX, TO, 32, 3);               \
  } while (0)
    if (dtype_x == PTV3_F32 && dtype_out2 == PTV3_F32) LAUNCHC(float, float);
    else if (dtype_x == PTV3_F32 && dtype_out2 == PTV3_BF16) LAUNCHC(float, __nv_bfloat16);
    else if (dtype_x == PTV3_BF16 && dtype_out2 == PTV3_F32) LAUNCHC(__nv_bfloat16, float);
    else if (dtype_x == PTV3_BF16 && dtype_out2 == PTV3_BF16) LAUNCHC(__nv_bfloat16, __nv_bfloat16);
    else return PTV3_EINVAL;
#undef LAUNCHC
#undef LAUNCHVEC
    return check_launch();
  }
#define LAUNCHV(TX, TO, V)                                                                                   \
  ln_residual_kernel<TX, TO, V><<<g, 256, 0, st>>>((const TX*)x, gamma, beta, eps, residual, res_out, gamma2, \
                                                   beta2, (TO*)out2, n, c, row_scale)
#define LAUNCH(TX, TO)                                                                   \
  do {                                                                                   \
    if (c <= 32) LAUNCHV(TX, TO, 1);                                                     \
    else if (c <= 64) LAUNCHV(TX, TO, 2);                                                \
    else if (c <= 128) LAUNCHV(TX, TO, 4);                                               \
    else if (c <= 256) LAUNCHV(TX, TO, 8);                                               \
    else if (c <= 512) LAUNCHV(TX, TO, 16);                                              \
    else LAUNCHV(TX, TO, 32);                                                            \
  } while (0)
  if (dtype_x == PTV3_F32 && dtype_out2 == PTV3_F32) LAUNCH(float, float);
  else if (dtype_x == PTV3_F32 && dtype_out2 == PTV3_BF16) LAUNCH(float, __nv_bfloat16);
  else if (dtype_x == PTV3_BF16 && dtype_out2 == PTV3_F32) LAUNCH(__nv_bfloat16, float);
  else if (dtype_x == PTV3_BF16 && dtype_out2 == PTV3_BF16) LAUNCH(__nv_bfloat16, __nv_bfloat16);
  else return PTV3_EINVAL;
#undef LAUNCH
#undef LAUNCHV
  return check_launch();
}

}  // extern "C"
