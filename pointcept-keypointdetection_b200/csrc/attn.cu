// A3/A4: window records + CUDA-core window attention (fp32 math; the parity
// path for fp32 models and the checker for the tensor-core kernel in
// attn_tc.cu).  Semantics: oracle/ops.py::windows_from_offset / window_attention.
#include "common.cuh"

namespace ptv3 {

// one thread per scene (scenes are few); windows are written in scene order
__global__ void build_windows_kernel(const int64_t* __restrict__ offset, int64_t n_scenes, int64_t K,
                                     int64_t* __restrict__ windows, int64_t n_windows_max,
                                     int64_t* __restrict__ n_windows) {
  // serial prefix over scenes by thread 0 of the block (n_scenes is tiny), then parallel fill
  __shared__ int64_t total;
  if (threadIdx.x == 0) {
    int64_t w = 0, s = 0;
    for (int64_t b = 0; b < n_scenes; ++b) {
      int64_t e = offset[b], cnt = e - s;
      int64_t nw = cnt <= 0 ? 0 : (cnt <= K ? 1 : (cnt + K - 1) / K);
      for (int64_t i = 0; i < nw && w < n_windows_max; ++i, ++w) {
        int64_t q0 = s + i * K;
        int64_t ql = (e - q0 < K) ? (e - q0) : K;
        int64_t k0, kl;
        if (cnt <= K) {
          k0 = s;
          kl = cnt;
        } else {
          kl = K;
          k0 = (ql == K) ? q0 : e - K;
        }
        windows[4 * w + 0] = q0;
        windows[4 * w + 1] = ql;
        windows[4 * w + 2] = k0;
        windows[4 * w + 3] = kl;
      }
      s = e;
    }
    total = w;
    n_windows[0] = w;
  }
  __syncthreads();
  for (int64_t w = total + threadIdx.x; w < n_windows_max; w += blockDim.x) {
    windows[4 * w + 0] = 0;
    windows[4 * w + 1] = 0;
    windows[4 * w + 2] = 0;
    windows[4 * w + 3] = 0;
  }
}

// CUDA-core kernel: grid (window, head, query block of 128); one thread per query row,
// keys/values streamed through shared memory in chunks of 128 rows, online softmax.
constexpr int kAttnD = 16;
constexpr int kAttnThreads = 128;
constexpr int kAttnChunk = 128;

template <typename T>
__global__ void __launch_bounds__(kAttnThreads) attn_window_simt_kernel(
    const T* __restrict__ qkv, const int64_t* __restrict__ order, const int64_t* __restrict__ windows,
    const int64_t* __restrict__ n_windows, int h, float scale, T* __restrict__ out) {
  __shared__ float ks[kAttnChunk][kAttnD];
  __shared__ float vs[kAttnChunk][kAttnD];
  const int w = blockIdx.x, head = blockIdx.y, qb = blockIdx.z;
  if (w >= n_windows[0]) return;
  const int64_t q0 = windows[4 * w + 0], ql = windows[4 * w + 1], k0 = windows[4 * w + 2], kl = windows[4 * w + 3];
  if ((int64_t)qb * kAttnThreads >= ql) return;
  const int64_t rs = (int64_t)3 * h * kAttnD;  // row stride of qkv
  const int64_t qi = (int64_t)qb * kAttnThreads + threadIdx.x;
  const bool active = qi < ql;
  int64_t qrow = 0;
  float q[kAttnD], o[kAttnD];
  float m = -INFINITY, l = 0.f;
#pragma unroll
  for (int i = 0; i < kAttnD; ++i) {
    q[i] = 0.f;
    o[i] = 0.f;
  }
  if (active) {
    qrow = order ? order[q0 + qi] : q0 + qi;
    const T* qp = qkv + qrow * rs + head * kAttnD;
#pragma unroll
    for (int i = 0; i < kAttnD; ++i) q[i] = Elem<T>::load(qp + i) * scale;
  }
  for (int64_t c0 = 0; c0 < kl; c0 += kAttnChunk) {
    int cl = (int)((kl - c0 < kAttnChunk) ? (kl - c0) : kAttnChunk);
    __syncthreads();
    for (int e = threadIdx.x; e < cl * kAttnD; e += kAttnThreads) {
      int j = e / kAttnD, i = e % kAttnD;
      int64_t krow = order ? order[k0 + c0 + j] : k0 + c0 + j;
      const T* base = qkv + krow * rs + head * kAttnD + i;
      ks[j][i] = Elem<T>::load(base + (int64_t)h * kAttnD);
      vs[j][i] = Elem<T>::load(base + (int64_t)2 * h * kAttnD);
    }
    __syncthreads();
    if (active) {
      for (int j = 0; j < cl; ++j) {
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < kAttnD; ++i) s = fmaf(q[i], ks[j][i], s);
        float mn = fmaxf(m, s);
        float alpha = __expf(m - mn), p = __expf(s - mn);
        l = l * alpha + p;
#pragma unroll
        for (int i = 0; i < kAttnD; ++i) o[i] = fmaf(o[i], alpha, p * vs[j][i]);
        m = mn;
      }
    }
  }
  if (active) {
    float inv = 1.f / l;
    T* op = out + qrow * (int64_t)h * kAttnD + head * kAttnD;
#pragma unroll
    for (int i = 0; i < kAttnD; ++i) Elem<T>::store(op + i, o[i] * inv);
  }
}

// implemented in attn_tc.cu
int attn_window_tc05(const __nv_bfloat16* qkv, const int64_t* order, const int64_t* windows,
                     const int64_t* n_windows, int64_t n_windows_max, int64_t max_len, int h, float scale,
                     __nv_bfloat16* out, cudaStream_t st);
extern int g_attn_tc05_variant;
extern int g_attn_qsplit;
extern int g_conv_async;
extern int g_attn_bwd_simt;
extern int g_wgrad_simt;
extern int g_attn_bwd_warps;
int attn_window_tc(const __nv_bfloat16* qkv, const int64_t* order, const int64_t* windows, const int64_t* n_windows,
                   int64_t n_windows_max, int64_t max_len, int64_t n, int h, float scale, __nv_bfloat16* out,
                   cudaStream_t st, float* lse = nullptr);
// implemented in attn_f16.cu (the default bf16 kernel)
int attn_window_f16(const __nv_bfloat16* qkv, const int64_t* order, const int64_t* windows, const int64_t* n_windows,
                    int64_t n_windows_max, int64_t max_len, int h, float scale, __nv_bfloat16* out, cudaStream_t st,
                    float* lse, int* err, const int* only_if = nullptr);
// implemented in attn_tm.cu (tcgen05 / TMEM, single pass); redo = (n_windows_max * h) int flags, zeroed by the caller
int attn_window_tm(const __nv_bfloat16* qkv, const int64_t* order, const int64_t* windows, const int64_t* n_windows,
                   int64_t n_windows_max, int64_t max_len, int h, float scale, __nv_bfloat16* out, int* redo,
                   cudaStream_t st);
extern int g_attn_f16_variant;
extern int g_attn_tm_poly;
extern int g_attn_lse_algo;

}  // namespace ptv3

using namespace ptv3;

extern "C" {

int ptv3_build_windows(const int64_t* offset, int64_t n_scenes, int64_t patch, int64_t* windows, int64_t n_windows_max,
                       int64_t* n_windows, void* stream) {
  if (n_scenes < 1 || patch < 1 || n_windows_max < 0 || !offset || !n_windows) return PTV3_EINVAL;
  build_windows_kernel<<<1, 256, 0, as_stream(stream)>>>(offset, n_scenes, patch, windows, n_windows_max, n_windows);
  return check_launch();
}

size_t ptv3_attn_window_workspace(int64_t n_windows_max, int h) {
  return align_up((size_t)(n_windows_max > 0 ? n_windows_max : 1) * (size_t)(h > 0 ? h : 1) * sizeof(int));
}

int ptv3_attn_window_fwd(const void* qkv, int dtype, const int64_t* order, const int64_t* windows,
                         const int64_t* n_windows, int64_t n_windows_max, int64_t max_len, int64_t n, int h, int d,
                         float scale, int algo, void* out, void* workspace, size_t workspace_bytes, void* stream) {
  if (n < 0 || h < 1 || h > 65535 || d != kAttnD || n_windows_max < 0 || max_len < 1 || algo < 0 || algo > 5)
    return PTV3_EINVAL;
  if (n == 0 || n_windows_max == 0) return PTV3_OK;
  if (!qkv || !windows || !n_windows || !out) return PTV3_EINVAL;
  cudaStream_t st = as_stream(stream);
  if (dtype == PTV3_BF16 && algo == 5) {
    // tcgen05 / TMEM kernel; (window, head) units it flags (records outside its contract, rows whose scores spread over
    // more than ~100 octaves) are recomputed by the register kernel in a second, normally empty, launch
    if (max_len > 1024) return PTV3_EUNSUPPORTED;
    const size_t need = ptv3_attn_window_workspace(n_windows_max, h);
    if (!workspace || workspace_bytes < need) return PTV3_EWORKSPACE;
    int* redo = static_cast<int*>(workspace);
    if (cudaMemsetAsync(redo, 0, need, st) != cudaSuccess) return PTV3_ELAUNCH;
    int rc = attn_window_tm((const __nv_bfloat16*)qkv, order, windows, n_windows, n_windows_max, max_len, h, scale,
                            (__nv_bfloat16*)out, redo, st);
    if (rc != PTV3_OK) return rc;
    return attn_window_f16((const __nv_bfloat16*)qkv, order, windows, n_windows, n_windows_max, max_len, h, scale,
                           (__nv_bfloat16*)out, st, nullptr, device_error_flag(), redo);
  }
  if (dtype == PTV3_BF16 && algo == 3) {
    if (max_len > 1024) return PTV3_EUNSUPPORTED;
    return attn_window_tc05((const __nv_bfloat16*)qkv, order, windows, n_windows, n_windows_max, max_len, h, scale,
                            (__nv_bfloat16*)out, st);
  }
  // 0 (auto) / 2: the first-generation mma.sync kernel -- inside the model it is still the fastest one (bench step:
  // attention 11.5 ms vs 12.3 ms with kernel 4, profiles/r02_attention.md)
  if (dtype == PTV3_BF16 && (algo == 2 || algo == 0))
    return attn_window_tc((const __nv_bfloat16*)qkv, order, windows, n_windows, n_windows_max, max_len, n, h, scale,
                          (__nv_bfloat16*)out, st);
  if (dtype == PTV3_BF16 && algo != 1)      // 4: the f16-staged register kernel (the most accurate bf16 path)
    return attn_window_f16((const __nv_bfloat16*)qkv, order, windows, n_windows, n_windows_max, max_len, h, scale,
                           (__nv_bfloat16*)out, st, nullptr, device_error_flag());
  if (algo >= 2) return PTV3_EUNSUPPORTED;
  int64_t qblocks = cdiv(max_len, kAttnThreads);
  if (qblocks > 65535) return PTV3_EINVAL;
  dim3 grid((unsigned)n_windows_max, (unsigned)h, (unsigned)qblocks);
  if (dtype == PTV3_F32)
    attn_window_simt_kernel<float><<<grid, kAttnThreads, 0, st>>>((const float*)qkv, order, windows, n_windows, h, scale,
                                                                  (float*)out);
  else if (dtype == PTV3_BF16)
    attn_window_simt_kernel<__nv_bfloat16><<<grid, kAttnThreads, 0, st>>>(
        (const __nv_bfloat16*)qkv, order, windows, n_windows, h, scale, (__nv_bfloat16*)out);
  else
    return PTV3_EINVAL;
  return check_launch();
}

int ptv3_attn_window_fwd_lse(const void* qkv, int dtype, const int64_t* order, const int64_t* windows,
                             const int64_t* n_windows, int64_t n_windows_max, int64_t max_len, int64_t n, int h, int d,
                             float scale, void* out, float* lse, void* stream) {
  if (n < 0 || h < 1 || h > 65535 || d != kAttnD || n_windows_max < 0 || max_len < 1) return PTV3_EINVAL;
  if (dtype != PTV3_BF16) return PTV3_EUNSUPPORTED;
  if (n == 0 || n_windows_max == 0) return PTV3_OK;
  if (!qkv || !windows || !n_windows || !out || !lse) return PTV3_EINVAL;
  if (g_attn_lse_algo != 2)
    return attn_window_f16((const __nv_bfloat16*)qkv, order, windows, n_windows, n_windows_max, max_len, h, scale,
                           (__nv_bfloat16*)out, as_stream(stream), lse, device_error_flag());
  return attn_window_tc((const __nv_bfloat16*)qkv, order, windows, n_windows, n_windows_max, max_len, n, h, scale,
                        (__nv_bfloat16*)out, as_stream(stream), lse);
}

int ptv3_debug_set(int key, int value) {
  if (key == 1) { g_attn_tc05_variant = value; return PTV3_OK; }
  if (key == 2) { g_attn_qsplit = value; return PTV3_OK; }
  if (key == 3) { g_conv_async = value; return PTV3_OK; }
  if (key == 4) { g_attn_bwd_simt = value; return PTV3_OK; }
  if (key == 5) { g_wgrad_simt = value; return PTV3_OK; }
  if (key == 6) { g_attn_bwd_warps = value; return PTV3_OK; }
  if (key == 7) { g_attn_f16_variant = value; return PTV3_OK; }
  if (key == 8) { g_attn_lse_algo = value; return PTV3_OK; }
  if (key == 9) { g_attn_tm_poly = value; return PTV3_OK; }
  return PTV3_EINVAL;
}


}  // extern "C"
