// SURVEY 8f rank 1 (first correct path): backward kernels of the three custom ops that carry gradients --
// SubMConv3d weight gradient (the data gradient re-uses the forward kernels over the mirrored neighbour map),
// window attention, and the CSR segment reductions of pooling.  CUDA-core kernels with fp32 accumulation,
// templated on the storage type so that the fp32 parity path and the bf16 path share one implementation.
// Semantics = torch.autograd through oracle/ops.py::{subm_conv, window_attention, segment_csr}.
#include "common.cuh"

namespace ptv3 {

// ---------------------------------------------------------------------------------------------------------------
// SubMConv3d weight gradient:  dW[k][o][c] += sum_i dY[i][o] * X[nbr[i,k]][c]      (absent neighbours skipped)
// grid (k, row chunk, 64x64 tile of (o,c)); 256 threads, each a 4x4 register tile; rows staged 16 at a time.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kWgTO = 64, kWgTC = 64, kWgR = 16;

template <typename T>
__global__ void __launch_bounds__(256) subm_wgrad_simt_kernel(const T* __restrict__ x, const T* __restrict__ dy,
                                                              const int32_t* __restrict__ nbr, int64_t n, int k3, int cin,
                                                              int cout, int64_t rows_per_cta, float* __restrict__ dw) {
  __shared__ float dys[kWgR][kWgTO + 4];
  __shared__ float xs[kWgR][kWgTC + 4];
  __shared__ int js[kWgR];
  const int k = blockIdx.x;
  const int n_ct = (cin + kWgTC - 1) / kWgTC;
  const int o0 = ((int)blockIdx.z / n_ct) * kWgTO, c0 = ((int)blockIdx.z % n_ct) * kWgTC;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;  // tx -> input channel, ty -> output channel
  float acc[4][4];
#pragma unroll
  for (int u = 0; u < 4; ++u)
#pragma unroll
    for (int v = 0; v < 4; ++v) acc[u][v] = 0.f;
  const int64_t r_begin = (int64_t)blockIdx.y * rows_per_cta;
  const int64_t r_end = (r_begin + rows_per_cta < n) ? r_begin + rows_per_cta : n;
  for (int64_t r0 = r_begin; r0 < r_end; r0 += kWgR) {
    __syncthreads();
    if (threadIdx.x < kWgR) {
      int64_t i = r0 + threadIdx.x;
      js[threadIdx.x] = (i < r_end) ? nbr[i * k3 + k] : -1;
    }
    __syncthreads();
    bool any = false;
#pragma unroll
    for (int r = 0; r < kWgR; ++r) any |= js[r] >= 0;
    if (!any) continue;  // uniform over the block
    for (int e = threadIdx.x; e < kWgR * kWgTO; e += 256) {
      int r = e / kWgTO, o = e % kWgTO;
      float v = 0.f;
      if (js[r] >= 0 && o0 + o < cout) v = Elem<T>::load(dy + (r0 + r) * cout + o0 + o);
      dys[r][o] = v;
    }
    for (int e = threadIdx.x; e < kWgR * kWgTC; e += 256) {
      int r = e / kWgTC, c = e % kWgTC;
      float v = 0.f;
      if (js[r] >= 0 && c0 + c < cin) v = Elem<T>::load(x + (int64_t)js[r] * cin + c0 + c);
      xs[r][c] = v;
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kWgR; ++r) {
      float a[4], b[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) a[u] = dys[r][ty * 4 + u];
#pragma unroll
      for (int v = 0; v < 4; ++v) b[v] = xs[r][tx * 4 + v];
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int v = 0; v < 4; ++v) acc[u][v] = fmaf(a[u], b[v], acc[u][v]);
    }
  }
#pragma unroll
  for (int u = 0; u < 4; ++u)
#pragma unroll
    for (int v = 0; v < 4; ++v) {
      int o = o0 + ty * 4 + u, c = c0 + tx * 4 + v;
      if (o < cout && c < cin && acc[u][v] != 0.f) atomicAdd(dw + ((int64_t)k * cout + o) * cin + c, acc[u][v]);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Window attention backward (head_dim 16).  With S = (scale q) k^T, P = softmax(S), D_i = dO_i . O_i:
//   dV_j = sum_i P_ij dO_i      dS_ij = P_ij (dO_i . V_j - D_i)      dQ_i = scale sum_j dS_ij K_j
//   dK_j = sum_i dS_ij (scale Q_i)
// Kernel 1 (thread = query): recomputes the log-sum-exp of its row, then dQ; writes lse and D for kernel 2.
// Kernel 2 (thread = key): dK, dV over the window's queries; a row can be a key of two windows (the last window of
// a scene re-uses the tail of the previous one), so dK / dV are accumulated with atomicAdd into a zeroed buffer.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kBD = 16, kBThreads = 128, kBChunk = 128;

template <typename T>
__global__ void __launch_bounds__(kBThreads) attn_bwd_dq_kernel(const T* __restrict__ qkv, const T* __restrict__ out,
                                                                const T* __restrict__ dout,
                                                                const int64_t* __restrict__ order,
                                                                const int64_t* __restrict__ windows,
                                                                const int64_t* __restrict__ n_windows, int h, float scale,
                                                                float* __restrict__ dqkv, float* __restrict__ lse,
                                                                float* __restrict__ dvec) {
  __shared__ float ks[kBChunk][kBD];
  __shared__ float vs[kBChunk][kBD];
  const int w = blockIdx.x, head = blockIdx.y, qb = blockIdx.z;
  if (w >= n_windows[0]) return;
  const int64_t q0 = windows[4 * w + 0], ql = windows[4 * w + 1], k0 = windows[4 * w + 2], kl = windows[4 * w + 3];
  if ((int64_t)qb * kBThreads >= ql) return;
  const int64_t rs = (int64_t)3 * h * kBD;
  const int64_t qi = (int64_t)qb * kBThreads + threadIdx.x;
  const bool active = qi < ql;
  int64_t qrow = 0;
  float q[kBD], dO[kBD], dq[kBD];
  float D = 0.f, m = -INFINITY, l = 0.f;
#pragma unroll
  for (int i = 0; i < kBD; ++i) q[i] = dO[i] = dq[i] = 0.f;
  if (active) {
    qrow = order ? order[q0 + qi] : q0 + qi;
    const T* qp = qkv + qrow * rs + head * kBD;
    const T* op = out + qrow * (int64_t)h * kBD + head * kBD;
    const T* dp = dout + qrow * (int64_t)h * kBD + head * kBD;
#pragma unroll
    for (int i = 0; i < kBD; ++i) {
      q[i] = Elem<T>::load(qp + i) * scale;
      dO[i] = Elem<T>::load(dp + i);
      D = fmaf(dO[i], Elem<T>::load(op + i), D);
    }
  }
  for (int pass = 0; pass < 2; ++pass) {
    float lse_row = (pass == 1) ? m + __logf(l) : 0.f;
    for (int64_t c0 = 0; c0 < kl; c0 += kBChunk) {
      int cl = (int)((kl - c0 < kBChunk) ? (kl - c0) : kBChunk);
      __syncthreads();
      for (int e = threadIdx.x; e < cl * kBD; e += kBThreads) {
        int j = e / kBD, i = e % kBD;
        int64_t krow = order ? order[k0 + c0 + j] : k0 + c0 + j;
        const T* base = qkv + krow * rs + head * kBD + i;
        ks[j][i] = Elem<T>::load(base + (int64_t)h * kBD);
        vs[j][i] = Elem<T>::load(base + (int64_t)2 * h * kBD);
      }
      __syncthreads();
      if (!active) continue;
      for (int j = 0; j < cl; ++j) {
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < kBD; ++i) s = fmaf(q[i], ks[j][i], s);
        if (pass == 0) {
          float mn = fmaxf(m, s);
          l = l * __expf(m - mn) + __expf(s - mn);
          m = mn;
        } else {
          float p = __expf(s - lse_row), dpv = 0.f;
#pragma unroll
          for (int i = 0; i < kBD; ++i) dpv = fmaf(dO[i], vs[j][i], dpv);
          float ds = p * (dpv - D);
#pragma unroll
          for (int i = 0; i < kBD; ++i) dq[i] = fmaf(ds, ks[j][i], dq[i]);
        }
      }
    }
  }
  if (active) {
    float* o = dqkv + qrow * rs + head * kBD;
#pragma unroll
    for (int i = 0; i < kBD; ++i) o[i] = dq[i] * scale;
    lse[qrow * h + head] = m + __logf(l);
    dvec[qrow * h + head] = D;
  }
}

template <typename T>
__global__ void __launch_bounds__(kBThreads) attn_bwd_dkv_kernel(const T* __restrict__ qkv, const T* __restrict__ dout,
                                                                 const int64_t* __restrict__ order,
                                                                 const int64_t* __restrict__ windows,
                                                                 const int64_t* __restrict__ n_windows, int h, float scale,
                                                                 const float* __restrict__ lse,
                                                                 const float* __restrict__ dvec,
                                                                 float* __restrict__ dqkv) {
  __shared__ float qs[kBChunk][kBD];
  __shared__ float dos[kBChunk][kBD];
  __shared__ float ls[kBChunk];
  __shared__ float dvs[kBChunk];
  const int w = blockIdx.x, head = blockIdx.y, kb = blockIdx.z;
  if (w >= n_windows[0]) return;
  const int64_t q0 = windows[4 * w + 0], ql = windows[4 * w + 1], k0 = windows[4 * w + 2], kl = windows[4 * w + 3];
  if ((int64_t)kb * kBThreads >= kl) return;
  const int64_t rs = (int64_t)3 * h * kBD;
  const int64_t kj = (int64_t)kb * kBThreads + threadIdx.x;
  const bool active = kj < kl;
  int64_t krow = 0;
  float k[kBD], v[kBD], dk[kBD], dv[kBD];
#pragma unroll
  for (int i = 0; i < kBD; ++i) k[i] = v[i] = dk[i] = dv[i] = 0.f;
  if (active) {
    krow = order ? order[k0 + kj] : k0 + kj;
    const T* base = qkv + krow * rs + head * kBD;
#pragma unroll
    for (int i = 0; i < kBD; ++i) {
      k[i] = Elem<T>::load(base + (int64_t)h * kBD + i);
      v[i] = Elem<T>::load(base + (int64_t)2 * h * kBD + i);
    }
  }
  for (int64_t c0 = 0; c0 < ql; c0 += kBChunk) {
    int cl = (int)((ql - c0 < kBChunk) ? (ql - c0) : kBChunk);
    __syncthreads();
    for (int e = threadIdx.x; e < cl * kBD; e += kBThreads) {
      int j = e / kBD, i = e % kBD;
      int64_t qrow = order ? order[q0 + c0 + j] : q0 + c0 + j;
      qs[j][i] = Elem<T>::load(qkv + qrow * rs + head * kBD + i) * scale;
      dos[j][i] = Elem<T>::load(dout + qrow * (int64_t)h * kBD + head * kBD + i);
      if (i == 0) {
        ls[j] = lse[qrow * h + head];
        dvs[j] = dvec[qrow * h + head];
      }
    }
    __syncthreads();
    if (!active) continue;
    for (int j = 0; j < cl; ++j) {
      float s = 0.f, dpv = 0.f;
#pragma unroll
      for (int i = 0; i < kBD; ++i) {
        s = fmaf(qs[j][i], k[i], s);
        dpv = fmaf(dos[j][i], v[i], dpv);
      }
      float p = __expf(s - ls[j]);
      float ds = p * (dpv - dvs[j]);
#pragma unroll
      for (int i = 0; i < kBD; ++i) {
        dv[i] = fmaf(p, dos[j][i], dv[i]);
        dk[i] = fmaf(ds, qs[j][i], dk[i]);
      }
    }
  }
  if (active) {
    float* o = dqkv + krow * rs + head * kBD;
#pragma unroll
    for (int i = 0; i < kBD; ++i) {
      atomicAdd(o + (int64_t)h * kBD + i, dk[i]);
      atomicAdd(o + (int64_t)2 * h * kBD + i, dv[i]);
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// segment_csr backward.  One thread per (segment, channel): sum -> copy, mean -> 1/len, max / min -> the FIRST row of
// the segment (CSR order) whose value equals the reduced value takes the gradient (torch_scatter's arg semantics).
// Segments partition the rows, so every element of dsrc that belongs to a segment is written exactly once.
// ---------------------------------------------------------------------------------------------------------------
template <typename T>
__global__ void segment_csr_bwd_kernel(const float* __restrict__ dout, const T* __restrict__ src,
                                       const T* __restrict__ out, const int64_t* __restrict__ gather,
                                       const int64_t* __restrict__ indptr, int64_t m, int64_t c, int reduce,
                                       float* __restrict__ dsrc) {
  const int64_t total = m * c;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t seg = e / c, ch = e % c;
    const int64_t b = indptr[seg], en = indptr[seg + 1];
    if (en <= b) continue;
    const float g = dout[e];
    if (reduce == 0 || reduce == 1) {
      const float v = reduce == 1 ? g / (float)(en - b) : g;
      for (int64_t j = b; j < en; ++j) dsrc[(gather ? gather[j] : j) * c + ch] = v;
    } else {
      const float target = Elem<T>::load(out + e);
      bool given = false;
      for (int64_t j = b; j < en; ++j) {
        const int64_t row = gather ? gather[j] : j;
        const bool hit = !given && Elem<T>::load(src + row * c + ch) == target;
        dsrc[row * c + ch] = hit ? g : 0.f;
        given |= hit;
      }
    }
  }
}

// csrc/attn_bwd_tc.cu
int attn_window_bwd_tc(const __nv_bfloat16* qkv, const __nv_bfloat16* out, const __nv_bfloat16* dout,
                       const int64_t* order, const int64_t* windows, const int64_t* n_windows, int64_t n_windows_max,
                       int64_t max_len, int h, float scale, float* dqkv, cudaStream_t st, const float* lse);
int g_attn_bwd_simt = 0;   // debug slot 4: 1 forces the CUDA-core kernels for bf16 too
// csrc/wgrad_tc.cu
int subm_wgrad_fast(const void* in, const void* dout, const int32_t* nbr, int64_t n, int c_in, int c_out, int k3, int dtype,
                    float* dweight, cudaStream_t st);
int g_wgrad_simt = 0;      // debug slot 5: 1 forces the generic CUDA-core kernel

}  // namespace ptv3

using namespace ptv3;

extern "C" {

int ptv3_subm_wgrad(const void* in, const void* dout, const int32_t* nbr, int64_t n, int c_in, int c_out, int k3, int dtype,
                    float* dweight, void* stream) {
  if (n < 0 || c_in < 1 || c_out < 1 || k3 < 1 || k3 > 65535) return PTV3_EINVAL;
  if (n == 0) return PTV3_OK;
  if (!in || !dout || !nbr || !dweight) return PTV3_EINVAL;
  if (!g_wgrad_simt) {
    int rc = subm_wgrad_fast(in, dout, nbr, n, c_in, c_out, k3, dtype, dweight, as_stream(stream));
    if (rc != PTV3_EUNSUPPORTED) return rc;
  }
  const int64_t tiles = cdiv(c_out, kWgTO) * cdiv(c_in, kWgTC);
  if (tiles > 65535) return PTV3_EINVAL;
  // enough row chunks to fill the machine a few times, at least 1024 rows per CTA to amortise the atomics
  int64_t chunks = cdiv((int64_t)kNumSMs * 8, (int64_t)k3 * tiles);
  if (chunks < 1) chunks = 1;
  int64_t rows_per_cta = cdiv(cdiv(n, chunks), kWgR) * kWgR;
  if (rows_per_cta < 1024) rows_per_cta = 1024;
  chunks = cdiv(n, rows_per_cta);
  if (chunks > 65535) return PTV3_EINVAL;
  dim3 grid((unsigned)k3, (unsigned)chunks, (unsigned)tiles);
  cudaStream_t st = as_stream(stream);
  if (dtype == PTV3_F32)
    subm_wgrad_simt_kernel<float><<<grid, 256, 0, st>>>((const float*)in, (const float*)dout, nbr, n, k3, c_in, c_out,
                                                        rows_per_cta, dweight);
  else if (dtype == PTV3_BF16)
    subm_wgrad_simt_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)in, (const __nv_bfloat16*)dout, nbr,
                                                                n, k3, c_in, c_out, rows_per_cta, dweight);
  else
    return PTV3_EINVAL;
  return check_launch();
}

size_t ptv3_attn_window_bwd_workspace(int64_t n, int h) { return align_up((size_t)n * h * 4) * 2; }

int ptv3_attn_window_bwd(const void* qkv, const void* out, const void* dout, int dtype, const int64_t* order,
                         const int64_t* windows, const int64_t* n_windows, int64_t n_windows_max, int64_t max_len,
                         int64_t n, int h, int d, float scale, const float* lse, float* dqkv, void* workspace,
                         size_t workspace_bytes, void* stream) {
  if (n < 0 || h < 1 || h > 65535 || d != kBD || n_windows_max < 0 || max_len < 1) return PTV3_EINVAL;
  if (n == 0 || n_windows_max == 0) return PTV3_OK;
  if (!qkv || !out || !dout || !windows || !n_windows || !dqkv) return PTV3_EINVAL;
  if (dtype == PTV3_BF16 && !g_attn_bwd_simt) {
    int rc = attn_window_bwd_tc((const __nv_bfloat16*)qkv, (const __nv_bfloat16*)out, (const __nv_bfloat16*)dout, order,
                                windows, n_windows, n_windows_max, max_len, h, scale, dqkv, as_stream(stream), lse);
    if (rc != PTV3_EUNSUPPORTED) return rc;
  }
  Workspace ws(workspace, workspace_bytes);
  float* lse_ws = ws.take<float>((size_t)n * h);
  float* dvec = ws.take<float>((size_t)n * h);
  if (!ws.ok()) return PTV3_EWORKSPACE;
  int64_t blocks = cdiv(max_len, kBThreads);
  if (blocks > 65535) return PTV3_EINVAL;
  dim3 grid((unsigned)n_windows_max, (unsigned)h, (unsigned)blocks);
  cudaStream_t st = as_stream(stream);
  if (dtype == PTV3_F32) {
    attn_bwd_dq_kernel<float><<<grid, kBThreads, 0, st>>>((const float*)qkv, (const float*)out, (const float*)dout, order,
                                                          windows, n_windows, h, scale, dqkv, lse_ws, dvec);
    attn_bwd_dkv_kernel<float><<<grid, kBThreads, 0, st>>>((const float*)qkv, (const float*)dout, order, windows,
                                                           n_windows, h, scale, lse_ws, dvec, dqkv);
  } else if (dtype == PTV3_BF16) {
    attn_bwd_dq_kernel<__nv_bfloat16><<<grid, kBThreads, 0, st>>>(
        (const __nv_bfloat16*)qkv, (const __nv_bfloat16*)out, (const __nv_bfloat16*)dout, order, windows, n_windows, h,
        scale, dqkv, lse_ws, dvec);
    attn_bwd_dkv_kernel<__nv_bfloat16><<<grid, kBThreads, 0, st>>>((const __nv_bfloat16*)qkv, (const __nv_bfloat16*)dout,
                                                                   order, windows, n_windows, h, scale, lse_ws, dvec, dqkv);
  } else {
    return PTV3_EINVAL;
  }
  return check_launch(2);
}

int ptv3_segment_csr_bwd(const float* dout, const void* src, const void* out, int dtype, const int64_t* gather,
                         const int64_t* indptr, int64_t m, int64_t c, int reduce, float* dsrc, void* stream) {
  if (m < 0 || c < 1 || reduce < 0 || reduce > 3) return PTV3_EINVAL;
  if (m == 0) return PTV3_OK;
  if (!dout || !indptr || !dsrc || (reduce >= 2 && (!src || !out))) return PTV3_EINVAL;
  cudaStream_t st = as_stream(stream);
  int grid = grid_for(m * c, 256);
  if (dtype == PTV3_F32)
    segment_csr_bwd_kernel<float><<<grid, 256, 0, st>>>(dout, (const float*)src, (const float*)out, gather, indptr, m, c,
                                                        reduce, dsrc);
  else if (dtype == PTV3_BF16)
    segment_csr_bwd_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(dout, (const __nv_bfloat16*)src,
                                                                (const __nv_bfloat16*)out, gather, indptr, m, c, reduce,
                                                                dsrc);
  else
    return PTV3_EINVAL;
  return check_launch();
}

}  // extern "C"
