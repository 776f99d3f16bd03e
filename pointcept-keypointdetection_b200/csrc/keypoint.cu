// A9: per-scene keypoint aggregation of the offset head (tools/infer_offset.py:366-399), on the device:
//   votes[i,k] = coord[i] + pred[i,k,:3],  conf[i,k] = pred[i,k,3]
//   argmax:   kp[b,k] = votes[i*,k],  i* = first index of max conf over the scene's rows
//   weighted: kp[b,k] = sum_i votes[i,k] w_i / sum_i w_i with w_i = conf > thresh ? conf : 0, argmax vote if no w_i
// One CTA per (scene, keypoint): coalesced 16-byte reads of the keypoint's (dx,dy,dz,conf), warp-shuffle + shared
// memory reduction in a fixed order (deterministic).  Replaces a per-scene, per-keypoint host loop with .item() syncs.
#include "common.cuh"

namespace ptv3 {

constexpr int kKpThreads = 512;

__global__ void __launch_bounds__(kKpThreads) keypoint_aggregate_kernel(const float* __restrict__ coord,
                                                                        const float* __restrict__ pred,
                                                                        const int64_t* __restrict__ offset, int k_pts,
                                                                        int weighted, float thresh,
                                                                        float* __restrict__ out) {
  const int b = blockIdx.x, k = blockIdx.y;
  const int64_t start = b ? offset[b - 1] : 0, end = offset[b];
  float best_v = -INFINITY;
  int64_t best_i = INT64_MAX;
  float sw = 0.f, sx = 0.f, sy = 0.f, sz = 0.f;
  for (int64_t i = start + threadIdx.x; i < end; i += kKpThreads) {
    const float4 p = *reinterpret_cast<const float4*>(pred + (i * k_pts + k) * 4);
    if (p.w > best_v) {   // strided ascending i per thread: the first maximum wins
      best_v = p.w;
      best_i = i;
    }
    if (weighted && p.w > thresh) {
      const float cx = coord[i * 3], cy = coord[i * 3 + 1], cz = coord[i * 3 + 2];
      sw += p.w;
      sx = fmaf(cx + p.x, p.w, sx);
      sy = fmaf(cy + p.y, p.w, sy);
      sz = fmaf(cz + p.z, p.w, sz);
    }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best_v, d);
    const int64_t oi = __shfl_xor_sync(0xffffffffu, best_i, d);
    if (ov > best_v || (ov == best_v && oi < best_i)) {
      best_v = ov;
      best_i = oi;
    }
    sw += __shfl_xor_sync(0xffffffffu, sw, d);
    sx += __shfl_xor_sync(0xffffffffu, sx, d);
    sy += __shfl_xor_sync(0xffffffffu, sy, d);
    sz += __shfl_xor_sync(0xffffffffu, sz, d);
  }
  __shared__ float s_v[kKpThreads / 32], s_w[kKpThreads / 32][4];
  __shared__ int64_t s_i[kKpThreads / 32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
    s_v[warp] = best_v;
    s_i[warp] = best_i;
    s_w[warp][0] = sw; s_w[warp][1] = sx; s_w[warp][2] = sy; s_w[warp][3] = sz;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < kKpThreads / 32; ++w) {
      if (s_v[w] > best_v || (s_v[w] == best_v && s_i[w] < best_i)) {
        best_v = s_v[w];
        best_i = s_i[w];
      }
      sw += s_w[w][0]; sx += s_w[w][1]; sy += s_w[w][2]; sz += s_w[w][3];
    }
    float* o = out + ((int64_t)b * k_pts + k) * 3;
    if (end <= start) {
      o[0] = o[1] = o[2] = 0.f;
    } else if (weighted && sw > 0.f) {
      o[0] = sx / sw; o[1] = sy / sw; o[2] = sz / sw;
    } else {
      const float* p = pred + (best_i * k_pts + k) * 4;
      o[0] = coord[best_i * 3] + p[0];
      o[1] = coord[best_i * 3 + 1] + p[1];
      o[2] = coord[best_i * 3 + 2] + p[2];
    }
  }
}

}  // namespace ptv3

using namespace ptv3;

extern "C" int ptv3_keypoint_aggregate(const float* coord, const float* pred, const int64_t* offset, int64_t n_scenes,
                                       int k_pts, int mode, float mask_thresh, float* out, void* stream) {
  if (n_scenes < 0 || k_pts < 1 || k_pts > 65535 || mode < 0 || mode > 1) return PTV3_EINVAL;
  if (n_scenes == 0) return PTV3_OK;
  if (!coord || !pred || !offset || !out) return PTV3_EINVAL;
  dim3 grid((unsigned)n_scenes, (unsigned)k_pts);
  keypoint_aggregate_kernel<<<grid, kKpThreads, 0, as_stream(stream)>>>(coord, pred, offset, k_pts, mode, mask_thresh,
                                                                        out);
  return check_launch();
}

// A10: validation metric of the offset head (hooks/offset_keypoint_evaluator.py:46-91) on the device.
//   gt[b,k]   = mean over the scene's rows with target[i,k,3] > 0 of coord[i] + target[i,k,:3]   (skipped if there are none)
//   pred[b,k] = coord[i*] + pred[i*,k,:3], i* = first argmax_i pred[i,k,3]
//   dist[b,k] = |pred - gt|_2 * scale[b]
// One CTA per (scene, keypoint) streams the keypoint's 16-byte (dx,dy,dz,mask) records of `pred` and `target` once
// (32 B / row algorithmic, coord rows only where the mask is set); fixed-order reduction.  keypoint_eval_accumulate then
// folds (scenes, K) distances into the running sums in the reference's order (scene-major, double precision), so a
// validation epoch needs one device->host read at its end instead of one .item() per scene and keypoint.
namespace ptv3 {

__global__ void __launch_bounds__(kKpThreads) keypoint_eval_kernel(const float* __restrict__ coord,
                                                                   const float* __restrict__ pred,
                                                                   const float* __restrict__ target,
                                                                   const int64_t* __restrict__ offset,
                                                                   const float* __restrict__ scale, int k_pts,
                                                                   float* __restrict__ dist, int32_t* __restrict__ valid) {
  const int b = blockIdx.x, k = blockIdx.y;
  const int64_t start = b ? offset[b - 1] : 0, end = offset[b];
  float best_v = -INFINITY;
  int64_t best_i = INT64_MAX;
  float gx = 0.f, gy = 0.f, gz = 0.f;
  int cnt = 0;
  for (int64_t i = start + threadIdx.x; i < end; i += kKpThreads) {
    const float4 p = *reinterpret_cast<const float4*>(pred + (i * k_pts + k) * 4);
    const float4 t = *reinterpret_cast<const float4*>(target + (i * k_pts + k) * 4);
    if (p.w > best_v) {
      best_v = p.w;
      best_i = i;
    }
    if (t.w > 0.f) {
      gx += coord[i * 3] + t.x;
      gy += coord[i * 3 + 1] + t.y;
      gz += coord[i * 3 + 2] + t.z;
      ++cnt;
    }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best_v, d);
    const int64_t oi = __shfl_xor_sync(0xffffffffu, best_i, d);
    if (ov > best_v || (ov == best_v && oi < best_i)) {
      best_v = ov;
      best_i = oi;
    }
    gx += __shfl_xor_sync(0xffffffffu, gx, d);
    gy += __shfl_xor_sync(0xffffffffu, gy, d);
    gz += __shfl_xor_sync(0xffffffffu, gz, d);
    cnt += __shfl_xor_sync(0xffffffffu, cnt, d);
  }
  __shared__ float s_v[kKpThreads / 32], s_g[kKpThreads / 32][3];
  __shared__ int64_t s_i[kKpThreads / 32];
  __shared__ int s_c[kKpThreads / 32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
    s_v[warp] = best_v;
    s_i[warp] = best_i;
    s_g[warp][0] = gx; s_g[warp][1] = gy; s_g[warp][2] = gz;
    s_c[warp] = cnt;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < kKpThreads / 32; ++w) {
      if (s_v[w] > best_v || (s_v[w] == best_v && s_i[w] < best_i)) {
        best_v = s_v[w];
        best_i = s_i[w];
      }
      gx += s_g[w][0]; gy += s_g[w][1]; gz += s_g[w][2];
      cnt += s_c[w];
    }
    const int64_t o = (int64_t)b * k_pts + k;
    if (cnt == 0 || best_i == INT64_MAX) {
      dist[o] = 0.f;
      valid[o] = 0;
    } else {
      const float* p = pred + (best_i * k_pts + k) * 4;
      const float dx = coord[best_i * 3] + p[0] - gx / (float)cnt;
      const float dy = coord[best_i * 3 + 1] + p[1] - gy / (float)cnt;
      const float dz = coord[best_i * 3 + 2] + p[2] - gz / (float)cnt;
      const float nrm = sqrtf(dx * dx + dy * dy + dz * dz);
      dist[o] = scale ? nrm * scale[b] : nrm;
      valid[o] = 1;
    }
  }
}

// acc (2 + 2K) doubles: [total_dist, total_samples, dist_per_kp[K], samples_per_kp[K]] -- the evaluator's running sums,
// updated in its own loop order (offset_keypoint_evaluator.py:47-91): warp 0 owns the per-keypoint sums (a lane per
// keypoint, scenes in order, sums in registers), one lane of warp 1 the per-scene means; (scenes x K) is a few hundred values.
__global__ void keypoint_eval_accumulate_kernel(const float* __restrict__ dist, const int32_t* __restrict__ valid,
                                                int64_t n_scenes, int k_pts, double* __restrict__ acc) {
  if (threadIdx.x < 32) {
    for (int k = threadIdx.x; k < k_pts; k += 32) {
      double sum = acc[2 + k], cnt = acc[2 + k_pts + k];
      for (int64_t b = 0; b < n_scenes; ++b) {
        if (!valid[b * k_pts + k]) continue;
        sum += (double)dist[b * k_pts + k];
        cnt += 1.0;
      }
      acc[2 + k] = sum;
      acc[2 + k_pts + k] = cnt;
    }
  } else if (threadIdx.x == 32) {
    double total = acc[0], samples = acc[1];
    for (int64_t b = 0; b < n_scenes; ++b) {
      double sum = 0.0;
      int nv = 0;
      for (int k = 0; k < k_pts; ++k) {
        if (!valid[b * k_pts + k]) continue;
        sum += (double)dist[b * k_pts + k];
        ++nv;
      }
      if (nv > 0) {
        total += sum / nv;
        samples += 1.0;
      }
    }
    acc[0] = total;
    acc[1] = samples;
  }
}

}  // namespace ptv3

// Split form of keypoint_eval_kernel for grids that fill the 148 SMs: CTA (scene, split) streams a contiguous slice of the
// scene's rows -- whole (K x 16 B) records of `pred` and `target`, consecutive threads on consecutive 16-byte records --
// and keeps one partial per keypoint; thread t always sees keypoint t % K because the block size is a multiple of K.
// Partials are combined in a fixed order (shared-memory tree inside the CTA, ascending splits in keypoint_eval_finish).
namespace ptv3 {

struct __align__(16) EvalPartial {
  float best_v, gx, gy, gz;
  int64_t best_i;
  int cnt, pad;
};

__device__ __forceinline__ void eval_merge(EvalPartial& a, const EvalPartial& o) {
  if (o.best_v > a.best_v || (o.best_v == a.best_v && o.best_i < a.best_i)) {
    a.best_v = o.best_v;
    a.best_i = o.best_i;
  }
  a.gx += o.gx; a.gy += o.gy; a.gz += o.gz;
  a.cnt += o.cnt;
}

__global__ void __launch_bounds__(kKpThreads, 2) keypoint_eval_split_kernel(const float* __restrict__ coord,
                                                                         const float* __restrict__ pred,
                                                                         const float* __restrict__ target,
                                                                         const int64_t* __restrict__ offset, int k_pts,
                                                                         EvalPartial* __restrict__ partials) {
  __shared__ EvalPartial s_p[kKpThreads];
  const int b = blockIdx.x, split = blockIdx.y, n_split = gridDim.y;
  const int64_t start = b ? offset[b - 1] : 0, end = offset[b], len = end > start ? end - start : 0;
  const int64_t r0 = start + len * split / n_split, r1 = start + len * (split + 1) / n_split;
  const int t = threadIdx.x, T = blockDim.x, G = T / k_pts;   // thread t: keypoint t % K, rows t / K + j * G
  EvalPartial me{-INFINITY, 0.f, 0.f, 0.f, INT64_MAX, 0, 0};
  const float4* p4 = reinterpret_cast<const float4*>(pred) + r0 * k_pts;
  const float4* t4 = reinterpret_cast<const float4*>(target) + r0 * k_pts;
  const int64_t n_rec = (r1 - r0) * k_pts;
  auto consume = [&](const float4& p, const float4& g, int64_t i) {
    if (p.w > me.best_v) {   // ascending rows per thread: the first maximum wins
      me.best_v = p.w;
      me.best_i = i;
    }
    if (g.w > 0.f) {
      me.gx += coord[i * 3] + g.x;
      me.gy += coord[i * 3 + 1] + g.y;
      me.gz += coord[i * 3 + 2] + g.z;
      ++me.cnt;
    }
  };
  int64_t r = t, row = r0 + t / k_pts;
  for (; r + 3 * (int64_t)T < n_rec; r += 4 * T, row += 4 * G) {   // 8 independent 16-byte loads in flight per thread
    float4 p[4], g[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      p[u] = __ldcs(p4 + r + u * T);
      g[u] = __ldcs(t4 + r + u * T);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) consume(p[u], g[u], row + u * G);
  }
  for (; r < n_rec; r += T, row += G) consume(__ldcs(p4 + r), __ldcs(t4 + r), row);
  s_p[t] = me;
  __syncthreads();
  const int g = t / k_pts;   // G threads per keypoint, tree over them
  for (int h = 256; h > 0; h >>= 1) {
    if (g < h && g + h < G) {
      eval_merge(me, s_p[t + h * k_pts]);
      s_p[t] = me;
    }
    __syncthreads();
  }
  if (t < k_pts) partials[((int64_t)b * n_split + split) * k_pts + t] = me;
}

constexpr int kEvalStage = 4096;   // (scene, keypoint) pairs staged in shared memory by the finish kernel

// One CTA of 32 warps: merges the splits of every (scene, keypoint) in a fixed order, writes dist / valid and, when `acc` is given,
// folds them into the running sums from shared memory (see keypoint_eval_accumulate_kernel for the order).
__global__ void __launch_bounds__(1024) keypoint_eval_finish_kernel(const float* __restrict__ coord,
                                                                   const float* __restrict__ pred,
                                                                   const float* __restrict__ scale,
                                                                   const EvalPartial* __restrict__ partials, int n_split,
                                                                   int64_t n_scenes, int k_pts, float* __restrict__ dist,
                                                                   int32_t* __restrict__ valid, double* __restrict__ acc) {
  __shared__ float s_d[kEvalStage];   // distance, < 0 = skipped keypoint
  const int64_t n_pairs = n_scenes * k_pts;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, n_warps = blockDim.x >> 5;
  for (int64_t o = warp; o < n_pairs; o += n_warps) {   // a warp per (scene, keypoint), a lane per split: one load latency
    const int64_t b = o / k_pts;
    const int k = (int)(o - b * k_pts);
    EvalPartial a{-INFINITY, 0.f, 0.f, 0.f, INT64_MAX, 0, 0};
    for (int s = lane; s < n_split; s += 32) eval_merge(a, partials[(b * n_split + s) * k_pts + k]);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {   // butterfly: every lane ends with the same merged partial, fixed order
      EvalPartial q;
      q.best_v = __shfl_xor_sync(0xffffffffu, a.best_v, d);
      q.best_i = __shfl_xor_sync(0xffffffffu, a.best_i, d);
      q.gx = __shfl_xor_sync(0xffffffffu, a.gx, d);
      q.gy = __shfl_xor_sync(0xffffffffu, a.gy, d);
      q.gz = __shfl_xor_sync(0xffffffffu, a.gz, d);
      q.cnt = __shfl_xor_sync(0xffffffffu, a.cnt, d);
      eval_merge(a, q);
    }
    if (lane) continue;
    float d = -1.f;
    if (a.cnt != 0 && a.best_i != INT64_MAX) {
      const float* p = pred + (a.best_i * k_pts + k) * 4;
      const float dx = coord[a.best_i * 3] + p[0] - a.gx / (float)a.cnt;
      const float dy = coord[a.best_i * 3 + 1] + p[1] - a.gy / (float)a.cnt;
      const float dz = coord[a.best_i * 3 + 2] + p[2] - a.gz / (float)a.cnt;
      const float nrm = sqrtf(dx * dx + dy * dy + dz * dz);
      d = scale ? nrm * scale[b] : nrm;
    }
    dist[o] = d < 0.f ? 0.f : d;
    valid[o] = d < 0.f ? 0 : 1;
    if (o < kEvalStage) s_d[o] = d;
  }
  if (!acc) return;
  __syncthreads();
  if (threadIdx.x < 32) {
    for (int k = threadIdx.x; k < k_pts; k += 32) {
      double sum = acc[2 + k], cnt = acc[2 + k_pts + k];
      for (int64_t b = 0; b < n_scenes; ++b) {
        const float d = s_d[b * k_pts + k];
        if (d < 0.f) continue;
        sum += (double)d;
        cnt += 1.0;
      }
      acc[2 + k] = sum;
      acc[2 + k_pts + k] = cnt;
    }
  } else if (threadIdx.x == 32) {
    double total = acc[0], samples = acc[1];
    for (int64_t b = 0; b < n_scenes; ++b) {
      double sum = 0.0;
      int nv = 0;
      for (int k = 0; k < k_pts; ++k) {
        const float d = s_d[b * k_pts + k];
        if (d < 0.f) continue;
        sum += (double)d;
        ++nv;
      }
      if (nv > 0) {
        total += sum / nv;
        samples += 1.0;
      }
    }
    acc[0] = total;
    acc[1] = samples;
  }
}

// splits per scene: one wave of 2 resident CTAs per SM over the whole grid (all CTAs run concurrently, no tail wave), at
// least ~1024 rows per CTA; 1 = single-pass kernel
static int eval_splits(int64_t n_rows, int64_t n_scenes, int k_pts) {
  if (n_scenes <= 0 || k_pts > kKpThreads) return 1;
  int64_t s = 2 * kNumSMs / n_scenes;
  const int64_t by_rows = n_rows / (n_scenes * 1024);
  if (s > by_rows) s = by_rows;
  return s < 2 ? 1 : (int)s;
}

}  // namespace ptv3

extern "C" size_t ptv3_keypoint_eval_workspace(int64_t n_rows, int64_t n_scenes, int k_pts) {
  const int s = eval_splits(n_rows, n_scenes, k_pts);
  return s > 1 ? (size_t)n_scenes * s * k_pts * sizeof(EvalPartial) : 0;
}

extern "C" int ptv3_keypoint_eval(const float* coord, const float* pred, const float* target, const int64_t* offset,
                                  const float* scale, int64_t n_rows, int64_t n_scenes, int k_pts, float* dist,
                                  int32_t* valid, double* acc, void* workspace, size_t workspace_bytes, void* stream) {
  if (n_scenes < 0 || n_rows < 0 || k_pts < 1 || k_pts > 65535) return PTV3_EINVAL;
  if (n_scenes == 0) return PTV3_OK;
  if (!coord || !pred || !target || !offset || !dist || !valid) return PTV3_EINVAL;
  const int splits = workspace ? eval_splits(n_rows, n_scenes, k_pts) : 1;
  if (splits > 1) {
    if (workspace_bytes < (size_t)n_scenes * splits * k_pts * sizeof(EvalPartial)) return PTV3_EWORKSPACE;
    EvalPartial* partials = static_cast<EvalPartial*>(workspace);
    dim3 grid((unsigned)n_scenes, (unsigned)splits);
    const int threads = k_pts * (kKpThreads / k_pts);
    keypoint_eval_split_kernel<<<grid, threads, 0, as_stream(stream)>>>(coord, pred, target, offset, k_pts, partials);
    const bool fold = acc && n_scenes * k_pts <= kEvalStage;
    keypoint_eval_finish_kernel<<<1, 1024, 0, as_stream(stream)>>>(coord, pred, scale, partials, splits, n_scenes, k_pts,
                                                                  dist, valid, fold ? acc : nullptr);
    if (fold) acc = nullptr;
  } else {
    dim3 grid((unsigned)n_scenes, (unsigned)k_pts);
    keypoint_eval_kernel<<<grid, kKpThreads, 0, as_stream(stream)>>>(coord, pred, target, offset, scale, k_pts, dist,
                                                                     valid);
  }
  if (acc) keypoint_eval_accumulate_kernel<<<1, 64, 0, as_stream(stream)>>>(dist, valid, n_scenes, k_pts, acc);
  return check_launch();
}
