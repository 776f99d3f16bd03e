// A4: serialized window attention, bf16 tensor-core kernel (head_dim 16).
//
// One CTA per (window, head): the window's K and V rows (<= 1024 x 16 bf16 each)
// are gathered ONCE into shared memory through the serialized order, then every
// warp walks 16-query tiles over 64-key chunks: S = Q K^T on the tensor cores,
// online softmax in registers (exp2 with the scale folded into one FFMA), P is
// re-used from the S accumulator registers as the A operand of P V, and the row
// sum comes from an extra MMA against a ones tile so it is the sum of exactly
// the bf16 P that multiplies V.  The [order] gather and the [inverse] scatter of
// the reference (point_transformer_v3m1_base.py:184-190, :216-218) are fused:
// rows are read from / written to storage order directly.
//
// head_dim 16 makes this kernel exp/issue bound, not tensor bound (SURVEY.md
// section 7): 64 tensor FLOPs per exp.
#include "common.cuh"

namespace ptv3 {

constexpr int kD = 16;
constexpr int kWarps = 8;
constexpr int kThreads = kWarps * 32;
constexpr int kChunk = 64;  // keys per softmax step

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

__device__ __forceinline__ void mma_bf16_16816(float c[4], const uint32_t a[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ void ldmatrix_x4(uint32_t r[4], uint32_t smem_addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(smem_addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t r[4], uint32_t smem_addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(smem_addr));
}

__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void* gptr) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr), "l"(gptr));
}

// smem: K [kpad][16] bf16, V [kpad][16] bf16, then the window's storage rows order_s [kpad] int32 (the queries
// of a window are a sub-range of its keys, so the Q gather and the output scatter index through it as well)
// kLse (training forward): also writes the rows' log-sum-exp, in log2 units of the scaled scores, for the backward
template <bool kLse>
__global__ void __launch_bounds__(kThreads, 3) attn_window_mma_kernel(
    const __nv_bfloat16* __restrict__ qkv, const int64_t* __restrict__ order, const int64_t* __restrict__ windows,
    const int64_t* __restrict__ n_windows, int h, float scale_log2, int kpad_max, __nv_bfloat16* __restrict__ out,
    float* __restrict__ lse, int* __restrict__ err) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  // heads of one window are neighbours in launch order: their 32-byte K / V / Q slices of the same 128-byte line
  // are fetched while the line is still in L2 (with the window index fastest, ncu showed 3.4 x the algorithmic bytes)
  const int head = (int)(blockIdx.x % (unsigned)h), w = (int)(blockIdx.x / (unsigned)h);
  if (w >= n_windows[0]) return;
  const int64_t q0 = windows[4 * w + 0], ql = windows[4 * w + 1], k0 = windows[4 * w + 2], kl = windows[4 * w + 3];
  if (ql <= 0 || kl <= 0) return;
  const int kpad = (int)((kl + kChunk - 1) / kChunk) * kChunk;
  // contract of the window records (ptv3_build_windows): keys fit the launch's shared memory and contain the queries
  if (kpad > kpad_max || q0 < k0 || q0 + ql > k0 + kl) {
    if (threadIdx.x == 0 && blockIdx.z == 0 && err) atomicOr(err, 1);
    return;
  }
  __nv_bfloat16* Ks = reinterpret_cast<__nv_bfloat16*>(smem_raw);
  __nv_bfloat16* Vs = Ks + (size_t)kpad * kD;
  int* order_s = reinterpret_cast<int*>(Vs + (size_t)kpad * kD);
  const uint32_t ks_addr = (uint32_t)__cvta_generic_to_shared(Ks);
  const uint32_t vs_addr = (uint32_t)__cvta_generic_to_shared(Vs);
  const int64_t rs = (int64_t)3 * h * kD;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  // ---- gather K and V rows of this (window, head): 4 x 16 B per key ----
  for (int idx = tid; idx < kpad * 4; idx += kThreads) {
    int key = idx >> 2, part = idx & 3;
    uint32_t dst = (part < 2 ? ks_addr : vs_addr) + (uint32_t)(key * 32 + (part & 1) * 16);
    if (key < kl) {
      int64_t row = order ? order[k0 + key] : k0 + key;
      if (part == 0) order_s[key] = (int)row;
      const __nv_bfloat16* src = qkv + row * rs + (int64_t)(part < 2 ? 1 : 2) * h * kD + head * kD + (part & 1) * 8;
      cp_async16(dst, src);
    } else {
      asm volatile("st.shared.v4.u32 [%0], {%1,%1,%1,%1};" ::"r"(dst), "r"(0u));
    }
  }
  asm volatile("cp.async.commit_group;");
  asm volatile("cp.async.wait_group 0;");
  __syncthreads();

  const int g = lane >> 2, t = lane & 3;
  const int n_qtiles = (int)((ql + 15) >> 4);
  const int n_chunks = kpad / kChunk;
  // ldmatrix lane addressing: lane supplies the row address of matrix (lane>>3), row (lane&7)
  //  K (non-trans): matrices {keys 0-7 dims 0-7, keys 0-7 dims 8-15, keys 8-15 dims 0-7, keys 8-15 dims 8-15}
  const uint32_t k_lane_off = (uint32_t)((((lane >> 4) << 3) + (lane & 7)) * 32 + ((lane >> 3) & 1) * 16);
  //  V (trans): matrices {keys 0-7 dims 0-7, keys 8-15 dims 0-7, keys 0-7 dims 8-15, keys 8-15 dims 8-15}
  const uint32_t v_lane_off = (uint32_t)(((((lane >> 3) & 1) << 3) + (lane & 7)) * 32 + (lane >> 4) * 16);
  const uint32_t ones = 0x3f803f80u;  // bf16 (1.0, 1.0)
  const float lazy_thr = 8.0f / scale_log2;   // raw-score distance that corresponds to a factor 2^8

  // blockIdx.z splits the query tiles of one (window, head) over gridDim.z CTAs (each stages K/V itself): finer
  // granularity for launches whose CTA count is a small non-integer number of waves
  // Q fragments (A operand, 16 x 16) come straight from global memory; the fragment of the warp's NEXT query tile is
  // requested before the key loop of the current one, so its L2 / HBM latency hides behind ~64 k scores of work
  const int q_rel = (int)(q0 - k0);   // queries are rows q_rel .. q_rel + ql - 1 of the staged key range
  auto load_q = [&](int qt, uint32_t (&qa)[4]) {
    qa[0] = qa[1] = qa[2] = qa[3] = 0u;
    const int qi0 = qt * 16 + g, qi1 = qi0 + 8;
    if (qi0 < ql) {
      const uint32_t* p = reinterpret_cast<const uint32_t*>(qkv + (int64_t)order_s[q_rel + qi0] * rs + head * kD);
      qa[0] = p[t];
      qa[2] = p[t + 4];
    }
    if (qi1 < ql) {
      const uint32_t* p = reinterpret_cast<const uint32_t*>(qkv + (int64_t)order_s[q_rel + qi1] * rs + head * kD);
      qa[1] = p[t];
      qa[3] = p[t + 4];
    }
  };
  const int qt_step = kWarps * (int)gridDim.z;
  uint32_t qa_next[4];
  {
    const int qt_first = warp * (int)gridDim.z + (int)blockIdx.z;
    if (qt_first < n_qtiles) load_q(qt_first, qa_next);
  }
  for (int qt = warp * (int)gridDim.z + (int)blockIdx.z; qt < n_qtiles; qt += qt_step) {
    uint32_t qa[4] = {qa_next[0], qa_next[1], qa_next[2], qa_next[3]};
    if (qt + qt_step < n_qtiles) load_q(qt + qt_step, qa_next);
    const int qi0 = qt * 16 + g, qi1 = qi0 + 8;
    const bool v0 = qi0 < ql, v1 = qi1 < ql;
    float o[2][4], l[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) o[0][i] = o[1][i] = l[i] = 0.f;
    float m0 = -INFINITY, m1 = -INFINITY;  // reference maxima (raw score units) of rows g and g+8

    for (int c = 0; c < n_chunks; ++c) {
      float s[8][4];
#pragma unroll
      for (int b = 0; b < 8; ++b) s[b][0] = s[b][1] = s[b][2] = s[b][3] = 0.f;
      const uint32_t kc = ks_addr + (uint32_t)(c * kChunk * 32);
#pragma unroll
      for (int b2 = 0; b2 < 4; ++b2) {
        uint32_t kb[4];
        ldmatrix_x4(kb, kc + (uint32_t)(b2 * 16 * 32) + k_lane_off);
        mma_bf16_16816(s[2 * b2], qa, kb[0], kb[1]);
        mma_bf16_16816(s[2 * b2 + 1], qa, kb[2], kb[3]);
      }
      if (c == n_chunks - 1 && kl < kpad) {
        const int base = c * kChunk + 2 * t;
#pragma unroll
        for (int b = 0; b < 8; ++b) {
          int key = base + b * 8;
          if (key >= kl) s[b][0] = s[b][2] = -INFINITY;
          if (key + 1 >= kl) s[b][1] = s[b][3] = -INFINITY;
        }
      }
      // Lazy rescaling: the running maxima m0 / m1 are only a reference point of the exponent, so they are moved
      // (quad shuffles, two exponentials, rescale of o and l) only when some row of the warp saw a score more than
      // 2^8 above its reference; otherwise P = 2^(s - m) <= 256 is as exact as before and the chunk skips all of it.
      float mx0 = s[0][0], mx1 = s[0][2];
#pragma unroll
      for (int b = 0; b < 8; ++b) {
        mx0 = fmaxf(mx0, fmaxf(s[b][0], s[b][1]));
        mx1 = fmaxf(mx1, fmaxf(s[b][2], s[b][3]));
      }
      if (__any_sync(0xffffffffu, mx0 > m0 + lazy_thr || mx1 > m1 + lazy_thr)) {
        mx0 = fmaxf(mx0, m0);
        mx1 = fmaxf(mx1, m1);
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        const float a0 = ex2((m0 - mx0) * scale_log2), a1 = ex2((m1 - mx1) * scale_log2);
        m0 = mx0;
        m1 = mx1;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          o[i][0] *= a0; o[i][1] *= a0; o[i][2] *= a1; o[i][3] *= a1;
        }
        l[0] *= a0; l[1] *= a0; l[2] *= a1; l[3] *= a1;
      }
      const float ms0 = m0 * scale_log2, ms1 = m1 * scale_log2;
      const uint32_t vc = vs_addr + (uint32_t)(c * kChunk * 32);
#pragma unroll
      for (int j = 0; j < 4; ++j) {  // 16 keys per step
        uint32_t pa[4];
        pa[0] = pack_bf16(ex2(fmaf(s[2 * j][0], scale_log2, -ms0)), ex2(fmaf(s[2 * j][1], scale_log2, -ms0)));
        pa[1] = pack_bf16(ex2(fmaf(s[2 * j][2], scale_log2, -ms1)), ex2(fmaf(s[2 * j][3], scale_log2, -ms1)));
        pa[2] = pack_bf16(ex2(fmaf(s[2 * j + 1][0], scale_log2, -ms0)), ex2(fmaf(s[2 * j + 1][1], scale_log2, -ms0)));
        pa[3] = pack_bf16(ex2(fmaf(s[2 * j + 1][2], scale_log2, -ms1)), ex2(fmaf(s[2 * j + 1][3], scale_log2, -ms1)));
        uint32_t vb[4];
        ldmatrix_x4_trans(vb, vc + (uint32_t)(j * 16 * 32) + v_lane_off);
        mma_bf16_16816(o[0], pa, vb[0], vb[1]);
        mma_bf16_16816(o[1], pa, vb[2], vb[3]);
        mma_bf16_16816(l, pa, ones, ones);
      }
    }
    // ---- epilogue: normalise, write bf16 pairs to storage rows ----
    if (v0) {
      float inv = 1.f / l[0];
      uint32_t* p = reinterpret_cast<uint32_t*>(out + (int64_t)order_s[q_rel + qi0] * h * kD + head * kD);
      p[t] = pack_bf16(o[0][0] * inv, o[0][1] * inv);
      p[t + 4] = pack_bf16(o[1][0] * inv, o[1][1] * inv);
      if (kLse && t == 0) lse[(int64_t)order_s[q_rel + qi0] * h + head] = m0 * scale_log2 + __log2f(l[0]);
    }
    if (v1) {
      float inv = 1.f / l[2];
      uint32_t* p = reinterpret_cast<uint32_t*>(out + (int64_t)order_s[q_rel + qi1] * h * kD + head * kD);
      p[t] = pack_bf16(o[0][2] * inv, o[0][3] * inv);
      p[t + 4] = pack_bf16(o[1][2] * inv, o[1][3] * inv);
      if (kLse && t == 0) lse[(int64_t)order_s[q_rel + qi1] * h + head] = m1 * scale_log2 + __log2f(l[2]);
    }
  }
}

int g_attn_qsplit = 0;   // 0 = auto; debug slot 2 forces 1 / 2 / 4

int attn_window_tc(const __nv_bfloat16* qkv, const int64_t* order, const int64_t* windows, const int64_t* n_windows,
                   int64_t n_windows_max, int64_t max_len, int64_t n, int h, float scale, __nv_bfloat16* out,
                   cudaStream_t st, float* lse) {
  (void)n;
  int kpad = (int)(cdiv(max_len, kChunk) * kChunk);
  size_t smem = (size_t)kpad * kD * 2 * 2 + (size_t)kpad * 4;
  if (smem > 200 * 1024) return PTV3_EUNSUPPORTED;  // patch > 3200
  static size_t configured_dev[kMaxDeviceSlots] = {};
  size_t& configured = configured_dev[device_slot()];
  if (smem > configured) {
    if (cudaFuncSetAttribute(attn_window_mma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) !=
            cudaSuccess ||
        cudaFuncSetAttribute(attn_window_mma_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) !=
            cudaSuccess)
      return PTV3_ELAUNCH;
    configured = smem;
  }
  // query split: pick s in {1,2,4} that minimises ceil(waves) / s-scaled work (3 CTAs per SM resident)
  int split = g_attn_qsplit;
  if (split <= 0) {
    const double slots = 3.0 * kNumSMs;
    double best = 1e30;
    split = 1;
    for (int s = 1; s <= 4; s *= 2) {
      const double ctas = (double)n_windows_max * h * s;
      // time ~ waves * (compute per CTA / s + K/V staging); staging is ~4 % of a full CTA's work
      const double t = ceil(ctas / slots) * (1.0 / s + 0.04);
      if (t < best - 1e-9) { best = t; split = s; }
    }
  }
  if ((int64_t)n_windows_max * h > 2147483647LL) return PTV3_EINVAL;
  dim3 grid((unsigned)(n_windows_max * h), 1u, (unsigned)split);
  if (lse)
    attn_window_mma_kernel<true><<<grid, kThreads, smem, st>>>(qkv, order, windows, n_windows, h,
                                                               scale * 1.4426950408889634f, kpad, out, lse,
                                                               device_error_flag());
  else
    attn_window_mma_kernel<false><<<grid, kThreads, smem, st>>>(qkv, order, windows, n_windows, h,
                                                                scale * 1.4426950408889634f, kpad, out, nullptr,
                                                                device_error_flag());
  return check_launch();
}

}  // namespace ptv3
