// A4: serialized window attention on the 5th-gen tensor cores (tcgen05 + TMEM), bf16, head_dim 16.
//
// One CTA per (window, head).  Q / K / V rows of the window are gathered ONCE through the serialized order (the
// reference's [order] gather, point_transformer_v3m1_base.py:184-190) with 16-byte cp.async into the shared-memory
// layouts UMMA reads directly:
//     Q, K : K-major, no swizzle   (8-row core matrices of 8x16 B; the two K halves 128 B apart, row groups 256 B)
//     V    : MN-major, no swizzle  (8-key core matrices; N blocks = [d 0-7 | d 8-15 | ones column | zeros])
// The ones column makes the P.V MMA also produce the softmax denominator (sum of exactly the bf16 P that multiplies V).
//
// Roles (576 threads): two "slots" work on two 128-query tiles at a time; a slot has 8 softmax warps (thread = one
// query row x 32 of the 64 keys of a chunk) and its own MMA-issuer warp (lane 0 issues tcgen05.mma), so the slots
// run decoupled and one slot's exps overlap the other's waits.  Per query tile and 64-key chunk:
//     S = Q K^T      tcgen05.mma M128 N64 K16  -> TMEM (two S buffers per slot)
//     pass 1         tcgen05.ld S, thread = query row: running max with 3-input max, no shuffles
//     pass 2         S again -> p = exp2(s*c - m*c) -> bf16 -> shared memory (128-byte swizzle, A operand)
//     O += P V       4 x tcgen05.mma M128 N32 K16 -> TMEM (32 columns: 16 outputs + denominator)
// Exact two-pass softmax (max first) instead of online rescaling: same instruction count per element, no
// correction step on the TMEM accumulator.  The kernel is bound by the exp (MUFU) pipe, not by the tensor pipe:
// head_dim 16 means 64 tensor FLOPs per exp.
// Output rows are written straight to storage order (the reference's [inverse] scatter, :216-218).
#include "common.cuh"
#include "tc05.cuh"

namespace ptv3 {

using namespace tc05;

namespace a5 {
constexpr int D = 16;
constexpr int QT = 128;        // queries per tile (UMMA M)
constexpr int KC = 64;         // keys per chunk (UMMA N of S, K extent of one P buffer)
constexpr int MAXQT = 8;       // <= 1024 queries per window
constexpr int MAXKC = 16;      // <= 1024 keys per window
constexpr int THREADS = 576;      // 16 softmax warps + 2 MMA-issuer warps
constexpr int Q_OFF = 0;                              // 8 tiles x 4096 B
constexpr int K_OFF = Q_OFF + MAXQT * QT * 32;        // 32768: 1024 keys x 32 B
constexpr int V_OFF = K_OFF + MAXKC * KC * 32;        // 65536: 128 key groups x 512 B
constexpr int P_OFF = V_OFF + MAXKC * KC * 64;        // 131072: [slot][buf] x 16384 B
constexpr int BAR_OFF = P_OFF + 4 * 16384;            // 196608
constexpr int MX_OFF = BAR_OFF + 256;                 // partial row maxima [2][2][128] fp32
constexpr int SMEM_BYTES = MX_OFF + 2048 + 1024;      // + alignment slack
// TMEM columns of slot s: S buffer b at s*256 + b*64, O at s*256 + 128 (32 columns)
constexpr int TMEM_COLS = 512;

__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float max3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
// no-swizzle descriptor: lbo = byte stride between core matrices along K, sbo = along M/N
__device__ __forceinline__ uint64_t desc_plain(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) |
         ((uint64_t)1 << 46);
}
}  // namespace a5

int g_attn_tc05_variant = 0;   // debug knob (tests): 1 swaps LBO/SBO of the MN-major V descriptor

__global__ void __launch_bounds__(a5::THREADS, 1) attn_window_tc05_kernel(
    const __nv_bfloat16* __restrict__ qkv, const int64_t* __restrict__ order, const int64_t* __restrict__ windows,
    const int64_t* __restrict__ n_windows, int h, float scale_log2, int variant, __nv_bfloat16* __restrict__ out) {
  using namespace a5;
  const int w = blockIdx.x, head = blockIdx.y;
  if (w >= n_windows[0]) return;
  const int64_t q0 = windows[4 * w + 0], k0 = windows[4 * w + 2];
  const int ql = (int)windows[4 * w + 1], kl = (int)windows[4 * w + 3];
  if (ql <= 0 || kl <= 0 || ql > MAXQT * QT || kl > MAXKC * KC) return;
  const int n_qt = (ql + QT - 1) / QT, nc = (kl + KC - 1) / KC;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t sb = smem_u32(smem);
  const uint32_t bar0 = sb + BAR_OFF;
  // barrier ids per slot s: +0,1 s_full[b]  +2,3 s_free[b]  +4,5 p_full[b]  +6,7 p_free[b]  +8 o_full  +9 o_free
  auto bar = [&](int s, int id) { return bar0 + 8u * (uint32_t)(s * 10 + id); };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + BAR_OFF + 200);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar(s, 0), 1); mbar_init(bar(s, 1), 1);
      mbar_init(bar(s, 2), 256); mbar_init(bar(s, 3), 256);
      mbar_init(bar(s, 4), 256); mbar_init(bar(s, 5), 256);
      mbar_init(bar(s, 6), 1); mbar_init(bar(s, 7), 1);
      mbar_init(bar(s, 8), 1); mbar_init(bar(s, 9), 128);
    }
    mbar_fence_init();
  }
  if (warp == 16) tmem_alloc<TMEM_COLS>(smem_u32(tmem_slot));

  // ---------------- stage Q, K, V of this (window, head) ----------------
  const int64_t rs = (int64_t)3 * h * D;
  const __nv_bfloat16* qbase = qkv + head * D;
  for (int j = tid; j < nc * KC; j += THREADS) {
    const uint32_t kd = sb + K_OFF + (j >> 3) * 256 + (j & 7) * 16;
    const uint32_t vd = sb + V_OFF + (j >> 3) * 512 + (j & 7) * 16;
    if (j < kl) {
      const int64_t row = order ? order[k0 + j] : k0 + j;
      const __nv_bfloat16* kp = qbase + row * rs + (int64_t)h * D;
      const __nv_bfloat16* vp = kp + (int64_t)h * D;
      cp_async16(kd, kp, 16u);
      cp_async16(kd + 128, kp + 8, 16u);
      cp_async16(vd, vp, 16u);
      cp_async16(vd + 128, vp + 8, 16u);
      asm volatile("st.shared.v4.u32 [%0], {%1, %2, %2, %2};" ::"r"(vd + 256), "r"(0x00003f80u), "r"(0u) : "memory");
    } else {
      st_shared_zero16(kd); st_shared_zero16(kd + 128);
      st_shared_zero16(vd); st_shared_zero16(vd + 128); st_shared_zero16(vd + 256);
    }
    st_shared_zero16(vd + 384);
  }
  for (int i = tid; i < n_qt * QT; i += THREADS) {
    const uint32_t qd = sb + Q_OFF + (i >> 7) * 4096 + ((i & 127) >> 3) * 256 + (i & 7) * 16;
    if (i < ql) {
      const int64_t row = order ? order[q0 + i] : q0 + i;
      const __nv_bfloat16* qp = qbase + row * rs;
      cp_async16(qd, qp, 16u);
      cp_async16(qd + 128, qp + 8, 16u);
    } else {
      st_shared_zero16(qd); st_shared_zero16(qd + 128);
    }
  }
  cp_async_commit();
  cp_async_wait<0>();
  fence_proxy_async();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = *tmem_slot;

  if (warp >= 16) {
    // =========================== MMA issuers: warp 16 -> slot 0, warp 17 -> slot 1 ===========================
    if (lane == 0) {
      const int s = warp - 16;
      constexpr uint32_t idesc_s = idesc_bf16_f32(QT, KC);                    // S: M128 N64, A and B K-major
      constexpr uint32_t idesc_o = idesc_bf16_f32(QT, 32) | (1u << 16);       // O: M128 N32, B (V) MN-major
      const uint32_t v_lbo = (variant & 1) ? 128u : 512u, v_sbo = (variant & 1) ? 512u : 128u;
      uint32_t gs = 0, gp = 0;
      auto issue_qk = [&](int t, int c) {
        const uint32_t b = gs & 1u;
        mbar_wait(bar(s, 2 + b), ((gs >> 1) & 1u) ^ 1u);
        fence_after_sync();
        mma_bf16_ss(tmem + s * 256 + b * 64, desc_plain(sb + Q_OFF + t * 4096, 128, 256),
                    desc_plain(sb + K_OFF + c * 2048, 128, 256), idesc_s, 0u);
        mma_commit(bar(s, b));
        ++gs;
      };
      auto issue_pv = [&](int r, int c) {
        const uint32_t b = gp & 1u;
        mbar_wait(bar(s, 4 + b), (gp >> 1) & 1u);
        if (c == 0) mbar_wait(bar(s, 9), ((uint32_t)r & 1u) ^ 1u);   // O of the previous tile has been read
        fence_after_sync();
        const uint32_t p_s = sb + P_OFF + (s * 2 + b) * 16384;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          mma_bf16_ss(tmem + s * 256 + 128, smem_desc_sw128(p_s) + 2 * kk,
                      desc_plain(sb + V_OFF + c * 4096 + kk * 1024, v_lbo, v_sbo), idesc_o, (c | kk) ? 1u : 0u);
        mma_commit(bar(s, 6 + b));
        if (c == nc - 1) mma_commit(bar(s, 8));
        ++gp;
      };
      for (int r = 0; 2 * r + s < n_qt; ++r) {
        const int t = 2 * r + s;
        if (!(variant & 2))
          for (int c = 0; c < nc; ++c) issue_qk(t, c);
        for (int c = 0; c <= nc; ++c) {
          if (c < nc) issue_qk(t, c);
          if (c >= 1) issue_pv(r, c - 1);
        }
      }
    }
  } else {
    // =========================== softmax: 8 warps per slot ===========================
    // warp w: slot w>>3, column half (w>>2)&1 of every 64-key chunk, TMEM lane quarter w&3 (the hardware rule:
    // a warp may only touch lanes 32*(warp%4)..+31); thread = one query row x 32 keys per chunk.
    const int s = warp >> 3;
    const int half = (warp >> 2) & 1;
    const int row = (warp & 3) * 32 + lane;
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    float* mx = reinterpret_cast<float*>(smem + MX_OFF);   // [slot][half][128] partial row maxima
    uint32_t gs = 0, gp = 0;
    const int rem_last = kl - (nc - 1) * KC - half * 32;   // valid keys of my half of the last chunk
    for (int r = 0; 2 * r + s < n_qt; ++r) {
      const int t = 2 * r + s;
      const int qi = t * QT + row;
      int64_t orow = -1;
      if (half == 0 && qi < ql) orow = order ? order[q0 + qi] : q0 + qi;
      uint32_t v[32];   // raw fp32 bits of my 32 columns of one S chunk
      auto load_chunk = [&](int c) {
        const uint32_t b = gs & 1u;
        mbar_wait(bar(s, b), (gs >> 1) & 1u);
        fence_after_sync();
        tmem_ld32(tmem + lane_base + s * 256 + b * 64 + half * 32, v);
        tmem_ld_wait();
        fence_before_sync();
        mbar_arrive(bar(s, 2 + b));
        ++gs;
        if (c == nc - 1 && rem_last < 32) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j >= rem_last) v[j] = 0xff800000u;   // -inf
        }
      };
      // ---- pass 1: row maximum (two independent chains per thread, halves combined through shared memory) ----
      float m0 = -INFINITY, m1 = -INFINITY;
      if (variant & 2) m0 = 30.f;
      for (int c = 0; c < nc && !(variant & 2); ++c) {
        load_chunk(c);
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          m0 = max3(m0, __uint_as_float(v[j]), __uint_as_float(v[j + 1]));
          m1 = max3(m1, __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
        }
      }
      float m = fmaxf(m0, m1);
      mx[(s * 2 + half) * 128 + row] = m;
      asm volatile("bar.sync %0, 256;" ::"r"(1 + s) : "memory");
      m = fmaxf(m, mx[(s * 2 + (half ^ 1)) * 128 + row]);
      const float msc = m * scale_log2;
      // ---- pass 2: P = exp2(s*c - m*c) as bf16 into the swizzled A-operand buffer ----
      for (int c = 0; c < nc; ++c) {
        load_chunk(c);
        uint32_t pk[16];
        if (variant & 4) {
#pragma unroll
          for (int j = 0; j < 16; ++j)
            pk[j] = pack2(fmaf(__uint_as_float(v[2 * j]), scale_log2, -msc),
                          fmaf(__uint_as_float(v[2 * j + 1]), scale_log2, -msc));
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j)
            pk[j] = pack2(ex2f(fmaf(__uint_as_float(v[2 * j]), scale_log2, -msc)),
                          ex2f(fmaf(__uint_as_float(v[2 * j + 1]), scale_log2, -msc)));
        }
        const uint32_t b = gp & 1u;
        mbar_wait(bar(s, 6 + b), ((gp >> 1) & 1u) ^ 1u);
        const uint32_t prow = sb + P_OFF + (s * 2 + b) * 16384 + (row >> 3) * 1024 + (row & 7) * 128;
#pragma unroll
        for (int i = 0; i < 4; ++i)
          asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(prow + (((half * 4 + i) ^ (row & 7)) << 4)),
                       "r"(pk[4 * i]), "r"(pk[4 * i + 1]), "r"(pk[4 * i + 2]), "r"(pk[4 * i + 3])
                       : "memory");
        fence_proxy_async();
        mbar_arrive(bar(s, 4 + b));
        ++gp;
      }
      // ---- epilogue (column-half 0 warps): O / denominator -> storage row ----
      if (half == 0) {
        mbar_wait(bar(s, 8), (uint32_t)r & 1u);
        fence_after_sync();
        uint32_t o[32];
        tmem_ld32(tmem + lane_base + s * 256 + 128, o);
        tmem_ld_wait();
        fence_before_sync();
        mbar_arrive(bar(s, 9));
        if (orow >= 0) {
          const float inv = 1.f / __uint_as_float(o[16]);
          uint32_t pk[8];
#pragma unroll
          for (int j = 0; j < 8; ++j)
            pk[j] = pack2(__uint_as_float(o[2 * j]) * inv, __uint_as_float(o[2 * j + 1]) * inv);
          uint4* dst = reinterpret_cast<uint4*>(out + orow * (int64_t)h * D + head * D);
          dst[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
          dst[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
        }
      }
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 16) tmem_dealloc<a5::TMEM_COLS>(tmem);
}

int attn_window_tc05(const __nv_bfloat16* qkv, const int64_t* order, const int64_t* windows,
                     const int64_t* n_windows, int64_t n_windows_max, int64_t max_len, int h, float scale,
                     __nv_bfloat16* out, cudaStream_t st) {
  if (max_len > a5::MAXKC * a5::KC) return PTV3_EUNSUPPORTED;
  static bool configured_dev[kMaxDeviceSlots] = {};
  bool& configured = configured_dev[device_slot()];
  if (!configured) {
    if (cudaFuncSetAttribute(attn_window_tc05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, a5::SMEM_BYTES) !=
        cudaSuccess)
      return PTV3_ELAUNCH;
    configured = true;
  }
  dim3 grid((unsigned)n_windows_max, (unsigned)h);
  attn_window_tc05_kernel<<<grid, a5::THREADS, a5::SMEM_BYTES, st>>>(qkv, order, windows, n_windows, h,
                                                                      scale * 1.4426950408889634f,
                                                                      g_attn_tc05_variant, out);
  return check_launch();
}

}  // namespace ptv3
