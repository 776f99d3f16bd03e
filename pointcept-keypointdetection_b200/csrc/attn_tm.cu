// A4: serialized window attention on the 5th-generation tensor cores (tcgen05 + TMEM), head_dim 16.
//
// Reference: SerializedAttention.forward, point_transformer_v3m1_base.py:172-222 (flash branch :208-215).
//
// Design (the measurements behind every choice: profiles/r02_attention.md):
//   * S is read out of tensor memory ONCE.  The exponent reference of a row is the maximum over its FIRST chunk of keys
//     and never moves: P = 2^(s - m) is stored as bf16 (8 exponent bits) and accumulated in fp32, so a later score
//     above the reference needs no rescaling of O -- nothing overflows until a score exceeds the reference by ~2^100,
//     nothing of weight is lost because the reference is a lower bound of the true maximum.  The epilogue checks every
//     denominator (finite, > 0); a (window, head) that fails is flagged and recomputed by the register kernel
//     (attn_f16.cu) inside the same C-ABI call -- scores that spread over more than 100 octaves inside one row.
//   * per step: S = Q K^T for 128 queries x 32 keys (tcgen05.mma M128 N32 K16) -> tcgen05.ld, one FFMA + one MUFU (or
//     an FMA-pipe polynomial) per score, bf16x2 -> tcgen05.st: P goes back to TENSOR MEMORY and is the A operand of
//     P V from there (tcgen05.mma with A in TMEM): no shared-memory round trip, no generic->async proxy fence.  The
//     denominator comes out of the same P V MMA through a ones column in V (N = 32: 16 dims | ones | zeros).
//   * MMA count is a resource: a tcgen05.mma M128 x N x K16 costs >= 45.6 clk for every N <= 64 (tools/ubench/pipes.cu),
//     so a step issues exactly three (Q K^T + two K = 16 slices of P V = 137 clk per 4 096 scores, a 30 scores / clk /
//     SM ceiling); a variant with a separate ones MMA (5 per step) was capped at 14 scores / clk / SM by the issue rate.
//   * one CTA per SM with 20 exponential warps = 5 step classes x 4 TMEM lane quarters; warp (q, i) owns rows
//     32q..32q+31 of every step g = i (mod 5), so the five exponential warps of a scheduler sit at different steps and
//     the control / TMEM phase of one overlaps the exponential phase of the others (a first version gave every warp a
//     corner of every tile: all warps marched in lockstep and the exponential unit idled, ncu XU 54 %; four classes
//     reached 58 - 70 %).
//   * ONE ISSUER THREAD PER STEP CLASS: wait -> tcgen05.mma -> tcgen05.commit costs a single thread ~400 clk per
//     round trip whatever the tile size, so one (or two) issuer threads for all classes capped the kernel at one step
//     per ~420 clk (ncu: 57 % of the stall samples = exponential warps waiting for their S tile).  Each class's issuer
//     refills the class's S tile as soon as its owner has read it (Q K^T of step g + 5 runs during the exponentials of
//     step g) and fires P V when the owner has written P.  O is zeroed by the epilogue and only ever accumulated,
//     because the issuers of one query tile are not ordered among themselves.
// Output rows are written straight to storage order (the reference's [inverse] scatter, :216-218).
#include "common.cuh"
#include "tc05.cuh"

namespace ptv3 {

using namespace tc05;

namespace tm {
constexpr int D = 16;
constexpr int QT = 128;          // queries per tile (UMMA M)
constexpr int KC = 32;           // keys per step (UMMA N of S, two K = 16 slices of P V)
constexpr int MAXK = 1024;       // keys / queries per window
constexpr int NCLS = 3;          // step classes = S / P tiles in flight (per CTA; two CTAs per SM)
constexpr int N_EXP = 4 * NCLS;  // exponential warps
constexpr int THREADS = (N_EXP + NCLS) * 32;   // + one MMA-issuer warp per step class
constexpr int Q_OFF = 0;                          // 8 tiles x 4096 B   (K-major, no swizzle, 8-row core matrices)
constexpr int K_OFF = Q_OFF + MAXK * 32;          // 32768
constexpr int V_OFF = K_OFF + MAXK * 32;          // 65536: 128 key groups x 256 B (MN-major: d0-7 | d8-15)
constexpr int ONES_OFF = V_OFF + MAXK * 32;       // 98304: 2 key groups x 256 B, column 0 = 1 (the denominator)
constexpr int BAR_OFF = ONES_OFF + 512;           // 98816
constexpr int MX_OFF = BAR_OFF + 256;             // exponent references of 4 query tiles in flight: [4][128 rows] fp32
constexpr int SMEM_BYTES = MX_OFF + 2048 + 64 + 1024;  // + the TMEM address slot + alignment slack
constexpr int TMEM_COLS = 256;                    // S: 3 x 32 | P: 3 x 16 at 96 | O: 2 x 32 at 144
constexpr int P_COL = NCLS * KC;
constexpr int O_COL = P_COL + NCLS * 16;

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
// D[tmem] (+)= A[tmem] * B[smem]: A = P, 128 lanes (rows) x 8 columns per 16 keys (two bf16 per 32-bit cell)
__device__ __forceinline__ void mma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t b_desc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::
          "r"(taddr),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
      "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// 2^x on the FMA pipe (see attn_f16.cu): degree-3 minimax after a round-to-nearest split, max relative error 7.5e-5
__device__ __forceinline__ float ex2_poly(float x) {
  x = fmaxf(x, -125.f);
  const float t = __fadd_rn(x, 12582912.f);
  const float r = __fsub_rn(x, __fsub_rn(t, 12582912.f));
  float p = __fmaf_rn(r, 0.0551716685295105f, 0.2426111251115799f);
  p = __fmaf_rn(p, r, 0.6932609677314758f);
  p = __fmaf_rn(p, r, 0.9999280571937561f);
  return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}
// no-swizzle descriptor: lbo = byte stride between core matrices along K, sbo = along M/N
__device__ __forceinline__ uint64_t desc_plain(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) |
         ((uint64_t)1 << 46);
}
}  // namespace tm

// POLYN: every 8 consecutive keys, POLYN exponentials go to the FMA-pipe polynomial instead of the MUFU unit
template <int POLYN>
__global__ void __launch_bounds__(tm::THREADS, 2) attn_window_tm_kernel(
    const __nv_bfloat16* __restrict__ qkv, const int64_t* __restrict__ order, const int64_t* __restrict__ windows,
    const int64_t* __restrict__ n_windows, int h, float scale_log2, __nv_bfloat16* __restrict__ out,
    int* __restrict__ redo /* (n_windows_max * h) flags: 1 = recompute with the register kernel */) {
  using namespace tm;
  const int head = (int)(blockIdx.x % (unsigned)h), w = (int)(blockIdx.x / (unsigned)h);
  if (w >= n_windows[0]) return;
  const int64_t q0 = windows[4 * w + 0], k0 = windows[4 * w + 2];
  const int ql = (int)windows[4 * w + 1], kl = (int)windows[4 * w + 3];
  if (ql <= 0 || kl <= 0) return;
  if (ql > MAXK || kl > MAXK || q0 < k0 || q0 + ql > k0 + kl) {     // outside this kernel's contract: register kernel
    if (threadIdx.x == 0) redo[blockIdx.x] = 1;
    return;
  }
  const int n_qt = (ql + QT - 1) / QT, nc = (kl + KC - 1) / KC;
  const int G = n_qt * nc;      // (query tile, key chunk) steps, query tile major

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t sb = smem_u32(smem);
  // barriers: s_full[NCLS] s_free[NCLS] p_full[NCLS] p_free[NCLS] m_full[4] o_full[2] o_free[2]
  const uint32_t bar0 = sb + BAR_OFF;
  auto s_full = [&](uint32_t b) { return bar0 + 8u * b; };
  auto s_free = [&](uint32_t b) { return bar0 + 8u * (NCLS + b); };
  auto p_full = [&](uint32_t b) { return bar0 + 8u * (2 * NCLS + b); };
  auto p_free = [&](uint32_t b) { return bar0 + 8u * (3 * NCLS + b); };
  auto m_full = [&](uint32_t b) { return bar0 + 8u * (4 * NCLS + b); };
  auto o_full = [&](uint32_t b) { return bar0 + 8u * (4 * NCLS + 4 + b); };
  auto o_free = [&](uint32_t b) { return bar0 + 8u * (4 * NCLS + 6 + b); };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + MX_OFF + 2048);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) {
    for (uint32_t b = 0; b < NCLS; ++b) {     // 4 = one arrival per lane quarter (the four warps that own a step)
      mbar_init(s_full(b), 1); mbar_init(s_free(b), 4);
      mbar_init(p_full(b), 4); mbar_init(p_free(b), 1);
    }
    for (uint32_t b = 0; b < 4; ++b) mbar_init(m_full(b), 4);
    // o_full: one commit per class that owns a chunk of the query tile (its last P V of the tile)
    for (uint32_t b = 0; b < 2; ++b) { mbar_init(o_full(b), (uint32_t)(nc < NCLS ? nc : NCLS)); mbar_init(o_free(b), 4); }
    mbar_fence_init();
  }
  if (warp == N_EXP) tmem_alloc<TMEM_COLS>(smem_u32(tmem_slot));

  // ---------------- stage Q, K, V of this (window, head) through the serialized order ----------------
  const int64_t rs = (int64_t)3 * h * D;
  const __nv_bfloat16* qbase = qkv + head * D;
  const int q_rel = (int)(q0 - k0);
  for (int j = tid; j < nc * KC; j += THREADS) {
    const uint32_t kd = sb + K_OFF + (j >> 3) * 256 + (j & 7) * 16;
    const uint32_t vd = sb + V_OFF + (j >> 3) * 256 + (j & 7) * 16;
    if (j < kl) {
      const int64_t row = order ? order[k0 + j] : k0 + j;
      const __nv_bfloat16* qp = qbase + row * rs;
      const __nv_bfloat16* kp = qp + (int64_t)h * D;
      const __nv_bfloat16* vp = kp + (int64_t)h * D;
      cp_async16(kd, kp, 16u);
      cp_async16(kd + 128, kp + 8, 16u);
      cp_async16(vd, vp, 16u);
      cp_async16(vd + 128, vp + 8, 16u);
      const int qi = j - q_rel;      // this key row is query qi of the window (queries are a sub-range of the keys)
      if (qi >= 0 && qi < ql) {
        const uint32_t qd = sb + Q_OFF + (qi >> 3) * 256 + (qi & 7) * 16;
        cp_async16(qd, qp, 16u);
        cp_async16(qd + 128, qp + 8, 16u);
      }
    } else {
      st_shared_zero16(kd); st_shared_zero16(kd + 128);
      st_shared_zero16(vd); st_shared_zero16(vd + 128);
    }
  }
  if (tid < 32) {     // the ones block: 16 keys x 16 columns, column 0 = 1.0 (bf16), same MN-major layout as V
    const uint32_t od = sb + ONES_OFF + (tid >> 4) * 256 + ((tid >> 1) & 7) * 16 + (tid & 1) * 128;
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %2, %2};" ::"r"(od), "r"((tid & 1) ? 0u : 0x00003f80u), "r"(0u) : "memory");
  }
  for (int i = ql + tid; i < n_qt * QT; i += THREADS) {     // rows of the last query tile past the last query
    const uint32_t qd = sb + Q_OFF + (i >> 3) * 256 + (i & 7) * 16;
    st_shared_zero16(qd); st_shared_zero16(qd + 128);
  }
  cp_async_commit();
  cp_async_wait<0>();
  fence_proxy_async();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = *tmem_slot;

  if (warp < 4) {     // O accumulators start at zero: the P V MMAs only ever accumulate (see the issuers)
    uint32_t z[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) z[j] = 0u;
    const uint32_t lb = (uint32_t)(warp * 32) << 16;
#pragma unroll
    for (int k = 0; k < 4; ++k) tmem_st16(tmem + lb + O_COL + k * 16, z);
    tmem_st_wait();
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();

  if (warp >= N_EXP) {
    // =========================== MMA issuer of step class cls = warp - N_EXP ===========================
    if (elect_one_sync()) {
      const int cls = warp - N_EXP;
      constexpr uint32_t idesc_s = idesc_bf16_f32(QT, KC);                    // S: M128 N32, A and B K-major
      constexpr uint32_t idesc_o = idesc_bf16_f32(QT, 16) | (1u << 16);       // M128 N16, B MN-major (V and the ones block)
      const uint64_t ones_desc = desc_plain(sb + ONES_OFF, 256, 128);
      const uint32_t b = (uint32_t)cls;
      int tq = 0, cq = cls;    // (query tile, chunk) of this class's next Q K^T
      while (cq >= nc) { cq -= nc; ++tq; }
      int t = tq, c = cq;      // ... and of its next P V
      auto issue_qk = [&]() {
        mma_bf16_ss(tmem + b * KC, desc_plain(sb + Q_OFF + tq * 4096, 128, 256),
                    desc_plain(sb + K_OFF + cq * 1024, 128, 256), idesc_s, 0u);
        mma_commit(s_full(b));
        cq += NCLS;
        while (cq >= nc) { cq -= nc; ++tq; }
      };
      if (cls < G) issue_qk();
      int t_seen = -1;
      uint32_t use = 0;
      for (int g = cls; g < G; g += NCLS, ++use) {
        if (g + NCLS < G) {     // the owner has read S(g): refill the tile with step g + NCLS while it computes
          mbar_wait(s_free(b), use & 1u);
          fence_after_sync();
          issue_qk();
        }
        mbar_wait(p_full(b), use & 1u);
        const uint32_t ob = (uint32_t)t & 1u;
        if (t != t_seen) {      // my first P V of this query tile: the epilogue of tile t-2 has read and zeroed O
          t_seen = t;
          mbar_wait(o_free(ob), (((uint32_t)t >> 1) & 1u) ^ 1u);
        }
        fence_after_sync();
#pragma unroll
        for (int kk = 0; kk < 2; ++kk) {     // 16 keys per MMA = 8 TMEM columns of P; always accumulate
          mma_bf16_ts(tmem + O_COL + ob * 32, tmem + P_COL + b * 16 + kk * 8,
                      desc_plain(sb + V_OFF + c * 1024 + kk * 512, 256, 128), idesc_o, 1u);
          mma_bf16_ts(tmem + O_COL + ob * 32 + 16, tmem + P_COL + b * 16 + kk * 8, ones_desc, idesc_o, 1u);
        }
        mma_commit(p_free(b));
        if (c + NCLS >= nc) mma_commit(o_full(ob));      // my last chunk of this tile
        c += NCLS;
        while (c >= nc) { c -= nc; ++t; }
      }
    }
  } else {
    // =========================== exponential warps ===========================
    // warp -> TMEM lane quarter q = warp % 4 (the hardware rule: a warp may only touch lanes 32*(warp%4)..+31) and
    // step class i = warp / 4: the warp owns rows 32q..32q+31 of every step g = i (mod NCLS); thread = one query row.
    const int q = warp & 3, i = warp >> 2;
    const int row = q * 32 + lane;
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    float* mx = reinterpret_cast<float*>(smem + MX_OFF);   // [tile & 3][row]: exponent reference of the row
    const uint32_t s_addr = tmem + lane_base + (uint32_t)i * KC, p_addr = tmem + lane_base + P_COL + (uint32_t)i * 16;
    bool bad = false;
    float msc = 0.f;       // reference maximum of my row of the current query tile, times scale * log2(e)
    int t = 0, c = i, t_ref = -1;
    while (c >= nc) { c -= nc; ++t; }
    uint32_t par = 0;
    for (int g = i; g < G; g += NCLS, par ^= 1u) {
      uint32_t v[32];
      mbar_wait_park(s_full((uint32_t)i), par, 4000);
      fence_after_sync();
      tmem_ld32(s_addr, v);
      tmem_ld_wait();
      fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(s_free((uint32_t)i));
      if (c == nc - 1) {
        const int rem = kl - c * KC;      // valid keys of the last chunk
        if (rem < KC) {
#pragma unroll
          for (int j = 0; j < KC; ++j)
            if (j >= rem) v[j] = 0xff800000u;   // -inf
        }
      }
      if (t != t_ref) {     // my first step in this query tile: fetch (or, for chunk 0, define) the row's reference
        t_ref = t;
        const uint32_t tb = (uint32_t)t & 3u;
        if (c == 0) {
          float m0 = __uint_as_float(v[0]), m1 = __uint_as_float(v[1]);
#pragma unroll
          for (int j = 2; j < KC; j += 4) {
            m0 = max3(m0, __uint_as_float(v[j]), __uint_as_float(v[j + 1]));
            if (j + 3 < KC) m1 = max3(m1, __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
          }
          const float m = fmaxf(m0, m1);
          mx[tb * 128 + row] = m;
          msc = m * scale_log2;
          __syncwarp();
          if (lane == 0) mbar_arrive(m_full(tb));      // default .release.cta: the stores above are visible to waiters
        } else {
          mbar_wait_park(m_full(tb), ((uint32_t)t >> 2) & 1u, 4000);
          msc = mx[tb * 128 + row] * scale_log2;
        }
      }
      uint32_t pk[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const float x0 = fmaf(__uint_as_float(v[2 * j]), scale_log2, -msc);
        const float x1 = fmaf(__uint_as_float(v[2 * j + 1]), scale_log2, -msc);
        // the odd key of a pair goes to the polynomial when its slot in the 8-key group is 1 (POLYN >= 1), 5 (>= 2),
        // 3 (>= 3): the two kinds of exponentials alternate evenly in the instruction stream
        const int slot = (2 * j + 1) & 7;
        const bool poly1 = (slot == 1 && POLYN >= 1) || (slot == 5 && POLYN >= 2) || (slot == 3 && POLYN >= 3);
        if (POLYN == 9) {   // TIMING ABLATION (wrong results): no exponentials at all
          pk[j] = pack2(fmaf(x0, 0.001f, 1.f), fmaf(x1, 0.001f, 1.f));
          continue;
        }
        pk[j] = pack2(ex2f(x0), poly1 ? ex2_poly(x1) : ex2f(x1));
      }
      mbar_wait_park(p_free((uint32_t)i), par ^ 1u, 4000);
      fence_after_sync();
      tmem_st16(p_addr, pk);
      tmem_st_wait();
      fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full((uint32_t)i));
      // ---- epilogue (the owner of a tile's last chunk): O / denominator -> storage row ----
      if (c == nc - 1) {
        const uint32_t ob = (uint32_t)t & 1u;
        mbar_wait_park(o_full(ob), ((uint32_t)t >> 1) & 1u, 4000);
        fence_after_sync();
        uint32_t o[16], lbits;
        tmem_ld16(tmem + lane_base + O_COL + ob * 32, o);
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(lbits) : "r"(tmem + lane_base + O_COL + ob * 32 + 16) : "memory");
        tmem_ld_wait();
        {     // leave the accumulator zeroed for tile t + 2 (the issuers only accumulate)
          uint32_t z[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) z[j] = 0u;
          tmem_st16(tmem + lane_base + O_COL + ob * 32, z);
          tmem_st16(tmem + lane_base + O_COL + ob * 32 + 16, z);
          tmem_st_wait();
        }
        fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(o_free(ob));
        const int qi = t * QT + row;
        if (qi < ql) {
          const float l = __uint_as_float(lbits);
          if (!(l > 0.f) || !(l < 3.0e38f)) bad = true;      // NaN, inf, 0: scores spread over > 100 octaves
          const float inv = 1.f / l;
          uint32_t ok[8];
#pragma unroll
          for (int j = 0; j < 8; ++j)
            ok[j] = pack2(__uint_as_float(o[2 * j]) * inv, __uint_as_float(o[2 * j + 1]) * inv);
          const int64_t orow = order ? order[q0 + qi] : q0 + qi;
          uint4* dst = reinterpret_cast<uint4*>(out + orow * (int64_t)h * D + head * D);
          dst[0] = make_uint4(ok[0], ok[1], ok[2], ok[3]);
          dst[1] = make_uint4(ok[4], ok[5], ok[6], ok[7]);
        }
      }
      c += NCLS;
      while (c >= nc) { c -= nc; ++t; }
    }
    if (__any_sync(0xffffffffu, bad) && lane == 0) redo[blockIdx.x] = 1;
  }
  fence_before_sync();
  __syncthreads();
  if (warp == tm::N_EXP) tmem_dealloc<tm::TMEM_COLS>(tmem);
}

int g_attn_tm_poly = 2;   // debug slot 9: eighths of the exponentials on the FMA-pipe polynomial (0 .. 3); 2 is the fastest

int attn_window_tm(const __nv_bfloat16* qkv, const int64_t* order, const int64_t* windows, const int64_t* n_windows,
                   int64_t n_windows_max, int64_t max_len, int h, float scale, __nv_bfloat16* out, int* redo,
                   cudaStream_t st) {
  if (max_len > tm::MAXK) return PTV3_EUNSUPPORTED;
  if ((int64_t)n_windows_max * h > 2147483647LL) return PTV3_EINVAL;
  static bool configured[16] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  dev &= 15;
  if (!configured[dev]) {
    if (cudaFuncSetAttribute(attn_window_tm_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, tm::SMEM_BYTES) !=
            cudaSuccess ||
        cudaFuncSetAttribute(attn_window_tm_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, tm::SMEM_BYTES) !=
            cudaSuccess ||
        cudaFuncSetAttribute(attn_window_tm_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, tm::SMEM_BYTES) !=
            cudaSuccess ||
        cudaFuncSetAttribute(attn_window_tm_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, tm::SMEM_BYTES) !=
            cudaSuccess ||
        cudaFuncSetAttribute(attn_window_tm_kernel<9>, cudaFuncAttributeMaxDynamicSharedMemorySize, tm::SMEM_BYTES) !=
            cudaSuccess)
      return PTV3_ELAUNCH;
    configured[dev] = true;
  }
  const unsigned grid = (unsigned)(n_windows_max * h);
  const float sl2 = scale * 1.4426950408889634f;
  switch (g_attn_tm_poly) {
    case 1: attn_window_tm_kernel<1><<<grid, tm::THREADS, tm::SMEM_BYTES, st>>>(qkv, order, windows, n_windows, h, sl2, out, redo); break;
    case 2: attn_window_tm_kernel<2><<<grid, tm::THREADS, tm::SMEM_BYTES, st>>>(qkv, order, windows, n_windows, h, sl2, out, redo); break;
    case 3: attn_window_tm_kernel<3><<<grid, tm::THREADS, tm::SMEM_BYTES, st>>>(qkv, order, windows, n_windows, h, sl2, out, redo); break;
    case 9: attn_window_tm_kernel<9><<<grid, tm::THREADS, tm::SMEM_BYTES, st>>>(qkv, order, windows, n_windows, h, sl2, out, redo); break;
    default: attn_window_tm_kernel<0><<<grid, tm::THREADS, tm::SMEM_BYTES, st>>>(qkv, order, windows, n_windows, h, sl2, out, redo); break;
  }
  return check_launch();
}

}  // namespace ptv3
