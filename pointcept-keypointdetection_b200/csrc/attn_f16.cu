// A4: serialized window attention, second-generation register-resident kernel (head_dim 16), the default bf16 path.
//
// Reference: SerializedAttention.forward, point_transformer_v3m1_base.py:172-222 (flash branch :208-215).
//
// Why this shape (profiles/r02_attention.md): at head_dim 16 a score costs 64 tensor FLOPs but one exponential, and the
// exponential unit (MUFU, 16 / clk / SM; the packed ex2.approx.f16x2 / bf16x2 forms are TWO MUFU ops each in SASS) is
// the roof.  Everything else has to get out of its way:
//   * operands are converted to f16 while they are staged (bf16 -> f16 is exact for |x| in [6e-5, 65504]): Q is
//     pre-multiplied by scale * log2(e) in fp32 and rounded ONCE to f16 (11-bit mantissa, no extra bf16 rounding), so the
//     MMA result is already in log2 units;
//   * the running reference maximum enters through the ACCUMULATOR: S = Q'K^T + (-m) comes out of the tensor core as
//     x = s - m, which removes the per-score FFMA of the classic formulation;
//   * P = 2^x is packed to f16x2 (more mantissa than the bf16 P of the first kernel) and multiplies an f16 V;
//   * lazy rescaling: m moves (quad shuffles, rescale of O and l) only when some score of the warp exceeds it by 2^8;
//   * K / V live in shared memory as 32-byte rows with the 32-byte swizzle (16-byte half h of key k at
//     k*32 + ((h ^ (k>>2)) & 1)*16): every ldmatrix phase touches 32 distinct banks (the un-swizzled layout of the
//     first kernel was 2-way conflicted: 79 % shared-memory pipe in ncu);
//   * every warp owns TWO 16-query tiles and shares each K / V fragment between them (half the ldmatrix traffic, two
//     independent dependency chains for the schedulers);
//   * one CTA per (window, head), no query split: K / V are staged exactly once.
// Output rows are written straight to storage order (the reference's [inverse] scatter, :216-218).
#include <cuda_fp16.h>

#include "common.cuh"

namespace ptv3 {

namespace af {
constexpr int kD = 16;
constexpr int kWarps = 8;
constexpr int kThreads = kWarps * 32;
constexpr float kLazy = 8.0f;   // log2 units: P <= 2^8 before the reference moves

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// 2^x on the FMA pipe (the MUFU unit is the roof of this kernel, the FMA pipe idles at 5 %): round-to-nearest split
// x = n + r through the 1.5 * 2^23 trick, degree-3 minimax of 2^r on [-0.5, 0.5] (max relative error 7.5e-5, below the
// 4.9e-4 rounding of the f16 result), exponent added as an integer.  1 ALU + 7 FMA-pipe instructions.  x <= 2^7.
__device__ __forceinline__ float ex2_poly(float x) {
  x = fmaxf(x, -125.f);
  const float t = __fadd_rn(x, 12582912.f);
  const float r = __fsub_rn(x, __fsub_rn(t, 12582912.f));
  float p = __fmaf_rn(r, 0.0551716685295105f, 0.2426111251115799f);
  p = __fmaf_rn(p, r, 0.6932609677314758f);
  p = __fmaf_rn(p, r, 0.9999280571937561f);
  return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}
// element (b, e) of a tile's score block goes to the polynomial when its slot (bit-reversed, so the two kinds
// interleave evenly) is below POLYN: fraction POLYN / 8 of all exponentials
// POLYN >= 16 are TIMING ABLATIONS (wrong results by design, tools/attn_bench2.py), a bit mask on top of 16:
// 1 = exponentials replaced by one FFMA, 2 = no P V / row-sum MMAs, 4 = no Q K^T MMAs, 8 = no row-maximum tracking
template <int POLYN>
__device__ __forceinline__ float ex2_sel(float x, int b, int e) {
  constexpr int kRev[8] = {0, 4, 2, 6, 1, 5, 3, 7};
  if (POLYN >= 16 && (POLYN & 1)) return fmaf(x, 0.001f, 1.f);
  return (POLYN > 0 && POLYN < 8 && kRev[((b & 1) << 2) | e] < POLYN) ? ex2_poly(x) : ex2(x);
}
__device__ __forceinline__ uint32_t pack_f16(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
// two bf16 (packed) -> two f16 (packed), optionally scaled; clamped to the finite f16 range
__device__ __forceinline__ uint32_t bf16x2_to_f16x2(uint32_t v, float mul) {
  float lo = __uint_as_float(v << 16) * mul, hi = __uint_as_float(v & 0xffff0000u) * mul;
  lo = fminf(fmaxf(lo, -60000.f), 60000.f);
  hi = fminf(fmaxf(hi, -60000.f), 60000.f);
  return pack_f16(lo, hi);
}
// D = A(f16) B(f16) + C, fp32 accumulate; C may be a different register set (the reference maximum)
__device__ __forceinline__ void mma_f16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1, float c0,
                                        float c1, float c2, float c3) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%11,%12,%13};"
      : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1), "f"(c0), "f"(c1), "f"(c2), "f"(c3));
}
__device__ __forceinline__ void mma_f16_acc(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ uint4 ldg_nc16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void sts16(uint32_t addr, uint4 v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ float quad_max(float v) {
  v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 1));
  return fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 2));
}
}  // namespace af

// smem: K [kpad][16] f16 (swizzled), V [kpad][16] f16 (swizzled), order_s [kpad] int32 (storage rows of the keys; the
// queries of a window are a sub-range of its keys, so the Q gather and the output scatter index through it as well).
// NB = 16-key blocks per softmax chunk (2 -> 32 keys, 4 -> 64 keys).
// kLse (training forward): also writes the rows' log-sum-exp (log2 units of the scaled scores) for the backward.
// POLYN: eighths of the exponentials evaluated on the FMA pipe (ex2_poly).
// PP ("ping-pong"): the two warps of a CTA that share a scheduler (warp w and w + 4; 2 CTAs / SM -> two such pairs per
// scheduler) take turns in the exponential phase through a pair of mbarriers.  Without it the round-robin issue makes
// the four warps of a scheduler walk through the MUFU phase TOGETHER (each at a quarter of the speed) and then through
// the MUFU-free phase together, where the exponential unit idles ("convoy": ncu XU 77 %, half of all stall samples
// sitting on MUFU instructions with mio_throttle / wait while the pipe still has 23 % air).  With the hand-over one
// warp of each pair computes exponentials while its partner runs S = Q'K^T and the row maxima of its next chunk.
template <int NB, int QT, bool kLse, int kMinCtas, int POLYN, bool PP>
__global__ void __launch_bounds__(af::kThreads, kMinCtas) attn_window_f16_kernel(
    const __nv_bfloat16* __restrict__ qkv, const int64_t* __restrict__ order, const int64_t* __restrict__ windows,
    const int64_t* __restrict__ n_windows, int h, float scale_log2, int kpad_max, __nv_bfloat16* __restrict__ out,
    float* __restrict__ lse, int* __restrict__ err, const int* __restrict__ only_if) {
  using namespace af;
  constexpr int kChunk = NB * 16;
  if (only_if != nullptr && only_if[blockIdx.x] == 0) return;   // fallback pass behind the tcgen05 kernel
  extern __shared__ __align__(128) unsigned char smem_raw[];
  // heads of one window are neighbours in launch order: their 32-byte K / V / Q slices of one 128-byte line are fetched
  // while the line is still in L2
  const int head = (int)(blockIdx.x % (unsigned)h), w = (int)(blockIdx.x / (unsigned)h);
  if (w >= n_windows[0]) return;
  const int64_t q0 = windows[4 * w + 0], ql = windows[4 * w + 1], k0 = windows[4 * w + 2], kl = windows[4 * w + 3];
  if (ql <= 0 || kl <= 0) return;
  const int kpad = (int)((kl + 63) >> 6) << 6;
  // contract of the window records (ptv3_build_windows): keys fit the launch's shared memory and contain the queries
  if (kpad > kpad_max || q0 < k0 || q0 + ql > k0 + kl) {
    if (threadIdx.x == 0 && err) atomicOr(err, 1);
    return;
  }
  const uint32_t ks_addr = (uint32_t)__cvta_generic_to_shared(smem_raw);
  const uint32_t vs_addr = ks_addr + (uint32_t)kpad * 32u;
  int* order_s = reinterpret_cast<int*>(smem_raw + (size_t)kpad * 64);
  const uint32_t bar0 = (uint32_t)__cvta_generic_to_shared(smem_raw + (size_t)kpad * 68);   // 8 mbarriers (PP)
  const int64_t rs = (int64_t)3 * h * kD;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  // ---- stage K and V of this (window, head): bf16 -> f16, 32-byte rows, 32-byte swizzle ----
  {
    const __nv_bfloat16* kbase = qkv + (int64_t)h * kD + head * kD;
#pragma unroll 1
    for (int key0 = 0; key0 < kpad; key0 += 2 * kThreads) {
      uint4 kv[2][4];
      int rowi[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int key = key0 + u * kThreads + tid;
        rowi[u] = -1;
        if (key < kl) rowi[u] = (int)(order ? order[k0 + key] : k0 + key);
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        if (rowi[u] >= 0) {
          const __nv_bfloat16* p = kbase + (int64_t)rowi[u] * rs;
          kv[u][0] = ldg_nc16(p);
          kv[u][1] = ldg_nc16(p + 8);
          kv[u][2] = ldg_nc16(p + (int64_t)h * kD);
          kv[u][3] = ldg_nc16(p + (int64_t)h * kD + 8);
        } else {
          kv[u][0] = kv[u][1] = kv[u][2] = kv[u][3] = make_uint4(0u, 0u, 0u, 0u);
        }
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int key = key0 + u * kThreads + tid;
        if (key < kpad) {
          if (key < kl) order_s[key] = rowi[u];
          const uint32_t sw = (uint32_t)((key >> 2) & 1) << 4;
#pragma unroll
          for (int part = 0; part < 4; ++part) {
            uint4 v = kv[u][part];
            v.x = bf16x2_to_f16x2(v.x, 1.f);
            v.y = bf16x2_to_f16x2(v.y, 1.f);
            v.z = bf16x2_to_f16x2(v.z, 1.f);
            v.w = bf16x2_to_f16x2(v.w, 1.f);
            sts16((part < 2 ? ks_addr : vs_addr) + (uint32_t)key * 32u + (((uint32_t)(part & 1) << 4) ^ sw), v);
          }
        }
      }
    }
  }
  if (PP && tid < kWarps) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8u * tid));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  const int g = lane >> 2, t = lane & 3;
  const int n_qtiles = (int)((ql + 15) >> 4);
  const int n_pairs = (n_qtiles + QT - 1) / QT;   // groups of QT query tiles, one group per warp and pass
  const int n_chunks = kpad / kChunk;
  // ldmatrix lane addressing: lane supplies the address of row (lane & 7) of matrix (lane >> 3)
  //  K (non-trans): matrices {keys 0-7 dims 0-7, keys 0-7 dims 8-15, keys 8-15 dims 0-7, keys 8-15 dims 8-15}
  //  V (trans):     matrices {keys 0-7 dims 0-7, keys 8-15 dims 0-7, keys 0-7 dims 8-15, keys 8-15 dims 8-15}
  const uint32_t rsw = (uint32_t)((lane >> 2) & 1);   // (key >> 2) & 1 of the lane's row (block bases are multiples of 16)
  const uint32_t k_lane_off = (uint32_t)((((lane >> 4) << 3) + (lane & 7)) * 32) + (((uint32_t)((lane >> 3) & 1) ^ rsw) << 4);
  const uint32_t v_lane_off = (uint32_t)(((((lane >> 3) & 1) << 3) + (lane & 7)) * 32) + (((uint32_t)(lane >> 4) ^ rsw) << 4);
  const uint32_t ones = 0x3c003c00u;  // f16 (1.0, 1.0)
  const int q_rel = (int)(q0 - k0);   // queries are rows q_rel .. q_rel + ql - 1 of the staged key range

  // raw bf16x2 Q words of one 16-query tile: [0] row g k 2t, [1] row g+8 k 2t, [2] row g k 2t+8, [3] row g+8 k 2t+8
  auto load_q = [&](int qt, uint32_t (&qr)[4]) {
    qr[0] = qr[1] = qr[2] = qr[3] = 0u;
    const int qi0 = qt * 16 + g, qi1 = qi0 + 8;
    if (qi0 < ql) {
      const uint32_t* p = reinterpret_cast<const uint32_t*>(qkv + (int64_t)order_s[q_rel + qi0] * rs + head * kD);
      qr[0] = __ldg(p + t);
      qr[2] = __ldg(p + t + 4);
    }
    if (qi1 < ql) {
      const uint32_t* p = reinterpret_cast<const uint32_t*>(qkv + (int64_t)order_s[q_rel + qi1] * rs + head * kD);
      qr[1] = __ldg(p + t);
      qr[3] = __ldg(p + t + 4);
    }
  };

  uint32_t q_next[QT][4];
#pragma unroll
  for (int u = 0; u < QT; ++u) load_q(QT * warp + u, q_next[u]);       // tiles past the last query load zeros
  // PP: both warps of a pair must run the same number of (pass, chunk) hand-overs: every warp walks
  // ceil(n_pairs / kWarps) passes, a warp without queries in the last pass computes on zero queries and stores nothing
  const uint32_t my_bar = bar0 + 8u * warp, partner_bar = bar0 + 8u * (warp ^ 4);
  uint32_t n_phase = 0;
  const int pr_end = PP ? ((n_pairs + kWarps - 1) / kWarps) * kWarps : n_pairs;
  for (int pr = warp; pr < pr_end; pr += kWarps) {
    uint32_t qa[QT][4];
#pragma unroll
    for (int u = 0; u < QT; ++u)
#pragma unroll
      for (int i = 0; i < 4; ++i) qa[u][i] = bf16x2_to_f16x2(q_next[u][i], scale_log2);
    if (pr + kWarps < pr_end) {
#pragma unroll
      for (int u = 0; u < QT; ++u) load_q(QT * (pr + kWarps) + u, q_next[u]);
    }
    float o[QT][2][4], l[QT][4];
    float nm[QT][2];   // minus the reference maximum (log2 units) of rows g / g+8 of both tiles
#pragma unroll
    for (int u = 0; u < QT; ++u) {
      nm[u][0] = nm[u][1] = 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i) o[u][0][i] = o[u][1][i] = l[u][i] = 0.f;
    }

#pragma unroll 1
    for (int c = 0; c < n_chunks; ++c) {
      float s[QT][2 * NB][4];
      const uint32_t kc = ks_addr + (uint32_t)(c * kChunk * 32);
#pragma unroll
      for (int b = 0; b < NB; ++b) {
        uint32_t kb[4];
        ldsm_x4(kb, kc + (uint32_t)(b * 16 * 32) + k_lane_off);
#pragma unroll
        for (int u = 0; u < QT; ++u) {
          if (POLYN >= 16 && (POLYN & 4)) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              s[u][2 * b][i] = nm[u][i >> 1] + __uint_as_float(kb[i] & 0x3fffffffu);
              s[u][2 * b + 1][i] = nm[u][i >> 1] - __uint_as_float(kb[i] & 0x3fffffffu);
            }
            continue;
          }
          mma_f16(s[u][2 * b], qa[u], kb[0], kb[1], nm[u][0], nm[u][0], nm[u][1], nm[u][1]);
          mma_f16(s[u][2 * b + 1], qa[u], kb[2], kb[3], nm[u][0], nm[u][0], nm[u][1], nm[u][1]);
        }
      }
      if (kl < (int64_t)(c + 1) * kChunk) {   // kpad is a multiple of 64: with 32-key chunks the last TWO may be ragged
        const int base = c * kChunk + 2 * t;
#pragma unroll
        for (int b = 0; b < 2 * NB; ++b) {
          const int key = base + b * 8;
#pragma unroll
          for (int u = 0; u < QT; ++u) {
            if (key >= kl) s[u][b][0] = s[u][b][2] = -INFINITY;
            if (key + 1 >= kl) s[u][b][1] = s[u][b][3] = -INFINITY;
          }
        }
      }
      // x = s - m already; the reference moves only when a score of the warp exceeds it by 2^kLazy (or at chunk 0,
      // where it is set to the chunk's row maximum)
      float mx[QT][2];
#pragma unroll
      for (int u = 0; u < QT; ++u) {
        mx[u][0] = fmaxf(s[u][0][0], s[u][0][1]);
        mx[u][1] = fmaxf(s[u][0][2], s[u][0][3]);
#pragma unroll
        for (int b = 1; b < 2 * NB; ++b) {
          mx[u][0] = fmaxf(mx[u][0], fmaxf(s[u][b][0], s[u][b][1]));
          mx[u][1] = fmaxf(mx[u][1], fmaxf(s[u][b][2], s[u][b][3]));
        }
      }
      float mall = fmaxf(mx[0][0], mx[0][1]);
#pragma unroll
      for (int u = 1; u < QT; ++u) mall = fmaxf(mall, fmaxf(mx[u][0], mx[u][1]));
      const bool move = !(POLYN >= 16 && (POLYN & 8)) && mall > kLazy;
      if (c == 0 || __any_sync(0xffffffffu, move)) {
#pragma unroll
        for (int u = 0; u < QT; ++u)
#pragma unroll
          for (int r = 0; r < 2; ++r) {
            float d = quad_max(mx[u][r]);
            if (c > 0) d = fmaxf(d, 0.f);                 // rows that stayed below their reference keep it
            const float a = (c > 0) ? ex2(-d) : 1.f;      // chunk 0: O = l = 0, any reference is valid
            nm[u][r] -= d;
#pragma unroll
            for (int b = 0; b < 2 * NB; ++b) {
              s[u][b][2 * r] -= d;
              s[u][b][2 * r + 1] -= d;
            }
            o[u][0][2 * r] *= a; o[u][0][2 * r + 1] *= a;
            o[u][1][2 * r] *= a; o[u][1][2 * r + 1] *= a;
            l[u][2 * r] *= a; l[u][2 * r + 1] *= a;
          }
      }
      if (PP) {   // my turn in the exponential phase: the first warp of a pair (w < 4) starts, then strict alternation
        const uint32_t parity = (warp < 4 ? n_phase + 1u : n_phase) & 1u;
        uint32_t ok;
        do {
          asm volatile(
              "{\n\t.reg .pred p;\n\t"
              "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
              "selp.u32 %0, 1, 0, p;\n\t}"
              : "=r"(ok)
              : "r"(my_bar), "r"(parity)
              : "memory");
        } while (!ok);
      }
      const uint32_t vc = vs_addr + (uint32_t)(c * kChunk * 32);
#pragma unroll
      for (int j = 0; j < NB; ++j) {  // 16 keys per step
        uint32_t vb[4];
        ldsm_x4_t(vb, vc + (uint32_t)(j * 16 * 32) + v_lane_off);
#pragma unroll
        for (int u = 0; u < QT; ++u) {
          uint32_t pa[4];
          pa[0] = pack_f16(ex2_sel<POLYN>(s[u][2 * j][0], 0, 0), ex2_sel<POLYN>(s[u][2 * j][1], 0, 1));
          pa[1] = pack_f16(ex2_sel<POLYN>(s[u][2 * j][2], 0, 2), ex2_sel<POLYN>(s[u][2 * j][3], 0, 3));
          pa[2] = pack_f16(ex2_sel<POLYN>(s[u][2 * j + 1][0], 1, 0), ex2_sel<POLYN>(s[u][2 * j + 1][1], 1, 1));
          pa[3] = pack_f16(ex2_sel<POLYN>(s[u][2 * j + 1][2], 1, 2), ex2_sel<POLYN>(s[u][2 * j + 1][3], 1, 3));
          if (POLYN >= 16 && (POLYN & 2)) {
            o[u][0][0] += __uint_as_float((pa[0] ^ vb[0]) & 0x3fffffffu);
            o[u][0][2] += __uint_as_float((pa[1] ^ vb[1]) & 0x3fffffffu);
            l[u][0] += __uint_as_float((pa[2] ^ vb[2]) & 0x3fffffffu);
            l[u][2] += __uint_as_float((pa[3] ^ vb[3]) & 0x3fffffffu);
            continue;
          }
          mma_f16_acc(o[u][0], pa, vb[0], vb[1]);
          mma_f16_acc(o[u][1], pa, vb[2], vb[3]);
          mma_f16_acc(l[u], pa, ones, ones);
        }
      }
      if (PP) {   // all exponentials of this chunk are issued: the partner may start its own
        __syncwarp();
        if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(partner_bar) : "memory");
        ++n_phase;
      }
    }
    // ---- epilogue: normalise, write bf16 pairs to storage rows ----
#pragma unroll
    for (int u = 0; u < QT; ++u)
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int qi = (QT * pr + u) * 16 + g + 8 * r;
        if (qi < ql) {
          const float inv = 1.f / l[u][2 * r];
          const int64_t row = order_s[q_rel + qi];
          uint32_t* p = reinterpret_cast<uint32_t*>(out + row * h * kD + head * kD);
          p[t] = pack_bf16(o[u][0][2 * r] * inv, o[u][0][2 * r + 1] * inv);
          p[t + 4] = pack_bf16(o[u][1][2 * r] * inv, o[u][1][2 * r + 1] * inv);
          if (kLse && t == 0) lse[row * h + head] = __log2f(l[u][2 * r]) - nm[u][r];
        }
      }
  }
}

int g_attn_lse_algo = 0;      // debug slot 8: 2 = training forward on the first-generation kernel
// debug slot 7 (ablation): 0 = auto; otherwise 100 * pingpong(1 = off, 2 = on) + 10 * (1 + eighths of the exponentials
// on the FMA-pipe polynomial, 0..3) + chunk variant (1 = 64-key chunks / 2 CTAs per SM, 2 = 32-key chunks / 3 CTAs)
int g_attn_f16_variant = 0;
constexpr int kDefaultVariant = 211;

template <int NB, int QT, bool kLse, int kMinCtas, int POLYN, bool PP>
static int launch_f16(const __nv_bfloat16* qkv, const int64_t* order, const int64_t* windows, const int64_t* n_windows,
                      int64_t n_windows_max, int kpad, size_t smem, int h, float scale, __nv_bfloat16* out, float* lse,
                      int* err, cudaStream_t st, const int* only_if) {
  static size_t configured[16] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  dev &= 15;
  if (smem > configured[dev]) {
    if (cudaFuncSetAttribute(attn_window_f16_kernel<NB, QT, kLse, kMinCtas, POLYN, PP>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
      return PTV3_ELAUNCH;
    configured[dev] = smem;
  }
  attn_window_f16_kernel<NB, QT, kLse, kMinCtas, POLYN, PP><<<(unsigned)(n_windows_max * h), af::kThreads, smem, st>>>(
      qkv, order, windows, n_windows, h, scale * 1.4426950408889634f, kpad, out, lse, err, only_if);
  return check_launch();
}

int attn_window_f16(const __nv_bfloat16* qkv, const int64_t* order, const int64_t* windows, const int64_t* n_windows,
                    int64_t n_windows_max, int64_t max_len, int h, float scale, __nv_bfloat16* out, cudaStream_t st,
                    float* lse, int* err, const int* only_if) {
  const int kpad = (int)(cdiv(max_len, 64) * 64);
  const size_t smem = (size_t)kpad * 68 + 64;
  if (smem > 200 * 1024) return PTV3_EUNSUPPORTED;  // patch > 3008
  if ((int64_t)n_windows_max * h > 2147483647LL) return PTV3_EINVAL;
  const int v = g_attn_f16_variant > 0 ? g_attn_f16_variant : kDefaultVariant;
  const int chunkv = v % 10, poly = (v / 10) % 10 - 1;
  const bool pp = v / 100 == 2;
#define PTV3_F16_ARGS qkv, order, windows, n_windows, n_windows_max, kpad, smem, h, scale, out, lse, err, st, only_if
  if (lse) return launch_f16<4, 2, true, 2, 0, true>(PTV3_F16_ARGS);
  if (v >= 1000) {   // timing ablations (results are wrong by design)
    switch (v - 1000) {   // bit mask: 1 no exp, 2 no PV MMAs, 4 no QK MMAs, 8 no max tracking
      case 1: return launch_f16<4, 2, false, 2, 17, false>(PTV3_F16_ARGS);
      case 2: return launch_f16<4, 2, false, 2, 18, false>(PTV3_F16_ARGS);
      case 3: return launch_f16<4, 2, false, 2, 19, false>(PTV3_F16_ARGS);
      case 6: return launch_f16<4, 2, false, 2, 22, false>(PTV3_F16_ARGS);
      case 7: return launch_f16<4, 2, false, 2, 23, false>(PTV3_F16_ARGS);
      case 14: return launch_f16<4, 2, false, 2, 30, false>(PTV3_F16_ARGS);
      default: return launch_f16<4, 2, false, 2, 31, false>(PTV3_F16_ARGS);
    }
  }
#define PTV3_F16_POLY(NB, QT, CT, PPV)                                       \
  switch (poly) {                                                            \
    case 1: return launch_f16<NB, QT, false, CT, 1, PPV>(PTV3_F16_ARGS);      \
    case 2: return launch_f16<NB, QT, false, CT, 2, PPV>(PTV3_F16_ARGS);      \
    case 3: return launch_f16<NB, QT, false, CT, 3, PPV>(PTV3_F16_ARGS);      \
    default: return launch_f16<NB, QT, false, CT, 0, PPV>(PTV3_F16_ARGS);     \
  }
  if (chunkv == 3) { PTV3_F16_POLY(4, 1, 3, false) }     // one query tile per warp, 64-key chunks, 3 CTAs / SM
  if (chunkv == 4) { PTV3_F16_POLY(4, 1, 4, false) }     // ... 4 CTAs / SM (64 registers)
  if (chunkv == 2) {
    if (pp) { PTV3_F16_POLY(2, 2, 3, true) }
    PTV3_F16_POLY(2, 2, 3, false)
  }
  if (pp) { PTV3_F16_POLY(4, 2, 2, true) }
  PTV3_F16_POLY(4, 2, 2, false)
#undef PTV3_F16_POLY
#undef PTV3_F16_ARGS
}

}  // namespace ptv3
