// A8: the Block's MLP  fc1 -> GELU(erf) -> fc2 (+ residual add) as ONE persistent tcgen05 kernel (bf16, C = 32 / 64).
// Reference: MLP.forward point_transformer_v3m1_base.py:225-248 and the residual wiring of Block.forward :330-337.
//
// The reference (and a cuBLAS chain) writes the (N, 4C) hidden activation to HBM, reads it for GELU, writes it again and
// reads it for fc2: 4 x N x 4C x 2 B.  Here the hidden tile never leaves the SM:
//   GEMM1  D1[128 x 128] = X[128 x C] . W1[hb]^T          tcgen05.mma, accumulator in TMEM (2 buffers)
//   epi 1  tcgen05.ld D1 -> + b1 -> GELU(erf) -> bf16 -> shared memory in the 128-byte-swizzled A-operand layout
//   GEMM2  D2[128 x C] += H[128 x 128] . W2[:, hb]^T      accumulator in TMEM (2 buffers, one per row tile in flight)
//   epi 2  tcgen05.ld D2 -> + b2 (+ fp32 residual) -> fp32 residual stream and its bf16 copy
// Both weight matrices (pre-packed into shared-memory images by pack_kmajor_kernel) stay resident in shared memory
// for the CTA's lifetime (one bulk copy each); X tiles are double-buffered by two loader warps with 16-byte cp.async.
// HBM traffic per row: read C x 2 B (+ C x 4 B residual), write C x (4 + 2) B  -- the hidden activation costs nothing.
#include <cstdlib>
#include "common.cuh"
#include "tc05.cuh"

namespace ptv3 {

using namespace tc05;

namespace mlp {
constexpr int TM = 128;
constexpr int HB = 128;            // hidden columns per GEMM1 / K extent per GEMM2 step
constexpr int THREADS = 640;       // warps 0-1 loaders, 2 MMA issuer (+TMEM alloc), 3 spare, 4-19 epilogue
constexpr int LOADERS = 64;
constexpr int EPI = 512;
constexpr int OUTW = 128;         // mlp_fused_kernel: four more warps (20-23) for the output epilogue

template <int C>
struct Smem {
  static constexpr int HID = 4 * C;
  static constexpr int NHB = HID / HB;
  static constexpr int W1_BYTES = NHB * HB * 128;            // [hb][128 rows][64 (C padded)] swizzled
  static constexpr int W2_TILE = C * 128;                    // one K-block of W2: [C rows][64]
  static constexpr int W2_BYTES = (HID / 64) * W2_TILE;
  static constexpr int A1_BYTES = TM * 128;                  // [128 rows][64]
  static constexpr int A2_BYTES = 2 * TM * 128;              // two K-blocks
  static constexpr int W1_OFF = 0;
  static constexpr int W2_OFF = W1_OFF + W1_BYTES;
  static constexpr int A1_OFF = W2_OFF + W2_BYTES;
  static constexpr int A2_OFF = A1_OFF + 2 * A1_BYTES;
  static constexpr int BIAS_OFF = A2_OFF + 2 * A2_BYTES;      // b1 (HID floats) then b2 (C floats)
  static constexpr int BAR_OFF = BIAS_OFF + (HID + C) * 4;
  static constexpr int DYN_BYTES = BAR_OFF + 256 + 1024;
};

using ptv3::gelu_erf;   // common.cuh
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
}  // namespace mlp

// generic K-major packer: w (rows, k) bf16 row-major -> [tile][kb][tile_rows][64] with the 128-byte swizzle (zero
// padded along K up to a multiple of 64)
__global__ void __launch_bounds__(256) pack_kmajor_kernel(const __nv_bfloat16* __restrict__ w, int rows, int k,
                                                          int tile_rows, int n_kb, __nv_bfloat16* __restrict__ packed) {
  const int64_t chunks = (int64_t)n_kb * rows * 8;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < chunks; t += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(t & 7);
    const int64_t rowid = t >> 3;
    const int rl = (int)(rowid % tile_rows);
    const int64_t r2 = rowid / tile_rows;
    const int kb = (int)(r2 % n_kb);
    const int tile = (int)(r2 / n_kb);
    const int row = tile * tile_rows + rl;
    const int kk = kb * 64 + c * 8;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (kk < k) v = *reinterpret_cast<const uint4*>(w + (int64_t)row * k + kk);
    char* base = reinterpret_cast<char*>(packed) + ((int64_t)tile * n_kb + kb) * tile_rows * 128;
    const int off = (rl >> 3) * 1024 + (rl & 7) * 128 + ((c ^ (rl & 7)) << 4);
    *reinterpret_cast<uint4*>(base + off) = v;
  }
}

template <int C>
__global__ void __launch_bounds__(mlp::THREADS + mlp::OUTW, 1) mlp_fused_kernel(
    const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ w1p, const float* __restrict__ b1,
    const __nv_bfloat16* __restrict__ w2p, const float* __restrict__ b2, const float* __restrict__ residual,
    const float* __restrict__ row_scale, int64_t n,
    int n_tiles, float* __restrict__ res_out, __nv_bfloat16* __restrict__ out_bf16, int dbg) {
  using namespace mlp;
  using L = Smem<C>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t sb = smem_u32(smem);
  auto bar = [&](int id) { return sb + L::BAR_OFF + 8u * (uint32_t)id; };
  // 0 w_full | 1,2 a1_full | 3,4 a1_free | 5,6 d1_full | 7,8 d1_free | 9,10 a2_full | 11,12 a2_free | 13,14 d2_full
  // | 15,16 d2_free
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + L::BAR_OFF + 200);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // provably warp-uniform for nvcc (role branches)
  const int my_tiles = (n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (tid == 0) {
    mbar_init(bar(0), 1);
    for (int b = 0; b < 2; ++b) {
      mbar_init(bar(1 + b), LOADERS); mbar_init(bar(3 + b), 1);
      mbar_init(bar(5 + b), 1);       mbar_init(bar(7 + b), EPI);
      mbar_init(bar(9 + b), EPI);     mbar_init(bar(11 + b), 1);
      mbar_init(bar(13 + b), 1);      mbar_init(bar(15 + b), OUTW);
    }
    mbar_fence_init();
  }
  for (int j = tid; j < L::HID + C; j += THREADS + OUTW)
    reinterpret_cast<float*>(smem + L::BIAS_OFF)[j] = j < L::HID ? b1[j] : b2[j - L::HID];
  if (warp == 2) tmem_alloc<512>(smem_u32(tmem_slot));
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = *tmem_slot;

  if (warp < 2) {
    // =========================== loaders ===========================
    if (tid == 0) {
      mbar_arrive_expect_tx(bar(0), L::W1_BYTES + L::W2_BYTES);
      bulk_g2s(sb + L::W1_OFF, w1p, L::W1_BYTES, bar(0));
      bulk_g2s(sb + L::W2_OFF, w2p, L::W2_BYTES, bar(0));
    }
    if (C == 32) {   // K is padded to 64: the upper four 16-byte chunks of every row stay zero for the kernel's life
      for (int idx = tid; idx < 2 * TM * 4; idx += LOADERS) {
        const int buf = idx / (TM * 4), r = (idx / 4) % TM, c = 4 + (idx & 3);
        st_shared_zero16(sb + L::A1_OFF + buf * L::A1_BYTES + (r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4));
      }
    }
    const int c = tid & 7, r0 = tid >> 3;
    for (int i = 0; i < my_tiles; ++i) {
      const int ab = i & 1;
      const int64_t row0 = (int64_t)(blockIdx.x + (int64_t)i * gridDim.x) * TM;
      mbar_wait(bar(3 + ab), ((uint32_t)(i >> 1) & 1u) ^ 1u);
      if (c < C / 8) {
        const uint32_t a1 = sb + L::A1_OFF + ab * L::A1_BYTES;
#pragma unroll
        for (int q = 0; q < 16; ++q) {
          const int r = r0 + 8 * q;
          const uint32_t dst = a1 + (r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4);
          if (row0 + r < n) cp_async16(dst, x + (row0 + r) * C + c * 8, 16u);
          else st_shared_zero16(dst);
        }
      }
      cp_async_commit();
      cp_async_wait<0>();
      fence_proxy_async();
      mbar_arrive(bar(1 + ab));
    }
  } else if (warp == 2) {
    // =========================== MMA issuer ===========================
    // One elected lane runs the issue schedule.  (Issuing from inside `if (lane == 0)` instead makes nvcc wrap every
    // tcgen05 instruction in an ELECT / R2UR / BRA.U.ANY loop.)
    // Flat schedule over hidden-block steps t = tile * NHB + hb with GEMM1 two steps ahead of GEMM2:
    //   fc1(0) fc1(1) | fc1(t+2) fc2(t) | ...   so the epilogue of step t+1 never waits for a GEMM1 that was queued
    // behind the GEMM2 of step t.
    {
      constexpr uint32_t idesc1 = idesc_bf16_f32(TM, HB), idesc2 = idesc_bf16_f32(TM, C);
      constexpr int NHB = L::NHB;
      mbar_wait(bar(0), 0);
      const int T = my_tiles * NHB;
      auto fc1 = [&](int t) {
        const int it = t / NHB, hb = t - it * NHB;
        const uint32_t ab = (uint32_t)it & 1u, bb = (uint32_t)t & 1u;
        if (hb == 0) mbar_wait(bar(1 + ab), (uint32_t)(it >> 1) & 1u);   // X tile landed
        mbar_wait(bar(7 + bb), (((uint32_t)t >> 1) & 1u) ^ 1u);           // D1[bb] drained by the epilogue of step t-2
        fence_after_sync();
        {
          const uint64_t a_desc = smem_desc_sw128(sb + L::A1_OFF + ab * L::A1_BYTES);
          const uint64_t b_desc = smem_desc_sw128(sb + L::W1_OFF + hb * HB * 128);
#pragma unroll
          for (int kk = 0; kk < C / 16 && kk < 4; ++kk)
            mma_bf16_ss(tmem + bb * HB, a_desc + 2 * kk, b_desc + 2 * kk, idesc1, kk ? 1u : 0u);
          mma_commit(bar(5 + bb));
          if (hb == NHB - 1) mma_commit(bar(3 + ab));   // X tile fully consumed
        }
      };
      auto fc2 = [&](int t) {
        const int it = t / NHB, h2 = t - it * NHB;
        const uint32_t ab = (uint32_t)it & 1u, bb = (uint32_t)t & 1u;
        mbar_wait(bar(9 + bb), ((uint32_t)t >> 1) & 1u);                  // GELU(h) block written by the epilogue
        if (h2 == 0) mbar_wait(bar(15 + ab), ((uint32_t)(it >> 1) & 1u) ^ 1u);   // D2[ab] drained (tile it-2)
        fence_after_sync();
        {
#pragma unroll
          for (int kb = 0; kb < 2; ++kb) {
            const uint64_t a_desc = smem_desc_sw128(sb + L::A2_OFF + bb * L::A2_BYTES + kb * 16384);
            const uint64_t b_desc = smem_desc_sw128(sb + L::W2_OFF + (2 * h2 + kb) * L::W2_TILE);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)
              mma_bf16_ss(tmem + 256 + ab * C, a_desc + 2 * kk, b_desc + 2 * kk, idesc2, (h2 | kb | kk) ? 1u : 0u);
          }
          mma_commit(bar(11 + bb));
          if (h2 == NHB - 1) mma_commit(bar(13 + ab));
        }
      };
      if (elect_one_sync()) {   // one thread runs the whole issue schedule (state and descriptors in uniform registers)
        if (T > 0) fc1(0);
        if (T > 1) fc1(1);
        for (int t = 0; t < T; ++t) {
          if (t + 2 < T) fc1(t + 2);
          fc2(t);
        }
      }
      __syncwarp();
    }
  } else if (warp >= 4) {
    // =========================== epilogue warps ===========================
    const int q4 = warp & 3, quarter = (warp - 4) >> 2;   // TMEM lane quarter, column quarter of the 128-wide block
    const int row = q4 * 32 + lane;
    const uint32_t lane_base = (uint32_t)(q4 * 32) << 16;
    auto gelu_step = [&](int i, int hb, uint32_t e1) {   // D1 -> + b1 -> GELU -> bf16 A operand of GEMM2
        const uint32_t bb = e1 & 1u;
      mbar_wait(bar(5 + bb), (e1 >> 1) & 1u);
      fence_after_sync();
      uint32_t v[32], pk[16];
      const int col0 = quarter * 32;
      tmem_ld32(tmem + lane_base + bb * HB + col0, v);
      tmem_ld_wait();
      fence_before_sync();
      mbar_arrive(bar(7 + bb));
      const float4* bp = reinterpret_cast<const float4*>(smem + L::BIAS_OFF) + (hb * HB + col0) / 4;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 bv = bp[j];
        pk[2 * j] = pack2(gelu_erf(__uint_as_float(v[4 * j]) + bv.x), gelu_erf(__uint_as_float(v[4 * j + 1]) + bv.y));
        pk[2 * j + 1] =
            pack2(gelu_erf(__uint_as_float(v[4 * j + 2]) + bv.z), gelu_erf(__uint_as_float(v[4 * j + 3]) + bv.w));
      }
      mbar_wait(bar(11 + bb), ((e1 >> 1) & 1u) ^ 1u);
      // my 32 hidden columns = four 16-byte chunks of K-block (quarter >> 1), chunks (quarter & 1) * 4 ..
      const uint32_t prow = sb + L::A2_OFF + bb * L::A2_BYTES + (quarter >> 1) * 16384 + (row >> 3) * 1024 +
                            (row & 7) * 128;
#pragma unroll
      for (int c4 = 0; c4 < 4; ++c4)
        asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(
                         prow + ((((quarter & 1) * 4 + c4) ^ (row & 7)) << 4)),
                     "r"(pk[4 * c4]), "r"(pk[4 * c4 + 1]), "r"(pk[4 * c4 + 2]), "r"(pk[4 * c4 + 3])
                     : "memory");
      fence_proxy_async();
      mbar_arrive(bar(9 + bb));
    };
    // ---- output epilogue (4 separate warps, 20-23): D2 -> + b2 (+ residual) -> fp32 stream and bf16 copy.  On the GELU warps
    // its HBM round trip stalled the GELU steps of the next tile (timing ablation: 0.301 -> 0.218 ms without it at C = 64).
    // This thread owns row `row` and all C columns, in pieces of 32.
    auto epilogue2 = [&](int i) {
      constexpr int PW = C < 32 ? C : 32, PIECES = C / PW;
      const uint32_t ab = (uint32_t)i & 1u, tph = (uint32_t)(i >> 1) & 1u;
      const int64_t grow = (int64_t)(blockIdx.x + (int64_t)i * gridDim.x) * TM + row;
      float rres[4][8];      // one 32-column piece of the residual row (256-bit accesses); the first piece is requested
#pragma unroll            // BEFORE waiting for GEMM2, the next one while the current piece is being stored
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int q = 0; q < 8; ++q) rres[j][q] = 0.f;
      float rs = 1.f;      // DropPath on the branch: 0 or 1 / keep for the whole row (NULL = none)
      if (grow < n && !(dbg & 8)) {
        if (residual) {
#pragma unroll
          for (int j = 0; j < 4; ++j) ldg256_cs(residual + grow * C + 8 * j, rres[j]);
        }
        if (row_scale) rs = row_scale[grow];
      }
      mbar_wait(bar(13 + ab), tph);
      fence_after_sync();
      if (dbg & 8) {      // timing ablation: no output epilogue
        fence_before_sync();
        mbar_arrive(bar(15 + ab));
        return;
      }
#pragma unroll 1
      for (int pc = 0; pc < PIECES; ++pc) {
        const int col0 = 32 * pc;
        uint32_t v[32];
        tmem_ld32(tmem + lane_base + 256 + ab * C + col0, v);
        tmem_ld_wait();
        if (pc == PIECES - 1) {
          fence_before_sync();
          mbar_arrive(bar(15 + ab));
        }
        if (grow < n) {
          const float* b2s = reinterpret_cast<const float*>(smem + L::BIAS_OFF) + L::HID + col0;
          uint32_t pk[16];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float f[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) f[q] = fmaf(__uint_as_float(v[8 * j + q]) + b2s[8 * j + q], rs, rres[j][q]);
            if (res_out) stg256(res_out + grow * C + col0 + 8 * j, f);
#pragma unroll
            for (int q = 0; q < 4; ++q) pk[4 * j + q] = pack2(f[2 * q], f[2 * q + 1]);
          }
          if (residual && pc + 1 < PIECES) {
#pragma unroll
            for (int j = 0; j < 4; ++j) ldg256_cs(residual + grow * C + col0 + 32 + 8 * j, rres[j]);
          }
          if (out_bf16) {
            uint32_t lo[8], hi[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) { lo[q] = pk[q]; hi[q] = pk[8 + q]; }
            stg256_u32(out_bf16 + grow * C + col0, lo);
            stg256_u32(out_bf16 + grow * C + col0 + 16, hi);
          }
        }
      }
    };
    if (warp < 20) {
      uint32_t e1 = 0;
      for (int i = 0; i < my_tiles; ++i)
        for (int hb = 0; hb < L::NHB; ++hb) gelu_step(i, hb, e1++);
    } else {
      for (int i = 0; i < my_tiles; ++i) epilogue2(i);
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 2) tmem_dealloc<512>(tmem);
}

// ---------------------------------------------------------------------------------------------------------------
// proj + LayerNorm + MLP (round 2): the second half of Block.forward (:329-337) as ONE kernel for C = 32 / 64
//   x2 = x1 + rs_a * (a Wp^T + bp)            attention projection + DropPath + shortcut
//   h  = LayerNorm(x2; g2, be2)                pre-norm of the MLP
//   out = x2 + rs_m * (GELU(h W1^T + b1) W2^T + b2)
// = mlp_fused_kernel with one more GEMM in front: the loaders bring the attention output tile `a` (A0), the issuer runs
// GEMM0 (D0 = A0 . Wp^T, two tiles ahead), and the epilogue warps turn D0 into the A operand of GEMM1 themselves:
// + bp, x rs_a, + x1 (read from HBM), the row statistics of the 4 column quarters exchanged through shared memory,
// normalised, bf16 into the swizzled A1 stage.  x2 is parked in `res_out` (same thread writes it, reads it back in
// epilogue 2 and overwrites it with the final value), so it costs no registers across the tile.
// Against the separate launches (cuBLASLt proj, ln_residual, mlp_fused) this removes the write + read of proj's
// output, of x2 and of h: 28 C -> 16 C bytes per row (12 C from HBM: the x2 round trip stays in L2), and two launches.
// ---------------------------------------------------------------------------------------------------------------
namespace mlp {
template <int C>
struct PSmem {
  static constexpr int HID = 4 * C;
  static constexpr int NHB = HID / HB;
  static constexpr int W1_BYTES = NHB * HB * 128;
  static constexpr int W2_TILE = C * 128;
  static constexpr int W2_BYTES = (HID / 64) * W2_TILE;
  static constexpr int WP_BYTES = C * 128;                   // [C rows][64] one K-block
  static constexpr int A1_BYTES = TM * 128;
  static constexpr int A2_BYTES = 2 * TM * 128;
  static constexpr int W1_OFF = 0;
  static constexpr int W2_OFF = W1_OFF + W1_BYTES;
  static constexpr int WP_OFF = W2_OFF + W2_BYTES;
  static constexpr int A0_OFF = WP_OFF + WP_BYTES;
  static constexpr int A1_OFF = A0_OFF + 2 * A1_BYTES;
  static constexpr int A2_OFF = A1_OFF + 2 * A1_BYTES;
  static constexpr int BIAS_OFF = A2_OFF + 2 * A2_BYTES;      // b1 (HID) | b2 (C) | bp (C) | g2 (C) | be2 (C)
  static constexpr int STAT_OFF = BIAS_OFF + (HID + 4 * C) * 4;   // [2 parities][128 rows][4 quarters] float2
  static constexpr int BAR_OFF = STAT_OFF + 2 * TM * 4 * 8;
  static constexpr int DYN_BYTES = BAR_OFF + 256 + 1024;
  static constexpr int D2_COL = 256, D0_COL = 256 + 2 * C;    // TMEM columns: D1 2 x 128 | D2 2 x C | D0 2 x C
};
__device__ __forceinline__ void epi_sync() { asm volatile("bar.sync 2, %0;" ::"n"(EPI) : "memory"); }
}  // namespace mlp

template <int C>
__global__ void __launch_bounds__(mlp::THREADS, 1) proj_mlp_fused_kernel(
    const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ wpp, const float* __restrict__ bp,
    const float* __restrict__ x1, const float* __restrict__ rs_attn, const float* __restrict__ g2,
    const float* __restrict__ be2, float eps, const __nv_bfloat16* __restrict__ w1p, const float* __restrict__ b1,
    const __nv_bfloat16* __restrict__ w2p, const float* __restrict__ b2, const float* __restrict__ rs_mlp, int64_t n,
    int n_tiles, float* __restrict__ res_out, __nv_bfloat16* __restrict__ out_bf16) {
  using namespace mlp;
  using L = PSmem<C>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t sb = smem_u32(smem);
  auto bar = [&](int id) { return sb + L::BAR_OFF + 8u * (uint32_t)id; };
  // 0 w_full | 1,2 a1_full | 3,4 a1_free | 5,6 d1_full | 7,8 d1_free | 9,10 a2_full | 11,12 a2_free | 13,14 d2_full
  // | 15,16 d2_free | 17,18 a0_full | 19,20 a0_free | 21,22 d0_full | 23,24 d0_free
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + L::BAR_OFF + 240);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int my_tiles = (n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (tid == 0) {
    mbar_init(bar(0), 1);
    for (int b = 0; b < 2; ++b) {
      mbar_init(bar(1 + b), EPI);     mbar_init(bar(3 + b), 1);
      mbar_init(bar(5 + b), 1);       mbar_init(bar(7 + b), EPI);
      mbar_init(bar(9 + b), EPI);     mbar_init(bar(11 + b), 1);
      mbar_init(bar(13 + b), 1);      mbar_init(bar(15 + b), EPI);
      mbar_init(bar(17 + b), LOADERS); mbar_init(bar(19 + b), 1);
      mbar_init(bar(21 + b), 1);      mbar_init(bar(23 + b), EPI);
    }
    mbar_fence_init();
  }
  {
    float* bs = reinterpret_cast<float*>(smem + L::BIAS_OFF);
    for (int j = tid; j < L::HID + 4 * C; j += THREADS) {
      const int k = j - L::HID;
      bs[j] = j < L::HID ? b1[j] : k < C ? b2[k] : k < 2 * C ? bp[k - C] : k < 3 * C ? g2[k - 2 * C] : be2[k - 3 * C];
    }
  }
  if (warp == 2) tmem_alloc<512>(smem_u32(tmem_slot));
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = *tmem_slot;

  if (warp < 2) {
    // =========================== loaders: attention output tiles -> A0 ===========================
    if (tid == 0) {
      mbar_arrive_expect_tx(bar(0), L::W1_BYTES + L::W2_BYTES + L::WP_BYTES);
      bulk_g2s(sb + L::W1_OFF, w1p, L::W1_BYTES, bar(0));
      bulk_g2s(sb + L::W2_OFF, w2p, L::W2_BYTES, bar(0));
      bulk_g2s(sb + L::WP_OFF, wpp, L::WP_BYTES, bar(0));
    }
    if (C == 32) {   // K padded to 64: the upper four 16-byte chunks of every A0 / A1 row stay zero
      for (int idx = tid; idx < 4 * TM * 4; idx += LOADERS) {
        const int buf = idx / (TM * 4), r = (idx / 4) % TM, c = 4 + (idx & 3);
        st_shared_zero16(sb + L::A0_OFF + buf * L::A1_BYTES + (r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4));
      }
    }
    const int c = tid & 7, r0 = tid >> 3;
    for (int i = 0; i < my_tiles; ++i) {
      const int ab = i & 1;
      const int64_t row0 = (int64_t)(blockIdx.x + (int64_t)i * gridDim.x) * TM;
      mbar_wait(bar(19 + ab), ((uint32_t)(i >> 1) & 1u) ^ 1u);
      if (c < C / 8) {
        const uint32_t a0 = sb + L::A0_OFF + ab * L::A1_BYTES;
#pragma unroll
        for (int q = 0; q < 16; ++q) {
          const int r = r0 + 8 * q;
          const uint32_t dst = a0 + (r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4);
          if (row0 + r < n) cp_async16(dst, a + (row0 + r) * C + c * 8, 16u);
          else st_shared_zero16(dst);
        }
      }
      cp_async_commit();
      cp_async_wait<0>();
      fence_proxy_async();
      mbar_arrive(bar(17 + ab));
    }
  } else if (warp == 2) {
    // =========================== MMA issuer ===========================
    // mlp_fused_kernel's flat schedule (GEMM1 two hidden-block steps ahead of GEMM2) with GEMM0 of tile it + 2 issued
    // right after the first GEMM1 of tile it: by then the epilogue warps have long consumed D0 of tile it.
    {
      constexpr uint32_t idesc0 = idesc_bf16_f32(TM, C), idesc1 = idesc_bf16_f32(TM, HB), idesc2 = idesc_bf16_f32(TM, C);
      constexpr int NHB = L::NHB;
      mbar_wait(bar(0), 0);
      const int T = my_tiles * NHB;
      auto proj = [&](int it) {
        const uint32_t ab = (uint32_t)it & 1u, ph = (uint32_t)(it >> 1) & 1u;
        mbar_wait(bar(17 + ab), ph);            // attention-output tile landed
        mbar_wait(bar(23 + ab), ph ^ 1u);       // D0[ab] drained by E0 of tile it - 2
        fence_after_sync();
        const uint64_t a_desc = smem_desc_sw128(sb + L::A0_OFF + ab * L::A1_BYTES);
        const uint64_t b_desc = smem_desc_sw128(sb + L::WP_OFF);
#pragma unroll
        for (int kk = 0; kk < C / 16; ++kk)
          mma_bf16_ss(tmem + L::D0_COL + ab * C, a_desc + 2 * kk, b_desc + 2 * kk, idesc0, kk ? 1u : 0u);
        mma_commit(bar(21 + ab));
        mma_commit(bar(19 + ab));               // A0[ab] may be refilled
      };
      auto fc1 = [&](int t) {
        const int it = t / NHB, hb = t - it * NHB;
        const uint32_t ab = (uint32_t)it & 1u, bb = (uint32_t)t & 1u;
        if (hb == 0) mbar_wait(bar(1 + ab), (uint32_t)(it >> 1) & 1u);   // LayerNorm'd tile written by the epilogue warps
        mbar_wait(bar(7 + bb), (((uint32_t)t >> 1) & 1u) ^ 1u);
        fence_after_sync();
        const uint64_t a_desc = smem_desc_sw128(sb + L::A1_OFF + ab * L::A1_BYTES);
        const uint64_t b_desc = smem_desc_sw128(sb + L::W1_OFF + hb * HB * 128);
#pragma unroll
        for (int kk = 0; kk < C / 16 && kk < 4; ++kk)
          mma_bf16_ss(tmem + bb * HB, a_desc + 2 * kk, b_desc + 2 * kk, idesc1, kk ? 1u : 0u);
        mma_commit(bar(5 + bb));
        if (hb == NHB - 1) mma_commit(bar(3 + ab));
        if (hb == 0 && it + 2 < my_tiles) proj(it + 2);
      };
      auto fc2 = [&](int t) {
        const int it = t / NHB, h2 = t - it * NHB;
        const uint32_t ab = (uint32_t)it & 1u, bb = (uint32_t)t & 1u;
        mbar_wait(bar(9 + bb), ((uint32_t)t >> 1) & 1u);
        if (h2 == 0) mbar_wait(bar(15 + ab), ((uint32_t)(it >> 1) & 1u) ^ 1u);
        fence_after_sync();
#pragma unroll
        for (int kb = 0; kb < 2; ++kb) {
          const uint64_t a_desc = smem_desc_sw128(sb + L::A2_OFF + bb * L::A2_BYTES + kb * 16384);
          const uint64_t b_desc = smem_desc_sw128(sb + L::W2_OFF + (2 * h2 + kb) * L::W2_TILE);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            mma_bf16_ss(tmem + L::D2_COL + ab * C, a_desc + 2 * kk, b_desc + 2 * kk, idesc2, (h2 | kb | kk) ? 1u : 0u);
        }
        mma_commit(bar(11 + bb));
        if (h2 == NHB - 1) mma_commit(bar(13 + ab));
      };
      if (elect_one_sync()) {
        if (my_tiles > 0) proj(0);
        if (my_tiles > 1) proj(1);
        if (T > 0) fc1(0);
        if (T > 1) fc1(1);
        for (int t = 0; t < T; ++t) {
          if (t + 2 < T) fc1(t + 2);
          fc2(t);
        }
      }
      __syncwarp();
    }
  } else if (warp >= 4) {
    // =========================== epilogue warps ===========================
    const int q4 = warp & 3, quarter = (warp - 4) >> 2;   // TMEM lane quarter, column quarter
    const int row = q4 * 32 + lane;
    const uint32_t lane_base = (uint32_t)(q4 * 32) << 16;
    constexpr int NC = C / 4;                               // columns of a row this thread owns in E0 / E2
    const int col0 = quarter * NC;
    const float* bs = reinterpret_cast<const float*>(smem + L::BIAS_OFF);
    float2* stats = reinterpret_cast<float2*>(smem + L::STAT_OFF);

    // ---- E0: D0 -> + bp, x rs_a, + x1 -> x2 (parked in res_out) -> LayerNorm -> bf16 A operand of GEMM1 ----
    auto e0_step = [&](int i) {
      const uint32_t ab = (uint32_t)i & 1u, tph = (uint32_t)(i >> 1) & 1u;
      const int64_t grow = (int64_t)(blockIdx.x + (int64_t)i * gridDim.x) * TM + row;
      float xr[NC / 8][8];
#pragma unroll
      for (int j = 0; j < NC / 8; ++j)
#pragma unroll
        for (int q = 0; q < 8; ++q) xr[j][q] = 0.f;
      float rs = 1.f;
      if (grow < n) {
#pragma unroll
        for (int j = 0; j < NC / 8; ++j) ldg256_cs(x1 + grow * C + col0 + 8 * j, xr[j]);
        if (rs_attn) rs = rs_attn[grow];
      }
      mbar_wait(bar(21 + ab), tph);
      fence_after_sync();
      uint32_t v[NC];
      if constexpr (NC == 16) tmem_ld16(tmem + lane_base + L::D0_COL + ab * C + col0, v);
      else tmem_ld8(tmem + lane_base + L::D0_COL + ab * C + col0, v);
      tmem_ld_wait();
      fence_before_sync();
      mbar_arrive(bar(23 + ab));
      float f[NC];
      const float* bps = bs + L::HID + C;
#pragma unroll
      for (int j = 0; j < NC; ++j) f[j] = (__uint_as_float(v[j]) + bps[col0 + j]) * rs;
#pragma unroll
      for (int j = 0; j < NC / 8; ++j)
#pragma unroll
        for (int q = 0; q < 8; ++q) f[8 * j + q] += xr[j][q];
      if (grow < n) {
#pragma unroll
        for (int j = 0; j < NC / 8; ++j) {
          float o8[8];
#pragma unroll
          for (int q = 0; q < 8; ++q) o8[q] = f[8 * j + q];
          stg256(res_out + grow * C + col0 + 8 * j, o8);
        }
      }
      float sum = 0.f, sq = 0.f;
#pragma unroll
      for (int j = 0; j < NC; ++j) {
        sum += f[j];
        sq = fmaf(f[j], f[j], sq);
      }
      float2* st = stats + ((int)ab * TM + row) * 4;
      st[quarter] = make_float2(sum, sq);
      epi_sync();
      const float4 s01 = reinterpret_cast<const float4*>(st)[0], s23 = reinterpret_cast<const float4*>(st)[1];
      const float mean = (s01.x + s01.z + s23.x + s23.z) * (1.f / C);
      const float var = fmaxf((s01.y + s01.w + s23.y + s23.w) * (1.f / C) - mean * mean, 0.f);
      const float rstd = rsqrtf(var + eps);
      const float* gs = bs + L::HID + 2 * C;
      const float* bes = bs + L::HID + 3 * C;
      uint32_t pk[NC / 2];
#pragma unroll
      for (int j = 0; j < NC / 2; ++j)
        pk[j] = pack2((f[2 * j] - mean) * rstd * gs[col0 + 2 * j] + bes[col0 + 2 * j],
                      (f[2 * j + 1] - mean) * rstd * gs[col0 + 2 * j + 1] + bes[col0 + 2 * j + 1]);
      mbar_wait(bar(3 + ab), tph ^ 1u);       // GEMM1 of tile i - 2 no longer reads A1[ab]
      const uint32_t prow = sb + L::A1_OFF + ab * L::A1_BYTES + (row >> 3) * 1024 + (row & 7) * 128;
#pragma unroll
      for (int c4 = 0; c4 < NC / 8; ++c4)
        asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(prow + (((col0 / 8 + c4) ^ (row & 7)) << 4)),
                     "r"(pk[4 * c4]), "r"(pk[4 * c4 + 1]), "r"(pk[4 * c4 + 2]), "r"(pk[4 * c4 + 3])
                     : "memory");
      fence_proxy_async();
      mbar_arrive(bar(1 + ab));
    };

    auto gelu_step = [&](int hb, uint32_t e1) {   // D1 -> + b1 -> GELU -> bf16 A operand of GEMM2
      const uint32_t bb = e1 & 1u;
      mbar_wait(bar(5 + bb), (e1 >> 1) & 1u);
      fence_after_sync();
      uint32_t v[32], pk[16];
      const int gc0 = quarter * 32;
      tmem_ld32(tmem + lane_base + bb * HB + gc0, v);
      tmem_ld_wait();
      fence_before_sync();
      mbar_arrive(bar(7 + bb));
      const float4* bp4 = reinterpret_cast<const float4*>(smem + L::BIAS_OFF) + (hb * HB + gc0) / 4;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 bv = bp4[j];
        pk[2 * j] = pack2(gelu_erf(__uint_as_float(v[4 * j]) + bv.x), gelu_erf(__uint_as_float(v[4 * j + 1]) + bv.y));
        pk[2 * j + 1] =
            pack2(gelu_erf(__uint_as_float(v[4 * j + 2]) + bv.z), gelu_erf(__uint_as_float(v[4 * j + 3]) + bv.w));
      }
      mbar_wait(bar(11 + bb), ((e1 >> 1) & 1u) ^ 1u);
      const uint32_t prow = sb + L::A2_OFF + bb * L::A2_BYTES + (quarter >> 1) * 16384 + (row >> 3) * 1024 +
                            (row & 7) * 128;
#pragma unroll
      for (int c4 = 0; c4 < 4; ++c4)
        asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(
                         prow + ((((quarter & 1) * 4 + c4) ^ (row & 7)) << 4)),
                     "r"(pk[4 * c4]), "r"(pk[4 * c4 + 1]), "r"(pk[4 * c4 + 2]), "r"(pk[4 * c4 + 3])
                     : "memory");
      fence_proxy_async();
      mbar_arrive(bar(9 + bb));
    };

    auto epilogue2 = [&](int i) {   // D2 -> + b2, x rs_m, + x2 (read back from res_out) -> fp32 stream + bf16 copy
      const uint32_t ab = (uint32_t)i & 1u, tph = (uint32_t)(i >> 1) & 1u;
      const int64_t grow = (int64_t)(blockIdx.x + (int64_t)i * gridDim.x) * TM + row;
      float rres[NC / 8][8];
#pragma unroll
      for (int j = 0; j < NC / 8; ++j)
#pragma unroll
        for (int q = 0; q < 8; ++q) rres[j][q] = 0.f;
      float rs = 1.f;
      if (grow < n) {
#pragma unroll
        for (int j = 0; j < NC / 8; ++j) ldg256_cs(res_out + grow * C + col0 + 8 * j, rres[j]);   // this thread's own E0 store
        if (rs_mlp) rs = rs_mlp[grow];
      }
      mbar_wait(bar(13 + ab), tph);
      fence_after_sync();
      uint32_t v[NC];
      if constexpr (NC == 16) tmem_ld16(tmem + lane_base + L::D2_COL + ab * C + col0, v);
      else tmem_ld8(tmem + lane_base + L::D2_COL + ab * C + col0, v);
      tmem_ld_wait();
      fence_before_sync();
      mbar_arrive(bar(15 + ab));
      if (grow < n) {
        float f[NC];
        const float* b2s = bs + L::HID;
#pragma unroll
        for (int j = 0; j < NC; ++j) f[j] = (__uint_as_float(v[j]) + b2s[col0 + j]) * rs;
#pragma unroll
        for (int j = 0; j < NC / 8; ++j) {
          float o8[8];
#pragma unroll
          for (int q = 0; q < 8; ++q) o8[q] = f[8 * j + q] += rres[j][q];
          stg256(res_out + grow * C + col0 + 8 * j, o8);
        }
        if (out_bf16) {
          if constexpr (NC == 16) {
            uint32_t pk8[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) pk8[q] = pack2(f[2 * q], f[2 * q + 1]);
            stg256_u32(out_bf16 + grow * C + col0, pk8);
          } else {
            *reinterpret_cast<uint4*>(out_bf16 + grow * C + col0) =
                make_uint4(pack2(f[0], f[1]), pack2(f[2], f[3]), pack2(f[4], f[5]), pack2(f[6], f[7]));
          }
        }
      }
    };
    // order per tile i: G(i,0)  E0(i+1)  E2(i-1)  G(i,1..)  -- the LayerNorm of the NEXT tile is ready before its first
    // GEMM1 is due, and the wait for GEMM2 of tile i-1 hides behind the first GELU block of tile i.  E0(i+1) must come
    // BEFORE E2(i-1): the issuer queues GEMM1 of tile i+1 (which waits for E0(i+1)) ahead of the last GEMM2 of tile i-1
    // when a tile has a single hidden block (C = 32), so the other order deadlocks.
    uint32_t e1 = 0;
    if (my_tiles > 0) e0_step(0);
    for (int i = 0; i <= my_tiles; ++i) {
      if (i < my_tiles) gelu_step(0, e1++);
      if (i + 1 < my_tiles) e0_step(i + 1);
      if (i >= 1) epilogue2(i - 1);
      if (i < my_tiles)
        for (int hb = 1; hb < L::NHB; ++hb) gelu_step(hb, e1++);
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 2) tmem_dealloc<512>(tmem);
}

template <int C>
static int launch_proj_mlp(const __nv_bfloat16* a, const __nv_bfloat16* wpp, const float* bp, const float* x1,
                           const float* rs_attn, const float* g2, const float* be2, float eps, const __nv_bfloat16* w1p,
                           const float* b1, const __nv_bfloat16* w2p, const float* b2, const float* rs_mlp, int64_t n,
                           float* res_out, __nv_bfloat16* out_bf16, cudaStream_t st) {
  using L = mlp::PSmem<C>;
  static_assert(L::DYN_BYTES <= 232448, "shared memory budget");
  static_assert(L::D0_COL + 2 * C <= 512, "TMEM columns");
  static bool configured_dev[kMaxDeviceSlots] = {};
  bool& configured = configured_dev[device_slot()];
  if (!configured) {
    if (cudaFuncSetAttribute(proj_mlp_fused_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::DYN_BYTES) !=
        cudaSuccess)
      return PTV3_ELAUNCH;
    configured = true;
  }
  const int n_tiles = (int)cdiv(n, mlp::TM);
  const int g = n_tiles < kNumSMs ? n_tiles : kNumSMs;
  proj_mlp_fused_kernel<C><<<g, mlp::THREADS, L::DYN_BYTES, st>>>(a, wpp, bp, x1, rs_attn, g2, be2, eps, w1p, b1, w2p,
                                                                   b2, rs_mlp, n, n_tiles, res_out, out_bf16);
  return check_launch();
}

// ---------------------------------------------------------------------------------------------------------------
// The same MLP for C = 128 (hidden 512), round 2: the two weight matrices (2 x 128 KB) do not fit in shared memory, so
// they are STREAMED: per hidden block hb the 32 KB images W1[hb] (128 hidden rows x K = 128) and W2[:, hb] (128 output
// rows x K = 128) travel through a 3-slot ring, fetched from L2 by one bulk copy each in exactly the order the issuer
// consumes them (fc1(0) fc1(1) | fc1(t+2) fc2(t) ...).  X tile single-buffered (2 K-blocks), GELU tile double-buffered,
// D1 2 x 128 + D2 2 x 128 = all 512 TMEM columns.  Replaces cuBLASLt fc1 + gelu_bf16_kernel + cuBLASLt fc2 +
// ln_residual (40 C bytes per row through HBM) by 12 C.
// ---------------------------------------------------------------------------------------------------------------
namespace mlps {
constexpr int TM = 128, THREADS = 768, LOADERS = 64, EPI = 512, OUTW = 128, WBLK = 32768, NSLOT = 3;
// warps: 0-1 X loaders | 2 MMA issuer (+ TMEM alloc) | 3 weight streamer | 4-19 GELU epilogue | 20-23 output epilogue
// C = 128: hidden blocks of 128 (GEMM1 N = 128, K = 128; GEMM2 N = 128, K = 128), D1 2 x 128 + D2 2 x 128 TMEM columns.
// C = 256: hidden blocks of  64 (GEMM1 N =  64, K = 256; GEMM2 N = 256, K =  64), D1 2 x 64 + ONE D2 of 256 columns.
// Either way a weight block is 32 KB and the ring holds three.
template <int C>
struct Cfg {
  static constexpr int HID = 4 * C;
  static constexpr int HB = C == 128 ? 128 : 64;
  static constexpr int NHB = HID / HB;
  static constexpr int KB1 = C / 64;                       // K-blocks of the X tile
  static constexpr int KB2 = HB / 64;                      // K-blocks of the GELU tile
  static constexpr bool D2_DOUBLE = C == 128;
  static constexpr int A1_BYTES = TM * C * 2;
  static constexpr int A2_BYTES = TM * HB * 2;
  static constexpr int RING_OFF = 0;
  static constexpr int A1_OFF = RING_OFF + NSLOT * WBLK;
  static constexpr int A2_OFF = A1_OFF + A1_BYTES;
  static constexpr int BIAS_OFF = A2_OFF + 2 * A2_BYTES;   // b1 (HID) | b2 (C)
  static constexpr int BAR_OFF = BIAS_OFF + (HID + C) * 4;
  static constexpr int DYN_BYTES = BAR_OFF + 256;          // no alignment slack: the dynamic array is declared __align__(1024)
  static constexpr int GC = HB / 4;                        // hidden columns a thread handles per GELU step (32 / 16)
  static_assert(HB * C * 2 == WBLK, "weight block size");
};
// barriers: 0-3 w_full | 4-7 w_free | 8 a1_full | 9 a1_free | 10,11 d1_full | 12,13 d1_free | 14,15 a2_full | 16,17 a2_free
//           | 18,19 d2_full | 20,21 d2_free
}  // namespace mlps

template <int C>
__global__ void __launch_bounds__(mlps::THREADS, 1) mlp_stream_kernel(
    const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ w1p, const float* __restrict__ b1,
    const __nv_bfloat16* __restrict__ w2p, const float* __restrict__ b2, const float* __restrict__ residual,
    const float* __restrict__ row_scale, int64_t n, int n_tiles, float* __restrict__ res_out,
    __nv_bfloat16* __restrict__ out_bf16, int dbg) {
  using namespace mlps;
  using L = Cfg<C>;
  using mlp::gelu_erf;
  using mlp::pack2;
  constexpr int HB = L::HB, NHB = L::NHB;
  extern __shared__ __align__(1024) uint8_t smem_aligned[];
  uint8_t* smem = smem_aligned;
  const uint32_t sb = smem_u32(smem);
  if (sb & 1023u) __trap();      // the swizzled operand layouts need the 1024-byte alignment the declaration asks for
  auto bar = [&](int id) { return sb + L::BAR_OFF + 8u * (uint32_t)id; };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + L::BAR_OFF + 240);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int my_tiles = (n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const int T = my_tiles * NHB;
  // D2 buffer / phase of tile `it`
  auto d2_buf = [](int it) -> uint32_t { return L::D2_DOUBLE ? (uint32_t)it & 1u : 0u; };
  auto d2_phase = [](int it) -> uint32_t { return L::D2_DOUBLE ? (uint32_t)(it >> 1) & 1u : (uint32_t)it & 1u; };

  if (tid == 0) {
    for (int s = 0; s < NSLOT; ++s) { mbar_init(bar(s), 1); mbar_init(bar(4 + s), 1); }
    mbar_init(bar(8), LOADERS); mbar_init(bar(9), 1);
    for (int b = 0; b < 2; ++b) {
      mbar_init(bar(10 + b), 1);   mbar_init(bar(12 + b), EPI);
      mbar_init(bar(14 + b), EPI); mbar_init(bar(16 + b), 1);
      mbar_init(bar(18 + b), 1);   mbar_init(bar(20 + b), OUTW);
    }
    mbar_fence_init();
  }
  for (int j = tid; j < L::HID + C; j += THREADS)
    reinterpret_cast<float*>(smem + L::BIAS_OFF)[j] = j < L::HID ? b1[j] : b2[j - L::HID];
  if (warp == 2) tmem_alloc<512>(smem_u32(tmem_slot));
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = *tmem_slot;

  if (warp < 2) {
    // =========================== X tile loaders ===========================
    constexpr int CH = C / 8;                       // 16-byte chunks per row
    constexpr int RSTEP = LOADERS / CH;             // rows covered per pass
    const int c = tid % CH, r0 = tid / CH;
    for (int i = 0; i < my_tiles; ++i) {
      const int64_t row0 = (int64_t)(blockIdx.x + (int64_t)i * gridDim.x) * TM;
      mbar_wait(bar(9), ((uint32_t)i & 1u) ^ 1u);
      const uint32_t a1 = sb + L::A1_OFF + (c >> 3) * 16384;
#pragma unroll 8
      for (int q = 0; q < TM / RSTEP; ++q) {
        const int r = r0 + RSTEP * q;
        const uint32_t dst = a1 + (r >> 3) * 1024 + (r & 7) * 128 + (((c & 7) ^ (r & 7)) << 4);
        if (dbg & 4) continue;
        if (row0 + r < n) cp_async16(dst, x + (row0 + r) * C + c * 8, 16u);
        else st_shared_zero16(dst);
      }
      cp_async_commit();
      cp_async_wait<0>();
      fence_proxy_async();
      mbar_arrive(bar(8));
    }
  } else if (warp == 3) {
    // =========================== weight streamer (one thread) ===========================
    // same sequence as the issuer: fc1(0) fc1(1) | fc1(t+2) fc2(t) ...
    if (lane == 0) {
      uint32_t gi = 0;
      auto fetch = [&](const __nv_bfloat16* base, int hb) {
        const uint32_t slot = gi % NSLOT, ph = (gi / NSLOT) & 1u;
        mbar_wait(bar(4 + slot), ph ^ 1u);
        const uint32_t bytes = (dbg & 2) ? 16u : (uint32_t)WBLK;
        mbar_arrive_expect_tx(bar(slot), bytes);
        bulk_g2s(sb + L::RING_OFF + slot * WBLK, reinterpret_cast<const char*>(base) + (size_t)hb * WBLK, bytes, bar(slot));
        ++gi;
      };
      if (T > 0) fetch(w1p, 0);
      if (T > 1) fetch(w1p, 1 % NHB);
      for (int t = 0; t < T; ++t) {
        if (t + 2 < T) fetch(w1p, (t + 2) % NHB);
        fetch(w2p, t % NHB);
      }
    }
  } else if (warp == 2) {
    // =========================== MMA issuer ===========================
    constexpr uint32_t idesc1 = idesc_bf16_f32(TM, HB), idesc2 = idesc_bf16_f32(TM, C);
    uint32_t gi = 0;
    auto fc1 = [&](int t) {
      const int it = t / NHB, hb = t - it * NHB;
      const uint32_t bb = (uint32_t)t & 1u;
      const uint32_t slot = gi % NSLOT, ph = (gi / NSLOT) & 1u;
      ++gi;
      if (hb == 0) mbar_wait(bar(8), (uint32_t)it & 1u);
      mbar_wait(bar(12 + bb), (((uint32_t)t >> 1) & 1u) ^ 1u);
      mbar_wait(bar(slot), ph);
      fence_after_sync();
#pragma unroll
      for (int kb = 0; kb < L::KB1; ++kb) {      // W1 block: [KB1 K-blocks][HB rows][64]
        const uint64_t a_desc = smem_desc_sw128(sb + L::A1_OFF + kb * 16384);
        const uint64_t b_desc = smem_desc_sw128(sb + L::RING_OFF + slot * WBLK + kb * (HB * 128));
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          mma_bf16_ss(tmem + bb * HB, a_desc + 2 * kk, b_desc + 2 * kk, idesc1, (kb | kk) ? 1u : 0u);
      }
      mma_commit(bar(10 + bb));
      mma_commit(bar(4 + slot));
      if (hb == NHB - 1) mma_commit(bar(9));
    };
    auto fc2 = [&](int t) {
      const int it = t / NHB, h2 = t - it * NHB;
      const uint32_t ab = d2_buf(it), bb = (uint32_t)t & 1u;
      const uint32_t slot = gi % NSLOT, ph = (gi / NSLOT) & 1u;
      ++gi;
      mbar_wait(bar(14 + bb), ((uint32_t)t >> 1) & 1u);
      if (h2 == 0) mbar_wait(bar(20 + ab), d2_phase(it) ^ 1u);
      mbar_wait(bar(slot), ph);
      fence_after_sync();
#pragma unroll
      for (int kb = 0; kb < L::KB2; ++kb) {      // W2 block: [KB2 K-blocks][C rows][64]
        const uint64_t a_desc = smem_desc_sw128(sb + L::A2_OFF + bb * L::A2_BYTES + kb * 16384);
        const uint64_t b_desc = smem_desc_sw128(sb + L::RING_OFF + slot * WBLK + kb * (C * 128));
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          mma_bf16_ss(tmem + 256 + ab * C, a_desc + 2 * kk, b_desc + 2 * kk, idesc2, (h2 | kb | kk) ? 1u : 0u);
      }
      mma_commit(bar(16 + bb));
      mma_commit(bar(4 + slot));
      if (h2 == NHB - 1) mma_commit(bar(18 + ab));
    };
    if (elect_one_sync()) {
      if (T > 0) fc1(0);
      if (T > 1) fc1(1);
      for (int t = 0; t < T; ++t) {
        if (t + 2 < T) fc1(t + 2);
        fc2(t);
      }
    }
    __syncwarp();
  } else {
    // =========================== epilogue warps ===========================
    // 16 GELU warps (D1 -> A2) and 4 OUTPUT warps (D2 -> HBM).  On the same warps the output epilogue's HBM round trip
    // (residual read, stores) stalled the GELU steps of the next tile: 36 % of the kernel (timing ablation 8).
    const int q4 = warp & 3, quarter = (warp - 4) >> 2;
    const int row = q4 * 32 + lane;
    const uint32_t lane_base = (uint32_t)(q4 * 32) << 16;
    auto gelu_step = [&](int hb, uint32_t e1) {
      constexpr int GC = L::GC;
      const uint32_t bb = e1 & 1u;
      mbar_wait(bar(10 + bb), (e1 >> 1) & 1u);
      fence_after_sync();
      uint32_t v[GC], pk[GC / 2];
      const int col0 = quarter * GC;
      if constexpr (GC == 32) tmem_ld32(tmem + lane_base + bb * HB + col0, v);
      else tmem_ld16(tmem + lane_base + bb * HB + col0, v);
      tmem_ld_wait();
      fence_before_sync();
      mbar_arrive(bar(12 + bb));
      const float4* bp = reinterpret_cast<const float4*>(smem + L::BIAS_OFF) + (hb * HB + col0) / 4;
#pragma unroll
      for (int j = 0; j < GC / 4; ++j) {
        const float4 bv = bp[j];
        if (dbg & 1) {
          pk[2 * j] = pack2(__uint_as_float(v[4 * j]) + bv.x, __uint_as_float(v[4 * j + 1]) + bv.y);
          pk[2 * j + 1] = pack2(__uint_as_float(v[4 * j + 2]) + bv.z, __uint_as_float(v[4 * j + 3]) + bv.w);
          continue;
        }
        pk[2 * j] = pack2(gelu_erf(__uint_as_float(v[4 * j]) + bv.x), gelu_erf(__uint_as_float(v[4 * j + 1]) + bv.y));
        pk[2 * j + 1] =
            pack2(gelu_erf(__uint_as_float(v[4 * j + 2]) + bv.z), gelu_erf(__uint_as_float(v[4 * j + 3]) + bv.w));
      }
      mbar_wait(bar(16 + bb), ((e1 >> 1) & 1u) ^ 1u);
      // my GC hidden columns = GC / 8 16-byte chunks, chunk index col0 / 8 + c4 inside the 64-column K-blocks of the tile
      const uint32_t prow = sb + L::A2_OFF + bb * L::A2_BYTES + (row >> 3) * 1024 + (row & 7) * 128;
#pragma unroll
      for (int c4 = 0; c4 < GC / 8; ++c4) {
        const int chunk = col0 / 8 + c4;
        asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(prow + (chunk >> 3) * 16384 +
                                                                     (((chunk & 7) ^ (row & 7)) << 4)),
                     "r"(pk[4 * c4]), "r"(pk[4 * c4 + 1]), "r"(pk[4 * c4 + 2]), "r"(pk[4 * c4 + 3])
                     : "memory");
      }
      fence_proxy_async();
      mbar_arrive(bar(14 + bb));
    };
    auto epilogue2 = [&](int i) {   // output warps: this thread owns row `row` and walks all C columns in pieces of 32
      constexpr int PIECES = C / 32;
      const uint32_t ab = d2_buf(i), tph = d2_phase(i);
      const int64_t grow = (int64_t)(blockIdx.x + (int64_t)i * gridDim.x) * TM + row;
      // the first piece of the residual row is requested BEFORE waiting for GEMM2 (its latency hides behind the wait)
      float rres[4][8];
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int q = 0; q < 8; ++q) rres[j][q] = 0.f;
      float rs = 1.f;
      if (grow < n && !(dbg & 8)) {
        if (residual) {
#pragma unroll
          for (int j = 0; j < 4; ++j) ldg256_cs(residual + grow * C + 8 * j, rres[j]);
        }
        if (row_scale) rs = row_scale[grow];
      }
      mbar_wait(bar(18 + ab), tph);
      fence_after_sync();
      if (dbg & 8) {      // timing ablation: no output epilogue at all
        fence_before_sync();
        mbar_arrive(bar(20 + ab));
        return;
      }
#pragma unroll 1
      for (int pc = 0; pc < PIECES; ++pc) {
        const int col0 = 32 * pc;
        uint32_t v[32];
        tmem_ld32(tmem + lane_base + 256 + ab * C + col0, v);
        tmem_ld_wait();
        if (pc == PIECES - 1) {
          fence_before_sync();
          mbar_arrive(bar(20 + ab));
        }
        uint32_t pk[16];
        if (grow < n) {
          const float* b2s = reinterpret_cast<const float*>(smem + L::BIAS_OFF) + L::HID + col0;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float f[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) f[q] = fmaf(__uint_as_float(v[8 * j + q]) + b2s[8 * j + q], rs, rres[j][q]);
            if (res_out) stg256(res_out + grow * C + col0 + 8 * j, f);
#pragma unroll
            for (int q = 0; q < 4; ++q) pk[4 * j + q] = pack2(f[2 * q], f[2 * q + 1]);
          }
          // next piece of the residual row: in flight while this piece's bf16 copy is stored and the next tcgen05.ld runs
          if (residual && pc + 1 < PIECES) {
#pragma unroll
            for (int j = 0; j < 4; ++j) ldg256_cs(residual + grow * C + col0 + 32 + 8 * j, rres[j]);
          }
          if (out_bf16) {
            uint32_t lo[8], hi[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) { lo[q] = pk[q]; hi[q] = pk[8 + q]; }
            stg256_u32(out_bf16 + grow * C + col0, lo);
            stg256_u32(out_bf16 + grow * C + col0 + 16, hi);
          }
        }
      }
    };
    if (warp < 20) {
      uint32_t e1 = 0;
      for (int i = 0; i < my_tiles; ++i)
        for (int hb = 0; hb < NHB; ++hb) gelu_step(hb, e1++);
    } else {
      for (int i = 0; i < my_tiles; ++i) epilogue2(i);
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 2) tmem_dealloc<512>(tmem);
}

template <int C>
static int launch_mlp_stream(const __nv_bfloat16* x, const __nv_bfloat16* w1p, const float* b1, const __nv_bfloat16* w2p,
                             const float* b2, const float* residual, const float* row_scale, int64_t n, float* res_out,
                             __nv_bfloat16* out_bf16, cudaStream_t st) {
  using L = mlps::Cfg<C>;
  static_assert(L::DYN_BYTES <= 232448, "shared memory budget");
  static bool configured_dev[kMaxDeviceSlots] = {};
  bool& configured = configured_dev[device_slot()];
  if (!configured) {
    if (cudaFuncSetAttribute(mlp_stream_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::DYN_BYTES) !=
        cudaSuccess)
      return PTV3_ELAUNCH;
    configured = true;
  }
  const int n_tiles = (int)cdiv(n, mlps::TM);
  const int g = n_tiles < kNumSMs ? n_tiles : kNumSMs;
  static int dbg = -1;      // PTV3_MLP_STREAM_DBG: timing ablations (results wrong by design): 1 no GELU, 2 16-byte weight
  if (dbg < 0) {            // fetches, 4 no X loads
    const char* e = getenv("PTV3_MLP_STREAM_DBG");
    dbg = e ? atoi(e) : 0;
  }
  mlp_stream_kernel<C><<<g, mlps::THREADS, L::DYN_BYTES, st>>>(x, w1p, b1, w2p, b2, residual, row_scale, n, n_tiles,
                                                                 res_out, out_bf16, dbg);
  return check_launch();
}

// rows per packing tile of fc1's weight (ptv3_pack_kmajor tile_rows): the hidden-block width of the kernel that will read it
int mlp_fc1_tile_rows(int c) { return c == 256 ? 64 : 128; }

bool mlp_fused_supported(int c, int hidden) { return (c == 32 || c == 64 || c == 128 || c == 256) && hidden == 4 * c; }

template <int C>
static int launch_mlp(const __nv_bfloat16* x, const __nv_bfloat16* w1p, const float* b1, const __nv_bfloat16* w2p,
                      const float* b2, const float* residual, const float* row_scale, int64_t n, float* res_out,
                      __nv_bfloat16* out_bf16, cudaStream_t st) {
  using L = mlp::Smem<C>;
  static_assert(L::DYN_BYTES <= 232448, "shared memory budget");
  static bool configured_dev[kMaxDeviceSlots] = {};
  bool& configured = configured_dev[device_slot()];
  if (!configured) {
    cudaFuncSetAttribute(mlp_fused_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::DYN_BYTES);
    configured = true;
  }
  const int n_tiles = (int)cdiv(n, mlp::TM);
  const int g = n_tiles < kNumSMs ? n_tiles : kNumSMs;
  static int dbg = -1;      // PTV3_MLP_STREAM_DBG bit 8: timing ablation without the output epilogue (results wrong)
  if (dbg < 0) {
    const char* e = getenv("PTV3_MLP_STREAM_DBG");
    dbg = e ? atoi(e) : 0;
  }
  mlp_fused_kernel<C><<<g, mlp::THREADS + mlp::OUTW, L::DYN_BYTES, st>>>(x, w1p, b1, w2p, b2, residual, row_scale, n, n_tiles,
                                                              res_out, out_bf16, dbg);
  return check_launch();
}

}  // namespace ptv3

using namespace ptv3;

extern "C" {

size_t ptv3_pack_kmajor_bytes(int rows, int k) { return (size_t)rows * (size_t)((k + 63) / 64) * 128; }

int ptv3_pack_kmajor(const void* w, int rows, int k, int tile_rows, void* packed, void* stream) {
  if (!w || !packed || rows < 1 || k < 8 || (k % 8) != 0 || tile_rows < 8 || (tile_rows % 8) != 0 ||
      (rows % tile_rows) != 0)
    return PTV3_EINVAL;
  const int n_kb = (k + 63) / 64;
  pack_kmajor_kernel<<<grid_for((int64_t)n_kb * rows * 8, 256), 256, 0, as_stream(stream)>>>(
      (const __nv_bfloat16*)w, rows, k, tile_rows, n_kb, (__nv_bfloat16*)packed);
  return check_launch();
}

int ptv3_mlp_fused_supported(int c, int hidden) { return mlp_fused_supported(c, hidden) ? 1 : 0; }

int ptv3_mlp_fc1_tile_rows(int c) { return mlp_fc1_tile_rows(c); }

int ptv3_mlp_fused(const void* x, const void* w1_packed, const float* b1, const void* w2_packed, const float* b2,
                   const float* residual, int64_t n, int c, int hidden, float* res_out, void* out_bf16, void* stream) {
  return ptv3_mlp_fused_rs(x, w1_packed, b1, w2_packed, b2, residual, nullptr, n, c, hidden, res_out, out_bf16, stream);
}

int ptv3_mlp_fused_rs(const void* x, const void* w1_packed, const float* b1, const void* w2_packed, const float* b2,
                      const float* residual, const float* row_scale, int64_t n, int c, int hidden, float* res_out,
                      void* out_bf16, void* stream) {
  if (n < 0 || !mlp_fused_supported(c, hidden)) return n < 0 ? PTV3_EINVAL : PTV3_EUNSUPPORTED;
  if (n == 0) return PTV3_OK;
  if (!x || !w1_packed || !b1 || !w2_packed || !b2 || (!res_out && !out_bf16)) return PTV3_EINVAL;
  cudaStream_t st = as_stream(stream);
  if (c == 128)
    return launch_mlp_stream<128>((const __nv_bfloat16*)x, (const __nv_bfloat16*)w1_packed, b1,
                                  (const __nv_bfloat16*)w2_packed, b2, residual, row_scale, n, res_out,
                                  (__nv_bfloat16*)out_bf16, st);
  if (c == 256)
    return launch_mlp_stream<256>((const __nv_bfloat16*)x, (const __nv_bfloat16*)w1_packed, b1,
                                  (const __nv_bfloat16*)w2_packed, b2, residual, row_scale, n, res_out,
                                  (__nv_bfloat16*)out_bf16, st);
  if (c == 32)
    return launch_mlp<32>((const __nv_bfloat16*)x, (const __nv_bfloat16*)w1_packed, b1, (const __nv_bfloat16*)w2_packed,
                          b2, residual, row_scale, n, res_out, (__nv_bfloat16*)out_bf16, st);
  return launch_mlp<64>((const __nv_bfloat16*)x, (const __nv_bfloat16*)w1_packed, b1, (const __nv_bfloat16*)w2_packed,
                        b2, residual, row_scale, n, res_out, (__nv_bfloat16*)out_bf16, st);
}

int ptv3_proj_mlp_fused(const void* a, const void* wproj_packed, const float* bproj, const float* x1,
                        const float* rs_attn, const float* ln_gamma, const float* ln_beta, float eps,
                        const void* w1_packed, const float* b1, const void* w2_packed, const float* b2,
                        const float* rs_mlp, int64_t n, int c, int hidden, float* res_out, void* out_bf16,
                        void* stream) {
  if (n < 0 || !mlp_fused_supported(c, hidden) || c > 64) return n < 0 ? PTV3_EINVAL : PTV3_EUNSUPPORTED;
  if (n == 0) return PTV3_OK;
  if (!a || !wproj_packed || !bproj || !x1 || !ln_gamma || !ln_beta || !w1_packed || !b1 || !w2_packed || !b2 ||
      !res_out || res_out == x1)
    return PTV3_EINVAL;
  cudaStream_t st = as_stream(stream);
#define PTV3_PM(CC)                                                                                              \
  launch_proj_mlp<CC>((const __nv_bfloat16*)a, (const __nv_bfloat16*)wproj_packed, bproj, x1, rs_attn, ln_gamma, \
                      ln_beta, eps, (const __nv_bfloat16*)w1_packed, b1, (const __nv_bfloat16*)w2_packed, b2,     \
                      rs_mlp, n, res_out, (__nv_bfloat16*)out_bf16, st)
  const int rc = c == 32 ? PTV3_PM(32) : PTV3_PM(64);
#undef PTV3_PM
  return rc;
}

}  // extern "C"
