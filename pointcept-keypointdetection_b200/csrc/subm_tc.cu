// A5: bf16 submanifold convolution as an implicit GEMM on the 5th-gen tensor cores (tcgen05 + TMEM).
//
//   out[i, :] = bias + sum_k W[k] . in[nbr[i,k], :]      (nbr = cached neighbour map, -1 = absent)
//
// One CTA owns 128 output rows x NT output channels.  The K dimension is the concatenation (offset k, c_in) cut
// into K-blocks of 64 bf16 (128 bytes): for c_in >= 64 a block is one (offset, 64-channel slice), for c_in == 32 a
// block is two offsets.  Per K-block
//   A (128 x 64): rows gathered from HBM/L2 through the neighbour map by 4 producer warps with 16-byte cp.async
//                 (zero fill for absent neighbours), written in the 128-byte-swizzled K-major layout UMMA reads;
//   B (NT x 64):  the weight block, packed once by subm_pack_weight_kernel into exactly that shared-memory image,
//                 so one 1-D bulk copy (TMA unit, cp.async.bulk + mbarrier complete_tx) brings it in;
//   MMA:          one elected thread issues 4 x tcgen05.mma (M128, N=NT, K16), accumulator in TMEM (NT columns).
// K-blocks whose offsets have no present neighbour anywhere in the 128-row tile are skipped.
// Epilogue: tcgen05.ld (32 lanes x 32 columns per warp) -> + bias -> bf16 -> 64-byte row segments to HBM.
#include <cstdlib>
#include "common.cuh"
#include "tc05.cuh"

namespace ptv3 {

using namespace tc05;

constexpr int TM = 128;          // output rows per CTA (UMMA M)
constexpr int KBLK = 64;         // bf16 per K-block (one 128-byte swizzle row)
constexpr int A_BYTES = TM * KBLK * 2;
constexpr int MAX_K3 = 27;

static inline int n_kblocks(int c_in, int k3) { return (k3 * c_in + KBLK - 1) / KBLK; }
static inline int pick_nt(int c_out) { return c_out >= 256 ? 256 : c_out; }

bool subm_conv_tc_supported(int c_in, int c_out, int k3) {
  bool cin_ok = c_in == 32 || (c_in % 64 == 0 && c_in >= 64 && c_in <= 1024);      // PTv3-large: 768
  bool cout_ok = c_out == 32 || c_out == 64 || c_out == 128 || (c_out % 256 == 0 && c_out >= 256 && c_out <= 1024);
  return cin_ok && cout_ok && k3 >= 1 && k3 <= MAX_K3;
}

size_t subm_conv_tc_packed_bytes(int c_in, int c_out, int k3) {
  return (size_t)n_kblocks(c_in, k3) * c_out * KBLK * 2;
}
size_t subm_conv_tc_workspace(int64_t, int c_in, int c_out, int k3) {
  return align_up(subm_conv_tc_packed_bytes(c_in, c_out, k3));
}

// packed[nt_tile][kb][co_local][64] with the 128-byte swizzle applied: the exact shared-memory image of a B block.
__global__ void __launch_bounds__(256) subm_pack_weight_kernel(const __nv_bfloat16* __restrict__ w, int c_in,
                                                               int c_out, int k3, int nt, int n_kb,
                                                               __nv_bfloat16* __restrict__ packed) {
  // one thread per 16-byte chunk (8 bf16)
  const int64_t chunks = (int64_t)n_kb * c_out * 8;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < chunks; t += (int64_t)gridDim.x * blockDim.x) {
    int c = (int)(t & 7);
    int64_t rowid = t >> 3;               // (tile, kb, co_local) flattened
    int co_l = (int)(rowid % nt);
    int64_t r2 = rowid / nt;
    int kb = (int)(r2 % n_kb);
    int tile = (int)(r2 / n_kb);
    int co = tile * nt + co_l;
    int kk = kb * KBLK + c * 8;           // position in the concatenated K = (offset, c_in)
    int k = kk / c_in, ci = kk % c_in;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (k < k3) v = *reinterpret_cast<const uint4*>(w + ((int64_t)k * c_out + co) * c_in + ci);
    char* base = reinterpret_cast<char*>(packed) + ((int64_t)tile * n_kb + kb) * nt * 128;
    int off = (co_l >> 3) * 1024 + (co_l & 7) * 128 + ((c ^ (co_l & 7)) << 4);
    *reinterpret_cast<uint4*>(base + off) = v;
  }
}

template <int NT, int STAGES>
struct ConvSmem {
  static constexpr int B_BYTES = NT * KBLK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int NBR_OFF = STAGES * STAGE_BYTES;
  static constexpr int NBR_BYTES = TM * MAX_K3 * 4;              // one neighbour tile; two buffers
  static constexpr int MISC_OFF = NBR_OFF + 2 * NBR_BYTES;       // stage flags, masks, tmem pointer
  static constexpr int BAR_OFF = MISC_OFF + 256;                 // mbarriers
  static constexpr int TOTAL = BAR_OFF + 8 * (2 * STAGES + 6);
  static constexpr int DYN_BYTES = TOTAL + 1024;                 // slack for the 1024-byte alignment
};

int g_conv_async = 1;   // debug slot 3 / PTV3_CONV_ASYNC: 1 = gather copies arrive on the stage barrier themselves (default), 0 = wait + fence + arrive

constexpr int kGatherWarps = 16;
constexpr int kGatherGroups = 4;                       // K-block g is filled by the 4 warps of group g % 4
constexpr int kWarpsPerGroup = kGatherWarps / kGatherGroups;
constexpr int kRowsPerWarp = TM / kWarpsPerGroup;      // 32 rows of the A stage per warp
constexpr int kSlots = kRowsPerWarp / 4;               // 16-byte slots per lane and K-block
constexpr int PRODUCERS = kGatherWarps * 32;
constexpr int kMmaWarp = kGatherWarps;                 // warp 16: MMA issuer; warps 17-20: epilogue
constexpr int kConvThreads = (kGatherWarps + 5) * 32;  // 672

__device__ __forceinline__ void producers_sync() { asm volatile("bar.sync 1, %0;" ::"n"(PRODUCERS) : "memory"); }

// Persistent: CTA b handles row tiles b, b + gridDim.x, ...  The three roles run decoupled through mbarriers, so the
// gather of tile i+1 (and the bulk prefetch of its neighbour rows) overlaps the MMAs of tile i and the epilogue of
// tile i-1 (two accumulators in TMEM).
//
// Gather warps: K-block number g (a CTA-wide running count over tiles) belongs to warp group g % 4; its four warps
// fill 32 rows of the 128 x 64 A stage each (8 slots per lane), wait for their own copies and arrive.  Four K-blocks
// are therefore in flight, and the per-K-block bookkeeping (stage acquire, fence, arrive) is paid by 4 warps, not 16.
template <int NT, int STAGES, bool ASYNC>
__global__ void __launch_bounds__(kConvThreads, 1) subm_conv_tc_kernel(const __nv_bfloat16* __restrict__ in,
                                                                       const __nv_bfloat16* __restrict__ packed,
                                                                       const float* __restrict__ bias,
                                                                       const int32_t* __restrict__ nbr, int64_t n,
                                                                       int c_in, int c_out, int k3, int n_kb_total,
                                                                       int n_tiles, __nv_bfloat16* __restrict__ out) {
  using L = ConvSmem<NT, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t s_base = smem_u32(smem);
  uint32_t* masks = reinterpret_cast<uint32_t*>(smem + L::MISC_OFF + 64);        // [4]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + L::MISC_OFF + 128);
  volatile uint32_t* tile_nkb = reinterpret_cast<volatile uint32_t*>(smem + L::MISC_OFF + 160);   // [16] K-blocks per tile
  const uint32_t bar_base = s_base + L::BAR_OFF;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  auto nbr_full = [&](int b) { return bar_base + 8u * (2 * STAGES + b); };
  auto acc_full = [&](int b) { return bar_base + 8u * (2 * STAGES + 2 + b); };
  auto acc_free = [&](int b) { return bar_base + 8u * (2 * STAGES + 4 + b); };

  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // provably warp-uniform for nvcc (role branches)
  const int nt_tile = blockIdx.y;
  const int my_tiles = (n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) {
      // one arrive per warp of the owning group (per lane when the copies arrive themselves) + the weight tx bytes
      mbar_init(full_bar(s), ASYNC ? kWarpsPerGroup * 32 : kWarpsPerGroup);
      mbar_init(empty_bar(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(nbr_full(b), 1);
      mbar_init(acc_full(b), 1);
      mbar_init(acc_free(b), 128);
    }
    mbar_fence_init();
    masks[0] = masks[1] = masks[2] = masks[3] = 0;
  }
  if (warp == kMmaWarp) tmem_alloc<2 * NT>(smem_u32(tmem_slot));
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < kMmaWarp) {
    // =========================== gather warps ===========================
    const int c = lane & 7, rr = lane >> 3;   // lane -> (16-byte chunk, row within a group of 4)
    const int group = warp / kWarpsPerGroup, wg = warp % kWarpsPerGroup;
    // byte offsets of this lane's slot inside an 8-row / 1024-byte swizzle group, for even and odd passes
    const uint32_t off_even = (uint32_t)(rr * 128 + ((c ^ rr) << 4));
    const uint32_t off_odd = (uint32_t)((4 + rr) * 128 + ((c ^ (4 + rr)) << 4));
    const char* wtile = reinterpret_cast<const char*>(packed) + (int64_t)nt_tile * n_kb_total * L::B_BYTES;
    const uint32_t nbr_tile_bytes = (uint32_t)(TM * k3 * 4);
    const uint32_t nbr_step = (uint32_t)(16 * k3);   // 4 rows further in the neighbour tile
    const uint32_t row_bytes = (uint32_t)c_in * 2u;
    auto tile_full = [&](int i) { return ((int64_t)(blockIdx.x + (int64_t)i * gridDim.x) + 1) * TM <= n; };
    auto prefetch_nbr = [&](int i) {   // one elected thread of warp 0: bulk copy of a full tile's neighbour rows
      const int64_t row0 = (int64_t)(blockIdx.x + (int64_t)i * gridDim.x) * TM;
      mbar_arrive_expect_tx(nbr_full(i & 1), nbr_tile_bytes);
      bulk_g2s(s_base + L::NBR_OFF + (i & 1) * L::NBR_BYTES, nbr + row0 * k3, nbr_tile_bytes, nbr_full(i & 1));
    };
    if (warp == 0 && my_tiles > 0 && tile_full(0)) {
      if (elect_one_sync()) prefetch_nbr(0);
      __syncwarp();
    }
    uint32_t g0 = 0;  // K-blocks issued by this CTA before the current tile (identical in every gather warp)
    for (int i = 0; i < my_tiles; ++i) {
      const int tb = i & 1;
      const int64_t row0 = (int64_t)(blockIdx.x + (int64_t)i * gridDim.x) * TM;
      int32_t* nbr_s = reinterpret_cast<int32_t*>(smem + L::NBR_OFF + tb * L::NBR_BYTES);
      if (warp == 0) {
        if (elect_one_sync()) {
          masks[(i + 2) & 3] = 0;
          if (i + 1 < my_tiles && tile_full(i + 1)) prefetch_nbr(i + 1);   // buffer tb^1 was released by the
        }                                                                  // producers_sync() that ended tile i-1
        __syncwarp();
      }
      if (tile_full(i)) {
        mbar_wait(nbr_full(tb), (uint32_t)(i >> 1) & 1u);
      } else {  // ragged last tile: plain loads, -1 beyond n
        for (int idx = tid; idx < TM * k3; idx += PRODUCERS) {
          const int64_t row = row0 + idx / k3;
          nbr_s[idx] = row < n ? nbr[row0 * k3 + idx] : -1;
        }
        producers_sync();
      }
      // per-offset presence over the 128 rows of the tile
      const uint32_t nbr_sa = s_base + L::NBR_OFF + tb * L::NBR_BYTES;
      const int64_t tile_row0 = row0 < n ? row0 : 0;
      {   // thread -> (row, quarter of the offsets): 512 threads share the 128 x k3 presence test
        const int row = tid & (TM - 1), kq = tid >> 7, kper = (k3 + 3) >> 2;
        const int k_lo = kq * kper, k_hi = min(k3, k_lo + kper);
        uint32_t local = 0;
        for (int k = k_lo; k < k_hi; ++k)
          local |= (ld_shared_s32(nbr_sa + (uint32_t)(row * k3 + k) * 4u) >= 0 ? 1u : 0u) << k;
        local = __reduce_or_sync(0xffffffffu, local);
        if (lane == 0 && local) atomicOr(&masks[i & 3], local);
      }
      producers_sync();
      const uint32_t mask = masks[i & 3];
      const int per = c_in >= KBLK ? c_in / KBLK : 1;   // 1, 2, 4, 8 on the PTv3 path; 3 / 6 in SpUNet's decoder
      const int per_shift = 31 - __clz(per);
      const bool per_pow2 = (per & (per - 1)) == 0;
      uint32_t blk_mask;  // bit j = K-block group j is needed (offset j for c_in >= 64, offset pair j for c_in == 32)
      if (c_in >= KBLK) {
        blk_mask = mask;
      } else {
        blk_mask = 0;
        for (int kb = 0; kb < n_kb_total; ++kb) blk_mask |= (((mask >> (2 * kb)) & 3u) ? 1u : 0u) << kb;
      }
      const int n_kb = __popc(blk_mask) * per;
      // every gather warp stores the same value before it issues (and arrives for) any K-block of this tile, so the
      // MMA warp, which reads it after the tile's first full barrier, always sees it
      if (lane == 0) tile_nkb[i & 15] = (uint32_t)n_kb;
      __syncwarp();
      // this group's K-blocks of the tile: done = first, first + 4, ...  (done counts K-blocks in mask-bit order)
      int popped = 0;   // set bits of blk_mask already dropped
      for (int done = (int)((uint32_t)(group + kGatherGroups - (int)(g0 % kGatherGroups)) % kGatherGroups); done < n_kb;
           done += kGatherGroups) {
        const int jidx = per_pow2 ? done >> per_shift : done / per;
        const int cb = done - jidx * per;
        while (popped < jidx) {
          blk_mask &= blk_mask - 1;
          ++popped;
        }
        const int j = __ffs(blk_mask) - 1;
        const uint32_t g = g0 + (uint32_t)done;
        const int s = (int)(g % STAGES);
        mbar_wait(empty_bar(s), ((g / STAGES) & 1u) ^ 1u);
        const int kb = j * per + cb;
        const uint32_t a_s = s_base + s * L::STAGE_BYTES;
        if (wg == 0 && elect_one_sync()) {
          mbar_expect_tx(full_bar(s), L::B_BYTES);
          bulk_g2s(a_s + A_BYTES, wtile + (int64_t)kb * L::B_BYTES, L::B_BYTES, full_bar(s));
        }
        int k, ch;
        if (c_in >= KBLK) {
          k = j;
          ch = cb * KBLK + c * 8;
        } else {
          k = 2 * j + (c >> 2);
          ch = (c & 3) * 8;
        }
        const bool k_ok = k < k3;
        const char* in_ch = reinterpret_cast<const char*>(in) + ch * 2;
        const uint32_t np = nbr_sa + (uint32_t)((wg * kRowsPerWarp + rr) * k3 + k) * 4u;
        const uint32_t a_w = a_s + (uint32_t)(wg * (kRowsPerWarp / 8)) * 1024u;
        int src[kSlots];
#pragma unroll
        for (int u = 0; u < kSlots; ++u) src[u] = k_ok ? ld_shared_s32(np + (uint32_t)u * nbr_step) : -1;
        if constexpr (ASYNC) {
          // every slot is a cp.async (absent neighbours: zero source bytes = zero fill); the mbarrier arrival is
          // triggered by the copies themselves, so the warp moves on without waiting
#pragma unroll
          for (int u = 0; u < kSlots; ++u) {
            const uint32_t dst = a_w + (uint32_t)(u >> 1) * 1024u + ((u & 1) ? off_odd : off_even);
            const bool ok = src[u] >= 0;
            const uint32_t r = ok ? (uint32_t)src[u] : (uint32_t)tile_row0;   // zero-byte copies aim at a row of this tile
            cp_async16(dst, in_ch + (uint64_t)r * row_bytes, ok ? 16u : 0u);
          }
          cp_async_mbar_arrive_noinc(full_bar(s));
        } else {
#pragma unroll
          for (int u = 0; u < kSlots; ++u) {
            const uint32_t dst = a_w + (uint32_t)(u >> 1) * 1024u + ((u & 1) ? off_odd : off_even);
            if (src[u] >= 0)   // absent neighbours are zeroed with a plain shared store
              cp_async16(dst, in_ch + (uint64_t)(uint32_t)src[u] * row_bytes, 16u);
            else
              st_shared_zero16(dst);
          }
          cp_async_commit();
          // blocking hand-over: the other three groups keep their K-blocks in flight meanwhile, and a warp never
          // holds an un-arrived stage at a barrier or while it waits for another stage
          cp_async_wait<0>();
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) mbar_arrive(full_bar(s));
        }
      }
      g0 += (uint32_t)n_kb;
      producers_sync();   // everyone is done reading nbr_s[tb]: it may be refilled for tile i+2
    }
  } else if (warp == kMmaWarp) {
    // =========================== MMA issuer ===========================
    // The whole warp runs the loop (uniform control flow, values in uniform registers); one elected lane issues.
    // Issuing from inside `if (lane == 0)` instead makes nvcc wrap every tcgen05 instruction in an ELECT / R2UR /
    // BRA.U.ANY loop: ~135 instructions per K-block on a single thread, which bounded the whole kernel.
    // ONE elected thread runs the whole issue loop: inside `if (elect_one_sync())` nvcc knows a single lane is
    // active, keeps the ring state and descriptors in uniform registers and emits the tcgen05 instructions without
    // per-instruction ELECT / R2UR loops or per-K-block warp re-convergence.
    if (my_tiles > 0 && elect_one_sync()) {
      constexpr uint32_t idesc = idesc_bf16_f32(TM, NT);
      uint32_t s = 0, ph = 0;
      for (int i = 0; i < my_tiles; ++i) {
        const uint32_t ab = (uint32_t)i & 1u;
        const uint32_t d_tmem = tmem_base + ab * NT;
        mbar_wait(full_bar(s), ph);   // first K-block of the tile: tile_nkb[i] was written before it was issued
        fence_after_sync();
        const int n_kb = ld_shared_s32(s_base + L::MISC_OFF + 160 + 4u * ((uint32_t)i & 15u));
        mbar_wait(acc_free(ab), (((uint32_t)i >> 1) & 1u) ^ 1u);   // accumulator drained by the epilogue (tile i-2)
        fence_after_sync();
        for (int kbi = 0; kbi < n_kb; ++kbi) {
          if (kbi > 0) {
            mbar_wait(full_bar(s), ph);
            fence_after_sync();
          }
          const uint32_t a_s = s_base + s * L::STAGE_BYTES;
          const uint64_t a_desc = smem_desc_sw128(a_s);
          const uint64_t b_desc = smem_desc_sw128(a_s + A_BYTES);
#pragma unroll
          for (int kk = 0; kk < KBLK / 16; ++kk)
            mma_bf16_ss(d_tmem, a_desc + 2 * kk, b_desc + 2 * kk, idesc, (kbi > 0 || kk > 0) ? 1u : 0u);
          mma_commit(empty_bar(s));
          if (++s == STAGES) {
            s = 0;
            ph ^= 1u;
          }
        }
        mma_commit(acc_full(ab));
      }
    }
    __syncwarp();
  } else {
    // =========================== epilogue: TMEM -> +bias -> bf16 rows ===========================
    const int q4 = warp & 3;   // TMEM lane quarter this warp may access
    for (int i = 0; i < my_tiles; ++i) {
      const uint32_t ab = (uint32_t)i & 1u;
      mbar_wait_sleep(acc_full(ab), (uint32_t)(i >> 1) & 1u, 128);
      fence_after_sync();
      const int64_t row = (int64_t)(blockIdx.x + (int64_t)i * gridDim.x) * TM + q4 * 32 + lane;
#pragma unroll 1
      for (int cb = 0; cb < NT; cb += 32) {
        uint32_t v[32];
        tmem_ld32(tmem_base + ((uint32_t)(q4 * 32) << 16) + ab * NT + (uint32_t)cb, v);
        tmem_ld_wait();
        if (cb + 32 >= NT) {   // last read of this accumulator: hand it back to the MMA warp
          fence_before_sync();
          mbar_arrive(acc_free(ab));
        }
        if (row < n) {
          const int co0 = nt_tile * NT + cb;
          uint32_t pk[16];
#pragma unroll
          for (int q = 0; q < 16; ++q) {
            float a = __uint_as_float(v[2 * q]) + (bias ? bias[co0 + 2 * q] : 0.f);
            float b = __uint_as_float(v[2 * q + 1]) + (bias ? bias[co0 + 2 * q + 1] : 0.f);
            __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
            pk[q] = *reinterpret_cast<uint32_t*>(&h);
          }
          // two 256-bit stores: a whole 32-byte sector per lane and instruction (16-byte stores touch every sector twice)
          uint32_t lo[8], hi[8];
#pragma unroll
          for (int q = 0; q < 8; ++q) { lo[q] = pk[q]; hi[q] = pk[8 + q]; }
          stg256_u32(out + row * c_out + co0, lo);
          stg256_u32(out + row * c_out + co0 + 16, hi);
        }
      }
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == kMmaWarp) tmem_dealloc<2 * NT>(tmem_base);
}

template <int NT, int STAGES>
static int launch_conv(const __nv_bfloat16* in, const __nv_bfloat16* packed, const float* bias, const int32_t* nbr,
                       int64_t n, int c_in, int c_out, int k3, __nv_bfloat16* out, cudaStream_t st) {
  using L = ConvSmem<NT, STAGES>;
  static_assert(STAGES <= 8, "ring");
  static_assert(L::DYN_BYTES <= 232448, "shared memory budget");
  static bool configured_dev[kMaxDeviceSlots] = {};
  bool& configured = configured_dev[device_slot()];
  if (!configured) {
    if (const char* e = getenv("PTV3_CONV_ASYNC")) g_conv_async = atoi(e);
    cudaFuncSetAttribute(subm_conv_tc_kernel<NT, STAGES, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         L::DYN_BYTES);
    cudaFuncSetAttribute(subm_conv_tc_kernel<NT, STAGES, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         L::DYN_BYTES);
    configured = true;
  }
  const int n_tiles = (int)cdiv(n, TM);
  const int ny = c_out / NT;
  int gx = kNumSMs / ny;
  if (gx > n_tiles) gx = n_tiles;
  dim3 grid((unsigned)gx, (unsigned)ny);
  if (g_conv_async & 1)
    subm_conv_tc_kernel<NT, STAGES, true><<<grid, kConvThreads, L::DYN_BYTES, st>>>(
        in, packed, bias, nbr, n, c_in, c_out, k3, n_kblocks(c_in, k3), n_tiles, out);
  else
    subm_conv_tc_kernel<NT, STAGES, false><<<grid, kConvThreads, L::DYN_BYTES, st>>>(
        in, packed, bias, nbr, n, c_in, c_out, k3, n_kblocks(c_in, k3), n_tiles, out);
  return check_launch();
}

int subm_pack_weight_tc(const __nv_bfloat16* w, int c_in, int c_out, int k3, __nv_bfloat16* packed, cudaStream_t st) {
  if (!subm_conv_tc_supported(c_in, c_out, k3)) return PTV3_EUNSUPPORTED;
  const int n_kb = n_kblocks(c_in, k3), nt = pick_nt(c_out);
  int64_t chunks = (int64_t)n_kb * c_out * 8;
  subm_pack_weight_kernel<<<grid_for(chunks, 256), 256, 0, st>>>(w, c_in, c_out, k3, nt, n_kb, packed);
  return check_launch();
}

int subm_conv_tc_packed(const __nv_bfloat16* in, const __nv_bfloat16* packed, const float* bias, const int32_t* nbr,
                        int64_t n, int c_in, int c_out, int k3, __nv_bfloat16* out, cudaStream_t st) {
  if (!subm_conv_tc_supported(c_in, c_out, k3)) return PTV3_EUNSUPPORTED;
  switch (pick_nt(c_out)) {
    case 32: return launch_conv<32, 8>(in, packed, bias, nbr, n, c_in, c_out, k3, out, st);
    case 64: return launch_conv<64, 7>(in, packed, bias, nbr, n, c_in, c_out, k3, out, st);
    case 128: return launch_conv<128, 5>(in, packed, bias, nbr, n, c_in, c_out, k3, out, st);
    default: return launch_conv<256, 4>(in, packed, bias, nbr, n, c_in, c_out, k3, out, st);
  }
}

int subm_conv_tc(const __nv_bfloat16* in, const __nv_bfloat16* w, const float* bias, const int32_t* nbr, int64_t n,
                 int c_in, int c_out, int k3, __nv_bfloat16* out, void* ws, size_t ws_bytes, cudaStream_t st) {
  if (ws == nullptr || ws_bytes < subm_conv_tc_workspace(n, c_in, c_out, k3)) return PTV3_EWORKSPACE;
  __nv_bfloat16* packed = reinterpret_cast<__nv_bfloat16*>(ws);
  int rc = subm_pack_weight_tc(w, c_in, c_out, k3, packed, st);
  if (rc != PTV3_OK) return rc;
  return subm_conv_tc_packed(in, packed, bias, nbr, n, c_in, c_out, k3, out, st);
}

}  // namespace ptv3
