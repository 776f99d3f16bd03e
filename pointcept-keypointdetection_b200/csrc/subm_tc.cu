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
#include "common.cuh"
#include "tc05.cuh"

namespace ptv3 {

using namespace tc05;

constexpr int TM = 128;          // output rows per CTA (UMMA M)
constexpr int KBLK = 64;         // bf16 per K-block (one 128-byte swizzle row)
constexpr int A_BYTES = TM * KBLK * 2;
constexpr int MAX_K3 = 27;
constexpr int PRODUCERS = 128;
constexpr int LAG = 2;           // cp.async groups in flight per producer thread

static inline int n_kblocks(int c_in, int k3) { return (k3 * c_in + KBLK - 1) / KBLK; }
static inline int pick_nt(int c_out) { return c_out >= 256 ? 256 : c_out; }

bool subm_conv_tc_supported(int c_in, int c_out, int k3) {
  bool cin_ok = c_in == 32 || (c_in % 64 == 0 && c_in >= 64 && c_in <= 512);
  bool cout_ok = c_out == 32 || c_out == 64 || c_out == 128 || c_out == 256 || c_out == 512;
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
  static constexpr int NBR_BYTES = TM * MAX_K3 * 4;
  static constexpr int LIST_OFF = NBR_OFF + NBR_BYTES;          // kb_list: up to 216 entries (uint16)
  static constexpr int LIST_BYTES = 512;
  static constexpr int BAR_OFF = LIST_OFF + LIST_BYTES;          // full[STAGES], empty[STAGES], accum, then misc
  static constexpr int TOTAL = BAR_OFF + 256;
  static constexpr int DYN_BYTES = TOTAL + 1024;                 // slack for the 1024-byte alignment
};

template <int NT, int STAGES>
__global__ void __launch_bounds__(160) subm_conv_tc_kernel(const __nv_bfloat16* __restrict__ in,
                                                           const __nv_bfloat16* __restrict__ packed,
                                                           const float* __restrict__ bias,
                                                           const int32_t* __restrict__ nbr, int64_t n, int c_in,
                                                           int c_out, int k3, int n_kb_total,
                                                           __nv_bfloat16* __restrict__ out) {
  using L = ConvSmem<NT, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t s_base = smem_u32(smem);
  int32_t* nbr_s = reinterpret_cast<int32_t*>(smem + L::NBR_OFF);
  uint16_t* kb_list = reinterpret_cast<uint16_t*>(smem + L::LIST_OFF);
  const uint32_t bar_base = s_base + L::BAR_OFF;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  const uint32_t accum_bar = bar_base + 8u * (2 * STAGES);
  volatile uint32_t* misc = reinterpret_cast<volatile uint32_t*>(smem + L::BAR_OFF + 8 * (2 * STAGES + 1));
  // misc[0] = tmem base, misc[1] = number of K-blocks in this tile, misc[2..] = per-offset presence flags (27)

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t row0 = (int64_t)blockIdx.x * TM;
  const int nt_tile = blockIdx.y;

  if (tid < 32) misc[2 + tid] = 0;
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), PRODUCERS);
      mbar_init(empty_bar(s), 1);
    }
    mbar_init(accum_bar, 1);
    mbar_fence_init();
  }
  if (warp == 4) tmem_alloc<NT>(smem_u32((const void*)&misc[0]));
  __syncthreads();

  // ---- neighbour rows of this tile -> shared memory, per-offset presence flags
  for (int idx = tid; idx < TM * k3; idx += blockDim.x) {
    int r = idx / k3, k = idx - r * k3;
    int64_t row = row0 + r;
    int v = row < n ? nbr[row * k3 + k] : -1;
    nbr_s[idx] = v;
    if (v >= 0) misc[2 + k] = 1;
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  if (tid == 0) {
    int cnt = 0;
    if (c_in >= KBLK) {
      const int per = c_in / KBLK;
      for (int k = 0; k < k3; ++k)
        if (misc[2 + k])
          for (int cb = 0; cb < per; ++cb) kb_list[cnt++] = (uint16_t)(k * per + cb);
    } else {  // c_in == 32: two offsets per K-block
      for (int kb = 0; kb < n_kb_total; ++kb) {
        int k0 = 2 * kb, k1 = 2 * kb + 1;
        if (misc[2 + k0] || (k1 < k3 && misc[2 + k1])) kb_list[cnt++] = (uint16_t)kb;
      }
    }
    misc[1] = (uint32_t)cnt;
  }
  __syncthreads();
  const int n_kb = (int)misc[1];
  const uint32_t tmem_base = misc[0];

  if (warp < 4) {
    // =========================== producers: gather A, bulk-copy B ===========================
    const int c = tid & 7, r0 = tid >> 3;
    const char* wtile = reinterpret_cast<const char*>(packed) + (int64_t)nt_tile * n_kb_total * L::B_BYTES;
    for (int it = 0; it < n_kb; ++it) {
      const int s = it % STAGES;
      const uint32_t ph = (uint32_t)(it / STAGES) & 1u;
      mbar_wait(empty_bar(s), ph ^ 1u);
      const int kb = kb_list[it];
      const uint32_t a_s = s_base + s * L::STAGE_BYTES;
      if (tid == 0) {
        mbar_expect_tx(full_bar(s), L::B_BYTES);
        bulk_g2s(a_s + A_BYTES, wtile + (int64_t)kb * L::B_BYTES, L::B_BYTES, full_bar(s));
      }
      int k, ch;
      if (c_in >= KBLK) {
        const int per = c_in / KBLK;
        k = kb / per;
        ch = (kb - k * per) * KBLK + c * 8;
      } else {
        k = 2 * kb + (c >> 2);
        ch = (c & 3) * 8;
      }
      const bool k_ok = k < k3;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int r = r0 + 16 * j;
        const int src = k_ok ? nbr_s[r * k3 + k] : -1;
        const __nv_bfloat16* g = in + (src >= 0 ? (int64_t)src * c_in + ch : 0);
        const uint32_t dst = a_s + (r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4);
        cp_async16(dst, g, src >= 0 ? 16u : 0u);
      }
      cp_async_commit();
      if (it >= LAG) {
        cp_async_wait<LAG>();
        fence_proxy_async();
        mbar_arrive(full_bar((it - LAG) % STAGES));
      }
    }
    cp_async_wait<0>();
    fence_proxy_async();
    for (int it = (n_kb > LAG ? n_kb - LAG : 0); it < n_kb; ++it) mbar_arrive(full_bar(it % STAGES));

    // =========================== epilogue: TMEM -> bf16 rows ===========================
    mbar_wait(accum_bar, 0);
    fence_after_sync();
    const int r = warp * 32 + lane;
    const int64_t row = row0 + r;
#pragma unroll 1
    for (int cb = 0; cb < NT; cb += 32) {
      uint32_t v[32];
      tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)cb, v);
      tmem_ld_wait();
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
        uint4* dst = reinterpret_cast<uint4*>(out + row * c_out + co0);
#pragma unroll
        for (int q = 0; q < 4; ++q) dst[q] = make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
      }
    }
  } else if (lane == 0) {
    // =========================== MMA issuer (one thread) ===========================
    constexpr uint32_t idesc = idesc_bf16_f32(TM, NT);
    for (int it = 0; it < n_kb; ++it) {
      const int s = it % STAGES;
      const uint32_t ph = (uint32_t)(it / STAGES) & 1u;
      mbar_wait(full_bar(s), ph);
      fence_after_sync();
      const uint32_t a_s = s_base + s * L::STAGE_BYTES;
      const uint64_t a_desc = smem_desc_sw128(a_s);
      const uint64_t b_desc = smem_desc_sw128(a_s + A_BYTES);
#pragma unroll
      for (int kk = 0; kk < KBLK / 16; ++kk)   // +32 bytes per K=16 step -> +2 in the (addr >> 4) field
        mma_bf16_ss(tmem_base, a_desc + 2 * kk, b_desc + 2 * kk, idesc, (it | kk) ? 1u : 0u);
      mma_commit(empty_bar(s));
    }
    mma_commit(accum_bar);
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 4) tmem_dealloc<NT>(tmem_base);
}

template <int NT, int STAGES>
static int launch_conv(const __nv_bfloat16* in, const __nv_bfloat16* packed, const float* bias, const int32_t* nbr,
                       int64_t n, int c_in, int c_out, int k3, __nv_bfloat16* out, cudaStream_t st) {
  using L = ConvSmem<NT, STAGES>;
  static bool configured = false;
  if (!configured) {
    cudaFuncSetAttribute(subm_conv_tc_kernel<NT, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::DYN_BYTES);
    configured = true;
  }
  dim3 grid((unsigned)cdiv(n, TM), (unsigned)(c_out / NT));
  subm_conv_tc_kernel<NT, STAGES><<<grid, 160, L::DYN_BYTES, st>>>(in, packed, bias, nbr, n, c_in, c_out, k3,
                                                                    n_kblocks(c_in, k3), out);
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
    case 32: return launch_conv<32, 4>(in, packed, bias, nbr, n, c_in, c_out, k3, out, st);
    case 64: return launch_conv<64, 4>(in, packed, bias, nbr, n, c_in, c_out, k3, out, st);
    case 128: return launch_conv<128, 3>(in, packed, bias, nbr, n, c_in, c_out, k3, out, st);
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
