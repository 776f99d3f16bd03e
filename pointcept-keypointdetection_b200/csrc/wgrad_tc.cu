// SubMConv3d weight gradient, fast paths (the generic CUDA-core kernel is csrc/backward.cu):
//
//  (1) bf16, C_in and C_out multiples of the tile (32 or 64): one CTA per (kernel offset k, row chunk, 64x64 or 32x32
//      tile of dW[k]).  The CTA first compacts the chunk's PRESENT (row, neighbour) pairs of offset k into shared
//      memory (at stage 0 only ~6 of 27 offsets are occupied per site), then walks them 64 at a time: dY rows and the
//      gathered X rows arrive by cp.async (double buffered), both are used through ldmatrix.trans as the A (m = o,
//      k = row) and B (k = row, n = c) operands of mma.sync m16n8k16, fp32 accumulators, one atomicAdd per element.
//  (2) fp32 stem (C_out = 32, C_in <= 8, k^3 = 125): row-major instead of offset-major -- a warp owns a row (lane =
//      output channel), loops over the row's present offsets and accumulates into a CTA-wide shared-memory image of
//      the whole dW (k^3 x C_in x 32 floats) with shared atomics; dY is read once instead of once per offset.
#include "common.cuh"

namespace ptv3 {

namespace {

__device__ __forceinline__ void mma16816(float c[4], const uint32_t a[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm4t(uint32_t r[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void cp16z(uint32_t dst, const void* src, bool valid) {
  int sz = valid ? 16 : 0;   // src-size 0 = zero fill
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz));
}

constexpr int kWgThreads = 128;
constexpr int kWgGroup = 64;       // pairs per pipeline step (4 k16 steps)
constexpr int kWgRows = 2048;      // rows per CTA (pair list <= 16 KB)

}  // namespace

template <int T>
__global__ void __launch_bounds__(kWgThreads) subm_wgrad_mma_kernel(const __nv_bfloat16* __restrict__ x,
                                                                    const __nv_bfloat16* __restrict__ dy,
                                                                    const int32_t* __restrict__ nbr, int64_t n, int k3,
                                                                    int cin, int cout, float* __restrict__ dw) {
  constexpr int PITCH = T * 2 + 16;            // bytes; +16 keeps the 8 rows of an ldmatrix on different banks
  constexpr int TILE = kWgGroup * PITCH;       // one operand tile
  constexpr int MT = T / 32, NT = T / 16;      // m16 tiles / n8 blocks per warp (warp tile T/2 x T/2)
  constexpr int PIECES = T * 2 / 16;           // 16-byte pieces per row
  extern __shared__ __align__(128) unsigned char smem_raw[];
  int2* pairs = reinterpret_cast<int2*>(smem_raw);                       // (row, neighbour)
  const uint32_t tiles = (uint32_t)__cvta_generic_to_shared(smem_raw + kWgRows * 8);   // [2][dY tile, X tile]
  __shared__ int n_pairs_s;
  const int k = blockIdx.x;
  const int n_ct = cin / T;
  const int o0 = ((int)blockIdx.z / n_ct) * T, c0 = ((int)blockIdx.z % n_ct) * T;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t r_begin = (int64_t)blockIdx.y * kWgRows;
  const int rows = (int)((n - r_begin < kWgRows) ? (n - r_begin) : kWgRows);
  if (tid == 0) n_pairs_s = 0;
  __syncthreads();
  // ---- compaction of the present pairs (order inside the chunk does not matter for a sum) ----
  for (int r = tid; r < kWgRows; r += kWgThreads) {
    int j = (r < rows) ? nbr[(r_begin + r) * k3 + k] : -1;
    unsigned m = __ballot_sync(0xffffffffu, j >= 0);
    int base = 0;
    if (lane == 0 && m) base = atomicAdd(&n_pairs_s, __popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (j >= 0) pairs[base + __popc(m & ((1u << lane) - 1))] = make_int2((int)r, j);
  }
  __syncthreads();
  const int n_pairs = n_pairs_s;
  if (n_pairs == 0) return;
  const int n_groups = (n_pairs + kWgGroup - 1) / kWgGroup;

  auto issue = [&](int grp, int buf) {
    const uint32_t dyt = tiles + (uint32_t)(buf * 2 * TILE), xt = dyt + TILE;
    for (int e = tid; e < kWgGroup * PIECES; e += kWgThreads) {
      const int p = e / PIECES, piece = e % PIECES;
      const int pi = grp * kWgGroup + p;
      const bool valid = pi < n_pairs;
      const int2 pr = valid ? pairs[pi] : make_int2(0, 0);
      cp16z(dyt + (uint32_t)(p * PITCH + piece * 16), dy + (r_begin + pr.x) * cout + o0 + piece * 8, valid);
      cp16z(xt + (uint32_t)(p * PITCH + piece * 16), x + (int64_t)pr.y * cin + c0 + piece * 8, valid);
    }
    asm volatile("cp.async.commit_group;");
  };

  float acc[MT][NT][4];
#pragma unroll
  for (int a = 0; a < MT; ++a)
#pragma unroll
    for (int b = 0; b < NT; ++b) acc[a][b][0] = acc[a][b][1] = acc[a][b][2] = acc[a][b][3] = 0.f;
  const int wo = (warp >> 1) * (T / 2), wc = (warp & 1) * (T / 2);   // warp tile origin inside the CTA tile
  // ldmatrix.trans lane offsets (storage [row][channel], pitch PITCH):
  //  A (m = o, k = row): matrices {rows 0-7 o 0-7, rows 0-7 o 8-15, rows 8-15 o 0-7, rows 8-15 o 8-15}
  //  B (k = row, n = c): matrices {rows 0-7 c 0-7, rows 8-15 c 0-7, rows 0-7 c 8-15, rows 8-15 c 8-15}
  const uint32_t a_off = (uint32_t)((((lane >> 4) << 3) + (lane & 7)) * PITCH + ((lane >> 3) & 1) * 16);
  const uint32_t b_off = (uint32_t)(((((lane >> 3) & 1) << 3) + (lane & 7)) * PITCH + (lane >> 4) * 16);

  issue(0, 0);
  for (int grp = 0; grp < n_groups; ++grp) {
    const int buf = grp & 1;
    if (grp + 1 < n_groups) {
      issue(grp + 1, buf ^ 1);
      asm volatile("cp.async.wait_group 1;");
    } else {
      asm volatile("cp.async.wait_group 0;");
    }
    __syncthreads();
    const uint32_t dyt = tiles + (uint32_t)(buf * 2 * TILE), xt = dyt + TILE;
#pragma unroll
    for (int ks = 0; ks < kWgGroup / 16; ++ks) {
      uint32_t af[MT][4], bf[NT / 2][4];
#pragma unroll
      for (int a = 0; a < MT; ++a) ldsm4t(af[a], dyt + (uint32_t)(ks * 16 * PITCH + (wo + a * 16) * 2) + a_off);
#pragma unroll
      for (int b = 0; b < NT / 2; ++b) ldsm4t(bf[b], xt + (uint32_t)(ks * 16 * PITCH + (wc + b * 16) * 2) + b_off);
#pragma unroll
      for (int a = 0; a < MT; ++a)
#pragma unroll
        for (int b = 0; b < NT / 2; ++b) {
          mma16816(acc[a][2 * b], af[a], bf[b][0], bf[b][1]);
          mma16816(acc[a][2 * b + 1], af[a], bf[b][2], bf[b][3]);
        }
    }
    __syncthreads();   // the buffer is re-filled by the next iteration's issue
  }
  const int g = lane >> 2, t = lane & 3;
  float* base = dw + (int64_t)k * cout * cin;
#pragma unroll
  for (int a = 0; a < MT; ++a)
#pragma unroll
    for (int b = 0; b < NT; ++b) {
      const int o = o0 + wo + a * 16 + g, c = c0 + wc + b * 8 + 2 * t;
      atomicAdd(base + (int64_t)o * cin + c, acc[a][b][0]);
      atomicAdd(base + (int64_t)o * cin + c + 1, acc[a][b][1]);
      atomicAdd(base + (int64_t)(o + 8) * cin + c, acc[a][b][2]);
      atomicAdd(base + (int64_t)(o + 8) * cin + c + 1, acc[a][b][3]);
    }
}

// ---- (2) fp32 stem: cout == 32, cin <= 8 ----
constexpr int kStemWgThreads = 256;

__global__ void __launch_bounds__(kStemWgThreads) subm_wgrad_stem_kernel(const float* __restrict__ x,
                                                                         const float* __restrict__ dy,
                                                                         const int32_t* __restrict__ nbr, int64_t n, int k3,
                                                                         int cin, int64_t rows_per_cta,
                                                                         float* __restrict__ dw) {
  extern __shared__ float acc_s[];   // [k3][cin][32]
  const int total = k3 * cin * 32;
  for (int e = threadIdx.x; e < total; e += kStemWgThreads) acc_s[e] = 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t r_begin = (int64_t)blockIdx.x * rows_per_cta;
  const int64_t r_end = (r_begin + rows_per_cta < n) ? r_begin + rows_per_cta : n;
  for (int64_t i = r_begin + warp; i < r_end; i += kStemWgThreads / 32) {
    const float g = dy[i * 32 + lane];
    const int32_t* nb = nbr + i * k3;
    for (int kb = 0; kb < k3; kb += 32) {
      const int j = (kb + lane < k3) ? nb[kb + lane] : -1;
      unsigned m = __ballot_sync(0xffffffffu, j >= 0);
      while (m) {
        const int src = __ffs(m) - 1;
        m &= m - 1;
        const int jj = __shfl_sync(0xffffffffu, j, src);
        const float* xr = x + (int64_t)jj * cin;
        float* a = acc_s + ((kb + src) * cin) * 32 + lane;
        for (int c = 0; c < cin; ++c) atomicAdd(a + c * 32, g * xr[c]);
      }
    }
  }
  __syncthreads();
  // dW layout (k3, cout = 32, cin)
  for (int e = threadIdx.x; e < total; e += kStemWgThreads) {
    const float v = acc_s[e];
    if (v != 0.f) {
      const int o = e & 31, c = (e >> 5) % cin, k = (e >> 5) / cin;
      atomicAdd(dw + ((int64_t)k * 32 + o) * cin + c, v);
    }
  }
}

// returns PTV3_EUNSUPPORTED when no fast path applies
int subm_wgrad_fast(const void* in, const void* dout, const int32_t* nbr, int64_t n, int c_in, int c_out, int k3, int dtype,
                    float* dweight, cudaStream_t st) {
  if (dtype == PTV3_BF16 && ((c_in % 64 == 0 && c_out % 64 == 0) || (c_in % 32 == 0 && c_out % 32 == 0))) {
    const int T = (c_in % 64 == 0 && c_out % 64 == 0) ? 64 : 32;
    const int64_t tiles = (int64_t)(c_in / T) * (c_out / T);
    const int64_t chunks = cdiv(n, kWgRows);
    if (tiles > 65535 || chunks > 65535) return PTV3_EUNSUPPORTED;
    dim3 grid((unsigned)k3, (unsigned)chunks, (unsigned)tiles);
    const size_t smem = (size_t)kWgRows * 8 + (size_t)4 * kWgGroup * (T * 2 + 16);
    if (T == 64) {
      static bool configured_dev[kMaxDeviceSlots] = {};
      bool& configured = configured_dev[device_slot()];
      if (!configured) {
        if (cudaFuncSetAttribute(subm_wgrad_mma_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) !=
            cudaSuccess)
          return PTV3_ELAUNCH;
        configured = true;
      }
      subm_wgrad_mma_kernel<64><<<grid, kWgThreads, smem, st>>>((const __nv_bfloat16*)in, (const __nv_bfloat16*)dout, nbr,
                                                                n, k3, c_in, c_out, dweight);
    } else {
      subm_wgrad_mma_kernel<32><<<grid, kWgThreads, smem, st>>>((const __nv_bfloat16*)in, (const __nv_bfloat16*)dout, nbr,
                                                                n, k3, c_in, c_out, dweight);
    }
    return check_launch();
  }
  if (dtype == PTV3_F32 && c_out == 32 && c_in <= 8) {
    const size_t smem = (size_t)k3 * c_in * 32 * 4;
    if (smem > 200 * 1024) return PTV3_EUNSUPPORTED;
    static size_t configured_dev2[kMaxDeviceSlots] = {};
    size_t& configured = configured_dev2[device_slot()];
    if (smem > configured) {
      if (cudaFuncSetAttribute(subm_wgrad_stem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) !=
          cudaSuccess)
        return PTV3_ELAUNCH;
      configured = smem;
    }
    int64_t ctas = kNumSMs * 2;
    int64_t rows_per_cta = cdiv(n, ctas);
    if (rows_per_cta < 64) rows_per_cta = 64;
    ctas = cdiv(n, rows_per_cta);
    subm_wgrad_stem_kernel<<<(unsigned)ctas, kStemWgThreads, smem, st>>>((const float*)in, (const float*)dout, nbr, n, k3,
                                                                         c_in, rows_per_cta, dweight);
    return check_launch();
  }
  return PTV3_EUNSUPPORTED;
}

}  // namespace ptv3
