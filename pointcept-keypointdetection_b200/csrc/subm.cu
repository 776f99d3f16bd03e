// A5: submanifold convolution -- hashed neighbour map + CUDA-core kernels
// (fp32 parity path, small-C_in stem).  The bf16 tensor-core kernel lives in
// subm_tc.cu.  Semantics: oracle/ops.py::subm_neighbor_table / subm_conv.
#include "common.cuh"

namespace ptv3 {

// ------------------------------ neighbour map ----------------------------------
constexpr uint64_t kEmptyKey = ~0ull;
constexpr int kCoordBias = 8;  // keeps shifted coordinates non-negative for |delta| <= 8

// sites the 64-bit key can hold: batch id in [0, 1024), coordinates in [0, 2^18 - 2 * kCoordBias) (the bias keeps
// neighbour offsets |delta| <= 8 non-negative and inside the field).  PRECONDITION of the backward (dX through the
// mirrored weights relies on j = nbr[i,k] <=> i = nbr[j,k^3-1-k]): sites are unique, which GridSample / pooling
// guarantee; duplicates are tolerated by the forward (the lowest row wins) but not by that symmetry.
__device__ __forceinline__ bool site_in_range(int b, int x, int y, int z) {
  constexpr unsigned kMax = (1u << 18) - 2u * kCoordBias;
  return (unsigned)b < 1024u && (unsigned)x < kMax && (unsigned)y < kMax && (unsigned)z < kMax;
}
__device__ __forceinline__ uint64_t pack_site(int b, int x, int y, int z) {
  // 10 bits batch | 3 x 18 bits coordinates
  return ((uint64_t)(uint32_t)b << 54) | ((uint64_t)(uint32_t)(x + kCoordBias) << 36) |
         ((uint64_t)(uint32_t)(y + kCoordBias) << 18) | (uint64_t)(uint32_t)(z + kCoordBias);
}

__device__ __forceinline__ uint32_t hash_site(uint64_t k) {
  k ^= k >> 33;
  k *= 0xff51afd7ed558ccdull;
  k ^= k >> 33;
  k *= 0xc4ceb9fe1a85ec53ull;
  k ^= k >> 33;
  return (uint32_t)k;
}

__global__ void __launch_bounds__(256) subm_insert_kernel(const int32_t* __restrict__ indices, int64_t n,
                                                          unsigned long long* __restrict__ keys,
                                                          int* __restrict__ vals, uint32_t mask,
                                                          int* __restrict__ err) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int4 s = reinterpret_cast<const int4*>(indices)[i];
    if (!site_in_range(s.x, s.y, s.z, s.w)) {   // would alias another site's key: leave it out and say so
      if (err) atomicOr(err, 2);
      continue;
    }
    uint64_t key = pack_site(s.x, s.y, s.z, s.w);
    uint32_t slot = hash_site(key) & mask;
    while (true) {
      unsigned long long prev = atomicCAS(&keys[slot], (unsigned long long)kEmptyKey, (unsigned long long)key);
      if (prev == kEmptyKey || prev == key) {
        atomicMin(&vals[slot], (int)i);  // duplicate sites: lowest row wins
        break;
      }
      slot = (slot + 1) & mask;
    }
  }
}

__global__ void __launch_bounds__(256) subm_lookup_kernel(const int32_t* __restrict__ indices, int64_t n, int ksize,
                                                          const unsigned long long* __restrict__ keys,
                                                          const int* __restrict__ vals, uint32_t mask,
                                                          int32_t* __restrict__ nbr) {
  const int k3 = ksize * ksize * ksize, r = ksize / 2;
  const int64_t total = n * k3;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    int64_t i = t / k3;
    int k = (int)(t - i * k3);
    int dz = k % ksize - r, dy = (k / ksize) % ksize - r, dx = k / (ksize * ksize) - r;
    int4 s = reinterpret_cast<const int4*>(indices)[i];
    int x = s.y + dx, y = s.z + dy, z = s.w + dz;
    int found = -1;
    if (x >= -kCoordBias && y >= -kCoordBias && z >= -kCoordBias) {
      uint64_t key = pack_site(s.x, x, y, z);
      uint32_t slot = hash_site(key) & mask;
      while (true) {
        unsigned long long cur = keys[slot];
        if (cur == key) {
          found = vals[slot];
          break;
        }
        if (cur == kEmptyKey) break;
        slot = (slot + 1) & mask;
      }
    }
    nbr[t] = found;
  }
}

static uint32_t table_capacity(int64_t n) {
  uint64_t cap = 1024;
  while (cap < (uint64_t)(2 * n)) cap <<= 1;
  return (uint32_t)cap;
}

// ------------------------------ tiled CUDA-core conv ---------------------------
constexpr int BM = 64, BN = 64, BK = 16;

template <typename T>
__global__ void __launch_bounds__(256) subm_conv_tiled_kernel(const T* __restrict__ in, const T* __restrict__ w,
                                                              const float* __restrict__ bias,
                                                              const int32_t* __restrict__ nbr, int64_t n, int cin,
                                                              int cout, int k3, T* __restrict__ out) {
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  __shared__ int nb[BM];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int64_t row0 = (int64_t)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k = 0; k < k3; ++k) {
    int has = 0;
    if (tid < BM) {
      int64_t r = row0 + tid;
      int v = r < n ? nbr[r * k3 + k] : -1;
      nb[tid] = v;
      has = v >= 0;
    }
    if (!__syncthreads_or(has)) continue;
    for (int c0 = 0; c0 < cin; c0 += BK) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        int idx = e * 256 + tid;  // 0..1023
        int m = idx >> 4, kk = idx & 15;
        int src = nb[m];
        int ci = c0 + kk;
        As[kk][m] = (src >= 0 && ci < cin) ? Elem<T>::load(in + (int64_t)src * cin + ci) : 0.f;
        int co = n0 + m;
        Bs[kk][m] = (co < cout && ci < cin) ? Elem<T>::load(w + ((int64_t)k * cout + co) * cin + ci) : 0.f;
      }
      __syncthreads();
#pragma unroll
      for (int kk = 0; kk < BK; ++kk) {
        float a[4], b[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
        for (int j = 0; j < 4; ++j) b[j] = Bs[kk][tx * 4 + j];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
      }
      __syncthreads();
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int64_t r = row0 + ty * 4 + i;
    if (r >= n) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int co = n0 + tx * 4 + j;
      if (co < cout) Elem<T>::store(out + r * cout + co, acc[i][j] + (bias ? bias[co] : 0.f));
    }
  }
}

// ------------------------------ small-C_in "row per warp" conv (stem) -----------
// lane = output channel (c_out <= 32*CPL), loops only over the occupied offsets.
template <typename T, int CPL>
__global__ void __launch_bounds__(256) subm_conv_rowwarp_kernel(const T* __restrict__ in, const T* __restrict__ w,
                                                                const float* __restrict__ bias,
                                                                const int32_t* __restrict__ nbr, int64_t n, int cin,
                                                                int cout, int k3, T* __restrict__ out) {
  extern __shared__ float ws[];  // [k3][cin][cout]
  for (int idx = threadIdx.x; idx < k3 * cin * cout; idx += blockDim.x) {
    int co = idx % cout, ci = (idx / cout) % cin, k = idx / (cout * cin);
    ws[idx] = Elem<T>::load(w + ((int64_t)k * cout + co) * cin + ci);
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t stride = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (; row < n; row += stride) {
    float acc[CPL];
#pragma unroll
    for (int q = 0; q < CPL; ++q) acc[q] = 0.f;
    const int32_t* nb = nbr + row * k3;
    for (int kb = 0; kb < k3; kb += 32) {
      int mine = (kb + lane < k3) ? nb[kb + lane] : -1;
      uint32_t present = __ballot_sync(0xffffffffu, mine >= 0);
      // every lane gathers the features of ITS neighbour (cin <= 8) so the random loads of a row overlap ...
      float xin[8];
#pragma unroll
      for (int ci = 0; ci < 8; ++ci)
        xin[ci] = (mine >= 0 && ci < cin) ? Elem<T>::load(in + (int64_t)mine * cin + ci) : 0.f;
      // ... then the occupied offsets are consumed one by one through shuffles
      while (present) {
        int b = __ffs(present) - 1;
        present &= present - 1;
        int k = kb + b;
#pragma unroll
        for (int ci = 0; ci < 8; ++ci) {
          if (ci < cin) {
            float x = __shfl_sync(0xffffffffu, xin[ci], b);
#pragma unroll
            for (int q = 0; q < CPL; ++q) {
              int co = q * 32 + lane;
              if (co < cout) acc[q] = fmaf(x, ws[(k * cin + ci) * cout + co], acc[q]);
            }
          }
        }
      }
    }
#pragma unroll
    for (int q = 0; q < CPL; ++q) {
      int co = q * 32 + lane;
      if (co < cout) Elem<T>::store(out + row * cout + co, acc[q] + (bias ? bias[co] : 0.f));
    }
  }
}


// ------------------------------ stem specialisation: C_in = 4, C_out = 32 -------------------
// lane = output channel; weights as float4 [k][co] in shared memory (one LDS.128 per neighbour); the neighbour ids of
// the whole row (<= 128 offsets) and the lanes' own neighbour features are loaded up front so their latencies overlap.
template <typename T>
__device__ __forceinline__ float4 load_feat4(const T* p);
template <>
__device__ __forceinline__ float4 load_feat4<float>(const float* p) { return *reinterpret_cast<const float4*>(p); }
template <>
__device__ __forceinline__ float4 load_feat4<__nv_bfloat16>(const __nv_bfloat16* p) {
  uint2 u = *reinterpret_cast<const uint2*>(p);
  return make_float4(__uint_as_float(u.x << 16), __uint_as_float(u.x & 0xffff0000u), __uint_as_float(u.y << 16),
                     __uint_as_float(u.y & 0xffff0000u));
}

template <typename T>
__global__ void __launch_bounds__(256) subm_conv_stem4_kernel(const T* __restrict__ in, const T* __restrict__ w,
                                                              const float* __restrict__ bias,
                                                              const int32_t* __restrict__ nbr, int64_t n, int k3,
                                                              T* __restrict__ out) {
  extern __shared__ float4 ws4[];  // [k3][32]
  for (int idx = threadIdx.x; idx < k3 * 32; idx += blockDim.x) {
    const int co = idx & 31, k = idx >> 5;
    const T* p = w + ((int64_t)k * 32 + co) * 4;
    ws4[idx] = make_float4(Elem<T>::load(p), Elem<T>::load(p + 1), Elem<T>::load(p + 2), Elem<T>::load(p + 3));
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t stride = (int64_t)gridDim.x * (blockDim.x >> 5);
  const float b = bias ? bias[lane] : 0.f;
  for (; row < n; row += stride) {
    const int32_t* nb = nbr + row * k3;
    int mine[4];
    float4 x[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) mine[q] = (q * 32 + lane < k3) ? nb[q * 32 + lane] : -1;
#pragma unroll
    for (int q = 0; q < 4; ++q)
      x[q] = mine[q] >= 0 ? load_feat4<T>(in + (int64_t)mine[q] * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
    float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      uint32_t present = __ballot_sync(0xffffffffu, mine[q] >= 0);
      while (present) {
        const int j = __ffs(present) - 1;
        present &= present - 1;
        const float4 wv = ws4[(q * 32 + j) * 32 + lane];
        acc0 = fmaf(__shfl_sync(0xffffffffu, x[q].x, j), wv.x, acc0);
        acc1 = fmaf(__shfl_sync(0xffffffffu, x[q].y, j), wv.y, acc1);
        acc0 = fmaf(__shfl_sync(0xffffffffu, x[q].z, j), wv.z, acc0);
        acc1 = fmaf(__shfl_sync(0xffffffffu, x[q].w, j), wv.w, acc1);
      }
    }
    Elem<T>::store(out + row * 32 + lane, acc0 + acc1 + b);
  }
}

// Stem (k = 5, C_in = 4, C_out = 32, fp32) with the neighbour search inside: the 125-column table of the stem is used
// by nothing else in inference, so writing it (583 MB at 1.17 M voxels) and reading it back is pure overhead.  A warp
// owns a row: lanes probe the hash table for 4 offsets each, gather their 16-byte input rows, then the present
// offsets are walked in the same order as subm_conv_stem4_kernel (bit-identical sums); lane = output channel.
// Epilogue: + bias (folded BatchNorm), optional exact GELU, fp32 and / or bf16 copies.
__global__ void __launch_bounds__(512, 2) subm_stem_fused_kernel(const float* __restrict__ in, const float* __restrict__ w,
                                                              const float* __restrict__ bias,
                                                              const int32_t* __restrict__ indices,
                                                              const unsigned long long* __restrict__ keys,
                                                              const int* __restrict__ vals, uint32_t mask, int64_t n,
                                                              int ksize, int act, float* __restrict__ out32,
                                                              __nv_bfloat16* __restrict__ out16) {
  extern __shared__ float4 ws4[];  // [k3][32]
  const int k3 = ksize * ksize * ksize, r = ksize / 2;
  for (int idx = threadIdx.x; idx < k3 * 32; idx += blockDim.x) {
    const int co = idx & 31, k = idx >> 5;
    const float* p = w + ((int64_t)k * 32 + co) * 4;
    ws4[idx] = make_float4(p[0], p[1], p[2], p[3]);
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t stride = (int64_t)gridDim.x * (blockDim.x >> 5);
  const float b = bias ? bias[lane] : 0.f;
  // this lane's four kernel offsets
  int dxs[4], dys[4], dzs[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int k = q * 32 + lane;
    dzs[q] = k % ksize - r;
    dys[q] = (k / ksize) % ksize - r;
    dxs[q] = k / (ksize * ksize) - r;
  }
  for (; row < n; row += stride) {
    const int4 s = reinterpret_cast<const int4*>(indices)[row];
    // the four first probes are issued together (independent loads), collisions are resolved afterwards
    int mine[4];
    uint64_t key[4];
    uint32_t slot[4];
    unsigned long long cur[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int x = s.y + dxs[q], y = s.z + dys[q], z = s.w + dzs[q];
      const bool ok = q * 32 + lane < k3 && x >= -kCoordBias && y >= -kCoordBias && z >= -kCoordBias;
      key[q] = ok ? pack_site(s.x, x, y, z) : kEmptyKey;
      slot[q] = hash_site(key[q]) & mask;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) cur[q] = key[q] != kEmptyKey ? keys[slot[q]] : (unsigned long long)kEmptyKey;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      while (cur[q] != key[q] && cur[q] != kEmptyKey) {
        slot[q] = (slot[q] + 1) & mask;
        cur[q] = keys[slot[q]];
      }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) mine[q] = (key[q] != kEmptyKey && cur[q] == key[q]) ? vals[slot[q]] : -1;
    float4 x[4];
#pragma unroll
    for (int q = 0; q < 4; ++q)
      x[q] = mine[q] >= 0 ? *reinterpret_cast<const float4*>(in + (int64_t)mine[q] * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
    float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      uint32_t present = __ballot_sync(0xffffffffu, mine[q] >= 0);
      while (present) {
        const int j = __ffs(present) - 1;
        present &= present - 1;
        const float4 wv = ws4[(q * 32 + j) * 32 + lane];
        acc0 = fmaf(__shfl_sync(0xffffffffu, x[q].x, j), wv.x, acc0);
        acc1 = fmaf(__shfl_sync(0xffffffffu, x[q].y, j), wv.y, acc1);
        acc0 = fmaf(__shfl_sync(0xffffffffu, x[q].z, j), wv.z, acc0);
        acc1 = fmaf(__shfl_sync(0xffffffffu, x[q].w, j), wv.w, acc1);
      }
    }
    float y = acc0 + acc1 + b;
    if (act == 1) y = 0.5f * y * (1.f + erff(y * 0.70710678118654752f));
    if (out32) out32[row * 32 + lane] = y;
    if (out16) out16[row * 32 + lane] = __float2bfloat16_rn(y);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Stem with a BLOCK-OCCUPANCY neighbour search (round 2).  subm_stem_fused_kernel spends its time on 125 random hash
// probes per voxel (one 32-byte sector each: ~5 KB of L2 -> L1 traffic per row).  The z-order serialization that
// Point.serialization() has already computed groups every 4 x 4 x 4 block of cells into one contiguous run of the
// sorted order (code >> 6 is the block, code & 63 the Morton cell inside it), so:
//   stem_blocks_kernel   one pass over the sorted codes: per run head -> hash insert of (block key -> run start, 64-bit
//                        occupancy mask); also writes the sorted rows as int32
//   subm_stem_z_kernel   a 5^3 window touches at most 2 x 2 x 2 blocks: lanes 0-7 probe one block each (8 probes per
//                        row instead of 125); a neighbour cell is present iff its mask bit is set and its row is
//                        rows_sorted[start + popcount(mask below the cell)] -- one dependent 4-byte load, only for
//                        the ~16 present neighbours.  Same walking order of the present offsets as the kernel above:
//                        bit-identical sums.
// Requires unique sites (GridSample / pooling guarantee it; a block whose run holds the same cell twice sets bit 8 of
// the device error word) and depth >= 2.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) stem_blocks_kernel(const int64_t* __restrict__ code, const int64_t* __restrict__ order,
                                                          int64_t n, unsigned long long* __restrict__ keys,
                                                          uint4* __restrict__ payload, uint32_t mask,
                                                          int32_t* __restrict__ rows_sorted, int* __restrict__ err) {
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = order[j];
    rows_sorted[j] = (int32_t)row;
    const uint64_t c = (uint64_t)code[row];
    const uint64_t kb = c >> 6;
    if (j > 0 && ((uint64_t)code[order[j - 1]] >> 6) == kb) continue;      // not a run head
    uint64_t m = 1ull << (c & 63);
    int cnt = 1;
    for (int64_t t = j + 1; t < n; ++t) {
      const uint64_t ct = (uint64_t)code[order[t]];
      if ((ct >> 6) != kb) break;
      m |= 1ull << (ct & 63);
      ++cnt;
    }
    if (cnt != __popcll(m) && err) atomicOr(err, 8);                        // duplicate sites: ranks would be wrong
    uint32_t slot = hash_site(kb) & mask;
    while (true) {
      const unsigned long long prev = atomicCAS(&keys[slot], (unsigned long long)kEmptyKey, (unsigned long long)kb);
      if (prev == kEmptyKey) break;      // block keys are unique among run heads (the codes are sorted)
      slot = (slot + 1) & mask;
    }
    payload[slot] = make_uint4((uint32_t)m, (uint32_t)(m >> 32), (uint32_t)j, (uint32_t)cnt);
  }
}

__device__ __forceinline__ uint64_t spread3_u64(uint64_t v) {
  v &= 0xffffull;
  v = (v | (v << 32)) & 0x001f00000000ffffull;
  v = (v | (v << 16)) & 0x001f0000ff0000ffull;
  v = (v | (v << 8)) & 0x100f00f00f00f00full;
  v = (v | (v << 4)) & 0x10c30c30c30c30c3ull;
  v = (v | (v << 2)) & 0x1249249249249249ull;
  return v;
}

__global__ void __launch_bounds__(512, 2) subm_stem_z_kernel(const float* __restrict__ in, const float* __restrict__ w,
                                                          const float* __restrict__ bias,
                                                          const int32_t* __restrict__ indices,
                                                          const unsigned long long* __restrict__ keys,
                                                          const uint4* __restrict__ payload, uint32_t mask,
                                                          const int32_t* __restrict__ rows_sorted, int depth, int64_t n,
                                                          int ksize, int act, float* __restrict__ out32,
                                                          __nv_bfloat16* __restrict__ out16) {
  extern __shared__ float4 ws4[];  // [k3][32] weights, then [warps][128] gathered input rows of the row in flight
  const int k3 = ksize * ksize * ksize, r = ksize / 2;
  float4* xs = ws4 + k3 * 32 + (threadIdx.x >> 5) * 128;
  for (int idx = threadIdx.x; idx < k3 * 32; idx += blockDim.x) {
    const int co = idx & 31, k = idx >> 5;
    const float* p = w + ((int64_t)k * 32 + co) * 4;
    ws4[idx] = make_float4(p[0], p[1], p[2], p[3]);
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t stride = (int64_t)gridDim.x * (blockDim.x >> 5);
  const float b = bias ? bias[lane] : 0.f;
  const int nblk_axis = 1 << (depth - 2);          // blocks per axis
  const int bshift = 3 * depth - 6;                // batch id sits above the block Morton code
  int dxs[4], dys[4], dzs[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int k = q * 32 + lane;
    dzs[q] = k % ksize - r;
    dys[q] = (k / ksize) % ksize - r;
    dxs[q] = k / (ksize * ksize) - r;
  }
  for (; row < n; row += stride) {
    const int4 s = reinterpret_cast<const int4*>(indices)[row];
    const int bx0 = (s.y - r) >> 2, by0 = (s.z - r) >> 2, bz0 = (s.w - r) >> 2;      // arithmetic shifts: -1 below the grid
    // lanes 0-7: one candidate block each
    uint32_t m_lo = 0, m_hi = 0, start = 0;
    if (lane < 8) {
      const int bx = (lane & 4) ? (s.y + r) >> 2 : bx0, by = (lane & 2) ? (s.z + r) >> 2 : by0,
                bz = (lane & 1) ? (s.w + r) >> 2 : bz0;
      if (bx >= 0 && by >= 0 && bz >= 0 && bx < nblk_axis && by < nblk_axis && bz < nblk_axis) {
        const uint64_t kb = ((uint64_t)(uint32_t)s.x << bshift) | (spread3_u64((uint32_t)bx) << 2) |
                            (spread3_u64((uint32_t)by) << 1) | spread3_u64((uint32_t)bz);
        uint32_t slot = hash_site(kb) & mask;
        unsigned long long cur = keys[slot];
        while (cur != kb && cur != kEmptyKey) {
          slot = (slot + 1) & mask;
          cur = keys[slot];
        }
        if (cur == kb) {
          const uint4 pl = payload[slot];
          m_lo = pl.x; m_hi = pl.y; start = pl.z;
        }
      }
    }
    int mine[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int x = s.y + dxs[q], y = s.z + dys[q], z = s.w + dzs[q];
      const int sel = (((x >> 2) != bx0) ? 4 : 0) | (((y >> 2) != by0) ? 2 : 0) | (((z >> 2) != bz0) ? 1 : 0);
      const uint32_t lo = __shfl_sync(0xffffffffu, m_lo, sel), hi = __shfl_sync(0xffffffffu, m_hi, sel);
      const uint32_t st = __shfl_sync(0xffffffffu, start, sel);
      const uint64_t m = ((uint64_t)hi << 32) | lo;
      const int cell = ((x & 2) << 4) | ((y & 2) << 3) | ((z & 2) << 2) | ((x & 1) << 2) | ((y & 1) << 1) | (z & 1);
      const bool present = q * 32 + lane < k3 && ((m >> cell) & 1ull);
      mine[q] = present ? rows_sorted[st + __popcll(m & ((1ull << cell) - 1ull))] : -1;
    }
    // the gathered input rows go through shared memory: one broadcast 16-byte load per present neighbour instead of
    // four shuffles (the walk below was shuffle-issue bound)
    __syncwarp();
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (mine[q] >= 0) xs[q * 32 + lane] = *reinterpret_cast<const float4*>(in + (int64_t)mine[q] * 4);
    __syncwarp();
    float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      uint32_t present = __ballot_sync(0xffffffffu, mine[q] >= 0);
      while (present) {
        const int j = __ffs(present) - 1;
        present &= present - 1;
        const float4 wv = ws4[(q * 32 + j) * 32 + lane];
        const float4 xv = xs[q * 32 + j];
        acc0 = fmaf(xv.x, wv.x, acc0);
        acc1 = fmaf(xv.y, wv.y, acc1);
        acc0 = fmaf(xv.z, wv.z, acc0);
        acc1 = fmaf(xv.w, wv.w, acc1);
      }
    }
    float y = acc0 + acc1 + b;
    if (act == 1) y = 0.5f * y * (1.f + erff(y * 0.70710678118654752f));
    if (out32) out32[row * 32 + lane] = y;
    if (out16) out16[row * 32 + lane] = __float2bfloat16_rn(y);
  }
}

// (N, k^3) neighbour table through the same block table (k = 3 or 5): 8 lanes per row probe one candidate block each,
// then share the k^3 offsets (lane l takes k = l, l + 8, ...).  Same table as subm_lookup_kernel, bit for bit.
__global__ void __launch_bounds__(256) subm_lookup_z_kernel(const int32_t* __restrict__ indices, int64_t n, int ksize,
                                                            const unsigned long long* __restrict__ keys,
                                                            const uint4* __restrict__ payload, uint32_t mask,
                                                            const int32_t* __restrict__ rows_sorted, int depth,
                                                            int32_t* __restrict__ nbr) {
  const int k3 = ksize * ksize * ksize, r = ksize / 2;
  const int l8 = threadIdx.x & 7;
  const int nblk_axis = 1 << (depth - 2), bshift = 3 * depth - 6;
  const int64_t groups = ((int64_t)gridDim.x * blockDim.x) >> 3;
  const int64_t n_pad = (n + 3) & ~3ll;      // whole warps stay in the loop: the shuffles below need all 32 lanes
  for (int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3; row < n_pad; row += groups) {
    const bool live = row < n;
    const int4 s = live ? reinterpret_cast<const int4*>(indices)[row] : make_int4(0, 0, 0, 0);
    const int bx0 = (s.y - r) >> 2, by0 = (s.z - r) >> 2, bz0 = (s.w - r) >> 2;
    uint32_t m_lo = 0, m_hi = 0, start = 0;
    {
      const int bx = (l8 & 4) ? (s.y + r) >> 2 : bx0, by = (l8 & 2) ? (s.z + r) >> 2 : by0,
                bz = (l8 & 1) ? (s.w + r) >> 2 : bz0;
      if (live && bx >= 0 && by >= 0 && bz >= 0 && bx < nblk_axis && by < nblk_axis && bz < nblk_axis) {
        const uint64_t kb = ((uint64_t)(uint32_t)s.x << bshift) | (spread3_u64((uint32_t)bx) << 2) |
                            (spread3_u64((uint32_t)by) << 1) | spread3_u64((uint32_t)bz);
        uint32_t slot = hash_site(kb) & mask;
        unsigned long long cur = keys[slot];
        while (cur != kb && cur != kEmptyKey) {
          slot = (slot + 1) & mask;
          cur = keys[slot];
        }
        if (cur == kb) {
          const uint4 pl = payload[slot];
          m_lo = pl.x; m_hi = pl.y; start = pl.z;
        }
      }
    }
    for (int k0 = 0; k0 < k3; k0 += 8) {
      const int k = k0 + l8;
      const int dz = k % ksize - r, dy = (k / ksize) % ksize - r, dx = k / (ksize * ksize) - r;
      const int x = s.y + dx, y = s.z + dy, z = s.w + dz;
      const int sel = (((x >> 2) != bx0) ? 4 : 0) | (((y >> 2) != by0) ? 2 : 0) | (((z >> 2) != bz0) ? 1 : 0);
      const uint32_t lo = __shfl_sync(0xffffffffu, m_lo, sel, 8), hi = __shfl_sync(0xffffffffu, m_hi, sel, 8);
      const uint32_t st = __shfl_sync(0xffffffffu, start, sel, 8);
      if (k < k3 && live) {
        const uint64_t m = ((uint64_t)hi << 32) | lo;
        const int cell = ((x & 2) << 4) | ((y & 2) << 3) | ((z & 2) << 2) | ((x & 1) << 2) | ((y & 1) << 1) | (z & 1);
        nbr[row * k3 + k] = ((m >> cell) & 1ull) ? rows_sorted[st + __popcll(m & ((1ull << cell) - 1ull))] : -1;
      }
    }
  }
}

template <typename T>
static int conv_cuda_core(const T* in, const T* w, const float* bias, const int32_t* nbr, int64_t n, int cin, int cout,
                          int k3, T* out, cudaStream_t st) {
  size_t smem = (size_t)k3 * cin * cout * sizeof(float);
  if (cin == 4 && cout == 32 && k3 <= 128) {
    const size_t sm4 = (size_t)k3 * 32 * sizeof(float4);
    int g = (int)cdiv(n, 8);
    const int cap = kNumSMs * 3 * 4;
    if (g > cap) g = cap;
    cudaFuncSetAttribute(subm_conv_stem4_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm4);
    subm_conv_stem4_kernel<T><<<g, 256, sm4, st>>>(in, w, bias, nbr, n, k3, out);
    return check_launch();
  }
  if (cin <= 8 && cout <= 64 && smem <= 200 * 1024) {
    int g = (int)cdiv(n, 8);
    int cap = kNumSMs * 2 * 8;
    if (g > cap) g = cap;
    if (cout <= 32) {
      cudaFuncSetAttribute(subm_conv_rowwarp_kernel<T, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      subm_conv_rowwarp_kernel<T, 1><<<g, 256, smem, st>>>(in, w, bias, nbr, n, cin, cout, k3, out);
    } else {
      cudaFuncSetAttribute(subm_conv_rowwarp_kernel<T, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      subm_conv_rowwarp_kernel<T, 2><<<g, 256, smem, st>>>(in, w, bias, nbr, n, cin, cout, k3, out);
    }
    return check_launch();
  }
  dim3 grid((unsigned)cdiv(n, BM), (unsigned)cdiv(cout, BN));
  subm_conv_tiled_kernel<T><<<grid, 256, 0, st>>>(in, w, bias, nbr, n, cin, cout, k3, out);
  return check_launch();
}

// implemented in subm_tc.cu
size_t subm_conv_tc_workspace(int64_t n, int c_in, int c_out, int k3);
int subm_conv_tc(const __nv_bfloat16* in, const __nv_bfloat16* w, const float* bias, const int32_t* nbr, int64_t n,
                 int c_in, int c_out, int k3, __nv_bfloat16* out, void* ws, size_t ws_bytes, cudaStream_t st);
bool subm_conv_tc_supported(int c_in, int c_out, int k3);
size_t subm_conv_tc_packed_bytes(int c_in, int c_out, int k3);
int subm_pack_weight_tc(const __nv_bfloat16* w, int c_in, int c_out, int k3, __nv_bfloat16* packed, cudaStream_t st);
int subm_conv_tc_packed(const __nv_bfloat16* in, const __nv_bfloat16* packed, const float* bias, const int32_t* nbr,
                        int64_t n, int c_in, int c_out, int k3, __nv_bfloat16* out, cudaStream_t st);

}  // namespace ptv3

using namespace ptv3;

extern "C" {

size_t ptv3_subm_map_workspace(int64_t n) {
  uint32_t cap = table_capacity(n > 0 ? n : 1);
  return align_up((size_t)cap * 8) + align_up((size_t)cap * 4) + 512;
}

int ptv3_subm_build_map(const int32_t* indices, int64_t n, int ksize, int32_t* nbr, void* workspace,
                        size_t workspace_bytes, void* stream) {
  if (n < 0 || ksize < 1 || ksize > 7 || (ksize & 1) == 0) return PTV3_EINVAL;
  if (n == 0) return PTV3_OK;
  if (!indices || !nbr || n >= (1ll << 30)) return PTV3_EINVAL;
  uint32_t cap = table_capacity(n);
  Workspace ws(workspace, workspace_bytes);
  unsigned long long* keys = ws.take<unsigned long long>(cap);
  int* vals = ws.take<int>(cap);
  if (!ws.ok()) return PTV3_EWORKSPACE;
  cudaStream_t st = as_stream(stream);
  cudaMemsetAsync(keys, 0xff, (size_t)cap * 8, st);
  cudaMemsetAsync(vals, 0x7f, (size_t)cap * 4, st);
  subm_insert_kernel<<<grid_for(n, 256), 256, 0, st>>>(indices, n, keys, vals, cap - 1, device_error_flag());
  int64_t total = n * ksize * ksize * ksize;
  subm_lookup_kernel<<<grid_for(total, 256, 16), 256, 0, st>>>(indices, n, ksize, keys, vals, cap - 1, nbr);
  return check_launch(2);
}

int ptv3_subm_stem_fused(const int32_t* indices, int64_t n, int ksize, const float* in, const float* weight,
                         const float* bias, int act, float* out32, void* out16, void* workspace, size_t workspace_bytes,
                         void* stream) {
  if (n < 0 || ksize < 1 || ksize > 5 || (ksize & 1) == 0 || act < 0 || act > 1) return PTV3_EINVAL;
  if (n == 0) return PTV3_OK;
  if (!indices || !in || !weight || (!out32 && !out16) || n >= (1ll << 30)) return PTV3_EINVAL;
  const int k3 = ksize * ksize * ksize;
  uint32_t cap = table_capacity(n);
  Workspace ws(workspace, workspace_bytes);
  unsigned long long* keys = ws.take<unsigned long long>(cap);
  int* vals = ws.take<int>(cap);
  if (!ws.ok()) return PTV3_EWORKSPACE;
  cudaStream_t st = as_stream(stream);
  cudaMemsetAsync(keys, 0xff, (size_t)cap * 8, st);
  cudaMemsetAsync(vals, 0x7f, (size_t)cap * 4, st);
  subm_insert_kernel<<<grid_for(n, 256), 256, 0, st>>>(indices, n, keys, vals, cap - 1, device_error_flag());
  const size_t smem = (size_t)k3 * 32 * sizeof(float4);
  static bool configured_dev[kMaxDeviceSlots] = {};
  bool& configured = configured_dev[device_slot()];
  if (!configured) {
    if (cudaFuncSetAttribute(subm_stem_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 125 * 32 * 16) !=
        cudaSuccess)
      return PTV3_ELAUNCH;
    configured = true;
  }
  int64_t blocks = cdiv(n, 16);
  if (blocks > kNumSMs * 2) blocks = kNumSMs * 2;
  subm_stem_fused_kernel<<<(unsigned)blocks, 512, smem, st>>>(in, weight, bias, indices, keys, vals, cap - 1, n, ksize,
                                                               act, out32, (__nv_bfloat16*)out16);
  return check_launch(2);
}

size_t ptv3_subm_stem_z_workspace(int64_t n) {
  uint32_t cap = table_capacity(n > 0 ? n : 1);
  return align_up((size_t)cap * 8) + align_up((size_t)cap * 16) + align_up((size_t)(n > 0 ? n : 1) * 4) + 1024;
}

int ptv3_subm_stem_fused_z(const int32_t* indices, int64_t n, int ksize, const int64_t* code_z, const int64_t* order_z,
                           int depth, const float* in, const float* weight, const float* bias, int act, float* out32,
                           void* out16, void* workspace, size_t workspace_bytes, void* stream) {
  if (n < 0 || ksize < 1 || ksize > 5 || (ksize & 1) == 0 || act < 0 || act > 1 || depth < 2 || depth > 16)
    return PTV3_EINVAL;
  if (n == 0) return PTV3_OK;
  if (!indices || !in || !weight || !code_z || !order_z || (!out32 && !out16) || n >= (1ll << 30)) return PTV3_EINVAL;
  const int k3 = ksize * ksize * ksize;
  uint32_t cap = table_capacity(n);
  Workspace ws(workspace, workspace_bytes);
  unsigned long long* keys = ws.take<unsigned long long>(cap);
  uint4* payload = ws.take<uint4>(cap);
  int32_t* rows_sorted = ws.take<int32_t>((size_t)n);
  if (!ws.ok()) return PTV3_EWORKSPACE;
  cudaStream_t st = as_stream(stream);
  cudaMemsetAsync(keys, 0xff, (size_t)cap * 8, st);
  stem_blocks_kernel<<<grid_for(n, 256), 256, 0, st>>>(code_z, order_z, n, keys, payload, cap - 1, rows_sorted,
                                                       device_error_flag());
  const size_t smem = (size_t)k3 * 32 * sizeof(float4);
  static bool configured_dev[kMaxDeviceSlots] = {};
  bool& configured = configured_dev[device_slot()];
  if (!configured) {
    if (cudaFuncSetAttribute(subm_stem_z_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (125 * 32 + 16 * 128) * 16) != cudaSuccess)
      return PTV3_ELAUNCH;
    configured = true;
  }
  int64_t blocks = cdiv(n, 16);
  if (blocks > kNumSMs * 2) blocks = kNumSMs * 2;
  subm_stem_z_kernel<<<(unsigned)blocks, 512, smem + 16 * 128 * sizeof(float4), st>>>(in, weight, bias, indices, keys, payload, cap - 1, rows_sorted,
                                                          depth, n, ksize, act, out32, (__nv_bfloat16*)out16);
  return check_launch(2);
}

int ptv3_subm_build_map_z(const int32_t* indices, int64_t n, int ksize, const int64_t* code_z, const int64_t* order_z,
                          int depth, int32_t* nbr, void* workspace, size_t workspace_bytes, void* stream) {
  if (n < 0 || (ksize != 3 && ksize != 5) || depth < 2 || depth > 16) return PTV3_EINVAL;
  if (n == 0) return PTV3_OK;
  if (!indices || !nbr || !code_z || !order_z || n >= (1ll << 30)) return PTV3_EINVAL;
  uint32_t cap = table_capacity(n);
  Workspace ws(workspace, workspace_bytes);
  unsigned long long* keys = ws.take<unsigned long long>(cap);
  uint4* payload = ws.take<uint4>(cap);
  int32_t* rows_sorted = ws.take<int32_t>((size_t)n);
  if (!ws.ok()) return PTV3_EWORKSPACE;
  cudaStream_t st = as_stream(stream);
  cudaMemsetAsync(keys, 0xff, (size_t)cap * 8, st);
  stem_blocks_kernel<<<grid_for(n, 256), 256, 0, st>>>(code_z, order_z, n, keys, payload, cap - 1, rows_sorted,
                                                       device_error_flag());
  subm_lookup_z_kernel<<<grid_for(n * 8, 256, 16), 256, 0, st>>>(indices, n, ksize, keys, payload, cap - 1, rows_sorted,
                                                                 depth, nbr);
  return check_launch(2);
}

size_t ptv3_subm_conv_workspace(int64_t n, int c_in, int c_out, int k3, int dtype, int algo) {
  if (dtype == PTV3_BF16 && algo != 1 && subm_conv_tc_supported(c_in, c_out, k3))
    return subm_conv_tc_workspace(n, c_in, c_out, k3);
  return 0;
}

int ptv3_subm_conv(const void* in, const void* weight, const float* bias, const int32_t* nbr, int64_t n, int c_in,
                   int c_out, int k3, int dtype, int algo, void* out, void* workspace, size_t workspace_bytes,
                   void* stream) {
  if (n < 0 || c_in < 1 || c_out < 1 || k3 < 1 || algo < 0 || algo > 2) return PTV3_EINVAL;
  if (n == 0) return PTV3_OK;
  if (!in || !weight || !nbr || !out) return PTV3_EINVAL;
  cudaStream_t st = as_stream(stream);
  if (dtype == PTV3_BF16) {
    bool tc_ok = subm_conv_tc_supported(c_in, c_out, k3);
    if (algo == 2 && !tc_ok) return PTV3_EUNSUPPORTED;
    if (algo != 1 && tc_ok)
      return subm_conv_tc((const __nv_bfloat16*)in, (const __nv_bfloat16*)weight, bias, nbr, n, c_in, c_out, k3,
                          (__nv_bfloat16*)out, workspace, workspace_bytes, st);
    return conv_cuda_core<__nv_bfloat16>((const __nv_bfloat16*)in, (const __nv_bfloat16*)weight, bias, nbr, n, c_in,
                                         c_out, k3, (__nv_bfloat16*)out, st);
  }
  if (dtype == PTV3_F32) {
    if (algo == 2) return PTV3_EUNSUPPORTED;
    return conv_cuda_core<float>((const float*)in, (const float*)weight, bias, nbr, n, c_in, c_out, k3, (float*)out,
                                 st);
  }
  return PTV3_EINVAL;
}

int ptv3_subm_conv_tc_supported(int c_in, int c_out, int k3) { return subm_conv_tc_supported(c_in, c_out, k3) ? 1 : 0; }

size_t ptv3_subm_pack_weight_bytes(int c_in, int c_out, int k3) {
  return subm_conv_tc_supported(c_in, c_out, k3) ? subm_conv_tc_packed_bytes(c_in, c_out, k3) : 0;
}

int ptv3_subm_pack_weight(const void* weight, int c_in, int c_out, int k3, void* packed, void* stream) {
  if (!weight || !packed) return PTV3_EINVAL;
  return subm_pack_weight_tc((const __nv_bfloat16*)weight, c_in, c_out, k3, (__nv_bfloat16*)packed, as_stream(stream));
}

int ptv3_subm_conv_packed(const void* in, const void* packed, const float* bias, const int32_t* nbr, int64_t n,
                          int c_in, int c_out, int k3, void* out, void* stream) {
  if (n < 0) return PTV3_EINVAL;
  if (n == 0) return PTV3_OK;
  if (!in || !packed || !nbr || !out) return PTV3_EINVAL;
  return subm_conv_tc_packed((const __nv_bfloat16*)in, (const __nv_bfloat16*)packed, bias, nbr, n, c_in, c_out, k3,
                             (__nv_bfloat16*)out, as_stream(stream));
}

}  // extern "C"
