// A0: GridSample -- voxel floor, min-shift, FNV/ravel hash, stable sort by key,
// run-length unique, per-voxel representative.  Bit-exact against
// oracle/gridsample.py (pinned on the reference's GridSample, tests/golden/gridsample.npz).
#include "common.cuh"

namespace ptv3 {

int radix_argsort_rows(const uint64_t* keys, int64_t rows, int64_t n, int key_bits, int64_t* order, int64_t* inverse,
                       uint64_t* sorted_keys, void* ws_ptr, size_t ws_bytes, cudaStream_t stream);
size_t radix_argsort_workspace(int64_t rows, int64_t n);

__global__ void gs_init_minmax(long long* mm) {
  int i = threadIdx.x;
  if (i < 3) mm[i] = INT64_MAX;        // min
  else if (i < 6) mm[i] = INT64_MIN;   // max
}

// The quotient `coord / np.array(grid_size)` of transform.py:850 is a float32 division in the reference's pinned
// numpy 1.26.4 (uv_requirements.txt:84: value-based casting treats the 0-d float64 array like a scalar) and a float64
// division under numpy >= 2 (NEP 50).  Points within ~1e-6 cells of a voxel face land in different voxels, so the
// precision is selectable (f32div): hash_type 0 / 1 = fnv / ravel with the float64 quotient, 3 / 2 = fnv / ravel with
// the float32 quotient.  The inference-time voxelisers (tools/OffsetKeyPointPrediction_Qt.py:284,
// tools/pigseg_shared.py:133) divide by a Python float, which every numpy keeps in float32.
__global__ void __launch_bounds__(256) gs_floor_kernel(const float* __restrict__ coord, int64_t r, double grid_size,
                                                       int f32div, int64_t* __restrict__ grid,
                                                       long long* __restrict__ mm) {
  const float gs32 = (float)grid_size;
  long long mn[3] = {INT64_MAX, INT64_MAX, INT64_MAX}, mx[3] = {INT64_MIN, INT64_MIN, INT64_MIN};
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < r; i += (int64_t)gridDim.x * blockDim.x) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      double q = f32div ? (double)__fdiv_rn(coord[3 * i + a], gs32)
                        : (double)coord[3 * i + a] / grid_size;  // IEEE div.rn.f64
      long long g = (long long)floor(q);
      grid[3 * i + a] = g;
      mn[a] = g < mn[a] ? g : mn[a];
      mx[a] = g > mx[a] ? g : mx[a];
    }
  }
#pragma unroll
  for (int a = 0; a < 3; ++a) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
      long long o = __shfl_xor_sync(0xffffffffu, mn[a], d);
      mn[a] = o < mn[a] ? o : mn[a];
      o = __shfl_xor_sync(0xffffffffu, mx[a], d);
      mx[a] = o > mx[a] ? o : mx[a];
    }
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      atomicMin(&mm[a], mn[a]);
      atomicMax(&mm[3 + a], mx[a]);
    }
  }
}

// grid -= min ; key = hash(grid)
__global__ void __launch_bounds__(256) gs_hash_kernel(int64_t* __restrict__ grid, int64_t r,
                                                      const long long* __restrict__ mm, int hash_type,
                                                      uint64_t* __restrict__ key, int64_t* __restrict__ min_out) {
  const long long m0 = mm[0], m1 = mm[1], m2 = mm[2];
  const uint64_t a1 = (uint64_t)(mm[4] - m1) + 1ull, a2 = (uint64_t)(mm[5] - m2) + 1ull;
  if (blockIdx.x == 0 && threadIdx.x < 3) min_out[threadIdx.x] = mm[threadIdx.x];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < r; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t x = grid[3 * i + 0] - m0, y = grid[3 * i + 1] - m1, z = grid[3 * i + 2] - m2;
    grid[3 * i + 0] = x;
    grid[3 * i + 1] = y;
    grid[3 * i + 2] = z;
    uint64_t h;
    if (hash_type == 0) {
      // transform.py:949-963: multiply-then-xor over whole uint64 words
      h = 14695981039346656037ull;
      h *= 1099511628211ull;
      h ^= (uint64_t)x;
      h *= 1099511628211ull;
      h ^= (uint64_t)y;
      h *= 1099511628211ull;
      h ^= (uint64_t)z;
    } else {
      // transform.py:930-946
      h = ((uint64_t)x * a1 + (uint64_t)y) * a2 + (uint64_t)z;
    }
    key[i] = h;
  }
}

__global__ void __launch_bounds__(256) gs_flags_kernel(const uint64_t* __restrict__ skey, int64_t r,
                                                       uint32_t* __restrict__ flags) {
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < r; j += (int64_t)gridDim.x * blockDim.x)
    flags[j] = (j == 0 || skey[j] != skey[j - 1]) ? 1u : 0u;
}

__global__ void __launch_bounds__(256) gs_segments_kernel(const int64_t* __restrict__ idx_sort,
                                                          const uint32_t* __restrict__ flags,
                                                          const uint32_t* __restrict__ ex,
                                                          const uint32_t* __restrict__ total, int64_t r,
                                                          int64_t* __restrict__ inverse, int64_t* __restrict__ seg_start,
                                                          int64_t* __restrict__ n_voxels) {
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < r; j += (int64_t)gridDim.x * blockDim.x) {
    uint32_t fl = flags[j];
    int64_t v = (int64_t)ex[j] + fl - 1;
    inverse[idx_sort[j]] = v;
    if (fl) seg_start[v] = j;
    if (j == 0) {
      int64_t m = (int64_t)total[0];
      seg_start[m] = r;
      n_voxels[0] = m;
    }
  }
}

__global__ void __launch_bounds__(256) gs_select_kernel(const int64_t* __restrict__ idx_sort,
                                                        const int64_t* __restrict__ seg_start,
                                                        const int64_t* __restrict__ n_voxels,
                                                        const int64_t* __restrict__ rnd, int64_t* __restrict__ count,
                                                        int64_t* __restrict__ idx_unique) {
  const int64_t m = n_voxels[0];
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < m; v += (int64_t)gridDim.x * blockDim.x) {
    int64_t s = seg_start[v], c = seg_start[v + 1] - s;
    count[v] = c;
    int64_t pick = rnd ? (rnd[v] % c) : 0;
    idx_unique[v] = idx_sort[s + pick];
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Batched GridSample: every scene of a step in ONE launch chain (the per-scene chain cost ~20
// launches + one host read per scene).  Scenes are contiguous row ranges (cumulative `offset`); the result is the
// per-scene result of ptv3_gridsample concatenated: voxels ordered by (scene, ascending key), point indices global.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ int scene_of(const int64_t* __restrict__ offset, int b, int64_t i) {
  int lo = 0, hi = b - 1;          // first scene whose end offset is > i
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (offset[mid] > i) hi = mid; else lo = mid + 1;
  }
  return lo;
}

__global__ void gsb_init_minmax(long long* mm, int b) {
  for (int i = threadIdx.x; i < 6 * b; i += blockDim.x) mm[i] = (i % 6) < 3 ? INT64_MAX : INT64_MIN;
}

__global__ void __launch_bounds__(256) gsb_floor_kernel(const float* __restrict__ coord, int64_t r,
                                                        const int64_t* __restrict__ offset, int b, double grid_size,
                                                        int f32div, int64_t* __restrict__ grid, int* __restrict__ sid,
                                                        long long* __restrict__ mm) {
  const float gs32 = (float)grid_size;
  for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x; i0 < r; i0 += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = i0 + threadIdx.x;
    const bool live = i < r;
    int s = -1;
    long long g[3] = {0, 0, 0};
    if (live) {
      s = scene_of(offset, b, i);
      sid[i] = s;
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        const double q = f32div ? (double)__fdiv_rn(coord[3 * i + a], gs32) : (double)coord[3 * i + a] / grid_size;
        g[a] = (long long)floor(q);
        grid[3 * i + a] = g[a];
      }
    }
    // one atomic pair per warp and axis when the whole warp sits in one scene (the common case), per lane otherwise
    const unsigned act = __ballot_sync(0xffffffffu, live);
    const int s0 = __shfl_sync(0xffffffffu, s, __ffs(act) - 1);
    const bool uniform = __all_sync(0xffffffffu, !live || s == s0);
    if (uniform) {
      if (act == 0) continue;
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        long long mn = live ? g[a] : INT64_MAX, mx = live ? g[a] : INT64_MIN;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
          long long o = __shfl_xor_sync(0xffffffffu, mn, d);
          mn = o < mn ? o : mn;
          o = __shfl_xor_sync(0xffffffffu, mx, d);
          mx = o > mx ? o : mx;
        }
        if ((threadIdx.x & 31) == 0) {
          atomicMin(&mm[6 * s0 + a], mn);
          atomicMax(&mm[6 * s0 + 3 + a], mx);
        }
      }
    } else if (live) {
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        atomicMin(&mm[6 * s + a], g[a]);
        atomicMax(&mm[6 * s + 3 + a], g[a]);
      }
    }
  }
}

__global__ void __launch_bounds__(256) gsb_hash_kernel(int64_t* __restrict__ grid, int64_t r, const int* __restrict__ sid,
                                                       const long long* __restrict__ mm, int b, int hash_type,
                                                       uint64_t* __restrict__ key, int64_t* __restrict__ min_out) {
  if (blockIdx.x == 0)
    for (int i = threadIdx.x; i < 3 * b; i += blockDim.x) min_out[i] = mm[6 * (i / 3) + i % 3];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < r; i += (int64_t)gridDim.x * blockDim.x) {
    const long long* m = mm + 6 * sid[i];
    const int64_t x = grid[3 * i + 0] - m[0], y = grid[3 * i + 1] - m[1], z = grid[3 * i + 2] - m[2];
    grid[3 * i + 0] = x;
    grid[3 * i + 1] = y;
    grid[3 * i + 2] = z;
    uint64_t h;
    if (hash_type == 0) {
      h = 14695981039346656037ull;
      h *= 1099511628211ull;
      h ^= (uint64_t)x;
      h *= 1099511628211ull;
      h ^= (uint64_t)y;
      h *= 1099511628211ull;
      h ^= (uint64_t)z;
    } else {
      const uint64_t a1 = (uint64_t)(m[4] - m[1]) + 1ull, a2 = (uint64_t)(m[5] - m[2]) + 1ull;
      h = ((uint64_t)x * a1 + (uint64_t)y) * a2 + (uint64_t)z;
    }
    key[i] = h;
  }
}

// second sort key: scene id of the j-th point in key order
__global__ void __launch_bounds__(256) gsb_gather_sid_kernel(const int64_t* __restrict__ idx, const int* __restrict__ sid,
                                                             int64_t r, uint64_t* __restrict__ out) {
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < r; j += (int64_t)gridDim.x * blockDim.x)
    out[j] = (uint64_t)sid[idx[j]];
}

// compose the two sorts (idx_sort = idx1[order2], sorted key likewise) and flag the first point of every voxel
__global__ void __launch_bounds__(256) gsb_compose_kernel(const int64_t* __restrict__ idx1, const uint64_t* __restrict__ skey1,
                                                          const int64_t* __restrict__ order2,
                                                          const uint64_t* __restrict__ ssid, int64_t r,
                                                          int64_t* __restrict__ idx_sort, uint32_t* __restrict__ flags) {
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < r; j += (int64_t)gridDim.x * blockDim.x) {
    const int64_t o = order2[j];
    idx_sort[j] = idx1[o];
    bool head = j == 0;
    if (!head) {
      const int64_t op = order2[j - 1];
      head = ssid[j] != ssid[j - 1] || skey1[o] != skey1[op];
    }
    flags[j] = head ? 1u : 0u;
  }
}

// voxels per scene (cumulative): the exclusive voxel scan at every scene boundary of the sorted order (scenes keep their
// row ranges under the stable (scene, key) sort)
__global__ void gsb_voxel_offset_kernel(const int64_t* __restrict__ offset, int b, const uint32_t* __restrict__ ex,
                                        const uint32_t* __restrict__ total, int64_t r, int64_t* __restrict__ voxel_offset) {
  for (int i = threadIdx.x; i < b; i += blockDim.x) {
    const int64_t e = offset[i];
    voxel_offset[i] = e >= r ? (int64_t)total[0] : (int64_t)ex[e];
  }
}


// ---------------------------------------------------------------------------------------------------------------
// PT-v3m2 GridPooling front end (point_transformer_v3m2_sonata.py:393-417).  The reference runs
// torch.unique(dim=0) over rows [x/stride | batch << 48, y/stride, z/stride]: clusters come out in ascending
// (batch, x, y, z) order.  Here one 64-bit key (batch, x, y, z) with `bits` bits per axis is radix-sorted (stable) and
// run-length encoded; the pooled grid / batch / spconv indices are decoded from the run heads and the per-axis maximum
// of the pooled grid (the next serialization depth, structure.py:73) is reduced in the same pass.
// ---------------------------------------------------------------------------------------------------------------
template <typename TG>
__global__ void __launch_bounds__(256) gp_key_kernel(const TG* __restrict__ grid, const int64_t* __restrict__ batch,
                                                     int64_t n, int stride, int bits, uint64_t* __restrict__ key,
                                                     long long* __restrict__ meta, int* __restrict__ err) {
  if (blockIdx.x == 0 && threadIdx.x < 4) meta[threadIdx.x] = threadIdx.x == 0 ? 0 : -1;
  const uint64_t lim = 1ull << bits;
  bool bad = false;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const long long gx = (long long)grid[3 * i], gy = (long long)grid[3 * i + 1], gz = (long long)grid[3 * i + 2];
    const uint64_t x = (uint64_t)(gx / stride), y = (uint64_t)(gy / stride), z = (uint64_t)(gz / stride);
    bad |= gx < 0 || gy < 0 || gz < 0 || x >= lim || y >= lim || z >= lim;
    key[i] = ((((uint64_t)batch[i] << bits | x) << bits | y) << bits) | z;
  }
  if (bad) atomicOr(err, 4);
}

__global__ void __launch_bounds__(256) gp_heads_kernel(const uint64_t* __restrict__ skey,
                                                       const int64_t* __restrict__ seg_start,
                                                       const int64_t* __restrict__ n_pooled, int bits,
                                                       int64_t* __restrict__ grid, int64_t* __restrict__ batch,
                                                       int32_t* __restrict__ indices4, long long* __restrict__ meta) {
  const int64_t m = n_pooled[0];
  const uint64_t msk = (1ull << bits) - 1ull;
  long long mx[3] = {-1, -1, -1};
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < m; v += (int64_t)gridDim.x * blockDim.x) {
    const uint64_t k = skey[seg_start[v]];
    const long long z = (long long)(k & msk), y = (long long)((k >> bits) & msk), x = (long long)((k >> (2 * bits)) & msk);
    const long long b = (long long)(k >> (3 * bits));
    grid[3 * v] = x;
    grid[3 * v + 1] = y;
    grid[3 * v + 2] = z;
    batch[v] = b;
    *reinterpret_cast<int4*>(indices4 + 4 * v) = make_int4((int)b, (int)x, (int)y, (int)z);
    mx[0] = x > mx[0] ? x : mx[0];
    mx[1] = y > mx[1] ? y : mx[1];
    mx[2] = z > mx[2] ? z : mx[2];
  }
#pragma unroll
  for (int a = 0; a < 3; ++a) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
      long long o = __shfl_xor_sync(0xffffffffu, mx[a], d);
      mx[a] = o > mx[a] ? o : mx[a];
    }
    if ((threadIdx.x & 31) == 0 && mx[a] >= 0) atomicMax(&meta[1 + a], mx[a]);
  }
}

}  // namespace ptv3

using namespace ptv3;

extern "C" {

size_t ptv3_gridsample_workspace(int64_t r) {
  int64_t rr = r > 0 ? r : 1;
  return radix_argsort_workspace(1, rr) + align_up((size_t)rr * 8) /*sorted keys*/ +
         align_up((size_t)(rr + 1) * 8) /*seg_start*/ + 2 * align_up((size_t)rr * 4) +
         align_up(scan_workspace_elems(1, rr) * 4) + 3 * 256 + 1024;
}

int ptv3_gridsample(const float* coord, int64_t r, double grid_size, int hash_type, const int64_t* rnd,
                    int64_t* grid_coord, uint64_t* key, int64_t* idx_sort, int64_t* inverse, int64_t* count,
                    int64_t* idx_unique, int64_t* min_grid, int64_t* n_voxels, void* workspace,
                    size_t workspace_bytes, void* stream) {
  if (r < 0 || !(grid_size > 0.0) || hash_type < 0 || hash_type > 3 || !n_voxels) return PTV3_EINVAL;
  const int f32div = hash_type >= 2;
  if (f32div) hash_type = 3 - hash_type;      // 2 -> ravel (1), 3 -> fnv (0)
  cudaStream_t st = as_stream(stream);
  if (r == 0) {
    cudaMemsetAsync(n_voxels, 0, 8, st);
    return PTV3_OK;
  }
  if (!coord || !grid_coord || !key || !idx_sort || !inverse || !count || !idx_unique || !min_grid) return PTV3_EINVAL;
  Workspace ws(workspace, workspace_bytes);
  size_t rs_bytes = radix_argsort_workspace(1, r);
  char* rs_ws = ws.take<char>(rs_bytes);
  uint64_t* skey = ws.take<uint64_t>((size_t)r);
  int64_t* seg_start = ws.take<int64_t>((size_t)r + 1);
  uint32_t* flags = ws.take<uint32_t>((size_t)r);
  uint32_t* ex = ws.take<uint32_t>((size_t)r);
  uint32_t* partials = ws.take<uint32_t>(scan_workspace_elems(1, r));
  long long* mm = ws.take<long long>(6);
  uint32_t* total = ws.take<uint32_t>(1);
  if (!ws.ok()) return PTV3_EWORKSPACE;
  int g = grid_for(r, 256);
  gs_init_minmax<<<1, 32, 0, st>>>(mm);
  gs_floor_kernel<<<g, 256, 0, st>>>(coord, r, grid_size, f32div, grid_coord, mm);
  gs_hash_kernel<<<g, 256, 0, st>>>(grid_coord, r, mm, hash_type, key, min_grid);
  g_launch_count += 3;
  int rc = radix_argsort_rows(key, 1, r, 64, idx_sort, nullptr, skey, rs_ws, rs_bytes, st);
  if (rc != PTV3_OK) return rc;
  gs_flags_kernel<<<g, 256, 0, st>>>(skey, r, flags);
  g_launch_count += 1;
  rc = exclusive_scan_u32(flags, ex, 1, r, partials, total, st);
  if (rc != PTV3_OK) return rc;
  gs_segments_kernel<<<g, 256, 0, st>>>(idx_sort, flags, ex, total, r, inverse, seg_start, n_voxels);
  gs_select_kernel<<<g, 256, 0, st>>>(idx_sort, seg_start, n_voxels, rnd, count, idx_unique);
  return check_launch(2);
}

size_t ptv3_gridsample_batch_workspace(int64_t r, int64_t n_scenes) {
  int64_t rr = r > 0 ? r : 1;
  return radix_argsort_workspace(1, rr) + 3 * align_up((size_t)rr * 8) /* skey1, ssid keys, ssid sorted */ +
         2 * align_up((size_t)rr * 8) /* idx1, order2 */ + align_up((size_t)(rr + 1) * 8) /* seg_start */ +
         align_up((size_t)rr * 4) /* sid */ + 2 * align_up((size_t)rr * 4) /* flags, ex */ +
         align_up(scan_workspace_elems(1, rr) * 4) + align_up((size_t)(n_scenes > 0 ? n_scenes : 1) * 48) + 2048;
}

int ptv3_gridsample_batch(const float* coord, int64_t r, const int64_t* offset, int64_t n_scenes, double grid_size,
                          int hash_type, const int64_t* rnd, int64_t* grid_coord, uint64_t* key, int64_t* idx_sort,
                          int64_t* inverse, int64_t* count, int64_t* idx_unique, int64_t* min_grid,
                          int64_t* voxel_offset, int64_t* n_voxels, void* workspace, size_t workspace_bytes,
                          void* stream) {
  if (r < 0 || n_scenes < 1 || n_scenes > (1 << 20) || !(grid_size > 0.0) || hash_type < 0 || hash_type > 3 || !n_voxels ||
      !offset)
    return PTV3_EINVAL;
  const int f32div = hash_type >= 2;
  if (f32div) hash_type = 3 - hash_type;
  cudaStream_t st = as_stream(stream);
  if (r == 0) {
    cudaMemsetAsync(n_voxels, 0, 8, st);
    if (voxel_offset) cudaMemsetAsync(voxel_offset, 0, 8 * (size_t)n_scenes, st);
    return PTV3_OK;
  }
  if (!coord || !grid_coord || !key || !idx_sort || !inverse || !count || !idx_unique || !min_grid || !voxel_offset)
    return PTV3_EINVAL;
  const int b = (int)n_scenes;
  Workspace ws(workspace, workspace_bytes);
  size_t rs_bytes = radix_argsort_workspace(1, r);
  char* rs_ws = ws.take<char>(rs_bytes);
  uint64_t* skey1 = ws.take<uint64_t>((size_t)r);
  uint64_t* sidk = ws.take<uint64_t>((size_t)r);
  uint64_t* ssid = ws.take<uint64_t>((size_t)r);
  int64_t* idx1 = ws.take<int64_t>((size_t)r);
  int64_t* order2 = ws.take<int64_t>((size_t)r);
  int64_t* seg_start = ws.take<int64_t>((size_t)r + 1);
  int* sid = ws.take<int>((size_t)r);
  uint32_t* flags = ws.take<uint32_t>((size_t)r);
  uint32_t* ex = ws.take<uint32_t>((size_t)r);
  uint32_t* partials = ws.take<uint32_t>(scan_workspace_elems(1, r));
  long long* mm = ws.take<long long>((size_t)6 * b);
  uint32_t* total = ws.take<uint32_t>(1);
  if (!ws.ok()) return PTV3_EWORKSPACE;
  const int g = grid_for(r, 256);
  gsb_init_minmax<<<1, 256, 0, st>>>(mm, b);
  gsb_floor_kernel<<<g, 256, 0, st>>>(coord, r, offset, b, grid_size, f32div, grid_coord, sid, mm);
  gsb_hash_kernel<<<g, 256, 0, st>>>(grid_coord, r, sid, mm, b, hash_type, key, min_grid);
  g_launch_count += 3;
  // stable LSD order: by key (64 bits), then by scene id
  int rc = radix_argsort_rows(key, 1, r, 64, idx1, nullptr, skey1, rs_ws, rs_bytes, st);
  if (rc != PTV3_OK) return rc;
  int scene_bits = 1;
  while ((1ll << scene_bits) < n_scenes) ++scene_bits;
  gsb_gather_sid_kernel<<<g, 256, 0, st>>>(idx1, sid, r, sidk);
  g_launch_count += 1;
  rc = radix_argsort_rows(sidk, 1, r, scene_bits, order2, nullptr, ssid, rs_ws, rs_bytes, st);
  if (rc != PTV3_OK) return rc;
  gsb_compose_kernel<<<g, 256, 0, st>>>(idx1, skey1, order2, ssid, r, idx_sort, flags);
  g_launch_count += 1;
  rc = exclusive_scan_u32(flags, ex, 1, r, partials, total, st);
  if (rc != PTV3_OK) return rc;
  gs_segments_kernel<<<g, 256, 0, st>>>(idx_sort, flags, ex, total, r, inverse, seg_start, n_voxels);
  gs_select_kernel<<<g, 256, 0, st>>>(idx_sort, seg_start, n_voxels, rnd, count, idx_unique);
  gsb_voxel_offset_kernel<<<1, 256, 0, st>>>(offset, b, ex, total, r, voxel_offset);
  return check_launch(3);
}

size_t ptv3_grid_pool_workspace(int64_t n) {
  int64_t rr = n > 0 ? n : 1;
  return radix_argsort_workspace(1, rr) + 2 * align_up((size_t)rr * 8) /* keys, sorted keys */ +
         2 * align_up((size_t)rr * 4) /* flags, ex */ + align_up(scan_workspace_elems(1, rr) * 4) + 1024;
}

int ptv3_grid_pool(const void* grid_coord, int grid_is_i64, const int64_t* batch, int64_t n, int stride, int bits,
                   int scene_bits, int64_t* cluster, int64_t* indices, int64_t* idx_ptr, int64_t* pooled_grid,
                   int64_t* pooled_batch, int32_t* indices4, int64_t* meta, void* workspace, size_t workspace_bytes,
                   void* stream) {
  if (n < 0 || stride < 1 || bits < 1 || scene_bits < 0 || 3 * bits + scene_bits > 63 || !meta) return PTV3_EINVAL;
  cudaStream_t st = as_stream(stream);
  if (n == 0) {
    cudaMemsetAsync(meta, 0, 8, st);
    cudaMemsetAsync(meta + 1, 0xff, 24, st);
    return PTV3_OK;
  }
  if (!grid_coord || !batch || !cluster || !indices || !idx_ptr || !pooled_grid || !pooled_batch || !indices4)
    return PTV3_EINVAL;
  Workspace ws(workspace, workspace_bytes);
  size_t rs_bytes = radix_argsort_workspace(1, n);
  char* rs_ws = ws.take<char>(rs_bytes);
  uint64_t* key = ws.take<uint64_t>((size_t)n);
  uint64_t* skey = ws.take<uint64_t>((size_t)n);
  uint32_t* flags = ws.take<uint32_t>((size_t)n);
  uint32_t* ex = ws.take<uint32_t>((size_t)n);
  uint32_t* partials = ws.take<uint32_t>(scan_workspace_elems(1, n));
  uint32_t* total = ws.take<uint32_t>(1);
  if (!ws.ok()) return PTV3_EWORKSPACE;
  const int g = grid_for(n, 256);
  long long* m = reinterpret_cast<long long*>(meta);
  if (grid_is_i64)
    gp_key_kernel<int64_t><<<g, 256, 0, st>>>((const int64_t*)grid_coord, batch, n, stride, bits, key, m,
                                              device_error_flag());
  else
    gp_key_kernel<int32_t><<<g, 256, 0, st>>>((const int32_t*)grid_coord, batch, n, stride, bits, key, m,
                                              device_error_flag());
  g_launch_count += 1;
  int rc = radix_argsort_rows(key, 1, n, 3 * bits + scene_bits, indices, nullptr, skey, rs_ws, rs_bytes, st);
  if (rc != PTV3_OK) return rc;
  gs_flags_kernel<<<g, 256, 0, st>>>(skey, n, flags);
  g_launch_count += 1;
  rc = exclusive_scan_u32(flags, ex, 1, n, partials, total, st);
  if (rc != PTV3_OK) return rc;
  gs_segments_kernel<<<g, 256, 0, st>>>(indices, flags, ex, total, n, cluster, idx_ptr, meta);
  gp_heads_kernel<<<g, 256, 0, st>>>(skey, idx_ptr, meta, bits, pooled_grid, pooled_batch, indices4, m);
  return check_launch(2);
}

}  // extern "C"
