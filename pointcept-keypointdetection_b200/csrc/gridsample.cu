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

}  // extern "C"
