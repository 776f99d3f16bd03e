// A6: SerializedPooling integer front end (no sort: run-length over the parent's
// serialized orders), pooled integer attributes, batch -> offset.
// Bit-exact against oracle/ptv3.py::_pool (pinned on the reference forward).
#include "common.cuh"

namespace ptv3 {

__global__ void __launch_bounds__(256) pool_flags_kernel(const int64_t* __restrict__ code,
                                                         const int64_t* __restrict__ order, int64_t n, int shift,
                                                         uint32_t* __restrict__ flags) {
  int row = blockIdx.y;
  const int64_t* c = code + (int64_t)row * n;
  const int64_t* o = order + (int64_t)row * n;
  uint32_t* f = flags + (int64_t)row * n;
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
    int64_t cur = c[o[j]] >> shift;
    uint32_t fl = 1u;
    if (j > 0) fl = (cur != (c[o[j - 1]] >> shift)) ? 1u : 0u;
    f[j] = fl;
  }
}

// row 0 defines the clusters
__global__ void __launch_bounds__(256) pool_cluster_kernel(const int64_t* __restrict__ order0,
                                                           const uint32_t* __restrict__ flags0,
                                                           const uint32_t* __restrict__ ex0,
                                                           const uint32_t* __restrict__ total0, int64_t n,
                                                           int64_t* __restrict__ cluster, int64_t* __restrict__ indices,
                                                           int64_t* __restrict__ idx_ptr, int64_t* __restrict__ head,
                                                           int64_t* __restrict__ n_pooled) {
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
    int64_t p = order0[j];
    uint32_t fl = flags0[j];
    int64_t k = (int64_t)ex0[j] + fl - 1;
    cluster[p] = k;
    indices[j] = p;
    if (fl) {
      idx_ptr[k] = j;
      head[k] = p;
    }
    if (j == 0) {
      int64_t m = (int64_t)total0[0];
      idx_ptr[m] = n;
      n_pooled[0] = m;
    }
  }
}

// every row: the pooled points in the order their runs appear along the parent's order
__global__ void __launch_bounds__(256) pool_orders_kernel(const int64_t* __restrict__ code,
                                                          const int64_t* __restrict__ order,
                                                          const uint32_t* __restrict__ flags,
                                                          const uint32_t* __restrict__ ex,
                                                          const int64_t* __restrict__ cluster, int64_t n, int shift,
                                                          int64_t* __restrict__ new_code,
                                                          int64_t* __restrict__ new_order,
                                                          int64_t* __restrict__ new_inverse) {
  int row = blockIdx.y;
  const int64_t* c = code + (int64_t)row * n;
  const int64_t* o = order + (int64_t)row * n;
  const uint32_t* f = flags + (int64_t)row * n;
  const uint32_t* e = ex + (int64_t)row * n;
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
    if (f[j]) {
      int64_t p = o[j];
      int64_t k = (int64_t)e[j];
      int64_t cl = cluster[p];
      new_order[(int64_t)row * n + k] = cl;
      new_inverse[(int64_t)row * n + cl] = k;
      new_code[(int64_t)row * n + cl] = c[p] >> shift;
    }
  }
}

__global__ void __launch_bounds__(256) pool_gather_ints_kernel(const int64_t* __restrict__ grid,
                                                               const int64_t* __restrict__ batch,
                                                               const int64_t* __restrict__ head, int64_t m,
                                                               int pool_depth, int64_t* __restrict__ new_grid,
                                                               int64_t* __restrict__ new_batch,
                                                               int32_t* __restrict__ new_indices) {
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < m; k += (int64_t)gridDim.x * blockDim.x) {
    int64_t p = head ? head[k] : k;
    int64_t x = grid[3 * p + 0] >> pool_depth, y = grid[3 * p + 1] >> pool_depth, z = grid[3 * p + 2] >> pool_depth;
    int64_t b = batch[p];
    if (new_grid) {
      new_grid[3 * k + 0] = x;
      new_grid[3 * k + 1] = y;
      new_grid[3 * k + 2] = z;
    }
    if (new_batch) new_batch[k] = b;
    if (new_indices) {
      int4 v = make_int4((int)b, (int)x, (int)y, (int)z);
      reinterpret_cast<int4*>(new_indices)[k] = v;
    }
  }
}

__global__ void __launch_bounds__(256) batch_to_offset_kernel(const int64_t* __restrict__ batch, int64_t n,
                                                              int64_t n_scenes, int64_t* __restrict__ offset) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t b = batch[i];
    if (i == 0)
      for (int64_t s = 0; s < b && s < n_scenes; ++s) offset[s] = 0;
    int64_t nb = (i + 1 < n) ? batch[i + 1] : n_scenes;
    for (int64_t s = b; s < nb && s < n_scenes; ++s) offset[s] = i + 1;
  }
}

__global__ void fill_i64_kernel(int64_t* p, int64_t v, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// offset2batch (utils/misc.py:13-22) without the host sync of repeat_interleave: row i belongs to the first scene
// whose cumulative offset exceeds i (binary search over the few scene offsets); rows beyond offset[-1] get the
// last scene
__global__ void __launch_bounds__(256) offset_to_batch_kernel(const int64_t* __restrict__ offset, int64_t n_scenes,
                                                              int64_t n, int64_t* __restrict__ batch) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t lo = 0, hi = n_scenes - 1;
    while (lo < hi) {
      const int64_t mid = (lo + hi) >> 1;
      if (offset[mid] > i) hi = mid;
      else lo = mid + 1;
    }
    batch[i] = lo;
  }
}

__global__ void __launch_bounds__(256) sparse_pair_maps_kernel(const int64_t* __restrict__ grid,
                                                               const int64_t* __restrict__ cluster, int64_t n,
                                                               int32_t* __restrict__ nbr_fwd,
                                                               int32_t* __restrict__ nbr_inv) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int k = (int)(((grid[3 * i] & 1) * 2 + (grid[3 * i + 1] & 1)) * 2 + (grid[3 * i + 2] & 1));
    const int64_t c = cluster[i];
    nbr_fwd[c * 8 + k] = (int32_t)i;
    nbr_inv[i * 8 + k] = (int32_t)c;
  }
}

}  // namespace ptv3

using namespace ptv3;

extern "C" {

size_t ptv3_pool_front_workspace(int64_t rows, int64_t n) {
  int64_t nn = n > 0 ? n : 1;
  return 2 * align_up((size_t)rows * nn * 4) + align_up(scan_workspace_elems(rows, nn) * 4) +
         align_up((size_t)rows * 4) + 1024;
}

int ptv3_pool_front(const int64_t* code, const int64_t* order, int64_t rows, int64_t n, int shift, int64_t* cluster,
                    int64_t* indices, int64_t* idx_ptr, int64_t* head, int64_t* new_code, int64_t* new_order,
                    int64_t* new_inverse, int64_t* n_pooled, void* workspace, size_t workspace_bytes, void* stream) {
  if (rows < 1 || n < 0 || shift < 0 || shift > 62 || !n_pooled || !idx_ptr) return PTV3_EINVAL;
  cudaStream_t st = as_stream(stream);
  if (n == 0) {
    fill_i64_kernel<<<1, 32, 0, st>>>(n_pooled, 0, 1);
    fill_i64_kernel<<<1, 32, 0, st>>>(idx_ptr, 0, 1);
    return check_launch(2);
  }
  Workspace ws(workspace, workspace_bytes);
  uint32_t* flags = ws.take<uint32_t>((size_t)rows * n);
  uint32_t* ex = ws.take<uint32_t>((size_t)rows * n);
  uint32_t* partials = ws.take<uint32_t>(scan_workspace_elems(rows, n));
  uint32_t* totals = ws.take<uint32_t>((size_t)rows);
  if (!ws.ok()) return PTV3_EWORKSPACE;
  dim3 grid(grid_for(n, 256), (unsigned)rows);
  pool_flags_kernel<<<grid, 256, 0, st>>>(code, order, n, shift, flags);
  g_launch_count += 1;
  int rc = exclusive_scan_u32(flags, ex, rows, n, partials, totals, st);
  if (rc != PTV3_OK) return rc;
  pool_cluster_kernel<<<grid.x, 256, 0, st>>>(order, flags, ex, totals, n, cluster, indices, idx_ptr, head, n_pooled);
  pool_orders_kernel<<<grid, 256, 0, st>>>(code, order, flags, ex, cluster, n, shift, new_code, new_order,
                                           new_inverse);
  return check_launch(2);
}

int ptv3_pool_gather_ints(const int64_t* grid, const int64_t* batch, const int64_t* head, int64_t m, int pool_depth,
                          int64_t* new_grid, int64_t* new_batch, int32_t* new_indices, void* stream) {
  if (m < 0 || pool_depth < 0) return PTV3_EINVAL;
  if (m == 0) return PTV3_OK;
  pool_gather_ints_kernel<<<grid_for(m, 256), 256, 0, as_stream(stream)>>>(grid, batch, head, m, pool_depth, new_grid,
                                                                           new_batch, new_indices);
  return check_launch();
}

int ptv3_offset_to_batch(const int64_t* offset, int64_t n_scenes, int64_t n, int64_t* batch, void* stream) {
  if (n < 0 || n_scenes < 1 || !offset) return PTV3_EINVAL;
  if (n == 0) return PTV3_OK;
  if (!batch) return PTV3_EINVAL;
  offset_to_batch_kernel<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(offset, n_scenes, n, batch);
  return check_launch();
}

int ptv3_sparse_pair_maps(const int64_t* grid, const int64_t* cluster, int64_t n, int64_t m, int32_t* nbr_fwd,
                          int32_t* nbr_inv, void* stream) {
  if (n < 0 || m < 0) return PTV3_EINVAL;
  if (n == 0) return PTV3_OK;
  if (!grid || !cluster || !nbr_fwd || !nbr_inv) return PTV3_EINVAL;
  cudaStream_t st = as_stream(stream);
  cudaMemsetAsync(nbr_fwd, 0xff, (size_t)m * 8 * 4, st);
  cudaMemsetAsync(nbr_inv, 0xff, (size_t)n * 8 * 4, st);
  sparse_pair_maps_kernel<<<grid_for(n, 256), 256, 0, st>>>(grid, cluster, n, nbr_fwd, nbr_inv);
  return check_launch();
}

int ptv3_batch_to_offset(const int64_t* batch, int64_t n, int64_t n_scenes, int64_t* offset, void* stream) {
  if (n < 0 || n_scenes < 1 || !offset) return PTV3_EINVAL;
  if (n == 0) {
    fill_i64_kernel<<<(unsigned)cdiv(n_scenes, 256), 256, 0, as_stream(stream)>>>(offset, 0, n_scenes);
    return check_launch();
  }
  batch_to_offset_kernel<<<grid_for(n, 256), 256, 0, as_stream(stream)>>>(batch, n, n_scenes, offset);
  return check_launch();
}

}  // extern "C"
