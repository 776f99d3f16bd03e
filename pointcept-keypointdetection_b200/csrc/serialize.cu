// A2: Morton / Hilbert codes + row-wise LSD radix argsort with fused inverse.
// Integer work, bit-exact against oracle/serialization.py (pinned on the
// reference's encode(), tests/golden/serialization.npz).
#include <mutex>
#include "common.cuh"

namespace ptv3 {

// ------------------------------ encode ---------------------------------------
// spread the low 16 bits of v so that bit i lands on bit 3i
__device__ __forceinline__ uint64_t spread3(uint64_t v) {
  v &= 0xffffull;
  v = (v | (v << 32)) & 0x001f00000000ffffull;
  v = (v | (v << 16)) & 0x001f0000ff0000ffull;
  v = (v | (v << 8)) & 0x100f00f00f00f00full;
  v = (v | (v << 4)) & 0x10c30c30c30c30c3ull;
  v = (v | (v << 2)) & 0x1249249249249249ull;
  return v;
}

// x -> bit 3i+2, y -> 3i+1, z -> 3i   (z_order.py:40-52)
__device__ __forceinline__ uint64_t morton3(uint32_t x, uint32_t y, uint32_t z) {
  return (spread3(x) << 2) | (spread3(y) << 1) | spread3(z);
}

// Skilling transpose -> interleave -> gray decode (hilbert.py:143-198)
__device__ __forceinline__ uint64_t hilbert3(uint32_t x0, uint32_t x1, uint32_t x2, int depth) {
  uint32_t X[3] = {x0, x1, x2};
  for (int bit = depth - 1; bit >= 0; --bit) {
    uint32_t q = 1u << bit, p = q - 1u;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      if (X[d] & q) {
        X[0] ^= p;  // invert the low bits of dim 0
      } else {
        uint32_t t = (X[0] ^ X[d]) & p;  // exchange the low bits of dim 0 and dim d
        X[0] ^= t;
        X[d] ^= t;
      }
    }
  }
  uint64_t h = morton3(X[0], X[1], X[2]);
  h ^= h >> 1;
  h ^= h >> 2;
  h ^= h >> 4;
  h ^= h >> 8;
  h ^= h >> 16;
  h ^= h >> 32;
  return h;
}

struct OrderList {
  int ids[8];
  int n;
};

template <typename G>
__global__ void __launch_bounds__(256) encode_kernel(const G* __restrict__ grid, const int64_t* __restrict__ batch,
                                                     int64_t n, int depth, OrderList orders,
                                                     int64_t* __restrict__ code) {
  const uint32_t mask = (depth >= 32) ? 0xffffffffu : ((1u << depth) - 1u);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    uint32_t x = (uint32_t)((int64_t)grid[3 * i + 0]) & mask;
    uint32_t y = (uint32_t)((int64_t)grid[3 * i + 1]) & mask;
    uint32_t z = (uint32_t)((int64_t)grid[3 * i + 2]) & mask;
    int64_t hi = batch ? (int64_t)((uint64_t)batch[i] << (3 * depth)) : 0;
    uint64_t cz = 0, czt = 0, ch = 0, cht = 0;
    bool need[4] = {false, false, false, false};
    for (int r = 0; r < orders.n; ++r) need[orders.ids[r] & 3] = true;
    if (need[0]) cz = morton3(x, y, z);
    if (need[1]) czt = morton3(y, x, z);
    if (need[2]) ch = hilbert3(x, y, z, depth);
    if (need[3]) cht = hilbert3(y, x, z, depth);
    for (int r = 0; r < orders.n; ++r) {
      int id = orders.ids[r] & 3;
      uint64_t c = id == 0 ? cz : id == 1 ? czt : id == 2 ? ch : cht;
      code[(int64_t)r * n + i] = hi | (int64_t)c;
    }
  }
}

// per-axis maximum of an (n,3) grid: out[a] = max_i grid[i][a]
template <typename G>
__global__ void __launch_bounds__(256) grid_max_kernel(const G* __restrict__ grid, int64_t n,
                                                       unsigned long long* __restrict__ out) {
  long long m[3] = {INT64_MIN, INT64_MIN, INT64_MIN};
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      long long v = (long long)grid[3 * i + a];
      m[a] = v > m[a] ? v : m[a];
    }
  }
#pragma unroll
  for (int a = 0; a < 3; ++a) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
      long long o = __shfl_xor_sync(0xffffffffu, m[a], d);
      m[a] = o > m[a] ? o : m[a];
    }
    if ((threadIdx.x & 31) == 0) atomicMax(reinterpret_cast<long long*>(out) + a, m[a]);
  }
}

__global__ void init_i64_kernel(int64_t* p, int64_t v, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// ------------------------------ radix argsort ----------------------------------
constexpr int kRsThreads = 256;
constexpr int kRsItems = 8;
constexpr int kRsTile = kRsThreads * kRsItems;  // 2048 keys per CTA
constexpr int kRadix = 256;

// hist[row][digit][block]
__global__ void __launch_bounds__(kRsThreads) rs_hist_kernel(const uint64_t* __restrict__ keys, int64_t n, int shift,
                                                             uint32_t* __restrict__ hist, int nblk) {
  __shared__ uint32_t sh[kRadix];
  int row = blockIdx.y;
  const uint64_t* k = keys + (int64_t)row * n;
  sh[threadIdx.x] = 0;
  __syncthreads();
  int64_t base = (int64_t)blockIdx.x * kRsTile;
#pragma unroll
  for (int i = 0; i < kRsItems; ++i) {
    int64_t idx = base + (int64_t)i * kRsThreads + threadIdx.x;
    if (idx < n) atomicAdd(&sh[(uint32_t)(k[idx] >> shift) & 255u], 1u);
  }
  __syncthreads();
  hist[((int64_t)row * kRadix + threadIdx.x) * nblk + blockIdx.x] = sh[threadIdx.x];
}

// stable scatter.  Tile order = warp-major, round, lane = memory order.
// FIRST: values are the implicit indices.  LAST: write order (int64) + inverse.
template <bool FIRST, bool LAST>
__global__ void __launch_bounds__(kRsThreads) rs_scatter_kernel(
    const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in, uint64_t* __restrict__ keys_out,
    uint32_t* __restrict__ vals_out, int64_t* __restrict__ order, int64_t* __restrict__ inverse, int64_t n, int shift,
    const uint32_t* __restrict__ hist_scanned, int nblk) {
  __shared__ uint32_t cnt[kRsThreads / 32][kRadix];
  const int row = blockIdx.y;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint64_t* kin = keys_in + (int64_t)row * n;
  for (int i = threadIdx.x; i < (kRsThreads / 32) * kRadix; i += kRsThreads) (&cnt[0][0])[i] = 0;
  __syncthreads();

  uint64_t key[kRsItems];
  uint32_t rank[kRsItems];
  const int64_t wbase = (int64_t)blockIdx.x * kRsTile + (int64_t)warp * (kRsItems * 32);
  const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
  for (int r = 0; r < kRsItems; ++r) {
    int64_t idx = wbase + r * 32 + lane;
    bool valid = idx < n;
    key[r] = valid ? kin[idx] : ~0ull;
    uint32_t digit = valid ? ((uint32_t)(key[r] >> shift) & 255u) : 256u;
    uint32_t peers = __match_any_sync(0xffffffffu, digit);
    int leader = __ffs(peers) - 1;
    uint32_t old = 0;
    if (valid && lane == leader) {
      old = cnt[warp][digit];
      cnt[warp][digit] = old + __popc(peers);
    }
    old = __shfl_sync(0xffffffffu, old, leader);
    rank[r] = old + __popc(peers & lt_mask);
    __syncwarp();
  }
  __syncthreads();
  {
    // digit d = threadIdx.x: exclusive scan over warps + global base of (row, d, block)
    uint32_t run = hist_scanned[((int64_t)row * kRadix + threadIdx.x) * nblk + blockIdx.x];
#pragma unroll
    for (int w = 0; w < kRsThreads / 32; ++w) {
      uint32_t c = cnt[w][threadIdx.x];
      cnt[w][threadIdx.x] = run;
      run += c;
    }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < kRsItems; ++r) {
    int64_t idx = wbase + r * 32 + lane;
    if (idx < n) {
      uint32_t digit = (uint32_t)(key[r] >> shift) & 255u;
      int64_t pos = (int64_t)cnt[warp][digit] + rank[r];
      uint32_t v = FIRST ? (uint32_t)idx : vals_in[(int64_t)row * n + idx];
      if (LAST) {
        order[(int64_t)row * n + pos] = (int64_t)v;
        if (inverse) inverse[(int64_t)row * n + v] = pos;
        if (keys_out) keys_out[(int64_t)row * n + pos] = key[r];
      } else {
        keys_out[(int64_t)row * n + pos] = key[r];
        vals_out[(int64_t)row * n + pos] = v;
      }
    }
  }
}

struct RsBuffers {
  uint64_t *kA, *kB;
  uint32_t *vA, *vB, *hist, *partials;
  int nblk;
};

static size_t rs_workspace_bytes(int64_t rows, int64_t n) {
  int64_t nn = n > 0 ? n : 1;
  int64_t nblk = cdiv(nn, kRsTile);
  size_t b = 0;
  b += 2 * align_up((size_t)rows * nn * 8);
  b += 2 * align_up((size_t)rows * nn * 4);
  b += align_up((size_t)rows * kRadix * nblk * 4);
  b += align_up(scan_workspace_elems(rows, (int64_t)kRadix * nblk) * 4);
  return b + 1024;
}

// sorts (rows,n) keys; returns order/inverse as int64; optional sorted keys out.
int radix_argsort_rows(const uint64_t* keys, int64_t rows, int64_t n, int key_bits, int64_t* order, int64_t* inverse,
                       uint64_t* sorted_keys, void* ws_ptr, size_t ws_bytes, cudaStream_t stream) {
  if (rows <= 0 || n <= 0) return PTV3_OK;
  if (n >= (1ll << 31) || key_bits < 1 || key_bits > 64) return PTV3_EINVAL;
  Workspace ws(ws_ptr, ws_bytes);
  int nblk = (int)cdiv(n, kRsTile);
  uint64_t* kA = ws.take<uint64_t>((size_t)rows * n);
  uint64_t* kB = ws.take<uint64_t>((size_t)rows * n);
  uint32_t* vA = ws.take<uint32_t>((size_t)rows * n);
  uint32_t* vB = ws.take<uint32_t>((size_t)rows * n);
  uint32_t* hist = ws.take<uint32_t>((size_t)rows * kRadix * nblk);
  uint32_t* partials = ws.take<uint32_t>(scan_workspace_elems(rows, (int64_t)kRadix * nblk));
  if (!ws.ok()) return PTV3_EWORKSPACE;
  int passes = (key_bits + 7) / 8;
  dim3 grid(nblk, (unsigned)rows);
  const uint64_t* kin = keys;
  const uint32_t* vin = nullptr;
  for (int p = 0; p < passes; ++p) {
    int shift = 8 * p;
    bool first = p == 0, last = p == passes - 1;
    uint64_t* kout = (p & 1) ? kB : kA;
    uint32_t* vout = (p & 1) ? vB : vA;
    rs_hist_kernel<<<grid, kRsThreads, 0, stream>>>(kin, n, shift, hist, nblk);
    g_launch_count += 1;
    int rc = exclusive_scan_u32(hist, hist, rows, (int64_t)kRadix * nblk, partials, nullptr, stream);
    if (rc != PTV3_OK) return rc;
    if (first && last)
      rs_scatter_kernel<true, true><<<grid, kRsThreads, 0, stream>>>(kin, vin, sorted_keys, nullptr, order, inverse, n,
                                                                      shift, hist, nblk);
    else if (first)
      rs_scatter_kernel<true, false><<<grid, kRsThreads, 0, stream>>>(kin, vin, kout, vout, nullptr, nullptr, n, shift,
                                                                       hist, nblk);
    else if (last)
      rs_scatter_kernel<false, true><<<grid, kRsThreads, 0, stream>>>(kin, vin, sorted_keys, nullptr, order, inverse,
                                                                       n, shift, hist, nblk);
    else
      rs_scatter_kernel<false, false><<<grid, kRsThreads, 0, stream>>>(kin, vin, kout, vout, nullptr, nullptr, n,
                                                                        shift, hist, nblk);
    g_launch_count += 1;
    kin = kout;
    vin = vout;
  }
  return cudaGetLastError() == cudaSuccess ? PTV3_OK : PTV3_ELAUNCH;
}

size_t radix_argsort_workspace(int64_t rows, int64_t n) { return rs_workspace_bytes(rows, n); }

}  // namespace ptv3

using namespace ptv3;

extern "C" {

int ptv3_encode(const void* grid_coord, int grid_is_i64, const int64_t* batch, int64_t n, int depth,
                const int* order_ids_host, int n_orders, int64_t* code, void* stream) {
  if (n < 0 || depth < 1 || depth > 16 || n_orders < 1 || n_orders > 8 || !order_ids_host) return PTV3_EINVAL;
  if (n == 0) return PTV3_OK;
  if (!grid_coord || !code) return PTV3_EINVAL;
  OrderList ol;
  ol.n = n_orders;
  for (int i = 0; i < 8; ++i) ol.ids[i] = 0;
  for (int i = 0; i < n_orders; ++i) {
    if (order_ids_host[i] < 0 || order_ids_host[i] > 3) return PTV3_EINVAL;
    ol.ids[i] = order_ids_host[i];
  }
  int g = grid_for(n, 256);
  if (grid_is_i64)
    encode_kernel<int64_t><<<g, 256, 0, as_stream(stream)>>>((const int64_t*)grid_coord, batch, n, depth, ol, code);
  else
    encode_kernel<int32_t><<<g, 256, 0, as_stream(stream)>>>((const int32_t*)grid_coord, batch, n, depth, ol, code);
  return check_launch();
}

int ptv3_grid_max(const void* grid_coord, int grid_is_i64, int64_t n, int64_t* out_max3, void* stream) {
  if (n < 0 || !out_max3) return PTV3_EINVAL;
  init_i64_kernel<<<1, 32, 0, as_stream(stream)>>>(out_max3, INT64_MIN, 3);
  g_launch_count += 1;
  if (n == 0) return check_launch(0);
  int g = grid_for(n, 256, 2);
  if (grid_is_i64)
    grid_max_kernel<int64_t><<<g, 256, 0, as_stream(stream)>>>((const int64_t*)grid_coord, n,
                                                               (unsigned long long*)out_max3);
  else
    grid_max_kernel<int32_t><<<g, 256, 0, as_stream(stream)>>>((const int32_t*)grid_coord, n,
                                                               (unsigned long long*)out_max3);
  return check_launch();
}

size_t ptv3_argsort_rows_workspace(int64_t rows, int64_t n) { return radix_argsort_workspace(rows, n); }

int ptv3_argsort_rows(const int64_t* code, int64_t rows, int64_t n, int key_bits, int64_t* order, int64_t* inverse,
                      void* workspace, size_t workspace_bytes, void* stream) {
  if (rows < 0 || n < 0) return PTV3_EINVAL;
  if (rows == 0 || n == 0) return PTV3_OK;
  if (!order || !code) return PTV3_EINVAL;
  return radix_argsort_rows((const uint64_t*)code, rows, n, key_bits, order, inverse, nullptr, workspace,
                            workspace_bytes, as_stream(stream));
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------
// ptv3_argsort_rows_static: ptv3_argsort_rows without a caller workspace and without ANY allocation (no counterpart in
// the reference; added for callers that cannot call cudaMalloc: CUDA graph
// capture, tight serving loops, harnesses that forbid it).  The radix sort's scratch (two key buffers, two index
// buffers, the per-tile histograms and the scan partials) lives in the file-scope device array below: 8 MiB that every
// GPU which loads the library holds until exit.  In host code `g_argsort_static_scratch` is only a host-side stand-in
// for the symbol -- never a device address -- so the device address is resolved ONCE PER DEVICE, with that device
// current, by cudaGetSymbolAddress, and that pointer is what the kernels receive.
// Every call on a device shares the array and finds the previous call's contents in it.  Nothing needs to start at
// zero: each pass writes every histogram slot, scan partial, key and index it later reads (rs_hist_kernel stores all
// 256 x nblk counters of a row; the scatter of pass p writes the buffers pass p + 1 reads).  Calls that overlap on
// the GPU (different streams, or graph branches that run concurrently) RACE on the array: serialise them.
// ---------------------------------------------------------------------------------------------------------------
namespace ptv3 {
constexpr size_t kArgsortStaticBytes = (size_t)PTV3_ARGSORT_STATIC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_argsort_static_scratch[kArgsortStaticBytes];

constexpr int kMaxStaticDevices = 64;
static void* g_argsort_static_ptr[kMaxStaticDevices] = {};
static std::mutex g_argsort_static_mutex;

// Device address of the scratch array on the CURRENT device (resolved at the first call on that device).
static int argsort_static_scratch_ptr(void** out) {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxStaticDevices) return PTV3_ELAUNCH;
  std::lock_guard<std::mutex> lock(g_argsort_static_mutex);
  if (g_argsort_static_ptr[dev] == nullptr) {
    void* ptr = nullptr;
    if (cudaGetSymbolAddress((void**)&ptr, g_argsort_static_scratch) != cudaSuccess || ptr == nullptr) {
      cudaGetLastError();   // leave no sticky error behind
      return PTV3_ELAUNCH;
    }
    g_argsort_static_ptr[dev] = ptr;
  }
  *out = g_argsort_static_ptr[dev];
  return PTV3_OK;
}
}  // namespace ptv3

extern "C" {

int64_t ptv3_argsort_rows_static_max_elems(void) { return (int64_t)PTV3_ARGSORT_STATIC_MAX_ELEMS; }

int ptv3_argsort_rows_static(const int64_t* code, int64_t rows, int64_t n, int key_bits, int64_t* order,
                             int64_t* inverse, void* stream) {
  if (rows < 0 || n < 0) return PTV3_EINVAL;
  if (rows == 0 || n == 0) return PTV3_OK;
  if (!order || !code) return PTV3_EINVAL;
  // the stated limit first (a clear error, independent of how the scratch is laid out), then the exact byte check
  if (rows > PTV3_ARGSORT_STATIC_MAX_ELEMS || n > PTV3_ARGSORT_STATIC_MAX_ELEMS ||
      rows * n > (int64_t)PTV3_ARGSORT_STATIC_MAX_ELEMS)
    return PTV3_ETOOLARGE;
  if (radix_argsort_workspace(rows, n) > kArgsortStaticBytes) return PTV3_ETOOLARGE;
  void* scratch = nullptr;
  int rc = argsort_static_scratch_ptr(&scratch);
  if (rc != PTV3_OK) return rc;
  return radix_argsort_rows((const uint64_t*)code, rows, n, key_bits, order, inverse, nullptr, scratch,
                            kArgsortStaticBytes, as_stream(stream));
}

}  // extern "C"
