// Shared helpers for libptv3_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stddef.h>

#include "ptv3_b200.h"

namespace ptv3 {

extern int64_t g_launch_count;  // host-side count of kernel launches (abi.cu)

// Index of the current device for per-device one-time set-up flags (cudaFuncSetAttribute applies to the current device only).
constexpr int kMaxDeviceSlots = 64;
inline int device_slot() {
  int dev = 0;
  cudaGetDevice(&dev);
  return dev >= 0 && dev < kMaxDeviceSlots ? dev : 0;
}
int* device_error_flag();   // abi.cu: device address of the sticky error word (bit mask, see ptv3_device_error)

inline int check_launch(int n_kernels = 1) {
  g_launch_count += n_kernels;
  return cudaGetLastError() == cudaSuccess ? PTV3_OK : PTV3_ELAUNCH;
}

inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

// bump allocator over the caller's workspace
struct Workspace {
  char* base;
  size_t cap, used;
  Workspace(void* p, size_t bytes) : base(static_cast<char*>(p)), cap(bytes), used(0) {}
  template <typename T>
  T* take(size_t count) {
    size_t bytes = align_up(count * sizeof(T));
    if (used + bytes > cap || base == nullptr) { used = cap + 1; return nullptr; }
    T* r = reinterpret_cast<T*>(base + used);
    used += bytes;
    return r;
  }
  bool ok() const { return used <= cap; }
};

constexpr int kNumSMs = 148;

inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }

// grid for a grid-stride elementwise kernel: enough CTAs to fill 148 SMs a few times over
inline int grid_for(int64_t work_items, int threads, int max_waves = 8) {
  int64_t g = cdiv(work_items, threads);
  int64_t cap = (int64_t)kNumSMs * (2048 / threads) * max_waves;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

// ---------------------------------------------------------------------------
// exclusive prefix sum of uint32 arrays, `rows` independent arrays of length n
// (row stride n).  Three launches: chunk sums, scan of chunk sums, scan+add.
// total[row] (optional, device) receives the sum of every row.
// workspace: scan_workspace_u32(rows, n) uint32 elements.
// ---------------------------------------------------------------------------
constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanChunk = kScanThreads * kScanItems;  // 2048

inline size_t scan_workspace_elems(int64_t rows, int64_t n) { return (size_t)(rows * cdiv(n > 0 ? n : 1, kScanChunk)); }

int exclusive_scan_u32(const uint32_t* in, uint32_t* out, int64_t rows, int64_t n, uint32_t* partials,
                       uint32_t* total /* (rows) or nullptr */, cudaStream_t stream);

// ---------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t warp_inclusive_scan(uint32_t v, int lane) {
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, v, d);
    if (lane >= d) v += t;
  }
  return v;
}

// exclusive scan across a 256-thread block of one value per thread; returns the
// exclusive prefix, *block_total receives the block sum.  smem: 8 uint32.
__device__ __forceinline__ uint32_t block_exclusive_scan_256(uint32_t v, uint32_t* smem8, uint32_t* block_total) {
  int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t inc = warp_inclusive_scan(v, lane);
  if (lane == 31) smem8[warp] = inc;
  __syncthreads();
  uint32_t warp_off = 0, tot = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) {
    uint32_t s = smem8[w];
    if (w < warp) warp_off += s;
    tot += s;
  }
  *block_total = tot;
  __syncthreads();
  return warp_off + inc - v;
}

__device__ __forceinline__ float bf16_bits_to_float(uint16_t b) { return __uint_as_float(((uint32_t)b) << 16); }

template <typename T>
struct Elem;
template <>
struct Elem<float> {
  static __device__ __forceinline__ float load(const float* p) { return *p; }
  static __device__ __forceinline__ void store(float* p, float v) { *p = v; }
};
template <>
struct Elem<__nv_bfloat16> {
  static __device__ __forceinline__ float load(const __nv_bfloat16* p) { return __bfloat162float(*p); }
  static __device__ __forceinline__ void store(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
};

// exact-erf GELU, 0.5 x (1 + erf(x / sqrt 2)), with erf from Abramowitz-Stegun 7.1.26 (|error| <= 1.5e-7, far below
// the bf16 rounding of the hidden activation): 13 FMA-pipe instructions + one rcp and one ex2 on the MUFU pipe
// instead of the ~30 of erff().  gelu = 0.5 (x + |x|) - 0.5 |x| p(t) exp(-x^2 / 2),  t = 1 / (1 + 0.3275911 |x| / sqrt 2)
__device__ __forceinline__ float gelu_erf(float x) {
  const float ax = fabsf(x);
  float t;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(ax, 0.23164189f, 1.0f)));   // 0.3275911 / sqrt(2)
  float p = fmaf(t, 1.061405429f, -1.453152027f);
  p = fmaf(p, t, 1.421413741f);
  p = fmaf(p, t, -0.284496736f);
  p = fmaf(p, t, 0.254829592f);
  p *= t;
  float e;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * x * -0.72134752f));            // exp(-x^2/2) = 2^(-x^2 log2(e)/2)
  return fmaf(-0.5f * ax, p * e, 0.5f * (x + ax));
}

}  // namespace ptv3
