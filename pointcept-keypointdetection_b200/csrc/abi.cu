// libptv3_b200.so: version / error strings / shared prefix-sum helper.
#include "common.cuh"

namespace ptv3 {

int64_t g_launch_count = 0;

// Sticky device-side error word (bit mask): a kernel that meets input outside its contract ORs a bit in and skips the
// offending work instead of producing silent garbage; ptv3_device_error() reads and clears it (one synchronising copy).
//   1 = attention window record violates the ptv3_build_windows contract (k_len > max_len, queries outside the keys)
//   2 = sparse-conv site outside the packed key range (batch id >= 1024 or a coordinate outside [0, 262128))
__device__ int g_device_error_flag = 0;
int* device_error_flag() {
  int* p = nullptr;
  return cudaGetSymbolAddress((void**)&p, g_device_error_flag) == cudaSuccess ? p : nullptr;
}
int device_error_fetch_and_clear() {
  int v = 0, zero = 0;
  if (cudaMemcpyFromSymbol(&v, g_device_error_flag, sizeof(int)) != cudaSuccess) return PTV3_ELAUNCH;
  if (v != 0) cudaMemcpyToSymbol(g_device_error_flag, &zero, sizeof(int));
  return v;
}

// ---- phase 1: per-chunk sums ------------------------------------------------
__global__ void __launch_bounds__(kScanThreads) scan_chunk_sums(const uint32_t* __restrict__ in, int64_t n,
                                                                uint32_t* __restrict__ partials, int chunks) {
  __shared__ uint32_t sm[8];
  int row = blockIdx.y;
  const uint32_t* src = in + (int64_t)row * n;
  int64_t base = (int64_t)blockIdx.x * kScanChunk;
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    int64_t idx = base + (int64_t)i * kScanThreads + threadIdx.x;
    if (idx < n) s += src[idx];
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int w = 0; w < 8; ++w) t += sm[w];
    partials[(int64_t)row * chunks + blockIdx.x] = t;
  }
}

// ---- phase 2: one block per row scans the chunk sums in place ----------------
__global__ void __launch_bounds__(kScanThreads) scan_partials(uint32_t* __restrict__ partials, int chunks,
                                                              uint32_t* __restrict__ total) {
  __shared__ uint32_t sm[8];
  uint32_t* p = partials + (int64_t)blockIdx.x * chunks;
  uint32_t carry = 0;
  for (int base = 0; base < chunks; base += kScanThreads) {
    int idx = base + threadIdx.x;
    uint32_t v = idx < chunks ? p[idx] : 0u;
    uint32_t tot;
    uint32_t ex = block_exclusive_scan_256(v, sm, &tot);
    if (idx < chunks) p[idx] = carry + ex;
    carry += tot;
  }
  if (total != nullptr && threadIdx.x == 0) total[blockIdx.x] = carry;
}

// ---- phase 3: scan inside each chunk, add the chunk offset -------------------
__global__ void __launch_bounds__(kScanThreads) scan_chunks(const uint32_t* __restrict__ in, uint32_t* __restrict__ out,
                                                            int64_t n, const uint32_t* __restrict__ partials,
                                                            int chunks) {
  __shared__ uint32_t sm[8];
  int row = blockIdx.y;
  const uint32_t* src = in + (int64_t)row * n;
  uint32_t* dst = out + (int64_t)row * n;
  // thread t owns kScanItems consecutive elements -> one exclusive scan of the per-thread sums
  int64_t base = (int64_t)blockIdx.x * kScanChunk + (int64_t)threadIdx.x * kScanItems;
  uint32_t v[kScanItems];
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    v[i] = (base + i < n) ? src[base + i] : 0u;
    s += v[i];
  }
  uint32_t tot;
  uint32_t ex = block_exclusive_scan_256(s, sm, &tot) + partials[(int64_t)row * chunks + blockIdx.x];
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    if (base + i < n) dst[base + i] = ex;
    ex += v[i];
  }
}

int exclusive_scan_u32(const uint32_t* in, uint32_t* out, int64_t rows, int64_t n, uint32_t* partials, uint32_t* total,
                       cudaStream_t stream) {
  if (rows <= 0) return PTV3_OK;
  int chunks = (int)cdiv(n > 0 ? n : 1, kScanChunk);
  dim3 grid(chunks, (unsigned)rows);
  scan_chunk_sums<<<grid, kScanThreads, 0, stream>>>(in, n, partials, chunks);
  scan_partials<<<(unsigned)rows, kScanThreads, 0, stream>>>(partials, chunks, total);
  scan_chunks<<<grid, kScanThreads, 0, stream>>>(in, out, n, partials, chunks);
  return check_launch(3);
}

}  // namespace ptv3

extern "C" {

int ptv3_abi_version(void) { return 1; }

const char* ptv3_error_string(int code) {
  switch (code) {
    case PTV3_OK: return "ok";
    case PTV3_EINVAL: return "invalid argument or unsupported shape";
    case PTV3_EWORKSPACE: return "workspace too small";
    case PTV3_ELAUNCH: return "CUDA launch failure";
    case PTV3_EUNSUPPORTED: return "unsupported configuration";
    case PTV3_ETOOLARGE: return "size beyond the fixed capacity of this allocation-free entry point";
    default: return "unknown error";
  }
}

int64_t ptv3_launch_count(void) { return ptv3::g_launch_count; }

int ptv3_device_error(void) { return ptv3::device_error_fetch_and_clear(); }

}  // extern "C"
