// Pipe-rate micro-benchmarks behind profiles/r02_attention.md: what bounds a head_dim-16 softmax on one B200 SM.
// Each kernel runs ONE CTA per SM (148 CTAs) and times its own loop with clock64(), so the result is in
// operations per SM clock and does not depend on the clock the box happens to run at.
//   mufu_f32      ex2.approx.ftz.f32                      (MUFU.EX2)
//   mufu_f16x2    ex2.approx.f16x2  -> 2 exps per PTX op  (SASS: two MUFU.EX2.F16 + PRMT)
//   hmma          mma.sync.m16n8k16 f16 -> fp32           (legacy tensor path, HMMA)
//   ldsm          ldmatrix.x4 (conflict-free)             (shared-memory pipe)
//   ldtm          tcgen05.ld.32x32b.x32                   (TMEM read path), 4 / 8 / 16 warps
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench/pipes tools/ubench/pipes.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define CHECK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

constexpr int ITERS = 4096;

__global__ void mufu_f32(float* out, long long* cyc, float seed) {
  float v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = seed * (threadIdx.x + i) * 1e-6f;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(v[i]));
  }
  long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += v[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

__global__ void mufu_f16x2(uint32_t* out, long long* cyc, uint32_t seed) {
  uint32_t v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = seed + threadIdx.x + i;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(v[i]));
  }
  long long t1 = clock64();
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s ^= v[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

__global__ void hmma(float* out, long long* cyc, uint32_t seed) {
  float c[4][4];
  uint32_t a[4] = {seed, seed + 1, seed + 2, seed + 3}, b0 = seed + 4, b1 = seed + 5;
#pragma unroll
  for (int i = 0; i < 4; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.f;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3])
                   : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
  long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

__global__ void ldsm(uint32_t* out, long long* cyc, int conflict) {
  __shared__ __align__(128) unsigned char sm[32768];
  for (int i = threadIdx.x; i < 8192; i += blockDim.x) reinterpret_cast<uint32_t*>(sm)[i] = i;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // conflict = 0: 32-byte rows with the 32-byte swizzle (attn_f16.cu); 1: plain 32-byte rows (the round-1 layout)
  const uint32_t rsw = conflict ? 0u : (uint32_t)((lane >> 2) & 1);
  uint32_t addr = (uint32_t)__cvta_generic_to_shared(sm) + (warp & 15) * 1024 +
                  (uint32_t)((((lane >> 4) << 3) + (lane & 7)) * 32) + ((((uint32_t)(lane >> 3) & 1) ^ rsw) << 4);
  uint32_t acc = 0;
  long long t0 = clock64();
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint32_t r0, r1, r2, r3;
      asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr + ((i & 1) << 9)));
      acc ^= r0 ^ r1 ^ r2 ^ r3;
    }
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

__global__ void ldtm(uint32_t* out, long long* cyc) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = slot + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)((warp >> 2) & 3) * 64;
  uint32_t acc = 0;
  long long t0 = clock64();
  for (int it = 0; it < ITERS / 4; ++it) {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      uint32_t v[32];
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
          "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
          "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
          : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
            "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
            "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
            "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
          : "r"(tmem + i * 32)
          : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
      for (int j = 0; j < 32; ++j) acc ^= v[j];
    }
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(slot) : "memory");
}

// Independent streams of the instruction kinds of one softmax chunk, in selectable proportions, to see which
// companion steals MUFU issue: per loop trip NM ex2, NC cvt.rn.f16x2.f32, NX 3-input max, NH mma.m16n8k16.
template <int NM, int NC, int NX, int NH>
__global__ void mix(float* out, long long* cyc, float seed) {
  float v[8], mx[4];
  uint32_t pk[4] = {0u, 0u, 0u, 0u};
  float c[4][4];
  uint32_t a[4] = {1u, 2u, 3u, 4u}, b0 = 5u, b1 = 6u;
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = seed * (threadIdx.x + i) * 1e-6f;
#pragma unroll
  for (int i = 0; i < 4; ++i) { mx[i] = seed * i; c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.f; }
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < ITERS / 4; ++it) {
#pragma unroll
    for (int i = 0; i < NM; ++i) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(v[i & 7]));
#pragma unroll
    for (int i = 0; i < NC; ++i) asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(pk[i & 3]) : "f"(mx[i & 3]), "f"(mx[(i + 1) & 3]));
#pragma unroll
    for (int i = 0; i < NX; ++i) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(mx[i & 3]) : "f"(seed), "f"(mx[(i + 2) & 3]));
#pragma unroll
    for (int i = 0; i < NH; ++i)
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+f"(c[i & 3][0]), "+f"(c[i & 3][1]), "+f"(c[i & 3][2]), "+f"(c[i & 3][3])
                   : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
  long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += v[i];
#pragma unroll
  for (int i = 0; i < 4; ++i) s += mx[i] + __uint_as_float(pk[i]) + c[i][0] + c[i][1] + c[i][2] + c[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// tcgen05.mma issue rate for small tiles: one thread issues `reps` MMAs M128 x N x K16 (bf16, fp32 accumulate in TMEM) back
// to back into the same accumulator, A from shared memory (SS) or from tensor memory (TS), then commits and waits.
// Operands are whatever the shared / tensor memory holds (timing only).
template <int N, bool TS, int M = 128>
__global__ void tc_mma(long long* cyc, int reps) {
  extern __shared__ __align__(1024) unsigned char sm[];
  __shared__ uint32_t slot;
  __shared__ __align__(8) unsigned long long bar;
  const int warp = threadIdx.x >> 5;
  const uint32_t sb = (uint32_t)__cvta_generic_to_shared(sm);
  const uint32_t barp = (uint32_t)__cvta_generic_to_shared(&bar);
  for (int i = threadIdx.x; i < 16384 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sm)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(barp));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = slot;
  if (threadIdx.x == 0) {
    // K-major no-swizzle descriptors (8-row core matrices, the two K halves 128 B apart, row groups 256 B apart)
    const uint64_t adesc = (uint64_t)((sb >> 4) & 0x3FFF) | ((uint64_t)(128 >> 4) << 16) | ((uint64_t)(256 >> 4) << 32) | ((uint64_t)1 << 46);
    const uint64_t bdesc = (uint64_t)(((sb + 8192) >> 4) & 0x3FFF) | ((uint64_t)(128 >> 4) << 16) | ((uint64_t)(256 >> 4) << 32) | ((uint64_t)1 << 46);
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
    long long t0 = clock64();
    for (int i = 0; i < reps; ++i) {
      if (TS)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem), "r"(tmem + 256), "l"(bdesc), "r"(idesc), "r"(1u) : "memory");
      else
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(1u) : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(barp) : "memory");
    uint32_t ok = 0;
    while (!ok)
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(barp), "r"(0u) : "memory");
    long long t1 = clock64();
    cyc[blockIdx.x] = t1 - t0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}

static double mean_cycles(long long* d_cyc, int n) {
  static long long h[256];
  cudaMemcpy(h, d_cyc, n * sizeof(long long), cudaMemcpyDeviceToHost);
  double s = 0;
  for (int i = 0; i < n; ++i) s += (double)h[i];
  return s / n;
}

int main() {
  const int G = 148;
  void* out;
  long long* cyc;
  CHECK(cudaMalloc(&out, G * 1024 * 4));
  CHECK(cudaMalloc(&cyc, G * sizeof(long long)));
  for (int threads : {256, 512, 1024}) {
    for (int rep = 0; rep < 2; ++rep) mufu_f32<<<G, threads>>>((float*)out, cyc, 1.f);
    CHECK(cudaDeviceSynchronize());
    printf("mufu_f32    %4d threads/SM: %6.2f exp/clk/SM\n", threads, (double)threads * ITERS * 8 / mean_cycles(cyc, G));
    for (int rep = 0; rep < 2; ++rep) mufu_f16x2<<<G, threads>>>((uint32_t*)out, cyc, 7u);
    CHECK(cudaDeviceSynchronize());
    printf("mufu_f16x2  %4d threads/SM: %6.2f exp/clk/SM (2 per PTX op)\n", threads, (double)threads * ITERS * 16 / mean_cycles(cyc, G));
    for (int rep = 0; rep < 2; ++rep) hmma<<<G, threads>>>((float*)out, cyc, 3u);
    CHECK(cudaDeviceSynchronize());
    double mm = (double)(threads / 32) * ITERS * 4 / mean_cycles(cyc, G);
    printf("hmma        %4d threads/SM: %6.3f m16n8k16/clk/SM = %7.1f dense FLOP/clk/SM\n", threads, mm, mm * 4096);
    for (int conflict = 0; conflict < 2; ++conflict) {
      for (int rep = 0; rep < 2; ++rep) ldsm<<<G, threads>>>((uint32_t*)out, cyc, conflict);
      CHECK(cudaDeviceSynchronize());
      printf("ldsm.x4 %s %4d threads/SM: %6.1f B/clk/SM\n", conflict ? "plain 32B rows " : "32B-swizzled   ", threads,
             (double)(threads / 32) * ITERS * 4 * 512 / mean_cycles(cyc, G));
    }
  }
#define MIX(NM, NC, NX, NH, label)                                                                                  \
  for (int threads : {512, 768}) {                                                                                  \
    for (int rep = 0; rep < 2; ++rep) mix<NM, NC, NX, NH><<<G, threads>>>((float*)out, cyc, 1.f);                   \
    CHECK(cudaDeviceSynchronize());                                                                                 \
    printf("mix %-44s %4d threads/SM: %6.2f exp/clk/SM\n", label, threads,                                          \
           (double)threads * (ITERS / 4) * NM / mean_cycles(cyc, G));                                               \
  }
  MIX(16, 0, 0, 0, "16 ex2")
  MIX(16, 8, 0, 0, "16 ex2 + 8 cvt.f16x2")
  MIX(16, 0, 8, 0, "16 ex2 + 8 max3")
  MIX(16, 0, 0, 10, "16 ex2 + 10 hmma")
  MIX(16, 8, 8, 0, "16 ex2 + 8 cvt + 8 max3")
  MIX(16, 8, 8, 10, "16 ex2 + 8 cvt + 8 max3 + 10 hmma (softmax mix)")
  MIX(16, 8, 8, 6, "16 ex2 + 8 cvt + 8 max3 + 6 hmma")
#define TCMMA(N, TS)                                                                                             \
  {                                                                                                              \
    cudaFuncSetAttribute(tc_mma<N, TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768);                     \
    for (int rep = 0; rep < 2; ++rep) tc_mma<N, TS><<<G, 128, 32768>>>(cyc, 2000);                               \
    CHECK(cudaDeviceSynchronize());                                                                              \
    printf("tcgen05.mma M128 N%-3d K16 %s: %6.1f clk per MMA (ideal %5.1f at 8192 FLOP/clk/SM)\n", N,            \
           TS ? "A in TMEM" : "A in smem", mean_cycles(cyc, G) / 2000.0, 128.0 * N * 16 * 2 / 8192.0);            \
  }
  TCMMA(16, false) TCMMA(32, false) TCMMA(64, false) TCMMA(128, false) TCMMA(256, false)
  TCMMA(16, true) TCMMA(32, true) TCMMA(64, true) TCMMA(128, true)
#define TCMMA64(N)                                                                                               \
  {                                                                                                              \
    cudaFuncSetAttribute(tc_mma<N, false, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768);              \
    for (int rep = 0; rep < 2; ++rep) tc_mma<N, false, 64><<<G, 128, 32768>>>(cyc, 2000);                        \
    CHECK(cudaDeviceSynchronize());                                                                              \
    printf("tcgen05.mma M64  N%-3d K16 A in smem: %6.1f clk per MMA (ideal %5.1f at 8192 FLOP/clk/SM)\n", N,      \
           mean_cycles(cyc, G) / 2000.0, 64.0 * N * 16 * 2 / 8192.0);                                            \
  }
  TCMMA64(64) TCMMA64(128) TCMMA64(256)
  for (int threads : {128, 256, 512}) {
    for (int rep = 0; rep < 2; ++rep) ldtm<<<G, threads>>>((uint32_t*)out, cyc);
    CHECK(cudaDeviceSynchronize());
    printf("ldtm.x32    %4d threads/SM: %6.1f B/clk/SM (%5.1f fp32 scores/clk/SM)\n", threads,
           (double)threads * (ITERS / 4) * 2 * 128 / mean_cycles(cyc, G), (double)threads * (ITERS / 4) * 2 * 32 / mean_cycles(cyc, G));
  }
  return 0;
}
