// Window attention backward, bf16 tensor-core kernel (head_dim 16); the fp32 / generic path is csrc/backward.cu.
//
// One CTA per (window, head), 8 warps.  K, V, Q, dO of the window are gathered once into shared memory through the
// serialized order (4 x 32 B per row), D_i = dO_i . O_i and the row's log-sum-exp live next to them.
//   phase A  a warp owns 16 queries: sweep 1 recomputes the log-sum-exp of its rows (S = Q K^T, online max / sum),
//            sweep 2 walks the keys 16 at a time: P = 2^(S' - lse), dP = dO V^T, dS = P (dP - D), dQ += dS K.
//   phase B  a warp owns 16 keys and walks the queries 16 at a time on the TRANSPOSED problem, so that P^T and dS^T
//            come out of the accumulators in the A-operand layout: S^T = K Q^T, dP^T = V dO^T, dV += P^T dO,
//            dK += dS^T Q.  dK / dV go out with atomicAdd (a row can be a key of two windows).
// Nothing of size window x window is stored; the price is recomputing S three times (exp count = 3 x forward).
#include "common.cuh"

namespace ptv3 {

namespace {

constexpr int kD = 16;

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ void mma16816(float c[4], const uint32_t a[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm4(uint32_t r[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void ldsm4t(uint32_t r[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void cp16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src));
}
__device__ __forceinline__ void zero16(uint32_t dst) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%1,%1,%1};" ::"r"(dst), "r"(0u));
}

}  // namespace

// smem: K, V, Q, dO [kpad][16] bf16 each; lse_s, d_s [kpad] fp32; order_s [kpad] int32
// lse_g (optional): the forward's log-sum-exp per (row, head), log2 units of the scaled scores -> sweep 1 is skipped
template <int kWarps, int kT>
__global__ void __launch_bounds__(kWarps * 32, 1) attn_bwd_mma_kernel(
    const __nv_bfloat16* __restrict__ qkv, const __nv_bfloat16* __restrict__ out, const __nv_bfloat16* __restrict__ dout,
    const int64_t* __restrict__ order, const int64_t* __restrict__ windows, const int64_t* __restrict__ n_windows, int h,
    float scale, float sl2, int kpad_max, float* __restrict__ dqkv, const float* __restrict__ lse_g) {
  constexpr int kThreads = kWarps * 32;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int head = (int)(blockIdx.x % (unsigned)h), w = (int)(blockIdx.x / (unsigned)h);   // heads adjacent: L2 line reuse
  if (w >= n_windows[0]) return;
  const int64_t q0 = windows[4 * w + 0], ql = windows[4 * w + 1], k0 = windows[4 * w + 2], kl = windows[4 * w + 3];
  if (ql <= 0 || kl <= 0) return;
  const int kpad = (int)((kl + 63) / 64) * 64;
  if (kpad > kpad_max) return;
  __nv_bfloat16* Ks = reinterpret_cast<__nv_bfloat16*>(smem_raw);
  const uint32_t ks = (uint32_t)__cvta_generic_to_shared(Ks);
  const uint32_t vs = ks + (uint32_t)kpad * 32, qs = vs + (uint32_t)kpad * 32, dos = qs + (uint32_t)kpad * 32;
  float* lse_s = reinterpret_cast<float*>(smem_raw + (size_t)kpad * 128);
  float* d_s = lse_s + kpad;
  int* order_s = reinterpret_cast<int*>(d_s + kpad);
  const int64_t rs = (int64_t)3 * h * kD, ro = (int64_t)h * kD;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int q_rel = (int)(q0 - k0);

  // ---- stage K, V (rows of the key range) ----
  for (int idx = tid; idx < kpad * 4; idx += kThreads) {
    int key = idx >> 2, part = idx & 3;
    uint32_t dst = (part < 2 ? ks : vs) + (uint32_t)(key * 32 + (part & 1) * 16);
    if (key < kl) {
      int64_t row = order ? order[k0 + key] : k0 + key;
      if (part == 0) order_s[key] = (int)row;
      cp16(dst, qkv + row * rs + (int64_t)(part < 2 ? 1 : 2) * h * kD + head * kD + (part & 1) * 8);
    } else {
      zero16(dst);
    }
  }
  for (int i = tid; i < kpad; i += kThreads) {
    lse_s[i] = 1e30f;   // padded queries: P = 2^(0 - 1e30) = 0
    d_s[i] = 0.f;
  }
  __syncthreads();      // order_s visible
  // ---- stage Q, dO (rows of the query range, indexed from 0) ----
  for (int idx = tid; idx < kpad * 4; idx += kThreads) {
    int qi = idx >> 2, part = idx & 3;
    uint32_t dst = (part < 2 ? qs : dos) + (uint32_t)(qi * 32 + (part & 1) * 16);
    if (qi < ql) {
      int64_t row = order_s[q_rel + qi];
      const __nv_bfloat16* src = part < 2 ? qkv + row * rs + head * kD + (part & 1) * 8
                                          : dout + row * ro + head * kD + (part & 1) * 8;
      cp16(dst, src);
    } else {
      zero16(dst);
    }
  }
  asm volatile("cp.async.commit_group;");
  asm volatile("cp.async.wait_group 0;");
  __syncthreads();
  // ---- D_i = dO_i . O_i ----
  for (int qi = tid; qi < ql; qi += kThreads) {
    const int64_t row = order_s[q_rel + qi];
    const uint4* op = reinterpret_cast<const uint4*>(out + row * ro + head * kD);
    const uint4* dp = reinterpret_cast<const uint4*>(smem_raw + (size_t)kpad * 96 + (size_t)qi * 32);
    float acc = 0.f;
#pragma unroll
    for (int hf = 0; hf < 2; ++hf) {
      uint4 a = op[hf], b = dp[hf];
      const uint32_t av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        acc = fmaf(__uint_as_float(av[e] << 16), __uint_as_float(bv[e] << 16), acc);
        acc = fmaf(__uint_as_float(av[e] & 0xffff0000u), __uint_as_float(bv[e] & 0xffff0000u), acc);
      }
    }
    d_s[qi] = acc;
    if (lse_g) lse_s[qi] = lse_g[row * h + head];
  }
  __syncthreads();

  const int g = lane >> 2, t = lane & 3;
  const int n_qtiles = (int)((ql + 15) >> 4);
  const int n_ktiles = (int)((kl + 15) >> 4);
  // ldmatrix lane offsets inside a 16-row x 16-col bf16 tile (row pitch 32 B)
  //  b_off: matrices {rows 0-7 cols 0-7, rows 0-7 cols 8-15, rows 8-15 cols 0-7, rows 8-15 cols 8-15}
  //         non-trans -> B operands (n = row, k = col) of two n8 blocks
  //  a_off: matrices {rows 0-7 cols 0-7, rows 8-15 cols 0-7, rows 0-7 cols 8-15, rows 8-15 cols 8-15}
  //         non-trans -> the A operand (m = row, k = col); trans -> B operands (k = row, n = col) of two n8 blocks
  const uint32_t b_off = (uint32_t)((((lane >> 4) << 3) + (lane & 7)) * 32 + ((lane >> 3) & 1) * 16);
  const uint32_t a_off = (uint32_t)(((((lane >> 3) & 1) << 3) + (lane & 7)) * 32 + (lane >> 4) * 16);

  // =========================== phase A: dQ (and the rows' log-sum-exp) ===========================
  // kT query tiles per warp and pass: the K, V and K^T fragments of a key tile serve all of them.  Q / dO rows are zero up
  // to kpad, and padded queries have lse = 1e30 (P = 0).
  for (int qt = warp * kT; qt < n_qtiles; qt += kWarps * kT) {
    uint32_t qa[kT][4], da[kT][4];
    float lse[kT][2], D[kT][2];
#pragma unroll
    for (int j = 0; j < kT; ++j) {
      ldsm4(qa[j], qs + (uint32_t)((qt + j) * 512) + a_off);
      ldsm4(da[j], dos + (uint32_t)((qt + j) * 512) + a_off);
      const int qi0 = (qt + j) * 16 + g, qi1 = qi0 + 8;
      // sweep 1 (only without the forward's log-sum-exp): log2 units of the scaled scores
      float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
      for (int c = 0; c < (lse_g ? 0 : kpad / 64); ++c) {
        float s[8][4];
#pragma unroll
        for (int b2 = 0; b2 < 4; ++b2) {
          uint32_t kb[4];
          ldsm4(kb, ks + (uint32_t)((c * 4 + b2) * 512) + b_off);
#pragma unroll
          for (int i = 0; i < 4; ++i) s[2 * b2][i] = s[2 * b2 + 1][i] = 0.f;
          mma16816(s[2 * b2], qa[j], kb[0], kb[1]);
          mma16816(s[2 * b2 + 1], qa[j], kb[2], kb[3]);
        }
        if ((c + 1) * 64 > kl) {
          const int base = c * 64 + 2 * t;
#pragma unroll
          for (int b = 0; b < 8; ++b) {
            int key = base + b * 8;
            if (key >= kl) s[b][0] = s[b][2] = -INFINITY;
            if (key + 1 >= kl) s[b][1] = s[b][3] = -INFINITY;
          }
        }
        float mx0 = m0, mx1 = m1;
#pragma unroll
        for (int b = 0; b < 8; ++b) {
          mx0 = fmaxf(mx0, fmaxf(s[b][0], s[b][1]) * sl2);
          mx1 = fmaxf(mx1, fmaxf(s[b][2], s[b][3]) * sl2);
        }
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        l0 *= ex2(m0 - mx0);
        l1 *= ex2(m1 - mx1);
        m0 = mx0;
        m1 = mx1;
#pragma unroll
        for (int b = 0; b < 8; ++b) {
          l0 += ex2(fmaf(s[b][0], sl2, -m0)) + ex2(fmaf(s[b][1], sl2, -m0));
          l1 += ex2(fmaf(s[b][2], sl2, -m1)) + ex2(fmaf(s[b][3], sl2, -m1));
        }
      }
      l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
      l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
      l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
      l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
      lse[j][0] = lse_g ? lse_s[qi0] : (qi0 < ql ? m0 + __log2f(l0) : 1e30f);
      lse[j][1] = lse_g ? lse_s[qi1] : (qi1 < ql ? m1 + __log2f(l1) : 1e30f);
      D[j][0] = d_s[qi0];
      D[j][1] = d_s[qi1];
    }
    // sweep 2: dQ
    float dq[kT][2][4];
#pragma unroll
    for (int j = 0; j < kT; ++j)
#pragma unroll
      for (int i = 0; i < 4; ++i) dq[j][0][i] = dq[j][1][i] = 0.f;
    for (int kt = 0; kt < n_ktiles; ++kt) {
      const uint32_t kb_addr = ks + (uint32_t)(kt * 512), vb_addr = vs + (uint32_t)(kt * 512);
      uint32_t kb[4], vb[4], kt_[4];
      ldsm4(kb, kb_addr + b_off);
      ldsm4(vb, vb_addr + b_off);
      ldsm4t(kt_, kb_addr + a_off);
      const bool edge = (kt + 1) * 16 > kl;   // keys past the window end
      const int key = kt * 16 + 2 * t;
#pragma unroll
      for (int j = 0; j < kT; ++j) {
        float s0[4] = {0.f, 0.f, 0.f, 0.f}, s1[4] = {0.f, 0.f, 0.f, 0.f};
        float p0[4] = {0.f, 0.f, 0.f, 0.f}, p1[4] = {0.f, 0.f, 0.f, 0.f};
        mma16816(s0, qa[j], kb[0], kb[1]);
        mma16816(s1, qa[j], kb[2], kb[3]);
        mma16816(p0, da[j], vb[0], vb[1]);   // dP
        mma16816(p1, da[j], vb[2], vb[3]);
        float ds0[4], ds1[4];
        ds0[0] = ex2(fmaf(s0[0], sl2, -lse[j][0])) * (p0[0] - D[j][0]);
        ds0[1] = ex2(fmaf(s0[1], sl2, -lse[j][0])) * (p0[1] - D[j][0]);
        ds0[2] = ex2(fmaf(s0[2], sl2, -lse[j][1])) * (p0[2] - D[j][1]);
        ds0[3] = ex2(fmaf(s0[3], sl2, -lse[j][1])) * (p0[3] - D[j][1]);
        ds1[0] = ex2(fmaf(s1[0], sl2, -lse[j][0])) * (p1[0] - D[j][0]);
        ds1[1] = ex2(fmaf(s1[1], sl2, -lse[j][0])) * (p1[1] - D[j][0]);
        ds1[2] = ex2(fmaf(s1[2], sl2, -lse[j][1])) * (p1[2] - D[j][1]);
        ds1[3] = ex2(fmaf(s1[3], sl2, -lse[j][1])) * (p1[3] - D[j][1]);
        if (edge) {
          if (key >= kl) ds0[0] = ds0[2] = 0.f;
          if (key + 1 >= kl) ds0[1] = ds0[3] = 0.f;
          if (key + 8 >= kl) ds1[0] = ds1[2] = 0.f;
          if (key + 9 >= kl) ds1[1] = ds1[3] = 0.f;
        }
        uint32_t pa[4] = {pack_bf16(ds0[0], ds0[1]), pack_bf16(ds0[2], ds0[3]), pack_bf16(ds1[0], ds1[1]),
                          pack_bf16(ds1[2], ds1[3])};
        mma16816(dq[j][0], pa, kt_[0], kt_[1]);
        mma16816(dq[j][1], pa, kt_[2], kt_[3]);
      }
    }
#pragma unroll
    for (int j = 0; j < kT; ++j) {
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int qi = (qt + j) * 16 + g + 8 * r;
        if (qi < ql) {
          float* o = dqkv + (int64_t)order_s[q_rel + qi] * rs + head * kD;
          *reinterpret_cast<float2*>(o + 2 * t) = make_float2(dq[j][0][2 * r] * scale, dq[j][0][2 * r + 1] * scale);
          *reinterpret_cast<float2*>(o + 8 + 2 * t) = make_float2(dq[j][1][2 * r] * scale, dq[j][1][2 * r + 1] * scale);
          if (t == 0 && !lse_g) lse_s[qi] = lse[j][r];
        }
      }
    }
  }
  __syncthreads();

  // =========================== phase B: dK, dV on the transposed problem ===========================
  // kT key tiles per warp and pass: the four Q / dO fragment loads and the lse / D reads of a query tile serve all of them
  // (shared-memory bandwidth, not MUFU, bounded this phase with one tile per warp: 4 ldmatrix.x4 + 4 LDS.64 per
  // 8 exp2 per lane).  K / V rows are zero up to kpad (a multiple of 64), so tiles kt .. kt+kT-1 (kT <= 4) are always readable.
  for (int kt = warp * kT; kt < n_ktiles; kt += kWarps * kT) {
    uint32_t ka[kT][4], va[kT][4];
    float dk[kT][2][4], dv[kT][2][4];
#pragma unroll
    for (int j = 0; j < kT; ++j) {
      ldsm4(ka[j], ks + (uint32_t)((kt + j) * 512) + a_off);
      ldsm4(va[j], vs + (uint32_t)((kt + j) * 512) + a_off);
#pragma unroll
      for (int i = 0; i < 4; ++i) dk[j][0][i] = dk[j][1][i] = dv[j][0][i] = dv[j][1][i] = 0.f;
    }
    for (int qt = 0; qt < n_qtiles; ++qt) {
      const uint32_t q_addr = qs + (uint32_t)(qt * 512), d_addr = dos + (uint32_t)(qt * 512);
      uint32_t qb[4], db[4], dt[4], qt_[4];
      ldsm4(qb, q_addr + b_off);
      ldsm4(db, d_addr + b_off);
      ldsm4t(dt, d_addr + a_off);
      ldsm4t(qt_, q_addr + a_off);
      const int qc = qt * 16 + 2 * t;
      const float2 ls0 = *reinterpret_cast<const float2*>(lse_s + qc), ls1 = *reinterpret_cast<const float2*>(lse_s + qc + 8);
      const float2 dd0 = *reinterpret_cast<const float2*>(d_s + qc), dd1 = *reinterpret_cast<const float2*>(d_s + qc + 8);
#pragma unroll
      for (int j = 0; j < kT; ++j) {
        float s0[4] = {0.f, 0.f, 0.f, 0.f}, s1[4] = {0.f, 0.f, 0.f, 0.f};
        float p0[4] = {0.f, 0.f, 0.f, 0.f}, p1[4] = {0.f, 0.f, 0.f, 0.f};
        mma16816(s0, ka[j], qb[0], qb[1]);   // S^T: rows = keys, cols = queries qt*16 + (0..7 | 8..15)
        mma16816(s1, ka[j], qb[2], qb[3]);
        mma16816(p0, va[j], db[0], db[1]);   // dP^T
        mma16816(p1, va[j], db[2], db[3]);
        float pt0[4], pt1[4];
        pt0[0] = ex2(fmaf(s0[0], sl2, -ls0.x));
        pt0[1] = ex2(fmaf(s0[1], sl2, -ls0.y));
        pt0[2] = ex2(fmaf(s0[2], sl2, -ls0.x));
        pt0[3] = ex2(fmaf(s0[3], sl2, -ls0.y));
        pt1[0] = ex2(fmaf(s1[0], sl2, -ls1.x));
        pt1[1] = ex2(fmaf(s1[1], sl2, -ls1.y));
        pt1[2] = ex2(fmaf(s1[2], sl2, -ls1.x));
        pt1[3] = ex2(fmaf(s1[3], sl2, -ls1.y));
        uint32_t pa[4] = {pack_bf16(pt0[0], pt0[1]), pack_bf16(pt0[2], pt0[3]), pack_bf16(pt1[0], pt1[1]),
                          pack_bf16(pt1[2], pt1[3])};
        uint32_t sa[4] = {pack_bf16(pt0[0] * (p0[0] - dd0.x), pt0[1] * (p0[1] - dd0.y)),
                          pack_bf16(pt0[2] * (p0[2] - dd0.x), pt0[3] * (p0[3] - dd0.y)),
                          pack_bf16(pt1[0] * (p1[0] - dd1.x), pt1[1] * (p1[1] - dd1.y)),
                          pack_bf16(pt1[2] * (p1[2] - dd1.x), pt1[3] * (p1[3] - dd1.y))};
        mma16816(dv[j][0], pa, dt[0], dt[1]);
        mma16816(dv[j][1], pa, dt[2], dt[3]);
        mma16816(dk[j][0], sa, qt_[0], qt_[1]);
        mma16816(dk[j][1], sa, qt_[2], qt_[3]);
      }
    }
#pragma unroll
    for (int j = 0; j < kT; ++j) {
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int kj = (kt + j) * 16 + g + 8 * r;
        if (kj < kl) {
          float* o = dqkv + (int64_t)order_s[kj] * rs + (int64_t)h * kD + head * kD;
          atomicAdd(o + 2 * t, dk[j][0][2 * r] * scale);
          atomicAdd(o + 2 * t + 1, dk[j][0][2 * r + 1] * scale);
          atomicAdd(o + 8 + 2 * t, dk[j][1][2 * r] * scale);
          atomicAdd(o + 9 + 2 * t, dk[j][1][2 * r + 1] * scale);
          o += (int64_t)h * kD;
          atomicAdd(o + 2 * t, dv[j][0][2 * r]);
          atomicAdd(o + 2 * t + 1, dv[j][0][2 * r + 1]);
          atomicAdd(o + 8 + 2 * t, dv[j][1][2 * r]);
          atomicAdd(o + 9 + 2 * t, dv[j][1][2 * r + 1]);
        }
      }
    }
  }
}

int g_attn_bwd_warps = 16;   // debug slot 6

// returns PTV3_EUNSUPPORTED when the window does not fit in shared memory (caller falls back to the CUDA-core path)
int attn_window_bwd_tc(const __nv_bfloat16* qkv, const __nv_bfloat16* out, const __nv_bfloat16* dout,
                       const int64_t* order, const int64_t* windows, const int64_t* n_windows, int64_t n_windows_max,
                       int64_t max_len, int h, float scale, float* dqkv, cudaStream_t st, const float* lse) {
  int kpad = (int)(cdiv(max_len, 64) * 64);
  size_t smem = (size_t)kpad * 128 + (size_t)kpad * 12;
  if (smem > 220 * 1024) return PTV3_EUNSUPPORTED;
  static size_t configured_dev[kMaxDeviceSlots] = {};
  size_t& configured = configured_dev[device_slot()];
  if (smem > configured) {
    if (cudaFuncSetAttribute(attn_bwd_mma_kernel<8, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) !=
            cudaSuccess ||
        cudaFuncSetAttribute(attn_bwd_mma_kernel<16, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) !=
            cudaSuccess)
      return PTV3_ELAUNCH;
    configured = smem;
  }
  if ((int64_t)n_windows_max * h > 2147483647LL) return PTV3_EINVAL;
  dim3 grid((unsigned)(n_windows_max * h));
  // one CTA per SM (140 KB of staged operands): 16 warps hide the ldmatrix / MUFU latency better than 8
  if (g_attn_bwd_warps == 8)
    attn_bwd_mma_kernel<8, 4><<<grid, 256, smem, st>>>(qkv, out, dout, order, windows, n_windows, h, scale,
                                                    scale * 1.4426950408889634f, kpad, dqkv, lse);
  else
    attn_bwd_mma_kernel<16, 2><<<grid, 512, smem, st>>>(qkv, out, dout, order, windows, n_windows, h, scale,
                                                     scale * 1.4426950408889634f, kpad, dqkv, lse);
  return check_launch();
}

}  // namespace ptv3
