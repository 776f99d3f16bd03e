// A8: one C-ABI call per PT-v3m1 Block (inference, bf16 operands): the launch chain of Block.forward
// (point_transformer_v3m1_base.py:318-338: xCPE conv -> Linear -> LN -> +x -> LN -> qkv -> window attention -> proj -> +x
// -> LN -> MLP -> +x) issued natively instead of ~15 Python / ctypes round trips (a single scan cost
// 7.4 ms on the host for ~2 ms of device work).  The kernels are the library's own entry points; the dense Linears are
// plain library GEMMs (cuBLASLt, bias in the epilogue -- what torch.nn.functional.linear runs), resolved with dlopen so
// that the library has no link-time dependency on cuBLAS (it picks up the copy torch already loaded).
#include <cublasLt.h>
#include <dlfcn.h>

#include <map>
#include <mutex>
#include <tuple>

#include "common.cuh"

extern "C" {
int ptv3_subm_conv_packed(const void* feat, const void* packed, const float* bias, const int32_t* nbr, int64_t n, int cin,
                          int cout, int k3, void* out, void* stream);
int ptv3_ln_residual(const void* x, int dtype_x, const float* gamma, const float* beta, float eps,
                     const float* residual, float* res_out, const float* gamma2, const float* beta2, void* out2,
                     int dtype_out2, int64_t n, int c, void* stream);
size_t ptv3_attn_window_workspace(int64_t n_windows_max, int h);
int ptv3_attn_window_fwd(const void* qkv, int dtype, const int64_t* order, const int64_t* windows,
                         const int64_t* n_windows, int64_t n_windows_max, int64_t max_len, int64_t n, int h, int d,
                         float scale, int algo, void* out, void* workspace, size_t workspace_bytes, void* stream);
int ptv3_mlp_fused(const void* x, const void* w1_packed, const float* b1, const void* w2_packed, const float* b2,
                   const float* residual, int64_t n, int c, int hidden, float* res_out, void* out_bf16, void* stream);
int ptv3_affine_act(const void* x, int dtype_x, const float* scale, const float* shift, int act, int64_t n, int c,
                    float* out32, void* out16, void* stream);
}

namespace ptv3 {

namespace {

struct LtApi {
  void* lib = nullptr;
  decltype(&cublasLtCreate) create = nullptr;
  decltype(&cublasLtMatmulDescCreate) desc_create = nullptr;
  decltype(&cublasLtMatmulDescSetAttribute) desc_set = nullptr;
  decltype(&cublasLtMatrixLayoutCreate) layout_create = nullptr;
  decltype(&cublasLtMatmulPreferenceCreate) pref_create = nullptr;
  decltype(&cublasLtMatmulPreferenceSetAttribute) pref_set = nullptr;
  decltype(&cublasLtMatmulAlgoGetHeuristic) heuristic = nullptr;
  decltype(&cublasLtMatmul) matmul = nullptr;
  bool ok = false;
};

struct Plan {
  cublasLtMatmulDesc_t op = nullptr;
  cublasLtMatrixLayout_t a = nullptr, b = nullptr, c = nullptr;
  cublasLtMatmulAlgo_t algo;
  bool valid = false;
};

constexpr size_t kLtWorkspace = 32u << 20;

std::mutex g_mu;
LtApi g_api;
cublasLtHandle_t g_handle[16] = {nullptr};
std::map<std::tuple<int, int64_t, int, int>, Plan> g_plans;   // (device, rows, in, out)

bool load_api() {
  if (g_api.ok) return true;
  void* lib = dlopen("libcublasLt.so.12", RTLD_NOW | RTLD_NOLOAD);      // the copy torch loaded, if any
  if (!lib) lib = dlopen("libcublasLt.so.12", RTLD_NOW | RTLD_GLOBAL);
  if (!lib) lib = dlopen("libcublasLt.so", RTLD_NOW | RTLD_GLOBAL);
  if (!lib) return false;
  g_api.lib = lib;
#define PTV3_SYM(field, name)                                              \
  g_api.field = reinterpret_cast<decltype(g_api.field)>(dlsym(lib, name)); \
  if (!g_api.field) return false;
  PTV3_SYM(create, "cublasLtCreate")
  PTV3_SYM(desc_create, "cublasLtMatmulDescCreate")
  PTV3_SYM(desc_set, "cublasLtMatmulDescSetAttribute")
  PTV3_SYM(layout_create, "cublasLtMatrixLayoutCreate")
  PTV3_SYM(pref_create, "cublasLtMatmulPreferenceCreate")
  PTV3_SYM(pref_set, "cublasLtMatmulPreferenceSetAttribute")
  PTV3_SYM(heuristic, "cublasLtMatmulAlgoGetHeuristic")
  PTV3_SYM(matmul, "cublasLtMatmul")
#undef PTV3_SYM
  g_api.ok = true;
  return true;
}

// y (rows, out) = x (rows, in) W^T (out, in) + bias, all bf16 row-major, fp32 accumulation: C^T = W x^T in cuBLAS's
// column-major terms (A = W with op T, B = x, bias along the rows of C^T)
int linear_bf16(const void* x, const void* w, const void* bias, void* y, int64_t rows, int in, int out, void* lt_ws,
                cudaStream_t st) {
  std::lock_guard<std::mutex> lock(g_mu);
  if (!load_api()) return PTV3_EUNSUPPORTED;
  int dev = 0;
  cudaGetDevice(&dev);
  dev &= 15;
  if (!g_handle[dev] && g_api.create(&g_handle[dev]) != CUBLAS_STATUS_SUCCESS) return PTV3_ELAUNCH;
  Plan& p = g_plans[std::make_tuple(dev, rows, in, out)];
  if (!p.valid) {
    if (g_api.desc_create(&p.op, CUBLAS_COMPUTE_32F, CUDA_R_32F) != CUBLAS_STATUS_SUCCESS) return PTV3_ELAUNCH;
    cublasOperation_t ta = CUBLAS_OP_T, tb = CUBLAS_OP_N;
    cublasLtEpilogue_t epi = CUBLASLT_EPILOGUE_BIAS;
    cudaDataType_t bt = CUDA_R_16BF;
    g_api.desc_set(p.op, CUBLASLT_MATMUL_DESC_TRANSA, &ta, sizeof(ta));
    g_api.desc_set(p.op, CUBLASLT_MATMUL_DESC_TRANSB, &tb, sizeof(tb));
    g_api.desc_set(p.op, CUBLASLT_MATMUL_DESC_EPILOGUE, &epi, sizeof(epi));
    g_api.desc_set(p.op, CUBLASLT_MATMUL_DESC_BIAS_DATA_TYPE, &bt, sizeof(bt));
    if (g_api.layout_create(&p.a, CUDA_R_16BF, in, out, in) != CUBLAS_STATUS_SUCCESS ||
        g_api.layout_create(&p.b, CUDA_R_16BF, in, rows, in) != CUBLAS_STATUS_SUCCESS ||
        g_api.layout_create(&p.c, CUDA_R_16BF, out, rows, out) != CUBLAS_STATUS_SUCCESS)
      return PTV3_ELAUNCH;
    cublasLtMatmulPreference_t pref = nullptr;
    if (g_api.pref_create(&pref) != CUBLAS_STATUS_SUCCESS) return PTV3_ELAUNCH;
    size_t ws = kLtWorkspace;
    g_api.pref_set(pref, CUBLASLT_MATMUL_PREF_MAX_WORKSPACE_BYTES, &ws, sizeof(ws));
    // the bias pointer takes part in the heuristic (alignment): set a representative one
    g_api.desc_set(p.op, CUBLASLT_MATMUL_DESC_BIAS_POINTER, &bias, sizeof(bias));
    cublasLtMatmulHeuristicResult_t res;
    int found = 0;
    if (g_api.heuristic(g_handle[dev], p.op, p.a, p.b, p.c, p.c, pref, 1, &res, &found) != CUBLAS_STATUS_SUCCESS ||
        found < 1)
      return PTV3_EUNSUPPORTED;
    p.algo = res.algo;
    p.valid = true;
  }
  g_api.desc_set(p.op, CUBLASLT_MATMUL_DESC_BIAS_POINTER, &bias, sizeof(bias));
  const float one = 1.f, zero = 0.f;
  if (g_api.matmul(g_handle[dev], p.op, &one, w, p.a, x, p.b, &zero, y, p.c, y, p.c, &p.algo, lt_ws, kLtWorkspace, st) !=
      CUBLAS_STATUS_SUCCESS)
    return PTV3_ELAUNCH;
  g_launch_count += 1;
  return PTV3_OK;
}

struct Arena {
  char* base;
  size_t cap, used = 0;
  void* take(size_t bytes) {
    bytes = align_up(bytes);
    if (!base || used + bytes > cap) { used = cap + 1; return nullptr; }
    void* r = base + used;
    used += bytes;
    return r;
  }
  bool ok() const { return used <= cap; }
};

}  // namespace
}  // namespace ptv3

using namespace ptv3;

extern "C" {

size_t ptv3_block_workspace(const ptv3_block_params* p, int64_t n, int64_t n_windows_max) {
  if (!p || n < 0) return 0;
  const size_t nc2 = align_up((size_t)(n > 0 ? n : 1) * p->c * 2), nc4 = align_up((size_t)(n > 0 ? n : 1) * p->c * 4);
  size_t total = 6 * nc2 /* y1 y2 h a pr h2 */ + 3 * nc2 /* qkv */ + 2 * nc4 /* x1 x2 */ + align_up(kLtWorkspace) +
                 align_up(ptv3_attn_window_workspace(n_windows_max, p->heads));
  if (!p->mlp_fused) total += align_up((size_t)(n > 0 ? n : 1) * p->hidden * 2) + nc2;   // f1, f2
  return total + 1024;
}

int ptv3_block_fwd(const ptv3_block_params* p, const float* x, const void* conv_in, const int32_t* nbr,
                   const int64_t* order, const int64_t* windows, const int64_t* n_windows, int64_t n_windows_max,
                   int64_t n, float* x_out, void* x_out_bf16, void* workspace, size_t workspace_bytes, void* stream) {
  if (!p || n < 0 || p->c < 32 || p->c % 16 || p->heads * 16 != p->c || p->hidden < p->c) return PTV3_EINVAL;
  if (n == 0) return PTV3_OK;
  if (!x || !conv_in || !nbr || !windows || !n_windows || !x_out || !x_out_bf16) return PTV3_EINVAL;
  const int c = p->c;
  Arena ar{static_cast<char*>(workspace), workspace_bytes};
  const size_t nc2 = (size_t)n * c * 2, nc4 = (size_t)n * c * 4;
  void* y1 = ar.take(nc2);
  void* y2 = ar.take(nc2);
  void* h = ar.take(nc2);
  void* qkv = ar.take(3 * nc2);
  void* a = ar.take(nc2);
  void* pr = ar.take(nc2);
  void* h2 = ar.take(nc2);
  float* x1 = static_cast<float*>(ar.take(nc4));
  float* x2 = static_cast<float*>(ar.take(nc4));
  void* lt_ws = ar.take(kLtWorkspace);
  const size_t attn_ws_bytes = ptv3_attn_window_workspace(n_windows_max, p->heads);
  void* attn_ws = ar.take(attn_ws_bytes);
  void *f1 = nullptr, *f2 = nullptr;
  if (!p->mlp_fused) {
    f1 = ar.take((size_t)n * p->hidden * 2);
    f2 = ar.take(nc2);
  }
  if (!ar.ok()) return PTV3_EWORKSPACE;
  cudaStream_t st = as_stream(stream);
  int rc;
#define PTV3_TRY(call) \
  if ((rc = (call)) != PTV3_OK) return rc
  // xCPE: SubMConv3d -> Linear -> LayerNorm, + shortcut, then the pre-norm of the attention (:318-326)
  PTV3_TRY(ptv3_subm_conv_packed(conv_in, p->conv_w_packed, p->conv_bias, nbr, n, c, c, p->k3, y1, stream));
  PTV3_TRY(linear_bf16(y1, p->cpe_lin_w, p->cpe_lin_b, y2, n, c, c, lt_ws, st));
  PTV3_TRY(ptv3_ln_residual(y2, PTV3_BF16, p->cpe_ln_g, p->cpe_ln_b, p->eps_cpe, x, x1, p->norm1_g, p->norm1_b, h,
                            PTV3_BF16, n, c, stream));
  // attention (:327-331)
  PTV3_TRY(linear_bf16(h, p->qkv_w, p->qkv_b, qkv, n, c, 3 * c, lt_ws, st));
  PTV3_TRY(ptv3_attn_window_fwd(qkv, PTV3_BF16, order, windows, n_windows, n_windows_max, p->patch, n, p->heads, 16,
                                p->scale, p->attn_algo, a, attn_ws, attn_ws_bytes, stream));
  if (p->mlp_fused && p->proj_w_packed && p->proj_b_f32) {
    // proj -> + shortcut -> LayerNorm -> MLP -> + shortcut in ONE kernel (:329-337), C = 32 / 64
    PTV3_TRY(ptv3_proj_mlp_fused(a, p->proj_w_packed, p->proj_b_f32, x1, p->drop_attn, p->norm2_g, p->norm2_b,
                                 p->eps_norm2, p->fc1_w, static_cast<const float*>(p->fc1_b), p->fc2_w,
                                 static_cast<const float*>(p->fc2_b), p->drop_mlp, n, c, p->hidden, x_out, x_out_bf16,
                                 stream));
    return PTV3_OK;
  }
  PTV3_TRY(linear_bf16(a, p->proj_w, p->proj_b, pr, n, c, c, lt_ws, st));
  PTV3_TRY(ptv3_ln_residual_rs(pr, PTV3_BF16, nullptr, nullptr, p->eps_norm2, p->drop_attn, x1, x2, p->norm2_g,
                               p->norm2_b, h2, PTV3_BF16, n, c, stream));
  // MLP + shortcut (:332-337)
  if (p->mlp_fused) {
    PTV3_TRY(ptv3_mlp_fused_rs(h2, p->fc1_w, static_cast<const float*>(p->fc1_b), p->fc2_w,
                               static_cast<const float*>(p->fc2_b), x2, p->drop_mlp, n, c, p->hidden, x_out, x_out_bf16,
                               stream));
  } else {
    PTV3_TRY(linear_bf16(h2, p->fc1_w, p->fc1_b, f1, n, c, p->hidden, lt_ws, st));
    PTV3_TRY(ptv3_affine_act(f1, PTV3_BF16, nullptr, nullptr, 1, n, p->hidden, nullptr, f1, stream));   // exact GELU
    PTV3_TRY(linear_bf16(f1, p->fc2_w, p->fc2_b, f2, n, p->hidden, c, lt_ws, st));
    PTV3_TRY(ptv3_ln_residual_rs(f2, PTV3_BF16, nullptr, nullptr, 1e-5f, p->drop_mlp, x2, x_out, nullptr, nullptr,
                                 x_out_bf16, PTV3_BF16, n, c, stream));
  }
#undef PTV3_TRY
  return PTV3_OK;
}

}  // extern "C"
