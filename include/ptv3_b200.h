/*
 * ptv3_b200.h -- flat C ABI of libptv3_b200.so: the PTv3 backbone hot path as
 * hand-written sm_100a CUDA.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in `_host`;
 *   - the library never allocates or frees device memory, never synchronises
 *     the device and never touches a stream other than the one passed in
 *     (`stream` is a cudaStream_t passed as void*; 0 = legacy default stream);
 *   - scratch memory comes from the caller: `ptv3_<op>_workspace()` returns the
 *     bytes `ptv3_<op>()` needs for the same sizes (256-byte aligned pointer);
 *   - return value: 0 on success, a negative PTV3_E* code otherwise
 *     (`ptv3_error_string`).  The Python host raises RuntimeError on != 0,
 *     mirroring the reference's assert/exception convention
 *     (pointcept/models/utils/structure.py:59,76,81);
 *   - integer index arrays use int64 where the reference's tensors are int64
 *     (serialized_code/order/inverse, cluster, idx_ptr), int32 for the spconv
 *     style indices / neighbour tables (spconv uses int32 indices).
 * All reference citations are relative to /root/reference.
 */
#ifndef PTV3_B200_H
#define PTV3_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PTV3_OK 0
#define PTV3_EINVAL (-1)     /* bad argument / unsupported shape */
#define PTV3_EWORKSPACE (-2) /* workspace too small */
#define PTV3_ELAUNCH (-3)    /* cudaGetLastError() != cudaSuccess after a launch */
#define PTV3_EUNSUPPORTED (-4)
#define PTV3_ETOOLARGE (-5)   /* size beyond the fixed capacity of an allocation-free (`_static`) entry point */

/* element types of floating point feature buffers */
#define PTV3_F32 0
#define PTV3_BF16 1

/* serialization orders, pointcept/models/utils/serialization/default.py:9-24 */
#define PTV3_ORDER_Z 0
#define PTV3_ORDER_Z_TRANS 1
#define PTV3_ORDER_HILBERT 2
#define PTV3_ORDER_HILBERT_TRANS 3

int ptv3_abi_version(void);
const char* ptv3_error_string(int code);
/* number of kernels this library has launched since load (host counter; bench.py's gpu_launches) */
int64_t ptv3_launch_count(void);

/* ---------------------------------------------------------------------------
 * A2  space-filling-curve codes.
 * Replaces pointcept/models/utils/serialization/default.py:9-24 (encode),
 * z_order.py:66-101 (xyz2key) and hilbert.py:91-198 (encode).
 * grid_coord: (n,3) int32 or int64 (grid_is_i64); batch: (n) int64 or NULL.
 * code[r*n + i] = (batch[i] << 3*depth) | curve(order_ids_host[r])(grid[i] & (2^depth-1)).
 * 1 <= depth <= 16, 1 <= n_orders <= 8.
 * ------------------------------------------------------------------------- */
int ptv3_encode(const void* grid_coord, int grid_is_i64, const int64_t* batch, int64_t n, int depth,
                const int* order_ids_host, int n_orders, int64_t* code, void* stream);

/* per-axis max of an (n,3) grid (int64 x 3, device) -- feeds depth = bit_length(max+1),
 * structure.py:73, and sparse_shape = max + 96, structure.py:131-134 */
int ptv3_grid_max(const void* grid_coord, int grid_is_i64, int64_t n, int64_t* out_max3, void* stream);

/* ---------------------------------------------------------------------------
 * A2  row-wise argsort of (rows,n) int64 codes + inverse permutation.
 * Replaces torch.argsort + scatter_ at structure.py:92-99.  LSD radix sort over
 * the low `key_bits` bits only (codes must be >= 0 and < 2^key_bits), stable,
 * so order is defined for duplicate codes as well (ascending index).
 * order[r][j] = index of the j-th smallest code of row r; inverse[r][order[r][j]] = j.
 * ------------------------------------------------------------------------- */
size_t ptv3_argsort_rows_workspace(int64_t rows, int64_t n);
int ptv3_argsort_rows(const int64_t* code, int64_t rows, int64_t n, int key_bits, int64_t* order,
                      int64_t* inverse /* may be NULL */, void* workspace, size_t workspace_bytes, void* stream);

/* ptv3_argsort_rows without a workspace argument and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same inputs, outputs and result (bit-identical) as ptv3_argsort_rows; the
 * scratch is a file-scope `static __device__` array of PTV3_ARGSORT_STATIC_SCRATCH_BYTES (8 MiB) inside the library,
 * held by every GPU that loads it until the process exits.  Limit: rows * n <= PTV3_ARGSORT_STATIC_MAX_ELEMS (262 144
 * codes per call, e.g. the 4 orders of a 65 536-voxel scene) AND ptv3_argsort_rows_workspace(rows, n) <= the array size
 * (only violated by thousands of very short rows, whose per-row histograms dominate); otherwise the call returns
 * PTV3_ETOOLARGE and enqueues nothing.
 * The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine), only enqueues on `stream`.
 * All calls on a device share the one array (it keeps the previous call's contents; nothing in it has to start at
 * zero): calls that OVERLAP on the GPU -- different streams, concurrent graph branches -- race; serialise them. */
#define PTV3_ARGSORT_STATIC_SCRATCH_BYTES (8u << 20)
#define PTV3_ARGSORT_STATIC_MAX_ELEMS 262144
int64_t ptv3_argsort_rows_static_max_elems(void);
int ptv3_argsort_rows_static(const int64_t* code, int64_t rows, int64_t n, int key_bits, int64_t* order,
                             int64_t* inverse /* may be NULL */, void* stream);

/* ---------------------------------------------------------------------------
 * A0  GridSample voxel hashing / unique / representative selection.
 * Replaces pointcept/datasets/transform.py:848-927 (+ fnv_hash_vec :949-963,
 * ravel_hash_vec :930-946).  Contract: within a voxel points are
 * ordered by ascending original index (stable sort); the per-voxel random draw
 * is the input array `rand` (NULL = first point of every voxel).
 *   coord      (r,3) float32
 *   grid_coord (r,3) int64   floor(coord/grid_size) - min   (all points)
 *   key        (r)   uint64  hash of grid_coord.  hash_type selects the hash AND the precision of the quotient
 *                            coord / grid_size (transform.py:850): 0 = fnv, float64 quotient (numpy >= 2);
 *                            1 = ravel, float64; 2 = ravel, float32; 3 = fnv, float32.  float32 is what the
 *                            reference's pinned numpy 1.26.4 computes for float32 coordinates
 *                            (uv_requirements.txt:84) and what the A0' inference voxelisers compute under any numpy
 *                            (tools/OffsetKeyPointPrediction_Qt.py:276-296, tools/pigseg_shared.py:132-146)
 *   idx_sort   (r)   int64   stable argsort of key
 *   inverse    (r)   int64   voxel id of every point (ids in ascending-key order)
 *   count      (r)   int64   first *n_voxels entries valid: points per voxel
 *   idx_unique (r)   int64   first *n_voxels entries valid: selected point per voxel
 *   min_grid   (3)   int64   the subtracted minimum
 *   n_voxels   (1)   int64
 * ------------------------------------------------------------------------- */
size_t ptv3_gridsample_workspace(int64_t r);
int ptv3_gridsample(const float* coord, int64_t r, double grid_size, int hash_type, const int64_t* rand /* NULL ok */,
                    int64_t* grid_coord, uint64_t* key, int64_t* idx_sort, int64_t* inverse, int64_t* count,
                    int64_t* idx_unique, int64_t* min_grid, int64_t* n_voxels, void* workspace,
                    size_t workspace_bytes, void* stream);

/* The same for EVERY scene of a step in one launch chain (scenes = contiguous row ranges, `offset` cumulative, device):
 * the per-scene results of ptv3_gridsample concatenated.  grid_coord / key / min_grid (n_scenes,3) are per scene (each
 * scene shifted by its own minimum); idx_sort / idx_unique hold GLOBAL row indices ordered by (scene, ascending key);
 * inverse (r) = global voxel id; voxel_offset (n_scenes) = cumulative voxel counts; n_voxels (1) = their total.
 * `rand` is indexed by global voxel id.  One stable 64-bit key sort + one scene-id pass instead of a sort per scene. */
size_t ptv3_gridsample_batch_workspace(int64_t r, int64_t n_scenes);
int ptv3_gridsample_batch(const float* coord, int64_t r, const int64_t* offset, int64_t n_scenes, double grid_size,
                          int hash_type, const int64_t* rand /* NULL ok */, int64_t* grid_coord, uint64_t* key,
                          int64_t* idx_sort, int64_t* inverse, int64_t* count, int64_t* idx_unique, int64_t* min_grid,
                          int64_t* voxel_offset, int64_t* n_voxels, void* workspace, size_t workspace_bytes,
                          void* stream);

/* PT-v3m2 GridPooling integer front end (replaces point_transformer_v3m2_sonata.py:393-417: torch.unique(dim=0) over
 * rows [x/stride | batch << 48, y/stride, z/stride], torch.sort(cluster), cumsum(counts)).  grid_coord (n,3) int32 or
 * int64, non-negative; batch (n) int64; `bits` = bits per axis that hold every grid_coord / stride (the parent's
 * serialization depth), `scene_bits` = bits that hold every batch id; 3*bits + scene_bits <= 63.  Outputs: cluster (n)
 * = pooling_inverse, ids in ascending (batch, x, y, z) order like torch.unique; indices (n) = rows sorted by cluster
 * (stable); idx_ptr (n+1 allocated, n_pooled+1 valid) CSR pointer; pooled_grid (n,3) / pooled_batch (n) / indices4
 * (n,4) int32 [b,x,y,z] (first n_pooled rows valid); meta (4) int64 = {n_pooled, max x, max y, max z of the pooled grid}
 * (the one host read of the level: row count + next serialization depth, structure.py:73).  A coordinate outside
 * [0, 2^bits * stride) sets bit 4 of ptv3_device_error(). */
size_t ptv3_grid_pool_workspace(int64_t n);
int ptv3_grid_pool(const void* grid_coord, int grid_is_i64, const int64_t* batch, int64_t n, int stride, int bits,
                   int scene_bits, int64_t* cluster, int64_t* indices, int64_t* idx_ptr, int64_t* pooled_grid,
                   int64_t* pooled_batch, int32_t* indices4, int64_t* meta, void* workspace, size_t workspace_bytes,
                   void* stream);

/* ---------------------------------------------------------------------------
 * A6  SerializedPooling integer front end.
 * Replaces point_transformer_v3m1_base.py:385-412: code >> 3*pool_depth,
 * torch.unique(code[0]) -> cluster/counts, torch.sort(cluster) -> indices,
 * idx_ptr, head_indices, and the argsort of the pooled codes.  No sort is
 * needed: the parent rows are already sorted by every order (prefix property,
 * tests/test_oracle.py::test_code_prefix_property).
 *   code, order (rows,n) int64 of the parent;   shift = 3*pool_depth
 *   cluster (n)      pooled id of every parent point  (= torch.unique inverse)
 *   indices (n)      parent points grouped by cluster (= order[0])
 *   idx_ptr (n+1)    first *n_pooled+1 entries valid (CSR over `indices`)
 *   head    (n)      first *n_pooled valid: indices[idx_ptr[k]]
 *   new_code/new_order/new_inverse (rows, n) with row stride n; first *n_pooled
 *                    entries of each row valid
 *   n_pooled (1) int64
 * ------------------------------------------------------------------------- */
size_t ptv3_pool_front_workspace(int64_t rows, int64_t n);
int ptv3_pool_front(const int64_t* code, const int64_t* order, int64_t rows, int64_t n, int shift, int64_t* cluster,
                    int64_t* indices, int64_t* idx_ptr, int64_t* head, int64_t* new_code, int64_t* new_order,
                    int64_t* new_inverse, int64_t* n_pooled, void* workspace, size_t workspace_bytes, void* stream);

/* gather of the per-point integer attributes of the pooled point set (:426-433):
 * new_grid[k] = grid[head[k]] >> pool_depth ; new_batch[k] = batch[head[k]];
 * spconv indices (m,4) int32 = [batch, x, y, z] (structure.py:137-141). */
int ptv3_pool_gather_ints(const int64_t* grid /* (n,3) */, const int64_t* batch, const int64_t* head, int64_t m,
                          int pool_depth, int64_t* new_grid, int64_t* new_batch, int32_t* new_indices, void* stream);

/* batch id per row from the cumulative offsets (offset2batch, models/utils/misc.py:13-22); n = number of rows
 * (known to the caller from feat / coord, so no host sync is needed) */
int ptv3_offset_to_batch(const int64_t* offset, int64_t n_scenes, int64_t n, int64_t* batch, void* stream);

/* offset (cumulative scene sizes, int64 (n_scenes)) from a sorted batch vector */
int ptv3_batch_to_offset(const int64_t* batch, int64_t n, int64_t n_scenes, int64_t* offset, void* stream);

/* ---------------------------------------------------------------------------
 * torch_scatter.segment_csr / scatter_mean (call sites
 * point_transformer_v3m1_base.py:416-421, keypoint_ptv3.py:44).
 * out[k] = reduce_{j in [indptr[k], indptr[k+1])} src[gather ? gather[j] : j]
 * reduce: 0 sum, 1 mean, 2 max, 3 min; empty segments give 0.
 * src (n_src, c), out (m, c), dtype PTV3_F32 or PTV3_BF16 (fp32 accumulate).
 * ------------------------------------------------------------------------- */
int ptv3_segment_csr(const void* src, int dtype, const int64_t* gather /* NULL ok */, const int64_t* indptr, int64_t m,
                     int64_t c, int reduce, void* out, void* stream);

/* A7  SerializedUnpooling gather-add (:471-482) with the two GELUs fused:
 * skip_act[i] = act(skip[i]); out[i] = skip_act[i] + act(up[inverse[i]]); act = GELU(erf) if gelu else identity.
 * skip (n,c), up (m,c), skip_act (n,c) or NULL in `dtype`; out (n,c) is always fp32 (the residual stream).
 * `gelu` bit 0: apply GELU to both operands; bit 1: the second output holds the SUM in `dtype` instead of act(skip)
 * (PT-v3m2 GridUnpooling, point_transformer_v3m2_sonata.py:498-499, refreshes the sparse tensor with the sum). */
int ptv3_unpool_add(const void* skip, const void* up, int dtype, const int64_t* inverse, int64_t n, int64_t c, int gelu,
                    void* out, void* skip_act, void* stream);

/* ---------------------------------------------------------------------------
 * A5  submanifold convolution (spconv 2.3.6 SubMConv3d semantics restated,
 * call sites point_transformer_v3m1_base.py:278-284,499-506).
 * Neighbour map: indices (n,4) int32 [b,x,y,z]; nbr (n, ksize^3) int32, -1 =
 * no active site at grid+delta; kernel offset index ((dx+r)*k+(dy+r))*k+(dz+r).
 * The map is what the spconv indice_key cache holds (structure.py:137-146).
 * ------------------------------------------------------------------------- */
size_t ptv3_subm_map_workspace(int64_t n);
int ptv3_subm_build_map(const int32_t* indices, int64_t n, int ksize, int32_t* nbr, void* workspace,
                        size_t workspace_bytes, void* stream);

/* out[i,:] = bias + sum_k W[k] in[nbr[i,k],:]
 * weight layout (k^3, c_out, c_in) (= spconv's (out,kD,kH,kW,in) permuted once on the host side),
 * dtype of in/weight/out given by `dtype`; bias fp32 or NULL; accumulation fp32.
 * algo: 0 = auto, 1 = CUDA-core reference kernel, 2 = tcgen05 implicit GEMM (bf16 only). */
size_t ptv3_subm_conv_workspace(int64_t n, int c_in, int c_out, int k3, int dtype, int algo);
int ptv3_subm_conv(const void* in, const void* weight, const float* bias, const int32_t* nbr, int64_t n, int c_in,
                   int c_out, int k3, int dtype, int algo, void* out, void* workspace, size_t workspace_bytes,
                   void* stream);

/* Stem SubMConv3d (:499-511: k = 5, C_in = 4 -> 32, followed by BatchNorm + GELU) with the neighbour search fused
 * into the convolution: no (n, k^3) table is materialised.  in (n,4) fp32, weight (k^3, 32, 4) fp32 (eval BatchNorm
 * already folded in), bias fp32 (32) or NULL, act 0 none / 1 GELU (erf); out32 fp32 (n,32) and / or out16 bf16 (n,32).
 * workspace: ptv3_subm_map_workspace(n) bytes (the hash table). */
int ptv3_subm_stem_fused(const int32_t* indices, int64_t n, int ksize, const float* in, const float* weight,
                         const float* bias, int act, float* out32, void* out16, void* workspace, size_t workspace_bytes,
                         void* stream);

/* The same result (bit-identical) with a block-occupancy neighbour search instead of 125 hash probes per voxel.  Uses
 * what Point.serialization() (structure.py:52-109) has already computed: code_z (n) = the "z" codes of the rows
 * (batch << 3*depth | Morton), order_z (n) = their argsort, depth = the serialization depth (>= 2).  Every 4x4x4 block
 * of cells is one run of the sorted order; a 5^3 window touches at most 8 blocks, so a row costs 8 hash probes of
 * (block -> run start, 64-bit occupancy mask) + one 4-byte load per PRESENT neighbour.  Sites must be unique (bit 8 of
 * ptv3_device_error() otherwise).  workspace: ptv3_subm_stem_z_workspace(n). */
size_t ptv3_subm_stem_z_workspace(int64_t n);
/* ptv3_subm_build_map (k = 3 or 5) through the same block table: identical (n, k^3) table, 8 hash probes per row
 * instead of k^3.  workspace: ptv3_subm_stem_z_workspace(n). */
int ptv3_subm_build_map_z(const int32_t* indices, int64_t n, int ksize, const int64_t* code_z, const int64_t* order_z,
                          int depth, int32_t* nbr, void* workspace, size_t workspace_bytes, void* stream);
int ptv3_subm_stem_fused_z(const int32_t* indices, int64_t n, int ksize, const int64_t* code_z, const int64_t* order_z,
                           int depth, const float* in, const float* weight, const float* bias, int act, float* out32,
                           void* out16, void* workspace, size_t workspace_bytes, void* stream);

/* bf16 tensor-core path with the weight packed once (eval: weights are constants).  `packed` is the
 * shared-memory image of the K-blocked, 128-byte-swizzled weight (ptv3_subm_pack_weight_bytes() bytes, 1024-byte
 * aligned); supported shapes: c_in == 32 or a multiple of 64 (<= 512), c_out in {32,64,128,256,512}, k3 <= 27. */
int ptv3_subm_conv_tc_supported(int c_in, int c_out, int k3);
size_t ptv3_subm_pack_weight_bytes(int c_in, int c_out, int k3);
int ptv3_subm_pack_weight(const void* weight /* (k3,c_out,c_in) bf16 */, int c_in, int c_out, int k3, void* packed,
                          void* stream);
int ptv3_subm_conv_packed(const void* in, const void* packed, const float* bias, const int32_t* nbr, int64_t n,
                          int c_in, int c_out, int k3, void* out, void* stream);

/* ---------------------------------------------------------------------------
 * A3/A4  serialized window attention.
 * Window records replace get_padding_and_inverse (:115-170): int64 (w,4) =
 * (q_start, q_len, k_start, k_len) in serialized-row space; the ragged last
 * window of a scene with n > K attends to the last K rows of the scene.
 * n_windows_max = sum_b max(1, ceil(n_b / K)) is what the caller allocates.
 * ------------------------------------------------------------------------- */
int ptv3_build_windows(const int64_t* offset, int64_t n_scenes, int64_t patch, int64_t* windows,
                       int64_t n_windows_max, int64_t* n_windows, void* stream);

/* out[row(p), h, :] = softmax_j(scale * q[row(p),h,:] . k[row(j),h,:]) v[row(j),h,:]
 * for p in the window's queries, j in the window's keys, row(p) = order ? order[p] : p.
 * qkv (n,3,h,d) packed rows (= Linear(C,3C) output, :184-190 / flash layout :208-215);
 * out (n,h,d) in storage order (the reference's feat[inverse], :216-218). d == 16.
 * algo: 0 auto (bf16 -> 2, fp32 -> 1), 1 CUDA-core fp32-math kernel, 2 first-generation mma.sync kernel (bf16 P),
 * 3 round-1 tcgen05/TMEM kernel (bf16, windows <= 1024, two-pass), 4 the register kernel (operands staged as f16 with
 * the 32-byte swizzle, reference maximum folded into the accumulator, two query tiles per warp),
 * 5 the single-pass tcgen05/TMEM kernel (bf16, windows <= 1024; needs ptv3_attn_window_workspace bytes; units it
 * cannot finish safely are recomputed by kernel 4 inside the same call).
 * A window record that breaks the contract of ptv3_build_windows (k_len > max_len, or queries outside the key range)
 * makes its CTA return without writing and raises the sticky flag read by ptv3_device_error(). */
size_t ptv3_attn_window_workspace(int64_t n_windows_max, int h);
int ptv3_attn_window_fwd(const void* qkv, int dtype, const int64_t* order /* NULL = identity */, const int64_t* windows,
                         const int64_t* n_windows /* device scalar */, int64_t n_windows_max,
                         int64_t max_len /* upper bound of q_len and k_len */, int64_t n, int h, int d, float scale,
                         int algo, void* out, void* workspace /* algo 5 only, else NULL ok */, size_t workspace_bytes,
                         void* stream);

/* ---------------------------------------------------------------------------
 * A8  fused MLP of Block.forward (MLP.forward :225-248 + the residual add :337): one tcgen05 kernel for
 *   h = GELU_erf(x W1^T + b1);  f = h W2^T + b2;  r = residual ? residual + f : f
 *   res_out = r (fp32, optional), out_bf16 = bf16(r) (optional).  x (n,c) bf16; c in {32, 64, 128, 256}, hidden = 4c
 *   (c <= 64: both weights resident in shared memory; c >= 128: 32 KB weight blocks streamed through a 3-slot ring).
 * Weights are packed once with ptv3_pack_kmajor into K-blocked, 128-byte-swizzled shared-memory images:
 *   w1 (hidden, c) with tile_rows = ptv3_mlp_fc1_tile_rows(c);  w2 (c, hidden) with tile_rows = c.
 * ------------------------------------------------------------------------- */
size_t ptv3_pack_kmajor_bytes(int rows, int k);
int ptv3_pack_kmajor(const void* w /* (rows,k) bf16 */, int rows, int k, int tile_rows, void* packed, void* stream);
int ptv3_mlp_fused_supported(int c, int hidden);
/* tile_rows to pack fc1's weight with (ptv3_pack_kmajor): the hidden-block width of the kernel for this c (128; 64 at c = 256) */
int ptv3_mlp_fc1_tile_rows(int c);
int ptv3_mlp_fused(const void* x, const void* w1_packed, const float* b1, const void* w2_packed, const float* b2,
                   const float* residual /* NULL ok */, int64_t n, int c, int hidden, float* res_out /* NULL ok */,
                   void* out_bf16 /* NULL ok */, void* stream);
/* The same with DropPath on the MLP branch: (fc2 out + b2) * row_scale[row] before the residual add (NULL = none). */
int ptv3_mlp_fused_rs(const void* x, const void* w1_packed, const float* b1, const void* w2_packed, const float* b2,
                      const float* residual, const float* row_scale, int64_t n, int c, int hidden, float* res_out,
                      void* out_bf16, void* stream);
/* The second half of Block.forward (:329-337) in one tcgen05 kernel (c in {32, 64}, hidden = 4c, bf16 operands):
 *   x2 = x1 + rs_attn * (a Wproj^T + bproj);  h = LayerNorm(x2; ln_gamma, ln_beta, eps);
 *   res_out = x2 + rs_mlp * (GELU_erf(h W1^T + b1) W2^T + b2);  out_bf16 = bf16(res_out) (optional)
 * a (n,c) bf16 = attention output, x1 (n,c) fp32 = residual stream after the xCPE branch, rs_* fp32 (n) DropPath row
 * scales or NULL; wproj_packed = ptv3_pack_kmajor of (c,c) with tile_rows = c, bproj / LayerNorm parameters fp32.
 * res_out (n,c) fp32 is required and must not alias x1 (x2 is parked in it between the two epilogues). */
int ptv3_proj_mlp_fused(const void* a, const void* wproj_packed, const float* bproj, const float* x1,
                        const float* rs_attn, const float* ln_gamma, const float* ln_beta, float eps,
                        const void* w1_packed, const float* b1, const void* w2_packed, const float* b2,
                        const float* rs_mlp, int64_t n, int c, int hidden, float* res_out, void* out_bf16,
                        void* stream);

/* ---------------------------------------------------------------------------
 * A8  one call per Block (inference, bf16 operands): the whole launch chain of Block.forward
 * (point_transformer_v3m1_base.py:318-338) issued natively -- xCPE SubMConv3d (tcgen05 kernel) -> Linear -> LayerNorm ->
 * + shortcut -> LayerNorm -> qkv Linear -> window attention -> proj Linear -> + shortcut -> LayerNorm -> MLP -> + shortcut.
 * The Linears are plain library GEMMs (cuBLASLt, bias epilogue, resolved with dlopen at first use); everything else is
 * this library's kernels.  Replaces ~15 host round trips per Block by one (a single scan was host-bound).
 *   x          (n,c) fp32   residual stream entering the Block
 *   conv_in    (n,c) bf16   what the xCPE conv reads (the sparse tensor's features: equals bf16(x) except right after
 *                           an unpooling, where the reference leaves the skip branch there, :476-477)
 *   x_out      (n,c) fp32, x_out_bf16 (n,c) bf16   residual stream leaving the Block and its bf16 copy
 * Weights: bf16 row-major (out, in) with bf16 biases for the Linears; LayerNorm parameters fp32; conv_w_packed from
 * ptv3_subm_pack_weight; with mlp_fused = 1 (c <= 64) fc1_w / fc2_w are ptv3_pack_kmajor images and fc1_b / fc2_b fp32.
 * ------------------------------------------------------------------------- */
typedef struct ptv3_block_params {
  int c, heads, hidden, k3, patch, mlp_fused, attn_algo;
  float scale, eps_cpe, eps_norm2;
  const void* conv_w_packed;
  const float* conv_bias;
  const void *cpe_lin_w, *cpe_lin_b;
  const float *cpe_ln_g, *cpe_ln_b, *norm1_g, *norm1_b, *norm2_g, *norm2_b;
  const void *qkv_w, *qkv_b, *proj_w, *proj_b;
  const void *fc1_w, *fc1_b, *fc2_w, *fc2_b;
  /* DropPath of the two residual branches for THIS call (timm DropPath in train() mode, :329-335): fp32 (n) row scales,
   * 0 or 1 / keep_prob; NULL = none (eval).  Used by a frozen backbone inside a training step. */
  const float *drop_attn, *drop_mlp;
  /* with mlp_fused = 1: ptv3_pack_kmajor image of the (c,c) projection weight (tile_rows = c) and its fp32 bias; when both
   * are set, proj + shortcut + LayerNorm + MLP + shortcut run as ONE kernel (ptv3_proj_mlp_fused).  NULL = separate
   * launches (cuBLASLt proj, ptv3_ln_residual, ptv3_mlp_fused). */
  const void* proj_w_packed;
  const float* proj_b_f32;
} ptv3_block_params;
size_t ptv3_block_workspace(const ptv3_block_params* p, int64_t n, int64_t n_windows_max);
int ptv3_block_fwd(const ptv3_block_params* p, const float* x, const void* conv_in, const int32_t* nbr,
                   const int64_t* order, const int64_t* windows, const int64_t* n_windows, int64_t n_windows_max,
                   int64_t n, float* x_out, void* x_out_bf16, void* workspace, size_t workspace_bytes, void* stream);

/* development knob (tests only): key 1 = operand-descriptor variant of the tcgen05 attention kernel, 7 = chunk /
 * occupancy variant of the default attention kernel, 8 = kernel of the training forward */
int ptv3_debug_set(int key, int value);

/* Reads AND clears the library's sticky device-side error word, a bit mask (0 = none): 1 = an attention window record
 * violated the ptv3_build_windows contract (its CTA wrote nothing); 2 = a sparse-conv site lies outside the packed
 * key range (batch id >= 1024 or a coordinate outside [0, 262128)) and was left out of the neighbour map.
 * 4 = ptv3_grid_pool saw a coordinate outside [0, 2^bits * stride); 8 = ptv3_subm_stem_fused_z found duplicate sites.
 * Synchronises the device (one 4-byte copy): meant for tests and for shims that forward caller-built inputs. */
int ptv3_device_error(void);

/* ---------------------------------------------------------------------------
 * Dense glue kernels of Block.forward (:318-338): fused residual + LayerNorm.
 *   t = gamma ? LayerNorm(x; gamma, beta, eps) : x          (fp32 math)
 *   r = residual ? residual + t : t ;  res_out = r  (fp32, optional)
 *   out2 = gamma2 ? LayerNorm(r; gamma2, beta2, eps) : r    (dtype_out2, optional: pre-norm /
 *          bf16 copy for the next GEMM).  x: dtype_x; residual, res_out: fp32.
 * ------------------------------------------------------------------------- */
int ptv3_ln_residual(const void* x, int dtype_x, const float* gamma, const float* beta, float eps,
                     const float* residual, float* res_out, const float* gamma2, const float* beta2, void* out2,
                     int dtype_out2, int64_t n, int c, void* stream);
/* The same with DropPath on the branch (timm.layers.DropPath as used by Block, point_transformer_v3m1_base.py:329-335):
 * t *= row_scale[row] (fp32 (n): 0 or 1 / keep_prob, NULL = none) before the residual add.  Used when a frozen backbone
 * runs in train() mode (DefaultSegmentorV2(freeze_backbone=True), default.py:58-60) through the fused kernels. */
int ptv3_ln_residual_rs(const void* x, int dtype_x, const float* gamma, const float* beta, float eps,
                        const float* row_scale, const float* residual, float* res_out, const float* gamma2,
                        const float* beta2, void* out2, int dtype_out2, int64_t n, int c, void* stream);

/* ---------------------------------------------------------------------------
 * A6 glue: y = act(x * scale[c] + shift[c]) in one pass (eval BatchNorm + GELU after the pooled segment-max,
 * point_transformer_v3m1_base.py:433-437).  act: 0 none, 1 GELU (erf), 2 ReLU.  x (n,c) fp32 / bf16, c % 4 == 0;
 * scale / shift fp32 (c) or NULL; out32 fp32 (n,c) and / or out16 bf16 (n,c).
 * ------------------------------------------------------------------------- */
int ptv3_affine_act(const void* x, int dtype_x, const float* scale, const float* shift, int act, int64_t n, int c,
                    float* out32, void* out16, void* stream);

/* ---------------------------------------------------------------------------
 * A9  keypoint aggregation of the offset head (tools/infer_offset.py:366-399, the evaluator's
 * hooks/offset_keypoint_evaluator.py:46-91 uses the same rule): per scene b and keypoint k
 *   votes[i,k] = coord[i] + pred[i,k,0:3],  conf[i,k] = pred[i,k,3]   (model units; the caller applies scale/centroid)
 *   mode 0 (argmax):   out[b,k] = votes[i*,k], i* = first argmax_i conf[i,k] over the scene's rows
 *   mode 1 (weighted): confidence-weighted mean of the votes with conf > mask_thresh; argmax vote if there is none.
 * coord (n,3) fp32, pred (n,k_pts,4) fp32, offset int64 (n_scenes) cumulative, out (n_scenes,k_pts,3) fp32.
 * ------------------------------------------------------------------------- */
int ptv3_keypoint_aggregate(const float* coord, const float* pred, const int64_t* offset, int64_t n_scenes, int k_pts,
                            int mode, float mask_thresh, float* out, void* stream);

/* ---------------------------------------------------------------------------
 * A10  validation metric of the offset head (hooks/offset_keypoint_evaluator.py:46-91): per scene b
 * and keypoint k
 *   gt[b,k]    = mean over the scene's rows with target[i,k,3] > 0 of coord[i] + target[i,k,0:3]
 *   kp[b,k]    = coord[i*] + pred[i*,k,0:3], i* = first argmax_i pred[i,k,3]
 *   dist[b,k]  = |kp - gt|_2 * scale[b]   (scale NULL = 1),  valid[b,k] = 1;  no masked row: dist = 0, valid = 0 (skipped).
 * coord (n,3), pred / target (n,k_pts,4) fp32, offset int64 (n_scenes) cumulative, scale fp32 (n_scenes) or NULL,
 * dist fp32 / valid int32 (n_scenes,k_pts).
 * acc: NULL, or (2 + 2*k_pts) doubles [total_dist, total_samples, dist_per_kp[k_pts], samples_per_kp[k_pts]] that the
 * call UPDATES in the evaluator's own order (scene-major; per scene total_dist += mean of its valid distances) -- the
 * running sums of a validation epoch stay on the device (the reference reads every distance with .item()).
 * workspace: NULL (one CTA per scene and keypoint) or ptv3_keypoint_eval_workspace(n_rows, n_scenes, k_pts) bytes: each
 * scene is then split over enough CTAs to fill the 148 SMs (contiguous whole-row reads, partials combined in a fixed order;
 * the query returns 0 when the batch is too small to split).  Both forms give the same argmax row and valid flags.
 * ------------------------------------------------------------------------- */
size_t ptv3_keypoint_eval_workspace(int64_t n_rows, int64_t n_scenes, int k_pts);
int ptv3_keypoint_eval(const float* coord, const float* pred, const float* target, const int64_t* offset,
                       const float* scale, int64_t n_rows, int64_t n_scenes, int k_pts, float* dist, int32_t* valid,
                       double* acc, void* workspace, size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------
 * backward of the ops that carry gradients (training step, engines/train.py:184-245 drives
 * loss.backward() through spconv / flash_attn / torch_scatter autograd functions in the reference).
 * All gradients are fp32; forward operands are `dtype` (PTV3_F32 / PTV3_BF16).
 *
 * SubMConv3d weight gradient: dweight[k][o][c] += sum_i dout[i][o] * in[nbr[i,k]][c]  (accumulates; the caller
 * zeroes it).  The data gradient is the forward conv of `dout` with weights W'[k] = W[k3-1-k]^T (the neighbour
 * map is symmetric), i.e. ptv3_subm_conv / ptv3_subm_conv_packed again.
 * ------------------------------------------------------------------------- */
int ptv3_subm_wgrad(const void* in, const void* dout, const int32_t* nbr, int64_t n, int c_in, int c_out, int k3,
                    int dtype, float* dweight /* (k3,c_out,c_in) */, void* stream);

/* Window attention backward (flash_attn_varlen_qkvpacked_func's backward, point_transformer_v3m1_base.py:208-215).
 * qkv (n,3,h,16), out / dout (n,h,16) in `dtype`; dqkv (n,3,h,16) fp32, ZEROED by the caller (the K / V parts are
 * accumulated: the last window of a scene shares keys with the one before it).
 * `lse` (n,h) fp32 or NULL: the rows' log-sum-exp in log2 units of the scaled scores as written by
 * ptv3_attn_window_fwd_lse (bf16 only; PTV3_EUNSUPPORTED otherwise); NULL = recomputed inside the backward. */
int ptv3_attn_window_fwd_lse(const void* qkv, int dtype, const int64_t* order, const int64_t* windows,
                             const int64_t* n_windows, int64_t n_windows_max, int64_t max_len, int64_t n, int h, int d,
                             float scale, void* out, float* lse, void* stream);
size_t ptv3_attn_window_bwd_workspace(int64_t n, int h);
int ptv3_attn_window_bwd(const void* qkv, const void* out, const void* dout, int dtype, const int64_t* order,
                         const int64_t* windows, const int64_t* n_windows, int64_t n_windows_max, int64_t max_len,
                         int64_t n, int h, int d, float scale, const float* lse /* NULL ok */, float* dqkv,
                         void* workspace, size_t workspace_bytes, void* stream);

/* torch_scatter.segment_csr backward (:416-421): dsrc[gather[j]] for every j of segment k from dout[k];
 * max / min give the gradient to the first row (CSR order) that equals the reduced value. */
int ptv3_segment_csr_bwd(const float* dout, const void* src, const void* out, int dtype,
                         const int64_t* gather /* NULL ok */, const int64_t* indptr, int64_t m, int64_t c, int reduce,
                         float* dsrc, void* stream);

/* LayerNorm backward (Block.forward :318-338): x (n,c) in dtype_x, dy (n,c) in dtype_dy, gamma fp32 (c);
 * dx fp32 (n,c); dgamma / dbeta fp32 (c) are ACCUMULATED (the caller zeroes them).  c in {32,64,128,256,512},
 * otherwise PTV3_EUNSUPPORTED (the host falls back to torch's LayerNorm for that module). */
int ptv3_layer_norm_bwd(const void* x, int dtype_x, const void* dy, int dtype_dy, const float* gamma, float eps,
                        int64_t n, int c, float* dx, float* dgamma, float* dbeta, void* stream);

/* ---------------------------------------------------------------------------
 * SURVEY 8f rank 2 (SpUNet-v1m1, sparse_unet/spconv_unet_v1m1_base.py:137-213): index pairs of
 * spconv.SparseConv3d(kernel 2, stride 2) and of the SparseInverseConv3d that shares its indice_key.
 * The coarse sites come from ptv3_pool_front on a z-order code (code >> 3 partitions like (batch, grid >> 1));
 * this call turns `cluster` (fine row -> coarse row) into the two gather tables the conv kernels consume:
 *   nbr_fwd (m,8): nbr_fwd[c][k] = fine row of coarse site c at kernel offset k = ((x&1)*2 + (y&1))*2 + (z&1), -1 absent
 *   nbr_inv (n,8): nbr_inv[i][k(i)] = cluster[i], the other seven entries -1
 * grid (n,3) int64 fine grid coordinates, cluster (n) int64.
 * ------------------------------------------------------------------------- */
int ptv3_sparse_pair_maps(const int64_t* grid, const int64_t* cluster, int64_t n, int64_t m, int32_t* nbr_fwd,
                          int32_t* nbr_inv, void* stream);

/* Column sums out[j] = sum_i x[i][j] (fp32 accumulate; the bias gradient of a Linear).  x (n,c) fp32 / bf16 with
 * 16-byte rows pieces: c % 8 == 0 (bf16) or c % 4 == 0 (fp32) and at most 256 vectors per row, else PTV3_EUNSUPPORTED.
 * `out` (c) fp32 is overwritten. */
int ptv3_column_sum(const void* x, int dtype, int64_t n, int c, float* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PTV3_B200_H */
