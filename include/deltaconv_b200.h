/* deltaconv_b200 — C ABI of the B200-native (sm_100a) DeltaConv hot path.
 *
 * Conventions (all entry points):
 *   - plain C, no torch types; every pointer is a DEVICE pointer to a contiguous row-major array
 *     unless a leading dimension ("ld", in elements) is given, in which case rows are ld apart;
 *   - float = fp32, indices int32 (ELL) or int64 (edge_index, batch, as PyTorch-Geometric holds them);
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it, nothing synchronises;
 *   - the library allocates nothing: the caller supplies outputs and workspaces (sizes from dc_*_ws_bytes);
 *   - return value: 0 ok, <0 invalid argument, >0 a cudaError_t; dc_last_error() holds the message
 *     (thread-local).
 *
 * Layouts:
 *   idx [N,k] int32   neighbour of point i in slot e (slot 0 = i itself, ascending distance);
 *                     the value N is the padding sentinel of the reference (grad_div_mls.py:226-227)
 *   G, D [N,k,2] f32  grad[2i+a, idx[i,e]] = G[i,e,a];  div[i, 2*idx[i,e]+a] = D[i,e,a]
 *   vector fields [2N,C]: rows (2i, 2i+1) are the (x, y) components of point i (operators.py:7,12-16)
 *   in_ptr [N+1], in_edge [nnz] int32: incoming edges of every point (edge id = i*k+e), ascending.
 *
 * Each declaration cites the reference interface it replaces (paths relative to the reference
 * repository root, rubenwiersma/deltaconv @186fec3).
 */
#ifndef DELTACONV_B200_H
#define DELTACONV_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

const char* dc_last_error(void);
int dc_abi_version(void);
int dc_device_sm_count(void);
/* kernels launched by the library since the last reset */
long long dc_launch_count(void);
void dc_reset_launch_count(void);

/* ---- graph ------------------------------------------------------------------------------- */

/* cloud_ptr[B+1] from the sorted PyG batch vector (deltanet_base.py:44 `data.batch`). */
int dc_cloud_ptr(const int64_t* batch, int N, int B, int32_t* cloud_ptr, void* stream);

/* knn_graph(pos, k, batch, loop=True, flow='target_to_source') — deltanet_base.py:52,63 (torch_cluster).
 * Neighbours ascending by (d2, index), d2 = (dx*dx+dy*dy)+dz*dz in fp32 without contraction.
 * edge_index_or_null: [2, N*k] int64 in PyG layout; d2_or_null: [N,k]. k <= 128. */
int dc_knn(const float* pos, const int32_t* cloud_ptr, int B, int N, int k, int32_t* idx,
           int64_t* edge_index_or_null, float* d2_or_null, void* stream);

/* The same graph for large clouds: a uniform grid per cloud (counting sort by cell) and a ring search per query instead of the
 * O(P^2) scan; exact, so bit-identical to dc_knn.  Worth it from a few thousand points per cloud (deltaconv_b200.ops.knn switches at
 * 1536 points per cloud, at 1000 for k >= 20).  Workspace from dc_knn_grid_ws_bytes(N, B); no host synchronisation. */
size_t dc_knn_grid_ws_bytes(int N, int B);
int dc_set_knn_grid_density(float points_per_bounding_box_cell);   /* tuning knob; 0 (default) = by cloud size: 6 / 3 / 1.5 */
/* tuning: survivors per round from which the top-k list merges instead of inserting (one / two list slots per lane; <= 0 keeps) */
int dc_set_knn_merge_min(int one_slot, int two_slots);   /* tuning knob, defaults 12 and 8 */
int dc_knn_grid(const float* pos, const int32_t* cloud_ptr, int B, int N, int k, int32_t* idx,
                int64_t* edge_index_or_null, float* d2_or_null, void* ws, size_t ws_bytes, void* stream);

/* Sort key for a cache-friendly point order of large clouds: (cloud << 30) | 30-bit Morton code inside the cloud's bounding box.
 * No counterpart in the reference (its point order is whatever the dataset has); ws: B * 24 bytes. */
int dc_morton_keys(const float* pos, const int32_t* cloud_ptr, int B, int N, int64_t* keys, void* ws, size_t ws_bytes, void* stream);

/* geodesicFPS(points, n_samples) — deltaconv/cpp/core.cpp:17-27, sampling.cpp:21-81 (pybind11 module deltaconv_bindings;
 * Python face deltaconv/geometry/fps.py:5-19, transform deltaconv/transforms/geodesic_fps.py:14-43) for a whole batch of clouds:
 * farthest-point sampling by shortest-path distance on the 10-nearest-neighbour graph of every cloud, fp64 like the reference
 * (pos is [N,3] DOUBLE), the same sample sequence bit for bit when started from the same point.
 * start[B]: first sample of every cloud (index inside the cloud; the reference draws it at random, sampling.cpp:33-40);
 * samples[B, n_samples] int32: indices inside the cloud (-1 for an empty cloud); n_samples may exceed a cloud's size (the
 * surplus repeats index 0 as in the reference); dist_or_null[N]: distance of every point to its nearest sample.
 * order_or_null[N]: a permutation of the points that keeps every cloud in its own range, order[t] = the point visited t-th.
 * The result does not depend on it; the time of the neighbour search does: in a space-filling-curve order (dc_morton_keys)
 * most of a cloud is skipped per query, in an arbitrary order (or NULL) every pair of points of a cloud is examined.
 * max_cloud_points: an upper bound on the cloud sizes (N is always valid); clouds up to 18,432 points are sampled from
 * shared memory, larger ones from the workspace.  Clouds of fewer than 11 points use the neighbours they have (the
 * reference throws there).  Workspace from dc_geodesic_fps_ws_bytes(N). */
size_t dc_geodesic_fps_ws_bytes(int N);
int dc_geodesic_fps(const double* pos, const int32_t* cloud_ptr, int B, int N, int max_cloud_points, int n_samples,
                    const int32_t* order_or_null, const int32_t* start, int32_t* samples, double* dist_or_null,
                    void* ws, size_t ws_bytes, void* stream);

/* [2,N*k] int64 edge_index (rows grouped by centre, fixed k — grad_div_mls.py:24-25,224) <-> ELL idx. */
int dc_ell_from_edge_index(const int64_t* edge_index, int N, int k, int32_t* idx, void* stream);
int dc_edge_index_from_ell(const int32_t* idx, int N, int k, int64_t* edge_index, void* stream);

/* Incoming-edge CSR of the graph: the structure torch_sparse caches as CSC for the backward SpMM. */
size_t dc_ell_transpose_ws_bytes(int N);
int dc_ell_transpose(const int32_t* idx, int N, int k, int32_t* in_ptr, int32_t* in_edge,
                     void* ws, size_t ws_bytes, void* stream);

/* ---- tangent frames ------------------------------------------------------------------------ */

/* build_tangent_basis(normal) — grad_div_mls.py:50-69. */
int dc_tangent_basis(const float* normal, int N, float* xb, float* yb, void* stream);

/* estimate_basis(pos, edge_index, k, orientation) — grad_div_mls.py:10-47 (batched 3xk SVD). */
int dc_estimate_basis(const float* pos, const int32_t* idx, int N, int k, const float* orient_or_null,
                      float* normal, float* xb, float* yb, void* stream);

/* ---- operator build ------------------------------------------------------------------------ */

/* build_grad_div(pos, normal, x_basis, y_basis, edge_index, batch, kernel_width, regularizer,
 * normalized, shape_regularizer) — grad_div_mls.py:197-277 with coords_projected (:72-97),
 * gaussian_weights (:100-116), weighted_least_squares (:119-152), fit_vector_mapping (:155-194).
 * Optional debug outputs expose those helpers' results: coords [N,k,2], weights [N,k], mapping [N,k,2,2].
 * d2_or_null [N,k]: the squared neighbour distances of dc_knn / dc_knn_grid (+inf in padded slots); given, the per-point mean
 * edge length (:233, :112) is taken from them instead of gathering the neighbour positions again.
 * normalized: 0 = no per-cloud normalisation (:258-260), 1 = G and D hold the normalised entries, 2 = deferred: G and D hold
 * the entries BEFORE the normalisation and point_scale [N] the factor of every point's cloud (1 / S_b, or 1 where S_b <= 1e-5);
 * the operator applications take it as `row_scale` and multiply their result rows (every neighbour of a point lies in its
 * cloud, and D = g M is linear in g), which saves a read-modify-write pass over all 16 k bytes per point of entries. */
size_t dc_build_ws_bytes(int N, int B);       /* recommended: includes 48 B per point of repacked gather records */
size_t dc_build_ws_min_bytes(int N, int B);   /* accepted minimum (4 B per point + 8 B per cloud): slower scalar gathers */
int dc_set_build_chunk(int points);    /* tuning: points per fit/normalise chunk, 0 = automatic; no reference counterpart */
int dc_build_grad_div(const float* pos, const float* normal, const float* xb, const float* yb,
                      const int32_t* idx, const float* d2_or_null, const int32_t* cloud_ptr, int B, int N, int k,
                      float kernel_width, float lam, float lam_shape, int use_shape, int normalized,
                      float* G, float* D, float* point_scale_or_null,
                      float* dbg_coords, float* dbg_weights, float* dbg_mapping,
                      void* ws, size_t ws_bytes, void* stream);

/* Same build for callers that must not allocate device memory (a host whose memory is partitioned once
 * at start-up, a loop that cannot afford cudaMalloc's device synchronisation): the scratch lives in a
 * fixed array inside the library (~1.03 MB held by every GPU that loads it, until exit).  No counterpart
 * in the reference.  Limits: N <= dc_build_static_max_points() (262144), B <= dc_build_static_max_clouds()
 * (4096); larger batches return an error.  All calls on a GPU share the array: calls that can overlap on
 * the GPU race; nothing has to be cleared between calls.  The first call on each GPU looks the array's
 * address up (cudaGetSymbolAddress) — make it before any stream capture.  The array used is the CURRENT
 * device's: buffers and stream must belong to it (the output buffer is checked).  No debug outputs. */
int dc_build_static_max_points(void);
int dc_build_static_max_clouds(void);
int dc_build_grad_div_static(const float* pos, const float* normal, const float* xb, const float* yb,
                             const int32_t* idx, const int32_t* cloud_ptr, int B, int N, int k,
                             float kernel_width, float lam, float lam_shape, int use_shape, int normalized,
                             float* G, float* D, void* stream);

/* ---- operator application (SparseTensor @ dense and operators.py) --------------------------- */

/* All six entry points below take (cloud_ptr_or_null, B, max_cloud_points).  They are used by an alternative kernel that
 * stages one cloud's rows in shared memory so the k-fold gathers never leave the SM (dc_set_spmm_staged(1); needs the
 * cloud layout and a cloud that fits).  It gives bitwise identical results but measured slower than the default L2-gather
 * kernel on B200, so it is off by default. */
int dc_set_spmm_staged(int v);
/* grad @ x — deltanet_base.py:78, deltaconv.py:66:  out[2i+a,:] = sum_e G[i,e,a] x[idx[i,e],:] */
int dc_grad_fwd(const int32_t* idx, const float* G, int N, int k, int C, const float* x, int64_t ldx,
                float* out, int64_t ldo, const float* row_scale_or_null, const int32_t* cloud_ptr_or_null, int B, int max_cloud_points, void* stream);
/* its transpose (autograd of the above w.r.t. x): dX[j,:] = sum_{(i,e)->j} G[i,e,:] . dV[2i:2i+2,:] */
int dc_grad_bwd(const int32_t* in_ptr, const int32_t* in_edge, const float* G, int N, int k, int C,
                const float* dV, int64_t ld_dv, float* dX, int64_t ld_dx, const float* row_scale_or_null, const int32_t* cloud_ptr_or_null, int B, int max_cloud_points, void* stream);

/* [div @ v | curl(v, div) | norm(v)] in one pass — deltaconv.py:57, operators.py:4-7,23-27.  With x given, x [N,C] is also
 * copied to x_copy (the first column block of the same torch.cat) by the same threads. */
int dc_divcurlnorm_fwd(const int32_t* idx, const float* D, int N, int k, int C, const float* v, int64_t ldv,
                       float* div_or_null, int64_t ld_div, float* curl_or_null, int64_t ld_curl,
                       float* norm_or_null, int64_t ld_norm, const float* x_or_null, int64_t ldx, float* x_copy, int64_t ld_copy,
                       const float* row_scale_or_null, const int32_t* cloud_ptr_or_null, int B, int max_cloud_points, void* stream);
int dc_divcurlnorm_bwd(const int32_t* in_ptr, const int32_t* in_edge, const float* D, int N, int k, int C,
                       const float* d_div_or_null, int64_t ld_ddiv, const float* d_curl_or_null, int64_t ld_dcurl,
                       const float* d_norm_or_null, int64_t ld_dnorm, const float* v, int64_t ldv,
                       float* dV, int64_t ld_dv, const float* row_scale_or_null, const int32_t* cloud_ptr_or_null, int B, int max_cloud_points, void* stream);

/* hodge_laplacian(v, grad, div) given div v and curl v — operators.py:35-46:
 * out = -(grad(div_v) + J grad(curl_v)), rows interleaved like every vector field.  With v given, v [2N,C] is also copied to
 * v_copy (the first column block of torch.cat([v, hodge, grad x]), deltaconv.py:66). */
int dc_hodge_fwd(const int32_t* idx, const float* G, int N, int k, int C, const float* div_v, int64_t ld_a,
                 const float* curl_v, int64_t ld_b, float* out, int64_t ldo, const float* v_or_null, int64_t ldv,
                 float* v_copy, int64_t ld_copy, const float* row_scale_or_null, const int32_t* cloud_ptr_or_null, int B, int max_cloud_points, void* stream);
int dc_hodge_bwd(const int32_t* in_ptr, const int32_t* in_edge, const float* G, int N, int k, int C,
                 const float* d_out, int64_t ld_do, float* d_div, int64_t ld_da, float* d_curl, int64_t ld_db,
                 const float* row_scale_or_null, const int32_t* cloud_ptr_or_null, int B, int max_cloud_points, void* stream);

/* ---- scalar stream: kNN max aggregation ------------------------------------------------------ */

/* scatter(s_mlp_max(x)[edge_index[1]], edge_index[0], reduce='max') — deltaconv.py:54 (torch_scatter).
 * arg [N,C] u8 = lowest slot attaining the maximum (the element that receives the gradient). */
int dc_gathermax_fwd(const int32_t* idx, int N, int k, int C, const float* y, int64_t ldy, float* out, int64_t ldo,
                     uint8_t* arg, void* stream);
int dc_gathermax_bwd(const int32_t* in_ptr, const int32_t* in_edge, int N, int k, int C, const uint8_t* arg,
                     const float* dout, int64_t ldd, float* dy, int64_t ld_dy, void* stream);
/* The same aggregation over leaky_relu(BatchNorm1d(y)) taken from the pre-activations y — scatter(reduce='max') of
 * deltaconv.py:54 with the last block of s_mlp_max (mlp.py:9) folded in; the activation is monotone per channel, so the
 * maximum is the activation of the per-channel max (gamma >= 0) or min (gamma < 0) of y.  Float4 layout only (C % 4 == 0). */
int dc_gathermax_bn_fwd(const int32_t* idx, int N, int k, int C, const float* y, int64_t ldy, const float* mean,
                        const float* invstd, const float* gamma_or_null, const float* beta_or_null, float slope,
                        float* out, int64_t ldo, uint8_t* arg, void* stream);

/* centralized first layer, deltaconv.py:50-52: edge features x[col]-x[row] ([N*k,C], dense) and the
 * max over the k rows of every point after the edge MLP. */
int dc_edge_diff_fwd(const int32_t* idx, int N, int k, int C, const float* x, int64_t ldx, float* out, void* stream);
int dc_edge_diff_bwd(const int32_t* in_ptr, const int32_t* in_edge, int N, int k, int C, const float* d_edge,
                     float* dx, int64_t ld_dx, void* stream);
int dc_segmax_fwd(const float* in, int N, int k, int C, const int32_t* idx_or_null, float* out, int64_t ldo,
                  uint8_t* arg, void* stream);
int dc_segmax_bwd(const float* dout, int64_t ldd, int N, int k, int C, const uint8_t* arg, float* din, void* stream);
/* The last block of an edge MLP folded into the max (deltaconv.py:50-52 with depth >= 2): y [N*k,C] are the pre-activations of
 * that block, out = max over each point's k rows of leaky_relu(BatchNorm1d(y)), selected through the per-channel extreme of y
 * like dc_gathermax_bn_fwd; the activation is never written.  Backward: dout [N,C] -> dy [N*k,C], d_gamma, d_beta, with the
 * BatchNorm sums taken over the arg-max rows only (training != 0: batch statistics over N*k rows).  Float4 layout only;
 * workspace dc_colreduce_ws_bytes(C). */
int dc_segmax_bn_fwd(const float* y, int N, int k, int C, const int32_t* idx_or_null, const float* mean, const float* invstd,
                     const float* gamma_or_null, const float* beta_or_null, float slope, float* out, int64_t ldo, uint8_t* arg,
                     void* stream);
int dc_segmax_bn_bwd(const float* y, int N, int k, int C, const uint8_t* arg, const float* dout, int64_t ldd, const float* mean,
                     const float* invstd, const float* gamma_or_null, const float* beta_or_null, float slope, int training,
                     float* d_gamma, float* d_beta, float* dy, void* ws, size_t ws_bytes, void* stream);

/* Fused form of the same three lines for Cin <= dc_central_max_cin() (4), k <= 64 and a depth-1 edge MLP:
 * scatter_max(LeakyReLU(BatchNorm(Linear(x[col]-x[row]))), row) without any [E,*] temporary.  dc_central_stats gives the
 * batch statistics over the E = N*k edge rows from the moments of the edge offsets (training mode; in eval mode pass the
 * running statistics to dc_central_fwd directly); moments = 20 doubles kept for dc_central_bwd. */
int dc_central_max_cin(void);
size_t dc_central_ws_bytes(int N, int k, int Cin, int Cout);
int dc_central_stats(const int32_t* idx, int N, int k, int Cin, int Cout, const float* x, int64_t ldx, const float* W,
                     float eps, float momentum, double* moments, float* mean, float* invstd,
                     float* running_mean_or_null, float* running_var_or_null, int64_t* num_batches_tracked_or_null,
                     void* ws, size_t ws_bytes, void* stream);
int dc_central_fwd(const int32_t* idx, int N, int k, int Cin, int Cout, const float* x, int64_t ldx, const float* W,
                   const float* gamma_or_null, const float* beta_or_null, const float* mean, const float* invstd, float slope,
                   float* out, int64_t ldo, uint8_t* arg, void* stream);
int dc_central_bwd(const int32_t* idx, const int32_t* in_ptr_or_null, const int32_t* in_edge_or_null, int N, int k, int Cin, int Cout,
                   const float* x, int64_t ldx, const float* W, const float* gamma_or_null, const float* beta_or_null,
                   const float* mean, const float* invstd, const double* moments, float slope, int training,
                   const uint8_t* arg, const float* dout, int64_t ldd, float* dW, float* dgamma, float* dbeta,
                   float* dx_or_null, int64_t ld_dx, void* ws, size_t ws_bytes, void* stream);
/* Depth >= 2 edge MLPs (train_shapenet.py:79-89 builds DeltaNetSegmentation with mlp_depth = 2): the FIRST block's activation
 * h [N*k,Cout] = LeakyReLU(BatchNorm(Linear(x[col]-x[row]))) written once from the coordinates (statistics from dc_central_stats,
 * no [E,Cin] offsets, no pre-activation tensor), and its backward from the dense gradient dh [N*k,Cout] in one pass over dh
 * (dW, dgamma, dbeta; no d/dx).  Same limits and workspace as dc_central_fwd / dc_central_bwd. */
int dc_central_edge_fwd(const int32_t* idx, int N, int k, int Cin, int Cout, const float* x, int64_t ldx, const float* W,
                        const float* gamma_or_null, const float* beta_or_null, const float* mean, const float* invstd, float slope,
                        float* out, void* stream);
int dc_central_edge_bwd(const int32_t* idx, int N, int k, int Cin, int Cout, const float* x, int64_t ldx, const float* W,
                        const float* gamma_or_null, const float* beta_or_null, const float* mean, const float* invstd,
                        const double* moments, float slope, int training, const float* dh, int64_t ldd, float* dW, float* dgamma,
                        float* dbeta, void* ws, size_t ws_bytes, void* stream);

/* ---- BatchNorm1d wrapper + LeakyReLU (nonlin.py:11-35, mlp.py:7-11) -------------------------- */

size_t dc_colreduce_ws_bytes(int C);
/* mode 0: batch statistics of x [M,C]; mode 1: of the norms of a vector field x [2M,C] (nonlin.py:68,77).
 * Writes mean/invstd (biased variance, eps) and updates running stats like torch.nn.BatchNorm1d. */
int dc_bn_stats(const float* x, int64_t ldx, int64_t M, int C, int mode, float eps, float momentum,
                float* mean, float* invstd, float* running_mean_or_null, float* running_var_or_null,
                int64_t* num_batches_tracked_or_null, void* ws, size_t ws_bytes, void* stream);
/* out = leaky_relu(gamma*(y-mean)*invstd+beta, slope) + addend   (addend: `x_max +`, deltaconv.py:59) */
int dc_bn_act_fwd(const float* y, int64_t ldy, int64_t M, int C, const float* mean, const float* invstd,
                  const float* gamma_or_null, const float* beta_or_null, float slope,
                  const float* addend_or_null, int64_t ld_add, float* out, int64_t ldo, void* stream);
int dc_bn_act_bwd(const float* y, int64_t ldy, const float* dout, int64_t ldd, int64_t M, int C,
                  const float* mean, const float* invstd, const float* gamma_or_null, const float* beta_or_null,
                  float slope, int training, float* d_gamma, float* d_beta, float* dy, int64_t ld_dy,
                  void* ws, size_t ws_bytes, void* stream);
int dc_col_sum(const float* x, int64_t ldx, int64_t M, int C, float* out, void* ws, size_t ws_bytes, void* stream);

/* ---- VectorNonLin (nonlin.py:38-86) ------------------------------------------------------------ */
int dc_vecnonlin_fwd(const float* v, int64_t ldv, int64_t M, int C, const float* mean_or_null,
                     const float* invstd_or_null, const float* gamma_or_null, const float* beta_or_null,
                     const float* bias_or_null, float* out, int64_t ldo, void* stream);
int dc_vecnonlin_bwd(const float* v, int64_t ldv, const float* dout, int64_t ldd, int64_t M, int C,
                     const float* mean_or_null, const float* invstd_or_null, const float* gamma_or_null,
                     const float* beta_or_null, const float* bias_or_null, int training,
                     float* d_gamma, float* d_beta, float* dbias_rows_or_null, float* dv, int64_t ld_dv,
                     void* ws, size_t ws_bytes, void* stream);

/* ---- global pools (deltanet_classification.py:46-47, deltanet_segmentation.py:61; PyG scatter) --- */
int dc_pool_fwd(const float* x, int64_t ldx, const int32_t* cloud_ptr, int B, int C, float* out_max_or_null,
                int64_t ld_max, float* out_mean_or_null, int64_t ld_mean, int32_t* arg, void* stream);
int dc_pool_bwd(const int32_t* cloud_ptr, int B, int N, int C, const int32_t* arg, const float* d_max_or_null,
                int64_t ld_dmax, const float* d_mean_or_null, int64_t ld_dmean, float* dx, int64_t ld_dx, void* stream);
/* The embedding layer's BatchNorm + LeakyReLU fused with the pools that consume it (deltanet_classification.py:45-49,
 * deltanet_segmentation.py:58-61: `lin_embedding`/`lin_global` -> global_max_pool / global_mean_pool): the pooled value is
 * leaky_relu(gamma*(y-mean)*invstd+beta) evaluated while y is read, and the backward rebuilds the pooled gradient of each row from
 * the [B,C] gradients, so neither the [N,C] activation nor its gradient is stored.  Training-mode statistics (dc_bn_stats).
 * bwd needs C % 4 == 0 and B <= 65535. */
int dc_bn_act_pool_fwd(const float* y, int64_t ldy, const int32_t* cloud_ptr, int B, int C, const float* mean, const float* invstd,
                       const float* gamma_or_null, const float* beta_or_null, float slope, float* out_max_or_null, int64_t ld_max,
                       float* out_mean_or_null, int64_t ld_mean, int32_t* arg, void* stream);
int dc_bn_act_pool_bwd(const float* y, int64_t ldy, int64_t M, int C, const int32_t* cloud_ptr, int B,
                       const int32_t* arg, const float* d_max_or_null, int64_t ld_dmax, const float* d_mean_or_null, int64_t ld_dmean,
                       const float* mean, const float* invstd, const float* gamma_or_null, const float* beta_or_null, float slope,
                       float* d_gamma, float* d_beta, float* dy, int64_t ld_dy, void* ws, size_t ws_bytes, void* stream);
/* One SGD step (momentum, weight decay; dampening 0, no Nesterov — torch.optim.SGD as experiments/train_*.py configure it) on flat
 * parameter / gradient / momentum buffers of n floats: buf = momentum*buf + grad + weight_decay*p; p -= lr*buf.  lr is a device scalar. */
int dc_sgd_step(float* params, const float* grads, float* momentum_buf, int64_t n, const float* lr_device,
                float momentum, float weight_decay, float grad_scale, void* stream);
/* One Adam step (torch.optim.Adam as experiments/train_shapeseg.py:82 configures it: L2 weight decay, no amsgrad) on flat buffers;
 * the step counter (int32, starts at 0, incremented by the call) and lr are device scalars. */
int dc_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, const float* lr_device,
                 int32_t* step_device, float beta1, float beta2, float eps, float weight_decay, float grad_scale, void* stream);
/* out = a + b on [M,C] column slices of wider row-major buffers: the sum of the two gradients of [div v | curl v], which feeds both
 * the scalar concatenation (deltaconv.py:57) and the Hodge Laplacian of the vector stream (deltaconv.py:66) */
int dc_add_rows(const float* a, int64_t lda, const float* b, int64_t ldb, int64_t M, int C, float* out, int64_t ldo, void* stream);
/* out[i,:] = src[cloud(i),:] — `global_max_pool(x, batch)[batch]`, deltanet_segmentation.py:61,64 */
int dc_bcast_rows(const float* src, int64_t lds, const int32_t* cloud_ptr, int B, int N, int C, float* out, int64_t ldo,
                  void* stream);
/* out[i,:] += src[cloud of i,:]: a Linear over [x_max[batch] | per-point features] (deltanet_segmentation.py:61-66) is the product
 * with the per-point block plus the product of the B pooled rows with their weight columns, added here. */
int dc_bcast_rows_add(const float* src, int64_t lds, const int32_t* cloud_ptr, int B, int N, int C, float* out, int64_t ldo,
                      void* stream);

/* ---- Linear (mlp.py:9,15; heads) ---------------------------------------------------------------- */
/* Precision of the tensor-core (tcgen05) path behind dc_linear_*: 3 = 3xTF32 split, fp32-grade accuracy (default);
 * 1 = plain TF32 (~1e-3 relative); 0 = tensor cores off, exact fp32 FFMA kernel.  Shapes the tensor-core
 * kernel does not take (row pitch not a multiple of 4 floats, M < 64, N < 16, K < 8) always use the FFMA kernel. */
int dc_set_gemm_mode(int mode);
int dc_get_gemm_mode(void);
/* tuning switch: 1 = keep the split A operand in shared memory (SS MMAs) instead of tensor memory (default 0) */
int dc_set_gemm_no_ts(int v);
int dc_set_gemm_keep_raw_hi(int v);
/* tuning switch: 1 (default) = GEMM epilogue through shared memory and TMA stores, 0 = row-per-lane direct stores */
int dc_set_gemm_tma_store(int v);
/* tuning switch: 1 (default) = CTA pairs (tcgen05 cta_group::2: 256-row tiles, each CTA holds half of the B tile), 0 = one CTA per tile */
int dc_set_gemm_pair(int v);
/* debug: device buffer of 8 x 256 int64 that CTA 0 of the pair kernel fills with clock stamps of its pipeline events (NULL = off) */
int dc_set_gemm_trace(void* device_buffer_or_null);
/* measurement only: extra copies of every k-block's MMAs (wide: N = 256 per instruction); products are garbage while set */
int dc_set_gemm_debug(int extra_mma, int wide);
size_t dc_gemm_ws_bytes(int64_t M, int64_t N, int64_t K);
/* The tensor-core product itself, C[M,N] = op(A) op(B)^T: a_mn/b_mn = 0 when the reduction index is contiguous
 * ([M,K] / [N,K] storage), 1 when the M / N index is contiguous ([K,M] / [K,N] storage).  Returns 1 if the shape is
 * not eligible.  Used by dc_linear_fwd (0,0), dc_linear_bwd_input (0,1) and dc_linear_bwd_weight (1,1). */
int dc_gemm_tc_raw(const float* A, int64_t lda, int a_mn, const float* B, int64_t ldb, int b_mn,
                   const float* bias_or_null, int64_t M, int N, int64_t K, float* C, int64_t ldc,
                   void* ws, size_t ws_bytes, void* stream);
int dc_linear_fwd(const float* X, int64_t ldx, const float* W, int64_t ldw, const float* bias_or_null,
                  int64_t M, int K, int Nout, float* Y, int64_t ldy, void* ws, size_t ws_bytes, void* stream);
int dc_linear_bwd_input(const float* dY, int64_t ld_dy, const float* W, int64_t ldw, int64_t M, int K, int Nout,
                        float* dX, int64_t ld_dx, void* ws, size_t ws_bytes, void* stream);
int dc_linear_bwd_weight(const float* dY, int64_t ld_dy, const float* X, int64_t ldx, int64_t M, int K, int Nout,
                         float* dW, int64_t ld_dw, void* ws, size_t ws_bytes, void* stream);
/* I_J folded into the weights (operators.py:19-21, deltaconv.py:68): W [Cout,2K] -> [[W1,-W2],[W2,W1]] as [2Cout, 2*K_padded].
 * K_padded > K: the operand's column blocks are padded to multiples of 4 floats; col_map [K] (device, nullable = identity)
 * gives the padded column of every source column, the remaining columns of W_expanded are zero. */
int dc_expand_vec_weight(const float* W, int Cout, int K, int K_padded, const int32_t* col_map_or_null,
                         float* W_expanded, void* stream);
int dc_fold_vec_weight_grad(const float* dW_expanded, int Cout, int K, int K_padded, const int32_t* col_map_or_null,
                            float* dW, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DELTACONV_B200_H */
