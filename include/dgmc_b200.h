/* libdgmc_b200 — C ABI of the B200 (sm_100a) kernels behind `dgmc.models.DGMC`.
 *
 * This is the drop-in boundary for the reference's correspondence-and-consensus path
 * (/root/reference/dgmc/models/dgmc.py:114-311 and the psi backbones rel.py / spline.py).
 * The reference has no native code of its own: every entry point below replaces a
 * THIRD-PARTY native routine the reference reaches from Python (ATen, torch_scatter,
 * torch_spline_conv, KeOps); the comment on each one cites the reference call site.
 *
 * Conventions
 *  - all pointers are DEVICE pointers on the current device unless marked "host";
 *  - values are fp32, indices int32 (the Python host converts int64 at the boundary),
 *    masks are uint8 (0/1); matrices are row-major, `ld*` = row stride in elements;
 *  - every call is asynchronous on `stream` (a cudaStream_t), performs no hidden sync and
 *    NEVER allocates or frees: outputs and workspaces belong to the caller
 *    (`*_workspace()` returns the bytes a call needs);
 *  - return value 0 on success, negative on failure (DGMC_ERR_*); `dgmc_last_error()`
 *    returns a thread-local description.  No exceptions cross the boundary.
 */
#ifndef DGMC_B200_H
#define DGMC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DGMC_B200_ABI_VERSION 1

typedef struct CUstream_st* dgmc_stream_t; /* == cudaStream_t */

const char* dgmc_last_error(void);
int dgmc_abi_version(void);
/* number of CUDA kernels this library has launched in this process (bench.py's gpu_launches) */
uint64_t dgmc_launch_count(void);

/* ---- K1: streaming top-k on the tensor cores -----------------------------------------------
 * idx_out[i, 0:k] = columns j of the k largest <h_s[i,:], h_t[j,:]>, descending by exact fp32
 * score, ties by ascending j; score_out (nullable) those fp32 scores.  One graph pair per call.
 * replaces DGMC.__top_k__ (dgmc.py:85-94: KeOps argKmin, else broadcast-mul-sum + topk).
 * bf16 tcgen05 GEMM fed by TMA with a fused per-row selection of k' = 16 or 32 candidates in the
 * epilogue (the N_s x N_t scores never reach HBM), stream-K work split, then an fp32 re-scoring of
 * the k' survivors.  Supported: c <= 512, 1 <= k <= 26, k <= n_t (dgmc_topk_tc_supported() != 0).
 * The workspace (bf16 operand copies + partial candidate lists) must be 256-byte aligned. */
int dgmc_topk_tc_supported(int n_s, int n_t, int c, int k);
size_t dgmc_topk_tc_workspace(int n_s, int n_t, int c, int k);
int dgmc_topk_tc(const float* h_s, int64_t lds, const float* h_t, int64_t ldt, int n_s, int n_t, int c,
                 int k, int32_t* idx_out, float* score_out, void* workspace, size_t workspace_bytes,
                 dgmc_stream_t stream);

/* ---- K1 (small shapes): exact fp32 top-k, one warp per source row ------------------------
 * idx_out[b, i, 0:k] = columns j of the k largest <h_s[b,i,:], h_t[b,j,:]>, descending, ties by
 * ascending j (graph-local column ids); score_out (nullable) the fp32 scores.
 * replaces DGMC.__top_k__ (dgmc.py:85-94: KeOps argKmin, else broadcast-mul-sum + topk) when
 * the pair is too small for a tensor-core tile.  h_s [B, n_s, c], h_t [B, n_t, c] contiguous. */
int dgmc_topk_simt(const float* h_s, const float* h_t, int B, int n_s, int n_t, int c, int k,
                   int32_t* idx_out, float* score_out, dgmc_stream_t stream);

/* ---- K3: row softmax over the k candidates -------------------------------------------
 * replaces ATen softmax at dgmc.py:202,205,225 (sparse branch: un-masked). */
int dgmc_softmax_k_fwd(const float* x, float* y, int64_t rows, int k, dgmc_stream_t stream);
int dgmc_softmax_k_bwd(const float* y, const float* dy, float* dx, int64_t rows, int k,
                       dgmc_stream_t stream);

/* ---- K3 (dense branch): masked row softmax -------------------------------------------
 * replaces masked_fill + softmax + masked_fill, dgmc.py:15-19 as used at :165,168,181.
 * x,y: [B, n_s, n_t]; s_mask [B, n_s], t_mask [B, n_t].  Entries outside
 * s_mask (x) t_mask come out as exactly 0 (the reference's NaN rows are re-zeroed too). */
int dgmc_masked_softmax_fwd(const float* x, const uint8_t* s_mask, const uint8_t* t_mask, float* y,
                            int B, int n_s, int n_t, dgmc_stream_t stream);
int dgmc_masked_softmax_bwd(const float* y, const float* dy, float* dx, int B, int n_s, int n_t,
                            dgmc_stream_t stream);

/* ---- loss / acc / hits_at_k on the sparse side-cars, one pass ---------------------------------
 * For every ground-truth pair g = (y_row[g], y_col[g]) over idx / val [rows, k] (int64 / fp32, the
 * `__idx__` / `__val__` attributes of the sparse outputs):
 *   nll[g] = sum over listed copies of -log(val + eps), hits[g] = number of copies (0: pair absent),
 *   correct[g] = [idx at the row's first arg-max == y_col[g]], in_top[g] = copies among the top_k values.
 * replaces the boolean-mask / argmax / argsort + gather chains of DGMC.loss / acc / hits_at_k
 * (dgmc.py:262-266, 281-284, 302-308); the host adds the arrays up on the device. */
int dgmc_sparse_metrics(const int64_t* idx, const float* val, const int64_t* y_row, const int64_t* y_col,
                        int64_t pairs, int k, int top_k, float eps, float* nll, float* hits, float* correct,
                        float* in_top, dgmc_stream_t stream);

/* ---- K2: candidate gather-dot ---------------------------------------------------------
 * out[m, j] = < a[m, :], b[idx[m, j], :] >.  replaces gather + mul + sum at dgmc.py:198-201
 * (S_hat initialisation) and is the backward of the indicator transport (dS = <r_s, dr_t>).
 * idx holds FLAT row numbers of b (graph offset already added). */
int dgmc_gather_dot_fwd(const float* a, int64_t lda, const float* b, int64_t ldb,
                        const int32_t* idx, float* out, int64_t m, int k, int c,
                        dgmc_stream_t stream);

/* ---- CSR construction: stable counting (radix) sort of `keys` --------------------------
 * ptr[s] .. ptr[s+1] delimit, in `perm`, the positions e (ascending) with keys[e] == s.
 * Used for (i) the by-target view of the candidate lists S_idx, rebuilt once per forward,
 * which turns torch_scatter.scatter_add (dgmc.py:209-212) into a deterministic segmented
 * gather, and (ii) the in-/out-adjacency of each graph (PyG propagate, rel.py:28,30), built
 * once per edge_index and cached by the host. */
size_t dgmc_csr_build_workspace(int64_t nnz, int64_t n_seg);
int dgmc_csr_build(const int32_t* keys, int64_t nnz, int64_t n_seg, int32_t* ptr, int32_t* perm,
                   int32_t* sorted_keys /* nullable: keys[perm], the segment id of every CSR slot */,
                   void* workspace, size_t workspace_bytes, dgmc_stream_t stream);

/* ---- K4 / K5: segmented weighted gather-reduce -----------------------------------------
 * out[s, 0:r] = seg_scale[s] * sum_{e in seg(s)} w_e * X[row_e, 0:r]
 *   seg(s)  = [ptr[s], ptr[s+1])            if ptr != NULL, else [s*fixed_len, (s+1)*fixed_len)
 *   row_e   = src[e] / src_div              (src_div >= 1; lets `perm` of entry ids m*k+j address row m)
 *   w_e     = (w ? w[w_index ? w_index[e] : e] : 1) * (src_scale ? src_scale[row_e] : 1)
 *   seg_scale may be NULL (= 1).  accumulate != 0 adds into `out` instead of overwriting it.
 * replaces: scatter_add of r_s (x) S (dgmc.py:209-212) as a by-target gather [K4]; its
 * backward dr_s; PyG mean aggregation x_j -> i in both flow directions (rel.py:27-30) and the
 * transposed aggregation of their backward [K5]; the scatter_add_ backward of the gathers at
 * dgmc.py:200,221. */
int dgmc_segment_gather(const int32_t* ptr, int fixed_len, const int32_t* src, int src_div,
                        const int32_t* w_index, const float* w, const float* src_scale,
                        const float* seg_scale, const float* x, int64_t ldx, float* out,
                        int64_t ldo, int64_t n_seg, int r, int accumulate, dgmc_stream_t stream);

/* ---- load-balanced variants for CSRs with hub segments ---------------------------------------
 * Same results as dgmc_segment_gather (ptr mode) / dgmc_diff_mlp_bwd_targets, but the unit of work
 * is a block of DGMC_CSR_BLOCK consecutive ENTRIES rather than a segment, so a target node that sits
 * in thousands of candidate lists no longer serialises one lane group.  block_seg[b] = segment of
 * seg_of[e] = segment id of CSR slot e (`sorted_keys` of dgmc_csr_build).  Needs r % 4 == 0 and
 * 16-byte aligned rows; workspace from dgmc_balanced_workspace(nnz, r).  Deterministic (ordered
 * partial sums). */
#define DGMC_CSR_BLOCK 32
size_t dgmc_balanced_workspace(int64_t nnz, int r);
int dgmc_segment_gather_balanced(const int32_t* ptr, const int32_t* seg_of, const int32_t* src,
                                 int src_div, const int32_t* w_index, const float* w,
                                 const float* src_scale, const float* seg_scale, const float* x,
                                 int64_t ldx, float* out, int64_t ldo, int64_t n_seg, int64_t nnz, int r,
                                 int accumulate, void* workspace, size_t workspace_bytes,
                                 dgmc_stream_t stream);
/* Test hook for the balanced kernels' second pass: a segment cut into more than 18 pieces (a hub) is put
 * on its CTA's list (64 entries) and reduced by the whole CTA; hubs that find the list full are reduced by
 * their own lane group in the same association.  entries = 0 .. 64 sets the list size in use (default 64);
 * results must be bit-identical for every value (tests/test_gpu_kernels.py). */
int dgmc_set_fixup_hub_list(int entries);

/* dgmc_segment_gather_balanced in accumulate mode with a ReLU backward fused into the store:
 *   out[s] = (out[s] + sum_{e in seg(s)} (w ? w[e] : 1) * X[src[e]]) * [relu_mask[s] > 0]   (element-wise)
 * relu_mask rows have the pitch of out (ldo).  Every segment must have at least one entry (the transposed
 * RelConv operator has: each node's own row).  r % 4 == 0, r <= 128, 16-byte aligned rows.
 * replaces: the scatter of PyG's propagate backward (rel.py:27-30) followed by the threshold_backward of
 * the ReLU after the previous conv (rel.py:84-85). */
int dgmc_segment_gather_balanced_relu(const int32_t* ptr, const int32_t* seg_of, const int32_t* src,
                                      const float* w, const float* x, int64_t ldx, float* out, int64_t ldo,
                                      const float* relu_mask, int64_t n_seg, int64_t nnz, int r,
                                      void* workspace, size_t workspace_bytes, dgmc_stream_t stream);
int dgmc_diff_mlp_bwd_targets_balanced(const float* g, const float* ps, const float* pt,
                                       const int32_t* tptr, const int32_t* seg_of,
                                       const int32_t* perm, const float* w2, float* d_pt, int64_t n_t,
                                       int64_t nnz, int k, int r, void* workspace,
                                       size_t workspace_bytes, dgmc_stream_t stream);

/* ---- K6: difference gather + consensus MLP update (hoisted first layer) -----------------
 * s_out[m, j] = s_in[m, j] + b2 + sum_c w2[c] * relu(ps[m, c] - pt[idx[m, j], c])
 * with ps = o_s W1^T + b1 and pt = o_t W1^T computed per NODE by the host (library GEMM):
 * Lin1(o_s - o_t) = W1 o_s - W1 o_t + b1.  replaces gather, sub and mlp(D) at dgmc.py:219-223
 * without materialising D [B, N_s, k, R].  b2 is a device pointer to one float. */
int dgmc_diff_mlp_fwd(const float* s_in, const float* ps, const float* pt, const int32_t* idx,
                      const float* w2, const float* b2, float* s_out, int64_t m, int k, int r,
                      dgmc_stream_t stream);
/* backward, row side: d_ps[m, c] = w2[c] * sum_j g[m, j] * [ps[m,c] > pt[idx[m,j],c]];
 * d_w2[c] = sum_{m,j} g[m,j] * relu(...); d_b2 = sum g.  Two-stage deterministic reduction
 * through `workspace`. */
size_t dgmc_diff_mlp_bwd_workspace(int64_t m, int r);
int dgmc_diff_mlp_bwd_rows(const float* g, const float* ps, const float* pt, const int32_t* idx,
                           const float* w2, float* d_ps, float* d_w2, float* d_b2, int64_t m,
                           int k, int r, void* workspace, size_t workspace_bytes,
                           dgmc_stream_t stream);
/* Same computation and outputs as dgmc_diff_mlp_bwd_rows, for callers that cannot supply a
 * workspace (a host process that owns all device memory through its own arena and gives plug-ins
 * no allocator, or a loop where cudaMalloc/cudaFree are banned because they serialise the device).
 * No counterpart in the reference.  The partial sums live in a 528,384-byte array inside the
 * library (held by every GPU that loads it until exit); r must be <= 128, larger r returns
 * DGMC_ERR_UNSUPPORTED.  All calls on a device share that array: calls that overlap on the GPU
 * race.  The first call on each device resolves the array's device address — make it outside
 * any stream capture. */
int dgmc_diff_mlp_bwd_rows_noalloc(const float* g, const float* ps, const float* pt,
                                   const int32_t* idx, const float* w2, float* d_ps, float* d_w2,
                                   float* d_b2, int64_t m, int k, int r, dgmc_stream_t stream);
/* backward, target side (by-target CSR of idx: tptr/perm from dgmc_csr_build):
 * d_pt[t, c] = -w2[c] * sum_{e in seg(t)} g[perm[e]] * [ps[perm[e]/k, c] > pt[t, c]]. */
int dgmc_diff_mlp_bwd_targets(const float* g, const float* ps, const float* pt,
                              const int32_t* tptr, const int32_t* perm, const float* w2,
                              float* d_pt, int64_t n_t, int k, int r, dgmc_stream_t stream);

/* ---- K5 (SplineConv): degree-1 open B-spline convolution -----------------------------------
 * replaces torch_spline_conv's spline_basis / spline_weighting and the scatter-mean of PyG
 * SplineConv(kernel_size=5) as constructed at spline.py:21 ([3p]; forward used at spline.py:49).
 * basis/widx: [n_edges, 2^dim]; pseudo in [0, 1]^dim.  The host computes
 * Y[n, kk, :] = x[n] @ W[kk] with one library GEMM; gather_fwd reduces it along the in-edges
 * (CSR by destination: in_ptr / in_nbr = source node / in_edge = edge id), gather_bwd produces
 * dY along the out-edges (CSR by source).  inv_deg = 1 / max(in-degree, 1).  Gradients with
 * respect to pseudo are not provided (edge attributes are data in every reference script). */
int dgmc_spline_basis(const float* pseudo, int64_t n_edges, int dim, int kernel_size, float* basis,
                      int32_t* widx, dgmc_stream_t stream);
int dgmc_spline_gather_fwd(const float* y, const int32_t* in_ptr, const int32_t* in_nbr,
                           const int32_t* in_edge, const float* basis, const int32_t* widx,
                           int n_corners, const float* inv_deg, float* out, int64_t ldo, int64_t n,
                           int n_kernels, int c, int accumulate, dgmc_stream_t stream);
int dgmc_spline_gather_bwd(const float* g, int64_t ldg, const int32_t* out_ptr,
                           const int32_t* out_nbr, const int32_t* out_edge, const float* basis,
                           const int32_t* widx, int n_corners, const float* inv_deg, float* dy,
                           int64_t n, int n_kernels, int c, dgmc_stream_t stream);

/* ---- weight gradient of a per-node linear map (split-K) -----------------------------------------
 * dw[m, n] = sum_k g[k, m] * x[k, n];  db[m] = sum_k g[k, m] (db may be NULL).  g [k, m], x [k, n].
 * replaces the autograd backward of the Linear layers in rel.py:28-31 / :89 and dgmc.py:74-78 when
 * the node count k is large and m x n is small: a library SGEMM tiles only the 64 x 192 output and
 * runs on 3 CTAs; this kernel splits the node dimension over the whole chip.  fp32, deterministic. */
size_t dgmc_wgrad_workspace(int64_t k, int m, int n);
/* Two-stage form: dgmc_wgrad_partial writes dgmc_wgrad_splits(k, m, n) partial tiles
 * (partial [splits, m, n], partial_b [splits, m], partial_b may be NULL); dgmc_wgrad_reduce adds
 * n_partials consecutive tiles.  Lets the backward passes of all consensus steps, which share their
 * weights (dgmc.py:204-223 calls psi_2 and the MLP num_steps times), pay for one reduction and hand
 * autograd one gradient per parameter instead of num_steps. */
int dgmc_wgrad_splits(int64_t k, int m, int n);
int dgmc_wgrad_partial(const float* g, int64_t ldg, const float* x, int64_t ldx, int64_t k, int m, int n,
                       float* partial, float* partial_b, dgmc_stream_t stream);
int dgmc_wgrad_reduce(const float* partial, const float* partial_b, int n_partials, int m, int n,
                      float* dw, float* db, dgmc_stream_t stream);
int dgmc_wgrad(const float* g, int64_t ldg, const float* x, int64_t ldx, int64_t k, int m, int n,
               float* dw, float* db, void* workspace, size_t workspace_bytes, dgmc_stream_t stream);

/* Tensor-core variant of dgmc_wgrad_partial (same partial layout, same dgmc_wgrad_reduce): tcgen05
 * kind::tf32 on TMA-loaded MN-major tiles, every operand split into hi + lo (rounded to tf32) and
 * accumulated as hi*hi + hi*lo + lo*hi in fp32 TMEM accumulators — fp32-level accuracy.
 * m <= 128 output channels; g and x need 16-byte aligned bases and pitches (dgmc_wgrad_tc_supported). */
int dgmc_wgrad_tc_supported(int m, int n, int64_t ldg, int64_t ldx, const float* g, const float* x);
int dgmc_wgrad_tc_splits(int64_t k, int m, int n);
int dgmc_wgrad_tc_partial(const float* g, int64_t ldg, const float* x, int64_t ldx, int64_t k, int m, int n,
                          float* partial, float* partial_b, dgmc_stream_t stream);
/* Same with the bias gradient summed over rows [0, bias_rows) of g only (see dgmc_linear_tc_rows). */
int dgmc_wgrad_tc_partial_rows(const float* g, int64_t ldg, const float* x, int64_t ldx, int64_t k, int m,
                               int n, int64_t bias_rows, float* partial, float* partial_b,
                               dgmc_stream_t stream);

/* out[i, 0:c] = y[i, 0:c] > 0 ? g[i, 0:c] : 0 on row-strided operands (column blocks of wider
 * buffers): the ReLU backward between two RelConv layers (rel.py:84-85) when the activations live in
 * the concatenated layer buffer.  c % 4 == 0, 16-byte aligned rows. */
int dgmc_relu_bwd(const float* g, int64_t ldg, const float* y, int64_t ldy, float* out, int64_t ldo,
                  int64_t rows, int c, dgmc_stream_t stream);

/* ---- per-node linear map on the tensor cores ------------------------------------------------
 * y[rows, n_out] = relu?(x[rows, k_in] w[n_out, k_in]^T + bias): the GEMM inside every RelConv
 * (rel.py:27-30 after hoisting the aggregation), SplineConv (spline.py:40-47), Lin (rel.py:66 /
 * spline.py:71) and the first MLP layer of dgmc.py:62-66; with the transposed weight it is also
 * their input gradient.  tcgen05 kind::tf32 with hi/lo operand splitting (three MMAs per product,
 * fp32 TMEM accumulator; ~2^-21 relative, i.e. fp32-level).  Two calls: the weight is split once
 * into the workspace (transposed = 1 reads w as [k_in, n_out], for dx = g w), then any number of
 * dgmc_linear_tc calls use that workspace.  x needs a 16-byte aligned base and pitch
 * (dgmc_linear_tc_supported); Inf inputs produce NaN (Inf - Inf in the split). */
int dgmc_linear_tc_supported(int64_t rows, int n_out, int k_in, int64_t ldx, const float* x);
size_t dgmc_linear_tc_workspace(int n_out, int k_in);
int dgmc_linear_tc_split_weight(const float* w, int64_t ldw, int transposed, int n_out, int k_in,
                                void* workspace, size_t workspace_bytes, dgmc_stream_t stream);
int dgmc_linear_tc(const float* x, int64_t ldx, const void* split_weight, const float* bias, int relu,
                   float* y, int64_t ldy, int64_t rows, int n_out, int k_in, dgmc_stream_t stream);
/* Same with the bias added to rows [0, bias_rows) only: the first consensus-MLP layer on the merged
 * node table [source rows ; target rows], whose bias belongs to the source side (Lin1(o_s - o_t) =
 * (W1 o_s + b1) - W1 o_t, dgmc.py:219-221). */
int dgmc_linear_tc_rows(const float* x, int64_t ldx, const void* split_weight, const float* bias,
                        int64_t bias_rows, int relu, float* y, int64_t ldy, int64_t rows, int n_out, int k_in,
                        dgmc_stream_t stream);

/* ---- fused RelConv layer: aggregation -> tensor-core GEMM in one kernel -------------------------
 * y[i, :] = relu?( [ x_i | sum_{e in seg(3i+1)} w_e x[src_e] | sum_{e in seg(3i+2)} w_e x[src_e] ] W^T + bias )
 *           (+ addend[i, :])
 * replaces RelConv.forward (rel.py:25-31: root(x) + mean-propagate(lin1 x) + mean-propagate(lin2 x),
 * the two PyG propagate calls and three Linear maps) and, with the operator of the transposed
 * aggregation and the block-transposed weight, its input gradient.  ptr [3 rows + 1] / src / w: CSR
 * with three segments per node — 3i: the node itself (one entry, weight 1; read through TMA, the list
 * is not consulted), 3i+1 / 3i+2: the two weighted neighbour lists (mean aggregation: w = 1 / degree).
 * order [ceil(rows / 32) * 32]: for every group of 32 consecutive rows a permutation of 0..31, rows
 * with the longest lists first (any permutation is correct; this one balances the gather rounds of a
 * warp).
 * split_weight: dgmc_linear_tc_split_weight(W [n_out, 3c]).  stacked (nullable) receives the fp32
 * operand rows [rows, 3c] for the weight gradient; addend (nullable, may alias y) is added before the
 * store.  c = 32 or 64 (the split weight stays resident in shared memory), n_out <= 64, x / stacked /
 * addend with 16-byte aligned bases and pitches
 * (dgmc_relconv_tc_supported).  3xTF32 arithmetic as dgmc_linear_tc. */
int dgmc_relconv_tc_supported(int64_t rows, int c, int n_out, int64_t ldx, const float* x);
int dgmc_relconv_tc(const int32_t* ptr, const int32_t* src, const float* w, const int32_t* order,
                    const float* x, int64_t ldx,
                    const void* split_weight, const float* bias, int relu, const float* addend, int64_t lda,
                    float* stacked, int64_t lds, float* y, int64_t ldy, int64_t rows, int c, int n_out,
                    dgmc_stream_t stream);

/* ---- dense branch: correspondence scores + masked row softmax on the tensor cores -------------------
 * s_hat[b] = h_s[b] h_t[b]^T, s_0[b] = masked_softmax(s_hat[b]) with S_mask = s_mask (x) t_mask.
 * replaces `h_s @ h_t.transpose(-1, -2)` and masked_softmax at dgmc.py:163-165 (ATen bmm + softmax).
 * The dgmc_linear_tc kernel in block-diagonal mode: tcgen05 kind::tf32 with hi/lo splitting (fp32-level),
 * one 128 x 128 tile per group of graphs, the softmax in the epilogue straight out of TMEM.  h_s
 * [B, n_s, c], h_t [B, n_t, c] zero-padded batches; graphs of up to 128 nodes, c % 4 == 0.  The
 * workspace receives the split copy of h_t. */
int dgmc_dense_scores_tc_supported(int B, int n_s, int n_t, int c, const float* h_s, const float* h_t);
size_t dgmc_dense_scores_tc_workspace(int B, int n_t, int c);
int dgmc_dense_scores_tc(const float* h_s, const float* h_t, const uint8_t* s_mask, const uint8_t* t_mask,
                         float* s_hat, float* s_0, int B, int n_s, int n_t, int c, void* workspace,
                         size_t workspace_bytes, dgmc_stream_t stream);

/* ---- dense branch: batched products of small per-graph matrices (exact fp32) ------------------------
 * nn: out[b] [m, n] = op(a[b]) [m, k] * bm[b] [k, n], op = transpose when trans_a (a[b] stored [k, m]);
 * nt: out[b] [m, n] = a[b] [m, k] * bm[b]^T (bm[b] stored [n, k]).  m, k <= 128 (nn).
 * replaces ATen bmm for r_t = S^T r_s (dgmc.py:170) and for the adjoints of that product and of the
 * score product (dS = r_s g^T, dr_s = S g, dh_s = dS h_t, dh_t = dS^T h_s). */
int dgmc_small_bmm_nn(const float* a, int trans_a, const float* bm, float* out, int batch, int m, int k, int n,
                      dgmc_stream_t stream);
int dgmc_small_bmm_nt(const float* a, const float* bm, float* out, int batch, int m, int n, int k,
                      dgmc_stream_t stream);

/* ---- dense branch: consensus update without materialising D [B, N_s, N_t, R] ----------------
 * s_out[b,i,j] = s_in[b,i,j] + [s_mask[b,i] & t_mask[b,j]] * (b2 + sum_c w2[c] relu(ps[b,i,c] - pt[b,j,c]))
 * replaces D = o_s[:, :, None] - o_t[:, None], mlp(D).masked_fill(~S_mask, 0) at dgmc.py:178-179
 * (first Linear hoisted to the nodes as in K6).  ps [B, n_s, r], pt [B, n_t, r]. */
int dgmc_dense_consensus_fwd(const float* s_in, const float* ps, const float* pt, const float* w2,
                             const float* b2, const uint8_t* s_mask, const uint8_t* t_mask,
                             float* s_out, int B, int n_s, int n_t, int r, dgmc_stream_t stream);
size_t dgmc_dense_consensus_bwd_workspace(int r);
int dgmc_dense_consensus_bwd(const float* g, const float* ps, const float* pt, const float* w2,
                             const uint8_t* s_mask, const uint8_t* t_mask, float* d_ps, float* d_pt,
                             float* d_w2, float* d_b2, int B, int n_s, int n_t, int r,
                             void* workspace, size_t workspace_bytes, dgmc_stream_t stream);

/* ---- multi-GPU: collectives of the row-sharded path over NVLink peer memory ---------------------
 * No counterpart in the reference (single process).  One process per GPU; the ranks share ONE symmetric
 * allocation (same size and layout on every rank, every rank's copy mapped into every process, e.g.
 * a CUDA VMM allocation whose handles the processes exchanged).  peer_bases: HOST array of `world` device pointers, entry q =
 * base of rank q's copy as mapped in THIS process; *_off: byte offsets inside the allocation, equal on
 * all ranks.  pad_off: dgmc_sym_pad_bytes() bytes of flag words, zeroed once before the first call.
 * Every rank must issue the same calls in the same order (they handshake block by block; a rank that
 * waits ~2 s gives up and counts a timeout instead of hanging, see dgmc_sym_timeouts()).  At most
 * DGMC_SYM_MAX_WORLD ranks.  Sums are taken in rank order 0, 1, ... by exactly one rank per element, so
 * all ranks receive bit-identical results.
 *   all_reduce:      out_r[0:n] = sum_q in_q[0:n] on every rank r            (the r_t all-reduce)
 *   all_gather:      out_r[this rank's runs] = in (local pointer)             (layer features of psi / o_t)
 *   reduce_scatter:  out (local pointer) = sum_q in_q[this rank's runs]
 * A rank's part of the gathered tensor is up to two runs [at0, at0 + len0) and [at1, at1 + len1) (floats,
 * relative to out_off / in_off; the local buffer holds run 0 then run 1): equal blocks in rank order are
 * {rank * block, block, 0, 0}; the node-sharded psi networks use two runs — the rank's source nodes and
 * its target nodes in the natural [source ; target] order, so that no permutation is needed on either
 * side.  max_len: the largest len0 + len1 over the ranks (every rank must launch the same grid).
 * n, at*, len*, max_len in floats, multiples of 4.  multicast_base (nullable): the allocation's NVLink-SHARP
 * multicast mapping, if the fabric has one — the switch then does the additions (multimem.ld_reduce)
 * and the broadcast (multimem.st).
 * Handshakes: all_reduce opens ("all partials complete") and closes ("all results landed") with one.
 * all_gather always closes with one; its opening one ("the peers no longer read the output region")
 * may be skipped (sync_before = 0) by a caller that alternates between TWO output regions and consumes
 * each result before its next collective call: a peer can then be at most one call ahead, i.e. writing
 * the other region.  reduce_scatter always opens with one; its closing one ("the peers have finished
 * reading my input") may be skipped (sync_after = 0) under the same two-region discipline for inputs. */
#define DGMC_SYM_MAX_BLOCKS 128
#define DGMC_SYM_MAX_WORLD 8
size_t dgmc_sym_pad_bytes(void);
int dgmc_sym_all_reduce(const void* const* peer_bases, const void* multicast_base, int rank, int world,
                        int64_t in_off, int64_t out_off, int64_t pad_off, int64_t n, dgmc_stream_t stream);
int dgmc_sym_all_gather(const void* const* peer_bases, int rank, int world, const float* in, int64_t out_off,
                        int64_t pad_off, int64_t at0, int64_t len0, int64_t at1, int64_t len1, int64_t max_len,
                        int sync_before, dgmc_stream_t stream);
int dgmc_sym_reduce_scatter(const void* const* peer_bases, int rank, int world, int64_t in_off, float* out,
                            int64_t pad_off, int64_t at0, int64_t len0, int64_t at1, int64_t len1,
                            int64_t max_len, int sync_after, dgmc_stream_t stream);
/* number of handshakes that gave up so far on this device (synchronises the device); 0 in a healthy run */
int dgmc_sym_timeouts(void);

#ifdef __cplusplus
}
#endif
#endif /* DGMC_B200_H */
