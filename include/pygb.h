/*
 * pygb.h — C ABI of libpygb.so: B200 (sm_100a) kernels for the message-passing
 * hot path of ggzhang0071/pytorch_geometric-1.
 *
 * What this boundary replaces.  The reference reaches native code for this
 * path only through the pinned packages torch-scatter==2.0.4
 * (/root/reference/requirement.txt:3) and torch-geometric==1.4.3
 * (requirement.txt:7): `torch.ops.torch_scatter.scatter_{sum,mean,max}(src,
 * index, dim, out, dim_size)` plus ATen `index_select`/`matmul`, called from
 * `MessagePassing.propagate` (reference call sites: gmm_conv.py:131,
 * README.md:43, examples/gcn.py:25,27, ConvexPruning.py:196,221).  Each entry
 * point below names the reference interface it stands in for (SURVEY.md
 * Appendix A section + in-tree call site).
 *
 * Conventions
 *  - extern "C"; plain pointers and sizes; no torch / C++ types.
 *  - Every pointer is a DEVICE pointer owned by the caller (torch's caching
 *    allocator on the Python side).  The library never allocates or frees
 *    device memory and never synchronises the stream; scratch space is passed
 *    in as `ws` with the size reported by the matching *_workspace() call.
 *  - Every call takes the CUDA stream to launch on (cudaStream_t passed as
 *    void*), is re-entrant, and returns 0 on success or a negative PYGB_E_*
 *    code; pygb_last_error() returns the thread-local message.
 *  - Index arrays coming from the reference API (`edge_index`, `index`) are
 *    int64 as in PyG; the compact CSR the kernels consume is int32
 *    (E', N < 2^31).
 *  - dtype: PYGB_F32 or PYGB_BF16 for feature matrices; accumulation is
 *    always fp32.
 */
#ifndef PYGB_H_
#define PYGB_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PYGB_ABI_VERSION 4

enum { PYGB_OK = 0, PYGB_E_ARG = -1, PYGB_E_CUDA = -2, PYGB_E_WORKSPACE = -3,
       PYGB_E_UNSUPPORTED = -4 };
enum { PYGB_F32 = 0, PYGB_BF16 = 1 };
enum { PYGB_SUM = 0, PYGB_MEAN = 1 };
/* self-loop handling, SURVEY A.2 */
enum { PYGB_LOOPS_REMOVE = 0,        /* remove_self_loops                       */
       PYGB_LOOPS_ADD = 1,           /* add_self_loops (keeps existing loops)   */
       PYGB_LOOPS_REMOVE_ADD = 2,    /* GATConv.forward: remove then add        */
       PYGB_LOOPS_ADD_REMAINING = 3  /* GCNConv.norm: add_remaining_self_loops  */ };

typedef void* pygb_stream_t; /* cudaStream_t */

int pygb_abi_version(void);
const char* pygb_last_error(void);
/* Fails (PYGB_E_UNSUPPORTED) unless the current device is compute capability 10.x. */
int pygb_check_device(int* sm_count, int* cc_major, int* cc_minor);

/* ---- K4: graph build (integer outputs bit-exact vs the oracle) ------------ */

/* Number of edges with row == col -> *d_count (device int64).  Replaces the
 * `mask = row != col` + masked-select size sync of remove_self_loops /
 * add_remaining_self_loops (A.2; ref site examples/qm9_nn_conv.py:11). */
int pygb_count_self_loops(const int64_t* edge_index, int64_t E, int64_t* d_count,
                          pygb_stream_t s);

/* One pass over edge_index: d_out[0] = edges with row == col, d_out[1] = edges with an
 * id outside [0, N) (device int64[2]).  The count the self-loop utilities need plus the
 * range check GCNConv.norm owes its callers, behind a single device->host read. */
int pygb_scan_edge_index(const int64_t* edge_index, int64_t E, int64_t N, int64_t* d_out,
                         pygb_stream_t s);

/* Self-loop handling (A.2).  `n_loops` is the value of pygb_count_self_loops
 * (ignored for PYGB_LOOPS_ADD).  E' = E - n_loops (REMOVE), E + N (ADD),
 * E - n_loops + N (REMOVE_ADD, ADD_REMAINING).  Output order: surviving edges
 * in original order, then loops 0..N-1.  edge_weight/out_edge_weight may be
 * NULL together.  ADD_REMAINING keeps an existing loop's weight; if a node has
 * several, the one with the highest edge id wins (the CPU reference's
 * last-write order, Q1). */
size_t pygb_apply_loops_workspace(int64_t E, int64_t N);
int pygb_apply_loops(const int64_t* edge_index, int64_t E, int64_t N, int mode,
                     int64_t n_loops, const float* edge_weight, float fill_value,
                     int64_t* out_edge_index, float* out_edge_weight, void* ws,
                     size_t ws_bytes, pygb_stream_t s);

/* Stable sort of E edges by key[e] in [0, n_rows): rowptr[n_rows+1],
 * perm[E] = original edge ids grouped by key, ascending inside a group
 * (so sequential reductions see the reference's CPU edge order), and, if
 * `other` != NULL, col[p] = (int32) other[perm[p]].  This is the structure that
 * replaces the per-element atomics of scatter_cuda.cu (A.4-A.6): all
 * reductions below walk it without atomics.  Keys outside [0, n_rows) set
 * *d_status (device int32, may be NULL) to 1 and are dropped from rowptr;
 * `other` values outside [0, n_cols) set it to 2 and are stored as 0. */
size_t pygb_csr_build_workspace(int64_t E, int64_t n_rows);
int pygb_csr_build(const int64_t* key, const int64_t* other, int64_t E,
                   int64_t n_rows, int64_t n_cols, int32_t* rowptr, int32_t* perm,
                   int32_t* col, int32_t* d_status, void* ws, size_t ws_bytes,
                   pygb_stream_t s);

/* out[p] = src[perm[p]] for a float edge attribute (weights into CSR order). */
int pygb_permute_f32(const float* src, const int32_t* perm, int64_t E, float* out,
                     pygb_stream_t s);

/* utils.degree (A.2; ref site examples/faust.py:19): counts per key as int64
 * or float32 (out_is_float).  Exact (integer atomics). */
int pygb_degree(const int64_t* index, int64_t E, int64_t N, int out_is_float,
                void* out, pygb_stream_t s);

/* deg[j] = sum_p w[perm[p]] over row j of a CSR keyed on the index to count
 * (GCNConv.norm uses `row`, A.9 Q8), added sequentially in ascending edge id —
 * the CPU reference's scatter_add_ order, hence bit-identical fp32 degrees. */
int pygb_weighted_degree(const float* w, const int32_t* rowptr, const int32_t* perm,
                         int64_t N, float* deg, pygb_stream_t s);

/* GCNConv.norm (A.9; ref sites examples/gcn.py:18-19, ConvexPruning.py:180-185):
 * dis = deg^-1/2 with inf -> 0, norm[e] = (dis[row[e]]*w[e])*dis[col[e]];
 * `w` NULL means all ones.  edge_index is the loop-augmented [2,E'] array and
 * deg comes from pygb_degree (unit weights) or pygb_weighted_degree. */
int pygb_gcn_norm(const int64_t* edge_index, const float* w, int64_t E, int64_t N,
                  const float* deg, float* norm, pygb_stream_t s);

/* The unit-weight normalisation in the slot order of a sorted graph: out[p] =
 * deg[col[p]]^-1/2 * deg[i]^-1/2 for slot p of row i — bit-identical to
 * pygb_permute_f32(pygb_gcn_norm(...), perm) when every edge weight is 1, without the
 * random gather (first call of GCNConv(cached=True), examples/gcn.py:18-19). */
int pygb_gcn_norm_csr(const int32_t* rowptr, const int32_t* col, const float* deg,
                      int64_t n_rows, float* out, pygb_stream_t s);

/* ---- K1: segmented gather-reduce (aggr = add | mean) ---------------------- */

/* out[i, :] = reduce_{p in [rowptr[i], rowptr[i+1])} w_p * x[idx[p], :]
 *   w_p = 1 if edge_w == NULL, edge_w[p] if wperm == NULL, else edge_w[wperm[p]].
 *   reduce = PYGB_SUM or PYGB_MEAN (sum / max(count,1), A.5);  bias (fp32[F],
 *   may be NULL) is added after the reduction (GCNConv.update).
 * Fuses index_select + message (norm * x_j) + scatter_add of
 * MessagePassing.propagate (A.1/A.4; ref sites examples/gcn.py:25,27).  With
 * idx = perm it reduces a materialised [E,F] message tensor given in original
 * edge order (the generic `scatter_` path, A.3).  One warp per row, 128-bit
 * loads when x/out/ld* are 16-byte aligned, fixed summation order, no atomics.
 * ldx/ldo are row strides in elements. */
int pygb_seg_gather_reduce(int reduce, int dtype, const void* x, int64_t ldx,
                           int64_t F, const int32_t* rowptr, const int32_t* idx,
                           const float* edge_w, const int32_t* wperm,
                           const float* bias, int64_t n_rows, void* out,
                           int64_t ldo, pygb_stream_t s);

/* Streaming form of pygb_seg_gather_reduce (K1 v2; same result contract, weights
 * must already be in CSR order).  The E slots of the CSR are cut into n_sub
 * sub-items of Ts slots (pygb_seg_plan picks Ts from E and the SM count); lane
 * groups of persistent warps stream one sub-item each, staging the gathered rows
 * through shared memory with cp.async.  Per CSR, pygb_seg_tables builds
 *   nz_rows[n_rows+2]  the non-empty rows in order, and
 *   sub_info[n_sub]    rank (in nz_rows) of the row owning slot s*Ts, bit 31 set
 *                      when that row began in an earlier sub-item;
 * per index array, pygb_seg_flag_idx makes a copy with bit 31 set on the last
 * slot of every row (idx = col for the fused gather, perm for a message tensor
 * in original edge order).  Rows cut by a sub-item boundary are combined from
 * fp32 partial rows in `ws` in slot order: no atomics, deterministic.  Requires
 * x/out 16-byte aligned with ldx/ldo padded to a 16-byte multiple >= F.
 * E = rowptr[n_rows]. */
int pygb_seg_plan(int64_t E, int32_t* Ts, int64_t* n_sub);
size_t pygb_seg_tables_workspace(int64_t n_rows);
int pygb_seg_tables(const int32_t* rowptr, int64_t n_rows, int64_t E, int32_t Ts,
                    int64_t n_sub, int32_t* nz_rows, int32_t* sub_info, void* ws,
                    size_t ws_bytes, pygb_stream_t s);
int pygb_seg_flag_idx(const int32_t* rowptr, int64_t n_rows, const int32_t* idx, int64_t E,
                      int32_t* idx_flagged, pygb_stream_t s);
size_t pygb_seg_stream_workspace(int dtype, int64_t n_sub, int64_t F);
int pygb_seg_gather_reduce_stream(int reduce, int dtype, const void* x, int64_t ldx,
                                  int64_t F, const int32_t* rowptr,
                                  const int32_t* idx_flagged, const float* edge_w,
                                  const float* bias, int64_t n_rows, int64_t E,
                                  const int32_t* nz_rows, const int32_t* sub_info,
                                  int32_t Ts, int64_t n_sub, void* out, int64_t ldo,
                                  void* ws, size_t ws_bytes, pygb_stream_t s);

/* L2 residency hints for the gather (no counterpart in the reference; the aggregation is HBM-bound
 * and on a power-law graph a few percent of the source rows take a third of the gathers).
 * pygb_seg_mark_hot sets bit 30 of every slot of a flagged index stream whose source row is
 * gathered at least `threshold` times (freq[j] = number of slots that name source j; ids must fit
 * 30 bits); pygb_seg_gather_reduce_stream_hinted is pygb_seg_gather_reduce_stream for such a stream:
 * marked gathers carry an L2 evict_last policy, the others evict_normal.  Results are bit-identical
 * to the unhinted call. */
int pygb_seg_mark_hot(int32_t* idx_flagged, int64_t E, const int32_t* freq, int64_t n_src,
                      int32_t threshold, pygb_stream_t s);
int pygb_seg_gather_reduce_stream_hinted(int reduce, int dtype, const void* x, int64_t ldx,
                                         int64_t F, const int32_t* rowptr,
                                         const int32_t* idx_flagged, const float* edge_w,
                                         const float* bias, int64_t n_rows, int64_t E,
                                         const int32_t* nz_rows, const int32_t* sub_info,
                                         int32_t Ts, int64_t n_sub, void* out, int64_t ldo,
                                         void* ws, size_t ws_bytes, pygb_stream_t s);

/* pygb_seg_gather_reduce_stream_routed (below) for a stream marked by pygb_seg_mark_hot. */
int pygb_seg_gather_reduce_stream_routed_hinted(int reduce, int dtype, const void* x, int64_t ldx,
                                                int64_t F, const int32_t* rowptr,
                                                const int32_t* idx_flagged, const float* edge_w,
                                                const float* bias, int64_t n_rows, int64_t E,
                                                const int32_t* nz_rows, const int32_t* sub_info,
                                                int32_t Ts, int64_t n_sub, void* const* out_tab,
                                                int64_t blk_rows, int64_t sub_rot, int64_t ldo,
                                                int64_t max_ctas, void* ws, size_t ws_bytes,
                                                pygb_stream_t s);

/* pygb_seg_gather_reduce_stream WITHOUT a workspace argument and without any device
 * allocation (no counterpart in the reference: added
 * for callers that cannot call cudaMalloc — code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  Same arguments, same kernels and
 * bit-identical results; the fix-up scratch (partial rows + tail table) lives in ONE
 * file-scope __device__ array of PYGB_SEG_STREAM_STATIC_WS_BYTES that every GPU loading
 * the library holds until exit.  Returns PYGB_E_WORKSPACE (message names both sizes) when
 * pygb_seg_stream_workspace(dtype, n_sub, F) exceeds that limit.  All calls on a device
 * share the array, which keeps the previous call's contents (the entry point resets, on
 * the stream, the part that must start clean): calls that overlap on the GPU — two
 * streams, two threads — RACE; serialise them on one stream. */
#define PYGB_SEG_STREAM_STATIC_WS_BYTES (16u << 20)
int pygb_seg_gather_reduce_stream_static(int reduce, int dtype, const void* x, int64_t ldx,
                                         int64_t F, const int32_t* rowptr,
                                         const int32_t* idx_flagged, const float* edge_w,
                                         const float* bias, int64_t n_rows, int64_t E,
                                         const int32_t* nz_rows, const int32_t* sub_info,
                                         int32_t Ts, int64_t n_sub, void* out, int64_t ldo,
                                         pygb_stream_t s);

/* The same kernel with a routed output (feature-sliced multi-GPU aggregation,
 * DESIGN.md §4): destination row r is stored at
 *   (T*)out_tab[r / blk_rows] + (r % blk_rows) * ldo,
 * where out_tab is a DEVICE array of ceil(n_rows / blk_rows) base pointers, one
 * per block of destination rows — local memory or peer-GPU memory opened with
 * pygb_peer_open (plain stores over NVLink; the caller orders them against the
 * consumer with a stream-ordered collective).  All bases 16-byte aligned.
 * sub_rot rotates the work order: the launch starts at sub-item sub_rot (slot
 * sub_rot * Ts) and wraps around, so that ranks writing to each other begin at
 * different row blocks and no peer receives from everybody at once; the result
 * does not depend on it.  max_ctas > 0 caps the persistent grid, max_ctas < 0
 * leaves |max_ctas| SMs free (the chunked pipeline of dist/sliced.py runs the routed
 * transform concurrently on another stream); 0: every SM. */
int pygb_seg_gather_reduce_stream_routed(int reduce, int dtype, const void* x, int64_t ldx,
                                         int64_t F, const int32_t* rowptr,
                                         const int32_t* idx_flagged, const float* edge_w,
                                         const float* bias, int64_t n_rows, int64_t E,
                                         const int32_t* nz_rows, const int32_t* sub_info,
                                         int32_t Ts, int64_t n_sub, void* const* out_tab,
                                         int64_t blk_rows, int64_t sub_rot, int64_t ldo,
                                         int64_t max_ctas, void* ws, size_t ws_bytes,
                                         pygb_stream_t s);

/* Same walk with a per-(edge, head) weight: out[i,h,:] = sum_p
 * head_w[wp(p)*H + h] * x[idx[p],h,:], wp(p) = wperm ? wperm[p] : p — GAT's
 * `x_j * alpha.view(-1,H,1)` + scatter_add (A.10) and its transpose. */
int pygb_seg_gather_reduce_heads(int dtype, const void* x, int64_t ldx, int H, int C,
                                 const int32_t* rowptr, const int32_t* idx,
                                 const float* head_w, const int32_t* wperm,
                                 const float* bias, int64_t n_rows, void* out,
                                 int64_t ldo, pygb_stream_t s);

/* out[e, :] = x[index[e], :] — the x_i / x_j gathers of propagate for generic
 * message() functions (A.1 step 2; ref site README.md:45-49). index is int64. */
int pygb_gather_rows(int dtype, const void* x, int64_t ldx, int64_t F,
                     const int64_t* index, int64_t E, void* out, int64_t ldo,
                     pygb_stream_t s);

/* out[perm[p]] = <g[row(p), :], x[col[p], :]> for every slot p of the sorted
 * graph (fp32, original edge order): the gradient of the fused weighted
 * aggregation with respect to the edge weight (autograd of GCNConv.message
 * `norm.view(-1, 1) * x_j`, A.9; ref site examples/gcn.py:25 with edge weights,
 * SURVEY 7.3 hard part 4).  No [E, F] temporary. */
int pygb_edge_dot(int dtype, const void* g, int64_t ldg, const void* x, int64_t ldx,
                  int64_t F, const int32_t* rowptr, const int32_t* col,
                  const int32_t* perm, int64_t n_rows, int64_t E, float* out,
                  pygb_stream_t s);

/* ---- K3: segmented max with argmax (aggr = max) --------------------------- */

/* out[i,f] = max_p src[idx[p], f] over row i, arg[i,f] = eid[p] of the FIRST
 * maximal entry in ascending edge-id order (strict '>', NaN never wins; A.6
 * Q4), rows without a winner: value 0, arg = E_fill (Q5).  Values below
 * `zero_below` are then replaced by 0 (utils.scatter_'s `out[out < -10000] = 0`,
 * A.3 Q2; pass -INFINITY to disable); with mask_arg != 0 their arg becomes
 * E_fill too, which is what the autograd of that in-place overwrite needs.
 * idx = eid = perm reduces a message tensor; idx = col, eid = perm is the fused
 * gather-max.  arg may be NULL.
 * Replaces scatter_kernel<max> + scatter_arg_kernel of scatter_cuda.cu. */
int pygb_seg_max(int dtype, const void* src, int64_t lds, int64_t F,
                 const int32_t* rowptr, const int32_t* idx, const int32_t* eid,
                 int64_t n_rows, int64_t E_fill, float zero_below, int mask_arg,
                 void* out, int64_t ldo, int64_t* arg, pygb_stream_t s);

/* Backward of K3: grad_src[e,f] = grad_out[i,f] where arg[i,f] == e, else 0.
 * grad_src must be zero-filled by the caller; one writer per element. */
int pygb_seg_max_backward(const float* grad_out, int64_t ldg, const int64_t* arg,
                          int64_t n_rows, int64_t F, int64_t E, float* grad_src,
                          int64_t lds, pygb_stream_t s);

/* ---- K2: edge softmax and fused GAT -------------------------------------- */

/* utils.softmax (A.7): out[e,h] = exp(src[e,h] - max_row) /
 * (sum_row exp(..) + 1e-16), rows given by rowptr/perm over `index`.  src/out
 * are [E,H] fp32 in original edge order. */
int pygb_edge_softmax(const float* src, int64_t H, const int32_t* rowptr,
                      const int32_t* perm, int64_t n_rows, float* out,
                      pygb_stream_t s);
/* grad_src[e,h] = out[e,h] * (g[e,h] - sum_row(out*g)). */
int pygb_edge_softmax_backward(const float* out, const float* grad_out, int64_t H,
                               const int32_t* rowptr, const int32_t* perm,
                               int64_t n_rows, float* grad_src, pygb_stream_t s);

/* a_dst[n,h] = <xw[n,h,:], att[h,:C]>,  a_src[n,h] = <xw[n,h,:], att[h,C:]>
 * (the two halves of GATConv.message's `(cat[x_i,x_j]*att).sum(-1)`, A.10).
 * a_dst may be NULL. */
int pygb_gat_scores(int dtype, const void* xw, int64_t ldx, const float* att,
                    int64_t N, int H, int C, float* a_src, float* a_dst,
                    pygb_stream_t s);

/* Fused GATConv propagate (A.10; ref sites ConvexPruning.py:209-221,
 * examples/ppi.py:22-28): for row i, head h
 *   l_p = leaky_relu(a_dst[i,h] + a_src[col[p],h]);  m = max_p l_p
 *   out[i,h,:] = sum_p exp(l_p - m) * xw[col[p],h,:] / (sum_p exp(l_p - m) + 1e-16)
 * a_dst may be NULL (bipartite x_i = None).  row_max/row_sum [n_rows,H] are
 * written for the backward (may be NULL).  bias fp32[H*C] may be NULL. */
int pygb_gat_forward(int dtype, const void* xw, int64_t ldx, const float* a_src,
                     const float* a_dst, int H, int C, float negative_slope,
                     const int32_t* rowptr, const int32_t* col, int64_t n_rows,
                     const float* bias, void* out, int64_t ldo, float* row_max,
                     float* row_sum, pygb_stream_t s);

/* Streaming form of pygb_gat_forward (K2 v2): one pass with an online softmax per
 * (row, head) over the flagged index stream / sub-item tables of pygb_seg_tables
 * and pygb_seg_flag_idx; same outputs (out, row_max, row_sum) up to fp32 rounding.
 * Needs C % (16/sizeof(T)) == 0 and 16-byte aligned rows, else PYGB_E_UNSUPPORTED
 * (the caller then uses pygb_gat_forward). */
size_t pygb_gat_stream_workspace(int64_t n_sub, int H, int C);
int pygb_gat_forward_stream(int dtype, const void* xw, int64_t ldx, const float* a_src,
                            const float* a_dst, int H, int C, float negative_slope,
                            const int32_t* rowptr, const int32_t* idx_flagged,
                            int64_t n_rows, int64_t E, const int32_t* nz_rows,
                            const int32_t* sub_info, int32_t Ts, int64_t n_sub,
                            const float* bias, void* out, int64_t ldo, float* row_max,
                            float* row_sum, void* ws, size_t ws_bytes, pygb_stream_t s);

/* alpha[e,h] of the fused forward, recomputed from the saved row statistics,
 * written in ORIGINAL edge order (e = perm[p], or p if perm == NULL). */
int pygb_gat_alpha(const float* a_src, const float* a_dst, int H, float negative_slope,
                   const int32_t* rowptr, const int32_t* col, const int32_t* perm,
                   int64_t n_rows, const float* row_max, const float* row_sum,
                   float* alpha, pygb_stream_t s);

/* GAT backward, per-edge part (dst-sorted rows, fp32): for slot p of row i
 *   alpha_p recomputed;  dalpha_p = <g[i,h,:], xw[col[p],h,:]>
 *   D[i,h] = <g[i,h,:], out[i,h,:] - bias[h,:]>   (= sum_p alpha_p dalpha_p)
 *   ds_p = alpha_p (dalpha_p - D[i,h]) * leaky_relu'(a_dst[i,h] + a_src[col[p],h])
 * alpha and ds are written as [E,H] in original edge order.  d_a_dst / d_a_src
 * follow as segmented sums of ds (pygb_seg_gather_reduce with idx = perm) and
 * d_xw as pygb_seg_gather_reduce_heads over the source-sorted CSR. */
int pygb_gat_edge_grad(const float* xw, int64_t ldx, const float* a_src,
                       const float* a_dst, int H, int C, float negative_slope,
                       const int32_t* rowptr, const int32_t* col, const int32_t* perm,
                       int64_t n_rows, const float* out, int64_t ldo, const float* bias,
                       const float* grad_out, int64_t ldg, const float* row_max,
                       const float* row_sum, float* alpha, float* ds, pygb_stream_t s);

/* Streaming GAT backward (autograd of A.10 / A.7, fp32; what `loss.backward()` at
 * examples/gcn.py:39 / ConvexPruning.py:386 runs through GATConv — the transposes of
 * index_select and scatter_add plus the softmax backward): alpha and ds are RECOMPUTED
 * from per-node records in both passes, so no [E,H] tensor is written or read.
 *   pygb_gat_backward_prep: rec[i,h] = (a_dst[i,h] (0 if a_dst == NULL),
 *       row_max[i,h] * log2(e), 1 / (row_sum[i,h] + 1e-16), <g[i,h,:], out[i,h,:] - bias[h,:]>)
 *       as float[n_dst, H, 4] (16-byte aligned).
 *   pygb_gat_backward_stream, mode 0: destination-sorted walk (rowptr / idx_flagged /
 *       nz_rows / sub_info of the dst CSR; row_feat = grad_out [n_dst, H*C], gat_feat = xw):
 *       d_a[i,h] = d a_dst = sum over the in-edges of i of ds;  d_feat unused (NULL).
 *   mode 1: source-sorted walk (tables of the src CSR; row_feat = xw [n_src, H*C],
 *       gat_feat = grad_out): d_feat[j,:] = sum_i alpha_ij g[i,:],  d_a[j,h] = d a_src.
 * ds = alpha (<g_i, xw_j>_h - rec.w) * leaky_relu'(a_dst + a_src).  Sums run in slot
 * order, cut rows are combined by a fix-up kernel: deterministic, no atomics.  Rows
 * without a slot get 0.  Needs C in {4,8,...,128}, H*C <= 512, 16-byte aligned rows
 * (else PYGB_E_UNSUPPORTED: the caller uses pygb_gat_edge_grad). */
int pygb_gat_backward_prep(const float* a_dst, const float* row_max, const float* row_sum,
                           const float* out, int64_t ldo, const float* bias,
                           const float* grad_out, int64_t ldg, int64_t n_dst, int H, int C,
                           float* rec, pygb_stream_t s);
size_t pygb_gat_backward_stream_workspace(int mode, int64_t n_sub, int H, int C);
int pygb_gat_backward_stream(int mode, const float* row_feat, int64_t ld_row,
                             const float* gat_feat, int64_t ld_gat, const float* a_src,
                             const float* rec, int H, int C, float negative_slope,
                             const int32_t* rowptr, const int32_t* idx_flagged, int64_t n_rows,
                             int64_t E, const int32_t* nz_rows, const int32_t* sub_info,
                             int32_t Ts, int64_t n_sub, float* d_feat, int64_t ldd, float* d_a,
                             void* ws, size_t ws_bytes, pygb_stream_t s);

/* ---- k-NN graph (the step before EdgeConv's propagate) ------------------------ */

/* nbr[q, r] = index of the r-th nearest point of q inside q's example (squared Euclidean
 * distance in fp32, ties to the lower index; -1 when the example has fewer candidates) —
 * `torch_cluster.knn_graph` of the reference (README.md:149-150, examples/pointnet++.py:39-41;
 * DynamicEdgeConv).  x [n, D] fp32 (row stride ldx), batch int64 [n] sorted, ptr int64 [B+1]
 * the first point of every example.  loop != 0 keeps q itself as a candidate.
 * k <= 64, D <= 128 (else PYGB_E_UNSUPPORTED). */
int pygb_knn_graph(const float* x, int64_t ldx, int64_t n, int D, const int64_t* batch,
                   const int64_t* ptr, int k, int loop, int32_t* nbr, pygb_stream_t s);

/* Farthest point sampling, `torch_cluster.fps` (examples/pointnet++.py:39): for every
 * example b (points ptr[b]..ptr[b+1] of x [n, D] fp32), out[out_ptr[b] + t], t <
 * out_ptr[b+1] - out_ptr[b], is the global index of the t-th sample: the first is
 * ptr[b] + start[b] (start int64 [B], local index), every next one the point with the largest
 * squared distance to the samples so far (ties to the lower index).  dmin: float scratch [n]. */
int pygb_fps(const float* x, int64_t ldx, int64_t n, int D, const int64_t* ptr,
             const int64_t* out_ptr, const int64_t* start, int64_t n_examples, float* dmin,
             int64_t* out, pygb_stream_t s);

/* Fixed-radius search, `torch_cluster.radius` (examples/pointnet++.py:40-41,46-48): for query
 * q (y [n_y, D], example batch_y[q]) nbr[q, 0..count[q]) are the first max_num_neighbors
 * points j of x in the same example (ptr_x) with |x_j - y_q|^2 <= r^2, in ascending j; the
 * rest of the row is -1.  skip_same_index != 0 drops j == q (radius_graph without loops). */
int pygb_radius(const float* x, int64_t ldx, const float* y, int64_t ldy, int D,
                const int64_t* ptr_x, const int64_t* batch_y, int64_t n_y, float r,
                int max_num_neighbors, int skip_same_index, int32_t* nbr, int32_t* count,
                pygb_stream_t s);

/* ---- K5: dense node-feature transform on the tensor cores ------------------ */

/* C[M,N] = A[M,K] . W[K,N] (+ bias[N]), fp32 in / fp32 out — the
 * `torch.matmul(x, self.weight)` of GCNConv.forward / GATConv.forward (A.9/A.10;
 * ref sites examples/gcn.py:25,27, ConvexPruning.py:196,221).  tcgen05.mma
 * kind::tf32 with TMA-staged operands and the accumulator in TMEM.
 * passes = 3: 3xTF32 split (Ah.Wh + Al.Wh + Ah.Wl), fp32-faithful (~1e-6 relative);
 * passes = 1: plain TF32 (10-bit mantissa inputs).  A must be 16-byte aligned with
 * lda % 4 == 0 (else PYGB_E_UNSUPPORTED: the caller uses its library GEMM).  `ws`
 * holds the split, transposed copy of W (pygb_gemm_weight_workspace bytes).
 * relu != 0 applies max(x, 0) after the bias in the epilogue (the activation
 * between two layers when the caller owns the whole forward, dist/sliced.py).
 * a_slices > 1: A is stored as a_slices planes of [M, K/a_slices] columns (row
 * stride lda, plane stride a_plane elements; K/a_slices a multiple of 32) — the
 * slice-major row buffer the aggregation kernels of the column-slice ranks fill
 * with contiguous stores; a_slices = 1: ordinary row-major A. */
size_t pygb_gemm_weight_workspace(int64_t K, int64_t N);
int pygb_gemm_xw(const float* A, int64_t lda, int64_t a_slices, int64_t a_plane, int64_t M,
                 int64_t K, const float* W, int64_t ldw, int64_t N, const float* bias,
                 int relu, float* C, int64_t ldc, int passes, void* ws, size_t ws_bytes,
                 pygb_stream_t s);

/* The same transform for bf16 operands and a bf16 result (north_star's 1e-2 bf16 mode;
 * SURVEY 8b `pygb_gemm_tcgen05(in_dtype tf32|bf16)`): tcgen05.mma kind::f16 with fp32
 * accumulation in TMEM, one pass, k-steps of 64 (or 32) elements.  A must be 16-byte
 * aligned with lda % 8 == 0; bias stays fp32; `ws` as above. */
int pygb_gemm_xw_bf16(const void* A, int64_t lda, int64_t M, int64_t K, const void* W,
                      int64_t ldw, int64_t N, const float* bias, int relu, void* C,
                      int64_t ldc, void* ws, size_t ws_bytes, pygb_stream_t s);

/* dW[Kx,Ng] = X[R,Kx]^T . G[R,Ng], fp32 — the weight gradient of `torch.matmul(x,
 * self.weight)` (what `loss.backward()` at examples/gcn.py:39 / ConvexPruning.py:386 computes
 * for GCNConv.weight / GATConv.weight).  The reduction runs over the R rows: split-K over
 * persistent CTAs, each accumulating its whole partial in TMEM with MN-major tcgen05.mma
 * kind::tf32 operands (no transposed copies), partials added in CTA order (deterministic).
 * passes as pygb_gemm_xw.  Needs Kx, Ng <= 256 and 16-byte aligned rows of X and G
 * (else PYGB_E_UNSUPPORTED: the caller uses its library GEMM).  `ws`: the partials. */
size_t pygb_gemm_xtg_workspace(int64_t R, int64_t Kx, int64_t Ng);
int pygb_gemm_xtg(const float* X, int64_t ldx, const float* G, int64_t ldg, int64_t R,
                  int64_t Kx, int64_t Ng, float* dW, int64_t ldw, int passes, void* ws,
                  size_t ws_bytes, pygb_stream_t s);

/* pygb_gemm_xw with the row layout -> column layout redistribution of the
 * feature-sliced multi-GPU layout fused into the epilogue: N is
 * N/Fc column blocks of Fc columns; element (r, q*Fc + c) of the result is stored at
 *   dst_tab[q * n_dup + d] + (dst_row0 + r) * ldd + c      for every d < n_dup,
 * dst_tab a DEVICE array of (N/Fc) * n_dup base pointers, local or peer-GPU memory
 * (pygb_peer_open): the tiles leave over NVLink while the main loop of the next
 * tile runs.  Fc and ldd multiples of 4, bases 16-byte aligned, n_dup <= 4.
 * max_ctas > 0 caps the persistent grid: the epilogue is NVLink-bound, a fraction of
 * the SMs saturates the links and the rest stay free for a concurrent aggregation. */
int pygb_gemm_xw_routed(const float* A, int64_t lda, int64_t a_slices, int64_t a_plane,
                        int64_t M, int64_t K, const float* W, int64_t ldw, int64_t N,
                        const float* bias, int relu,
                        float* const* dst_tab, int64_t Fc, int32_t n_dup, int64_t dst_row0,
                        int64_t ldd, int passes, int64_t max_ctas, void* ws, size_t ws_bytes,
                        pygb_stream_t s);

/* ---- multi-GPU: peer memory and the row -> column redistribution ------------ */

/* One process per GPU.  pygb_peer_alloc returns a zeroed device allocation whose
 * CUDA IPC handle (PYGB_PEER_HANDLE_BYTES bytes) pygb_peer_export writes; another
 * process of the same box maps it with pygb_peer_open (peer access over
 * NVLink/NVSwitch is enabled lazily) and unmaps it with pygb_peer_close.  These
 * replace nothing in the reference (single-device only); they carry
 * the exchange step of the partitioned configuration. */
#define PYGB_PEER_HANDLE_BYTES 64
int pygb_peer_alloc(size_t bytes, void** ptr);
int pygb_peer_free(void* ptr);
int pygb_peer_export(const void* ptr, void* handle, size_t handle_bytes);
int pygb_peer_open(const void* handle, size_t handle_bytes, void** ptr);
int pygb_peer_close(void* ptr);

/* Barrier of the ranks of one box over peer memory (one process per GPU, every rank on its
 * own device).  flag_tab: DEVICE array of `world` pointers, entry q = the block of `world`
 * zero-initialised 64-bit flags that lives in rank q's arena (as mapped in this process).
 * Every rank calls it with the same, strictly increasing `epoch` (1, 2, ...).  Kernels launched
 * on the stream before the call have completed — their stores into peer memory included —
 * before any rank's kernels after the call start.  A rank that waits longer than timeout_ns
 * traps.  Replaces a one-element NCCL all-reduce (≈4x lower latency). */
int pygb_peer_barrier(void* const* flag_tab, int32_t rank, int32_t world, int64_t epoch,
                      int64_t timeout_ns, pygb_stream_t s);

/* Row layout -> column layout: src is this rank's rows [n_rows, n_parts*Fc]
 * (row stride lds); column block q goes to
 *   (T*)dst_tab[q * n_dup + d] + (dst_row0 + r) * ldd + c,   c < Fc, d < n_dup,
 * dst_tab a DEVICE array of n_parts * n_dup base pointers (local or peer).  relu != 0
 * applies max(x, 0) on the way (the activation between two layers).  Fc, lds and
 * ldd are multiples of 16 bytes. */
int pygb_cols_scatter(int dtype, const void* src, int64_t lds, int64_t n_rows, int64_t Fc,
                      int32_t n_parts, int32_t n_dup, void* const* dst_tab, int64_t dst_row0,
                      int64_t ldd, int relu, pygb_stream_t s);

#ifdef __cplusplus
}
#endif
#endif /* PYGB_H_ */
