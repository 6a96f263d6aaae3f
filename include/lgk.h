/*
 * lgk.h -- C ABI of liblgk.so: lesion-gnn message-passing kernels for NVIDIA B200 (sm_100a).
 *
 * This is the drop-in boundary for the hot path named by BASELINE.json:north_star.  The reference
 * (zacharielegault/lesion-gnn) has no FFI of its own: its hot path is the set of PyG / torch-scatter /
 * torch-sparse / torch-cluster operators that its models call.  Each entry point below
 * cites the reference call site (file:line under /root/reference) whose operator it replaces.
 *
 * Contract
 *   - plain pointers and sizes only; every pointer is a DEVICE pointer unless the name ends in _host
 *   - the library never allocates or frees device memory and keeps no global mutable state except a
 *     thread-local error string; workspaces are caller-provided (sizes from the *_workspace_bytes calls)
 *   - every call only ENQUEUES work on `stream` (a cudaStream_t passed as void*); it never synchronises
 *     and never reads device data on the host, so all calls are CUDA-graph-capture safe
 *   - returns 0 on success, a negative lgk_status on failure; lgk_last_error() describes the failure
 *   - indices inside the library are int32 (N, E < 2^31); the int64 edge_index / batch tensors of the
 *     PyG API are converted once by lgk_csr_build / lgk_batch_ptr
 *   - `dtype`: LGK_F32 (fp32 storage, fp32 math) or LGK_BF16 (bf16 storage, fp32 accumulation)
 *   - there is no CPU fallback: the functions fail if no sm_100 device kernel image can run
 */
#ifndef LGK_H_
#define LGK_H_

#include <stdint.h>

#if defined(__GNUC__)
#define LGK_API __attribute__((visibility("default")))
#else
#define LGK_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  LGK_OK = 0,
  LGK_ERR_BAD_ARG = -1,     /* null pointer, negative size, unsupported enum value        */
  LGK_ERR_UNSUPPORTED = -2, /* shape outside what the kernels are instantiated for        */
  LGK_ERR_WORKSPACE = -3,   /* workspace too small                                        */
  LGK_ERR_LAUNCH = -4       /* cudaGetLastError() after the launch reported a failure     */
} lgk_status;

enum { LGK_F32 = 0, LGK_BF16 = 1 };
enum { LGK_SUM = 0, LGK_MEAN = 1, LGK_MAX = 2 };
/* self-loop handling inside an aggregation over a CSR row i:
 *   LGK_SELF_NONE    use the stored entries as they are                (SAGEConv, GraphConv)
 *   LGK_SELF_REPLACE skip stored (i,i) entries, add exactly one (i,i)  (GATConv remove+add; GCNConv add_remaining)
 *   LGK_SELF_ADD     keep stored entries, add one extra (i,i) term     (GINConv's (1+eps) x_i)              */
enum { LGK_SELF_NONE = 0, LGK_SELF_REPLACE = 1, LGK_SELF_ADD = 2 };
enum { LGK_ACT_NONE = 0, LGK_ACT_ELU = 1, LGK_ACT_RELU = 2 };

LGK_API const char* lgk_last_error(void);
LGK_API int lgk_version(void);
/* number of kernel launches enqueued through the library (by any thread of the process: autograd runs the backward on
 * its own) since the last reset */
LGK_API int64_t lgk_launch_count(void);
LGK_API void lgk_reset_launch_count(void);

/* ------------------------------------------------------------------------------------------------
 * (1) CSR / CSC / segment builder
 * ------------------------------------------------------------------------------------------------ */

/* Replaces torch_sparse.SparseTensor(row=edge_index[1], col=edge_index[0]) built by ToSparseTensor()
 * (reference src/lesion_gnn/datasets/datamodule.py:45; consumed at models/gat.py:87-88, gin.py:59-60,
 * drgnet.py:98-99), plus the transpose index the backward pass needs.
 *   edge_index : int64 [2, E] row-major (row 0 = src j, row 1 = dst i), values in [0, N)
 *   rowptr[N+1], col[E], perm[E]      CSR by dst, entries ordered by (dst, src, original edge id);
 *                                     col[p] = src of slot p, perm[p] = original edge id of slot p
 *   colptr[N+1], row[E], csr2csc[E]   CSC by src, entries ordered by CSR slot; row[q] = dst,
 *                                     csr2csc[q] = CSR slot of CSC slot q.  Pass NULL x3 to skip.
 * Result is bit-exact equal to a stable sort on key dst*N+src.                                      */
LGK_API int64_t lgk_csr_workspace_bytes(int64_t num_nodes, int64_t num_edges);
LGK_API int lgk_csr_build(const int64_t* edge_index, int64_t num_edges, int64_t num_nodes,
                  int32_t* rowptr, int32_t* col, int32_t* perm,
                  int32_t* colptr, int32_t* row, int32_t* csr2csc,
                  void* workspace, int64_t workspace_bytes, void* stream);

/* The same build for edges ALREADY GROUPED BY DESTINATION (edge_index[1] non-decreasing): the output order of
 * lgk_knn_graph / lgk_radius_graph and of the reference's per-sample KNNGraph + Batch.from_data_list collation
 * (datasets/datamodule.py:45-48, transforms.py:23).  rowptr comes from boundary detection on the sorted row -- no histogram,
 * scan or bucket fill for the CSR side -- and both in-bucket sorts run on 32-bit ids.  Outputs are bit-identical to
 * lgk_csr_build's.  `violation` (int32 device scalar, may be NULL) is set to 1 when the input is not sorted by destination
 * or holds an index outside [0, N) (the outputs are then unspecified but every access stays in bounds), else 0.
 * Workspace: lgk_csr_workspace_bytes.                                                                              */
LGK_API int lgk_csr_build_sorted(const int64_t* edge_index, int64_t num_edges, int64_t num_nodes, int32_t* rowptr, int32_t* col,
                         int32_t* perm, int32_t* colptr, int32_t* row, int32_t* csr2csc, int32_t* violation,
                         void* workspace, int64_t workspace_bytes, void* stream);

/* Graph segment pointer from the sorted `batch` vector of PyG collation:
 * ptr[g] = first node of graph g, ptr[B] = N; empty graphs get empty segments.
 * Replaces the implicit segmentation done by scatter(x, batch) in global_*_pool
 * (reference models/gat.py:56, gin.py:33, pointnet.py:38).                                          */
LGK_API int lgk_batch_ptr(const int64_t* batch, int64_t num_nodes, int64_t num_graphs, int32_t* ptr, void* stream);

/* gather rows of a per-edge array into CSR order: out[p] = in[perm[p]] (edge_weight -> SparseTensor value,
 * reference datamodule.py:45 / drgnet.py:55) */
LGK_API int lgk_gather_f32(const float* in, const int32_t* index, float* out, int64_t n, void* stream);

/* ------------------------------------------------------------------------------------------------
 * kNN graph (torch_cluster.knn_graph behind torch_geometric.transforms.KNNGraph;
 * reference src/lesion_gnn/transforms.py:19-23, configs/config.py:47, datasets/datamodule.py:43-48)
 * ------------------------------------------------------------------------------------------------ */

/* For every centre i of graph g (nodes ptr[g]..ptr[g+1]): its min(k', n_g) nearest nodes of the same
 * graph by squared L2 distance, k' = k if loop else k+1 with the centre itself then dropped; ties go
 * to the lower node index.  Edges are written grouped by centre, ascending distance:
 *   edge_src[eoff(i) + r] = r-th neighbour, edge_dst[...] = i,  eoff(i) = eptr[g] + (i-ptr[g]) * d_g,
 *   d_g = min(k', n_g) - (loop ? 0 : 1),  eptr = exclusive scan of n_g * d_g  (lgk_knn_edge_ptr).
 *   pos : [N, dim] row-major, pos_dtype 0 = float32, 1 = float64 (the reference stores float64,
 *         datasets/nodes/lesions.py:175); dim <= 8; k <= 64
 *   edge_index_out : int64 [2, E_total] row-major (row 0 = src, row 1 = dst), E_total = eptr[B]      */
LGK_API int lgk_knn_edge_ptr(const int32_t* ptr, int64_t num_graphs, int k, int loop, int64_t* eptr, void* stream);
LGK_API int lgk_knn_graph(const void* pos, int pos_dtype, int dim, const int32_t* ptr, const int64_t* eptr,
                  int64_t num_graphs, int64_t num_nodes, int k, int loop,
                  int64_t* edge_index_out, int64_t num_edges_total, void* stream);

/* Radius graph (torch_cluster.radius_graph behind torch_geometric.transforms.RadiusGraph; reference
 * scripts/sweep.py:113-117): for every centre i the nodes j of the same graph with |p_j - p_i|^2 < r^2 (distance in
 * the dtype of pos), ascending node index, at most max_num_neighbors; the centre itself only if loop.
 * Pass 1 (edge_index_out == NULL): counts[i] = number of neighbours kept.  Pass 2: offsets = exclusive scan of the
 * counts (int64 [N]), edge_index_out int64 [2, E] (row 0 = src j, row 1 = dst i), E = total count.        */
LGK_API int lgk_radius_graph(const void* pos, int pos_dtype, int dim, const int32_t* ptr, int64_t num_graphs, int64_t num_nodes,
                     double r, int loop, int max_num_neighbors, int32_t* counts, const int64_t* offsets,
                     int64_t* edge_index_out, int64_t num_edges_total, void* stream);

/* PointNet++ set abstraction (reference models/pointnet.py:13-29).
 * lgk_fps = torch_cluster.fps: per graph g the out_ptr[g+1] - out_ptr[g] = ceil(ratio * n_g) farthest-point samples (squared
 * L2 in the dtype of pos, ties to the lower index), first sample = start[g] (index inside the graph; NULL: node 0);
 * out_index[out_ptr[g] + s] = global node index of sample s.  dist_ws: [N] elements of the dtype of pos.            */
LGK_API int lgk_fps(const void* pos, int pos_dtype, int dim, const int32_t* ptr, const int64_t* out_ptr, const int64_t* start,
            int64_t num_graphs, int64_t num_nodes, void* dist_ws, int64_t* out_index, void* stream);
/* lgk_radius = torch_cluster.radius(x, y, r, batch_x, batch_y, max_num_neighbors): for every query y_i the points x_j of
 * the same graph with |x_j - y_i|^2 < r^2, ascending j, at most max_num_neighbors.  Pass 1 (row_out == NULL): counts[i];
 * pass 2: offsets = exclusive scan of the counts, row_out[e] = i (query), col_out[e] = j (point).                  */
LGK_API int lgk_radius(const void* x, const void* y, int pos_dtype, int dim, const int32_t* ptr_x, const int32_t* ptr_y,
               int64_t num_graphs, int64_t num_x, int64_t num_y, double r, int max_num_neighbors, int32_t* counts,
               const int64_t* offsets, int64_t* row_out, int64_t* col_out, void* stream);
/* message input of torch_geometric.nn.PointNetConv: out[e, :] = cat[x[src[e], :F], pos_src[src[e], :D] - pos_dst[dst[e], :D]]
 * (fp32; F may be 0 with x NULL)                                                                                */
LGK_API int lgk_edge_concat(const float* x, const float* pos_src, const float* pos_dst, const int64_t* src, const int64_t* dst,
                    float* out, int64_t num_edges, int64_t feat, int64_t pos_dim, void* stream);

/* Gaussian edge weights exp(-d^2 / 2 sigma^2) / sqrt(2 pi sigma^2) (reference src/lesion_gnn/transforms.py:50-58) */
LGK_API int lgk_gaussian_distance(const void* pos, int pos_dtype, int dim, const int64_t* edge_index, int64_t num_edges,
                          double sigma, float* edge_weight, void* stream);

/* ------------------------------------------------------------------------------------------------
 * (2)+(4) fused gather + segment-reduce aggregation (SpMM) and its transpose
 * Replaces MessagePassing.propagate = index_select + [mul] + scatter(sum|mean|max), or
 * torch_sparse.matmul(adj_t, x, reduce) on the adj_t path, behind GINConv (reference models/gin.py:23,31),
 * GraphConv (drgnet.py:31,33,55), SAGEConv / GCNConv (BASELINE.json configs 1,3,4; SURVEY.md 8a a-6,a-7).
 * ------------------------------------------------------------------------------------------------ */

/* out[r, :] = post[r] * REDUCE_{p in ptr[r]..ptr[r+1]} ( w(p) * pre[idx[p]] * x[idx[p], :] )  (+ self term) (+ bias)
 *   w(p)   = ew ? ew[emap ? emap[p] : p] : 1
 *   self   = self_mode ? (selfw ? selfw[r] : self_fill) * pre[r] * x[r, :] : none
 *   reduce = LGK_SUM | LGK_MEAN (divide by max(#entries,1); self_mode must be NONE) | LGK_MAX (no weights;
 *            empty rows give 0; arg_out[r,f] = slot p of the winning entry or -1, may be NULL)
 *   pre, post, ew, emap, selfw, bias, arg_out may be NULL.  x, out: [n_rows, F] row-major, `dtype`.
 * The forward pass calls it with the CSR (rowptr, col); the backward pass (dX = A^T dY) calls it with the
 * CSC (colptr, row, emap = csr2csc), so both directions are atomic-free and deterministic.            */
LGK_API int lgk_spmm(const int32_t* ptr, const int32_t* idx, const int32_t* emap, const float* ew,
             const float* pre, const float* post, const float* selfw, float self_fill, int self_mode,
             const void* x, const float* bias, void* out, int32_t* arg_out,
             int64_t n_rows, int64_t feat, int reduce, int dtype, void* stream);

/* Row-blocked ("quad") form of a CSR / CSC for lgk_spmm_blocked: four consecutive rows 4g..4g+3 share one merged
 * neighbour list, ascending by source id, qent[ptr[4g] + t] = source | membership << 28 (bit r: row 4g+r has the entry;
 * a duplicate edge of one row is emitted as a second entry), t < qlen[2g]; qlen[2g+1] = number of stored (i, i)
 * entries of each of the four rows, 8 bits per row (int32 [n_quads, 2]).  cmin / cmax [lgk_quad_chunks(n_rows)]: smallest /
 * largest source id referenced by each chunk of 16 quads (INT32_MAX / -1 for a chunk without entries); qmeta int32[4] =
 * {widest chunk window in rows, number of chunks too wide for the 256-byte-slice ring, 0, 0}, read on the device only.  Built once per
 * batch next to the CSR / CSC (lgk_csr_build) and shared by every layer; n_rows < 2^28.  qent has room for E entries. */
LGK_API int64_t lgk_quad_chunks(int64_t n_rows);
LGK_API int lgk_quad_build(const int32_t* ptr, const int32_t* idx, int64_t n_rows, int32_t* qlen, uint32_t* qent,
                   int32_t* cmin, int32_t* cmax, int32_t* qmeta, void* stream);
/* lgk_spmm for large graphs whose neighbours lie in a window of node ids (collated lesion graphs: block diagonal,
 * raster node order): same result as lgk_spmm without edge values (`ew`), reduce = LGK_SUM | LGK_MEAN, fp32, feat % 4 == 0,
 * summation in ascending source order.  Source rows are streamed once per CTA into a shared-memory ring by TMA and every
 * gathered row feeds the up to four destination rows of its quad; chunks whose window does not fit the ring are reduced
 * with global gathers by the same launch.  Same self-loop / pre / post / bias semantics as lgk_spmm.              */
LGK_API int lgk_spmm_blocked(const int32_t* ptr, const int32_t* qlen, const uint32_t* qent, const int32_t* cmin,
                     const int32_t* cmax, const int32_t* qmeta, const float* pre, const float* post, const float* selfw, float self_fill,
                     int self_mode, const void* x, const float* bias, void* out, int64_t n_rows, int64_t feat,
                     int reduce, int dtype, void* stream);

/* backward of the max aggregation: dx[j,f] = sum over CSC entries q of column j with
 * arg[row[q], f] == csr2csc[q] of dy[row[q], f]                                                        */
LGK_API int lgk_spmm_max_bwd(const int32_t* colptr, const int32_t* row, const int32_t* csr2csc, const int32_t* arg,
                     const void* dy, void* dx, int64_t n_rows, int64_t feat, int dtype, void* stream);

/* gcn_norm (torch_geometric.nn.conv.gcn_conv.gcn_norm): with add_remaining_self_loops
 * semantics computes deg_i = sum_{non-loop p} w_p + selfw_i and writes dinv[i] = deg_i^-1/2 (0 if deg_i == 0);
 * selfw_out[i] = weight of the last stored (i,i) entry in original edge order, else fill (NULL if ew == NULL
 * => all self weights equal `fill`).  add_self_loops == 0: plain weighted in-degree.                     */
LGK_API int lgk_gcn_norm(const int32_t* rowptr, const int32_t* col, const float* ew /* CSR order */,
                 float fill, int add_self_loops, float* dinv, float* selfw_out, int64_t num_nodes, void* stream);

/* ------------------------------------------------------------------------------------------------
 * GAT edge-softmax attention (torch_geometric.nn.GATConv; reference models/gat.py:31,51)
 * ------------------------------------------------------------------------------------------------ */

/* a_src[n,h] = <h[n,h,:], att_src[h,:]>, a_dst likewise.  h: [N, H*C] */
LGK_API int lgk_gat_alpha(const void* h, const float* att_src, const float* att_dst, float* a_src, float* a_dst,
                  int64_t num_nodes, int heads, int chan, int dtype, void* stream);

/* out[i,h,:] = sum_{j in N(i)} alpha_ij^h * h[j,h,:] + bias, alpha = softmax_j(leaky_relu(a_src[j,h] + a_dst[i,h]))
 * with denominator + 1e-16; N(i) = stored non-loop entries + one self loop if add_self_loops (GATConv's
 * remove_self_loops + add_self_loops), else the stored entries.  Single pass, online softmax.
 * rowmax/rowsum [N,H]: saved softmax statistics for the backward pass.
 * dropout_p > 0: attention dropout with a counter-based RNG keyed on (seed, CSR slot, head).  seed_dev (may be NULL):
 * a device-side int64 counter mixed into the seed at run time -- a captured CUDA graph freezes `seed`, so the training
 * step passes its optimiser step counter here and every replay draws a new mask; the forward and both backward
 * launches of one step must see the same value.                                                         */
LGK_API int lgk_gat_fwd(const int32_t* rowptr, const int32_t* col, const void* h, const float* a_src, const float* a_dst,
                const float* bias, void* out, float* rowmax, float* rowsum,
                int64_t num_nodes, int heads, int chan, float negative_slope, int add_self_loops,
                float dropout_p, uint64_t seed, const int64_t* seed_dev, int dtype, void* stream);

/* lgk_gat_fwd with the activation that follows every GAT layer of the reference (models/gat.py:51 `F.elu(conv(x))`)
 * written by the same launch: out as above (the backward needs it), out_elu = elu(out).                        */
LGK_API int lgk_gat_fwd_elu(const int32_t* rowptr, const int32_t* col, const void* h, const float* a_src, const float* a_dst,
                    const float* bias, void* out, void* out_elu, float* rowmax, float* rowsum,
                    int64_t num_nodes, int heads, int chan, float negative_slope, int add_self_loops,
                    float dropout_p, uint64_t seed, const int64_t* seed_dev, int dtype, void* stream);

/* backward, dst side (CSR): da_dst[i,h] = sum_j dz_ij, and the packed per-(node, head) statistics
 * stats[i,h] = {a_dst, rowmax, 1/(rowsum + 1e-16), delta} with delta = <dout[i,h,:], out[i,h,:] - bias>
 * (float [N,H,4], 16-byte aligned) from which the src-side pass recomputes every alpha_ij and dz_ij.        */
LGK_API int lgk_gat_bwd_dst(const int32_t* rowptr, const int32_t* col, const void* h, const float* a_src, const float* a_dst,
                    const float* rowmax, const float* rowsum, const void* out, const float* bias, const void* dout,
                    float* stats, float* da_dst,
                    int64_t num_nodes, int heads, int chan, float negative_slope, int add_self_loops,
                    float dropout_p, uint64_t seed, const int64_t* seed_dev, int dtype, void* stream);
/* lgk_gat_bwd_dst for a layer run with lgk_gat_fwd_elu: dy is the gradient of elu(out); the kernel forms
 * dout = dy * elu'(out) on the fly, uses it as above and writes it to `dout` (same type and shape as out) for
 * lgk_gat_bwd_src and the bias gradient -- no separate lgk_elu_bwd launch.                                    */
LGK_API int lgk_gat_bwd_dst_elu(const int32_t* rowptr, const int32_t* col, const void* h, const float* a_src, const float* a_dst,
                        const float* rowmax, const float* rowsum, const void* out, const float* bias, const void* dy,
                        void* dout, float* stats, float* da_dst,
                        int64_t num_nodes, int heads, int chan, float negative_slope, int add_self_loops,
                        float dropout_p, uint64_t seed, const int64_t* seed_dev, int dtype, void* stream);
/* backward, src side (CSC): dh[j,h,:] = sum_i alpha_ij dout[i,h,:] + da_src[j,h] att_src[h,:] + da_dst[j,h] att_dst[h,:],
 * da_src[j,h] = sum_i dz_ij                                                                             */
LGK_API int lgk_gat_bwd_src(const int32_t* colptr, const int32_t* row, const int32_t* csr2csc, const void* h,
                    const float* a_src, const float* a_dst, const float* rowmax, const float* rowsum,
                    const float* stats, const float* da_dst, const float* att_src, const float* att_dst,
                    const void* dout, void* dh, float* da_src,
                    int64_t num_nodes, int heads, int chan, float negative_slope, int add_self_loops,
                    float dropout_p, uint64_t seed, const int64_t* seed_dev, int dtype, void* stream);
/* datt_src[h,c] (+)= sum_n da_src[n,h] h[n,h,c]; datt_dst likewise (two-stage, deterministic); accumulate != 0 adds to
 * the outputs (gradient buckets) instead of overwriting them */
LGK_API int64_t lgk_gat_datt_workspace_bytes(int64_t num_nodes, int heads, int chan);
LGK_API int lgk_gat_datt(const void* h, const float* da_src, const float* da_dst, float* datt_src, float* datt_dst,
                 int64_t num_nodes, int heads, int chan, int accumulate, void* workspace, int64_t workspace_bytes, int dtype,
                 void* stream);

/* ------------------------------------------------------------------------------------------------
 * (3) global pooling = segment reduce over the sorted batch vector
 * (torch_geometric.nn.global_mean_pool / global_max_pool / global_add_pool; reference models/gat.py:56,
 *  gin.py:33, pointnet.py:38)
 * ------------------------------------------------------------------------------------------------ */

/* out_sum[g,:] = sum (or mean if `mean`) of x[ptr[g]..ptr[g+1], :], out_max[g,:] = max (0 if empty),
 * arg[g,f] = node index of the max or -1.  Either output may be NULL (fused mean||max when both given). */
LGK_API int64_t lgk_pool_workspace_bytes(int64_t num_nodes, int64_t num_graphs, int64_t feat);
LGK_API int lgk_pool_fwd(const void* x, const int32_t* ptr, void* out_sum, int mean, void* out_max, int32_t* arg,
                 int64_t num_nodes, int64_t num_graphs, int64_t feat,
                 void* workspace, int64_t workspace_bytes, int dtype, void* stream);
/* dx[n,:] = dsum[g(n),:] * (mean ? 1/max(cnt_g,1) : 1) + (arg[g(n),f] == n ? dmax[g(n),f] : 0) */
LGK_API int lgk_pool_bwd(const void* dsum, int mean, const void* dmax, const int32_t* arg, const int32_t* ptr, void* dx,
                 int64_t num_nodes, int64_t num_graphs, int64_t feat, int dtype, void* stream);

/* SortAggregation(k) (torch_geometric.nn.aggr.SortAggregation; reference models/drgnet.py:37,59):
 * per graph, nodes ordered by their LAST channel, descending, stable (ties: lower node index first; NaN first, as
 * torch.sort); out[g, r*F:(r+1)*F] = x[node of rank r, :] for r < min(k, n_g), zeros beyond;
 * sel[g, r] = that node's index or -1.  The backward writes dx (zero-filled here) through `sel`.          */
LGK_API int lgk_sort_pool_fwd(const void* x, const int32_t* ptr, void* out, int32_t* sel,
                      int64_t num_graphs, int64_t feat, int64_t k, int dtype, void* stream);
LGK_API int lgk_sort_pool_bwd(const void* dy, const int32_t* sel, void* dx,
                      int64_t num_nodes, int64_t num_graphs, int64_t feat, int64_t k, int dtype, void* stream);

/* GATConv forward on the windowed, row-blocked structure (lgk_quad_build of the CSR) for large graphs: the same outputs as
 * lgk_gat_fwd / lgk_gat_fwd_elu (out, optional ELU(out), rowmax, rowsum -- the statistics lgk_gat_bwd_* read), attention
 * dropout 0 only, chan % 32 == 0 (a 128-byte feature slice then lies inside one head), fp32.  Source rows are staged once
 * per CTA in a shared-memory ring by TMA together with their logit halves a_src[j, head]; four destination rows share one
 * merged neighbour list; the soft-max is online at batch granularity (32 merged entries).  Summation order differs from
 * lgk_gat_fwd (ascending source id over the merged list, self loop last): results agree to fp32 rounding, not bit for
 * bit; deterministic run to run.  workspace: lgk_gat_blocked_workspace_bytes (the transposed a_src).          */
LGK_API int64_t lgk_gat_blocked_workspace_bytes(int64_t n_rows, int heads);
LGK_API int lgk_gat_fwd_blocked(const int32_t* ptr, const int32_t* qlen, const uint32_t* qent, const int32_t* cmin,
                        const int32_t* cmax, const int32_t* qmeta, const void* h, const float* a_src, const float* a_dst,
                        const float* bias, void* out, void* out_elu, float* rowmax, float* rowsum, int64_t n_rows,
                        int heads, int chan, float negative_slope, int add_self_loops, void* workspace,
                        int64_t workspace_bytes, int dtype, void* stream);

/* ------------------------------------------------------------------------------------------------
 * (5) node-feature transform (the torch_geometric Linear inside every conv: `lin`, `lin_l`, `lin_r`,
 *     `lin_rel`, `lin_root`, MLP.lins; reference models/gat.py:29,31,46, gin.py:21,23,25)
 * ------------------------------------------------------------------------------------------------ */

/* out[n,m] = act( rowscale[n] * sum_k x[n,k] W[m,k] + bias[m] + (accumulate ? out[n,m] : 0) )
 * `dtype` is the storage type of the ACTIVATIONS (x, out, dy, dx): LGK_F32 or LGK_BF16; weights, biases, their
 * gradients and all accumulators are fp32.  fp32: 3xTF32 on tcgen05 for tall products (single-pass TF32 would break the
 * 1e-5 bar), FFMA otherwise; bf16: FFMA on bf16 loads.  rowscale, bias may be NULL.                     */
LGK_API int lgk_linear_fwd(const void* x, const float* weight, const float* bias, const float* rowscale, void* out,
                   int64_t n, int64_t k, int64_t m, int act, int accumulate, int dtype, void* stream);
/* GATConv.lin fused with the attention-logit halves (reference models/gat.py:31,51; PyG GATConv.forward: x = lin(x),
 * alpha_src = (x * att_src).sum(-1)):  out[n, H*C] = x W^T (no bias) and a_src[n,H] = <out[n,h,:], att_src[h,:]>, a_dst
 * likewise, formed from the fp32 ACCUMULATORS in the transform's epilogue (before out is rounded to `dtype`) whenever the
 * head width allows it (chan in {16, 32, 64} on the tensor-core path, chan % 4 == 0 and chan | 64 on the FFMA path);
 * other widths run the transform and then lgk_gat_alpha on the stored out.                                   */
LGK_API int lgk_linear_gat_fused(int64_t n, int64_t k, int heads, int chan, int dtype); /* 1: epilogue form (aligned pointers assumed) */
LGK_API int lgk_linear_gat_fwd(const void* x, const float* weight, const float* att_src, const float* att_dst, void* out,
                       float* a_src, float* a_dst, int64_t n, int64_t k, int heads, int chan, int dtype, void* stream);
/* dx[n,k] = sum_m dy[n,m] W[m,k]  (+ dx if accumulate) */
LGK_API int lgk_linear_bwd_x(const void* dy, const float* weight, void* dx, int64_t n, int64_t k, int64_t m,
                     int accumulate, int dtype, void* stream);
/* dW[m,k] (+)= sum_n dy[n,m] x[n,k]; db[m] (+)= sum_n dy[n,m] (db may be NULL). Split over n, deterministic. */
LGK_API int64_t lgk_linear_bwd_w_workspace_bytes(int64_t n, int64_t k, int64_t m);
LGK_API int lgk_linear_bwd_w(const void* dy, const void* x, float* dw, float* db, int64_t n, int64_t k, int64_t m,
                     int accumulate, void* workspace, int64_t workspace_bytes, int dtype, void* stream);

/* db[m] (+)= sum_n dy[n,m] -- the bias-gradient column sum of lgk_linear_bwd_w(dw = NULL) -- WITHOUT a caller workspace
 * and without any allocation: the slab partials live in a device array inside the library.  This entry point has no
 * counterpart in the reference; it was added for callers that cannot call
 * cudaMalloc (CUDA-graph capture, tight serving loops, harnesses that forbid it).  It calls no cudaMalloc / cudaFree and
 * no allocating library routine.  Limit: m <= LGK_COLSUM_STATIC_MAX_M (any n); larger m returns LGK_ERR_UNSUPPORTED with
 * a message naming the limit.  Cost: 296 x 1024 x 4 B = 1.16 MiB of device memory on every GPU that loads the library,
 * held until the process exits.  All calls on one GPU share that scratch and it keeps the last call's contents (each call
 * overwrites everything it reads, nothing needs clearing): calls that OVERLAP on the GPU (different streams, concurrent
 * graph launches) race -- order them on one stream or with events.  dy: [n, m] row-major `dtype`; db: fp32 [m];
 * result bit-identical to lgk_linear_bwd_w(dy, NULL, NULL, db, ...).                                                  */
#define LGK_COLSUM_STATIC_MAX_M 1024
LGK_API int lgk_colsum_static(const void* dy, float* db, int64_t n, int64_t m, int accumulate, int dtype, void* stream);

/* ------------------------------------------------------------------------------------------------
 * elementwise helpers of the step (activation after each conv: F.elu at reference models/gat.py:51, gin.py:31,
 * drgnet.py:55; optimiser: torch.optim.Adam at models/base.py:162-170)
 * ------------------------------------------------------------------------------------------------ */
LGK_API int lgk_elu_fwd(const void* x, void* y, int64_t n, int dtype, void* stream);
LGK_API int lgk_elu_bwd(const void* y, const void* dy, void* dx, int64_t n, int dtype, void* stream);
/* fused multi-tensor Adam / AdamW over one flat fp32 buffer; grad is multiplied by grad_scale first
 * (1/world_size after the data-parallel all-reduce).  The 1-based step count is `step`, or -- so that a
 * captured CUDA graph can be replayed -- read from the device counter `step_dev` when that is non-NULL.
 * zero_grad != 0: every gradient is set to zero once it has been consumed (optimizer.zero_grad() of the next
 * step without its own launch).  advance_step != 0 (needs step_dev, an int64[2]: {steps completed, scratch = 0}):
 * the launch uses step_dev[0] + 1 as its step count and stores it back when its last CTA retires, so the caller
 * launches no increment.                                                                                      */
LGK_API int lgk_adam_step(float* param, float* grad, float* exp_avg, float* exp_avg_sq, int64_t n,
                  float lr, float beta1, float beta2, float eps, float weight_decay, int64_t step,
                  int64_t* step_dev, float grad_scale, int adamw, int zero_grad, int advance_step, void* stream);

/* Clamped regression loss of the reference's training step, forward and gradient in one launch
 * (models/base.py:196-199 `criterion(logits, y)` with the regression clamp of models/gat.py:94-95:
 * loss = mean_i L(clamp(logits[i], lo, hi) - y[i]), L = d^2 (kind 0, nn.MSELoss) or SmoothL1 with beta 1 (kind 1)).
 * loss_out[0] receives the loss, dlogits (may be NULL) grad_scale * d loss / d logits[i] (zero outside [lo, hi], as
 * torch.clamp's backward; grad_scale = the weight of this rank's loss under data parallelism, 1 otherwise).
 * Labels are int64 as the DataLoader delivers them.  One CTA: n is a batch of graphs.  */
LGK_API int lgk_clamped_regression_loss(const float* logits, const int64_t* y, int64_t n, float lo, float hi, int kind,
                                float grad_scale, float* loss_out, float* dlogits, void* stream);

/* out[i, :] = src[index[i], :], rows of row_bytes bytes (a multiple of 4), index int64 in [0, src_rows): collation of a
 * device-resident dataset by node index -- Batch.from_data_list of the reference's DataLoader (datasets/datamodule.py:64-69)
 * without the host round trip.  Rows with an out-of-range index are left untouched.                              */
LGK_API int lgk_gather_rows(const void* src, const int64_t* index, void* out, int64_t n, int64_t row_bytes,
                    int64_t src_rows, void* stream);

/* ------------------------------------------------------------------------------------------------
 * BatchNorm1d + activation + dropout of GIN's MLP (torch_geometric.nn.models.MLP = Linear -> BatchNorm1d -> ELU ->
 * Dropout -> Linear; reference models/gin.py:23) and the activation + dropout after every GIN layer (gin.py:31-32)
 * ------------------------------------------------------------------------------------------------ */
/* y = dropout(act(batch_norm(x))), x, y [n, d] fp32.  training != 0: batch statistics (biased variance), running_mean /
 * running_var updated in place with `momentum` (unbiased variance) and *num_batches_tracked += 1 (each may be NULL), n > 1;
 * training == 0: the running statistics normalise.  save_mean / save_invstd [d] receive the statistics used (the backward
 * reads them).  weight / bias may be NULL (affine=False).  dropout: counter RNG keyed on (seed, *seed_dev, element), the
 * same function in the backward -- no mask is stored.  workspace: lgk_bn_workspace_bytes(n, d).                    */
LGK_API int64_t lgk_bn_workspace_bytes(int64_t n, int64_t d);
LGK_API int lgk_bn_act_fwd(const void* x, const float* weight, const float* bias, float* running_mean, float* running_var,
                   int64_t* num_batches_tracked, void* y, float* save_mean, float* save_invstd, int64_t n, int64_t d,
                   float eps, float momentum, int training, int act, float dropout_p, uint64_t seed, const int64_t* seed_dev,
                   void* workspace, int64_t workspace_bytes, int dtype, void* stream);
/* dx, dweight (+)= sum g xhat, dbias (+)= sum g with g = dy * dropout keep * act'(z), recomputed from x and the saved
 * statistics; training != 0: the batch statistics take part in dx.  dweight / dbias may be NULL.                  */
LGK_API int lgk_bn_act_bwd(const void* x, const void* dy, const float* weight, const float* bias, const float* save_mean,
                   const float* save_invstd, void* dx, float* dweight, float* dbias, int accumulate, int64_t n, int64_t d,
                   int training, int act, float dropout_p, uint64_t seed, const int64_t* seed_dev,
                   void* workspace, int64_t workspace_bytes, int dtype, void* stream);
/* y = dropout(act(x)) and dx = dy * keep * act'(x) (x = the pre-activation input), n elements fp32 */
LGK_API int lgk_act_dropout_fwd(const void* x, void* y, int64_t n, int act, float dropout_p, uint64_t seed,
                        const int64_t* seed_dev, int dtype, void* stream);
LGK_API int lgk_act_dropout_bwd(const void* x, const void* dy, void* dx, int64_t n, int act, float dropout_p, uint64_t seed,
                        const int64_t* seed_dev, int dtype, void* stream);

/* Data-parallel step without NCCL: the gradient all-reduce over NVLink peer memory fused with Adam / AdamW.
 * Every rank owns a symmetric allocation of lgk_dp_buffer_bytes(n, world) bytes that is mapped into all peers (layout: 8 KB
 * of flags, the n fp32 gradients at lgk_dp_grad_offset_bytes(), then receive slots float [2][world][n]); peer_base_dev is
 * a DEVICE array of `world` pointers to those allocations, in rank order (entry `rank` is the caller's own).  One call
 * enqueues ONE kernel: every CTA pushes its chunk of the own gradients into its slot of every peer, raises a per-chunk
 * flag there, waits for the peers' flags of the same chunk, sums all ranks' chunks in rank order (bit-identical on every
 * rank), scales by 1/world, applies Adam to the caller's parameters and clears the own bucket.  epoch_dev is int64[2],
 * zero at start, advanced by the call itself and never rewound; *step_dev = the Adam step count for the bias corrections: of
 * THIS update (>= 1) when advance_step = 0, of the updates completed so far when advance_step = 1 (the kernel adds one
 * and stores it back).  n % 4 == 0, world <= 16.  Capturable in a CUDA graph; all ranks must enqueue the call once per
 * step with the same n.                                                                                              */
LGK_API int64_t lgk_dp_buffer_bytes(int64_t n, int world);
LGK_API int64_t lgk_dp_grad_offset_bytes(void);
LGK_API int lgk_dp_allreduce_adam(const uint64_t* peer_base_dev, int rank, int world, int64_t n, float* param, float* exp_avg,
                          float* exp_avg_sq, float lr, float beta1, float beta2, float eps, float weight_decay,
                          int64_t* step_dev, int64_t* epoch_dev, int adamw, int advance_step, void* stream);

/* torch.optim.SGD over one flat fp32 buffer (reference models/base.py:167-168: lr and weight_decay; momentum /
 * nesterov with dampening 0 are supported through a zero-initialised momentum_buf, NULL when momentum == 0):
 * g = grad_scale * grad + weight_decay * p; [b = momentum * b + g; g = nesterov ? g + momentum * b : b;] p -= lr * g.
 * zero_grad != 0 clears every gradient once consumed.                                                            */
LGK_API int lgk_sgd_step(float* param, float* grad, float* momentum_buf, int64_t n, float lr, float momentum,
                 float weight_decay, int nesterov, float grad_scale, int zero_grad, void* stream);

/* nn.CrossEntropyLoss(weight=class_weights) of the reference's classification mode (models/base.py:93-94,196-199),
 * forward and gradient in one launch: loss = sum_i w[y_i] * (-log softmax(logits[i])[y_i]) / sum_i w[y_i]
 * (class_weight may be NULL: w = 1).  logits [n, num_classes] fp32 row-major, y int64 [n]; targets outside
 * [0, num_classes) contribute nothing.  dlogits (may be NULL) = grad_scale * d loss / d logits.  One CTA.        */
LGK_API int lgk_cross_entropy_loss(const float* logits, const int64_t* y, const float* class_weight, int64_t n,
                           int64_t num_classes, float grad_scale, float* loss_out, float* dlogits, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* LGK_H_ */
