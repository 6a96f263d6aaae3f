// kNN graph construction for a whole collated batch (torch_cluster.knn_graph behind
// torch_geometric.transforms.KNNGraph; reference transforms.py:19-23, configs/config.py:47, applied per sample
// on the CPU inside Dataset.__getitem__ at datasets/datamodule.py:43-48) and the Gaussian edge weights of
// reference transforms.py:50-58.
//
// Lesion graphs are small (1..600 nodes), so the search is brute force inside each graph: one
// thread per centre scans its graph's nodes in ascending index order and keeps the k best in a sorted register
// list (strict < on insertion => ties go to the lower index, the order torch_cluster's CUDA kernel yields).
// Distances are computed in the dtype of `pos` (the reference stores float64 centroids) without FMA
// contraction so that they round exactly like (p_i - p_j).pow(2).sum(-1).
#include "common.cuh"
#include <float.h>

namespace lgk {
namespace {

constexpr int kMaxDim = 8;

__device__ __forceinline__ float sq_add(float acc, float a, float b) {
  const float d = __fsub_rn(a, b);
  return __fadd_rn(acc, __fmul_rn(d, d));
}
__device__ __forceinline__ double sq_add(double acc, double a, double b) {
  const double d = __dsub_rn(a, b);
  return __dadd_rn(acc, __dmul_rn(d, d));
}
template <typename P> __device__ __forceinline__ P big();
template <> __device__ __forceinline__ float big<float>() { return FLT_MAX; }
template <> __device__ __forceinline__ double big<double>() { return DBL_MAX; }

__device__ __forceinline__ int graph_of(const int* __restrict__ ptr, int B, int i) {
  int lo = 0, hi = B;  // ptr[lo] <= i < ptr[hi]
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(ptr + mid) <= i) lo = mid; else hi = mid;
  }
  return lo;
}

__device__ __forceinline__ int edges_per_node(int n, int k, int loop) {
  return loop ? min(k, n) : min(k, n - 1);
}

// eptr[g] = sum_{g' < g} n_g' * d_g'   (single CTA, B is small)
__global__ void __launch_bounds__(1024) knn_edge_ptr_kernel(const int* __restrict__ ptr, int64_t B, int k, int loop,
                                                            int64_t* __restrict__ eptr) {
  __shared__ int64_t s_warp[32];
  __shared__ int64_t s_carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { s_carry = 0; eptr[0] = 0; }
  __syncthreads();
  for (int64_t base = 0; base < B; base += 1024) {
    const int64_t g = base + threadIdx.x;
    int64_t v = 0;
    if (g < B) {
      const int n = ptr[g + 1] - ptr[g];
      v = int64_t(n) * max(edges_per_node(n, k, loop), 0);
    }
    int64_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int64_t t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      int64_t w = s_warp[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int64_t t = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += t;
      }
      s_warp[lane] = w;
    }
    __syncthreads();
    const int64_t carry = s_carry;
    const int64_t incl = carry + (warp ? s_warp[warp - 1] : 0) + inc;
    if (g < B) eptr[g + 1] = incl;
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = incl;
    __syncthreads();
  }
}

template <typename P, int K>
__global__ void __launch_bounds__(128) knn_kernel(const P* __restrict__ pos, int dim, const int* __restrict__ ptr,
                                                  const int64_t* __restrict__ eptr, int B, int64_t N, int k, int loop,
                                                  int64_t* __restrict__ out_src, int64_t* __restrict__ out_dst) {
  const int64_t i64 = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i64 >= N) return;
  const int i = int(i64);
  const int g = graph_of(ptr, B, i);
  const int g0 = __ldg(ptr + g), g1 = __ldg(ptr + g + 1);
  const int n = g1 - g0;
  const int d = edges_per_node(n, k, loop);
  if (d <= 0) return;

  P ci[kMaxDim];
#pragma unroll
  for (int t = 0; t < kMaxDim; ++t) ci[t] = t < dim ? pos[int64_t(i) * dim + t] : P(0);

  P best_d[K];
  int best_i[K];
#pragma unroll
  for (int t = 0; t < K; ++t) { best_d[t] = big<P>(); best_i[t] = -1; }

  for (int c = g0; c < g1; ++c) {
    if (!loop && c == i) continue;
    P dist = P(0);
#pragma unroll
    for (int t = 0; t < kMaxDim; ++t)
      if (t < dim) dist = sq_add(dist, ci[t], pos[int64_t(c) * dim + t]);
    if (dist < best_d[K - 1]) {
      P cd = dist;
      int cidx = c;
      bool placed = false;
#pragma unroll
      for (int t = 0; t < K; ++t) {
        // strict <: an equal, earlier (lower-index) entry stays in front; once the candidate is placed the
        // displaced entries shift down unconditionally (a tie between two of THEM must not drop one)
        const bool sw = placed || cd < best_d[t];
        placed = sw;
        const P td = best_d[t];
        const int ti = best_i[t];
        best_d[t] = sw ? cd : td;
        best_i[t] = sw ? cidx : ti;
        cd = sw ? td : cd;
        cidx = sw ? ti : cidx;
      }
    }
  }
  const int64_t off = eptr[g] + int64_t(i - g0) * d;
#pragma unroll
  for (int t = 0; t < K; ++t) {
    if (t < d) {
      out_src[off + t] = best_i[t];
      out_dst[off + t] = i;
    }
  }
}

template <typename P>
__global__ void __launch_bounds__(256) gaussian_kernel(const P* __restrict__ pos, int dim, const int64_t* __restrict__ src,
                                                       const int64_t* __restrict__ dst, int64_t E, double sigma,
                                                       float* __restrict__ w) {
  const int64_t e = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (e >= E) return;
  const int64_t a = src[e], b = dst[e];
  P sq = P(0);
  for (int t = 0; t < dim; ++t) sq = sq_add(sq, pos[a * dim + t], pos[b * dim + t]);
  // reference: exp(-sq / (2 sigma^2)) / sqrt(2 pi sigma^2), evaluated in pos' dtype, then cast to float32
  const P two_s2 = P(2.0 * sigma * sigma);
  const P norm = P(sqrt(2.0 * 3.14159265358979323846 * sigma * sigma));
  w[e] = float(exp(-sq / two_s2) / norm);
}

template <typename P>
int launch_knn(const void* pos, int dim, const int* ptr, const int64_t* eptr, int64_t B, int64_t N, int k, int loop,
               int64_t* out, int64_t E, cudaStream_t stream) {
  const unsigned grid = unsigned(ceil_div(N, 128));
  const P* p = static_cast<const P*>(pos);
  int64_t* src = out;
  int64_t* dst = out + E;
  if (k <= 8) knn_kernel<P, 8><<<grid, 128, 0, stream>>>(p, dim, ptr, eptr, int(B), N, k, loop, src, dst);
  else if (k <= 16) knn_kernel<P, 16><<<grid, 128, 0, stream>>>(p, dim, ptr, eptr, int(B), N, k, loop, src, dst);
  else if (k <= 32) knn_kernel<P, 32><<<grid, 128, 0, stream>>>(p, dim, ptr, eptr, int(B), N, k, loop, src, dst);
  else knn_kernel<P, 64><<<grid, 128, 0, stream>>>(p, dim, ptr, eptr, int(B), N, k, loop, src, dst);
  LGK_CHECK_LAUNCH("knn_kernel");
  return LGK_OK;
}


// ---- PointNet++ set abstraction (reference models/pointnet.py:13-29): farthest point sampling,
// the bipartite radius query and the edge features cat[x_j, pos_j - pos_i] of PointNetConv ----------------------------

// torch_cluster.fps: one CTA per graph; sample s+1 = the node farthest (squared L2, dtype of pos) from the s samples
// chosen so far, ties to the lower index; the first sample is start[g] (local index) or node 0.
// out[out_ptr[g] + s] = global node index.  dist: [N] scratch in the dtype of pos.
template <typename P>
__global__ void __launch_bounds__(256) fps_kernel(const P* __restrict__ pos, int dim, const int* __restrict__ ptr,
                                                  const int64_t* __restrict__ out_ptr, const int64_t* __restrict__ start,
                                                  P* __restrict__ dist, int64_t* __restrict__ out) {
  __shared__ P s_val[256];
  __shared__ int s_idx[256];
  __shared__ int s_cur;
  const int g = blockIdx.x;
  const int lo = ptr[g], n = ptr[g + 1] - lo;
  const int64_t o0 = out_ptr[g];
  const int m = int(out_ptr[g + 1] - o0);
  if (n <= 0 || m <= 0) return;
  for (int i = threadIdx.x; i < n; i += 256) dist[lo + i] = big<P>();
  int cur = 0;
  if (start) { const int64_t st = start[g]; cur = st < 0 ? 0 : (st >= n ? n - 1 : int(st)); }
  for (int s = 0; s < m; ++s) {
    if (threadIdx.x == 0) out[o0 + s] = lo + cur;
    P pc[kMaxDim];
    for (int d = 0; d < dim; ++d) pc[d] = pos[int64_t(lo + cur) * dim + d];
    P best = P(-1);
    int bidx = 0x7fffffff;
    for (int i = threadIdx.x; i < n; i += 256) {
      P acc = P(0);
      for (int d = 0; d < dim; ++d) acc = sq_add(acc, pos[int64_t(lo + i) * dim + d], pc[d]);
      const P dm = min(dist[lo + i], acc);
      dist[lo + i] = dm;
      if (dm > best) { best = dm; bidx = i; }  // ascending i per thread: the first maximum stays
    }
    s_val[threadIdx.x] = best;
    s_idx[threadIdx.x] = bidx;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
      if (threadIdx.x < o) {
        const P v = s_val[threadIdx.x + o];
        const int j = s_idx[threadIdx.x + o];
        if (v > s_val[threadIdx.x] || (v == s_val[threadIdx.x] && j < s_idx[threadIdx.x])) { s_val[threadIdx.x] = v; s_idx[threadIdx.x] = j; }
      }
      __syncthreads();
    }
    if (threadIdx.x == 0) s_cur = s_idx[0];
    __syncthreads();
    cur = s_cur;
    __syncthreads();
  }
}

// torch_cluster.radius(x, y, r, batch_x, batch_y, max_num_neighbors): for every query y_i the points x_j of the same graph
// with |x_j - y_i|^2 < r^2, ascending j, at most max_nb.  Pass 1: counts[i]; pass 2 (offsets): row_out = i, col_out = j.
template <typename P, bool FILL>
__global__ void __launch_bounds__(128) radius_bipartite_kernel(const P* __restrict__ x, const P* __restrict__ y, int dim,
                                                               const int* __restrict__ ptr_x, const int* __restrict__ ptr_y, int B,
                                                               int64_t Ny, P r2, int max_nb, int* __restrict__ counts,
                                                               const int64_t* __restrict__ offsets, int64_t* __restrict__ row_out,
                                                               int64_t* __restrict__ col_out) {
  const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= Ny) return;
  const int g = graph_of(ptr_y, B, int(i));
  const int lo = ptr_x[g], hi = ptr_x[g + 1];
  P yi[kMaxDim];
  for (int d = 0; d < dim; ++d) yi[d] = y[i * dim + d];
  int cnt = 0;
  const int64_t base = FILL ? offsets[i] : 0;
  for (int j = lo; j < hi && cnt < max_nb; ++j) {
    P acc = P(0);
    for (int d = 0; d < dim; ++d) acc = sq_add(acc, x[int64_t(j) * dim + d], yi[d]);
    if (acc < r2) {
      if (FILL) {
        row_out[base + cnt] = i;
        col_out[base + cnt] = j;
      }
      ++cnt;
    }
  }
  if (!FILL) counts[i] = cnt;
}

// PointNetConv message input: out[e, :] = cat[x[src[e], :], pos_src[src[e], :] - pos_dst[dst[e], :]]  (fp32)
__global__ void __launch_bounds__(256) edge_concat_kernel(const float* __restrict__ x, const float* __restrict__ pos_src,
                                                          const float* __restrict__ pos_dst, const int64_t* __restrict__ src,
                                                          const int64_t* __restrict__ dst, float* __restrict__ out, int64_t E, int F,
                                                          int D) {
  const int lane = threadIdx.x & 31;
  const int64_t e = int64_t(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (e >= E) return;
  const int64_t j = src[e], i = dst[e];
  float* o = out + e * (F + D);
  for (int f = lane; f < F; f += 32) o[f] = __ldg(x + j * F + f);
  if (lane < D) o[F + lane] = __ldg(pos_src + j * D + lane) - __ldg(pos_dst + i * D + lane);
}
}  // namespace
}  // namespace lgk

using namespace lgk;

// ---- radius graph (torch_cluster.radius_graph behind torch_geometric.transforms.RadiusGraph; reference
// scripts/sweep.py:113-117): every j of the same graph with |p_j - p_i|^2 < r^2, in ascending index order, at most
// max_nb per centre.  Two passes over the same scan: count, then (after an exclusive scan of the counts) fill.
namespace lgk {
namespace {
template <typename P, bool FILL>
__global__ void __launch_bounds__(128) radius_kernel(const P* __restrict__ pos, int dim, const int* __restrict__ ptr, int B,
                                                     int64_t N, P r2, int loop, int max_nb, int* __restrict__ counts,
                                                     const int64_t* __restrict__ offsets, int64_t* __restrict__ src_out,
                                                     int64_t* __restrict__ dst_out) {
  const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const int g = graph_of(ptr, B, int(i));
  const int lo = ptr[g], hi = ptr[g + 1];
  P pi[kMaxDim];
  for (int d = 0; d < dim; ++d) pi[d] = pos[i * dim + d];
  int cnt = 0;
  const int64_t base = FILL ? offsets[i] : 0;
  for (int j = lo; j < hi && cnt < max_nb; ++j) {
    if (!loop && j == int(i)) continue;
    P acc = P(0);
    for (int d = 0; d < dim; ++d) acc = sq_add(acc, pi[d], pos[int64_t(j) * dim + d]);
    if (acc < r2) {
      if (FILL) {
        src_out[base + cnt] = j;
        dst_out[base + cnt] = i;
      }
      ++cnt;
    }
  }
  if (!FILL) counts[i] = cnt;
}
}  // namespace
}  // namespace lgk

extern "C" int lgk_radius_graph(const void* pos, int pos_dtype, int dim, const int32_t* ptr, int64_t B, int64_t N, double r,
                                int loop, int max_num_neighbors, int32_t* counts, const int64_t* offsets,
                                int64_t* edge_index_out, int64_t E, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(B >= 0 && N >= 0 && E >= 0 && r >= 0 && max_num_neighbors > 0, LGK_ERR_BAD_ARG, "lgk_radius_graph: bad argument");
  LGK_REQUIRE(dim > 0 && dim <= kMaxDim, LGK_ERR_UNSUPPORTED, "lgk_radius_graph: dim must be in [1, %d]", kMaxDim);
  LGK_REQUIRE(pos_dtype == 0 || pos_dtype == 1, LGK_ERR_BAD_ARG, "lgk_radius_graph: pos_dtype must be 0 (f32) or 1 (f64)");
  const bool fill = edge_index_out != nullptr;
  LGK_REQUIRE(fill ? offsets != nullptr : counts != nullptr, LGK_ERR_BAD_ARG, "lgk_radius_graph: counts (pass 1) or offsets + edge_index_out (pass 2)");
  if (N == 0 || (fill && E == 0)) return LGK_OK;
  LGK_REQUIRE(pos && ptr, LGK_ERR_BAD_ARG, "lgk_radius_graph: null pointer");
  const unsigned grid = unsigned(ceil_div(N, 128));
  const double r2 = r * r;
  if (pos_dtype == 0) {
    const float* p = static_cast<const float*>(pos);
    if (fill) radius_kernel<float, true><<<grid, 128, 0, stream>>>(p, dim, ptr, int(B), N, float(r2), loop, max_num_neighbors, nullptr, offsets, edge_index_out, edge_index_out + E);
    else radius_kernel<float, false><<<grid, 128, 0, stream>>>(p, dim, ptr, int(B), N, float(r2), loop, max_num_neighbors, counts, nullptr, nullptr, nullptr);
  } else {
    const double* p = static_cast<const double*>(pos);
    if (fill) radius_kernel<double, true><<<grid, 128, 0, stream>>>(p, dim, ptr, int(B), N, r2, loop, max_num_neighbors, nullptr, offsets, edge_index_out, edge_index_out + E);
    else radius_kernel<double, false><<<grid, 128, 0, stream>>>(p, dim, ptr, int(B), N, r2, loop, max_num_neighbors, counts, nullptr, nullptr, nullptr);
  }
  LGK_CHECK_LAUNCH("radius_kernel");
  return LGK_OK;
}

extern "C" int lgk_knn_edge_ptr(const int32_t* ptr, int64_t B, int k, int loop, int64_t* eptr, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(B >= 0 && k > 0 && ptr && eptr, LGK_ERR_BAD_ARG, "lgk_knn_edge_ptr: bad argument");
  knn_edge_ptr_kernel<<<1, 1024, 0, stream>>>(ptr, B, k, loop, eptr);
  LGK_CHECK_LAUNCH("knn_edge_ptr_kernel");
  return LGK_OK;
}

extern "C" int lgk_knn_graph(const void* pos, int pos_dtype, int dim, const int32_t* ptr, const int64_t* eptr, int64_t B,
                             int64_t N, int k, int loop, int64_t* edge_index_out, int64_t E, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(B >= 0 && N >= 0 && E >= 0, LGK_ERR_BAD_ARG, "lgk_knn_graph: negative size");
  LGK_REQUIRE(k > 0 && k <= 64, LGK_ERR_UNSUPPORTED, "lgk_knn_graph: k must be in [1, 64] (got %d)", k);
  LGK_REQUIRE(dim > 0 && dim <= kMaxDim, LGK_ERR_UNSUPPORTED, "lgk_knn_graph: dim must be in [1, %d]", kMaxDim);
  LGK_REQUIRE(pos_dtype == 0 || pos_dtype == 1, LGK_ERR_BAD_ARG, "lgk_knn_graph: pos_dtype must be 0 (f32) or 1 (f64)");
  if (N == 0 || E == 0) return LGK_OK;
  LGK_REQUIRE(pos && ptr && eptr && edge_index_out, LGK_ERR_BAD_ARG, "lgk_knn_graph: null pointer");
  return pos_dtype == 0 ? launch_knn<float>(pos, dim, ptr, eptr, B, N, k, loop, edge_index_out, E, stream)
                        : launch_knn<double>(pos, dim, ptr, eptr, B, N, k, loop, edge_index_out, E, stream);
}

extern "C" int lgk_gaussian_distance(const void* pos, int pos_dtype, int dim, const int64_t* edge_index, int64_t E,
                                     double sigma, float* edge_weight, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(E >= 0 && dim > 0 && sigma > 0, LGK_ERR_BAD_ARG, "lgk_gaussian_distance: bad argument");
  if (E == 0) return LGK_OK;
  LGK_REQUIRE(pos && edge_index && edge_weight, LGK_ERR_BAD_ARG, "lgk_gaussian_distance: null pointer");
  const unsigned grid = unsigned(ceil_div(E, 256));
  if (pos_dtype == 0)
    gaussian_kernel<float><<<grid, 256, 0, stream>>>(static_cast<const float*>(pos), dim, edge_index, edge_index + E, E, sigma, edge_weight);
  else if (pos_dtype == 1)
    gaussian_kernel<double><<<grid, 256, 0, stream>>>(static_cast<const double*>(pos), dim, edge_index, edge_index + E, E, sigma, edge_weight);
  else
    LGK_REQUIRE(false, LGK_ERR_BAD_ARG, "lgk_gaussian_distance: bad pos_dtype");
  LGK_CHECK_LAUNCH("gaussian_kernel");
  return LGK_OK;
}

extern "C" int lgk_fps(const void* pos, int pos_dtype, int dim, const int32_t* ptr, const int64_t* out_ptr, const int64_t* start,
                       int64_t B, int64_t N, void* dist_ws, int64_t* out_index, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(B >= 0 && N >= 0, LGK_ERR_BAD_ARG, "lgk_fps: negative size");
  LGK_REQUIRE(dim > 0 && dim <= kMaxDim, LGK_ERR_UNSUPPORTED, "lgk_fps: dim must be in [1, %d]", kMaxDim);
  LGK_REQUIRE(pos_dtype == 0 || pos_dtype == 1, LGK_ERR_BAD_ARG, "lgk_fps: pos_dtype must be 0 (f32) or 1 (f64)");
  if (B == 0 || N == 0) return LGK_OK;
  LGK_REQUIRE(pos && ptr && out_ptr && dist_ws && out_index, LGK_ERR_BAD_ARG, "lgk_fps: null pointer");
  if (pos_dtype == 0)
    fps_kernel<float><<<unsigned(B), 256, 0, stream>>>(static_cast<const float*>(pos), dim, ptr, out_ptr, start, static_cast<float*>(dist_ws), out_index);
  else
    fps_kernel<double><<<unsigned(B), 256, 0, stream>>>(static_cast<const double*>(pos), dim, ptr, out_ptr, start, static_cast<double*>(dist_ws), out_index);
  LGK_CHECK_LAUNCH("fps_kernel");
  return LGK_OK;
}

extern "C" int lgk_radius(const void* x, const void* y, int pos_dtype, int dim, const int32_t* ptr_x, const int32_t* ptr_y,
                          int64_t B, int64_t Nx, int64_t Ny, double r, int max_num_neighbors, int32_t* counts,
                          const int64_t* offsets, int64_t* row_out, int64_t* col_out, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(B >= 0 && Nx >= 0 && Ny >= 0 && r >= 0 && max_num_neighbors > 0, LGK_ERR_BAD_ARG, "lgk_radius: bad argument");
  LGK_REQUIRE(dim > 0 && dim <= kMaxDim, LGK_ERR_UNSUPPORTED, "lgk_radius: dim must be in [1, %d]", kMaxDim);
  LGK_REQUIRE(pos_dtype == 0 || pos_dtype == 1, LGK_ERR_BAD_ARG, "lgk_radius: pos_dtype must be 0 (f32) or 1 (f64)");
  const bool fill = row_out != nullptr;
  LGK_REQUIRE(fill ? (offsets != nullptr && col_out != nullptr) : counts != nullptr, LGK_ERR_BAD_ARG,
              "lgk_radius: counts (pass 1) or offsets + row_out + col_out (pass 2)");
  if (Ny == 0) return LGK_OK;
  LGK_REQUIRE(x && y && ptr_x && ptr_y, LGK_ERR_BAD_ARG, "lgk_radius: null pointer");
  const unsigned grid = unsigned(ceil_div(Ny, 128));
  if (pos_dtype == 0) {
    const float r2 = float(r) * float(r);
    if (fill) radius_bipartite_kernel<float, true><<<grid, 128, 0, stream>>>(static_cast<const float*>(x), static_cast<const float*>(y), dim, ptr_x, ptr_y, int(B), Ny, r2, max_num_neighbors, nullptr, offsets, row_out, col_out);
    else radius_bipartite_kernel<float, false><<<grid, 128, 0, stream>>>(static_cast<const float*>(x), static_cast<const float*>(y), dim, ptr_x, ptr_y, int(B), Ny, r2, max_num_neighbors, counts, nullptr, nullptr, nullptr);
  } else {
    const double r2 = r * r;
    if (fill) radius_bipartite_kernel<double, true><<<grid, 128, 0, stream>>>(static_cast<const double*>(x), static_cast<const double*>(y), dim, ptr_x, ptr_y, int(B), Ny, r2, max_num_neighbors, nullptr, offsets, row_out, col_out);
    else radius_bipartite_kernel<double, false><<<grid, 128, 0, stream>>>(static_cast<const double*>(x), static_cast<const double*>(y), dim, ptr_x, ptr_y, int(B), Ny, r2, max_num_neighbors, counts, nullptr, nullptr, nullptr);
  }
  LGK_CHECK_LAUNCH("radius_bipartite_kernel");
  return LGK_OK;
}

extern "C" int lgk_edge_concat(const float* x, const float* pos_src, const float* pos_dst, const int64_t* src, const int64_t* dst,
                               float* out, int64_t E, int64_t F, int64_t D, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(E >= 0 && F >= 0 && D >= 0 && D <= 32, LGK_ERR_BAD_ARG, "lgk_edge_concat: bad sizes");
  if (E == 0 || F + D == 0) return LGK_OK;
  LGK_REQUIRE((F == 0 || x) && (D == 0 || (pos_src && pos_dst)) && src && dst && out, LGK_ERR_BAD_ARG, "lgk_edge_concat: null pointer");
  edge_concat_kernel<<<unsigned(ceil_div(E, 8)), 256, 0, stream>>>(x, pos_src, pos_dst, src, dst, out, E, int(F), int(D));
  LGK_CHECK_LAUNCH("edge_concat_kernel");
  return LGK_OK;
}
