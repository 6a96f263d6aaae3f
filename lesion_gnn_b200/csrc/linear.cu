// (5) node-feature transform X W^T and its two backward products, fp32 FFMA.
//
// Every Linear inside the PyG convs (`lin`, `lin_l`, `lin_r`, `lin_rel`, `lin_root`, MLP.lins) and the
// reference's in_proj/out_proj (models/gat.py:29,46, gin.py:21,25) is a tall-skinny product: N = 10^4..10^6 rows
// against 16..1025 columns.  Arithmetic intensity is 6..57 flop/B, far below the tensor-core
// ridge, and the fp32 parity bar (1e-5) rules out single-pass TF32, so this path is plain FFMA with the
// epilogues the layers need fused in (row scale for GCN's D^-1/2, bias, ELU, accumulate for the
// root-weight term W_r x_i of SAGEConv / GraphConv).
//
// One kernel template covers the three products as C[i,j] = sum_r A(i,r) B(j,r):
//   forward   C = out [N,M]   A = x  (r contiguous)   B = W  (r contiguous)      R = K
//   grad x    C = dx  [N,K]   A = dy (r contiguous)   B = W  (j contiguous)      R = M
//   grad W    C = dW^T[K,M]   A = x  (i contiguous)   B = dy (j contiguous)      R = N, split over CTAs
// 128 x 64 output tile, 16-deep r tile, 256 threads, 8 x 4 micro-tile, register-prefetch double buffering.
#include "common.cuh"

#include <cooperative_groups.h>
#include <stdlib.h>

namespace lgk {
// tensor-core (3xTF32, tcgen05) forward transform, linear_tc.cu
int launch_linear_tc(const float* x, const float* w, const float* bias, const float* rowscale, float* out, int64_t n,
                     int64_t k, int64_t m, int act, int accumulate, cudaStream_t stream);
// TMA-fed tensor-core transform for 16-byte aligned rows (K % 4 == 0), linear_tma.cu
// tensor-core weight gradient of a wide transform, grad_w_tc.cu
int grad_w_tc_slabs(int64_t n, int64_t k, int64_t m, int64_t* slab_nodes);
bool grad_w_tc_fuses_db(int64_t k);
int launch_grad_w_tc(const float* dy, const float* x, float* partial, int64_t n, int64_t k, int64_t m, int fuse_db,
                     cudaStream_t stream);
int launch_linear_tma(const float* x, const float* w, const float* bias, const float* rowscale, float* out, int64_t n,
                      int64_t k, int64_t m, int act, int accumulate, cudaStream_t stream, bool* done, const GatAlpha* alpha = nullptr);
namespace {

// Products with at least this many rows and a reduction of at least 32 go to the tensor cores; below that the
// launch is latency-bound either way and the FFMA kernel has the cheaper prologue.  LGK_LINEAR_TC=0 forces the
// FFMA path (A/B measurements).
bool use_tensor_cores(int64_t n, int64_t k) {
  static const int mode = [] { const char* v = getenv("LGK_LINEAR_TC"); return (v && *v) ? atoi(v) : 1; }();
  return mode != 0 && n >= 512 && k >= 32;
}

constexpr int BI = 128, BJ = 64, BR = 16, PAD = 4, kThreads = 256;
constexpr int TI = 8, TJ = 4;

// activations (x, out, dy, dx) are float or bf16; weights, biases and every gradient of a parameter stay fp32, as do all
// accumulators.  4 consecutive elements per load: 16 bytes of fp32, 8 bytes of bf16.
template <typename T>
__device__ __forceinline__ float4 ld4(const T* p);
template <>
__device__ __forceinline__ float4 ld4<float>(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
template <>
__device__ __forceinline__ float4 ld4<__nv_bfloat16>(const __nv_bfloat16* p) {
  return bf16x4_to_f4(__ldg(reinterpret_cast<const uint2*>(p)));
}

struct GemmArgs {
  const void* A;
  const void* B;
  void* C;
  int64_t I, J, R;
  int64_t lda, ldb;   // stride of the non-contiguous index of A / B
  int64_t sCi, sCj;   // output strides
  const float* bias;      // [J]
  const float* rowscale;  // [I]
  int act;
  int accumulate;
  int64_t r_chunk;        // rows of R per blockIdx.z (split)
  int64_t split_stride;   // elements between split partials (0 when not split)
  GatAlpha alpha;         // a_src != nullptr: attention-logit halves from the accumulators (forward of GATConv.lin only)
};

// load one [rows x BR] tile of an operand whose element (i, r) sits at base[i * ld + r]  (r contiguous)
template <int ROWS, bool VEC, typename T>
__device__ __forceinline__ void load_rc(const T* __restrict__ base, int64_t ld, int64_t i0, int64_t imax, int64_t r0,
                                        int64_t rmax, float (&reg)[ROWS / 16]) {
  if (VEC) {
    const int r4 = threadIdx.x & 3, ii = threadIdx.x >> 2;  // 64 rows per pass, 4 floats each
#pragma unroll
    for (int u = 0; u < ROWS / 64; ++u) {
      const int64_t i = i0 + ii + 64 * u, r = r0 + r4 * 4;
      float4 v = f4_zero();
      if (i < imax && r < rmax) v = ld4<T>(base + i * ld + r);
      reg[u * 4 + 0] = v.x; reg[u * 4 + 1] = v.y; reg[u * 4 + 2] = v.z; reg[u * 4 + 3] = v.w;
    }
  } else {
    const int rr = threadIdx.x & 15, ii = threadIdx.x >> 4;  // 16 rows per pass
#pragma unroll
    for (int u = 0; u < ROWS / 16; ++u) {
      const int64_t i = i0 + ii + 16 * u, r = r0 + rr;
      reg[u] = (i < imax && r < rmax) ? Vec4<T>::load1(base, i * ld + r) : 0.f;
    }
  }
}
template <int ROWS, bool VEC>
__device__ __forceinline__ void store_rc(float (*s)[ROWS + PAD], const float (&reg)[ROWS / 16]) {
  if (VEC) {
    const int r4 = threadIdx.x & 3, ii = threadIdx.x >> 2;
#pragma unroll
    for (int u = 0; u < ROWS / 64; ++u)
#pragma unroll
      for (int c = 0; c < 4; ++c) s[r4 * 4 + c][ii + 64 * u] = reg[u * 4 + c];
  } else {
    const int rr = threadIdx.x & 15, ii = threadIdx.x >> 4;
#pragma unroll
    for (int u = 0; u < ROWS / 16; ++u) s[rr][ii + 16 * u] = reg[u];
  }
}
// operand whose element (i, r) sits at base[r * ld + i]  (i contiguous)
template <int ROWS, bool VEC, typename T>
__device__ __forceinline__ void load_ic(const T* __restrict__ base, int64_t ld, int64_t i0, int64_t imax, int64_t r0,
                                        int64_t rmax, float (&reg)[ROWS / 16]) {
  if (VEC) {
    constexpr int V = ROWS / 4;           // float4 per r
    constexpr int RP = kThreads / V;      // r per pass
    const int i4 = threadIdx.x % V, rr = threadIdx.x / V;
#pragma unroll
    for (int u = 0; u < BR / RP; ++u) {
      const int64_t i = i0 + i4 * 4, r = r0 + rr + RP * u;
      float4 v = f4_zero();
      if (i < imax && r < rmax) v = ld4<T>(base + r * ld + i);
      reg[u * 4 + 0] = v.x; reg[u * 4 + 1] = v.y; reg[u * 4 + 2] = v.z; reg[u * 4 + 3] = v.w;
    }
  } else {
    constexpr int RP = kThreads / ROWS;
    const int ii = threadIdx.x % ROWS, rr = threadIdx.x / ROWS;
#pragma unroll
    for (int u = 0; u < BR / RP; ++u) {
      const int64_t i = i0 + ii, r = r0 + rr + RP * u;
      reg[u] = (i < imax && r < rmax) ? Vec4<T>::load1(base, r * ld + i) : 0.f;
    }
  }
}
template <int ROWS, bool VEC>
__device__ __forceinline__ void store_ic(float (*s)[ROWS + PAD], const float (&reg)[ROWS / 16]) {
  if (VEC) {
    constexpr int V = ROWS / 4;
    constexpr int RP = kThreads / V;
    const int i4 = threadIdx.x % V, rr = threadIdx.x / V;
#pragma unroll
    for (int u = 0; u < BR / RP; ++u)
      *reinterpret_cast<float4*>(&s[rr + RP * u][i4 * 4]) =
          make_float4(reg[u * 4 + 0], reg[u * 4 + 1], reg[u * 4 + 2], reg[u * 4 + 3]);
  } else {
    constexpr int RP = kThreads / ROWS;
    const int ii = threadIdx.x % ROWS, rr = threadIdx.x / ROWS;
#pragma unroll
    for (int u = 0; u < BR / RP; ++u) s[rr + RP * u][ii] = reg[u];
  }
}

template <bool A_RC, bool B_RC, bool VEC, typename TA, typename TB, typename TC>
__global__ void __launch_bounds__(kThreads, 2) gemm_kernel(GemmArgs g) {
  const TA* gA = static_cast<const TA*>(g.A);
  const TB* gB = static_cast<const TB*>(g.B);
  __shared__ __align__(16) float As[BR][BI + PAD];
  __shared__ __align__(16) float Bs[BR][BJ + PAD];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int64_t i0 = int64_t(blockIdx.x) * BI, j0 = int64_t(blockIdx.y) * BJ;
  const int64_t rbeg = int64_t(blockIdx.z) * g.r_chunk;
  const int64_t rend = min(g.R, rbeg + g.r_chunk);

  float acc[TI][TJ];
#pragma unroll
  for (int a = 0; a < TI; ++a)
#pragma unroll
    for (int b = 0; b < TJ; ++b) acc[a][b] = 0.f;

  float ra[BI / 16], rb[BJ / 16];
  auto gload = [&](int64_t r0) {
    if (A_RC) load_rc<BI, VEC>(gA, g.lda, i0, g.I, r0, rend, ra); else load_ic<BI, VEC>(gA, g.lda, i0, g.I, r0, rend, ra);
    if (B_RC) load_rc<BJ, VEC>(gB, g.ldb, j0, g.J, r0, rend, rb); else load_ic<BJ, VEC>(gB, g.ldb, j0, g.J, r0, rend, rb);
  };
  auto sstore = [&]() {
    if (A_RC) store_rc<BI, VEC>(As, ra); else store_ic<BI, VEC>(As, ra);
    if (B_RC) store_rc<BJ, VEC>(Bs, rb); else store_ic<BJ, VEC>(Bs, rb);
  };

  if (rbeg < rend) gload(rbeg);
  for (int64_t r0 = rbeg; r0 < rend; r0 += BR) {
    sstore();
    __syncthreads();
    if (r0 + BR < rend) gload(r0 + BR);  // prefetch the next tile into registers while this one is multiplied
#pragma unroll
    for (int r = 0; r < BR; ++r) {
      const float4 a0 = *reinterpret_cast<const float4*>(&As[r][ty * TI]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[r][ty * TI + 4]);
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[r][tx * TJ]);
      const float av[TI] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float bv[TJ] = {b0.x, b0.y, b0.z, b0.w};
#pragma unroll
      for (int a = 0; a < TI; ++a)
#pragma unroll
        for (int b = 0; b < TJ; ++b) acc[a][b] = fmaf(av[a], bv[b], acc[a][b]);
    }
    __syncthreads();
  }

  TC* C = static_cast<TC*>(g.C) + int64_t(blockIdx.z) * g.split_stride;
  const bool split = g.split_stride != 0;
  if (g.alpha.a_src) {
    // a_src[i, head] = <acc row i, att_src> over the head's columns: chan % 4 == 0 and chan | BJ, so the chan / 4 threads
    // that hold a head's columns of a row are adjacent lanes of one 16-lane half warp (xor shuffles stay inside it)
    const int lanes = g.alpha.chan >> 2;
    const int64_t j = j0 + tx * TJ;
    float4 as4 = f4_zero(), ad4 = f4_zero();
    if (j < g.J) {
      as4 = __ldg(reinterpret_cast<const float4*>(g.alpha.att_src + j));
      ad4 = __ldg(reinterpret_cast<const float4*>(g.alpha.att_dst + j));
    }
#pragma unroll
    for (int a = 0; a < TI; ++a) {
      float s = fmaf(acc[a][0], as4.x, fmaf(acc[a][1], as4.y, fmaf(acc[a][2], as4.z, acc[a][3] * as4.w)));
      float d = fmaf(acc[a][0], ad4.x, fmaf(acc[a][1], ad4.y, fmaf(acc[a][2], ad4.z, acc[a][3] * ad4.w)));
      for (int off = 1; off < lanes; off <<= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, off);
        d += __shfl_xor_sync(0xffffffffu, d, off);
      }
      const int64_t i = i0 + ty * TI + a;
      if ((tx & (lanes - 1)) == 0 && i < g.I && j < g.J) {
        const int head = int(j) / g.alpha.chan;
        g.alpha.a_src[i * g.alpha.heads + head] = s;
        g.alpha.a_dst[i * g.alpha.heads + head] = d;
      }
    }
  }
#pragma unroll
  for (int a = 0; a < TI; ++a) {
    const int64_t i = i0 + ty * TI + a;
    if (i >= g.I) continue;
    const float rs = (!split && g.rowscale) ? __ldg(g.rowscale + i) : 1.f;
#pragma unroll
    for (int b = 0; b < TJ; ++b) {
      const int64_t j = j0 + tx * TJ + b;
      if (j >= g.J) continue;
      float v = acc[a][b];
      const int64_t at = i * g.sCi + j * g.sCj;
      if (!split) {
        v *= rs;
        if (g.bias) v += __ldg(g.bias + j);
        if (g.accumulate) v += Vec4<TC>::load1(C, at);
        if (g.act == LGK_ACT_ELU) v = elu1(v);
        else if (g.act == LGK_ACT_RELU) v = fmaxf(v, 0.f);
      }
      Vec4<TC>::store1(C, at, v);
    }
  }
}

// second stage of grad W: dW[m,k] (+)= sum_s partial[s][k*M + m]
__global__ void __launch_bounds__(256) splitk_reduce_kernel(const float* __restrict__ partial, float* __restrict__ out,
                                                            int64_t count, int splits, int accumulate) {
  const int64_t t = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (t >= count) return;
  float s = 0.f;
  for (int k = 0; k < splits; ++k) s += partial[int64_t(k) * count + t];
  if (accumulate) s += out[t];
  out[t] = s;
}

// ---- small-weight gradients: dW [m,k] and db [m] in one pass, k, m <= 128 ------------------------------------------
// Every Linear inside the convs has a tiny weight (64 x 64) and a very tall reduction (N nodes).  One launch of at
// most 2 CTAs per SM: a CTA reduces a contiguous slab of rows -- 32-row chunks of x and dy staged in shared memory,
// 4 x 4 register micro-tiles per 64 x 64 output block, db as a by-product of the dy tiles already in registers --
// and writes its partial [m*k + m] to the workspace; a second tiny launch adds the partials in slab order
// (deterministic, no atomics).  This replaced a generic split GEMM + a separate column-sum pass that together cost
// ~50 us per layer on the headline batch.
constexpr int kSlabChunk = 32;
template <int KB, int MB, bool VEC, typename T>  // 64-wide output blocks along k and m; VEC: 4-element loads (K, M % 4 == 0, aligned)
__global__ void __launch_bounds__(256) bwd_w_slab_kernel(const T* __restrict__ dy, const T* __restrict__ x,
                                                         float* __restrict__ partial, int64_t N, int K, int M,
                                                         int64_t slab_rows, int want_db) {
  extern __shared__ __align__(16) float sm[];
  const int kp = KB * 64, mp = MB * 64;
  float* xs = sm;                       // [kSlabChunk][kp]
  float* ds = sm + kSlabChunk * kp;     // [kSlabChunk][mp]
  const int tk = threadIdx.x >> 4, tm = threadIdx.x & 15;  // 16 x 16 threads, 4 x 4 outputs each per block
  const int64_t r0 = int64_t(blockIdx.x) * slab_rows;
  const int64_t r1 = min(N, r0 + slab_rows);
  float acc[KB][MB][4][4];
  float dbs[MB][4];
#pragma unroll
  for (int a = 0; a < KB; ++a)
#pragma unroll
    for (int b = 0; b < MB; ++b)
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[a][b][i][j] = 0.f;
#pragma unroll
  for (int b = 0; b < MB; ++b)
#pragma unroll
    for (int j = 0; j < 4; ++j) dbs[b][j] = 0.f;

  for (int64_t c0 = r0; c0 < r1; c0 += kSlabChunk) {
    const int rows = (r1 - c0 < kSlabChunk) ? int(r1 - c0) : kSlabChunk;
    __syncthreads();
    if (VEC) {
      for (int e = threadIdx.x; e < kSlabChunk * kp / 4; e += 256) {
        const int r = e / (kp / 4), c = (e - r * (kp / 4)) * 4;
        reinterpret_cast<float4*>(xs)[e] = (r < rows && c < K) ? ld4<T>(x + (c0 + r) * K + c) : f4_zero();
      }
      for (int e = threadIdx.x; e < kSlabChunk * mp / 4; e += 256) {
        const int r = e / (mp / 4), c = (e - r * (mp / 4)) * 4;
        reinterpret_cast<float4*>(ds)[e] = (r < rows && c < M) ? ld4<T>(dy + (c0 + r) * M + c) : f4_zero();
      }
    } else {
      for (int e = threadIdx.x; e < kSlabChunk * kp; e += 256) {
        const int r = e / kp, c = e - r * kp;
        xs[e] = (r < rows && c < K) ? Vec4<T>::load1(x, (c0 + r) * K + c) : 0.f;
      }
      for (int e = threadIdx.x; e < kSlabChunk * mp; e += 256) {
        const int r = e / mp, c = e - r * mp;
        ds[e] = (r < rows && c < M) ? Vec4<T>::load1(dy, (c0 + r) * M + c) : 0.f;
      }
    }
    __syncthreads();
#pragma unroll 4
    for (int r = 0; r < kSlabChunk; ++r) {
      float4 a4[KB], b4[MB];
#pragma unroll
      for (int a = 0; a < KB; ++a) a4[a] = *reinterpret_cast<const float4*>(xs + r * kp + a * 64 + tk * 4);
#pragma unroll
      for (int b = 0; b < MB; ++b) b4[b] = *reinterpret_cast<const float4*>(ds + r * mp + b * 64 + tm * 4);
#pragma unroll
      for (int a = 0; a < KB; ++a) {
        const float av[4] = {a4[a].x, a4[a].y, a4[a].z, a4[a].w};
#pragma unroll
        for (int b = 0; b < MB; ++b) {
          const float bv[4] = {b4[b].x, b4[b].y, b4[b].z, b4[b].w};
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[a][b][i][j] = fmaf(av[i], bv[j], acc[a][b][i][j]);
        }
      }
      if (tk == 0) {
#pragma unroll
        for (int b = 0; b < MB; ++b) { dbs[b][0] += b4[b].x; dbs[b][1] += b4[b].y; dbs[b][2] += b4[b].z; dbs[b][3] += b4[b].w; }
      }
    }
  }
  // partial layout = the outputs' own: dW[mi * K + ki] then db[mi]
  float* p = partial + int64_t(blockIdx.x) * (int64_t(K) * M + M);
#pragma unroll
  for (int a = 0; a < KB; ++a)
#pragma unroll
    for (int b = 0; b < MB; ++b)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int mi = b * 64 + tm * 4 + j;
        if (mi >= M) continue;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int ki = a * 64 + tk * 4 + i;
          if (ki < K) p[int64_t(mi) * K + ki] = acc[a][b][i][j];
        }
      }
  if (want_db && tk == 0) {
#pragma unroll
    for (int b = 0; b < MB; ++b)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int mi = b * 64 + tm * 4 + j;
        if (mi < M) p[int64_t(K) * M + mi] = dbs[b][j];
      }
  }
}

// The same slab reduction with the 32-row chunks of x and dy streamed through a three-stage cp.async ring (fp32, 16-byte
// aligned rows): chunk c + 2 is in flight while chunk c is multiplied.  r02: the single-buffered kernel above (load, barrier,
// multiply, barrier) moved 1.08 TB/s on config 3's 409 600-row products, 26 % of the GIN step.
constexpr int kPipeStages = 3;
__device__ __forceinline__ void cp_async_16_zfill(float* smem_dst, const float* gmem_src, bool valid) {
  const int bytes = valid ? 16 : 0;  // src-size 0: the 16 bytes are zero-filled, nothing is read
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(uint32_t(__cvta_generic_to_shared(smem_dst))), "l"(gmem_src), "r"(bytes)
               : "memory");
}
template <int KB, int MB>
__global__ void __launch_bounds__(256) bwd_w_slab_pipe_kernel(const float* __restrict__ dy, const float* __restrict__ x,
                                                              float* __restrict__ partial, int64_t N, int K, int M,
                                                              int64_t slab_rows, int want_db) {
  extern __shared__ __align__(16) float sm[];
  constexpr int kp = KB * 64, mp = MB * 64, stage_f = kSlabChunk * (kp + mp);
  const int tk = threadIdx.x >> 4, tm = threadIdx.x & 15;
  const int64_t r0 = int64_t(blockIdx.x) * slab_rows;
  const int64_t r1 = min(N, r0 + slab_rows);
  const int nchunks = r1 > r0 ? int((r1 - r0 + kSlabChunk - 1) / kSlabChunk) : 0;
  float acc[KB][MB][4][4];
  float dbs[MB][4];
#pragma unroll
  for (int a = 0; a < KB; ++a)
#pragma unroll
    for (int b = 0; b < MB; ++b)
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[a][b][i][j] = 0.f;
#pragma unroll
  for (int b = 0; b < MB; ++b)
#pragma unroll
    for (int j = 0; j < 4; ++j) dbs[b][j] = 0.f;

  auto issue = [&](int ci) {
    if (ci < nchunks) {
      float* xs = sm + (ci % kPipeStages) * stage_f;
      float* ds = xs + kSlabChunk * kp;
      const int64_t c0 = r0 + int64_t(ci) * kSlabChunk;
      const int rows = (r1 - c0 < kSlabChunk) ? int(r1 - c0) : kSlabChunk;
      for (int e = threadIdx.x; e < kSlabChunk * kp / 4; e += 256) {
        const int r = e / (kp / 4), c = (e - r * (kp / 4)) * 4;
        const bool ok = r < rows && c < K;
        cp_async_16_zfill(xs + e * 4, ok ? x + (c0 + r) * K + c : x, ok);
      }
      for (int e = threadIdx.x; e < kSlabChunk * mp / 4; e += 256) {
        const int r = e / (mp / 4), c = (e - r * (mp / 4)) * 4;
        const bool ok = r < rows && c < M;
        cp_async_16_zfill(ds + e * 4, ok ? dy + (c0 + r) * M + c : dy, ok);
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");  // an empty group past the end keeps the wait count uniform
  };
#pragma unroll
  for (int st = 0; st < kPipeStages - 1; ++st) issue(st);
  for (int ci = 0; ci < nchunks; ++ci) {
    issue(ci + kPipeStages - 1);
    asm volatile("cp.async.wait_group %0;" ::"n"(kPipeStages - 1) : "memory");
    __syncthreads();
    const float* xs = sm + (ci % kPipeStages) * stage_f;
    const float* ds = xs + kSlabChunk * kp;
#pragma unroll 4
    for (int r = 0; r < kSlabChunk; ++r) {
      float4 a4[KB], b4[MB];
#pragma unroll
      for (int a = 0; a < KB; ++a) a4[a] = *reinterpret_cast<const float4*>(xs + r * kp + a * 64 + tk * 4);
#pragma unroll
      for (int b = 0; b < MB; ++b) b4[b] = *reinterpret_cast<const float4*>(ds + r * mp + b * 64 + tm * 4);
#pragma unroll
      for (int a = 0; a < KB; ++a) {
        const float av[4] = {a4[a].x, a4[a].y, a4[a].z, a4[a].w};
#pragma unroll
        for (int b = 0; b < MB; ++b) {
          const float bv[4] = {b4[b].x, b4[b].y, b4[b].z, b4[b].w};
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[a][b][i][j] = fmaf(av[i], bv[j], acc[a][b][i][j]);
        }
      }
      if (tk == 0) {
#pragma unroll
        for (int b = 0; b < MB; ++b) { dbs[b][0] += b4[b].x; dbs[b][1] += b4[b].y; dbs[b][2] += b4[b].z; dbs[b][3] += b4[b].w; }
      }
    }
    __syncthreads();  // the stage is refilled by the next iteration's issue
  }
  float* p = partial + int64_t(blockIdx.x) * (int64_t(K) * M + M);
#pragma unroll
  for (int a = 0; a < KB; ++a)
#pragma unroll
    for (int b = 0; b < MB; ++b)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int mi = b * 64 + tm * 4 + j;
        if (mi >= M) continue;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int ki = a * 64 + tk * 4 + i;
          if (ki < K) p[int64_t(mi) * K + ki] = acc[a][b][i][j];
        }
      }
  if (want_db && tk == 0) {
#pragma unroll
    for (int b = 0; b < MB; ++b)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int mi = b * 64 + tm * 4 + j;
        if (mi < M) p[int64_t(K) * M + mi] = dbs[b][j];
      }
  }
}

// ---- small-weight gradient (k, m <= 64) in ONE launch: slab partials reduced through distributed shared memory ------
// r01/r02's two-launch form (bwd_w_slab_kernel + slab_reduce_kernel) wrote ~180 partials of 16.6 KB and read them back in a
// second launch: 1.48x the algorithmic DRAM traffic and two graph nodes per weight.  Here 8 consecutive slabs form a
// thread-block CLUSTER: every CTA leaves its partial in its own shared memory, the cluster's CTAs add the eight partials
// through DSMEM (CTA r owns every 8th block of 256 outputs, ranks added in rank order) and write ONE partial per cluster; the
// cluster that finishes last (device-scope counter) adds the ~23 cluster partials in cluster order into dW / db.  Fixed
// summation tree: deterministic.  fp32, any alignment (the VEC flag only selects the load width).
constexpr int kDwCluster = 8;
template <bool VEC>
__global__ void __cluster_dims__(kDwCluster, 1, 1) __launch_bounds__(256)
    bwd_w_cluster_kernel(const float* __restrict__ dy, const float* __restrict__ x, float* __restrict__ cpartial,
                         unsigned* __restrict__ counter, float* __restrict__ dw, float* __restrict__ db, int64_t N, int K, int M,
                         int64_t slab_rows, int want_db, int accumulate) {
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  __shared__ __align__(16) float xs[kSlabChunk * 64];
  __shared__ __align__(16) float ds[kSlabChunk * 64];
  __shared__ float part[64 * 64 + 64];
  __shared__ int is_last;
  const int tk = threadIdx.x >> 4, tm = threadIdx.x & 15;
  const int64_t r0 = int64_t(blockIdx.x) * slab_rows;
  const int64_t r1 = min(N, r0 + slab_rows);
  float acc[4][4], dbs[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    dbs[i] = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  }
  for (int64_t c0 = r0; c0 < r1; c0 += kSlabChunk) {
    const int rows = (r1 - c0 < kSlabChunk) ? int(r1 - c0) : kSlabChunk;
    __syncthreads();
    if (VEC) {
      for (int e = threadIdx.x; e < kSlabChunk * 16; e += 256) {
        const int r = e >> 4, c = (e & 15) * 4;
        reinterpret_cast<float4*>(xs)[e] = (r < rows && c < K) ? __ldg(reinterpret_cast<const float4*>(x + (c0 + r) * K + c)) : f4_zero();
        reinterpret_cast<float4*>(ds)[e] = (r < rows && c < M) ? __ldg(reinterpret_cast<const float4*>(dy + (c0 + r) * M + c)) : f4_zero();
      }
    } else {
      for (int e = threadIdx.x; e < kSlabChunk * 64; e += 256) {
        const int r = e >> 6, c = e & 63;
        xs[e] = (r < rows && c < K) ? __ldg(x + (c0 + r) * K + c) : 0.f;
        ds[e] = (r < rows && c < M) ? __ldg(dy + (c0 + r) * M + c) : 0.f;
      }
    }
    __syncthreads();
#pragma unroll 4
    for (int r = 0; r < kSlabChunk; ++r) {
      const float4 a4 = *reinterpret_cast<const float4*>(xs + r * 64 + tk * 4);
      const float4 b4 = *reinterpret_cast<const float4*>(ds + r * 64 + tm * 4);
      const float av[4] = {a4.x, a4.y, a4.z, a4.w}, bv[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
      if (tk == 0) { dbs[0] += b4.x; dbs[1] += b4.y; dbs[2] += b4.z; dbs[3] += b4.w; }
    }
  }
  // own partial, in the outputs' own layout: dW[mi * K + ki], then db[mi]
  const int total = K * M + (want_db ? M : 0);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int mi = tm * 4 + j;
    if (mi >= M) continue;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int ki = tk * 4 + i;
      if (ki < K) part[mi * K + ki] = acc[i][j];
    }
    if (want_db && tk == 0) part[K * M + mi] = dbs[j];
  }
  cluster.sync();
  const unsigned rank = cluster.block_rank();
  const int cid = blockIdx.x / kDwCluster, nclusters = gridDim.x / kDwCluster;
  const float* peer[kDwCluster];
#pragma unroll
  for (int c = 0; c < kDwCluster; ++c) peer[c] = cluster.map_shared_rank(part, c);
  for (int o = int(rank) * 256 + threadIdx.x; o < total; o += kDwCluster * 256) {
    float sum = 0.f;
#pragma unroll
    for (int c = 0; c < kDwCluster; ++c) sum += peer[c][o];
    if (nclusters > 1) {
      cpartial[int64_t(cid) * total + o] = sum;
    } else {
      float* dst = o < K * M ? dw + o : db + (o - K * M);
      *dst = accumulate ? *dst + sum : sum;
    }
  }
  if (nclusters == 1) {
    cluster.sync();  // nobody leaves while a peer still reads its shared memory
    return;
  }
  __threadfence();
  cluster.sync();    // the cluster's partial is complete (and fenced), all remote reads are done
  if (rank == 0 && threadIdx.x == 0) {
    const unsigned ticket = atomicAdd(counter, 1u);
    const int last = ticket == unsigned(nclusters - 1);
    if (last) *counter = 0u;  // ready for the next call on this workspace
#pragma unroll
    for (int c = 0; c < kDwCluster; ++c) *cluster.map_shared_rank(&is_last, c) = last;
  }
  cluster.sync();
  if (!is_last) return;
  __threadfence();
  for (int o = int(rank) * 256 + threadIdx.x; o < total; o += kDwCluster * 256) {
    float sum = 0.f;
    for (int c = 0; c < nclusters; ++c) sum += __ldcg(cpartial + int64_t(c) * total + o);  // cluster order
    float* dst = o < K * M ? dw + o : db + (o - K * M);
    *dst = accumulate ? *dst + sum : sum;
  }
}

// db[m] = sum_n dy[n,m] alone (bias gradients of the fused aggregation kernels): slab partials, same second stage.
// Threads = (row lane, float4 column); M % 4 == 0 fast path, scalar otherwise.
template <typename T>
__global__ void __launch_bounds__(256) colsum_slab_kernel(const T* __restrict__ dy, float* __restrict__ partial,
                                                          int64_t N, int M, int64_t slab_rows) {
  __shared__ float red[256 * 4];
  const int64_t r0 = int64_t(blockIdx.x) * slab_rows, r1 = min(N, r0 + slab_rows);
  float* p = partial + int64_t(blockIdx.x) * M;
  const bool vec = (M & 3) == 0 && ((reinterpret_cast<uintptr_t>(dy) & (4 * sizeof(T) - 1)) == 0);
  const int cols = vec ? M / 4 : M;  // column groups
  for (int c0 = 0; c0 < cols; c0 += 256) {
    const int width = min(256, cols - c0);   // column groups handled in this pass
    const int lanes = max(1, 256 / width);   // row lanes
    const int c = threadIdx.x % width, rl = threadIdx.x / width;
    float4 acc = f4_zero();
    if (rl < lanes) {
      if (vec) {
        for (int64_t r = r0 + rl; r < r1; r += lanes) {
          const float4 v = ld4<T>(dy + r * M + 4 * (c0 + c));
          acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
      } else {
        for (int64_t r = r0 + rl; r < r1; r += lanes) acc.x += Vec4<T>::load1(dy, r * M + c0 + c);
      }
    }
    __syncthreads();
    reinterpret_cast<float4*>(red)[threadIdx.x] = acc;
    __syncthreads();
    if (rl == 0) {
      for (int l = 1; l < lanes; ++l) {
        const float4 v = reinterpret_cast<const float4*>(red)[l * width + c];
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
      if (vec) reinterpret_cast<float4*>(p)[c0 + c] = acc;
      else p[c0 + c] = acc.x;
    }
  }
}

// slabs for a reduction over n rows: at most two CTAs per SM, at least 64 rows each
int slab_count(int64_t n) {
  // rows per slab (tuning knob LGK_SLAB_ROWS): fewer, taller slabs mean fewer partials for the second stage to read back
  static const int min_rows = [] { const char* v = getenv("LGK_SLAB_ROWS"); const int r = (v && *v) ? atoi(v) : 64; return r < 32 ? 32 : r; }();
  int64_t g = ceil_div(n, min_rows);
  if (g > 2 * kNumSMs) g = 2 * kNumSMs;
  return int(g < 1 ? 1 : g);
}
// ---- products the tiled kernel is the wrong shape for (measured in-graph on the headline batch, profiles/chain_bench.py:
//      out_proj forward [256 x 128] -> 1: 12.8 us, its dX 7.2 us, the 64 x 64 dX of a layer over 11 k rows 10.6 us -- eight
//      dependent 16-deep reduction steps in two CTAs, or 90 CTAs on 148 SMs) ---------------------------------------------
constexpr int kNarrowMax = 8;   // "narrow": at most this many output (forward) / input (dX) columns

// forward with m <= 8 outputs: a warp per row of x, lanes stride over k with 16-byte loads, one shuffle reduction per output
template <typename T, bool VEC>
__global__ void __launch_bounds__(256) linear_fwd_narrow_kernel(const T* __restrict__ x, const float* __restrict__ w,
                                                                const float* __restrict__ bias, const float* __restrict__ rowscale,
                                                                T* __restrict__ out, int64_t n, int k, int m, int act, int accumulate) {
  const int lane = threadIdx.x & 31;
  const int64_t i = int64_t(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (i >= n) return;
  float acc[kNarrowMax];
#pragma unroll
  for (int j = 0; j < kNarrowMax; ++j) acc[j] = 0.f;
  const T* xr = x + i * k;
  if (VEC) {
    for (int r = lane * 4; r < k; r += 128) {
      const float4 xv = ld4<T>(xr + r);
#pragma unroll
      for (int j = 0; j < kNarrowMax; ++j)
        if (j < m) acc[j] += f4_dot(xv, __ldg(reinterpret_cast<const float4*>(w + int64_t(j) * k + r)));
    }
  } else {
    for (int r = lane; r < k; r += 32) {
      const float xv = Vec4<T>::load1(xr, r);
#pragma unroll
      for (int j = 0; j < kNarrowMax; ++j)
        if (j < m) acc[j] = fmaf(xv, __ldg(w + int64_t(j) * k + r), acc[j]);
    }
  }
#pragma unroll
  for (int j = 0; j < kNarrowMax; ++j) {
    if (j >= m) break;  // m is uniform across the warp
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], o);
  }
  if (lane == 0) {
    const float rs = rowscale ? __ldg(rowscale + i) : 1.f;
#pragma unroll
    for (int j = 0; j < kNarrowMax; ++j) {
      if (j >= m) break;
      float v = acc[j] * rs;
      if (bias) v += __ldg(bias + j);
      if (accumulate) v += Vec4<T>::load1(out, i * m + j);
      if (act == LGK_ACT_ELU) v = elu1(v);
      else if (act == LGK_ACT_RELU) v = fmaxf(v, 0.f);
      Vec4<T>::store1(out, i * m + j, v);
    }
  }
}

// dX with m <= 8 columns of dy: dx[i, r] = sum_j dy[i, j] W[j, r] -- an outer-product shaped, element-wise kernel
template <typename T, bool VEC>
__global__ void __launch_bounds__(256) linear_bwd_x_narrow_kernel(const T* __restrict__ dy, const float* __restrict__ w,
                                                                  T* __restrict__ dx, int64_t n, int k, int m, int accumulate) {
  const int per_row = VEC ? (k >> 2) : k;
  const int64_t t = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (t >= n * per_row) return;
  const int64_t i = t / per_row;
  const int c = int(t - i * per_row);
  float dv[kNarrowMax];
#pragma unroll
  for (int j = 0; j < kNarrowMax; ++j) dv[j] = j < m ? Vec4<T>::load1(dy, i * m + j) : 0.f;
  if (VEC) {
    float4 acc = f4_zero();
#pragma unroll
    for (int j = 0; j < kNarrowMax; ++j)
      if (j < m) f4_fma(acc, dv[j], __ldg(reinterpret_cast<const float4*>(w + int64_t(j) * k) + c));
    if (accumulate) {
      const float4 o = ld4<T>(dx + i * k + c * 4);
      acc.x += o.x; acc.y += o.y; acc.z += o.z; acc.w += o.w;
    }
    Vec4<T>::store(dx, i * k + c * 4, acc);
  } else {
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < kNarrowMax; ++j)
      if (j < m) acc = fmaf(dv[j], __ldg(w + int64_t(j) * k + c), acc);
    if (accumulate) acc += Vec4<T>::load1(dx, i * k + c);
    Vec4<T>::store1(dx, i * k + c, acc);
  }
}

// dX against a small weight (k, m <= 128, both multiples of 4): dx[n, k] = dy[n, m] W[m, k].  The whole weight and a tile of
// dy rows go to shared memory with ONE batch of loads per thread (a single round trip instead of m / 16 dependent ones), a
// thread owns 4 rows x 4 columns, a CTA of 128 threads 4 * 128 / (k / 4 rounded up to a power of two) rows: 32 rows at k = 64,
// i.e. 358 CTAs for the 11 k nodes of the headline batch instead of 90.
constexpr int kDxThreads = 128;
int dx_small_col_groups(int64_t k) {
  int cg = 1;
  while (cg * 4 < k) cg <<= 1;
  return cg;  // 1 .. 32
}
size_t dx_small_smem(int64_t k, int64_t m) {
  const int rows = 4 * (kDxThreads / dx_small_col_groups(k));
  return size_t(m * k + rows * (m + 4)) * sizeof(float);
}
template <typename T>
__global__ void __launch_bounds__(kDxThreads) linear_bwd_x_small_kernel(const T* __restrict__ dy, const float* __restrict__ w,
                                                                        T* __restrict__ dx, int64_t n, int k, int m, int cgp,
                                                                        int accumulate) {
  extern __shared__ __align__(16) float dxs[];
  float* Ws = dxs;              // [m][k]
  float* Ds = dxs + m * k;      // [rows][m + 4]
  const int rows = 4 * (kDxThreads / cgp), ds = m + 4;
  const int64_t i0 = int64_t(blockIdx.x) * rows;
  for (int t = threadIdx.x; t < (m * k) >> 2; t += kDxThreads)
    reinterpret_cast<float4*>(Ws)[t] = __ldg(reinterpret_cast<const float4*>(w) + t);
  const int m4 = m >> 2;
  for (int t = threadIdx.x; t < rows * m4; t += kDxThreads) {
    const int rr = t / m4, c4 = t - rr * m4;
    const int64_t i = i0 + rr;
    const float4 v = i < n ? ld4<T>(dy + i * m + c4 * 4) : f4_zero();
    *reinterpret_cast<float4*>(Ds + rr * ds + c4 * 4) = v;
  }
  __syncthreads();
  const int cg = threadIdx.x & (cgp - 1), rg = threadIdx.x / cgp;
  if (cg * 4 >= k) return;
  float4 acc[4];
#pragma unroll
  for (int a = 0; a < 4; ++a) acc[a] = f4_zero();
  const float* d0 = Ds + (rg * 4) * ds;
  for (int r = 0; r < m; r += 4) {
    float4 dv[4], wv[4];
#pragma unroll
    for (int a = 0; a < 4; ++a) dv[a] = *reinterpret_cast<const float4*>(d0 + a * ds + r);
#pragma unroll
    for (int q = 0; q < 4; ++q) wv[q] = *reinterpret_cast<const float4*>(Ws + (r + q) * k + cg * 4);
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      f4_fma(acc[a], dv[a].x, wv[0]);
      f4_fma(acc[a], dv[a].y, wv[1]);
      f4_fma(acc[a], dv[a].z, wv[2]);
      f4_fma(acc[a], dv[a].w, wv[3]);
    }
  }
#pragma unroll
  for (int a = 0; a < 4; ++a) {
    const int64_t i = i0 + rg * 4 + a;
    if (i >= n) break;
    if (accumulate) {
      const float4 o = ld4<T>(dx + i * k + cg * 4);
      acc[a].x += o.x; acc[a].y += o.y; acc[a].z += o.z; acc[a].w += o.w;
    }
    Vec4<T>::store(dx, i * k + cg * 4, acc[a]);
  }
}

template <typename T>
int launch_bwd_x_small(const T* dy, const float* w, T* dx, int64_t n, int64_t k, int64_t m, int accumulate, cudaStream_t stream) {
  const size_t smem = dx_small_smem(k, m);
  if (smem > 48 * 1024) {
    static bool configured_dev[64] = {};
    int dev_ord = 0;
    cudaGetDevice(&dev_ord);
    if (!configured_dev[dev_ord & 63]) {
      if (cudaFuncSetAttribute(linear_bwd_x_small_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024) != cudaSuccess) {
        set_error("linear_bwd_x_small_kernel: cannot reserve shared memory");
        return LGK_ERR_LAUNCH;
      }
      configured_dev[dev_ord & 63] = true;
    }
  }
  const int cgp = dx_small_col_groups(k);
  const int rows = 4 * (kDxThreads / cgp);
  linear_bwd_x_small_kernel<T><<<unsigned(ceil_div(n, rows)), kDxThreads, smem, stream>>>(dy, w, dx, n, int(k), int(m), cgp, accumulate);
  LGK_CHECK_LAUNCH("linear_bwd_x_small_kernel");
  return LGK_OK;
}

bool small_weight(int64_t k, int64_t m) { return k <= 128 && m <= 128; }

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <bool A_RC, bool B_RC, typename TA, typename TB, typename TC>
int launch_gemm(const GemmArgs& g, int splits, bool vec, cudaStream_t stream, const char* name) {
  dim3 grid(unsigned(ceil_div(g.I, BI)), unsigned(ceil_div(g.J, BJ)), unsigned(splits));
  if (vec) gemm_kernel<A_RC, B_RC, true, TA, TB, TC><<<grid, kThreads, 0, stream>>>(g);
  else gemm_kernel<A_RC, B_RC, false, TA, TB, TC><<<grid, kThreads, 0, stream>>>(g);
  LGK_CHECK_LAUNCH(name);
  return LGK_OK;
}
using bf16 = __nv_bfloat16;
bool known_dtype(int dtype) { return dtype == LGK_F32 || dtype == LGK_BF16; }

// LGK_GRAD_W_TC=0 keeps the weight gradient of wide transforms on the FFMA split GEMM (profiling only)
bool use_grad_w_tc() {
  static const bool on = [] { const char* v = getenv("LGK_GRAD_W_TC"); return !(v && *v == '0'); }();
  return on;
}

int choose_wsplits(int64_t n, int64_t k, int64_t m) {
  const int64_t tiles = ceil_div(k, BI) * ceil_div(m, BJ);
  // all CTAs of the split GEMM in ONE wave: 2 resident CTAs per SM (launch bounds), so at most 2 * 148 of them --
  // rounding the split count up (9 tiles x 33 = 297 CTAs for in_proj) cost a second, nearly empty wave
  int64_t s = int64_t(kNumSMs) * 2 / tiles;
  const int64_t cap = ceil_div(n, 4 * BR);  // at least 64 rows of the reduction per split
  if (s > cap) s = cap;
  if (s < 1) s = 1;
  if (s > 1024) s = 1024;
  return int(s);
}

}  // namespace
}  // namespace lgk

using namespace lgk;

extern "C" int lgk_linear_fwd(const void* x, const float* weight, const float* bias, const float* rowscale, void* out,
                              int64_t n, int64_t k, int64_t m, int act, int accumulate, int dtype, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(n >= 0 && k > 0 && m > 0, LGK_ERR_BAD_ARG, "lgk_linear_fwd: bad sizes");
  LGK_REQUIRE(known_dtype(dtype), LGK_ERR_UNSUPPORTED, "lgk_linear_fwd: dtype must be LGK_F32 or LGK_BF16");
  if (n == 0) return LGK_OK;
  LGK_REQUIRE(x && weight && out, LGK_ERR_BAD_ARG, "lgk_linear_fwd: null pointer");
  if (m <= kNarrowMax) {  // out_proj and other heads with a handful of outputs: a warp per row
    const bool vec = (k % 4 == 0) && aligned16(x) && aligned16(weight);
    const unsigned grid = unsigned(ceil_div(n, 8));
#define LGK_NARROW_FWD(T, V)                                                                                                  \
  linear_fwd_narrow_kernel<T, V><<<grid, 256, 0, stream>>>(static_cast<const T*>(x), weight, bias, rowscale, static_cast<T*>(out), \
                                                           n, int(k), int(m), act, accumulate)
    if (dtype == LGK_BF16) { if (vec) LGK_NARROW_FWD(bf16, true); else LGK_NARROW_FWD(bf16, false); }
    else { if (vec) LGK_NARROW_FWD(float, true); else LGK_NARROW_FWD(float, false); }
#undef LGK_NARROW_FWD
    LGK_CHECK_LAUNCH("linear_fwd_narrow_kernel");
    return LGK_OK;
  }
  if (dtype == LGK_F32 && use_tensor_cores(n, k)) {
    bool done = false;
    if ((k & 3) == 0) {  // aligned rows: tensor-map TMA feeds the tensor core
      if (int rc = launch_linear_tma(static_cast<const float*>(x), weight, bias, rowscale, static_cast<float*>(out), n, k, m,
                                     act, accumulate, stream, &done))
        return rc;
      if (done) return LGK_OK;
    }
    return launch_linear_tc(static_cast<const float*>(x), weight, bias, rowscale, static_cast<float*>(out), n, k, m, act,
                            accumulate, stream);
  }
  GemmArgs g{};
  g.A = x; g.B = weight; g.C = out;
  g.I = n; g.J = m; g.R = k; g.lda = k; g.ldb = k; g.sCi = m; g.sCj = 1;
  g.bias = bias; g.rowscale = rowscale; g.act = act; g.accumulate = accumulate; g.r_chunk = k; g.split_stride = 0;
  const bool vec = (k % 4 == 0) && aligned16(x) && aligned16(weight);
  if (dtype == LGK_BF16) return launch_gemm<true, true, bf16, float, bf16>(g, 1, vec, stream, "gemm_kernel(fwd,bf16)");
  return launch_gemm<true, true, float, float, float>(g, 1, vec, stream, "gemm_kernel(fwd)");
}

extern "C" int lgk_gat_alpha(const void* h, const float* att_src, const float* att_dst, float* a_src, float* a_dst, int64_t N, int H,
                             int C, int dtype, void* stream_);

// 1 when lgk_linear_gat_fwd forms the logit halves in the transform's epilogue (from the fp32 accumulators) for this shape
extern "C" int lgk_linear_gat_fused(int64_t n, int64_t k, int heads, int chan, int dtype) {
  if (n <= 0 || k <= 0 || heads <= 0 || chan <= 0) return 0;
  if (dtype == LGK_F32 && use_tensor_cores(n, k))
    return (k & 3) == 0 && chan >= 16 && chan <= 64 && 64 % chan == 0 && n < (int64_t(1) << 31);
  return chan % 4 == 0 && BJ % chan == 0;
}

extern "C" int lgk_linear_gat_fwd(const void* x, const float* weight, const float* att_src, const float* att_dst, void* out,
                                  float* a_src, float* a_dst, int64_t n, int64_t k, int heads, int chan, int dtype, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int64_t m = int64_t(heads) * chan;
  LGK_REQUIRE(n >= 0 && k > 0 && heads > 0 && chan > 0, LGK_ERR_BAD_ARG, "lgk_linear_gat_fwd: bad sizes");
  LGK_REQUIRE(known_dtype(dtype), LGK_ERR_UNSUPPORTED, "lgk_linear_gat_fwd: dtype must be LGK_F32 or LGK_BF16");
  if (n == 0) return LGK_OK;
  LGK_REQUIRE(x && weight && att_src && att_dst && out && a_src && a_dst, LGK_ERR_BAD_ARG, "lgk_linear_gat_fwd: null pointer");
  GatAlpha alpha{att_src, att_dst, a_src, a_dst, heads, chan};
  const bool att_aligned = aligned16(att_src) && aligned16(att_dst);
  if (dtype == LGK_F32 && use_tensor_cores(n, k) && (k & 3) == 0 && att_aligned) {
    bool done = false;
    if (int rc = launch_linear_tma(static_cast<const float*>(x), weight, nullptr, nullptr, static_cast<float*>(out), n, k, m,
                                   LGK_ACT_NONE, 0, stream, &done, &alpha))
      return rc;
    if (done) return LGK_OK;
  }
  const bool fuse = (chan % 4 == 0) && (BJ % chan == 0) && att_aligned && !(dtype == LGK_F32 && use_tensor_cores(n, k));
  if (!fuse) {  // head widths the epilogues do not cover: the transform, then the logit halves from the stored h
    if (int rc = lgk_linear_fwd(x, weight, nullptr, nullptr, out, n, k, m, LGK_ACT_NONE, 0, dtype, stream_)) return rc;
    return lgk_gat_alpha(out, att_src, att_dst, a_src, a_dst, n, heads, chan, dtype, stream_);
  }
  GemmArgs g{};
  g.A = x; g.B = weight; g.C = out;
  g.I = n; g.J = m; g.R = k; g.lda = k; g.ldb = k; g.sCi = m; g.sCj = 1;
  g.act = LGK_ACT_NONE; g.r_chunk = k; g.split_stride = 0; g.alpha = alpha;
  const bool vec = (k % 4 == 0) && aligned16(x) && aligned16(weight);
  if (dtype == LGK_BF16) return launch_gemm<true, true, bf16, float, bf16>(g, 1, vec, stream, "gemm_kernel(gat fwd,bf16)");
  return launch_gemm<true, true, float, float, float>(g, 1, vec, stream, "gemm_kernel(gat fwd)");
}

extern "C" int lgk_linear_bwd_x(const void* dy, const float* weight, void* dx, int64_t n, int64_t k, int64_t m,
                                int accumulate, int dtype, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(n >= 0 && k > 0 && m > 0, LGK_ERR_BAD_ARG, "lgk_linear_bwd_x: bad sizes");
  LGK_REQUIRE(known_dtype(dtype), LGK_ERR_UNSUPPORTED, "lgk_linear_bwd_x: dtype must be LGK_F32 or LGK_BF16");
  if (n == 0) return LGK_OK;
  LGK_REQUIRE(dy && weight && dx, LGK_ERR_BAD_ARG, "lgk_linear_bwd_x: null pointer");
  if (m <= kNarrowMax) {  // dX of a head with a handful of outputs
    const bool nvec = (k % 4 == 0) && aligned16(weight) && aligned16(dx);
    const int64_t threads = n * (nvec ? k / 4 : k);
    const unsigned grid = unsigned(ceil_div(threads, 256));
#define LGK_NARROW_DX(T, V)                                                                                                    \
  linear_bwd_x_narrow_kernel<T, V><<<grid, 256, 0, stream>>>(static_cast<const T*>(dy), weight, static_cast<T*>(dx), n, int(k), int(m), \
                                                             accumulate)
    if (dtype == LGK_BF16) { if (nvec) LGK_NARROW_DX(bf16, true); else LGK_NARROW_DX(bf16, false); }
    else { if (nvec) LGK_NARROW_DX(float, true); else LGK_NARROW_DX(float, false); }
#undef LGK_NARROW_DX
    LGK_CHECK_LAUNCH("linear_bwd_x_narrow_kernel");
    return LGK_OK;
  }
  if (small_weight(k, m) && (k % 4 == 0) && (m % 4 == 0) && n <= 65536 && aligned16(dy) && aligned16(weight) && aligned16(dx)) {
    if (dtype == LGK_BF16)
      return launch_bwd_x_small(static_cast<const bf16*>(dy), weight, static_cast<bf16*>(dx), n, k, m, accumulate, stream);
    return launch_bwd_x_small(static_cast<const float*>(dy), weight, static_cast<float*>(dx), n, k, m, accumulate, stream);
  }
  GemmArgs g{};
  g.A = dy; g.B = weight; g.C = dx;
  g.I = n; g.J = k; g.R = m; g.lda = m; g.ldb = k; g.sCi = k; g.sCj = 1;
  g.accumulate = accumulate; g.r_chunk = m; g.split_stride = 0;
  const bool vec = (m % 4 == 0) && (k % 4 == 0) && aligned16(dy) && aligned16(weight);
  if (dtype == LGK_BF16) return launch_gemm<true, false, bf16, float, bf16>(g, 1, vec, stream, "gemm_kernel(bwd_x,bf16)");
  return launch_gemm<true, false, float, float, float>(g, 1, vec, stream, "gemm_kernel(bwd_x)");
}

extern "C" int64_t lgk_linear_bwd_w_workspace_bytes(int64_t n, int64_t k, int64_t m) {
  if (n <= 0 || k <= 0 || m <= 0) return 0;
  const int64_t slabs = slab_count(n);
  if (small_weight(k, m)) return align_up(slabs * (k * m + m) * 4, 256) + 256;  // + the cluster kernel's counter
  const int s = std::max(choose_wsplits(n, k, m), grad_w_tc_slabs(n, k, m, nullptr));
  return align_up(int64_t(s) * (k * m + m) * 4, 256) + align_up(slabs * m * 4, 256);
}

extern "C" int lgk_linear_bwd_w(const void* dy, const void* x, float* dw, float* db, int64_t n, int64_t k, int64_t m,
                                int accumulate, void* workspace, int64_t workspace_bytes, int dtype, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(n >= 0 && k > 0 && m > 0, LGK_ERR_BAD_ARG, "lgk_linear_bwd_w: bad sizes");
  LGK_REQUIRE(known_dtype(dtype), LGK_ERR_UNSUPPORTED, "lgk_linear_bwd_w: dtype must be LGK_F32 or LGK_BF16");
  const bool is_bf16 = dtype == LGK_BF16;
  LGK_REQUIRE(dw || db, LGK_ERR_BAD_ARG, "lgk_linear_bwd_w: nothing to compute");
  if (n == 0) {
    if (!accumulate) {
      if (dw) cudaMemsetAsync(dw, 0, 4 * k * m, stream);
      if (db) cudaMemsetAsync(db, 0, 4 * m, stream);
      LGK_CHECK_LAUNCH("lgk_linear_bwd_w(memset)");
    }
    return LGK_OK;
  }
  LGK_REQUIRE(dy && (x || !dw), LGK_ERR_BAD_ARG, "lgk_linear_bwd_w: null pointer");
  const int64_t need = lgk_linear_bwd_w_workspace_bytes(n, k, m);
  LGK_REQUIRE(workspace && workspace_bytes >= need, LGK_ERR_WORKSPACE, "lgk_linear_bwd_w: workspace %lld < %lld bytes",
              (long long)workspace_bytes, (long long)need);
  const int slabs = slab_count(n);
  const int64_t slab_rows = ceil_div(n, slabs);
  const float* dyf = static_cast<const float*>(dy);
  float* partial = static_cast<float*>(workspace);

  const bf16* dyh = static_cast<const bf16*>(dy);
  if (!dw) {  // bias gradient only
    if (is_bf16) colsum_slab_kernel<bf16><<<slabs, 256, 0, stream>>>(dyh, partial, n, int(m), slab_rows);
    else colsum_slab_kernel<float><<<slabs, 256, 0, stream>>>(dyf, partial, n, int(m), slab_rows);
    LGK_CHECK_LAUNCH("colsum_slab_kernel");
    slab_reduce_kernel<<<slab_reduce_grid(m), 256, 0, stream>>>(partial, nullptr, db, 0, m, slabs, accumulate);
    LGK_CHECK_LAUNCH("slab_reduce_kernel(db)");
    return LGK_OK;
  }

  if (small_weight(k, m)) {  // dW and db in one pass over slabs of rows
    const int kb = int(ceil_div(k, 64)), mb = int(ceil_div(m, 64));
    const size_t smem = size_t(kSlabChunk) * (kb + mb) * 64 * sizeof(float);
    const float* xf = static_cast<const float*>(x);
    const bf16* xh = static_cast<const bf16*>(x);
    const bool vec = (k % 4 == 0) && (m % 4 == 0) && aligned16(x) && aligned16(dy);
    const int wdb = db != nullptr;
#define LGK_SLAB(KB, MB)                                                                                                              \
  do {                                                                                                                                \
    if (is_bf16) {                                                                                                                    \
      if (vec) bwd_w_slab_kernel<KB, MB, true, bf16><<<slabs, 256, smem, stream>>>(dyh, xh, partial, n, int(k), int(m), slab_rows, wdb);  \
      else bwd_w_slab_kernel<KB, MB, false, bf16><<<slabs, 256, smem, stream>>>(dyh, xh, partial, n, int(k), int(m), slab_rows, wdb);     \
    } else {                                                                                                                          \
      if (vec) bwd_w_slab_kernel<KB, MB, true, float><<<slabs, 256, smem, stream>>>(dyf, xf, partial, n, int(k), int(m), slab_rows, wdb); \
      else bwd_w_slab_kernel<KB, MB, false, float><<<slabs, 256, smem, stream>>>(dyf, xf, partial, n, int(k), int(m), slab_rows, wdb);    \
    }                                                                                                                                 \
  } while (0)
    // k, m <= 64, fp32, not too tall: one cluster-reduced launch -- OPT-IN (LGK_DW_CLUSTER=1).  Measured on the headline step
    // (B200, 300 steps, A/B/A/B on one box): two-launch form 0.3219-0.3248 ms per step, 21.1 us per call; cluster form
    // 0.3526-0.3527 ms, 22.4 us per call, 49 instead of 53 launches.  The call itself is no faster (three cluster barriers
    // and the DSMEM pass cost what the second launch cost) and the step is 30 us SLOWER: a cluster of 8 CTAs needs 8 free
    // SMs of one GPC at the same moment, so these side-stream kernels no longer slip into the gaps of the main chain.
    static const bool cluster_on = [] { const char* v = getenv("LGK_DW_CLUSTER"); return v && *v == '1'; }();
    if (cluster_on && kb == 1 && mb == 1 && !is_bf16 && n < 65536) {
      const int cslabs = int(align_up(slabs, kDwCluster));
      const int64_t crows = ceil_div(n, cslabs);
      unsigned* counter = reinterpret_cast<unsigned*>(static_cast<char*>(workspace) + align_up(int64_t(slabs) * (k * m + m) * 4, 256));
      if (cslabs > kDwCluster) {
        cudaMemsetAsync(counter, 0, sizeof(unsigned), stream);  // the workspace arrives uninitialised
      }
      if (vec) bwd_w_cluster_kernel<true><<<cslabs, 256, 0, stream>>>(dyf, xf, partial, counter, dw, db, n, int(k), int(m), crows, wdb, accumulate);
      else bwd_w_cluster_kernel<false><<<cslabs, 256, 0, stream>>>(dyf, xf, partial, counter, dw, db, n, int(k), int(m), crows, wdb, accumulate);
      LGK_CHECK_LAUNCH("bwd_w_cluster_kernel");
      return LGK_OK;
    }
    // LGK_SLAB_PIPE=0: the single-buffered kernel for fp32 too (profiling knob)
    static const bool pipe_on = [] { const char* v = getenv("LGK_SLAB_PIPE"); return !(v && *v == '0'); }();
    // measured: 194 -> 128 us per call at n = 409 600 (config 3), but the headline step (n = 11 k, kernels of three
    // streams sharing the SMs) is 2.5 us slower with the 48 KB ring than with the 16 KB single buffer: tall products only
    if (vec && !is_bf16 && pipe_on && n >= 65536) {
      const size_t psm = size_t(kPipeStages) * smem;
#define LGK_SLAB_PIPE(KB, MB)                                                                                             \
  do {                                                                                                                    \
    static bool cfg[64] = {};                                                                                             \
    int dev_ord = 0;                                                                                                      \
    cudaGetDevice(&dev_ord);                                                                                              \
    if (!cfg[dev_ord & 63]) {                                                                                             \
      cudaFuncSetAttribute(bwd_w_slab_pipe_kernel<KB, MB>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kPipeStages) * kSlabChunk * (KB + MB) * 64 * 4); \
      cfg[dev_ord & 63] = true;                                                                                           \
    }                                                                                                                     \
    bwd_w_slab_pipe_kernel<KB, MB><<<slabs, 256, psm, stream>>>(dyf, xf, partial, n, int(k), int(m), slab_rows, wdb);       \
  } while (0)
      if (kb == 1 && mb == 1) LGK_SLAB_PIPE(1, 1);
      else if (kb == 1) LGK_SLAB_PIPE(1, 2);
      else if (mb == 1) LGK_SLAB_PIPE(2, 1);
      else LGK_SLAB_PIPE(2, 2);
#undef LGK_SLAB_PIPE
    } else if (kb == 1 && mb == 1) LGK_SLAB(1, 1);
    else if (kb == 1) LGK_SLAB(1, 2);
    else if (mb == 1) LGK_SLAB(2, 1);
    else LGK_SLAB(2, 2);
#undef LGK_SLAB
    LGK_CHECK_LAUNCH("bwd_w_slab_kernel");
    slab_reduce_kernel<<<slab_reduce_grid(k * m + m), 256, 0, stream>>>(partial, dw, db, k * m, m, slabs, accumulate);
    LGK_CHECK_LAUNCH("slab_reduce_kernel");
    return LGK_OK;
  }

  // wide weights (the 1025-wide in_proj): split GEMM over the rows + deterministic reduce, db from a slab pass
  const int tc_slabs = (use_grad_w_tc() && !is_bf16) ? grad_w_tc_slabs(n, k, m, nullptr) : 0;
  const int s = choose_wsplits(n, k, m);
  float* colpart = reinterpret_cast<float*>(static_cast<char*>(workspace) +
                                            align_up(int64_t(std::max(s, grad_w_tc_slabs(n, k, m, nullptr))) * (k * m + m) * 4, 256));
  GemmArgs g{};
  g.A = x; g.B = dy;
  g.I = k; g.J = m; g.R = n; g.lda = k; g.ldb = m; g.sCi = 1; g.sCj = k;  // C(i=k_idx, j=m_idx) -> dW[m*K + k]
  g.r_chunk = align_up(ceil_div(n, s), BR);
  const bool vec = (k % 4 == 0) && (m % 4 == 0) && aligned16(x) && aligned16(dy);
  if (tc_slabs > 0) {  // tensor cores: partial[slab][m*K + k | m], reduced in slab order; db rides along when it can
    const bool fuse_db = db && grad_w_tc_fuses_db(k);
    if (int rc = launch_grad_w_tc(dyf, static_cast<const float*>(x), partial, n, k, m, fuse_db, stream)) return rc;
    slab_reduce_kernel<<<slab_reduce_grid(k * m + (fuse_db ? m : 0)), 256, 0, stream>>>(partial, dw, fuse_db ? db : nullptr, k * m,
                                                                                       fuse_db ? m : 0, tc_slabs, accumulate);
    LGK_CHECK_LAUNCH("slab_reduce_kernel");
    if (fuse_db) return LGK_OK;
  } else if (s == 1) {
    g.C = dw; g.split_stride = 0; g.accumulate = accumulate;
    if (int rc = is_bf16 ? launch_gemm<false, false, bf16, bf16, float>(g, 1, vec, stream, "gemm_kernel(bwd_w,bf16)")
                         : launch_gemm<false, false, float, float, float>(g, 1, vec, stream, "gemm_kernel(bwd_w)"))
      return rc;
  } else {
    g.C = partial; g.split_stride = k * m;
    if (int rc = is_bf16 ? launch_gemm<false, false, bf16, bf16, float>(g, s, vec, stream, "gemm_kernel(bwd_w,split,bf16)")
                         : launch_gemm<false, false, float, float, float>(g, s, vec, stream, "gemm_kernel(bwd_w,split)"))
      return rc;
    splitk_reduce_kernel<<<unsigned(ceil_div(k * m, 256)), 256, 0, stream>>>(partial, dw, k * m, s, accumulate);
    LGK_CHECK_LAUNCH("splitk_reduce_kernel");
  }
  if (db) {
    if (is_bf16) colsum_slab_kernel<bf16><<<slabs, 256, 0, stream>>>(dyh, colpart, n, int(m), slab_rows);
    else colsum_slab_kernel<float><<<slabs, 256, 0, stream>>>(dyf, colpart, n, int(m), slab_rows);
    LGK_CHECK_LAUNCH("colsum_slab_kernel");
    slab_reduce_kernel<<<slab_reduce_grid(m), 256, 0, stream>>>(colpart, nullptr, db, 0, m, slabs, accumulate);
    LGK_CHECK_LAUNCH("slab_reduce_kernel(db)");
  }
  return LGK_OK;
}

// ---- lgk_colsum_static: the bias-gradient column sum with its scratch inside the library ---------------------------
// Same two kernels as lgk_linear_bwd_w(dw = NULL) -- slab partials, then the fixed-order second stage -- but the
// [slabs, m] partials live in the file-scope device array below instead of a caller-provided workspace, for callers that
// cannot allocate (CUDA-graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The array costs 296 x 1024 x 4 B = 1.16 MiB on every GPU that loads the
// library, until the process exits.  Every call shares it and it keeps the last call's contents: nothing has to be
// cleared, because each call's first kernel writes every partial [slab < slabs][column < m] that its second kernel reads;
// two calls that OVERLAP on one GPU (different streams, concurrent graphs) race on it -- serialise them.
#include <mutex>

constexpr int kStaticMaxSlabs = 2 * lgk::kNumSMs;   // what slab_count() can return at most
constexpr int kStaticMaxM = 1024;                    // LGK_COLSUM_STATIC_MAX_M in include/lgk.h
static __device__ __align__(256) float g_colsum_scratch[kStaticMaxSlabs * kStaticMaxM];

// The array's name is a host-side stand-in in host code, not a device address: the address a kernel may use comes from
// cudaGetSymbolAddress, asked once per device with that device current.
static float* colsum_scratch_of_current_device(int* rc) {
  static std::mutex mu;
  static float* ptr_of[64] = {};
  *rc = LGK_OK;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= 64) {
    set_error("lgk_colsum_static: cudaGetDevice failed or device ordinal %d out of range: %s", dev, cudaGetErrorString(e));
    *rc = LGK_ERR_LAUNCH;
    return nullptr;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr_of[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_colsum_scratch);
    if (e != cudaSuccess || !p) {
      set_error("lgk_colsum_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      *rc = LGK_ERR_LAUNCH;
      return nullptr;
    }
    ptr_of[dev] = static_cast<float*>(p);
  }
  return ptr_of[dev];
}

extern "C" int lgk_colsum_static(const void* dy, float* db, int64_t n, int64_t m, int accumulate, int dtype, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(n >= 0 && m > 0, LGK_ERR_BAD_ARG, "lgk_colsum_static: bad sizes");
  LGK_REQUIRE(m <= kStaticMaxM, LGK_ERR_UNSUPPORTED,
              "lgk_colsum_static: m = %lld exceeds the built-in scratch (LGK_COLSUM_STATIC_MAX_M = %d columns); use "
              "lgk_linear_bwd_w with a caller-provided workspace", (long long)m, kStaticMaxM);
  LGK_REQUIRE(known_dtype(dtype), LGK_ERR_UNSUPPORTED, "lgk_colsum_static: dtype must be LGK_F32 or LGK_BF16");
  LGK_REQUIRE(db, LGK_ERR_BAD_ARG, "lgk_colsum_static: null pointer");
  const int slabs = n > 0 ? slab_count(n) : 0;   // <= kStaticMaxSlabs by construction
  LGK_REQUIRE(slabs <= kStaticMaxSlabs, LGK_ERR_UNSUPPORTED, "lgk_colsum_static: %d slabs exceed the built-in scratch", slabs);
  int rc = LGK_OK;
  float* partial = colsum_scratch_of_current_device(&rc);
  if (rc != LGK_OK) return rc;
  if (n > 0) {
    LGK_REQUIRE(dy, LGK_ERR_BAD_ARG, "lgk_colsum_static: null pointer");
    const int64_t slab_rows = ceil_div(n, slabs);
    if (dtype == LGK_BF16) colsum_slab_kernel<bf16><<<slabs, 256, 0, stream>>>(static_cast<const bf16*>(dy), partial, n, int(m), slab_rows);
    else colsum_slab_kernel<float><<<slabs, 256, 0, stream>>>(static_cast<const float*>(dy), partial, n, int(m), slab_rows);
    LGK_CHECK_LAUNCH("colsum_slab_kernel(static)");
  }
  // slabs = 0 (no rows): the second stage reads nothing and writes the empty sum (0, or db unchanged when accumulating)
  slab_reduce_kernel<<<slab_reduce_grid(m), 256, 0, stream>>>(partial, nullptr, db, 0, m, slabs, accumulate);
  LGK_CHECK_LAUNCH("slab_reduce_kernel(static)");
  return LGK_OK;
}
