// (2)+(4) fused gather + segment-reduce aggregation (SpMM) over a CSR (forward) or CSC (backward).
//
// Replaces index_select + mul + scatter(sum|mean|max) of MessagePassing.propagate (reference gin.py:31,
// drgnet.py:55) and torch_sparse.matmul on the adj_t path.  No tensor cores: 0.25 flop per gathered byte.
//
// What bounds it (profiles/microbench/gather_bw.cu, measured on the B200): an SM pulls a gathered 512-byte row
// into registers in 4.2 cycles when it hits L1 (122 B/clk, the LSU data path = the shared-memory crossbar), 7.4
// cycles from L2 and 21 cycles from HBM.  With d neighbours per row the kernel moves d x F x 4 bytes through that
// path against 2 x F x 4 compulsory HBM bytes, so from d ~ 4 upwards the L1/L2 gather path, not HBM, is the
// ceiling; everything below is arranged to (a) keep the gathers L1 hits and (b) spend no other LSU or issue
// slots per edge.
//
// Mapping: a feature row is split into 128-bit pieces, one per lane; a row needs LPR lanes (F/4 rounded up to a
// power of two, at most 32; wider rows are cut into feature slices = grid.y), a warp reduces 32/LPR rows at
// once.  grid.x = bands of consecutive rows; a CTA walks its band in chunks of R rows.  Per chunk the CTA first
// stages the chunk's row pointers, neighbour ids and (if any) fully resolved edge weights
// ew[emap[p]] * pre[idx[p]] in shared memory with coalesced, CTA-wide loads -- so the dependent
// pointer -> index -> weight -> row chain of a CSR walk costs one latency per chunk instead of one per row -- and
// then every group reduces its rows reading ids/weights as 16-byte shared-memory broadcasts (1/4 wavefront per
// edge) and gathering 8 neighbour rows at a time (8 independent 16-byte loads in flight per lane).
#include "common.cuh"
#include <float.h>
#include <stdlib.h>
#include <algorithm>

namespace lgk {
namespace {

constexpr int kThreads = 256;

struct SpmmArgs {
  const int* ptr;
  const int* idx;
  const int* emap;
  const float* ew;
  const float* pre;
  const float* post;
  const float* selfw;
  const float* bias;
  int* arg_out;
  float self_fill;
  int self_mode;
  int64_t n_rows;
  int feat;
  int slice4;     // float4 columns per feature slice (grid.y slices)
  int64_t band;   // consecutive rows per CTA (grid.x bands)
  int ptr_words;  // shared-memory words reserved for the band's row pointers
};

// ---- vectorised path: feat % 4 == 0 ------------------------------------------------------------------
// A CTA owns a band of consecutive rows and walks it in chunks of R rows through a three-deep pipeline:
//   * the band's row pointers are read once, up front, into shared memory;
//   * while the groups reduce chunk c, the ids of chunk c+2 are already in flight into registers; when the
//   * while the groups reduce chunk c, the ids of chunk c+2 are in flight (cp.async) into the third staging
//     buffer; when the reduction is done they are rewritten in place as pre-multiplied row offsets, with the
//     weights fully resolved beside them;
//   * one barrier per chunk.
// So the ptr -> idx -> weight -> row dependency chain never stalls a reduction.  (Prefetching the neighbour rows
// of chunk c+2 into L2 -- per line through the LSU, per row with cp.async.bulk.prefetch, or the band's own rows in
// bulk -- was measured and does not pay: profiles/README.md, r01 v1.)
constexpr int kIdRegs = 3;       // staged edges per thread and chunk: a staging buffer holds kIdRegs * blockDim.x edges
constexpr int kStageBufs = 3;
constexpr int kMaxBand = 2044;   // rows per CTA (their row pointers: 8 KB of shared memory)

__device__ __forceinline__ void cp_async_4(void* smem_dst, const void* gmem_src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(uint32_t(__cvta_generic_to_shared(smem_dst))), "l"(gmem_src));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

__device__ __forceinline__ float edge_weight(const SpmmArgs& a, int p, int j) {
  float w = 1.f;
  if (a.ew) w = __ldg(a.ew + (a.emap ? __ldg(a.emap + p) : p));
  if (a.pre) w *= __ldg(a.pre + j);
  return w;
}

template <typename T, int LPR, int REDUCE, bool WEIGHTED, bool EPI>
__global__ void __launch_bounds__(512, 2) spmm_vec_kernel(SpmmArgs a, const T* __restrict__ x, T* __restrict__ out) {
  extern __shared__ __align__(16) int smem[];
  const int nthreads = blockDim.x;
  const int groups = nthreads / LPR;
  const int cap = nthreads * kIdRegs;
  int* s_ptr = smem;                                                   // [band + 1]
  unsigned* s_off = reinterpret_cast<unsigned*>(smem + a.ptr_words);   // [kStageBufs][cap + 8]: idx * F4
  float* s_w = reinterpret_cast<float*>(s_off + kStageBufs * (cap + 8));

  const int gl = threadIdx.x & (LPR - 1);
  const int grp = threadIdx.x / LPR;
  const int vec = blockIdx.y * a.slice4 + gl;  // this lane's float4 column
  const unsigned F4 = unsigned(a.feat) >> 2;
  const bool lane_on = gl < a.slice4 && vec < int(F4);
  // rows are addressed in float4 units with 32-bit offsets (n_rows * F4 < 2^32 is checked by the launcher)
  const T* xv = x + vec * 4;
  T* ov = out + vec * 4;
  const int64_t band_start = int64_t(blockIdx.x) * a.band;
  const int nb = int(min(a.n_rows, band_start + a.band) - band_start);
  const int skip_self = (WEIGHTED && a.self_mode == LGK_SELF_REPLACE) ? 1 : 0;

  for (int t = threadIdx.x; t <= nb; t += nthreads) s_ptr[t] = __ldg(a.ptr + band_start + t);
  __syncthreads();
  // rows per chunk: as many per group (1..8) as keep an average chunk within ~3/4 of a staging buffer
  const int band_edges = s_ptr[nb] - s_ptr[0];
  const int R = groups * max(1, min(8, int((int64_t(cap) * 3 / 4 * nb) / (int64_t(groups) * max(band_edges, 1)))));
  const int nchunks = (nb + R - 1) / R;

  // ids of chunk c -> staging buffer c % kStageBufs, asynchronously (cp.async: no registers held meanwhile)
  auto load_ids = [&](int c) {
    const int e0 = s_ptr[c * R], e1 = s_ptr[min(nb, (c + 1) * R)];
    const int base = e0 & ~3;
    if (e1 - base <= cap) {
      unsigned* so = s_off + (c % kStageBufs) * (cap + 8) - base;
      for (int p = e0 + threadIdx.x; p < e1; p += nthreads) cp_async_4(so + p, a.idx + p);
    }
    cp_async_commit();
  };
  // in place: raw ids -> pre-multiplied row offsets, weights resolved.
  // Every thread converts exactly the entries it copied itself.
  auto store_ids = [&](int c) {
    cp_async_wait_all();
    const int e0 = s_ptr[c * R], e1 = s_ptr[min(nb, (c + 1) * R)];
    const int base = e0 & ~3;
    if (e1 - base > cap) return;
    unsigned* so = s_off + (c % kStageBufs) * (cap + 8) - base;
    float* sw = s_w + (c % kStageBufs) * (cap + 8) - base;
    for (int p = e0 + threadIdx.x; p < e1; p += nthreads) {
      const int j = int(so[p]);
      so[p] = unsigned(j) * F4;
      if (WEIGHTED) {
        float w = edge_weight(a, p, j);
        if (skip_self) {  // the row of slot p: largest r with s_ptr[r] <= p (rows of this chunk)
          int lo = c * R, hi = min(nb, (c + 1) * R) - 1;
          while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (s_ptr[mid] <= p) lo = mid; else hi = mid - 1;
          }
          if (band_start + lo == j) w = 0.f;
        }
        sw[p] = w;
      }
    }
  };
  load_ids(0);
  store_ids(0);
  if (nchunks > 1) {
    load_ids(1);
    store_ids(1);
  }
  __syncthreads();

  for (int c = 0; c < nchunks; ++c) {
    if (c + 2 < nchunks) load_ids(c + 2);
    const int r_lo = c * R;
    const int nrows = min(R, nb - r_lo);
    const int e1 = s_ptr[r_lo + nrows];
    const int base = s_ptr[r_lo] & ~3;
    const bool staged = (e1 - base) <= cap;
    const unsigned* so = s_off + (c % kStageBufs) * (cap + 8);
    const float* sw = s_w + (c % kStageBufs) * (cap + 8);

    if (lane_on) {
      for (int lr = grp; lr < nrows; lr += groups) {
        const int64_t row = band_start + r_lo + lr;
        const int start = s_ptr[r_lo + lr], end = s_ptr[r_lo + lr + 1];
        const unsigned self_off = skip_self ? unsigned(row) * F4 : 0xffffffffu;
        float4 acc = (REDUCE == LGK_MAX) ? make_float4(-FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX) : f4_zero();
        int4 arg = make_int4(-1, -1, -1, -1);

        float4 acc2 = f4_zero();  // second accumulator chain of the sum
        // a row = head (up to the first 16-byte aligned slot) + blocks of eight aligned slots + tail
        const int q0 = min(end, (start + 3) & ~3);
        const int qe = staged ? q0 + ((end - q0) & ~7) : q0;
        // ---- main loop: eight staged entries per step, no predicates; a stored self loop that is to be dropped
        // already carries weight 0
#pragma unroll 1
        for (int q = q0; q < qe; q += 8) {
          const uint4 ja = *reinterpret_cast<const uint4*>(so + (q - base));
          const uint4 jb = *reinterpret_cast<const uint4*>(so + (q - base) + 4);
          const unsigned off[8] = {ja.x, ja.y, ja.z, ja.w, jb.x, jb.y, jb.z, jb.w};
          float4 v[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) v[k] = Vec4<T>::gather(xv + size_t(off[k]) * 4, 0);
          if (REDUCE == LGK_MAX) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              const int slot = q + k;
              if (v[k].x > acc.x) { acc.x = v[k].x; arg.x = slot; }
              if (v[k].y > acc.y) { acc.y = v[k].y; arg.y = slot; }
              if (v[k].z > acc.z) { acc.z = v[k].z; arg.z = slot; }
              if (v[k].w > acc.w) { acc.w = v[k].w; arg.w = slot; }
            }
          } else if (WEIGHTED) {
            const float4 wa = *reinterpret_cast<const float4*>(sw + (q - base));
            const float4 wb = *reinterpret_cast<const float4*>(sw + (q - base) + 4);
            const float w[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
#pragma unroll
            for (int k = 0; k < 8; k += 2) {
              f4_fma(acc, w[k], v[k]);
              f4_fma(acc2, w[k + 1], v[k + 1]);
            }
          } else {
#pragma unroll
            for (int k = 0; k < 8; k += 2) {
              acc.x += v[k].x; acc.y += v[k].y; acc.z += v[k].z; acc.w += v[k].w;
              acc2.x += v[k + 1].x; acc2.y += v[k + 1].y; acc2.z += v[k + 1].z; acc2.w += v[k + 1].w;
            }
          }
        }
        // ---- head and tail together (and every slot of a chunk that was too large to stage): one list of
        // nh + nt slots taken up to eight per step, predicated -- a step costs a full gather latency, so a head of
        // two and a tail of six are one step, not two
        const int nh = q0 - start, cnt = nh + (end - qe);
#pragma unroll 1
        for (int b0 = 0; b0 < cnt; b0 += 8) {
          unsigned off[8];
          float w[8];
          int slot[8];
          bool ok[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int t = b0 + k;
            ok[k] = t < cnt;
            slot[k] = t < nh ? start + t : qe + (t - nh);
          }
          if (staged) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              off[k] = ok[k] ? so[slot[k] - base] : 0u;
              w[k] = ok[k] ? (WEIGHTED ? sw[slot[k] - base] : 1.f) : 0.f;
            }
          } else {  // read through L1
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              const int j = ok[k] ? __ldg(a.idx + slot[k]) : 0;
              off[k] = unsigned(j) * F4;
              w[k] = ok[k] ? 1.f : 0.f;
              if (WEIGHTED && ok[k]) w[k] = (off[k] == self_off) ? 0.f : edge_weight(a, slot[k], j);
            }
          }
          float4 v[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            v[k] = f4_zero();
            if (ok[k]) v[k] = Vec4<T>::gather(xv + size_t(off[k]) * 4, 0);
          }
          if (REDUCE == LGK_MAX) {
            // the main loop ran first: on ties the LOWER slot must still win (first arg-max, torch-scatter's rule)
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              if (ok[k]) {
                const int sl = slot[k];
                if (v[k].x > acc.x || (v[k].x == acc.x && sl < arg.x)) { acc.x = v[k].x; arg.x = sl; }
                if (v[k].y > acc.y || (v[k].y == acc.y && sl < arg.y)) { acc.y = v[k].y; arg.y = sl; }
                if (v[k].z > acc.z || (v[k].z == acc.z && sl < arg.z)) { acc.z = v[k].z; arg.z = sl; }
                if (v[k].w > acc.w || (v[k].w == acc.w && sl < arg.w)) { acc.w = v[k].w; arg.w = sl; }
              }
            }
          } else {
#pragma unroll
            for (int k = 0; k < 8; k += 2) {
              f4_fma(acc, w[k], v[k]);
              f4_fma(acc2, w[k + 1], v[k + 1]);
            }
          }
        }
        if (REDUCE != LGK_MAX) { acc.x += acc2.x; acc.y += acc2.y; acc.z += acc2.z; acc.w += acc2.w; }

        float4 r = acc;
        const size_t ro = size_t(unsigned(row) * F4) * 4;
        if (REDUCE == LGK_MAX) {
          if (arg.x < 0) r.x = 0.f;
          if (arg.y < 0) r.y = 0.f;
          if (arg.z < 0) r.z = 0.f;
          if (arg.w < 0) r.w = 0.f;
          if (a.arg_out) st_stream_i4(reinterpret_cast<int4*>(a.arg_out + ro + vec * 4), arg);
        } else if (EPI) {  // any of: mean, post-scale, self term (bias below)
          float scale = 1.f;
          if (REDUCE == LGK_MEAN) scale = 1.f / float(max(end - start, 1));
          if (a.post) scale *= __ldg(a.post + row);
          if (a.self_mode != LGK_SELF_NONE) {
            float wself = a.selfw ? __ldg(a.selfw + row) : a.self_fill;
            if (a.pre) wself *= __ldg(a.pre + row);
            f4_fma(r, wself, Vec4<T>::gather(xv + ro, 0));
          }
          r.x *= scale; r.y *= scale; r.z *= scale; r.w *= scale;
        }
        if (EPI && a.bias) {
          float4 b = __ldg(reinterpret_cast<const float4*>(a.bias) + vec);
          r.x += b.x; r.y += b.y; r.z += b.z; r.w += b.w;
        }
        Vec4<T>::store(ov + ro, 0, r);
      }
    }
    if (c + 2 < nchunks) store_ids(c + 2);
    __syncthreads();
  }
}

// ---- scalar path: any feat (1025-wide raw lesion features, DRGNet's 1-wide last layer, ...) ----------
// warp per row, lanes stride over the feature dimension in chunks of 32 * UNR
template <typename T, int REDUCE>
__global__ void __launch_bounds__(kThreads) spmm_scalar_kernel(SpmmArgs a, const T* __restrict__ x, T* __restrict__ out) {
  constexpr int UNR = 4;
  const int lane = threadIdx.x & 31;
  const int64_t row = int64_t(blockIdx.x) * (kThreads / 32) + (threadIdx.x >> 5);
  if (row >= a.n_rows) return;
  const int F = a.feat;
  const int start = a.ptr[row], end = a.ptr[row + 1];
  float scale = 1.f;
  if (REDUCE == LGK_MEAN) scale = 1.f / float(max(end - start, 1));
  if (a.post) scale *= __ldg(a.post + row);
  float wself = 0.f;
  if (REDUCE != LGK_MAX && a.self_mode != LGK_SELF_NONE) {
    wself = a.selfw ? __ldg(a.selfw + row) : a.self_fill;
    if (a.pre) wself *= __ldg(a.pre + row);
  }
  for (int f0 = 0; f0 < F; f0 += 32 * UNR) {
    float acc[UNR];
    int arg[UNR];
#pragma unroll
    for (int u = 0; u < UNR; ++u) { acc[u] = (REDUCE == LGK_MAX) ? -FLT_MAX : 0.f; arg[u] = -1; }
    for (int base = start; base < end; base += 32) {
      const int p = base + lane;
      int my_j = int(row);
      float my_w = 0.f;
      if (p < end) {
        int j = ld_stream_i32(a.idx + p);
        float w = 1.f;
        if (a.ew) w = __ldg(a.ew + (a.emap ? ld_stream_i32(a.emap + p) : p));
        if (a.pre) w *= __ldg(a.pre + j);
        if (a.self_mode == LGK_SELF_REPLACE && j == int(row)) w = 0.f;
        my_j = j;
        my_w = w;
      }
      const int cnt = min(32, end - base);
      for (int t = 0; t < cnt; ++t) {
        const int j = __shfl_sync(0xffffffffu, my_j, t);
        const float w = __shfl_sync(0xffffffffu, my_w, t);
        const T* xr = x + int64_t(j) * F;
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
          const int f = f0 + u * 32 + lane;
          if (f < F) {
            float val = Vec4<T>::load1(xr, f);
            if (REDUCE == LGK_MAX) {
              if (val > acc[u]) { acc[u] = val; arg[u] = base + t; }
            } else {
              acc[u] = fmaf(w, val, acc[u]);
            }
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      const int f = f0 + u * 32 + lane;
      if (f < F) {
        float r = acc[u];
        if (REDUCE == LGK_MAX) {
          if (arg[u] < 0) r = 0.f;
          if (a.arg_out) a.arg_out[row * F + f] = arg[u];
        } else {
          if (a.self_mode != LGK_SELF_NONE) r = fmaf(wself, Vec4<T>::load1(x + row * F, f), r);
          r *= scale;
        }
        if (a.bias) r += __ldg(a.bias + f);
        Vec4<T>::store1(out + row * F, f, r);
      }
    }
  }
}

// ---- backward of max: CSC gather of dy where the forward's winning slot matches --------------------------
template <typename T>
__global__ void __launch_bounds__(kThreads) spmm_max_bwd_kernel(const int* __restrict__ colptr, const int* __restrict__ rowidx,
                                                                const int* __restrict__ csr2csc, const int* __restrict__ arg,
                                                                const T* __restrict__ dy, T* __restrict__ dx, int64_t n_rows, int F) {
  const int lane = threadIdx.x & 31;
  const int64_t j = int64_t(blockIdx.x) * (kThreads / 32) + (threadIdx.x >> 5);
  if (j >= n_rows) return;
  const int start = colptr[j], end = colptr[j + 1];
  for (int f = lane; f < F; f += 32) {
    float acc = 0.f;
    for (int q = start; q < end; ++q) {
      const int i = __ldg(rowidx + q), slot = __ldg(csr2csc + q);
      if (__ldg(arg + int64_t(i) * F + f) == slot) acc += Vec4<T>::load1(dy + int64_t(i) * F, f);
    }
    Vec4<T>::store1(dx + j * F, f, acc);
  }
}

__global__ void __launch_bounds__(256) gcn_norm_kernel(const int* __restrict__ rowptr, const int* __restrict__ col,
                                                       const float* __restrict__ ew, float fill, int add_self_loops,
                                                       float* __restrict__ dinv, float* __restrict__ selfw_out, int64_t N) {
  int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const int start = rowptr[i], end = rowptr[i + 1];
  float deg = 0.f, sw = fill;
  for (int p = start; p < end; ++p) {
    const int j = __ldg(col + p);
    const float w = ew ? __ldg(ew + p) : 1.f;
    if (add_self_loops && j == int(i)) {
      // stored loops sit next to each other in ascending original edge id: the last one wins
      if (ew) sw = w;
    } else {
      deg += w;
    }
  }
  if (add_self_loops) deg += sw;
  dinv[i] = deg == 0.f ? 0.f : 1.0f / sqrtf(deg);
  if (selfw_out) selfw_out[i] = sw;
}

// Launch geometry of the vectorised kernel.  Defaults were chosen on the B200 (profiles/README.md); the
// environment variables exist for that tuning sweep only and are read once.
struct SpmmTuning {
  int threads, max_slice4, waves;
  int64_t max_band;
  int carveout;  // preferred shared-memory carve-out in percent (-1: driver default)
};
int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return (v && *v) ? atoi(v) : dflt;
}
const SpmmTuning& tuning() {
  static const SpmmTuning t = {env_int("LGK_SPMM_THREADS", 512), env_int("LGK_SPMM_SLICE4", 32), env_int("LGK_SPMM_WAVES", 16),
                               int64_t(env_int("LGK_SPMM_MAX_BAND", kMaxBand)), env_int("LGK_SPMM_CARVEOUT", -1)};
  return t;
}

// The gathers live on L1 hits: ask for the smallest shared-memory carve-out that still holds the resident CTAs'
// staging buffers, once per instantiation (a function attribute, not a device-wide setting).
template <typename T, int LPR, int REDUCE, bool WEIGHTED, bool EPI>
void launch_vec(dim3 grid, int threads, size_t smem, cudaStream_t stream, const SpmmArgs& a, const T* x, T* out) {
  static const bool once = [] {
    const int pct = tuning().carveout;
    if (pct >= 0) cudaFuncSetAttribute(spmm_vec_kernel<T, LPR, REDUCE, WEIGHTED, EPI>, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
    return true;
  }();
  (void)once;
  spmm_vec_kernel<T, LPR, REDUCE, WEIGHTED, EPI><<<grid, threads, smem, stream>>>(a, x, out);
}

template <typename T, int REDUCE>
int launch_spmm(const SpmmArgs& a_in, const void* x_, void* out_, cudaStream_t stream) {
  SpmmArgs a = a_in;
  const T* x = static_cast<const T*>(x_);
  T* out = static_cast<T*>(out_);
  const int F = a.feat;
  const bool aligned = ((reinterpret_cast<uintptr_t>(x_) | reinterpret_cast<uintptr_t>(out_) | reinterpret_cast<uintptr_t>(a.idx) |
                         reinterpret_cast<uintptr_t>(a.bias) | reinterpret_cast<uintptr_t>(a.arg_out)) & 15) == 0;
  if (F % 4 == 0 && aligned && a.n_rows * (F / 4) < (int64_t(1) << 32)) {
    const SpmmTuning& t = tuning();
    const int F4 = F / 4;
    const int nslice = int(ceil_div(F4, min(32, max(1, t.max_slice4))));
    a.slice4 = int(ceil_div(F4, nslice));
    int lpr = 1;
    while (lpr < a.slice4) lpr <<= 1;
    // at most 64 groups per CTA, fewer threads on small inputs
    int threads = min(max(32, min(512, t.threads)) / 32 * 32, max(64, 64 * lpr));
    while (threads > 128 && ceil_div(a.n_rows, int64_t(threads / lpr) * 2) < 2 * kNumSMs) threads >>= 1;
    // bands: enough CTAs for `waves` rounds over the SMs
    int64_t band = ceil_div(a.n_rows, int64_t(kNumSMs) * t.waves);
    band = std::max<int64_t>(threads / lpr, std::min<int64_t>(band, std::min<int64_t>(t.max_band, kMaxBand)));
    a.band = band;
    a.ptr_words = int((band + 1 + 3) & ~int64_t(3));
    dim3 grid(unsigned(ceil_div(a.n_rows, band)), unsigned(nslice));
    const bool weighted = REDUCE != LGK_MAX && (a.ew || a.pre || a.self_mode == LGK_SELF_REPLACE);
    const bool epi = REDUCE == LGK_MEAN || a.post || a.bias || a.self_mode != LGK_SELF_NONE;
    const size_t smem = 4 * (size_t(a.ptr_words) + size_t(kStageBufs) * (threads * kIdRegs + 8) * (weighted ? 2 : 1));
#define LGK_SPMM_CASE(LPR)                                                                             \
  case LPR:                                                                                            \
    if (weighted) launch_vec<T, LPR, REDUCE, true, true>(grid, threads, smem, stream, a, x, out);        \
    else if (epi) launch_vec<T, LPR, REDUCE, false, true>(grid, threads, smem, stream, a, x, out);       \
    else launch_vec<T, LPR, REDUCE, false, false>(grid, threads, smem, stream, a, x, out);               \
    break;
    switch (lpr) {
      LGK_SPMM_CASE(1) LGK_SPMM_CASE(2) LGK_SPMM_CASE(4) LGK_SPMM_CASE(8) LGK_SPMM_CASE(16) LGK_SPMM_CASE(32)
    }
#undef LGK_SPMM_CASE
  } else {
    spmm_scalar_kernel<T, REDUCE><<<unsigned(ceil_div(a.n_rows, kThreads / 32)), kThreads, 0, stream>>>(a, x, out);
  }
  LGK_CHECK_LAUNCH("spmm_kernel");
  return LGK_OK;
}

template <typename T>
int dispatch_reduce(const SpmmArgs& a, int reduce, const void* x, void* out, cudaStream_t stream) {
  switch (reduce) {
    case LGK_SUM: return launch_spmm<T, LGK_SUM>(a, x, out, stream);
    case LGK_MEAN: return launch_spmm<T, LGK_MEAN>(a, x, out, stream);
    case LGK_MAX: return launch_spmm<T, LGK_MAX>(a, x, out, stream);
  }
  set_error("lgk_spmm: unknown reduce %d", reduce);
  return LGK_ERR_BAD_ARG;
}

}  // namespace
}  // namespace lgk

using namespace lgk;

extern "C" int lgk_spmm(const int32_t* ptr, const int32_t* idx, const int32_t* emap, const float* ew, const float* pre,
                        const float* post, const float* selfw, float self_fill, int self_mode, const void* x,
                        const float* bias, void* out, int32_t* arg_out, int64_t n_rows, int64_t feat, int reduce,
                        int dtype, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(n_rows >= 0 && feat >= 0, LGK_ERR_BAD_ARG, "lgk_spmm: negative size");
  if (n_rows == 0 || feat == 0) return LGK_OK;
  LGK_REQUIRE(ptr && x && out, LGK_ERR_BAD_ARG, "lgk_spmm: null pointer");
  LGK_REQUIRE(feat < (1 << 30) && n_rows < (int64_t(1) << 31), LGK_ERR_UNSUPPORTED, "lgk_spmm: size out of range");
  LGK_REQUIRE(self_mode >= LGK_SELF_NONE && self_mode <= LGK_SELF_ADD, LGK_ERR_BAD_ARG, "lgk_spmm: bad self_mode");
  LGK_REQUIRE(reduce == LGK_SUM || self_mode == LGK_SELF_NONE, LGK_ERR_BAD_ARG,
              "lgk_spmm: self-loop handling is defined for the sum reduction only");
  LGK_REQUIRE(reduce != LGK_MAX || (!ew && !pre && !post), LGK_ERR_BAD_ARG, "lgk_spmm: max takes no weights");
  LGK_REQUIRE(dtype == LGK_F32 || dtype == LGK_BF16, LGK_ERR_BAD_ARG, "lgk_spmm: bad dtype %d", dtype);
  LGK_REQUIRE(x != out, LGK_ERR_BAD_ARG, "lgk_spmm: in-place aggregation is not possible");
  SpmmArgs a{ptr, idx, emap, ew, pre, post, selfw, bias, arg_out, self_fill, self_mode, n_rows, int(feat), 0, 0};
  return dtype == LGK_F32 ? dispatch_reduce<float>(a, reduce, x, out, stream)
                          : dispatch_reduce<__nv_bfloat16>(a, reduce, x, out, stream);
}

extern "C" int lgk_spmm_max_bwd(const int32_t* colptr, const int32_t* row, const int32_t* csr2csc, const int32_t* arg,
                                const void* dy, void* dx, int64_t n_rows, int64_t feat, int dtype, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(n_rows >= 0 && feat >= 0, LGK_ERR_BAD_ARG, "lgk_spmm_max_bwd: negative size");
  if (n_rows == 0 || feat == 0) return LGK_OK;
  LGK_REQUIRE(colptr && arg && dy && dx, LGK_ERR_BAD_ARG, "lgk_spmm_max_bwd: null pointer");
  unsigned grid = unsigned(ceil_div(n_rows, kThreads / 32));
  if (dtype == LGK_F32)
    spmm_max_bwd_kernel<float><<<grid, kThreads, 0, stream>>>(colptr, row, csr2csc, arg, static_cast<const float*>(dy),
                                                              static_cast<float*>(dx), n_rows, int(feat));
  else if (dtype == LGK_BF16)
    spmm_max_bwd_kernel<__nv_bfloat16><<<grid, kThreads, 0, stream>>>(colptr, row, csr2csc, arg,
                                                                      static_cast<const __nv_bfloat16*>(dy),
                                                                      static_cast<__nv_bfloat16*>(dx), n_rows, int(feat));
  else
    LGK_REQUIRE(false, LGK_ERR_BAD_ARG, "lgk_spmm_max_bwd: bad dtype %d", dtype);
  LGK_CHECK_LAUNCH("spmm_max_bwd_kernel");
  return LGK_OK;
}

extern "C" int lgk_gcn_norm(const int32_t* rowptr, const int32_t* col, const float* ew, float fill, int add_self_loops,
                            float* dinv, float* selfw_out, int64_t num_nodes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(num_nodes >= 0, LGK_ERR_BAD_ARG, "lgk_gcn_norm: negative size");
  if (num_nodes == 0) return LGK_OK;
  LGK_REQUIRE(rowptr && dinv, LGK_ERR_BAD_ARG, "lgk_gcn_norm: null pointer");
  gcn_norm_kernel<<<unsigned(ceil_div(num_nodes, 256)), 256, 0, stream>>>(rowptr, col, ew, fill, add_self_loops, dinv,
                                                                          selfw_out, num_nodes);
  LGK_CHECK_LAUNCH("gcn_norm_kernel");
  return LGK_OK;
}
