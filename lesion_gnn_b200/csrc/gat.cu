// GAT edge-softmax attention, forward and backward (torch_geometric.nn.GATConv; reference models/gat.py:31,51).
//
// PyG runs ~15 ATen kernels per layer and materialises [E,H] logits three times and the [E,H,C] messages once.
// Here one kernel does leaky_relu + per-destination softmax + weighted aggregation in a single pass over the
// CSR with an online (running max / running sum) softmax, and the backward is two atomic-free passes: one over
// the CSR (destination side) and one over the CSC (source side) that recompute the attention coefficients
// from 4 per-(node, head) statistics instead of storing anything per edge.
//
// Mapping: one group of LPR lanes per (node, head); a lane owns 4 channels (VPL vectors if C > 128).  Adjacent
// groups of a warp are adjacent heads of the same node, so a warp still reads whole contiguous feature rows.
#include "common.cuh"
#include <math.h>
#include <stdlib.h>

#ifndef LGK_GAT_EDGE_BATCH
#define LGK_GAT_EDGE_BATCH 4
#endif

namespace lgk {
namespace {

constexpr int kThreads = 256;

struct GatArgs {
  const int* ptr;      // rowptr (dst side) or colptr (src side)
  const int* idx;      // col or row
  const int* slotmap;  // csr2csc on the src side (dropout key), null on the dst side
  const float* a_src;
  const float* a_dst;
  const float* bias;
  float* rowmax;
  float* rowsum;
  float4* stats;       // [N,H] {a_dst, rowmax, 1/(rowsum+1e-16), delta}
  const float* da_dst;
  float* da_out;       // da_dst (dst pass) or da_src (src pass)
  const float* att_src;
  const float* att_dst;
  void* act_out;       // forward only: second output elu(out), same type and shape as out (null: not wanted)
  int64_t n_nodes;
  int heads;
  int chan;
  float slope;
  int add_self;
  float drop_p;
  float drop_scale;
  uint64_t seed;
  const int64_t* seed_dev;  // optional device-side step counter mixed into the seed (CUDA-graph replays draw new masks)
};

template <int LPR>
__device__ __forceinline__ unsigned gmask() {
  if constexpr (LPR == 32) {
    return 0xffffffffu;
  } else {
    unsigned lane = threadIdx.x & 31;
    return ((1u << LPR) - 1u) << (lane & ~(LPR - 1));
  }
}
template <int LPR>
__device__ __forceinline__ float group_sum(float v, unsigned mask) {
#pragma unroll
  for (int o = LPR >> 1; o >= 1; o >>= 1) v += __shfl_xor_sync(mask, v, o, LPR);
  return v;
}
template <int LPR>
__device__ __forceinline__ float group_max(float v, unsigned mask) {
#pragma unroll
  for (int o = LPR >> 1; o >= 1; o >>= 1) v = fmaxf(v, __shfl_xor_sync(mask, v, o, LPR));
  return v;
}
__device__ __forceinline__ float lrelu(float z, float slope) { return z > 0.f ? z : slope * z; }
// seed of this launch: the host seed, advanced by the device-side counter when there is one -- a captured CUDA graph
// bakes the host value into its kernel nodes, the counter (the optimiser's step count) is what changes between replays;
// forward and both backward passes of a step read the same value
__device__ __forceinline__ uint64_t launch_seed(const GatArgs& a) {
  if (a.drop_p <= 0.f || !a.seed_dev) return a.seed;
  return a.seed + 0xD1B54A32D192ED03ull * uint64_t(__ldg(a.seed_dev));
}
__device__ __forceinline__ float keep_scale(const GatArgs& a, uint64_t seed, uint32_t slot, int head) {
  // attention dropout: 0 (dropped) or 1/(1-p) (kept); identical in forward and both backward passes
  if (a.drop_p <= 0.f) return 1.f;
  return uniform_hash(seed, slot, uint32_t(head)) < a.drop_p ? 0.f : a.drop_scale;
}
constexpr uint32_t kSelfSlotBit = 0x80000000u;
// backward passes: edges of a (node, head) group whose gathers are issued together; wide rows (more than one 16-byte vector
// per lane) keep the one-edge-at-a-time loop.  Measured on the headline batch (45 k groups of 4 lanes, ~7 edges each,
// profiles/r02/gat_edge_batch_ab.log, two interleaved runs per setting): the two kernels alone do not change (21.3 us per
// layer in-graph at 1, 2 or 4), the captured training step does: 0.3257-0.3267 ms at 1, 0.3277-0.3280 at 2, 0.3203-0.3204 at 4
// (62 / 78 registers instead of 40 / 46: fewer resident warps of these kernels leave room for the parameter-gradient
// kernels of the side streams that run beside them).
constexpr int kThinIds = 8, kThinRows = 2;  // thread-per-(node, head) mapping of the backward passes
template <int LPR, int VPL>
constexpr int kEdgeBatch = VPL > 1 ? 1 : (LPR < LGK_GAT_EDGE_BATCH ? LPR : LGK_GAT_EDGE_BATCH);

// ---- alpha = <h, att> ------------------------------------------------------------------------------------
template <typename T, int LPR, int VPL>
__global__ void __launch_bounds__(kThreads) gat_alpha_kernel(const T* __restrict__ h, const float* __restrict__ att_src,
                                                             const float* __restrict__ att_dst, float* __restrict__ a_src,
                                                             float* __restrict__ a_dst, int64_t n_items, int H, int C) {
  constexpr int GPW = 32 / LPR;
  const int gl = threadIdx.x & (LPR - 1);
  const int64_t item = (int64_t(blockIdx.x) * (kThreads / 32) + (threadIdx.x >> 5)) * GPW + ((threadIdx.x & 31) / LPR);
  if (item >= n_items) return;
  const unsigned mask = gmask<LPR>();
  const int head = int(item % H);
  const int C4 = C >> 2;
  float ps = 0.f, pd = 0.f;
#pragma unroll
  for (int v = 0; v < VPL; ++v) {
    const int vec = gl + v * LPR;
    if (vec < C4) {
      float4 hv = Vec4<T>::stream(h + item * C, vec * 4);
      float4 as = __ldg(reinterpret_cast<const float4*>(att_src + head * C) + vec);
      float4 ad = __ldg(reinterpret_cast<const float4*>(att_dst + head * C) + vec);
      ps += f4_dot(hv, as);
      pd += f4_dot(hv, ad);
    }
  }
  ps = group_sum<LPR>(ps, mask);
  pd = group_sum<LPR>(pd, mask);
  if (gl == 0) { a_src[item] = ps; a_dst[item] = pd; }
}

// ---- forward -----------------------------------------------------------------------------------------------
template <typename T, int LPR, int VPL>
__global__ void __launch_bounds__(kThreads) gat_fwd_kernel(GatArgs a, const T* __restrict__ h, T* __restrict__ out) {
  constexpr int GPW = 32 / LPR;
  const int gl = threadIdx.x & (LPR - 1);
  const int H = a.heads, C = a.chan, C4 = C >> 2;
  const int64_t item = (int64_t(blockIdx.x) * (kThreads / 32) + (threadIdx.x >> 5)) * GPW + ((threadIdx.x & 31) / LPR);
  if (item >= a.n_nodes * H) return;
  const unsigned mask = gmask<LPR>();
  const uint64_t seed = launch_seed(a);
  const int i = int(item / H), head = int(item - int64_t(i) * H);
  const int64_t F = int64_t(H) * C;
  const int start = a.ptr[i], end = a.ptr[i + 1];
  const float ad = __ldg(a.a_dst + item);

  float m = -INFINITY, s = 0.f;
  float4 acc[VPL];
#pragma unroll
  for (int v = 0; v < VPL; ++v) acc[v] = f4_zero();

  if (a.add_self) {  // GATConv: remove_self_loops + add_self_loops => exactly one (i,i) entry
    m = lrelu(__ldg(a.a_src + item) + ad, a.slope);
    s = 1.f;
    const float ks = keep_scale(a, seed, kSelfSlotBit | uint32_t(i), head);
#pragma unroll
    for (int v = 0; v < VPL; ++v) {
      const int vec = gl + v * LPR;
      if (vec < C4) f4_fma(acc[v], ks, Vec4<T>::gather(h + int64_t(i) * F + head * C, vec * 4));
    }
  }

  for (int base = start; base < end; base += LPR) {
    const int p = base + gl;
    int my_j = i;
    float my_e = -INFINITY;
    if (p < end) {
      const int j = ld_stream_i32(a.idx + p);
      if (!(a.add_self && j == i)) {
        my_j = j;
        my_e = lrelu(__ldg(a.a_src + int64_t(j) * H + head) + ad, a.slope);
      }
    }
    const float cm = group_max<LPR>(my_e, mask);
    if (cm == -INFINITY) continue;  // every entry of the chunk was a dropped stored loop
    const float m_new = fmaxf(m, cm);
    const float rescale = expf(m - m_new);  // m = -inf on the first chunk -> 0
    float my_p = expf(my_e - m_new);        // my_e = -inf -> 0
    s = s * rescale + group_sum<LPR>(my_p, mask);
    if (a.drop_p > 0.f && p < end) my_p *= keep_scale(a, seed, uint32_t(p), head);
#pragma unroll
    for (int v = 0; v < VPL; ++v) { acc[v].x *= rescale; acc[v].y *= rescale; acc[v].z *= rescale; acc[v].w *= rescale; }
    m = m_new;
    const int cnt = min(LPR, end - base);
#pragma unroll 4
    for (int u = 0; u < cnt; ++u) {
      const int j = __shfl_sync(mask, my_j, u, LPR);
      const float pu = __shfl_sync(mask, my_p, u, LPR);
      const T* hr = h + int64_t(j) * F + head * C;
#pragma unroll
      for (int v = 0; v < VPL; ++v) {
        const int vec = gl + v * LPR;
        if (vec < C4) f4_fma(acc[v], pu, Vec4<T>::gather(hr, vec * 4));
      }
    }
  }

  const float inv = 1.f / (s + 1e-16f);
#pragma unroll
  for (int v = 0; v < VPL; ++v) {
    const int vec = gl + v * LPR;
    if (vec < C4) {
      float4 r = acc[v];
      r.x *= inv; r.y *= inv; r.z *= inv; r.w *= inv;
      if (a.bias) {
        float4 b = __ldg(reinterpret_cast<const float4*>(a.bias + head * C) + vec);
        r.x += b.x; r.y += b.y; r.z += b.z; r.w += b.w;
      }
      Vec4<T>::store(out + int64_t(i) * F + head * C, vec * 4, r);
      if (a.act_out) {  // the ELU that follows every GAT layer of the reference (models/gat.py:51), same launch
        r.x = elu1(r.x); r.y = elu1(r.y); r.z = elu1(r.z); r.w = elu1(r.w);
        Vec4<T>::store(static_cast<T*>(a.act_out) + int64_t(i) * F + head * C, vec * 4, r);
      }
    }
  }
  if (gl == 0) {
    if (a.rowmax) a.rowmax[item] = m;
    if (a.rowsum) a.rowsum[item] = s;
  }
}

// ---- backward, destination side (CSR) -----------------------------------------------------------------------
// delta[i,h] = <dout, out - bias>;  da_dst[i,h] = sum_j dz_ij;  stats[i,h] = {a_dst, m, 1/(s+eps), delta}
template <typename T, int LPR, int VPL>
__global__ void __launch_bounds__(kThreads) gat_bwd_dst_kernel(GatArgs a, const T* __restrict__ h, const T* __restrict__ out,
                                                               const T* __restrict__ dout) {
  constexpr int GPW = 32 / LPR;
  const int gl = threadIdx.x & (LPR - 1);
  const int H = a.heads, C = a.chan, C4 = C >> 2;
  const int64_t item = (int64_t(blockIdx.x) * (kThreads / 32) + (threadIdx.x >> 5)) * GPW + ((threadIdx.x & 31) / LPR);
  if (item >= a.n_nodes * H) return;
  const unsigned mask = gmask<LPR>();
  const uint64_t seed = launch_seed(a);
  const int i = int(item / H), head = int(item - int64_t(i) * H);
  const int64_t F = int64_t(H) * C;
  const int start = a.ptr[i], end = a.ptr[i + 1];
  const float ad = __ldg(a.a_dst + item);
  const float m = __ldg(a.rowmax + item);
  const float inv = 1.f / (__ldg(a.rowsum + item) + 1e-16f);

  float4 dO[VPL];
  float part = 0.f;
#pragma unroll
  for (int v = 0; v < VPL; ++v) {
    const int vec = gl + v * LPR;
    dO[v] = f4_zero();
    if (vec < C4) {
      dO[v] = Vec4<T>::stream(dout + int64_t(i) * F + head * C, vec * 4);
      float4 o = Vec4<T>::stream(out + int64_t(i) * F + head * C, vec * 4);
      if (a.act_out) {  // `dout` is the gradient of elu(out): turn it into the gradient of out (elu' = y > 0 ? 1 : y + 1
                        // with y = elu(out), as lgk_elu_bwd computes it) and keep it for the src-side pass and db
        dO[v].x *= o.x > 0.f ? 1.f : elu1(o.x) + 1.f;
        dO[v].y *= o.y > 0.f ? 1.f : elu1(o.y) + 1.f;
        dO[v].z *= o.z > 0.f ? 1.f : elu1(o.z) + 1.f;
        dO[v].w *= o.w > 0.f ? 1.f : elu1(o.w) + 1.f;
        Vec4<T>::store(static_cast<T*>(a.act_out) + int64_t(i) * F + head * C, vec * 4, dO[v]);
      }
      if (a.bias) {
        float4 b = __ldg(reinterpret_cast<const float4*>(a.bias + head * C) + vec);
        o.x -= b.x; o.y -= b.y; o.z -= b.z; o.w -= b.w;
      }
      part += f4_dot(dO[v], o);
    }
  }
  const float delta = group_sum<LPR>(part, mask);
  float da = 0.f;

  // One edge: its source row and logit half arrive as arguments so that the caller can issue the loads of several edges
  // before the first dependent shuffle -- a (node, head) group walks only ~7 edges, the chain of L2 round trips is its time.
  auto edge = [&](const float4 (&hv)[VPL], float a_src_j, uint32_t slot) {
    float dotp = 0.f;
#pragma unroll
    for (int v = 0; v < VPL; ++v) dotp += f4_dot(dO[v], hv[v]);  // vectors past C4: dO = 0
    dotp = group_sum<LPR>(dotp, mask);
    const float z = a_src_j + ad;
    const float alpha = expf(lrelu(z, a.slope) - m) * inv;
    const float dalpha = keep_scale(a, seed, slot, head) * dotp;
    da += alpha * (dalpha - delta) * (z > 0.f ? 1.f : a.slope);
  };
  auto load_row = [&](int j, float4 (&hv)[VPL], float& a_src_j) {
    const T* hr = h + int64_t(j) * F + head * C;
#pragma unroll
    for (int v = 0; v < VPL; ++v) {
      const int vec = gl + v * LPR;
      hv[v] = vec < C4 ? Vec4<T>::gather(hr, vec * 4) : f4_zero();
    }
    a_src_j = __ldg(a.a_src + int64_t(j) * H + head);
  };

  if (a.add_self) {
    float4 hv[VPL];
    float as;
    load_row(i, hv, as);
    edge(hv, as, kSelfSlotBit | uint32_t(i));
  }
  if constexpr (LPR == 1 && VPL > 1) {
    // thread per (node, head) (narrow heads, see LGK_GAT_DISPATCH_BWD): no shuffles, the per-edge scalar work (logit, exp,
    // dropout hash) once per edge instead of once per lane of a group.  Ids in chunks of kThinIds (one round trip), rows of
    // kThinRows edges at a time (one round trip each).
    for (int base = start; base < end; base += kThinIds) {
      int js[kThinIds];
#pragma unroll
      for (int t = 0; t < kThinIds; ++t) {
        js[t] = -1;
        if (base + t < end) {
          const int j = ld_stream_i32(a.idx + base + t);
          if (!(a.add_self && j == i)) js[t] = j;
        }
      }
#pragma unroll
      for (int u0 = 0; u0 < kThinIds; u0 += kThinRows) {
        if (base + u0 >= end) break;
        float4 hv[kThinRows][VPL];
        float as[kThinRows];
#pragma unroll
        for (int t = 0; t < kThinRows; ++t) load_row(js[u0 + t] >= 0 ? js[u0 + t] : i, hv[t], as[t]);
#pragma unroll
        for (int t = 0; t < kThinRows; ++t)
          if (js[u0 + t] >= 0) edge(hv[t], as[t], uint32_t(base + u0 + t));
      }
    }
    a.da_out[item] = da;
    a.stats[item] = make_float4(ad, m, inv, delta);
    return;
  }
  constexpr int UB = kEdgeBatch<LPR, VPL>;  // edges whose loads are in flight together
  for (int base = start; base < end; base += LPR) {
    const int p = base + gl;
    int my_j = -1;
    if (p < end) {
      const int j = ld_stream_i32(a.idx + p);
      if (!(a.add_self && j == i)) my_j = j;
    }
    const int cnt = min(LPR, end - base);
    for (int u0 = 0; u0 < cnt; u0 += UB) {
      int js[UB];
      float4 hv[UB][VPL];
      float as[UB];
#pragma unroll
      for (int t = 0; t < UB; ++t) {
        js[t] = __shfl_sync(mask, my_j, u0 + t, LPR);  // lanes past the row's end and dropped stored loops hold -1
        load_row(js[t] >= 0 ? js[t] : i, hv[t], as[t]);
      }
#pragma unroll
      for (int t = 0; t < UB; ++t)
        if (js[t] >= 0) edge(hv[t], as[t], uint32_t(base + u0 + t));  // uniform across the group, same order as before
    }
  }
  if (gl == 0) {
    a.da_out[item] = da;
    a.stats[item] = make_float4(ad, m, inv, delta);
  }
}

// ---- backward, source side (CSC) ------------------------------------------------------------------------------
// dh[j,h,:] = sum_i alpha_ij keep_ij dout[i,h,:] + da_src att_src + da_dst att_dst;  da_src[j,h] = sum_i dz_ij
template <typename T, int LPR, int VPL>
__global__ void __launch_bounds__(kThreads) gat_bwd_src_kernel(GatArgs a, const T* __restrict__ h, const T* __restrict__ dout,
                                                               T* __restrict__ dh) {
  constexpr int GPW = 32 / LPR;
  const int gl = threadIdx.x & (LPR - 1);
  const int H = a.heads, C = a.chan, C4 = C >> 2;
  const int64_t item = (int64_t(blockIdx.x) * (kThreads / 32) + (threadIdx.x >> 5)) * GPW + ((threadIdx.x & 31) / LPR);
  if (item >= a.n_nodes * H) return;
  const unsigned mask = gmask<LPR>();
  const uint64_t seed = launch_seed(a);
  const int j = int(item / H), head = int(item - int64_t(j) * H);
  const int64_t F = int64_t(H) * C;
  const int start = a.ptr[j], end = a.ptr[j + 1];
  const float as = __ldg(a.a_src + item);

  float4 hj[VPL], acc[VPL];
#pragma unroll
  for (int v = 0; v < VPL; ++v) {
    const int vec = gl + v * LPR;
    hj[v] = f4_zero();
    acc[v] = f4_zero();
    if (vec < C4) hj[v] = Vec4<T>::stream(h + int64_t(j) * F + head * C, vec * 4);
  }
  float da = 0.f;

  auto edge = [&](const float4 st, const float4 (&dO)[VPL], uint32_t slot) {  // st = {a_dst, m, inv, delta} of the destination
    float dotp = 0.f;
#pragma unroll
    for (int v = 0; v < VPL; ++v) dotp += f4_dot(dO[v], hj[v]);  // vectors past C4: both 0
    dotp = group_sum<LPR>(dotp, mask);
    const float z = as + st.x;
    const float alpha = expf(lrelu(z, a.slope) - st.y) * st.z;
    const float ks = keep_scale(a, seed, slot, head);
    da += alpha * (ks * dotp - st.w) * (z > 0.f ? 1.f : a.slope);
    const float w = alpha * ks;
#pragma unroll
    for (int v = 0; v < VPL; ++v) f4_fma(acc[v], w, dO[v]);
  };
  auto load_row = [&](int i, float4& st, float4 (&dO)[VPL]) {
    st = __ldg(a.stats + int64_t(i) * H + head);  // same address for the group
    const T* dr = dout + int64_t(i) * F + head * C;
#pragma unroll
    for (int v = 0; v < VPL; ++v) {
      const int vec = gl + v * LPR;
      dO[v] = vec < C4 ? Vec4<T>::gather(dr, vec * 4) : f4_zero();
    }
  };

  if (a.add_self) {
    float4 st, dO[VPL];
    load_row(j, st, dO);
    edge(st, dO, kSelfSlotBit | uint32_t(j));
  }
  if constexpr (LPR == 1 && VPL > 1) {  // thread per (node, head): see gat_bwd_dst_kernel
    for (int base = start; base < end; base += kThinIds) {
      int is[kThinIds], slots[kThinIds];
#pragma unroll
      for (int t = 0; t < kThinIds; ++t) {
        is[t] = -1;
        slots[t] = 0;
        if (base + t < end) {
          const int i = ld_stream_i32(a.idx + base + t);
          if (!(a.add_self && i == j)) {
            is[t] = i;
            if (a.drop_p > 0.f) slots[t] = ld_stream_i32(a.slotmap + base + t);
          }
        }
      }
#pragma unroll
      for (int u0 = 0; u0 < kThinIds; u0 += kThinRows) {
        if (base + u0 >= end) break;
        float4 st[kThinRows], dO[kThinRows][VPL];
#pragma unroll
        for (int t = 0; t < kThinRows; ++t) load_row(is[u0 + t] >= 0 ? is[u0 + t] : j, st[t], dO[t]);
#pragma unroll
        for (int t = 0; t < kThinRows; ++t)
          if (is[u0 + t] >= 0) edge(st[t], dO[t], uint32_t(slots[u0 + t]));
      }
    }
  } else {
  constexpr int UB = kEdgeBatch<LPR, VPL>;  // edges whose loads are in flight together (see gat_bwd_dst_kernel)
  for (int base = start; base < end; base += LPR) {
    const int q = base + gl;
    int my_i = -1, my_slot = 0;
    if (q < end) {
      const int i = ld_stream_i32(a.idx + q);
      if (!(a.add_self && i == j)) {
        my_i = i;
        if (a.drop_p > 0.f) my_slot = ld_stream_i32(a.slotmap + q);
      }
    }
    const int cnt = min(LPR, end - base);
    for (int u0 = 0; u0 < cnt; u0 += UB) {
      int is[UB], slots[UB];
      float4 st[UB], dO[UB][VPL];
#pragma unroll
      for (int t = 0; t < UB; ++t) {
        is[t] = __shfl_sync(mask, my_i, u0 + t, LPR);
        slots[t] = __shfl_sync(mask, my_slot, u0 + t, LPR);
        load_row(is[t] >= 0 ? is[t] : j, st[t], dO[t]);
      }
#pragma unroll
      for (int t = 0; t < UB; ++t)
        if (is[t] >= 0) edge(st[t], dO[t], uint32_t(slots[t]));
    }
  }

  }  // group-of-lanes mapping
  const float dad = __ldg(a.da_dst + item);
#pragma unroll
  for (int v = 0; v < VPL; ++v) {
    const int vec = gl + v * LPR;
    if (vec < C4) {
      float4 r = acc[v];
      f4_fma(r, da, __ldg(reinterpret_cast<const float4*>(a.att_src + head * C) + vec));
      f4_fma(r, dad, __ldg(reinterpret_cast<const float4*>(a.att_dst + head * C) + vec));
      Vec4<T>::store(dh + int64_t(j) * F + head * C, vec * 4, r);
    }
  }
  if (gl == 0) a.da_out[item] = da;
}

// ---- datt = sum_n da[n,h] * h[n,h,c]: slab partials, then a deterministic second stage ----------------------------
// At most two CTAs per SM, each reducing a contiguous slab of nodes; threads = (row lane, float4 column) so a row is
// read with 16-byte loads and the two weights (da_src, da_dst of the column's head) are shared by the 4 channels.
int datt_slabs(int64_t N) {
  int64_t g = ceil_div(N, 64);
  if (g > 2 * kNumSMs) g = 2 * kNumSMs;
  return int(g < 1 ? 1 : g);
}
template <typename T>
__global__ void __launch_bounds__(256) gat_datt_partial_kernel(const T* __restrict__ h, const float* __restrict__ da_src,
                                                               const float* __restrict__ da_dst, float* __restrict__ partial,
                                                               int64_t N, int H, int C, int64_t slab_rows) {
  __shared__ float4 red_s[256], red_d[256];
  const int F = H * C, F4 = F >> 2;
  const int64_t r0 = int64_t(blockIdx.x) * slab_rows, r1 = min(N, r0 + slab_rows);
  float* p = partial + int64_t(blockIdx.x) * 2 * F;
  for (int c0 = 0; c0 < F4; c0 += 256) {
    const int width = min(256, F4 - c0);
    const int lanes = max(1, 256 / width);
    const int c = threadIdx.x % width, rl = threadIdx.x / width;
    float4 as = f4_zero(), ad = f4_zero();
    if (rl < lanes) {
      const int col = (c0 + c) * 4, head = col / C;
      for (int64_t n = r0 + rl; n < r1; n += lanes) {
        const float4 hv = Vec4<T>::stream(h + n * F, col);
        f4_fma(as, __ldg(da_src + n * H + head), hv);
        f4_fma(ad, __ldg(da_dst + n * H + head), hv);
      }
    }
    __syncthreads();
    red_s[threadIdx.x] = as;
    red_d[threadIdx.x] = ad;
    __syncthreads();
    if (rl == 0) {
      for (int l = 1; l < lanes; ++l) {
        const float4 u = red_s[l * width + c], v = red_d[l * width + c];
        as.x += u.x; as.y += u.y; as.z += u.z; as.w += u.w;
        ad.x += v.x; ad.y += v.y; ad.z += v.z; ad.w += v.w;
      }
      reinterpret_cast<float4*>(p)[c0 + c] = as;
      reinterpret_cast<float4*>(p + F)[c0 + c] = ad;
    }
  }
}
int check_shape(const char* who, int64_t N, int H, int C) {
  LGK_REQUIRE(N >= 0 && H > 0 && C > 0, LGK_ERR_BAD_ARG, "%s: bad sizes", who);
  LGK_REQUIRE(C % 4 == 0 && C <= 1024, LGK_ERR_UNSUPPORTED, "%s: channels per head must be a multiple of 4 and <= 1024 (got %d)", who, C);
  LGK_REQUIRE(N * H < (int64_t(1) << 31), LGK_ERR_UNSUPPORTED, "%s: N*H out of range", who);
  return LGK_OK;
}

// pick (LPR, VPL) from C/4 and launch KERNEL<T, LPR, VPL>
#define LGK_GAT_DISPATCH(KERNEL, T, n_items, C, stream, ...)                                         \
  do {                                                                                              \
    const int c4__ = (C) / 4;                                                                       \
    const int warps__ = kThreads / 32;                                                              \
    auto grid__ = [&](int lpr) { return unsigned(ceil_div((n_items), int64_t(warps__) * (32 / lpr))); }; \
    if (c4__ <= 1) KERNEL<T, 1, 1><<<grid__(1), kThreads, 0, stream>>>(__VA_ARGS__);                \
    else if (c4__ <= 2) KERNEL<T, 2, 1><<<grid__(2), kThreads, 0, stream>>>(__VA_ARGS__);           \
    else if (c4__ <= 4) KERNEL<T, 4, 1><<<grid__(4), kThreads, 0, stream>>>(__VA_ARGS__);           \
    else if (c4__ <= 8) KERNEL<T, 8, 1><<<grid__(8), kThreads, 0, stream>>>(__VA_ARGS__);           \
    else if (c4__ <= 16) KERNEL<T, 16, 1><<<grid__(16), kThreads, 0, stream>>>(__VA_ARGS__);        \
    else if (c4__ <= 32) KERNEL<T, 32, 1><<<grid__(32), kThreads, 0, stream>>>(__VA_ARGS__);        \
    else if (c4__ <= 64) KERNEL<T, 32, 2><<<grid__(32), kThreads, 0, stream>>>(__VA_ARGS__);        \
    else if (c4__ <= 128) KERNEL<T, 32, 4><<<grid__(32), kThreads, 0, stream>>>(__VA_ARGS__);       \
    else KERNEL<T, 32, 8><<<grid__(32), kThreads, 0, stream>>>(__VA_ARGS__);                        \
  } while (0)

// Backward passes with narrow heads (2 <= C / 4 <= 4, the reference's 4 x 16 and 2 x 16 layers): ONE THREAD per (node, head)
// instead of a group of C / 4 lanes.  The per-edge scalar work of these kernels (logit, exp, dropout hash, ~70 instructions) is
// the same for every lane of a group, so the group mapping executes it C / 4 times; a thread that owns the head's whole
// channel slice does it once, needs no shuffle, and issues its 2-4 16-byte gathers per edge back to back.  LGK_GAT_THIN=0 keeps
// the group mapping.
bool gat_thin() {
  static const bool on = [] { const char* v = getenv("LGK_GAT_THIN"); return !(v && *v == '0'); }();
  return on;
}
#define LGK_GAT_DISPATCH_BWD(KERNEL, T, n_items, C, stream, ...)                                                      \
  do {                                                                                                               \
    const int c4b__ = (C) / 4;                                                                                       \
    const unsigned gridb__ = unsigned(ceil_div((n_items), int64_t(kThreads)));                                       \
    if (gat_thin() && c4b__ == 2) KERNEL<T, 1, 2><<<gridb__, kThreads, 0, stream>>>(__VA_ARGS__);                   \
    else if (gat_thin() && (c4b__ == 3 || c4b__ == 4)) KERNEL<T, 1, 4><<<gridb__, kThreads, 0, stream>>>(__VA_ARGS__); \
    else LGK_GAT_DISPATCH(KERNEL, T, n_items, C, stream, __VA_ARGS__);                                               \
  } while (0)

GatArgs base_args(int64_t N, int H, int C, float slope, int add_self, float drop_p, uint64_t seed, const int64_t* seed_dev) {
  GatArgs a{};
  a.n_nodes = N; a.heads = H; a.chan = C; a.slope = slope; a.add_self = add_self;
  a.drop_p = drop_p; a.drop_scale = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f; a.seed = seed; a.seed_dev = seed_dev;
  return a;
}

}  // namespace
}  // namespace lgk

using namespace lgk;

extern "C" int lgk_gat_alpha(const void* h, const float* att_src, const float* att_dst, float* a_src, float* a_dst,
                             int64_t N, int H, int C, int dtype, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = check_shape("lgk_gat_alpha", N, H, C)) return rc;
  if (N == 0) return LGK_OK;
  LGK_REQUIRE(h && att_src && att_dst && a_src && a_dst, LGK_ERR_BAD_ARG, "lgk_gat_alpha: null pointer");
  const int64_t items = N * H;
  if (dtype == LGK_F32)
    LGK_GAT_DISPATCH(gat_alpha_kernel, float, items, C, stream, static_cast<const float*>(h), att_src, att_dst, a_src, a_dst, items, H, C);
  else if (dtype == LGK_BF16)
    LGK_GAT_DISPATCH(gat_alpha_kernel, __nv_bfloat16, items, C, stream, static_cast<const __nv_bfloat16*>(h), att_src, att_dst, a_src, a_dst, items, H, C);
  else
    LGK_REQUIRE(false, LGK_ERR_BAD_ARG, "lgk_gat_alpha: bad dtype");
  LGK_CHECK_LAUNCH("gat_alpha_kernel");
  return LGK_OK;
}

static int gat_fwd_impl(const int32_t* rowptr, const int32_t* col, const void* h, const float* a_src, const float* a_dst,
                           const float* bias, void* out, float* rowmax, float* rowsum, int64_t N, int H, int C,
                        float negative_slope, int add_self_loops, float dropout_p, uint64_t seed, const int64_t* seed_dev, int dtype,
                        void* stream_, void* out_elu) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = check_shape("lgk_gat_fwd", N, H, C)) return rc;
  if (N == 0) return LGK_OK;
  LGK_REQUIRE(rowptr && h && a_src && a_dst && out, LGK_ERR_BAD_ARG, "lgk_gat_fwd: null pointer");
  LGK_REQUIRE(dropout_p >= 0.f && dropout_p < 1.f, LGK_ERR_BAD_ARG, "lgk_gat_fwd: dropout must be in [0,1)");
  GatArgs a = base_args(N, H, C, negative_slope, add_self_loops, dropout_p, seed, seed_dev);
  a.ptr = rowptr; a.idx = col; a.a_src = a_src; a.a_dst = a_dst; a.bias = bias; a.rowmax = rowmax; a.rowsum = rowsum;
  a.act_out = out_elu;
  const int64_t items = N * H;
  if (dtype == LGK_F32)
    LGK_GAT_DISPATCH(gat_fwd_kernel, float, items, C, stream, a, static_cast<const float*>(h), static_cast<float*>(out));
  else if (dtype == LGK_BF16)
    LGK_GAT_DISPATCH(gat_fwd_kernel, __nv_bfloat16, items, C, stream, a, static_cast<const __nv_bfloat16*>(h), static_cast<__nv_bfloat16*>(out));
  else
    LGK_REQUIRE(false, LGK_ERR_BAD_ARG, "lgk_gat_fwd: bad dtype");
  LGK_CHECK_LAUNCH("gat_fwd_kernel");
  return LGK_OK;
}

extern "C" int lgk_gat_fwd(const int32_t* rowptr, const int32_t* col, const void* h, const float* a_src, const float* a_dst,
                           const float* bias, void* out, float* rowmax, float* rowsum, int64_t N, int H, int C,
                           float negative_slope, int add_self_loops, float dropout_p, uint64_t seed, const int64_t* seed_dev, int dtype,
                           void* stream_) {
  return gat_fwd_impl(rowptr, col, h, a_src, a_dst, bias, out, rowmax, rowsum, N, H, C, negative_slope, add_self_loops,
                      dropout_p, seed, seed_dev, dtype, stream_, nullptr);
}

extern "C" int lgk_gat_fwd_elu(const int32_t* rowptr, const int32_t* col, const void* h, const float* a_src,
                               const float* a_dst, const float* bias, void* out, void* out_elu, float* rowmax,
                               float* rowsum, int64_t N, int H, int C, float negative_slope, int add_self_loops,
                               float dropout_p, uint64_t seed, const int64_t* seed_dev, int dtype, void* stream_) {
  LGK_REQUIRE(out_elu || N == 0, LGK_ERR_BAD_ARG, "lgk_gat_fwd_elu: null pointer");
  return gat_fwd_impl(rowptr, col, h, a_src, a_dst, bias, out, rowmax, rowsum, N, H, C, negative_slope, add_self_loops,
                      dropout_p, seed, seed_dev, dtype, stream_, out_elu);
}

static int gat_bwd_dst_impl(const int32_t* rowptr, const int32_t* col, const void* h, const float* a_src,
                               const float* a_dst, const float* rowmax, const float* rowsum, const void* out,
                               const float* bias, const void* dout, float* delta, float* da_dst, int64_t N, int H, int C,
                               float negative_slope, int add_self_loops, float dropout_p, uint64_t seed, const int64_t* seed_dev, int dtype,
                               void* stream_, void* dpre_out) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (int rc = check_shape("lgk_gat_bwd_dst", N, H, C)) return rc;
  if (N == 0) return LGK_OK;
  LGK_REQUIRE(rowptr && h && a_src && a_dst && rowmax && rowsum && out && dout && delta && da_dst, LGK_ERR_BAD_ARG,
              "lgk_gat_bwd_dst: null pointer");
  GatArgs a = base_args(N, H, C, negative_slope, add_self_loops, dropout_p, seed, seed_dev);
  a.ptr = rowptr; a.idx = col; a.a_src = a_src; a.a_dst = a_dst; a.bias = bias;
  a.rowmax = const_cast<float*>(rowmax); a.rowsum = const_cast<float*>(rowsum);
  a.stats = reinterpret_cast<float4*>(delta); a.da_out = da_dst; a.act_out = dpre_out;
  const int64_t items = N * H;
  if (dtype == LGK_F32)
    LGK_GAT_DISPATCH_BWD(gat_bwd_dst_kernel, float, items, C, stream, a, static_cast<const float*>(h), static_cast<const float*>(out), static_cast<const float*>(dout));
  else if (dtype == LGK_BF16)
    LGK_GAT_DISPATCH_BWD(gat_bwd_dst_kernel, __nv_bfloat16, items, C, stream, a, static_cast<const __nv_bfloat16*>(h), static_cast<const __nv_bfloat16*>(out), static_cast<const __nv_bfloat16*>(dout));
  else
    LGK_REQUIRE(false, LGK_ERR_BAD_ARG, "lgk_gat_bwd_dst: bad dtype");
  LGK_CHECK_LAUNCH("gat_bwd_dst_kernel");
  return LGK_OK;
}

extern "C" int lgk_gat_bwd_dst(const int32_t* rowptr, const int32_t* col, const void* h, const float* a_src,
                               const float* a_dst, const float* rowmax, const float* rowsum, const void* out,
                               const float* bias, const void* dout, float* delta, float* da_dst, int64_t N, int H, int C,
                               float negative_slope, int add_self_loops, float dropout_p, uint64_t seed, const int64_t* seed_dev, int dtype,
                               void* stream_) {
  return gat_bwd_dst_impl(rowptr, col, h, a_src, a_dst, rowmax, rowsum, out, bias, dout, delta, da_dst, N, H, C,
                          negative_slope, add_self_loops, dropout_p, seed, seed_dev, dtype, stream_, nullptr);
}

extern "C" int lgk_gat_bwd_dst_elu(const int32_t* rowptr, const int32_t* col, const void* h, const float* a_src,
                                   const float* a_dst, const float* rowmax, const float* rowsum, const void* out,
                                   const float* bias, const void* dy, void* dout, float* delta, float* da_dst, int64_t N,
                                   int H, int C, float negative_slope, int add_self_loops, float dropout_p, uint64_t seed,
                                   const int64_t* seed_dev, int dtype, void* stream_) {
  LGK_REQUIRE(dout || N == 0, LGK_ERR_BAD_ARG, "lgk_gat_bwd_dst_elu: null pointer");
  return gat_bwd_dst_impl(rowptr, col, h, a_src, a_dst, rowmax, rowsum, out, bias, dy, delta, da_dst, N, H, C,
                          negative_slope, add_self_loops, dropout_p, seed, seed_dev, dtype, stream_, dout);
}

extern "C" int lgk_gat_bwd_src(const int32_t* colptr, const int32_t* row, const int32_t* csr2csc, const void* h,
                               const float* a_src, const float* a_dst, const float* rowmax, const float* rowsum,
                               const float* delta, const float* da_dst, const float* att_src, const float* att_dst,
                               const void* dout, void* dh, float* da_src, int64_t N, int H, int C, float negative_slope,
                               int add_self_loops, float dropout_p, uint64_t seed, const int64_t* seed_dev, int dtype, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  (void)a_dst; (void)rowmax; (void)rowsum;  // carried inside the packed statistics written by lgk_gat_bwd_dst
  if (int rc = check_shape("lgk_gat_bwd_src", N, H, C)) return rc;
  if (N == 0) return LGK_OK;
  LGK_REQUIRE(colptr && h && a_src && delta && da_dst && att_src && att_dst && dout && dh && da_src, LGK_ERR_BAD_ARG,
              "lgk_gat_bwd_src: null pointer");
  LGK_REQUIRE(dropout_p <= 0.f || csr2csc, LGK_ERR_BAD_ARG, "lgk_gat_bwd_src: dropout needs csr2csc");
  GatArgs a = base_args(N, H, C, negative_slope, add_self_loops, dropout_p, seed, seed_dev);
  a.ptr = colptr; a.idx = row; a.slotmap = csr2csc; a.a_src = a_src;
  a.stats = reinterpret_cast<float4*>(const_cast<float*>(delta)); a.da_dst = da_dst; a.da_out = da_src;
  a.att_src = att_src; a.att_dst = att_dst;
  const int64_t items = N * H;
  if (dtype == LGK_F32)
    LGK_GAT_DISPATCH_BWD(gat_bwd_src_kernel, float, items, C, stream, a, static_cast<const float*>(h), static_cast<const float*>(dout), static_cast<float*>(dh));
  else if (dtype == LGK_BF16)
    LGK_GAT_DISPATCH_BWD(gat_bwd_src_kernel, __nv_bfloat16, items, C, stream, a, static_cast<const __nv_bfloat16*>(h), static_cast<const __nv_bfloat16*>(dout), static_cast<__nv_bfloat16*>(dh));
  else
    LGK_REQUIRE(false, LGK_ERR_BAD_ARG, "lgk_gat_bwd_src: bad dtype");
  LGK_CHECK_LAUNCH("gat_bwd_src_kernel");
  return LGK_OK;
}

extern "C" int64_t lgk_gat_datt_workspace_bytes(int64_t N, int H, int C) {
  if (N <= 0 || H <= 0 || C <= 0) return 0;
  return int64_t(datt_slabs(N)) * 2 * H * C * 4;
}

extern "C" int lgk_gat_datt(const void* h, const float* da_src, const float* da_dst, float* datt_src, float* datt_dst,
                            int64_t N, int H, int C, int accumulate, void* workspace, int64_t workspace_bytes, int dtype,
                            void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(N >= 0 && H > 0 && C > 0, LGK_ERR_BAD_ARG, "lgk_gat_datt: bad sizes");
  LGK_REQUIRE(datt_src && datt_dst, LGK_ERR_BAD_ARG, "lgk_gat_datt: null pointer");
  const int F = H * C;
  if (N == 0) {
    if (!accumulate) {
      cudaMemsetAsync(datt_src, 0, 4 * F, stream);
      cudaMemsetAsync(datt_dst, 0, 4 * F, stream);
      LGK_CHECK_LAUNCH("lgk_gat_datt(memset)");
    }
    return LGK_OK;
  }
  LGK_REQUIRE(h && da_src && da_dst, LGK_ERR_BAD_ARG, "lgk_gat_datt: null pointer");
  const int64_t nslabs = datt_slabs(N);
  const int64_t slab_rows = ceil_div(N, nslabs);
  LGK_REQUIRE(C % 4 == 0, LGK_ERR_UNSUPPORTED, "lgk_gat_datt: channels per head must be a multiple of 4");
  const int64_t need = nslabs * 2 * F * 4;
  LGK_REQUIRE(workspace && workspace_bytes >= need, LGK_ERR_WORKSPACE, "lgk_gat_datt: workspace %lld < %lld bytes",
              (long long)workspace_bytes, (long long)need);
  float* partial = static_cast<float*>(workspace);
  const unsigned grid = unsigned(nslabs);
  if (dtype == LGK_F32)
    gat_datt_partial_kernel<float><<<grid, 256, 0, stream>>>(static_cast<const float*>(h), da_src, da_dst, partial, N, H, C, slab_rows);
  else if (dtype == LGK_BF16)
    gat_datt_partial_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>(static_cast<const __nv_bfloat16*>(h), da_src, da_dst, partial, N, H, C, slab_rows);
  else
    LGK_REQUIRE(false, LGK_ERR_BAD_ARG, "lgk_gat_datt: bad dtype");
  LGK_CHECK_LAUNCH("gat_datt_partial_kernel");
  slab_reduce_kernel<<<slab_reduce_grid(2 * F), 256, 0, stream>>>(partial, datt_src, datt_dst, F, F, int(nslabs), accumulate);
  LGK_CHECK_LAUNCH("slab_reduce_kernel(datt)");
  return LGK_OK;
}
