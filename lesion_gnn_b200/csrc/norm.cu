// BatchNorm1d + activation + dropout of GIN's MLP as three launches forward and three backward
// (torch_geometric.nn.models.MLP: Linear -> BatchNorm1d -> ELU -> Dropout -> Linear; reference models/gin.py:23,
// SURVEY.md 8a a-4), and the activation + dropout pair that follows every GIN layer (gin.py:31-32).
//
// HBM-bound column statistics over a tall [N, d] matrix (d = 32..512): a CTA owns a slab of rows and all columns, a
// thread a float4 of columns; per slab and column (mean, M2) are formed around the slab's first row (no cancellation),
// the slabs are combined with Chan's formula in double precision by a one-CTA finalise that also updates the running
// statistics (momentum, unbiased variance, num_batches_tracked) -- no atomics, deterministic.  The normalise + affine +
// ELU + dropout pass and the two backward passes recompute everything from x and the saved (mean, invstd): nothing of
// size [N, d] is stored besides the layer's own input and output.
#include "common.cuh"
#include <math.h>

namespace lgk {
namespace {

constexpr int kThreads = 256;

int norm_slabs(int64_t n) {
  int64_t s = ceil_div(n, 256);
  if (s > 2 * kNumSMs) s = 2 * kNumSMs;
  return int(s < 1 ? 1 : s);
}

// counter RNG for feature dropout: (seed, element index) -> uniform in [0,1)
__device__ __forceinline__ float uniform_hash64(uint64_t seed, uint64_t idx) {
  uint64_t z = seed + 0x9E3779B97F4A7C15ull * (idx + 1);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z = z ^ (z >> 31);
  return float(uint32_t(z >> 40)) * (1.0f / 16777216.0f);
}
__device__ __forceinline__ uint64_t launch_seed(uint64_t seed, const int64_t* seed_dev) {
  return seed_dev ? seed + 0xD1B54A32D192ED03ull * uint64_t(__ldg(seed_dev)) : seed;
}
__device__ __forceinline__ float act_fwd(float z, int act) {
  if (act == LGK_ACT_ELU) return elu1(z);
  if (act == LGK_ACT_RELU) return fmaxf(z, 0.f);
  return z;
}
// derivative of the activation at pre-activation z
__device__ __forceinline__ float act_grad(float z, int act) {
  if (act == LGK_ACT_ELU) return z > 0.f ? 1.f : expm1f(z) + 1.f;
  if (act == LGK_ACT_RELU) return z > 0.f ? 1.f : 0.f;
  return 1.f;
}

// activations are float or bf16 (statistics, parameters and every accumulator fp32)
template <typename T>
__device__ __forceinline__ float ldx(const T* p, int64_t i);
template <>
__device__ __forceinline__ float ldx<float>(const float* p, int64_t i) { return ld_stream_f32(p + i); }
template <>
__device__ __forceinline__ float ldx<__nv_bfloat16>(const __nv_bfloat16* p, int64_t i) { return __bfloat162float(p[i]); }

// ---- pass 1: per-slab column statistics around the slab's first row ------------------------------------------
// partial[slab][0][c] = sum (x - K_c), partial[slab][1][c] = sum (x - K_c)^2, partial[slab][2][c] = K_c
template <typename T>
__global__ void __launch_bounds__(kThreads) bn_stats_partial_kernel(const T* __restrict__ x, float* __restrict__ partial,
                                                                    int64_t n, int d, int64_t slab_rows) {
  __shared__ float red[2][kThreads];
  const int64_t r0 = int64_t(blockIdx.x) * slab_rows, r1 = min(n, r0 + slab_rows);
  float* p = partial + int64_t(blockIdx.x) * 3 * d;
  for (int c0 = 0; c0 < d; c0 += kThreads) {
    const int width = min(kThreads, d - c0);
    const int lanes = kThreads / width;  // row lanes per column
    const int c = threadIdx.x % width, rl = threadIdx.x / width;
    float s = 0.f, q = 0.f;
    const float K = (r0 < r1) ? ldx(x, r0 * d + c0 + c) : 0.f;
    if (rl < lanes) {
      for (int64_t r = r0 + rl; r < r1; r += lanes) {
        const float v = ldx(x, r * d + c0 + c) - K;
        s += v;
        q = fmaf(v, v, q);
      }
    }
    __syncthreads();
    red[0][threadIdx.x] = s;
    red[1][threadIdx.x] = q;
    __syncthreads();
    if (rl == 0) {
      for (int l = 1; l < lanes; ++l) { s += red[0][l * width + c]; q += red[1][l * width + c]; }
      p[c0 + c] = s;
      p[d + c0 + c] = q;
      p[2 * d + c0 + c] = K;
    }
  }
}

// ---- pass 2: combine the slabs (Chan), write mean / invstd, update the running statistics -----------------------
// 8 columns x 32 slab lanes per CTA: lane l combines slabs l, l + 32, ... in ascending order, lane 0 then combines the 32
// lane results in lane order -- a fixed tree, deterministic.  (One thread per column walking all ~300 slabs was a chain
// of ~300 dependent L2 round trips: 110 us for d = 64.)
constexpr int kFinCols = 8, kFinLanes = kThreads / kFinCols;
__device__ __forceinline__ void chan_combine(double& cnt, double& mean, double& m2, double nb, double mb, double m2b) {
  if (nb <= 0.0) return;
  const double delta = mb - mean, tot = cnt + nb;
  mean += delta * nb / tot;
  m2 += m2b + delta * delta * cnt * nb / tot;
  cnt = tot;
}
__global__ void __launch_bounds__(kThreads) bn_stats_finalize_kernel(const float* __restrict__ partial, int slabs, int64_t n, int d,
                                                                     int64_t slab_rows, float eps, float momentum,
                                                                     float* __restrict__ mean_out, float* __restrict__ invstd_out,
                                                                     float* __restrict__ running_mean, float* __restrict__ running_var,
                                                                     int64_t* __restrict__ num_batches_tracked) {
  __shared__ double s_cnt[kFinLanes][kFinCols], s_mean[kFinLanes][kFinCols], s_m2[kFinLanes][kFinCols];
  const int cl = threadIdx.x % kFinCols, lane = threadIdx.x / kFinCols;
  const int c = blockIdx.x * kFinCols + cl;
  double cnt = 0.0, mean = 0.0, m2 = 0.0;
  if (c < d) {
    for (int sl = lane; sl < slabs; sl += kFinLanes) {
      const int64_t rows = min(n, int64_t(sl + 1) * slab_rows) - int64_t(sl) * slab_rows;
      if (rows <= 0) break;
      const float* p = partial + int64_t(sl) * 3 * d;
      const double nb = double(rows), sb = double(p[c]), qb = double(p[d + c]), Kb = double(p[2 * d + c]);
      chan_combine(cnt, mean, m2, nb, Kb + sb / nb, qb - sb * sb / nb);
    }
  }
  s_cnt[lane][cl] = cnt; s_mean[lane][cl] = mean; s_m2[lane][cl] = m2;
  __syncthreads();
  if (lane == 0 && c < d) {
    for (int l = 1; l < kFinLanes; ++l) chan_combine(cnt, mean, m2, s_cnt[l][cl], s_mean[l][cl], s_m2[l][cl]);
    const double var = cnt > 0 ? m2 / cnt : 0.0;
    mean_out[c] = float(mean);
    invstd_out[c] = float(1.0 / sqrt(var + double(eps)));
    if (running_mean) running_mean[c] = float((1.0 - momentum) * double(running_mean[c]) + momentum * mean);
    if (running_var) running_var[c] = float((1.0 - momentum) * double(running_var[c]) + momentum * (cnt > 1 ? m2 / (cnt - 1) : var));
  }
  if (num_batches_tracked && blockIdx.x == 0 && threadIdx.x == 0) *num_batches_tracked += 1;
}

// eval mode: mean / invstd from the running statistics
__global__ void __launch_bounds__(kThreads) bn_eval_stats_kernel(const float* __restrict__ running_mean, const float* __restrict__ running_var,
                                                                 float eps, int d, float* __restrict__ mean_out, float* __restrict__ invstd_out) {
  const int c = blockIdx.x * kThreads + threadIdx.x;
  if (c < d) {
    mean_out[c] = running_mean[c];
    invstd_out[c] = 1.0f / sqrtf(running_var[c] + eps);
  }
}

// ---- pass 3: y = dropout(act((x - mean) * invstd * w + b)) --------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(kThreads) bn_act_fwd_kernel(const T* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                                                              const float* __restrict__ mean, const float* __restrict__ invstd,
                                                              T* __restrict__ y, int64_t total, int d, int act, float drop_p, uint64_t seed,
                                                              const int64_t* __restrict__ seed_dev) {
  const int64_t i = int64_t(blockIdx.x) * kThreads + threadIdx.x;
  if (i >= total) return;
  const int c = int(i % d);
  const float scale = __ldg(invstd + c) * (w ? __ldg(w + c) : 1.f);
  const float shift = (b ? __ldg(b + c) : 0.f) - __ldg(mean + c) * scale;
  float v = act_fwd(fmaf(ldx(x, i), scale, shift), act);
  if (drop_p > 0.f) v = uniform_hash64(launch_seed(seed, seed_dev), uint64_t(i)) < drop_p ? 0.f : v * (1.f / (1.f - drop_p));
  Vec4<T>::store1(y, i, v);
}

// ---- backward pass 1: per-slab sums of g and g * xhat, g = dy * keep * act'(z) ------------------------------------
template <typename T>
__global__ void __launch_bounds__(kThreads) bn_bwd_partial_kernel(const T* __restrict__ x, const T* __restrict__ dy, const float* __restrict__ w,
                                                                  const float* __restrict__ b, const float* __restrict__ mean,
                                                                  const float* __restrict__ invstd, float* __restrict__ partial, int64_t n,
                                                                  int d, int64_t slab_rows, int act, float drop_p, uint64_t seed,
                                                                  const int64_t* __restrict__ seed_dev) {
  __shared__ float red[2][kThreads];
  const int64_t r0 = int64_t(blockIdx.x) * slab_rows, r1 = min(n, r0 + slab_rows);
  float* p = partial + int64_t(blockIdx.x) * 2 * d;
  const uint64_t sd = drop_p > 0.f ? launch_seed(seed, seed_dev) : 0;
  const float keep = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
  for (int c0 = 0; c0 < d; c0 += kThreads) {
    const int width = min(kThreads, d - c0);
    const int lanes = kThreads / width;
    const int c = threadIdx.x % width, rl = threadIdx.x / width;
    const float m = __ldg(mean + c0 + c), is = __ldg(invstd + c0 + c);
    const float wc = w ? __ldg(w + c0 + c) : 1.f, bc = b ? __ldg(b + c0 + c) : 0.f;
    float sg = 0.f, sgx = 0.f;
    if (rl < lanes) {
      for (int64_t r = r0 + rl; r < r1; r += lanes) {
        const int64_t i = r * d + c0 + c;
        const float xh = (ldx(x, i) - m) * is;
        float g = ldx(dy, i) * act_grad(fmaf(xh, wc, bc), act);
        if (drop_p > 0.f) g = uniform_hash64(sd, uint64_t(i)) < drop_p ? 0.f : g * keep;
        sg += g;
        sgx = fmaf(g, xh, sgx);
      }
    }
    __syncthreads();
    red[0][threadIdx.x] = sg;
    red[1][threadIdx.x] = sgx;
    __syncthreads();
    if (rl == 0) {
      for (int l = 1; l < lanes; ++l) { sg += red[0][l * width + c]; sgx += red[1][l * width + c]; }
      p[c0 + c] = sg;          // -> d bias
      p[d + c0 + c] = sgx;     // -> d weight
    }
  }
}

// ---- backward pass 3: dx (training: batch statistics take part in the gradient; eval: plain scale) ----------------
// sums[0..d) = sum g, sums[d..2d) = sum g * xhat.  Block 0 also adds them into the parameter gradients when asked to.
template <typename T>
__global__ void __launch_bounds__(kThreads) bn_bwd_dx_kernel(const T* __restrict__ x, const T* __restrict__ dy, const float* __restrict__ w,
                                                             const float* __restrict__ b, const float* __restrict__ mean,
                                                             const float* __restrict__ invstd, const float* __restrict__ sums,
                                                             T* __restrict__ dx, int64_t n, int d, int training, int act, float drop_p,
                                                             uint64_t seed, const int64_t* __restrict__ seed_dev) {
  const int64_t i = int64_t(blockIdx.x) * kThreads + threadIdx.x;
  if (i >= n * d) return;
  const int c = int(i % d);
  const float m = __ldg(mean + c), is = __ldg(invstd + c);
  const float wc = w ? __ldg(w + c) : 1.f, bc = b ? __ldg(b + c) : 0.f;
  const float xh = (ldx(x, i) - m) * is;
  float g = ldx(dy, i) * act_grad(fmaf(xh, wc, bc), act);
  if (drop_p > 0.f) g = uniform_hash64(launch_seed(seed, seed_dev), uint64_t(i)) < drop_p ? 0.f : g * (1.f / (1.f - drop_p));
  float r;
  if (training) {
    const float inv_n = 1.f / float(n);
    r = wc * is * (g - (__ldg(sums + c) + xh * __ldg(sums + d + c)) * inv_n);
  } else {
    r = g * wc * is;
  }
  Vec4<T>::store1(dx, i, r);
}

// ---- activation + dropout (the ELU and Dropout after every GIN layer, gin.py:31-32) -------------------------------
template <typename T>
__global__ void __launch_bounds__(kThreads) act_dropout_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t total, int act,
                                                                   float drop_p, uint64_t seed, const int64_t* __restrict__ seed_dev) {
  const int64_t i = int64_t(blockIdx.x) * kThreads + threadIdx.x;
  if (i >= total) return;
  float v = act_fwd(ldx(x, i), act);
  if (drop_p > 0.f) v = uniform_hash64(launch_seed(seed, seed_dev), uint64_t(i)) < drop_p ? 0.f : v * (1.f / (1.f - drop_p));
  Vec4<T>::store1(y, i, v);
}
// dx = dy * keep * act'(x): x is the pre-activation input
template <typename T>
__global__ void __launch_bounds__(kThreads) act_dropout_bwd_kernel(const T* __restrict__ x, const T* __restrict__ dy, T* __restrict__ dx,
                                                                   int64_t total, int act, float drop_p, uint64_t seed,
                                                                   const int64_t* __restrict__ seed_dev) {
  const int64_t i = int64_t(blockIdx.x) * kThreads + threadIdx.x;
  if (i >= total) return;
  float g = ldx(dy, i) * act_grad(ldx(x, i), act);
  if (drop_p > 0.f) g = uniform_hash64(launch_seed(seed, seed_dev), uint64_t(i)) < drop_p ? 0.f : g * (1.f / (1.f - drop_p));
  Vec4<T>::store1(dx, i, g);
}


// ---- 4-wide forms (d % 4 == 0, 16-byte / 8-byte aligned rows): one thread = 4 consecutive columns of a row ------------
// (r02: the scalar kernels below moved 0.45-0.7 TB/s on config 3's [409 600, 64] activations -- one element per thread, a
// 64-bit modulo per element, 8 KB of loads in flight per SM in the statistics pass; these move 16 bytes per load and keep
// four row loads in flight per thread.)
__device__ __forceinline__ float4 f4_sub(float4 a, float4 b) { return make_float4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
__device__ __forceinline__ void f4_add(float4& a, float4 b) { a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }
__device__ __forceinline__ void f4_addsq(float4& a, float4 b) {
  a.x = fmaf(b.x, b.x, a.x); a.y = fmaf(b.y, b.y, a.y); a.z = fmaf(b.z, b.z, a.z); a.w = fmaf(b.w, b.w, a.w);
}
__device__ __forceinline__ float4 ldp4(const float* p, int c, float fill) {  // optional [d] parameter vector
  return p ? __ldg(reinterpret_cast<const float4*>(p + c)) : make_float4(fill, fill, fill, fill);
}
// dropout keep factor of element i: 0 or 1 / (1 - p)
__device__ __forceinline__ float keep_of(uint64_t sd, uint64_t i, float p, float inv_keep) {
  return uniform_hash64(sd, i) < p ? 0.f : inv_keep;
}

template <typename T>
__global__ void __launch_bounds__(kThreads) bn_stats_partial_vec_kernel(const T* __restrict__ x, float* __restrict__ partial,
                                                                        int64_t n, int d, int64_t slab_rows) {
  __shared__ float4 red[2][kThreads];
  const int w4 = d >> 2, lanes = kThreads / w4;
  const int c4 = threadIdx.x % w4, rl = threadIdx.x / w4;
  const int64_t r0 = int64_t(blockIdx.x) * slab_rows, r1 = min(n, r0 + slab_rows);
  float* p = partial + int64_t(blockIdx.x) * 3 * d;
  float4 s = f4_zero(), q = f4_zero();
  const float4 K = (r0 < r1 && rl < lanes) ? Vec4<T>::stream(x, r0 * d + 4 * c4) : f4_zero();
  if (rl < lanes) {
#pragma unroll 4
    for (int64_t r = r0 + rl; r < r1; r += lanes) {
      const float4 v = f4_sub(Vec4<T>::stream(x, r * d + 4 * c4), K);
      f4_add(s, v);
      f4_addsq(q, v);
    }
  }
  red[0][threadIdx.x] = s;
  red[1][threadIdx.x] = q;
  __syncthreads();
  if (rl == 0) {
    for (int l = 1; l < lanes; ++l) { f4_add(s, red[0][l * w4 + c4]); f4_add(q, red[1][l * w4 + c4]); }
    reinterpret_cast<float4*>(p)[c4] = s;
    reinterpret_cast<float4*>(p + d)[c4] = q;
    reinterpret_cast<float4*>(p + 2 * d)[c4] = K;
  }
}

template <typename T>
__global__ void __launch_bounds__(kThreads) bn_act_fwd_vec_kernel(const T* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                                                                  const float* __restrict__ mean, const float* __restrict__ invstd,
                                                                  T* __restrict__ y, int64_t total4, int d, int act, float drop_p,
                                                                  uint64_t seed, const int64_t* __restrict__ seed_dev) {
  const int64_t i4 = int64_t(blockIdx.x) * kThreads + threadIdx.x;
  if (i4 >= total4) return;
  const int64_t i = i4 * 4;
  const int c = int(i % d);
  const float4 is = __ldg(reinterpret_cast<const float4*>(invstd + c)), m = __ldg(reinterpret_cast<const float4*>(mean + c));
  const float4 wv = ldp4(w, c, 1.f), bv = ldp4(b, c, 0.f);
  const float4 v = Vec4<T>::stream(x, i);
  float4 r;
  r.x = act_fwd(fmaf(v.x, is.x * wv.x, bv.x - m.x * (is.x * wv.x)), act);
  r.y = act_fwd(fmaf(v.y, is.y * wv.y, bv.y - m.y * (is.y * wv.y)), act);
  r.z = act_fwd(fmaf(v.z, is.z * wv.z, bv.z - m.z * (is.z * wv.z)), act);
  r.w = act_fwd(fmaf(v.w, is.w * wv.w, bv.w - m.w * (is.w * wv.w)), act);
  if (drop_p > 0.f) {
    const uint64_t sd = launch_seed(seed, seed_dev);
    const float ik = 1.f / (1.f - drop_p);
    r.x *= keep_of(sd, uint64_t(i), drop_p, ik);
    r.y *= keep_of(sd, uint64_t(i + 1), drop_p, ik);
    r.z *= keep_of(sd, uint64_t(i + 2), drop_p, ik);
    r.w *= keep_of(sd, uint64_t(i + 3), drop_p, ik);
  }
  Vec4<T>::store(y, i, r);
}

// g = dy * keep * act'(z) for 4 consecutive elements; xh = normalised x
template <typename T>
__device__ __forceinline__ void bn_g4(const T* x, const T* dy, int64_t i, float4 m, float4 is, float4 wc, float4 bc, int act,
                                      float drop_p, uint64_t sd, float ik, float4& g, float4& xh) {
  const float4 xv = Vec4<T>::stream(x, i), dv = Vec4<T>::stream(dy, i);
  xh = make_float4((xv.x - m.x) * is.x, (xv.y - m.y) * is.y, (xv.z - m.z) * is.z, (xv.w - m.w) * is.w);
  g.x = dv.x * act_grad(fmaf(xh.x, wc.x, bc.x), act);
  g.y = dv.y * act_grad(fmaf(xh.y, wc.y, bc.y), act);
  g.z = dv.z * act_grad(fmaf(xh.z, wc.z, bc.z), act);
  g.w = dv.w * act_grad(fmaf(xh.w, wc.w, bc.w), act);
  if (drop_p > 0.f) {
    g.x *= keep_of(sd, uint64_t(i), drop_p, ik);
    g.y *= keep_of(sd, uint64_t(i + 1), drop_p, ik);
    g.z *= keep_of(sd, uint64_t(i + 2), drop_p, ik);
    g.w *= keep_of(sd, uint64_t(i + 3), drop_p, ik);
  }
}

template <typename T>
__global__ void __launch_bounds__(kThreads) bn_bwd_partial_vec_kernel(const T* __restrict__ x, const T* __restrict__ dy, const float* __restrict__ w,
                                                                      const float* __restrict__ b, const float* __restrict__ mean,
                                                                      const float* __restrict__ invstd, float* __restrict__ partial,
                                                                      int64_t n, int d, int64_t slab_rows, int act, float drop_p,
                                                                      uint64_t seed, const int64_t* __restrict__ seed_dev) {
  __shared__ float4 red[2][kThreads];
  const int w4 = d >> 2, lanes = kThreads / w4;
  const int c4 = threadIdx.x % w4, rl = threadIdx.x / w4;
  const int64_t r0 = int64_t(blockIdx.x) * slab_rows, r1 = min(n, r0 + slab_rows);
  float* p = partial + int64_t(blockIdx.x) * 2 * d;
  const uint64_t sd = drop_p > 0.f ? launch_seed(seed, seed_dev) : 0;
  const float ik = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
  float4 sg = f4_zero(), sgx = f4_zero();
  if (rl < lanes) {
    const int c = 4 * c4;
    const float4 m = __ldg(reinterpret_cast<const float4*>(mean + c)), is = __ldg(reinterpret_cast<const float4*>(invstd + c));
    const float4 wc = ldp4(w, c, 1.f), bc = ldp4(b, c, 0.f);
#pragma unroll 2
    for (int64_t r = r0 + rl; r < r1; r += lanes) {
      float4 g, xh;
      bn_g4(x, dy, r * d + c, m, is, wc, bc, act, drop_p, sd, ik, g, xh);
      f4_add(sg, g);
      sgx.x = fmaf(g.x, xh.x, sgx.x); sgx.y = fmaf(g.y, xh.y, sgx.y); sgx.z = fmaf(g.z, xh.z, sgx.z); sgx.w = fmaf(g.w, xh.w, sgx.w);
    }
  }
  red[0][threadIdx.x] = sg;
  red[1][threadIdx.x] = sgx;
  __syncthreads();
  if (rl == 0) {
    for (int l = 1; l < lanes; ++l) { f4_add(sg, red[0][l * w4 + c4]); f4_add(sgx, red[1][l * w4 + c4]); }
    reinterpret_cast<float4*>(p)[c4] = sg;          // -> d bias
    reinterpret_cast<float4*>(p + d)[c4] = sgx;     // -> d weight
  }
}

template <typename T>
__global__ void __launch_bounds__(kThreads) bn_bwd_dx_vec_kernel(const T* __restrict__ x, const T* __restrict__ dy, const float* __restrict__ w,
                                                                 const float* __restrict__ b, const float* __restrict__ mean,
                                                                 const float* __restrict__ invstd, const float* __restrict__ sums,
                                                                 T* __restrict__ dx, int64_t n, int d, int training, int act, float drop_p,
                                                                 uint64_t seed, const int64_t* __restrict__ seed_dev) {
  const int64_t i4 = int64_t(blockIdx.x) * kThreads + threadIdx.x;
  if (i4 * 4 >= n * d) return;
  const int64_t i = i4 * 4;
  const int c = int(i % d);
  const float4 m = __ldg(reinterpret_cast<const float4*>(mean + c)), is = __ldg(reinterpret_cast<const float4*>(invstd + c));
  const float4 wc = ldp4(w, c, 1.f), bc = ldp4(b, c, 0.f);
  const uint64_t sd = drop_p > 0.f ? launch_seed(seed, seed_dev) : 0;
  const float ik = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
  float4 g, xh, r;
  bn_g4(x, dy, i, m, is, wc, bc, act, drop_p, sd, ik, g, xh);
  if (training) {
    const float inv_n = 1.f / float(n);
    const float4 s0 = __ldg(reinterpret_cast<const float4*>(sums + c)), s1 = __ldg(reinterpret_cast<const float4*>(sums + d + c));
    r.x = wc.x * is.x * (g.x - (s0.x + xh.x * s1.x) * inv_n);
    r.y = wc.y * is.y * (g.y - (s0.y + xh.y * s1.y) * inv_n);
    r.z = wc.z * is.z * (g.z - (s0.z + xh.z * s1.z) * inv_n);
    r.w = wc.w * is.w * (g.w - (s0.w + xh.w * s1.w) * inv_n);
  } else {
    r = make_float4(g.x * wc.x * is.x, g.y * wc.y * is.y, g.z * wc.z * is.z, g.w * wc.w * is.w);
  }
  Vec4<T>::store(dx, i, r);
}

template <typename T>
__global__ void __launch_bounds__(kThreads) act_dropout_fwd_vec_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t total4, int act,
                                                                       float drop_p, uint64_t seed, const int64_t* __restrict__ seed_dev) {
  const int64_t i4 = int64_t(blockIdx.x) * kThreads + threadIdx.x;
  if (i4 >= total4) return;
  const int64_t i = i4 * 4;
  const float4 v = Vec4<T>::stream(x, i);
  float4 r = make_float4(act_fwd(v.x, act), act_fwd(v.y, act), act_fwd(v.z, act), act_fwd(v.w, act));
  if (drop_p > 0.f) {
    const uint64_t sd = launch_seed(seed, seed_dev);
    const float ik = 1.f / (1.f - drop_p);
    r.x *= keep_of(sd, uint64_t(i), drop_p, ik);
    r.y *= keep_of(sd, uint64_t(i + 1), drop_p, ik);
    r.z *= keep_of(sd, uint64_t(i + 2), drop_p, ik);
    r.w *= keep_of(sd, uint64_t(i + 3), drop_p, ik);
  }
  Vec4<T>::store(y, i, r);
}
template <typename T>
__global__ void __launch_bounds__(kThreads) act_dropout_bwd_vec_kernel(const T* __restrict__ x, const T* __restrict__ dy, T* __restrict__ dx,
                                                                       int64_t total4, int act, float drop_p, uint64_t seed,
                                                                       const int64_t* __restrict__ seed_dev) {
  const int64_t i4 = int64_t(blockIdx.x) * kThreads + threadIdx.x;
  if (i4 >= total4) return;
  const int64_t i = i4 * 4;
  const float4 v = Vec4<T>::stream(x, i), dv = Vec4<T>::stream(dy, i);
  float4 g = make_float4(dv.x * act_grad(v.x, act), dv.y * act_grad(v.y, act), dv.z * act_grad(v.z, act), dv.w * act_grad(v.w, act));
  if (drop_p > 0.f) {
    const uint64_t sd = launch_seed(seed, seed_dev);
    const float ik = 1.f / (1.f - drop_p);
    g.x *= keep_of(sd, uint64_t(i), drop_p, ik);
    g.y *= keep_of(sd, uint64_t(i + 1), drop_p, ik);
    g.z *= keep_of(sd, uint64_t(i + 2), drop_p, ik);
    g.w *= keep_of(sd, uint64_t(i + 3), drop_p, ik);
  }
  Vec4<T>::store(dx, i, g);
}

}  // namespace
}  // namespace lgk

using namespace lgk;

static bool all16(std::initializer_list<const void*> ps) {
  for (const void* q : ps)
    if (reinterpret_cast<uintptr_t>(q) & 15) return false;
  return true;
}

extern "C" int64_t lgk_bn_workspace_bytes(int64_t n, int64_t d) {
  if (n <= 0 || d <= 0) return 0;
  return (int64_t(norm_slabs(n)) * 3 * d + 2 * d) * 4;
}

extern "C" int lgk_bn_act_fwd(const void* x, const float* weight, const float* bias, float* running_mean, float* running_var,
                              int64_t* num_batches_tracked, void* y, float* save_mean, float* save_invstd, int64_t n, int64_t d,
                              float eps, float momentum, int training, int act, float dropout_p, uint64_t seed,
                              const int64_t* seed_dev, void* workspace, int64_t workspace_bytes, int dtype, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(n >= 0 && d > 0 && d < (1 << 24), LGK_ERR_BAD_ARG, "lgk_bn_act_fwd: bad sizes");
  LGK_REQUIRE(dtype == LGK_F32 || dtype == LGK_BF16, LGK_ERR_UNSUPPORTED, "lgk_bn_act_fwd: dtype must be LGK_F32 or LGK_BF16");
  const bool h = dtype == LGK_BF16;
  using bf16 = __nv_bfloat16;
  LGK_REQUIRE(dropout_p >= 0.f && dropout_p < 1.f, LGK_ERR_BAD_ARG, "lgk_bn_act_fwd: dropout must be in [0,1)");
  LGK_REQUIRE(act >= LGK_ACT_NONE && act <= LGK_ACT_RELU, LGK_ERR_BAD_ARG, "lgk_bn_act_fwd: bad activation");
  if (n == 0) return LGK_OK;
  LGK_REQUIRE(x && y && save_mean && save_invstd, LGK_ERR_BAD_ARG, "lgk_bn_act_fwd: null pointer");
  const bool vec = (d % 4 == 0) && d / 4 <= kThreads && all16({x, y, save_mean, save_invstd, weight, bias, workspace});
  if (training) {
    LGK_REQUIRE(n > 1, LGK_ERR_BAD_ARG, "lgk_bn_act_fwd: batch statistics need more than one row (as nn.BatchNorm1d)");
    LGK_REQUIRE(workspace && workspace_bytes >= lgk_bn_workspace_bytes(n, d), LGK_ERR_WORKSPACE, "lgk_bn_act_fwd: workspace too small");
    const int slabs = norm_slabs(n);
    const int64_t slab_rows = ceil_div(n, slabs);
    float* partial = static_cast<float*>(workspace);
    if (vec) {
      if (h) bn_stats_partial_vec_kernel<bf16><<<slabs, kThreads, 0, stream>>>(static_cast<const bf16*>(x), partial, n, int(d), slab_rows);
      else bn_stats_partial_vec_kernel<float><<<slabs, kThreads, 0, stream>>>(static_cast<const float*>(x), partial, n, int(d), slab_rows);
    } else {
      if (h) bn_stats_partial_kernel<bf16><<<slabs, kThreads, 0, stream>>>(static_cast<const bf16*>(x), partial, n, int(d), slab_rows);
      else bn_stats_partial_kernel<float><<<slabs, kThreads, 0, stream>>>(static_cast<const float*>(x), partial, n, int(d), slab_rows);
    }
    LGK_CHECK_LAUNCH("bn_stats_partial_kernel");
    bn_stats_finalize_kernel<<<unsigned(ceil_div(d, kFinCols)), kThreads, 0, stream>>>(
        partial, slabs, n, int(d), slab_rows, eps, momentum, save_mean, save_invstd, running_mean, running_var, num_batches_tracked);
    LGK_CHECK_LAUNCH("bn_stats_finalize_kernel");
  } else {
    LGK_REQUIRE(running_mean && running_var, LGK_ERR_BAD_ARG, "lgk_bn_act_fwd: eval mode needs the running statistics");
    bn_eval_stats_kernel<<<unsigned(ceil_div(d, kThreads)), kThreads, 0, stream>>>(running_mean, running_var, eps, int(d), save_mean, save_invstd);
    LGK_CHECK_LAUNCH("bn_eval_stats_kernel");
  }
  if (vec) {
    const int64_t t4 = n * d / 4;
    const unsigned grid = unsigned(ceil_div(t4, kThreads));
    if (h) bn_act_fwd_vec_kernel<bf16><<<grid, kThreads, 0, stream>>>(static_cast<const bf16*>(x), weight, bias, save_mean, save_invstd,
                                                                     static_cast<bf16*>(y), t4, int(d), act, dropout_p, seed, seed_dev);
    else bn_act_fwd_vec_kernel<float><<<grid, kThreads, 0, stream>>>(static_cast<const float*>(x), weight, bias, save_mean, save_invstd,
                                                                    static_cast<float*>(y), t4, int(d), act, dropout_p, seed, seed_dev);
    LGK_CHECK_LAUNCH("bn_act_fwd_vec_kernel");
    return LGK_OK;
  }
  const unsigned grid = unsigned(ceil_div(n * d, kThreads));
  if (h) bn_act_fwd_kernel<bf16><<<grid, kThreads, 0, stream>>>(static_cast<const bf16*>(x), weight, bias, save_mean, save_invstd,
                                                               static_cast<bf16*>(y), n * d, int(d), act, dropout_p, seed, seed_dev);
  else bn_act_fwd_kernel<float><<<grid, kThreads, 0, stream>>>(static_cast<const float*>(x), weight, bias, save_mean, save_invstd,
                                                              static_cast<float*>(y), n * d, int(d), act, dropout_p, seed, seed_dev);
  LGK_CHECK_LAUNCH("bn_act_fwd_kernel");
  return LGK_OK;
}

extern "C" int lgk_bn_act_bwd(const void* x, const void* dy, const float* weight, const float* bias, const float* save_mean,
                              const float* save_invstd, void* dx, float* dweight, float* dbias, int accumulate, int64_t n,
                              int64_t d, int training, int act, float dropout_p, uint64_t seed, const int64_t* seed_dev,
                              void* workspace, int64_t workspace_bytes, int dtype, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(n >= 0 && d > 0 && d < (1 << 24), LGK_ERR_BAD_ARG, "lgk_bn_act_bwd: bad sizes");
  LGK_REQUIRE(dtype == LGK_F32 || dtype == LGK_BF16, LGK_ERR_UNSUPPORTED, "lgk_bn_act_bwd: dtype must be LGK_F32 or LGK_BF16");
  const bool h = dtype == LGK_BF16;
  using bf16 = __nv_bfloat16;
  if (n == 0) return LGK_OK;
  LGK_REQUIRE(x && dy && save_mean && save_invstd && dx, LGK_ERR_BAD_ARG, "lgk_bn_act_bwd: null pointer");
  LGK_REQUIRE(workspace && workspace_bytes >= lgk_bn_workspace_bytes(n, d), LGK_ERR_WORKSPACE, "lgk_bn_act_bwd: workspace too small");
  const int slabs = norm_slabs(n);
  const int64_t slab_rows = ceil_div(n, slabs);
  float* partial = static_cast<float*>(workspace);
  float* sums = partial + int64_t(slabs) * 3 * d;  // [2 d]: sum g, sum g xhat
  const bool vec = (d % 4 == 0) && d / 4 <= kThreads && all16({x, dy, dx, save_mean, save_invstd, weight, bias, workspace});
  if (vec) {
    if (h) bn_bwd_partial_vec_kernel<bf16><<<slabs, kThreads, 0, stream>>>(static_cast<const bf16*>(x), static_cast<const bf16*>(dy), weight,
                                                                          bias, save_mean, save_invstd, partial, n, int(d), slab_rows, act,
                                                                          dropout_p, seed, seed_dev);
    else bn_bwd_partial_vec_kernel<float><<<slabs, kThreads, 0, stream>>>(static_cast<const float*>(x), static_cast<const float*>(dy), weight,
                                                                         bias, save_mean, save_invstd, partial, n, int(d), slab_rows, act,
                                                                         dropout_p, seed, seed_dev);
  } else if (h) bn_bwd_partial_kernel<bf16><<<slabs, kThreads, 0, stream>>>(static_cast<const bf16*>(x), static_cast<const bf16*>(dy), weight, bias,
                                                                    save_mean, save_invstd, partial, n, int(d), slab_rows, act, dropout_p, seed,
                                                                    seed_dev);
  else bn_bwd_partial_kernel<float><<<slabs, kThreads, 0, stream>>>(static_cast<const float*>(x), static_cast<const float*>(dy), weight, bias,
                                                                   save_mean, save_invstd, partial, n, int(d), slab_rows, act, dropout_p, seed,
                                                                   seed_dev);
  LGK_CHECK_LAUNCH("bn_bwd_partial_kernel");
  slab_reduce_kernel<<<slab_reduce_grid(2 * d), 256, 0, stream>>>(partial, sums, sums + d, d, d, slabs, 0);
  LGK_CHECK_LAUNCH("slab_reduce_kernel(bn)");
  if (dweight || dbias) {  // parameter gradients: (added to) the caller's buffers, same fixed order
    slab_reduce_kernel<<<slab_reduce_grid(2 * d), 256, 0, stream>>>(partial, dbias, dweight, d, d, slabs, accumulate);
    LGK_CHECK_LAUNCH("slab_reduce_kernel(bn params)");
  }
  if (vec) {
    const unsigned grid4 = unsigned(ceil_div(n * d / 4, kThreads));
    if (h) bn_bwd_dx_vec_kernel<bf16><<<grid4, kThreads, 0, stream>>>(static_cast<const bf16*>(x), static_cast<const bf16*>(dy), weight, bias,
                                                                     save_mean, save_invstd, sums, static_cast<bf16*>(dx), n, int(d), training,
                                                                     act, dropout_p, seed, seed_dev);
    else bn_bwd_dx_vec_kernel<float><<<grid4, kThreads, 0, stream>>>(static_cast<const float*>(x), static_cast<const float*>(dy), weight, bias,
                                                                    save_mean, save_invstd, sums, static_cast<float*>(dx), n, int(d), training,
                                                                    act, dropout_p, seed, seed_dev);
    LGK_CHECK_LAUNCH("bn_bwd_dx_vec_kernel");
    return LGK_OK;
  }
  const unsigned grid = unsigned(ceil_div(n * d, kThreads));
  if (h) bn_bwd_dx_kernel<bf16><<<grid, kThreads, 0, stream>>>(static_cast<const bf16*>(x), static_cast<const bf16*>(dy), weight, bias, save_mean,
                                                              save_invstd, sums, static_cast<bf16*>(dx), n, int(d), training, act, dropout_p,
                                                              seed, seed_dev);
  else bn_bwd_dx_kernel<float><<<grid, kThreads, 0, stream>>>(static_cast<const float*>(x), static_cast<const float*>(dy), weight, bias, save_mean,
                                                             save_invstd, sums, static_cast<float*>(dx), n, int(d), training, act, dropout_p,
                                                             seed, seed_dev);
  LGK_CHECK_LAUNCH("bn_bwd_dx_kernel");
  return LGK_OK;
}

extern "C" int lgk_act_dropout_fwd(const void* x, void* y, int64_t n, int act, float dropout_p, uint64_t seed,
                                   const int64_t* seed_dev, int dtype, void* stream_) {
  using bf16 = __nv_bfloat16;
  LGK_REQUIRE(dtype == LGK_F32 || dtype == LGK_BF16, LGK_ERR_UNSUPPORTED, "lgk_act_dropout_fwd: dtype must be LGK_F32 or LGK_BF16");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(n >= 0 && dropout_p >= 0.f && dropout_p < 1.f, LGK_ERR_BAD_ARG, "lgk_act_dropout_fwd: bad arguments");
  if (n == 0) return LGK_OK;
  LGK_REQUIRE(x && y, LGK_ERR_BAD_ARG, "lgk_act_dropout_fwd: null pointer");
  if (n % 4 == 0 && all16({x, y})) {
    const unsigned grid4 = unsigned(ceil_div(n / 4, kThreads));
    if (dtype == LGK_BF16) act_dropout_fwd_vec_kernel<bf16><<<grid4, kThreads, 0, stream>>>(static_cast<const bf16*>(x), static_cast<bf16*>(y),
                                                                                           n / 4, act, dropout_p, seed, seed_dev);
    else act_dropout_fwd_vec_kernel<float><<<grid4, kThreads, 0, stream>>>(static_cast<const float*>(x), static_cast<float*>(y), n / 4, act,
                                                                          dropout_p, seed, seed_dev);
    LGK_CHECK_LAUNCH("act_dropout_fwd_vec_kernel");
    return LGK_OK;
  }
  const unsigned grid = unsigned(ceil_div(n, kThreads));
  if (dtype == LGK_BF16) act_dropout_fwd_kernel<bf16><<<grid, kThreads, 0, stream>>>(static_cast<const bf16*>(x), static_cast<bf16*>(y), n, act,
                                                                                    dropout_p, seed, seed_dev);
  else act_dropout_fwd_kernel<float><<<grid, kThreads, 0, stream>>>(static_cast<const float*>(x), static_cast<float*>(y), n, act, dropout_p,
                                                                   seed, seed_dev);
  LGK_CHECK_LAUNCH("act_dropout_fwd_kernel");
  return LGK_OK;
}

extern "C" int lgk_act_dropout_bwd(const void* x, const void* dy, void* dx, int64_t n, int act, float dropout_p, uint64_t seed,
                                   const int64_t* seed_dev, int dtype, void* stream_) {
  using bf16 = __nv_bfloat16;
  LGK_REQUIRE(dtype == LGK_F32 || dtype == LGK_BF16, LGK_ERR_UNSUPPORTED, "lgk_act_dropout_bwd: dtype must be LGK_F32 or LGK_BF16");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(n >= 0 && dropout_p >= 0.f && dropout_p < 1.f, LGK_ERR_BAD_ARG, "lgk_act_dropout_bwd: bad arguments");
  if (n == 0) return LGK_OK;
  LGK_REQUIRE(x && dy && dx, LGK_ERR_BAD_ARG, "lgk_act_dropout_bwd: null pointer");
  if (n % 4 == 0 && all16({x, dy, dx})) {
    const unsigned grid4 = unsigned(ceil_div(n / 4, kThreads));
    if (dtype == LGK_BF16) act_dropout_bwd_vec_kernel<bf16><<<grid4, kThreads, 0, stream>>>(static_cast<const bf16*>(x), static_cast<const bf16*>(dy),
                                                                                           static_cast<bf16*>(dx), n / 4, act, dropout_p, seed,
                                                                                           seed_dev);
    else act_dropout_bwd_vec_kernel<float><<<grid4, kThreads, 0, stream>>>(static_cast<const float*>(x), static_cast<const float*>(dy),
                                                                          static_cast<float*>(dx), n / 4, act, dropout_p, seed, seed_dev);
    LGK_CHECK_LAUNCH("act_dropout_bwd_vec_kernel");
    return LGK_OK;
  }
  const unsigned grid = unsigned(ceil_div(n, kThreads));
  if (dtype == LGK_BF16) act_dropout_bwd_kernel<bf16><<<grid, kThreads, 0, stream>>>(static_cast<const bf16*>(x), static_cast<const bf16*>(dy),
                                                                                    static_cast<bf16*>(dx), n, act, dropout_p, seed, seed_dev);
  else act_dropout_bwd_kernel<float><<<grid, kThreads, 0, stream>>>(static_cast<const float*>(x), static_cast<const float*>(dy),
                                                                   static_cast<float*>(dx), n, act, dropout_p, seed, seed_dev);
  LGK_CHECK_LAUNCH("act_dropout_bwd_kernel");
  return LGK_OK;
}
