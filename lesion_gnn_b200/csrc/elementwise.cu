// Elementwise pieces of the training step: ELU (F.elu after every conv, reference models/gat.py:51, gin.py:31,
// drgnet.py:55), the fused multi-tensor Adam/AdamW step over one flat buffer (torch.optim.Adam at
// models/base.py:162-170), and the library-global error / launch bookkeeping.
#include "common.cuh"
#include <atomic>

namespace lgk {

static thread_local char g_error[512] = "";
// process-wide: autograd runs the backward launches on its own thread
static std::atomic<int64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

namespace {

template <typename T>
__global__ void __launch_bounds__(256) elu_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t n4, int64_t n) {
  const int64_t t = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (t < n4) {
    float4 v = Vec4<T>::stream(x, t * 4);
    v.x = elu1(v.x); v.y = elu1(v.y); v.z = elu1(v.z); v.w = elu1(v.w);
    Vec4<T>::store(y, t * 4, v);
  } else if (t == n4) {
    for (int64_t i = n4 * 4; i < n; ++i) Vec4<T>::store1(y, i, elu1(Vec4<T>::load1(x, i)));
  }
}
// dx = dy * (y > 0 ? 1 : y + 1)   (alpha = 1: elu'(x) = exp(x) = y + 1 on the negative side)
template <typename T>
__global__ void __launch_bounds__(256) elu_bwd_kernel(const T* __restrict__ y, const T* __restrict__ dy, T* __restrict__ dx,
                                                      int64_t n4, int64_t n) {
  const int64_t t = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (t < n4) {
    const float4 yv = Vec4<T>::stream(y, t * 4), g = Vec4<T>::stream(dy, t * 4);
    float4 r;
    r.x = g.x * (yv.x > 0.f ? 1.f : yv.x + 1.f);
    r.y = g.y * (yv.y > 0.f ? 1.f : yv.y + 1.f);
    r.z = g.z * (yv.z > 0.f ? 1.f : yv.z + 1.f);
    r.w = g.w * (yv.w > 0.f ? 1.f : yv.w + 1.f);
    Vec4<T>::store(dx, t * 4, r);
  } else if (t == n4) {
    for (int64_t i = n4 * 4; i < n; ++i) {
      const float yv = Vec4<T>::load1(y, i);
      Vec4<T>::store1(dx, i, Vec4<T>::load1(dy, i) * (yv > 0.f ? 1.f : yv + 1.f));
    }
  }
}

__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m,
                                                   float* __restrict__ v, int64_t n, float lr, float b1, float b2, float eps,
                                                   float wd, float bc1, float bc2_sqrt, int64_t* __restrict__ step_dev,
                                                   float gscale, int adamw, int zero_grad, int advance) {
  const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (step_dev) {  // graph-replay mode: bias corrections from the device-resident step counter
    // advance: the counter holds the steps COMPLETED so far and this launch moves it on -- by the CTA that finishes
    // last (step_dev[1] counts finished CTAs), i.e. after every CTA has read it: no separate increment launch
    const int64_t steps = *reinterpret_cast<volatile int64_t*>(step_dev) + (advance ? 1 : 0);
    const double t = double(steps);
    bc1 = float(1.0 - pow(double(b1), t));
    bc2_sqrt = float(sqrt(1.0 - pow(double(b2), t)));
    if (advance) {
      __syncthreads();  // every thread of this CTA has read the counter
      if (threadIdx.x == 0) {
        __threadfence();
        unsigned long long* done = reinterpret_cast<unsigned long long*>(step_dev + 1);
        if (atomicAdd(done, 1ull) == gridDim.x - 1) {
          *done = 0ull;
          step_dev[0] = steps;
        }
      }
    }
  }
  if (i >= n) return;
  float param = p[i];
  float grad = g[i] * gscale;
  if (zero_grad) g[i] = 0.f;  // the bucket is ready for the next backward: no separate fill launch
  if (adamw) param *= 1.f - lr * wd;        // decoupled decay (torch.optim.AdamW)
  else grad = fmaf(wd, param, grad);        // L2 penalty folded into the gradient (torch.optim.Adam)
  const float mi = fmaf(b1, m[i], (1.f - b1) * grad);
  const float vi = fmaf(b2, v[i], (1.f - b2) * grad * grad);
  m[i] = mi;
  v[i] = vi;
  const float denom = sqrtf(vi) / bc2_sqrt + eps;
  p[i] = param - (lr / bc1) * (mi / denom);
}

}  // namespace
}  // namespace lgk

using namespace lgk;

extern "C" const char* lgk_last_error(void) { return g_error; }
extern "C" int lgk_version(void) { return 100; }
extern "C" int64_t lgk_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
extern "C" void lgk_reset_launch_count(void) { g_launches.store(0, std::memory_order_relaxed); }

static bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

extern "C" int lgk_elu_fwd(const void* x, void* y, int64_t n, int dtype, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(n >= 0, LGK_ERR_BAD_ARG, "lgk_elu_fwd: negative size");
  if (n == 0) return LGK_OK;
  LGK_REQUIRE(x && y, LGK_ERR_BAD_ARG, "lgk_elu_fwd: null pointer");
  const int64_t n4 = (al16(x) && al16(y)) ? n / 4 : 0;
  const unsigned grid = unsigned(ceil_div(n4 + 1, 256));
  if (dtype == LGK_F32) elu_fwd_kernel<float><<<grid, 256, 0, stream>>>(static_cast<const float*>(x), static_cast<float*>(y), n4, n);
  else if (dtype == LGK_BF16) elu_fwd_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>(static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), n4, n);
  else LGK_REQUIRE(false, LGK_ERR_BAD_ARG, "lgk_elu_fwd: bad dtype");
  LGK_CHECK_LAUNCH("elu_fwd_kernel");
  return LGK_OK;
}

extern "C" int lgk_elu_bwd(const void* y, const void* dy, void* dx, int64_t n, int dtype, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(n >= 0, LGK_ERR_BAD_ARG, "lgk_elu_bwd: negative size");
  if (n == 0) return LGK_OK;
  LGK_REQUIRE(y && dy && dx, LGK_ERR_BAD_ARG, "lgk_elu_bwd: null pointer");
  const int64_t n4 = (al16(y) && al16(dy) && al16(dx)) ? n / 4 : 0;
  const unsigned grid = unsigned(ceil_div(n4 + 1, 256));
  if (dtype == LGK_F32) elu_bwd_kernel<float><<<grid, 256, 0, stream>>>(static_cast<const float*>(y), static_cast<const float*>(dy), static_cast<float*>(dx), n4, n);
  else if (dtype == LGK_BF16) elu_bwd_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>(static_cast<const __nv_bfloat16*>(y), static_cast<const __nv_bfloat16*>(dy), static_cast<__nv_bfloat16*>(dx), n4, n);
  else LGK_REQUIRE(false, LGK_ERR_BAD_ARG, "lgk_elu_bwd: bad dtype");
  LGK_CHECK_LAUNCH("elu_bwd_kernel");
  return LGK_OK;
}

extern "C" int lgk_adam_step(float* param, float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, float lr,
                             float beta1, float beta2, float eps, float weight_decay, int64_t step, int64_t* step_dev,
                             float grad_scale, int adamw, int zero_grad, int advance_step, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(n >= 0 && (step >= 1 || step_dev), LGK_ERR_BAD_ARG, "lgk_adam_step: bad size or step");
  if (step < 1) step = 1;
  if (n == 0) return LGK_OK;
  LGK_REQUIRE(param && grad && exp_avg && exp_avg_sq, LGK_ERR_BAD_ARG, "lgk_adam_step: null pointer");
  const double bc1 = 1.0 - pow(double(beta1), double(step));
  const double bc2 = 1.0 - pow(double(beta2), double(step));
  adam_kernel<<<unsigned(ceil_div(n, 256)), 256, 0, stream>>>(param, grad, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps,
                                                              weight_decay, float(bc1), float(sqrt(bc2)), step_dev, grad_scale, adamw, zero_grad, advance_step);
  LGK_CHECK_LAUNCH("adam_kernel");
  return LGK_OK;
}

// ---- clamped regression loss, forward + gradient ---------------------------------------------------------------
// One CTA of 1024 threads: a batch has a few hundred graphs.  Thread t adds its elements in index order, the block
// adds the thread sums in a fixed tree: deterministic.  Replaces a chain of ~12 elementwise torch launches (cast,
// clamp, mse, their backward with two compares, a logical_and and a where) between the forward and the backward.
namespace lgk {
namespace {
__global__ void __launch_bounds__(1024) clamped_loss_kernel(const float* __restrict__ logits, const int64_t* __restrict__ y,
                                                            int64_t n, float lo, float hi, int kind, float grad_scale,
                                                            float* __restrict__ loss_out, float* __restrict__ dlogits) {
  __shared__ float red[1024];
  const float inv_n = 1.f / float(n);
  float s = 0.f;
  for (int64_t i = threadIdx.x; i < n; i += 1024) {
    const float x = logits[i];
    const float c = fminf(fmaxf(x, lo), hi);
    const float d = c - float(y[i]);
    float l, g;
    if (kind == 0) { l = d * d; g = 2.f * d; }
    else if (fabsf(d) < 1.f) { l = 0.5f * d * d; g = d; }
    else { l = fabsf(d) - 0.5f; g = d > 0.f ? 1.f : -1.f; }
    s += l;
    if (dlogits) dlogits[i] = (x >= lo && x <= hi) ? g * inv_n * grad_scale : 0.f;
  }
  red[threadIdx.x] = s;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) loss_out[0] = red[0] * inv_n;
}
}  // namespace
}  // namespace lgk

extern "C" int lgk_clamped_regression_loss(const float* logits, const int64_t* y, int64_t n, float lo, float hi, int kind,
                                           float grad_scale, float* loss_out, float* dlogits, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(n > 0, LGK_ERR_BAD_ARG, "lgk_clamped_regression_loss: empty batch");
  LGK_REQUIRE(logits && y && loss_out, LGK_ERR_BAD_ARG, "lgk_clamped_regression_loss: null pointer");
  LGK_REQUIRE(kind == 0 || kind == 1, LGK_ERR_BAD_ARG, "lgk_clamped_regression_loss: kind must be 0 (MSE) or 1 (SmoothL1)");
  lgk::clamped_loss_kernel<<<1, 1024, 0, stream>>>(logits, y, n, lo, hi, kind, grad_scale, loss_out, dlogits);
  LGK_CHECK_LAUNCH("clamped_loss_kernel");
  return LGK_OK;
}

// ---- SGD (torch.optim.SGD; reference models/base.py:167-168 uses lr and weight_decay only) ---------------------------
namespace lgk {
namespace {
__global__ void __launch_bounds__(256) sgd_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ buf, int64_t n,
                                                  float lr, float momentum, float wd, int nesterov, float gscale, int zero_grad) {
  const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float param = p[i];
  float grad = fmaf(wd, param, g[i] * gscale);  // L2 penalty folded into the gradient, as torch.optim.SGD
  if (zero_grad) g[i] = 0.f;
  if (buf) {  // momentum buffer, zero-initialised: b = mu b + g reproduces torch's first step (b = g) with dampening 0
    const float b = fmaf(momentum, buf[i], grad);
    buf[i] = b;
    grad = nesterov ? fmaf(momentum, b, grad) : b;
  }
  p[i] = param - lr * grad;
}

// ---- (weighted) cross-entropy, forward + gradient (nn.CrossEntropyLoss(weight=class_weights), models/base.py:93-94) --
// One CTA: thread per graph (strided).  loss = sum_i w[y_i] nll_i / sum_i w[y_i]; fixed reduction tree: deterministic.
__global__ void __launch_bounds__(1024) cross_entropy_kernel(const float* __restrict__ logits, const int64_t* __restrict__ y,
                                                             const float* __restrict__ weight, int64_t n, int C, float grad_scale,
                                                             float* __restrict__ loss_out, float* __restrict__ dlogits) {
  __shared__ float red_l[1024], red_w[1024];
  float sl = 0.f, sw = 0.f;
  for (int64_t i = threadIdx.x; i < n; i += 1024) {
    const float* row = logits + i * C;
    float m = -INFINITY;
    for (int c = 0; c < C; ++c) m = fmaxf(m, row[c]);
    float s = 0.f;
    for (int c = 0; c < C; ++c) s += expf(row[c] - m);
    const int64_t t = y[i];
    const bool ok = t >= 0 && t < C;  // ignore_index-style targets contribute nothing
    const float w = ok ? (weight ? weight[t] : 1.f) : 0.f;
    if (ok) sl += w * (logf(s) + m - row[t]);
    sw += w;
  }
  red_l[threadIdx.x] = sl;
  red_w[threadIdx.x] = sw;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if (threadIdx.x < o) { red_l[threadIdx.x] += red_l[threadIdx.x + o]; red_w[threadIdx.x] += red_w[threadIdx.x + o]; }
    __syncthreads();
  }
  const float wsum = red_w[0];
  if (threadIdx.x == 0) loss_out[0] = red_l[0] / wsum;
  if (!dlogits) return;
  const float k = grad_scale / wsum;
  for (int64_t i = threadIdx.x; i < n; i += 1024) {
    const float* row = logits + i * C;
    float m = -INFINITY;
    for (int c = 0; c < C; ++c) m = fmaxf(m, row[c]);
    float s = 0.f;
    for (int c = 0; c < C; ++c) s += expf(row[c] - m);
    const int64_t t = y[i];
    const bool ok = t >= 0 && t < C;
    const float w = ok ? (weight ? weight[t] : 1.f) * k : 0.f;
    const float inv = 1.f / s;
    for (int c = 0; c < C; ++c) dlogits[i * C + c] = w * (expf(row[c] - m) * inv - (c == t ? 1.f : 0.f));
  }
}
}  // namespace
}  // namespace lgk

extern "C" int lgk_sgd_step(float* param, float* grad, float* momentum_buf, int64_t n, float lr, float momentum,
                            float weight_decay, int nesterov, float grad_scale, int zero_grad, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(n >= 0, LGK_ERR_BAD_ARG, "lgk_sgd_step: negative size");
  if (n == 0) return LGK_OK;
  LGK_REQUIRE(param && grad, LGK_ERR_BAD_ARG, "lgk_sgd_step: null pointer");
  LGK_REQUIRE(momentum == 0.f || momentum_buf, LGK_ERR_BAD_ARG, "lgk_sgd_step: momentum needs its buffer");
  LGK_REQUIRE(!nesterov || momentum > 0.f, LGK_ERR_BAD_ARG, "lgk_sgd_step: nesterov needs momentum");
  lgk::sgd_kernel<<<unsigned(ceil_div(n, 256)), 256, 0, stream>>>(param, grad, momentum == 0.f ? nullptr : momentum_buf, n, lr,
                                                                  momentum, weight_decay, nesterov, grad_scale, zero_grad);
  LGK_CHECK_LAUNCH("sgd_kernel");
  return LGK_OK;
}

extern "C" int lgk_cross_entropy_loss(const float* logits, const int64_t* y, const float* class_weight, int64_t n,
                                      int64_t num_classes, float grad_scale, float* loss_out, float* dlogits, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(n > 0 && num_classes > 0 && num_classes < (1 << 20), LGK_ERR_BAD_ARG, "lgk_cross_entropy_loss: bad sizes");
  LGK_REQUIRE(logits && y && loss_out, LGK_ERR_BAD_ARG, "lgk_cross_entropy_loss: null pointer");
  lgk::cross_entropy_kernel<<<1, 1024, 0, stream>>>(logits, y, class_weight, n, int(num_classes), grad_scale, loss_out, dlogits);
  LGK_CHECK_LAUNCH("cross_entropy_kernel");
  return LGK_OK;
}

// ---- row gather: out[i, :] = src[index[i], :] (collation of a device-resident dataset by node index: what
// Batch.from_data_list does on the host in the reference's DataLoader, datasets/datamodule.py:64-69) --------------------
namespace lgk {
namespace {
__global__ void __launch_bounds__(256) gather_rows_kernel(const uint32_t* __restrict__ src, const int64_t* __restrict__ index,
                                                          uint32_t* __restrict__ out, int64_t n, int64_t row_words, int64_t src_rows) {
  const int lane = threadIdx.x & 31;
  const int64_t i = int64_t(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (i >= n) return;
  const int64_t j = index[i];
  if (uint64_t(j) >= uint64_t(src_rows)) return;  // out of range: the row is left untouched
  const uint32_t* s = src + j * row_words;
  uint32_t* d = out + i * row_words;
  if ((row_words & 3) == 0 && ((reinterpret_cast<uintptr_t>(s) | reinterpret_cast<uintptr_t>(d)) & 15) == 0) {
    for (int64_t w = lane; w < row_words / 4; w += 32) reinterpret_cast<uint4*>(d)[w] = __ldg(reinterpret_cast<const uint4*>(s) + w);
  } else {
    for (int64_t w = lane; w < row_words; w += 32) d[w] = __ldg(s + w);
  }
}
}  // namespace
}  // namespace lgk

extern "C" int lgk_gather_rows(const void* src, const int64_t* index, void* out, int64_t n, int64_t row_bytes,
                               int64_t src_rows, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(n >= 0 && row_bytes >= 0 && src_rows >= 0, LGK_ERR_BAD_ARG, "lgk_gather_rows: negative size");
  if (n == 0 || row_bytes == 0) return LGK_OK;
  LGK_REQUIRE(src && index && out, LGK_ERR_BAD_ARG, "lgk_gather_rows: null pointer");
  LGK_REQUIRE(row_bytes % 4 == 0, LGK_ERR_UNSUPPORTED, "lgk_gather_rows: rows must be a multiple of 4 bytes");
  lgk::gather_rows_kernel<<<unsigned(ceil_div(n, 8)), 256, 0, stream>>>(static_cast<const uint32_t*>(src), index,
                                                                        static_cast<uint32_t*>(out), n, row_bytes / 4, src_rows);
  LGK_CHECK_LAUNCH("gather_rows_kernel");
  return LGK_OK;
}
