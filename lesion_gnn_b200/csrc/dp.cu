// Data-parallel gradient all-reduce FUSED with the optimiser step, over NVLink peer memory.
//
// The gradient bucket of a lesion-gnn model is 40 KB .. 4 MB: an NCCL all-reduce of that size is pure latency (18-37 us
// measured on 2 B200s, plus the eager launches around it) and the Adam launch waits behind it.  Here every rank owns a
// SYMMETRIC allocation (torch symmetric memory: the same buffer mapped into every peer over NVLink / NVSwitch) and ONE
// kernel per rank, captured in the step's CUDA graph, does the whole data-parallel tail:
//   push    CTA c copies chunk c of the own gradient bucket into slot [epoch parity][own rank] of EVERY peer's buffer
//           (posted 16-byte stores: one-way latency, nobody waits for a read round trip), then raises flag [own rank][c]
//           in every peer's pad (fence.sys + st.release.sys) -- chunk granularity, no grid-wide barrier;
//   wait    CTA c waits until every peer's flag [peer][c] has reached this epoch;
//   reduce  chunk c of all ranks' gradients is summed in RANK ORDER from the local slots (the own bucket stands in for the
//           own rank): every rank forms bit-identical sums, so the replicas stay bit-identical; times 1/world, Adam /
//           AdamW on the own parameters in the same pass, own bucket cleared for the next backward.
// The receive slots are double-buffered by epoch parity, which is what makes a departure barrier unnecessary: a peer can
// write parity p again only at epoch e + 2, i.e. after it has finished epoch e + 1, whose wait needed this rank's epoch
// e + 1 flags, which this rank raises only after its epoch-e kernel (the reader of parity p) has completed.
// Layout of the symmetric buffer (bytes): [0, 8192) flags uint64 [16 ranks][64 CTAs]; [8192, 8192 + 4 n) the fp32 gradient
// bucket; then the receive slots float [2][world][n].  epoch_dev = {epochs completed, CTA counter} is a counter of its own
// because the optimiser's step count is rewound after graph-capture warm-ups while flags and slots are not.
#include "common.cuh"
#include <math.h>

namespace lgk {
namespace {

constexpr int kMaxRanks = 16;
constexpr int kMaxCtas = 64;
constexpr int kPadBytes = kMaxRanks * kMaxCtas * 8;

__device__ __forceinline__ void st_release_sys(uint64_t* p, uint64_t v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_acquire_sys(const uint64_t* p) {
  uint64_t v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
// slots are written by the peers: system-scope loads (never served from a stale non-coherent line of two epochs ago)
__device__ __forceinline__ float4 ld_relaxed_sys_f4(const float4* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}

__global__ void __launch_bounds__(256) dp_allreduce_adam_kernel(const uint64_t* const* __restrict__ peer_base, int rank, int world,
                                                                int64_t n, float* __restrict__ p, float* __restrict__ m,
                                                                float* __restrict__ v, float lr, float b1, float b2, float eps,
                                                                float wd, int64_t* __restrict__ step_dev,
                                                                int64_t* __restrict__ epoch_dev, int adamw, int advance) {
  const uint64_t epoch = uint64_t(*reinterpret_cast<volatile int64_t*>(epoch_dev)) + 1;
  const int64_t n4 = n / 4;
  const int64_t per = (n4 + gridDim.x - 1) / gridDim.x;
  const int64_t lo = int64_t(blockIdx.x) * per, hi = min(lo + per, n4);
  char* own_base = reinterpret_cast<char*>(const_cast<uint64_t*>(peer_base[rank]));
  float4* own = reinterpret_cast<float4*>(own_base + kPadBytes);
  const int64_t slot_f4 = (int64_t(epoch & 1) * world) * n4;  // first slot of this parity, in float4 from the slot area

  // push: own chunk -> slot [parity][rank] of every peer
  for (int64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
    const float4 g = own[i];
    for (int r = 0; r < world; ++r) {
      if (r == rank) continue;
      float4* slots = reinterpret_cast<float4*>(reinterpret_cast<char*>(const_cast<uint64_t*>(peer_base[r])) + kPadBytes) + n4;
      slots[slot_f4 + int64_t(rank) * n4 + i] = g;
    }
  }
  __syncthreads();
  if (threadIdx.x < world && int(threadIdx.x) != rank) {
    __threadfence_system();  // the CTA's stores (ordered before this thread by the barrier) before the flag
    uint64_t* flags = const_cast<uint64_t*>(peer_base[threadIdx.x]);
    st_release_sys(flags + rank * kMaxCtas + blockIdx.x, epoch);
    // wait: the same chunk of that peer has landed here.  Bounded: a missing peer is a launch failure, not a hung GPU
    const uint64_t* mine = reinterpret_cast<const uint64_t*>(own_base) + threadIdx.x * kMaxCtas + blockIdx.x;
    unsigned spins = 0;
    while (ld_acquire_sys(mine) < epoch) {
      __nanosleep(100);
      if (++spins > (1u << 27)) __trap();  // ~15 s
    }
  }
  __syncthreads();

  // advance: step_dev[0] holds the steps COMPLETED and this launch moves it on (no separate increment launch)
  const int64_t step = *reinterpret_cast<volatile int64_t*>(step_dev) + (advance ? 1 : 0);
  const double t = double(step);
  const float bc1 = float(1.0 - pow(double(b1), t));
  const float bc2_sqrt = float(sqrt(1.0 - pow(double(b2), t)));
  const float gscale = 1.f / float(world);
  const float4* slots = own + n4 + slot_f4;
  for (int64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
    float4 g = f4_zero();
    for (int r = 0; r < world; ++r) {  // rank order: identical sums on every rank
      const float4 gr = r == rank ? own[i] : ld_relaxed_sys_f4(slots + int64_t(r) * n4 + i);
      g.x += gr.x; g.y += gr.y; g.z += gr.z; g.w += gr.w;
    }
    own[i] = f4_zero();  // ready for the next backward
    float4 pp = reinterpret_cast<float4*>(p)[i], mm = reinterpret_cast<float4*>(m)[i], vv = reinterpret_cast<float4*>(v)[i];
    float* pa = &pp.x; float* ma = &mm.x; float* va = &vv.x; float* ga = &g.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float param = pa[k], grad = ga[k] * gscale;
      if (adamw) param *= 1.f - lr * wd; else grad = fmaf(wd, param, grad);
      const float mi = fmaf(b1, ma[k], (1.f - b1) * grad);
      const float vi = fmaf(b2, va[k], (1.f - b2) * grad * grad);
      ma[k] = mi; va[k] = vi;
      pa[k] = param - (lr / bc1) * (mi / (sqrtf(vi) / bc2_sqrt + eps));
    }
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
  }
  // the CTA that finishes last moves the epoch (and the step count) on: every CTA of this launch has read them by then
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    unsigned long long* done = reinterpret_cast<unsigned long long*>(epoch_dev + 1);
    if (atomicAdd(done, 1ull) == gridDim.x - 1) {
      *done = 0ull;
      epoch_dev[0] = int64_t(epoch);
      if (advance) step_dev[0] = step;
    }
  }
}

}  // namespace
}  // namespace lgk

using namespace lgk;

extern "C" int64_t lgk_dp_buffer_bytes(int64_t n, int world) {
  if (n < 0 || world < 1) return 0;
  const int64_t n_pad = (n + 3) / 4 * 4;
  return kPadBytes + 4 * n_pad * (1 + 2 * int64_t(world > 1 ? world : 0));
}
extern "C" int64_t lgk_dp_grad_offset_bytes(void) { return kPadBytes; }

extern "C" int lgk_dp_allreduce_adam(const uint64_t* peer_base_dev, int rank, int world, int64_t n, float* param, float* exp_avg,
                                     float* exp_avg_sq, float lr, float beta1, float beta2, float eps, float weight_decay,
                                     int64_t* step_dev, int64_t* epoch_dev, int adamw, int advance_step, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(world >= 1 && world <= kMaxRanks && rank >= 0 && rank < world, LGK_ERR_BAD_ARG, "lgk_dp_allreduce_adam: bad rank / world");
  LGK_REQUIRE(n >= 0 && n % 4 == 0, LGK_ERR_BAD_ARG, "lgk_dp_allreduce_adam: the bucket length must be a multiple of 4");
  if (n == 0) return LGK_OK;
  LGK_REQUIRE(peer_base_dev && param && exp_avg && exp_avg_sq && step_dev && epoch_dev, LGK_ERR_BAD_ARG,
              "lgk_dp_allreduce_adam: null pointer");
  // the chunking must be the same on every rank: a function of n alone.  <= 64 CTAs: co-resident by a wide margin
  const unsigned grid = unsigned(std::min<int64_t>(ceil_div(n / 4, 256), kMaxCtas));
  dp_allreduce_adam_kernel<<<grid, 256, 0, stream>>>(reinterpret_cast<const uint64_t* const*>(peer_base_dev), rank, world, n, param,
                                                     exp_avg, exp_avg_sq, lr, beta1, beta2, eps, weight_decay, step_dev, epoch_dev, adamw,
                                                     advance_step);
  LGK_CHECK_LAUNCH("dp_allreduce_adam_kernel");
  return LGK_OK;
}
