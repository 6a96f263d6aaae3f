// (5) node-feature transform on the 5th-generation tensor cores: out = act(rowscale * x W^T + bias [+ out]).
//
// The reference's in_proj (models/gat.py:29, gin.py:21: Linear(1025 -> d)) is the one real dense GEMM of the
// path: 2*N*1025*d flops over N*1025*4 bytes of x, ~32 flop/B -- far too much for the FFMA pipe (it was the
// dominant kernel of the step, 7 % of the HBM roofline) and far below the tensor ridge, i.e. HBM-bound once
// the products run on tcgen05.  Plain TF32 (10-bit mantissa) would break the 1e-5 parity bar, so every operand
// is split on the fly into x = hi + lo (hi = x with the 13 low mantissa bits cleared, lo = x - hi, exact) and
// three MMAs accumulate lo*hi + hi*lo + hi*hi in fp32 in tensor memory ("3xTF32": the dropped lo*lo term and
// the truncation of lo are ~2^-22 relative).
//
// Persistent kernel, one CTA per SM, tiles of 128 rows of x against 64 rows of W.  21 warps in three roles:
//   stagers (16 warps)    plain 4-byte loads (lane = k, 128 bytes per warp load: the reference's 1025-float rows
//                         are not 16-byte aligned, so neither TMA tensor maps nor vector loads apply) of the K
//                         blocks into registers, two blocks ahead of their use; split, and the hi and lo operand
//                         tiles written in the K-major SWIZZLE_128B layout the tensor core reads (a warp writes one
//                         128-byte tile row per store: conflict free), fence.proxy.async, arrive.  (Until r01 v3
//                         the blocks went through a raw shared-memory ring with 4-byte cp.async: the LDGSTS issue
//                         rate, ~10 cycles per warp instruction, made that the cost of the kernel; per-row bulk
//                         copies of ~270 bytes measured 3x slower still, 16-byte chunk copies 1.5x slower.)
//   MMA (1 lane)          12 tcgen05.mma.kind::tf32 (M128 N64 K8) per 32-wide K block, commit -> frees the tile
//   epilogue (4 warps)    tcgen05.ld of the finished accumulators, row scale / bias / accumulate / ELU, store;
//                         overlaps the next tile's main loop (two accumulator buffers in tensor memory)
// The tensor core adds into its fp32 accumulator with truncation, which shows up as a relative error growing
// linearly with K (6e-6 at K = 1025): the four K steps of every block go to kAcc = 4 independent accumulators
// that the epilogue adds in fp32 (round to nearest) -- which also gives the tensor core four independent
// dependency chains instead of one.  Tensor memory: 2 buffers x 4 accumulators x 64 columns.
#include "common.cuh"
#include "umma.cuh"
#include <stdlib.h>
#include <algorithm>

namespace lgk {
namespace {

constexpr int BM = 128, BN = 64, BK = 32, kAcc = 4;
constexpr int kTileStages = 4;   // = stager groups
constexpr int kTileA = BM * BK * 4, kTileB = BN * BK * 4;      // bytes of one hi (or lo) operand tile
constexpr int kTileBytes = 2 * (kTileA + kTileB);              // A_hi, A_lo, B_hi, B_lo
constexpr int kSmemBytes = kTileStages * kTileBytes + 256 /*barriers*/;
constexpr int kConvWarps = 16;
constexpr int kWarpMma = 4, kWarpConv = 5;                     // warps 0-3: epilogue (TMEM lane quadrant = warp id)
constexpr int kThreadsTc = (kWarpConv + kConvWarps) * 32;
constexpr int kTmemCols = 2 * kAcc * BN;
static_assert(BK / 8 == kAcc, "one accumulator per K step of a block");
static_assert(kTmemCols == 512, "tensor memory allocation must be a power of two >= 32 columns");
static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");

struct TcArgs {
  const float* x;
  const float* w;
  const float* bias;
  const float* rowscale;
  float* out;
  int64_t n, k, m;
  int act, accumulate;
  int bm;               // rows of x per tile (multiple of 16, <= BM): the MMA is always M = 128, rows >= bm are not stored
  int64_t tiles_n;      // column tiles (fastest tile index)
  int64_t tiles_total;
  int debug;  // profiling only (LGK_TC_DEBUG): 1 = no MMAs, 2 = no conversion, 4 = no global copies (bit mask)
  int wide;   // 1: two MMAs per K step (A_hi x [B_hi; B_lo] as one N = 128 product, then A_lo x B_hi) instead of three N = 64 ones
};

struct Bars {
  uint64_t tile_full[kTileStages], tile_empty[kTileStages];
  uint64_t acc_full[2], acc_empty[2];
  uint32_t tmem_slot;
};

// hi = x with the 13 low mantissa bits cleared (exactly representable in TF32), lo = x - hi (exact)
__device__ __forceinline__ void split_mask(float x, float& hi, float& lo) {
  hi = __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);
  lo = x - hi;
}

// ---- epilogue -------------------------------------------------------------------------------------------------
__device__ __forceinline__ void epilogue_tile(const TcArgs& g, uint32_t tmem_acc, int nacc, int warp, int lane, int64_t i0,
                                              int64_t j0) {
  const int64_t i = (warp * 32 + lane < g.bm) ? i0 + warp * 32 + lane : g.n;  // rows of the MMA above bm: not stored
  const float rs = (g.rowscale && i < g.n) ? __ldg(g.rowscale + i) : 1.f;
#pragma unroll 1
  for (int c = 0; c < BN / 16; ++c) {
    float v[16];
    umma::tmem_ld16(tmem_acc + (uint32_t(warp * 32) << 16) + c * 16, v);
    for (int a = 1; a < nacc; ++a) {
      float u[16];
      umma::tmem_ld16(tmem_acc + (uint32_t(warp * 32) << 16) + a * BN + c * 16, u);
#pragma unroll
      for (int t = 0; t < 16; ++t) v[t] += u[t];
    }
    if (i >= g.n) continue;
    float* dst = g.out + i * g.m + j0 + c * 16;
    if (j0 + c * 16 + 16 <= g.m && (g.m & 3) == 0) {
#pragma unroll
      for (int t = 0; t < 16; t += 4) {
        float4 r = make_float4(v[t] * rs, v[t + 1] * rs, v[t + 2] * rs, v[t + 3] * rs);
        if (g.bias) {
          const float4 b = __ldg(reinterpret_cast<const float4*>(g.bias + j0 + c * 16 + t));
          r.x += b.x; r.y += b.y; r.z += b.z; r.w += b.w;
        }
        if (g.accumulate) {
          const float4 o = *reinterpret_cast<const float4*>(dst + t);
          r.x += o.x; r.y += o.y; r.z += o.z; r.w += o.w;
        }
        if (g.act == LGK_ACT_ELU) { r.x = elu1(r.x); r.y = elu1(r.y); r.z = elu1(r.z); r.w = elu1(r.w); }
        else if (g.act == LGK_ACT_RELU) { r.x = fmaxf(r.x, 0.f); r.y = fmaxf(r.y, 0.f); r.z = fmaxf(r.z, 0.f); r.w = fmaxf(r.w, 0.f); }
        *reinterpret_cast<float4*>(dst + t) = r;
      }
    } else {
#pragma unroll
      for (int t = 0; t < 16; ++t) {
        const int64_t j = j0 + c * 16 + t;
        if (j < g.m) {
          float r = v[t] * rs;
          if (g.bias) r += __ldg(g.bias + j);
          if (g.accumulate) r += dst[t];
          if (g.act == LGK_ACT_ELU) r = elu1(r);
          else if (g.act == LGK_ACT_RELU) r = fmaxf(r, 0.f);
          dst[t] = r;
        }
      }
    }
  }
}

// ---- kernel ---------------------------------------------------------------------------------------------------
template <int NA>
__global__ void __launch_bounds__(kThreadsTc, 1) linear_tc_kernel(TcArgs g) {
  extern __shared__ __align__(1024) uint8_t smem_dyn[];  // SWIZZLE_128B tiles need a 1024-byte aligned base
  const uint32_t tiles = umma::smem_u32(smem_dyn);        // operand tile ring
  Bars* bars = reinterpret_cast<Bars*>(smem_dyn + kTileStages * kTileBytes);
  if ((tiles & 1023u) != 0) __trap();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nkb = int((g.k + BK - 1) / BK);

  if (threadIdx.x == 0) {
    for (int s = 0; s < kTileStages; ++s) {
      umma::mbar_init(bars->tile_full + s, kConvWarps / kTileStages);  // the warps of one stager group
      umma::mbar_init(bars->tile_empty + s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      umma::mbar_init(bars->acc_full + s, 1);
      umma::mbar_init(bars->acc_empty + s, 4);
    }
    umma::mbar_init_fence();
  }
  if (warp == kWarpMma) umma::tmem_alloc(&bars->tmem_slot, kTmemCols);
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tmem_base = bars->tmem_slot;

  if (warp == kWarpMma) {
    // ---- MMA issuer ----
    const uint32_t idesc = umma::idesc_tf32(BM, BN), idesc_w = umma::idesc_tf32(BM, 2 * BN);
    const bool wide = g.wide != 0;
    uint32_t it = 0, li = 0;
    for (int64_t t = blockIdx.x; t < g.tiles_total; t += gridDim.x, ++li) {
      // wide: the four accumulators are 128 columns each ([hi*hi | hi*lo + lo*hi]) and fill the tensor memory: one buffer
      const uint32_t ab = wide ? 0u : (li & 1);
      umma::mbar_wait(bars->acc_empty + ab, (wide ? (li & 1) : ((li >> 1) & 1)) ^ 1);  // the epilogue has drained this buffer
      umma::tc_fence_after();
      for (int kb = 0; kb < nkb; ++kb, ++it) {
        const int s = it % kTileStages;
        umma::mbar_wait(bars->tile_full + s, (it / kTileStages) & 1);  // acquires the stagers' generic-proxy writes
        umma::fence_proxy_async_smem();                                // ... and hands them to the async proxy (the MMA)
        umma::tc_fence_after();
        if (lane == 0 && (g.debug & 1)) {
          umma::mma_commit(bars->tile_empty + s);
          if (kb == nkb - 1) umma::mma_commit(bars->acc_full + ab);
        } else if (lane == 0) {
          const uint32_t sa = tiles + s * kTileBytes;
          const uint64_t a_hi = umma::smem_desc_k_sw128(sa), a_lo = umma::smem_desc_k_sw128(sa + kTileA);
          const uint64_t b_hi = umma::smem_desc_k_sw128(sa + 2 * kTileA), b_lo = umma::smem_desc_k_sw128(sa + 2 * kTileA + kTileB);
          // Dependent MMAs on one accumulator serialise with their full latency (~240 cycles each at this tile
          // size), so the four K steps of a block go to four different accumulators AND the issue order is
          // term-major: consecutive instructions always target different accumulators, a dependent one follows
          // four issue slots later.  Per accumulator the order of additions (lo*hi, hi*lo, hi*hi) is unchanged.
          if (wide) {
            // B_hi and B_lo tiles are adjacent in the stage: [B_hi; B_lo] is ONE K-major operand of 128 rows.  Per K step
            // A_hi x [B_hi; B_lo] -> columns [hi*hi | hi*lo] (N = 128), then A_lo x B_hi added onto the second half (N = 64):
            // 8 MMAs per block instead of 12 -- the MMA issue chain, ~60 cycles per instruction whatever N, is what a tile
            // with 33 dependent K blocks waits for (LGK_TC_DEBUG ablation, profiles/README.md r02).
#pragma unroll
            for (int kk = 0; kk < BK / 8; ++kk) {
              const uint64_t adv = uint64_t(kk * 32 >> 4);
              umma::mma_tf32(tmem_base + uint32_t(kk) * 2 * BN, a_hi + adv, b_hi + adv, idesc_w, kb > 0);
            }
#pragma unroll
            for (int kk = 0; kk < BK / 8; ++kk) {
              const uint64_t adv = uint64_t(kk * 32 >> 4);
              umma::mma_tf32(tmem_base + uint32_t(kk) * 2 * BN + BN, a_lo + adv, b_hi + adv, idesc, true);
            }
          } else
#pragma unroll
          for (int term = 0; term < 3; ++term) {
#pragma unroll
            for (int kk = 0; kk < BK / 8; ++kk) {
              const uint32_t acc = tmem_base + (ab * kAcc + uint32_t(kk)) * BN;
              const uint64_t adv = uint64_t(kk * 32 >> 4);  // +32 bytes along K inside the 128-byte swizzled row
              if (term == 0) umma::mma_tf32(acc, a_lo + adv, b_hi + adv, idesc, kb > 0);  // small terms first
              else if (term == 1) umma::mma_tf32(acc, a_hi + adv, b_lo + adv, idesc, true);
              else umma::mma_tf32(acc, a_hi + adv, b_hi + adv, idesc, true);
            }
          }
          umma::mma_commit(bars->tile_empty + s);                      // frees the tile when these MMAs have read it
          if (kb == nkb - 1) umma::mma_commit(bars->acc_full + ab);    // all accumulators of the tile complete
        }
        __syncwarp();
      }
    }
  } else if (warp >= kWarpConv) {
    // ---- stagers: group grp (4 warps) stages the K blocks q = grp (mod 4) into tile stage grp ----
    // One batch of loads per block and thread (bm/4 x rows + 16 W rows of this warp, lane = k), consumed in the same
    // iteration: the four groups are in different phases of load / wait / convert, which is what overlaps the
    // memory latency (four blocks = 4 x 18-24 KB in flight per SM).
    const int cw = warp - kWarpConv;
    const int grp = cw >> 2, wg = cw & 3;
    constexpr int RA = 4 * NA, RB = BN / 4;
    const int64_t my_tiles = (g.tiles_total - blockIdx.x + gridDim.x - 1) / gridDim.x;
    const int64_t total = my_tiles * nkb;   // the K blocks of all tiles of this CTA form one sequence
    const uint32_t swz = uint32_t(lane) >> 2, lo2 = (uint32_t(lane) & 3u) << 2;
    const uint32_t tile = tiles + grp * kTileBytes;
#pragma unroll 1
    for (int64_t q = grp; q < total; q += kTileStages) {
      const int64_t lt = q / nkb, t = blockIdx.x + lt * gridDim.x;
      const int64_t i0 = (t / g.tiles_n) * g.bm + wg * RA, j0 = (t % g.tiles_n) * BN + wg * RB;  // first rows of this warp
      const int64_t k0 = (q - lt * nkb) * BK;
      const float* xa = g.x + i0 * g.k + k0 + lane;
      const float* wb = g.w + j0 * g.k + k0 + lane;
      float va[RA], vb[RB];
      if (g.debug & 4) {
#pragma unroll
        for (int r = 0; r < RA; ++r) va[r] = 0.f;
#pragma unroll
        for (int r = 0; r < RB; ++r) vb[r] = 0.f;
      } else if (i0 + RA <= g.n && j0 + RB <= g.m && k0 + BK <= g.k) {  // interior: no predicates
#pragma unroll
        for (int r = 0; r < RA; ++r) va[r] = __ldg(xa + r * g.k);
#pragma unroll
        for (int r = 0; r < RB; ++r) vb[r] = __ldg(wb + r * g.k);
      } else {
        const bool kin = k0 + lane < g.k;
#pragma unroll
        for (int r = 0; r < RA; ++r) va[r] = (kin && i0 + r < g.n) ? __ldg(xa + r * g.k) : 0.f;
#pragma unroll
        for (int r = 0; r < RB; ++r) vb[r] = (kin && j0 + r < g.m) ? __ldg(wb + r * g.k) : 0.f;
      }
      umma::mbar_wait(bars->tile_empty + grp, uint32_t((q / kTileStages) & 1) ^ 1);
      if (!(g.debug & 2)) {
#pragma unroll
        for (int r = 0; r < RA; ++r) {
          const uint32_t row = uint32_t(wg * RA + r);
          const uint32_t off = tile + row * 128u + (((swz ^ row) & 7u) << 4) + lo2;
          float hi, lo;
          split_mask(va[r], hi, lo);
          umma::sts_f32(off, hi);
          umma::sts_f32(off + kTileA, lo);
        }
#pragma unroll
        for (int r = 0; r < RB; ++r) {
          const uint32_t row = uint32_t(wg * RB + r);
          const uint32_t off = tile + 2 * kTileA + row * 128u + (((swz ^ row) & 7u) << 4) + lo2;
          float hi, lo;
          split_mask(vb[r], hi, lo);
          umma::sts_f32(off, hi);
          umma::sts_f32(off + kTileB, lo);
        }
      }
      // no proxy fence here: fence.proxy.async compiles to MEMBAR.ALL.CTA + FENCE.VIEW.ASYNC; the MMA thread fences
      // after it has acquired these writes through the barrier
      __syncwarp();
      if (lane == 0) umma::mbar_arrive(bars->tile_full + grp);
    }
  } else {
    // ---- epilogue: warp w < 4 owns TMEM lanes [32w, 32w+32) = rows i0 + 32w + lane ----
    const int nacc = kAcc;  // every K block feeds all four accumulators (one per K step of 8)
    uint32_t li = 0;
    for (int64_t t = blockIdx.x; t < g.tiles_total; t += gridDim.x, ++li) {
      const int64_t i0 = (t / g.tiles_n) * g.bm, j0 = (t % g.tiles_n) * BN;
      const bool wide = g.wide != 0;
      const uint32_t ab = wide ? 0u : (li & 1);
      umma::mbar_wait_relaxed(bars->acc_full + ab, wide ? (li & 1) : ((li >> 1) & 1));
      umma::tc_fence_after();
      epilogue_tile(g, tmem_base + ab * kAcc * BN, wide ? 2 * nacc : nacc, warp, lane, i0, j0);
      umma::tc_fence_before();
      __syncwarp();
      if (lane == 0) umma::mbar_arrive(bars->acc_empty + ab);
    }
  }
  umma::tc_fence_before();
  __syncthreads();
  if (warp == kWarpMma) umma::tmem_dealloc(tmem_base, kTmemCols);
}

}  // namespace

// used by lgk_linear_fwd (linear.cu) when the product is large enough to be worth a tensor-core launch
int launch_linear_tc(const float* x, const float* w, const float* bias, const float* rowscale, float* out, int64_t n,
                     int64_t k, int64_t m, int act, int accumulate, cudaStream_t stream) {
  // the attribute is per device: one flag per device ordinal
  static bool configured_dev[64] = {};
  int dev_ord = 0;
  cudaGetDevice(&dev_ord);
  bool& configured = configured_dev[dev_ord & 63];
  if (!configured) {
    cudaError_t e = cudaSuccess;
    for (auto* fn : {linear_tc_kernel<4>, linear_tc_kernel<5>, linear_tc_kernel<6>, linear_tc_kernel<7>, linear_tc_kernel<8>})
      if (e == cudaSuccess) e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
    if (e != cudaSuccess) {
      set_error("linear_tc_kernel: cannot reserve %d bytes of shared memory: %s", kSmemBytes, cudaGetErrorString(e));
      return LGK_ERR_LAUNCH;
    }
    configured = true;
  }
  TcArgs g{x, w, bias, rowscale, out, n, k, m, act, accumulate, BM, 0, 0, 0, 0};
  static const int debug = [] { const char* v = getenv("LGK_TC_DEBUG"); return (v && *v) ? atoi(v) : 0; }();
  g.debug = debug;
  static const int wide = [] { const char* v = getenv("LGK_TC_WIDE"); return (v && *v) ? atoi(v) : -1; }();
  g.tiles_n = ceil_div(m, BN);
  // tile height: the smallest multiple of 16 rows that still covers the matrix in the same number of waves, so that a
  // batch of 11 583 nodes becomes 145 tiles of 80 rows on 145 SMs instead of 91 tiles of 128 rows on 91 SMs (the
  // kernel is bound by what one SM can pull from memory, not by the MMAs, whose M stays 128)
  {
    const int64_t waves = ceil_div(ceil_div(n, BM) * g.tiles_n, kNumSMs);
    const int64_t row_tiles = std::max<int64_t>(1, waves * kNumSMs / g.tiles_n);
    g.bm = int(std::min<int64_t>(BM, std::max<int64_t>(64, align_up(ceil_div(n, row_tiles), 16))));
  }
  g.tiles_total = ceil_div(n, g.bm) * g.tiles_n;
  // two MMAs per K step (see the kernel) when every CTA has a single tile: 36.5 -> 33.5 us on the headline batch.  With several
  // tiles per CTA the wide accumulators leave no second tensor-memory buffer for the epilogue to overlap the next main loop:
  // 409 600 x 1025 -> 64 runs 578 us with three MMAs and 703 us with two.  LGK_TC_WIDE=0 / 1 forces either.
  g.wide = wide >= 0 ? wide : (g.tiles_total <= kNumSMs ? 1 : 0);
  const unsigned grid = unsigned(g.tiles_total < kNumSMs ? g.tiles_total : kNumSMs);
  switch (g.bm / kConvWarps) {  // x rows per stager warp
    case 4: linear_tc_kernel<4><<<grid, kThreadsTc, kSmemBytes, stream>>>(g); break;
    case 5: linear_tc_kernel<5><<<grid, kThreadsTc, kSmemBytes, stream>>>(g); break;
    case 6: linear_tc_kernel<6><<<grid, kThreadsTc, kSmemBytes, stream>>>(g); break;
    case 7: linear_tc_kernel<7><<<grid, kThreadsTc, kSmemBytes, stream>>>(g); break;
    default: linear_tc_kernel<8><<<grid, kThreadsTc, kSmemBytes, stream>>>(g); break;
  }
  LGK_CHECK_LAUNCH("linear_tc_kernel");
  return LGK_OK;
}

}  // namespace lgk
