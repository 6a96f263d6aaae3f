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
//   stagers (16 warps)    cp.async (4 bytes per lane, lane = k: the reference's 1025-float rows are not 16-byte
//                         aligned, so neither TMA tensor maps nor vector loads apply, and per-row bulk copies
//                         of ~270 bytes measured 3x slower, 16-byte chunk copies 1.5x slower) of the raw K
//                         blocks into a 3-deep ring in shared memory without holding a register; every thread reads
//                         back exactly the bytes it copied (cp.async.wait_group, no barrier), splits them and
//                         writes the hi and lo operand tiles in the K-major SWIZZLE_128B layout the tensor core
//                         reads (a warp writes one 128-byte tile row per store: conflict free),
//                         fence.proxy.async, arrive
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

namespace lgk {
namespace {

constexpr int BM = 128, BN = 64, BK = 32, kAcc = 4;
constexpr int kRawStages = 3, kTileStages = 3;
constexpr int kRawBytes = (BM + BN) * BK * 4;                  // one raw K block: x rows then W rows, 128 bytes each
constexpr int kTileA = BM * BK * 4, kTileB = BN * BK * 4;      // bytes of one hi (or lo) operand tile
constexpr int kTileBytes = 2 * (kTileA + kTileB);              // A_hi, A_lo, B_hi, B_lo
constexpr int kSmemBytes = kTileStages * kTileBytes + kRawStages * kRawBytes + 256 /*barriers*/;
constexpr int kConvWarps = 16;
constexpr int kWarpMma = 4, kWarpConv = 5;                     // warps 0-3: epilogue (TMEM lane quadrant = warp id)
constexpr int kThreadsTc = (kWarpConv + kConvWarps) * 32;
constexpr int kRowsA = BM / kConvWarps, kRowsB = BN / kConvWarps;  // tile rows staged per warp: 8 and 4
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
  int64_t tiles_n;      // column tiles (fastest tile index)
  int64_t tiles_total;
  int debug;  // profiling only (LGK_TC_DEBUG): 1 = no MMAs, 2 = no conversion, 4 = no global copies (bit mask)
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

// 4-byte asynchronous copy global -> shared (32-bit shared address); src_bytes 0 writes zeros without reading
__device__ __forceinline__ void cp_async4(uint32_t dst, const float* src, uint32_t src_bytes) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const float* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// ---- stager ---------------------------------------------------------------------------------------------------
// Per K block, lane = k and a warp owns 8 x rows and 4 W rows: raw element (slot, lane) lives at slot * 128 + lane * 4
// of the raw stage (a warp touches 128 consecutive bytes: conflict free both ways).  Every thread reads back exactly
// the bytes it copied, so cp.async.wait_group is the only synchronisation the raw ring needs.  All shared-memory
// addresses are 32-bit shared-window addresses.
struct StageCtx {
  const float* xa[kRowsA];   // &x[row, lane] (clamped to row 0 when the row is outside the matrix)
  const float* wb[kRowsB];
  uint32_t ta[kRowsA], tb[kRowsB];   // swizzled tile byte offsets of (row, lane)
  uint32_t raw_a, raw_b;             // raw byte offsets of this thread's first x / W element
  uint32_t va, vb;                   // row-valid bit masks
  bool rows_full;
};

__device__ __forceinline__ void issue_block(const TcArgs& g, const StageCtx& c, uint32_t raw, int lane, int64_t k0) {
  if (c.rows_full && k0 + BK <= g.k) {  // interior block: no predicates
#pragma unroll
    for (int r = 0; r < kRowsA; ++r) cp_async4(raw + c.raw_a + r * 128, c.xa[r] + k0);
#pragma unroll
    for (int r = 0; r < kRowsB; ++r) cp_async4(raw + c.raw_b + r * 128, c.wb[r] + k0);
  } else {
    const bool kin = k0 + lane < g.k;
    const int64_t ko = kin ? k0 : 0;
#pragma unroll
    for (int r = 0; r < kRowsA; ++r) cp_async4(raw + c.raw_a + r * 128, c.xa[r] + ko, (kin && ((c.va >> r) & 1)) ? 4u : 0u);
#pragma unroll
    for (int r = 0; r < kRowsB; ++r) cp_async4(raw + c.raw_b + r * 128, c.wb[r] + ko, (kin && ((c.vb >> r) & 1)) ? 4u : 0u);
  }
}

__device__ __forceinline__ void convert_block(const StageCtx& c, uint32_t raw, uint32_t tile) {
  float va[kRowsA], vb[kRowsB];
#pragma unroll
  for (int r = 0; r < kRowsA; ++r) va[r] = umma::lds_f32(raw + c.raw_a + r * 128);
#pragma unroll
  for (int r = 0; r < kRowsB; ++r) vb[r] = umma::lds_f32(raw + c.raw_b + r * 128);
#pragma unroll
  for (int r = 0; r < kRowsA; ++r) {
    float hi, lo;
    split_mask(va[r], hi, lo);
    umma::sts_f32(tile + c.ta[r], hi);
    umma::sts_f32(tile + kTileA + c.ta[r], lo);
  }
#pragma unroll
  for (int r = 0; r < kRowsB; ++r) {
    float hi, lo;
    split_mask(vb[r], hi, lo);
    umma::sts_f32(tile + 2 * kTileA + c.tb[r], hi);
    umma::sts_f32(tile + 2 * kTileA + kTileB + c.tb[r], lo);
  }
}

// ---- epilogue -------------------------------------------------------------------------------------------------
__device__ __forceinline__ void epilogue_tile(const TcArgs& g, uint32_t tmem_acc, int nacc, int warp, int lane, int64_t i0,
                                              int64_t j0) {
  const int64_t i = i0 + warp * 32 + lane;
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
__global__ void __launch_bounds__(kThreadsTc, 1) linear_tc_kernel(TcArgs g) {
  extern __shared__ __align__(1024) uint8_t smem_dyn[];  // SWIZZLE_128B tiles need a 1024-byte aligned base
  const uint32_t tiles = umma::smem_u32(smem_dyn);        // operand tile ring
  const uint32_t raws = tiles + kTileStages * kTileBytes;  // raw K block ring
  Bars* bars = reinterpret_cast<Bars*>(smem_dyn + kTileStages * kTileBytes + kRawStages * kRawBytes);
  if ((tiles & 1023u) != 0) __trap();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nkb = int((g.k + BK - 1) / BK);

  if (threadIdx.x == 0) {
    for (int s = 0; s < kTileStages; ++s) {
      umma::mbar_init(bars->tile_full + s, kConvWarps);
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
    const uint32_t idesc = umma::idesc_tf32(BM, BN);
    uint32_t it = 0, li = 0;
    for (int64_t t = blockIdx.x; t < g.tiles_total; t += gridDim.x, ++li) {
      const uint32_t ab = li & 1;
      umma::mbar_wait(bars->acc_empty + ab, ((li >> 1) & 1) ^ 1);  // the epilogue has drained this accumulator buffer
      umma::tc_fence_after();
      for (int kb = 0; kb < nkb; ++kb, ++it) {
        const int s = it % kTileStages;
        umma::mbar_wait(bars->tile_full + s, (it / kTileStages) & 1);
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
    // ---- stagers ----
    const int cw = warp - kWarpConv;
    StageCtx c;
#pragma unroll
    for (int r = 0; r < kRowsA; ++r) c.ta[r] = umma::sw128_offset_f32(cw * kRowsA + r, lane);
#pragma unroll
    for (int r = 0; r < kRowsB; ++r) c.tb[r] = umma::sw128_offset_f32(cw * kRowsB + r, lane);
    c.raw_a = uint32_t(cw * kRowsA) * 128 + lane * 4;
    c.raw_b = uint32_t(BM + cw * kRowsB) * 128 + lane * 4;
    // the K blocks of all tiles of this CTA form one sequence: block q = (tile q / nkb, K block q % nkb); copies
    // run kRawStages blocks ahead of the conversion, across tile boundaries
    const int64_t my_tiles = (g.tiles_total - blockIdx.x + gridDim.x - 1) / gridDim.x;
    const int64_t total = my_tiles * nkb;
    int64_t tile_of_ctx = -1;
    auto set_tile = [&](int64_t lt) {  // row pointers of local tile lt
      const int64_t t = blockIdx.x + lt * gridDim.x;
      const int64_t i0 = (t / g.tiles_n) * BM, j0 = (t % g.tiles_n) * BN;
      c.va = c.vb = 0;
#pragma unroll
      for (int r = 0; r < kRowsA; ++r) {
        const int64_t row = i0 + cw * kRowsA + r;
        const bool ok = row < g.n;
        c.va |= uint32_t(ok) << r;
        c.xa[r] = g.x + (ok ? row : 0) * g.k + lane;
      }
#pragma unroll
      for (int r = 0; r < kRowsB; ++r) {
        const int64_t row = j0 + cw * kRowsB + r;
        const bool ok = row < g.m;
        c.vb |= uint32_t(ok) << r;
        c.wb[r] = g.w + (ok ? row : 0) * g.k + lane;
      }
      c.rows_full = (i0 + BM <= g.n) && (j0 + BN <= g.m);
      tile_of_ctx = lt;
    };
    auto issue = [&](int64_t q) {
      if (q < total && !(g.debug & 4)) {
        const int64_t lt = q / nkb;
        if (lt != tile_of_ctx) set_tile(lt);
        issue_block(g, c, raws + uint32_t(q % kRawStages) * kRawBytes, lane, (q - lt * nkb) * BK);
      }
      cp_async_commit();  // an (empty) group per block keeps the wait_group arithmetic uniform
    };
    for (int q = 0; q < kRawStages; ++q) issue(q);
    for (int64_t q = 0; q < total; ++q) {
      cp_async_wait<kRawStages - 1>();  // this thread's copies of block q have landed
      const int ts = int(q % kTileStages);
      umma::mbar_wait(bars->tile_empty + ts, uint32_t((q / kTileStages) & 1) ^ 1);
      if (!(g.debug & 2)) convert_block(c, raws + uint32_t(q % kRawStages) * kRawBytes, tiles + ts * kTileBytes);
      umma::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) umma::mbar_arrive(bars->tile_full + ts);
      issue(q + kRawStages);  // refill the raw slot just consumed (the same thread wrote and read it)
    }
  } else {
    // ---- epilogue: warp w < 4 owns TMEM lanes [32w, 32w+32) = rows i0 + 32w + lane ----
    const int nacc = kAcc;  // every K block feeds all four accumulators (one per K step of 8)
    uint32_t li = 0;
    for (int64_t t = blockIdx.x; t < g.tiles_total; t += gridDim.x, ++li) {
      const int64_t i0 = (t / g.tiles_n) * BM, j0 = (t % g.tiles_n) * BN;
      const uint32_t ab = li & 1;
      umma::mbar_wait_relaxed(bars->acc_full + ab, (li >> 1) & 1);
      umma::tc_fence_after();
      epilogue_tile(g, tmem_base + ab * kAcc * BN, nacc, warp, lane, i0, j0);
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
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(linear_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
    if (e != cudaSuccess) {
      set_error("linear_tc_kernel: cannot reserve %d bytes of shared memory: %s", kSmemBytes, cudaGetErrorString(e));
      return LGK_ERR_LAUNCH;
    }
    configured = true;
  }
  TcArgs g{x, w, bias, rowscale, out, n, k, m, act, accumulate, 0, 0, 0};
  static const int debug = [] { const char* v = getenv("LGK_TC_DEBUG"); return (v && *v) ? atoi(v) : 0; }();
  g.debug = debug;
  g.tiles_n = ceil_div(m, BN);
  g.tiles_total = ceil_div(n, BM) * g.tiles_n;
  const unsigned grid = unsigned(g.tiles_total < kNumSMs ? g.tiles_total : kNumSMs);
  linear_tc_kernel<<<grid, kThreadsTc, kSmemBytes, stream>>>(g);
  LGK_CHECK_LAUNCH("linear_tc_kernel");
  return LGK_OK;
}

}  // namespace lgk
