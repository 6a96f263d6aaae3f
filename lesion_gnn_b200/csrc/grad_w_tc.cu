// (5c) weight (and bias) gradient of a wide node-feature transform on the tensor cores:
//   dW[m,k] = sum_i dy[i,m] x[i,k],   db[m] = sum_i dy[i,m].
//
// The gradient of the reference's in_proj (models/gat.py:29, gin.py:21: Linear(1025 -> d)) reduces over the nodes
// of the batch: 2*N*1025*d flops over N*(1025 + d)*4 bytes.  On the FFMA pipe it was the longest kernel of the
// training step (72 us on the headline batch, 9 % of the HBM roofline).  Here both operands are read exactly as
// they lie in memory: with the node index as the MMA's K dimension, x [N, k] IS an MN-major A operand (features
// contiguous) and dy [N, m] IS an MN-major B operand (m contiguous) -- no transposition anywhere, and a warp that
// copies 32 consecutive floats of one node writes exactly one 128-byte row of an MN-major atom, conflict free.
// D[feature, m] = dW^T accumulates in tensor memory (lane = feature, column = m), so the epilogue's stores to
// dW[m*K + feature] are coalesced over the lanes.  3xTF32 as in linear_tc.cu (hi/lo split on the fly, three MMAs
// per K step, four independent accumulators added in fp32 by the epilogue).
// The bias gradient rides along: when K is not a multiple of 128 the first unused feature row of the last tile is
// fed with ones, so D[K, m] = db[m] -- no separate column-sum pass over dy.
//
// Grid: feature tiles of 128 columns of x  x  slabs of nodes, ntiles * slabs <= 148 CTAs (one per SM, one wave).
// A CTA walks its slab in blocks of 32 nodes: 16 stager warps in four phase-shifted groups (group = block index
// mod 4) load their block (4 bytes per lane, 128 bytes per warp load: the reference's 1025-float rows are not
// 16-byte aligned) into registers, split it and write the hi/lo tiles of a 4-deep ring (4-byte cp.async into a
// raw ring measured 56 us against 38 us: the LDGSTS issue rate, ~10 cycles per warp instruction, was the cost);
// one lane issues 12 tcgen05.mma (M128 N=m K8) per block; at the end 4 warps drain the
// accumulators into partial[slab][m*K + k | m] (the layout of dW, then db), which lgk_linear_bwd_w reduces in
// slab order (deterministic).  m <= 128.
#include "common.cuh"
#include "umma.cuh"
#include <stdlib.h>

namespace lgk {
namespace {

constexpr int BF = 128;          // features (columns of x) per CTA = MMA M
constexpr int BNODE = 32;        // nodes per block = 4 MMA K steps of 8
constexpr int kAcc = 4;
constexpr int kAccCols = 128;    // tensor-memory columns reserved per accumulator (N = m rounded up to 16 <= 128)
constexpr int kAtom = 1024;                   // 8 nodes (one MMA K step) x 128 bytes
constexpr int kGroupStride = 4 * kAtom;       // one 32-column group: 4 node atoms back to back
constexpr int kTileA = 4 * kGroupStride;      // hi (or lo) tile of x: 128 features x 32 nodes, 16 KB
constexpr int kConvWarps = 16;
constexpr int kWarpMma = 4, kWarpConv = 5;
constexpr int kThreads = (kWarpConv + kConvWarps) * 32;
constexpr int kTmemCols = kAcc * kAccCols;
static_assert(kTmemCols == 512, "tensor memory allocation must be a power of two");

// MG = 32-column groups of dy (m <= 32 MG): the dy tiles take MG * 4 KB each
template <int MG> constexpr int tile_b() { return MG * kGroupStride; }
template <int MG> constexpr int tile_bytes() { return 2 * (kTileA + tile_b<MG>()); }
template <int MG> constexpr int tile_stages() { return MG <= 2 ? 4 : 3; }
template <int MG> constexpr int stager_groups() { return MG <= 2 ? 4 : 2; }
template <int MG> constexpr int smem_bytes() { return tile_stages<MG>() * tile_bytes<MG>() + 256; }
static_assert(smem_bytes<4>() <= 227 * 1024 && smem_bytes<2>() <= 227 * 1024, "shared memory budget");

struct GwArgs {
  const float* dy;
  const float* x;
  float* partial;
  int64_t n, k, m;
  int64_t slab_nodes;  // multiple of 32
  int fuse_db;         // feature row K of the last tile is fed with ones: D[K, :] = db
  int debug;           // profiling only (LGK_GW_DEBUG): 1 = no MMAs, 2 = no conversion, 4 = no global copies
  int wide;            // LGK_GW_WIDE (default 1): two MMAs per K step where the shape allows it
};

struct Bars {
  uint64_t tile_full[4], tile_empty[4];
  uint64_t acc_full;
  uint32_t tmem_slot;
};

__device__ __forceinline__ void split_mask(float x, float& hi, float& lo) {
  hi = __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);
  lo = x - hi;
}
// MN-major tf32 operands exist in one shared-memory layout only, SWIZZLE_128B with a 32-byte base (layout type 1):
// atoms of 4 K rows (nodes) x 128 bytes (32 columns), the four 32-byte chunks of a row XOR-permuted by the row index
// (address bits [5,7) ^= bits [7,9)); atoms of consecutive nodes kSbo bytes apart, the next 32 columns kGroupStride
// bytes further (the leading-dimension offset).  With the plain SWIZZLE_128B type the MMA returns zeros.
constexpr int kSbo = 512;
__device__ __forceinline__ uint64_t smem_desc_mn_sw128(uint32_t saddr) {
  return uint64_t((saddr & 0x3FFFFu) >> 4) | (uint64_t(kGroupStride >> 4) << 16) | (uint64_t(kSbo >> 4) << 32) |
         (uint64_t(1) << 46) | (uint64_t(1) << 61);
}
// element (node r of the block, column group cg, column c) of an operand tile lives at
//   cg * kGroupStride + r * 128 + (((c >> 3) ^ r) & 3) * 32 + (c & 7) * 4
__host__ __device__ constexpr uint32_t idesc_tf32_mn(int m, int n) { return umma::idesc_tf32(m, n) | (1u << 15) | (1u << 16); }  // A, B MN-major

template <int MG>
__global__ void __launch_bounds__(kThreads, 1) grad_w_tc_kernel(GwArgs g) {
  constexpr int kTileB = tile_b<MG>(), kTileBytes = tile_bytes<MG>(), kTileStages = tile_stages<MG>();
  extern __shared__ __align__(1024) uint8_t smem_dyn[];
  const uint32_t tiles = umma::smem_u32(smem_dyn);
  Bars* bars = reinterpret_cast<Bars*>(smem_dyn + kTileStages * kTileBytes);
  if ((tiles & 1023u) != 0) __trap();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t ntile = (g.k + BF - 1) / BF;
  const int64_t jt = blockIdx.x % ntile, sl = blockIdx.x / ntile;
  const int64_t j0 = jt * BF;
  const int64_t node0 = sl * g.slab_nodes;
  const int64_t slab = min(g.slab_nodes, g.n - node0);   // nodes of this slab (> 0 by construction of the grid)
  const int nkb = int((slab + BNODE - 1) / BNODE);
  const int ncols = int((g.m + 15) / 16) * 16;           // MMA N
  // two MMAs per K step instead of three when [dy_hi | dy_lo] is one contiguous operand (the dy tile is exactly MG groups
  // wide) that still fits an accumulator's 128 columns
  const bool wide = g.wide && ncols == 32 * MG && 2 * ncols <= kAccCols;

  // operand tiles start as zeros: columns of the dy tiles between m and N are never written
  for (int i = threadIdx.x; i < kTileStages * kTileBytes / 16; i += kThreads)
    *reinterpret_cast<uint4*>(smem_dyn + i * 16) = make_uint4(0, 0, 0, 0);
  umma::fence_proxy_async_smem();
  if (threadIdx.x == 0) {
    for (int s = 0; s < kTileStages; ++s) {
      umma::mbar_init(bars->tile_full + s, kConvWarps / stager_groups<MG>());  // the warps of one stager group
      umma::mbar_init(bars->tile_empty + s, 1);
    }
    umma::mbar_init(&bars->acc_full, 1);
    umma::mbar_init_fence();
  }
  if (warp == kWarpMma) umma::tmem_alloc(&bars->tmem_slot, kTmemCols);
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tmem_base = bars->tmem_slot;

  if (warp == kWarpMma) {
    const uint32_t idesc = idesc_tf32_mn(BF, ncols), idesc_w = idesc_tf32_mn(BF, 2 * ncols);
    for (int kb = 0; kb < nkb; ++kb) {
      const int s = kb % kTileStages;
      umma::mbar_wait(bars->tile_full + s, (kb / kTileStages) & 1);  // acquires the stagers' generic-proxy writes
      umma::fence_proxy_async_smem();                                // ... and hands them to the async proxy (the MMA)
      umma::tc_fence_after();
      if (lane == 0 && (g.debug & 1)) {
        umma::mma_commit(bars->tile_empty + s);
        if (kb == nkb - 1) umma::mma_commit(&bars->acc_full);
      } else if (lane == 0) {
        const uint32_t sa = tiles + s * kTileBytes;
        const uint64_t a_hi = smem_desc_mn_sw128(sa), a_lo = smem_desc_mn_sw128(sa + kTileA);
        const uint64_t b_hi = smem_desc_mn_sw128(sa + 2 * kTileA), b_lo = smem_desc_mn_sw128(sa + 2 * kTileA + kTileB);
        if (wide) {
          // the dy_lo tile follows the dy_hi tile group by group: [dy_hi | dy_lo] is ONE MN-major operand of 2 * ncols columns.
          // Per K step x_hi^T [dy_hi | dy_lo] -> columns [hi*hi | hi*lo] (N = 128), then x_lo^T dy_hi added onto the second
          // half: 8 MMAs per node block instead of 12 (the issue chain is what this kernel waits for, see linear_tc.cu)
#pragma unroll
          for (int kk = 0; kk < BNODE / 8; ++kk) {
            const uint64_t adv = uint64_t(kk * kAtom >> 4);
            umma::mma_tf32(tmem_base + uint32_t(kk) * kAccCols, a_hi + adv, b_hi + adv, idesc_w, kb > 0);
          }
#pragma unroll
          for (int kk = 0; kk < BNODE / 8; ++kk) {
            const uint64_t adv = uint64_t(kk * kAtom >> 4);
            umma::mma_tf32(tmem_base + uint32_t(kk) * kAccCols + uint32_t(ncols), a_lo + adv, b_hi + adv, idesc, true);
          }
        } else
#pragma unroll
        for (int term = 0; term < 3; ++term) {
#pragma unroll
          for (int kk = 0; kk < BNODE / 8; ++kk) {  // term-major: consecutive MMAs target different accumulators
            const uint32_t acc = tmem_base + uint32_t(kk) * kAccCols;
            const uint64_t adv = uint64_t(kk * kAtom >> 4);  // next 8 nodes
            if (term == 0) umma::mma_tf32(acc, a_lo + adv, b_hi + adv, idesc, kb > 0);
            else if (term == 1) umma::mma_tf32(acc, a_hi + adv, b_lo + adv, idesc, true);
            else umma::mma_tf32(acc, a_hi + adv, b_hi + adv, idesc, true);
          }
        }
        umma::mma_commit(bars->tile_empty + s);
        if (kb == nkb - 1) umma::mma_commit(&bars->acc_full);
      }
      __syncwarp();
    }
  } else if (warp >= kWarpConv) {
    // ---- stagers: group grp stages the node blocks q = grp (mod kGroups); inside a group a warp owns NPW nodes of
    // the block with all their column groups of x and dy.  One batch of loads per block and thread, consumed in the
    // same iteration: the groups are in different phases of load / wait / convert, which is what overlaps the memory
    // latency.  (Register prefetch across iterations does not: the compiler tracks all loads of the loop with one
    // scoreboard, so the first use of the oldest batch waits for the newest as well.)
    constexpr int kGroups = stager_groups<MG>(), WPG = kConvWarps / kGroups, NPW = BNODE / WPG;
    constexpr int RX = NPW * 4, RY = NPW * MG;
    const int cw = warp - kWarpConv;
    const int grp = cw / WPG, wg = cw % WPG;
    uint32_t okx = 0, oky = 0, onex = 0;  // per column group: inside the matrix / the ones column that yields db
#pragma unroll
    for (int cg = 0; cg < 4; ++cg) {
      const int64_t col = j0 + cg * 32 + lane;
      okx |= uint32_t(col < g.k) << cg;
      onex |= uint32_t(g.fuse_db && col == g.k) << cg;
    }
#pragma unroll
    for (int cg = 0; cg < MG; ++cg) oky |= uint32_t(cg * 32 + lane < g.m) << cg;
    const bool cols_full = (j0 + BF <= g.k) && (g.m == MG * 32);
    const uint32_t sw = uint32_t(lane) >> 3, lo3 = (uint32_t(lane) & 7u) << 2;
#pragma unroll 1
    for (int q = grp; q < nkb; q += kGroups) {
      const int64_t nb = int64_t(q) * BNODE + wg * NPW;  // first node of this warp, relative to the slab
      const float* xb = g.x + (node0 + nb) * g.k + j0 + lane;
      const float* yb = g.dy + (node0 + nb) * g.m + lane;
      float vx[RX], vy[RY];
      if (g.debug & 4) {
#pragma unroll
        for (int r = 0; r < RX; ++r) vx[r] = 0.f;
#pragma unroll
        for (int r = 0; r < RY; ++r) vy[r] = 0.f;
      } else if (cols_full && nb + NPW <= slab) {
#pragma unroll
        for (int r = 0; r < RX; ++r) vx[r] = __ldg(xb + (r >> 2) * g.k + (r & 3) * 32);
#pragma unroll
        for (int r = 0; r < RY; ++r) vy[r] = __ldg(yb + (r / MG) * g.m + (r % MG) * 32);
      } else {
#pragma unroll
        for (int r = 0; r < RX; ++r) {
          const bool nv = nb + (r >> 2) < slab;
          vx[r] = (((okx >> (r & 3)) & 1) && nv) ? __ldg(xb + (r >> 2) * g.k + (r & 3) * 32)
                                                 : (((onex >> (r & 3)) & 1) && nv) ? 1.f : 0.f;
        }
#pragma unroll
        for (int r = 0; r < RY; ++r)
          vy[r] = (((oky >> (r % MG)) & 1) && nb + r / MG < slab) ? __ldg(yb + (r / MG) * g.m + (r % MG) * 32) : 0.f;
      }
      const int ts = q % kTileStages;
      umma::mbar_wait(bars->tile_empty + ts, uint32_t((q / kTileStages) & 1) ^ 1);
      const uint32_t tile = tiles + ts * kTileBytes;
      if (!(g.debug & 2)) {
#pragma unroll
        for (int r = 0; r < RX; ++r) {
          const uint32_t node = uint32_t(wg * NPW + (r >> 2));
          const uint32_t to = tile + (r & 3) * kGroupStride + node * 128u + (((sw ^ node) & 3u) << 5) + lo3;
          float hi, lo;
          split_mask(vx[r], hi, lo);
          umma::sts_f32(to, hi);
          umma::sts_f32(to + kTileA, lo);
        }
#pragma unroll
        for (int r = 0; r < RY; ++r) {
          const uint32_t node = uint32_t(wg * NPW + r / MG);
          const uint32_t to = tile + 2 * kTileA + (r % MG) * kGroupStride + node * 128u + (((sw ^ node) & 3u) << 5) + lo3;
          float hi, lo;
          split_mask(vy[r], hi, lo);
          umma::sts_f32(to, hi);
          umma::sts_f32(to + kTileB, lo);
        }
      }
      // no proxy fence here (it would compile to a MEMBAR.ALL.CTA per block): the MMA thread fences after its wait
      __syncwarp();
      if (lane == 0) umma::mbar_arrive(bars->tile_full + ts);
    }
  } else {
    // ---- epilogue: warp w owns TMEM lanes [32w, 32w+32) = features j0 + 32w + lane; per output column the 32
    // lanes store 32 consecutive floats of a row of dW ----
    umma::mbar_wait_relaxed(&bars->acc_full, 0);
    umma::tc_fence_after();
    const int64_t f = j0 + warp * 32 + lane;
    float* dst = g.partial + sl * (g.k * g.m + (g.fuse_db ? g.m : 0));
#pragma unroll 1
    for (int c = 0; c < ncols / 16; ++c) {
      float v[16];
      umma::tmem_ld16(tmem_base + (uint32_t(warp * 32) << 16) + c * 16, v);
#pragma unroll
      for (int a = 1; a < kAcc; ++a) {
        float u[16];
        umma::tmem_ld16(tmem_base + (uint32_t(warp * 32) << 16) + a * kAccCols + c * 16, u);
#pragma unroll
        for (int t = 0; t < 16; ++t) v[t] += u[t];
      }
      if (wide) {  // the small terms: second half of every accumulator
#pragma unroll
        for (int a = 0; a < kAcc; ++a) {
          float u[16];
          umma::tmem_ld16(tmem_base + (uint32_t(warp * 32) << 16) + a * kAccCols + ncols + c * 16, u);
#pragma unroll
          for (int t = 0; t < 16; ++t) v[t] += u[t];
        }
      }
#pragma unroll
      for (int t = 0; t < 16; ++t) {
        const int64_t mc = c * 16 + t;
        if (mc < g.m) {
          if (f < g.k) dst[mc * g.k + f] = v[t];
          else if (f == g.k && g.fuse_db) dst[g.k * g.m + mc] = v[t];
        }
      }
    }
    umma::tc_fence_before();
  }
  umma::tc_fence_before();
  __syncthreads();
  if (warp == kWarpMma) umma::tmem_dealloc(tmem_base, kTmemCols);
}

}  // namespace

// number of node slabs the tensor-core weight gradient uses (0 = shape not supported by this kernel)
int grad_w_tc_slabs(int64_t n, int64_t k, int64_t m, int64_t* slab_nodes) {
  if (m > 128 || k < 2 * BF || n < 2048) return 0;
  const int64_t ntile = ceil_div(k, BF);
  if (ntile > kNumSMs) return 0;
  int64_t s = kNumSMs / ntile;
  const int64_t sn = align_up(ceil_div(n, s), BNODE);
  if (slab_nodes) *slab_nodes = sn;
  return int(ceil_div(n, sn));
}

// the bias gradient comes out of the same launch when the last feature tile has a free row
bool grad_w_tc_fuses_db(int64_t k) { return k % BF != 0; }

// partial[s][m*K + k | m] for s < grad_w_tc_slabs(n, k, m); the trailing m floats (db) only with fuse_db
int launch_grad_w_tc(const float* dy, const float* x, float* partial, int64_t n, int64_t k, int64_t m, int fuse_db,
                     cudaStream_t stream) {
  int64_t slab_nodes = 0;
  const int slabs = grad_w_tc_slabs(n, k, m, &slab_nodes);
  if (slabs <= 0 || (fuse_db && !grad_w_tc_fuses_db(k))) {
    set_error("grad_w_tc: unsupported shape n=%lld k=%lld m=%lld", (long long)n, (long long)k, (long long)m);
    return LGK_ERR_UNSUPPORTED;
  }
  // the attribute is per device: one flag per device ordinal
  static bool configured_dev[64] = {};
  int dev_ord = 0;
  cudaGetDevice(&dev_ord);
  bool& configured = configured_dev[dev_ord & 63];
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(grad_w_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes<2>());
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(grad_w_tc_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes<4>());
    if (e != cudaSuccess) {
      set_error("grad_w_tc_kernel: cannot reserve shared memory: %s", cudaGetErrorString(e));
      return LGK_ERR_LAUNCH;
    }
    configured = true;
  }
  static const int debug = [] { const char* v = getenv("LGK_GW_DEBUG"); return (v && *v) ? atoi(v) : 0; }();
  static const int wide = [] { const char* v = getenv("LGK_GW_WIDE"); return (v && *v) ? atoi(v) : 1; }();
  GwArgs g{dy, x, partial, n, k, m, slab_nodes, fuse_db, debug, wide};
  const unsigned grid = unsigned(ceil_div(k, BF) * slabs);
  if (m <= 64) grad_w_tc_kernel<2><<<grid, kThreads, smem_bytes<2>(), stream>>>(g);
  else grad_w_tc_kernel<4><<<grid, kThreads, smem_bytes<4>(), stream>>>(g);
  LGK_CHECK_LAUNCH("grad_w_tc_kernel");
  return LGK_OK;
}

}  // namespace lgk
