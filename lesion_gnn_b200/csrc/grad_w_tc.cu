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
// A CTA walks its slab in blocks of 32 nodes: 16 stager warps load the block (4 bytes per lane, 128 bytes per warp
// load: the reference's 1025-float rows are not 16-byte aligned) into registers two blocks ahead, split it and write
// the hi/lo tiles of a 3-deep ring (4-byte cp.async into a raw ring, as linear_tc.cu does, measured 56 us against
// 41 us: the LDGSTS issue rate, ~10 cycles per warp instruction, was the cost); one lane issues 12 tcgen05.mma
// (M128 N=m K8) per block; at the end 4 warps drain the
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
template <int MG> constexpr int tile_stages() { return 3; }
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
};

struct Bars {
  uint64_t tile_full[3], tile_empty[3];
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
// byte offset of (node r of the block, column group cg, column lane) inside an operand tile
__device__ __forceinline__ uint32_t mn_offset(uint32_t r, uint32_t cg, uint32_t lane) {
  return cg * kGroupStride + r * 128u + ((((lane >> 3) ^ r) & 3u) << 5) + ((lane & 7u) << 2);
}
__host__ __device__ constexpr uint32_t idesc_tf32_mn(int m, int n) { return umma::idesc_tf32(m, n) | (1u << 15) | (1u << 16); }  // A, B MN-major

template <int MG>
__global__ void __launch_bounds__(kThreads, 1) grad_w_tc_kernel(GwArgs g) {
  constexpr int kTileB = tile_b<MG>(), kTileBytes = tile_bytes<MG>(), kTileStages = tile_stages<MG>();
  constexpr int RX = 8, RY = 2 * MG;  // (node, column group) slots per stager warp: x and dy
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

  // operand tiles start as zeros: columns of the dy tiles between m and N are never written
  for (int i = threadIdx.x; i < kTileStages * kTileBytes / 16; i += kThreads)
    *reinterpret_cast<uint4*>(smem_dyn + i * 16) = make_uint4(0, 0, 0, 0);
  umma::fence_proxy_async_smem();
  if (threadIdx.x == 0) {
    for (int s = 0; s < kTileStages; ++s) {
      umma::mbar_init(bars->tile_full + s, kConvWarps);
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
    const uint32_t idesc = idesc_tf32_mn(BF, ncols);
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
    // ---- stagers: warp cw owns nodes 2cw and 2cw+1 of every block (all their column groups of x and dy) ----
    const int cw = warp - kWarpConv;
    // slot r of x: node 2cw + (r >> 2), column group r & 3; of dy: node 2cw + r / MG, column group r % MG.  One base
    // pointer and one tile offset per node parity instead of one per slot: the registers go to the staged values.
    uint32_t okx = 0, oky = 0, onex = 0;  // per column group: inside the matrix / the ones column that yields db
#pragma unroll
    for (int cg = 0; cg < 4; ++cg) {
      const int64_t col = j0 + cg * 32 + lane;
      okx |= uint32_t(col < g.k) << cg;
      onex |= uint32_t(g.fuse_db && col == g.k) << cg;
    }
#pragma unroll
    for (int cg = 0; cg < MG; ++cg) oky |= uint32_t(cg * 32 + lane < g.m) << cg;
    const float* xb = g.x + (node0 + 2 * cw) * g.k + j0 + lane;
    const float* yb = g.dy + (node0 + 2 * cw) * g.m + lane;
    const uint32_t t0 = mn_offset(2 * cw, 0, lane), t1 = mn_offset(2 * cw + 1, 0, lane);
    const bool cols_full = (j0 + BF <= g.k) && (g.m == MG * 32);
    const int64_t step_x = int64_t(BNODE) * g.k, step_y = int64_t(BNODE) * g.m;

    // global -> registers one block ahead of the conversion (coalesced 128-byte warp loads; zeros outside the matrix)
    auto load = [&](int q, float (&vx)[RX], float (&vy)[RY]) {
      const int64_t left = slab - int64_t(q) * BNODE;  // nodes of the slab from this block on
      if (q >= nkb || (g.debug & 4)) {
#pragma unroll
        for (int r = 0; r < RX; ++r) vx[r] = 0.f;
#pragma unroll
        for (int r = 0; r < RY; ++r) vy[r] = 0.f;
      } else if (cols_full && left >= BNODE) {
#pragma unroll
        for (int r = 0; r < RX; ++r) vx[r] = __ldg(xb + q * step_x + (r >> 2) * g.k + (r & 3) * 32);
#pragma unroll
        for (int r = 0; r < RY; ++r) vy[r] = __ldg(yb + q * step_y + (r / MG) * g.m + (r % MG) * 32);
      } else {
        const bool v0 = 2 * cw < left, v1 = 2 * cw + 1 < left;
#pragma unroll
        for (int r = 0; r < RX; ++r) {
          const bool nv = (r >> 2) ? v1 : v0;
          vx[r] = (((okx >> (r & 3)) & 1) && nv) ? __ldg(xb + q * step_x + (r >> 2) * g.k + (r & 3) * 32)
                                                 : (((onex >> (r & 3)) & 1) && nv) ? 1.f : 0.f;
        }
#pragma unroll
        for (int r = 0; r < RY; ++r)
          vy[r] = (((oky >> (r % MG)) & 1) && ((r / MG) ? v1 : v0)) ? __ldg(yb + q * step_y + (r / MG) * g.m + (r % MG) * 32) : 0.f;
      }
    };
    // Loads run two blocks ahead of the conversion; the loop is unrolled three times with the register sets changing
    // roles (copying a set whose loads are in flight would stall on them and serialise the loop on memory latency).
    auto step = [&](int q, const float (&vx)[RX], const float (&vy)[RY]) {
      if (q >= nkb) return;
      const int ts = q % kTileStages;
      umma::mbar_wait(bars->tile_empty + ts, uint32_t((q / kTileStages) & 1) ^ 1);
      const uint32_t tile = tiles + ts * kTileBytes;
      if (!(g.debug & 2)) {
#pragma unroll
        for (int r = 0; r < RX; ++r) {
          float hi, lo;
          split_mask(vx[r], hi, lo);
          const uint32_t to = ((r >> 2) ? t1 : t0) + (r & 3) * kGroupStride;
          umma::sts_f32(tile + to, hi);
          umma::sts_f32(tile + kTileA + to, lo);
        }
#pragma unroll
        for (int r = 0; r < RY; ++r) {
          float hi, lo;
          split_mask(vy[r], hi, lo);
          const uint32_t to = ((r / MG) ? t1 : t0) + (r % MG) * kGroupStride;
          umma::sts_f32(tile + 2 * kTileA + to, hi);
          umma::sts_f32(tile + 2 * kTileA + kTileB + to, lo);
        }
      }
      // no proxy fence here: fence.proxy.async compiles to MEMBAR.ALL.CTA + FENCE.VIEW.ASYNC, and the membar would
      // wait for this thread's prefetched loads of the next two blocks -- the MMA thread fences after its wait
      __syncwarp();
      if (lane == 0) umma::mbar_arrive(bars->tile_full + ts);
    };
    float x0[RX], y0[RY], x1[RX], y1[RY], x2[RX], y2[RY];
    load(0, x0, y0);
    load(1, x1, y1);
#pragma unroll 1
    for (int q = 0; q < nkb; q += 3) {
      load(q + 2, x2, y2);
      step(q, x0, y0);
      load(q + 3, x0, y0);
      step(q + 1, x1, y1);
      load(q + 4, x1, y1);
      step(q + 2, x2, y2);
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
  static bool configured = false;
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
  GwArgs g{dy, x, partial, n, k, m, slab_nodes, fuse_db, debug};
  const unsigned grid = unsigned(ceil_div(k, BF) * slabs);
  if (m <= 64) grad_w_tc_kernel<2><<<grid, kThreads, smem_bytes<2>(), stream>>>(g);
  else grad_w_tc_kernel<4><<<grid, kThreads, smem_bytes<4>(), stream>>>(g);
  LGK_CHECK_LAUNCH("grad_w_tc_kernel");
  return LGK_OK;
}

}  // namespace lgk
