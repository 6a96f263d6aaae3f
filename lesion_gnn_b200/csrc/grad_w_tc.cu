// (5c) weight gradient of a wide node-feature transform on the tensor cores: dW[m,k] = sum_i dy[i,m] x[i,k].
//
// The gradient of the reference's in_proj (models/gat.py:29, gin.py:21: Linear(1025 -> d)) reduces over the nodes
// of the batch: 2*N*1025*d flops over N*(1025 + d)*4 bytes.  On the FFMA pipe it was the longest kernel of the
// training step (72 us on the headline batch, 9 % of the HBM roofline).  Here both operands are read exactly as
// they lie in memory: with the node index as the MMA's K dimension, dy [N, m] IS the MN-major A operand
// (m contiguous) and x [N, k] IS the MN-major B operand (features contiguous) -- no transposition anywhere, and
// a warp that copies 32 consecutive features of one node writes exactly one 128-byte row of a SWIZZLE_128B
// MN-major atom (8 nodes x 32 features), conflict free.  3xTF32 as in linear_tc.cu (hi/lo split on the fly, three
// MMAs per K step, four independent accumulators in tensor memory added in fp32 by the epilogue).
//
// Grid: feature tiles of 128 columns of x  x  slabs of nodes, ntiles * slabs <= 148 CTAs (one per SM, one wave).
// A CTA walks its slab in blocks of 32 nodes: 16 stager warps cp.async (4 bytes per lane: the reference's
// 1025-float rows are not 16-byte aligned) the raw block into a 3-deep ring, read back their own bytes, split and
// write the hi/lo tiles; one lane issues 12 tcgen05.mma (M128 N128 K8) per block; at the end 4 warps drain the
// accumulators into partial[slab][m*K + k] (the layout of dW), which lgk_linear_bwd_w reduces in slab order
// (deterministic).
// m <= 128 (rows of dW above m are zero rows of the A tile, written once).
#include "common.cuh"
#include "umma.cuh"

namespace lgk {
namespace {

constexpr int BF = 128;          // features (columns of x) per CTA = MMA N
constexpr int BMM = 128;         // MMA M (rows of dW, zero padded above m)
constexpr int BNODE = 32;        // nodes per block = 4 MMA K steps of 8
constexpr int kAcc = 4;
constexpr int kRawStages = 3, kTileStages = 2;
constexpr int kAtom = 1024;                   // 8 nodes (one MMA K step) x 128 bytes
constexpr int kGroupStride = 4 * kAtom;       // one 32-column group: 4 node atoms back to back
constexpr int kTileA = 4 * kGroupStride, kTileB = 4 * kGroupStride;  // hi (or lo) operand tile: 16 KB each
constexpr int kTileBytes = 2 * (kTileA + kTileB);
constexpr int kConvWarps = 16;
constexpr int kWarpMma = 4, kWarpConv = 5;
constexpr int kThreads = (kWarpConv + kConvWarps) * 32;
constexpr int kTmemCols = kAcc * BF;
static_assert(kTmemCols == 512, "tensor memory allocation must be a power of two");

template <int MG> constexpr int raw_bytes() { return (BF / 32 + MG) * BNODE * 128; }
template <int MG> constexpr int smem_bytes() { return kTileStages * kTileBytes + kRawStages * raw_bytes<MG>() + 256; }
static_assert(smem_bytes<4>() <= 227 * 1024, "shared memory budget");

struct GwArgs {
  const float* dy;
  const float* x;
  float* partial;
  int64_t n, k, m;
  int64_t slab_nodes;  // multiple of 32
};

struct Bars {
  uint64_t tile_full[kTileStages], tile_empty[kTileStages];
  uint64_t acc_full;
  uint32_t tmem_slot;
};

__device__ __forceinline__ void split_mask(float x, float& hi, float& lo) {
  hi = __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);
  lo = x - hi;
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const float* src, uint32_t src_bytes) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const float* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

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
constexpr uint32_t idesc_tf32_mn(int m, int n) { return umma::idesc_tf32(m, n) | (1u << 15) | (1u << 16); }  // A, B MN-major

template <int MG>
__global__ void __launch_bounds__(kThreads, 1) grad_w_tc_kernel(GwArgs g) {
  constexpr int kRaw = raw_bytes<MG>();
  constexpr int RX = 8, RY = 2 * MG;  // (node, column group) slots per stager warp: x and dy
  extern __shared__ __align__(1024) uint8_t smem_dyn[];
  const uint32_t tiles = umma::smem_u32(smem_dyn);
  const uint32_t raws = tiles + kTileStages * kTileBytes;
  Bars* bars = reinterpret_cast<Bars*>(smem_dyn + kTileStages * kTileBytes + kRawStages * kRaw);
  if ((tiles & 1023u) != 0) __trap();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t ntile = (g.k + BF - 1) / BF;
  const int64_t jt = blockIdx.x % ntile, sl = blockIdx.x / ntile;
  const int64_t j0 = jt * BF;
  const int64_t node0 = sl * g.slab_nodes;
  const int64_t slab = min(g.slab_nodes, g.n - node0);   // nodes of this slab (> 0 by construction of the grid)
  const int nkb = int((slab + BNODE - 1) / BNODE);

  // operand tiles start as zeros: rows of the A tile above m are never written
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
    constexpr uint32_t idesc = idesc_tf32_mn(BMM, BF);
    for (int kb = 0; kb < nkb; ++kb) {
      const int s = kb % kTileStages;
      umma::mbar_wait(bars->tile_full + s, (kb / kTileStages) & 1);
      umma::tc_fence_after();
      if (lane == 0) {
        const uint32_t sa = tiles + s * kTileBytes;
        const uint64_t a_hi = smem_desc_mn_sw128(sa), a_lo = smem_desc_mn_sw128(sa + kTileA);
        const uint64_t b_hi = smem_desc_mn_sw128(sa + 2 * kTileA), b_lo = smem_desc_mn_sw128(sa + 2 * kTileA + kTileB);
#pragma unroll
        for (int term = 0; term < 3; ++term) {
#pragma unroll
          for (int kk = 0; kk < BNODE / 8; ++kk) {  // term-major: consecutive MMAs target different accumulators
            const uint32_t acc = tmem_base + uint32_t(kk) * BF;
            const uint64_t adv = uint64_t(kk * kAtom >> 4);  // next 8 nodes: next atom
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
    const float* xp[RX];
    const float* yp[RY];
    uint32_t tx[RX], ty[RY];
    uint32_t okx = 0, oky = 0;  // column-valid bits per slot
#pragma unroll
    for (int r = 0; r < RX; ++r) {
      const uint32_t node = 2 * cw + (r >> 2), cg = r & 3;
      const int64_t col = j0 + cg * 32 + lane;
      const bool ok = col < g.k;
      okx |= uint32_t(ok) << r;
      xp[r] = g.x + (node0 + node) * g.k + (ok ? col : 0);
      tx[r] = mn_offset(node, cg, lane);
    }
#pragma unroll
    for (int r = 0; r < RY; ++r) {
      const uint32_t node = 2 * cw + r / MG, cg = r % MG;
      const int64_t col = cg * 32 + lane;
      const bool ok = col < g.m;
      oky |= uint32_t(ok) << r;
      yp[r] = g.dy + (node0 + node) * g.m + (ok ? col : 0);
      ty[r] = mn_offset(node, cg, lane);
    }
    const uint32_t raw_x = uint32_t(cw * RX) * 128 + lane * 4;
    const uint32_t raw_y = uint32_t(BF / 32 * BNODE + cw * RY) * 128 + lane * 4;
    const bool cols_full = (j0 + BF <= g.k) && (g.m == MG * 32);
    const int64_t step_x = int64_t(BNODE) * g.k, step_y = int64_t(BNODE) * g.m;

    auto issue = [&](int q) {
      if (q < nkb) {
        const uint32_t raw = raws + uint32_t(q % kRawStages) * kRaw;
        const int64_t left = slab - int64_t(q) * BNODE;  // nodes of the slab from this block on
        if (cols_full && left >= BNODE) {
#pragma unroll
          for (int r = 0; r < RX; ++r) cp_async4(raw + raw_x + r * 128, xp[r] + q * step_x);
#pragma unroll
          for (int r = 0; r < RY; ++r) cp_async4(raw + raw_y + r * 128, yp[r] + q * step_y);
        } else {
          const bool v0 = 2 * cw < left, v1 = 2 * cw + 1 < left;
#pragma unroll
          for (int r = 0; r < RX; ++r) {
            const bool ok = ((okx >> r) & 1) && ((r >> 2) ? v1 : v0);
            cp_async4(raw + raw_x + r * 128, ok ? xp[r] + q * step_x : g.x, ok ? 4u : 0u);
          }
#pragma unroll
          for (int r = 0; r < RY; ++r) {
            const bool ok = ((oky >> r) & 1) && ((r / MG) ? v1 : v0);
            cp_async4(raw + raw_y + r * 128, ok ? yp[r] + q * step_y : g.dy, ok ? 4u : 0u);
          }
        }
      }
      cp_async_commit();
    };
    for (int q = 0; q < kRawStages; ++q) issue(q);
    for (int q = 0; q < nkb; ++q) {
      cp_async_wait<kRawStages - 1>();
      const int ts = q % kTileStages;
      umma::mbar_wait(bars->tile_empty + ts, uint32_t((q / kTileStages) & 1) ^ 1);
      const uint32_t raw = raws + uint32_t(q % kRawStages) * kRaw, tile = tiles + ts * kTileBytes;
      float vx[RX], vy[RY];
#pragma unroll
      for (int r = 0; r < RX; ++r) vx[r] = umma::lds_f32(raw + raw_x + r * 128);
#pragma unroll
      for (int r = 0; r < RY; ++r) vy[r] = umma::lds_f32(raw + raw_y + r * 128);
#pragma unroll
      for (int r = 0; r < RY; ++r) {
        float hi, lo;
        split_mask(vy[r], hi, lo);
        umma::sts_f32(tile + ty[r], hi);
        umma::sts_f32(tile + kTileA + ty[r], lo);
      }
#pragma unroll
      for (int r = 0; r < RX; ++r) {
        float hi, lo;
        split_mask(vx[r], hi, lo);
        umma::sts_f32(tile + 2 * kTileA + tx[r], hi);
        umma::sts_f32(tile + 2 * kTileA + kTileB + tx[r], lo);
      }
      umma::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) umma::mbar_arrive(bars->tile_full + ts);
      issue(q + kRawStages);
    }
  } else {
    // ---- epilogue: warp w owns TMEM lanes [32w, 32w+32) = rows of dW; a thread writes runs of 16 columns ----
    umma::mbar_wait_relaxed(&bars->acc_full, 0);
    umma::tc_fence_after();
    const int64_t mrow = warp * 32 + lane;
    if (warp * 32 < g.m) {
      float* dst = g.partial + sl * g.k * g.m;
#pragma unroll 1
      for (int c = 0; c < BF / 16; ++c) {
        float v[16];
        umma::tmem_ld16(tmem_base + (uint32_t(warp * 32) << 16) + c * 16, v);
#pragma unroll
        for (int a = 1; a < kAcc; ++a) {
          float u[16];
          umma::tmem_ld16(tmem_base + (uint32_t(warp * 32) << 16) + a * BF + c * 16, u);
#pragma unroll
          for (int t = 0; t < 16; ++t) v[t] += u[t];
        }
        if (mrow < g.m) {
#pragma unroll
          for (int t = 0; t < 16; ++t) {
            const int64_t col = j0 + c * 16 + t;
            if (col < g.k) dst[mrow * g.k + col] = v[t];
          }
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
  if (m > BMM || k < 2 * BF || n < 2048) return 0;
  const int64_t ntile = ceil_div(k, BF);
  if (ntile > kNumSMs) return 0;
  int64_t s = kNumSMs / ntile;
  const int64_t sn = align_up(ceil_div(n, s), BNODE);
  if (slab_nodes) *slab_nodes = sn;
  return int(ceil_div(n, sn));
}

// partial[s][m*K + k] for s < grad_w_tc_slabs(n, k, m)
int launch_grad_w_tc(const float* dy, const float* x, float* partial, int64_t n, int64_t k, int64_t m, cudaStream_t stream) {
  int64_t slab_nodes = 0;
  const int slabs = grad_w_tc_slabs(n, k, m, &slab_nodes);
  if (slabs <= 0) {
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
  GwArgs g{dy, x, partial, n, k, m, slab_nodes};
  const unsigned grid = unsigned(ceil_div(k, BF) * slabs);
  if (m <= 64) grad_w_tc_kernel<2><<<grid, kThreads, smem_bytes<2>(), stream>>>(g);
  else grad_w_tc_kernel<4><<<grid, kThreads, smem_bytes<4>(), stream>>>(g);
  LGK_CHECK_LAUNCH("grad_w_tc_kernel");
  return LGK_OK;
}

}  // namespace lgk
