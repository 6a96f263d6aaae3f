// (5) node-feature transform on the 5th-generation tensor cores, TMA-fed variant for 16-byte aligned rows (K % 4 == 0:
// every hidden transform of the reference's models -- 64 -> 64, 128 -> 128 -- and any aligned in_proj).
// out = act(rowscale * x W^T + bias [+ out]), 3xTF32 as in linear_tc.cu (x = hi + lo, three MMAs per K step into four
// TMEM accumulators, fp32 parity 2e-6).
//
// With aligned rows [N, K] is a legal tensor map: one TMA box {32 floats, 128 rows} of x and one {32 floats, 64 rows}
// of W per K block land in shared memory directly in the K-major SWIZZLE_128B layout the tensor core reads; rows
// beyond N / M and columns beyond K are zero-filled by the TMA unit (no masks anywhere).  The hi / lo split is then an
// elementwise pass at the SAME byte offset (the swizzle is a function of the address): 16 bytes per thread and step.
//
// Persistent kernel, one CTA per SM, 22 warps in four roles:
//   producer (1 lane)     2 tensor-map TMA loads per 32-wide K block into a 3-deep raw ring, expect_tx / complete_tx
//   splitters (16 warps)  raw tile -> hi and lo operand tiles, fence.proxy.async, arrive
//   MMA (1 lane)          12 tcgen05.mma.kind::tf32 (M128 N64 K8) per K block, commit -> frees the operand tiles
//   epilogue (4 warps)    tcgen05.ld of the finished accumulators, row scale / bias / accumulate / ELU, store;
//                         overlaps the next tile's main loop (two accumulator buffers in tensor memory)
// Measured (profiles/README.md): 1.28 M x 128 -> 128 in 592 us (cp.async kernel 803 us), of which 437 us remain with
// the MMAs, the split and the loads all switched off: the per-K-block barrier hand-offs, not the data path, are
// what is left to remove (wider K blocks with an in-place split).
#include "common.cuh"
#include "umma.cuh"

#include <cuda.h>  // CUtensorMap (types only: the encoder is fetched through cudaGetDriverEntryPoint)
#include <stdlib.h>

namespace lgk {
namespace {

constexpr int BM = 128, BN = 64, BK = 32, kAcc = 4;
constexpr int kRawStages = 3, kTileStages = 3;
constexpr int kRawA = BM * BK * 4, kRawB = BN * BK * 4;
constexpr int kRawBytes = kRawA + kRawB;                       // one raw K block: x tile then W tile
constexpr int kTileA = BM * BK * 4, kTileB = BN * BK * 4;      // bytes of one hi (or lo) operand tile
constexpr int kTileBytes = 2 * (kTileA + kTileB);              // A_hi, A_lo, B_hi, B_lo
constexpr int kSmemBytes = kTileStages * kTileBytes + kRawStages * kRawBytes + 256 /*barriers*/;
constexpr int kConvWarps = 16;
constexpr int kWarpMma = 4, kWarpProd = 5, kWarpConv = 6;      // warps 0-3: epilogue (TMEM lane quadrant = warp id)
constexpr int kThreadsTc = (kWarpConv + kConvWarps) * 32;
constexpr int kTmemCols = 2 * kAcc * BN;
static_assert(BK / 8 == kAcc, "one accumulator per K step of a block");
static_assert(kTmemCols == 512, "tensor memory allocation must be a power of two >= 32 columns");
static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");
static_assert((kRawA + kRawB) / 16 == 3 * kConvWarps * 32, "three 16-byte pieces per splitter thread and block");

struct TcArgs {
  const float* bias;
  const float* rowscale;
  float* out;
  int64_t n, k, m;
  int act, accumulate;
  int64_t tiles_n;      // column tiles (fastest tile index)
  int64_t tiles_total;
  int debug;            // profiling only (LGK_TC_DEBUG bit mask): 1 = no MMAs, 2 = no split, 4 = no TMA loads
  // GATConv.lin: the attention logit halves a_src[n, H] = <h[n, h, :], att_src[h, :]> (and a_dst) from the fp32
  // accumulators, in the epilogue -- no separate pass over h.  chan divides BN, so a tile holds whole heads.
  const float* att_src;
  const float* att_dst;
  float* a_src;
  float* a_dst;
  int heads, chan;
};

struct Bars {
  uint64_t raw_full[kRawStages], raw_empty[kRawStages];
  uint64_t tile_full[kTileStages], tile_empty[kTileStages];
  uint64_t acc_full[2], acc_empty[2];
  uint32_t tmem_slot;
};

// ---- small PTX helpers ----------------------------------------------------------------------------------------
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
      ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(umma::smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}

// hi = x with the 13 low mantissa bits cleared (exactly representable in TF32), lo = x - hi (exact)
__device__ __forceinline__ void split_mask(float x, float& hi, float& lo) {
  hi = __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);
  lo = x - hi;
}

__device__ __forceinline__ float4 lds_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_f4(uint32_t addr, float4 v) {
  asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// one 16-byte piece of a raw tile -> the same byte offset of the hi and the lo tile
__device__ __forceinline__ void split_piece(uint32_t raw, uint32_t hi_tile, uint32_t lo_tile, uint32_t off) {
  const float4 v = lds_f4(raw + off);
  float4 h, l;
  split_mask(v.x, h.x, l.x);
  split_mask(v.y, h.y, l.y);
  split_mask(v.z, h.z, l.z);
  split_mask(v.w, h.w, l.w);
  sts_f4(hi_tile + off, h);
  sts_f4(lo_tile + off, l);
}

// ---- epilogue -------------------------------------------------------------------------------------------------
__device__ __forceinline__ void epilogue_tile(const TcArgs& g, uint32_t tmem_acc, int nacc, int warp, int lane, int64_t i0,
                                              int64_t j0) {
  const int64_t i = i0 + warp * 32 + lane;
  const float rs = (g.rowscale && i < g.n) ? __ldg(g.rowscale + i) : 1.f;
  float as_acc = 0.f, ad_acc = 0.f;
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
    if (g.a_src) {  // chan = 16, 32 or 64: heads never straddle a tile or split a 16-column chunk; m = heads * chan
      const int jc = int(j0) + c * 16;
      if (jc < g.m) {
        if (jc % g.chan == 0) { as_acc = 0.f; ad_acc = 0.f; }
#pragma unroll
        for (int t = 0; t < 16; ++t) {
          as_acc = fmaf(v[t], __ldg(g.att_src + jc + t), as_acc);
          ad_acc = fmaf(v[t], __ldg(g.att_dst + jc + t), ad_acc);
        }
        if ((jc + 16) % g.chan == 0) {
          const int head = jc / g.chan;
          g.a_src[i * g.heads + head] = as_acc;
          g.a_dst[i * g.heads + head] = ad_acc;
        }
      }
    }
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
__global__ void __launch_bounds__(kThreadsTc, 1) linear_tma_kernel(TcArgs g, const __grid_constant__ CUtensorMap map_x,
                                                                  const __grid_constant__ CUtensorMap map_w) {
  extern __shared__ __align__(1024) uint8_t smem_dyn[];  // SWIZZLE_128B tiles need a 1024-byte aligned base
  const uint32_t tiles = umma::smem_u32(smem_dyn);        // operand tile ring
  const uint32_t raws = tiles + kTileStages * kTileBytes;  // raw K block ring
  Bars* bars = reinterpret_cast<Bars*>(smem_dyn + kTileStages * kTileBytes + kRawStages * kRawBytes);
  if ((tiles & 1023u) != 0) __trap();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nkb = int((g.k + BK - 1) / BK);
  // the K blocks of all tiles of this CTA form one sequence: block q = (tile q / nkb, K block q % nkb)
  const int64_t my_tiles = (g.tiles_total - blockIdx.x + gridDim.x - 1) / gridDim.x;
  const int64_t total = my_tiles * nkb;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kRawStages; ++s) {
      umma::mbar_init(bars->raw_full + s, 1);
      umma::mbar_init(bars->raw_empty + s, kConvWarps);
    }
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

  if (warp == kWarpProd) {
    // ---- TMA producer ----
    if (lane == 0) {
      if (!(g.debug & 8)) {
        tma_prefetch_desc(&map_x);
        tma_prefetch_desc(&map_w);
      }
      for (int64_t q = 0; q < total; ++q) {
        const int s = int(q % kRawStages);
        umma::mbar_wait(bars->raw_empty + s, uint32_t((q / kRawStages) & 1) ^ 1);
        const int64_t lt = q / nkb;
        const int k0 = int(q - lt * nkb) * BK;
        const int64_t t = blockIdx.x + lt * gridDim.x;
        const int r0 = int((t / g.tiles_n) * BM), c0 = int((t % g.tiles_n) * BN);
        const uint32_t raw = raws + uint32_t(s) * kRawBytes;
        if (g.debug & 4) {
          umma::mbar_arrive(bars->raw_full + s);
          continue;
        }
        umma::mbar_expect_tx(bars->raw_full + s, kRawBytes);
        umma::mbar_arrive(bars->raw_full + s);
        tma_load_2d(raw, &map_x, k0, r0, bars->raw_full + s);
        tma_load_2d(raw + kRawA, &map_w, k0, c0, bars->raw_full + s);
      }
    }
  } else if (warp == kWarpMma) {
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
          // The four K steps of a block go to four different accumulators and the issue order is term-major:
          // consecutive instructions target different accumulators, a dependent one follows four slots later.
          // Per accumulator the order of additions (lo*hi, hi*lo, hi*hi: small terms first) is fixed.
#pragma unroll
          for (int term = 0; term < 3; ++term) {
#pragma unroll
            for (int kk = 0; kk < BK / 8; ++kk) {
              const uint32_t acc = tmem_base + (ab * kAcc + uint32_t(kk)) * BN;
              const uint64_t adv = uint64_t(kk * 32 >> 4);  // +32 bytes along K inside the 128-byte swizzled row
              if (term == 0) umma::mma_tf32(acc, a_lo + adv, b_hi + adv, idesc, kb > 0);
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
    // ---- splitters: thread t handles the 16-byte pieces t and t + 512 of the x tile and piece t of the W tile ----
    const uint32_t t16 = uint32_t(threadIdx.x - kWarpConv * 32) * 16;
    for (int64_t q = 0; q < total; ++q) {
      const int rs = int(q % kRawStages), ts = int(q % kTileStages);
      umma::mbar_wait(bars->raw_full + rs, uint32_t((q / kRawStages) & 1));
      umma::mbar_wait(bars->tile_empty + ts, uint32_t((q / kTileStages) & 1) ^ 1);
      const uint32_t raw = raws + uint32_t(rs) * kRawBytes, tile = tiles + uint32_t(ts) * kTileBytes;
      if (!(g.debug & 2)) {
        split_piece(raw, tile, tile + kTileA, t16);
        split_piece(raw, tile, tile + kTileA, t16 + kConvWarps * 32 * 16);
        split_piece(raw + kRawA, tile + 2 * kTileA, tile + 2 * kTileA + kTileB, t16);
      }
      umma::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        umma::mbar_arrive(bars->tile_full + ts);
        umma::mbar_arrive(bars->raw_empty + rs);
      }
    }
  } else {
    // ---- epilogue: warp w < 4 owns TMEM lanes [32w, 32w+32) ----
    uint32_t li = 0;
    for (int64_t t = blockIdx.x; t < g.tiles_total; t += gridDim.x, ++li) {
      const int64_t i0 = (t / g.tiles_n) * BM, j0 = (t % g.tiles_n) * BN;
      const uint32_t ab = li & 1;
      umma::mbar_wait_relaxed(bars->acc_full + ab, (li >> 1) & 1);
      umma::tc_fence_after();
      epilogue_tile(g, tmem_base + ab * kAcc * BN, kAcc, warp, lane, i0, j0);
      umma::tc_fence_before();
      __syncwarp();
      if (lane == 0) umma::mbar_arrive(bars->acc_empty + ab);
    }
  }
  umma::tc_fence_before();
  __syncthreads();
  if (warp == kWarpMma) umma::tmem_dealloc(tmem_base, kTmemCols);
}

// ---- host side --------------------------------------------------------------------------------------------------
using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      p = nullptr;
    return reinterpret_cast<EncodeTiledFn>(p);
  }();
  return fn;
}

// [rows, k] fp32 row-major (k % 4 == 0, base 16-byte aligned) with boxes of {32 floats, box_rows rows}, SWIZZLE_128B,
// zero fill outside
bool make_map(CUtensorMap* map, const float* base, int64_t rows, int64_t k, int box_rows) {
  EncodeTiledFn enc = encode_tiled();
  if (!enc) return false;
  const cuuint64_t dims[2] = {cuuint64_t(k), cuuint64_t(rows)};
  const cuuint64_t strides[1] = {cuuint64_t(4 * k)};
  const cuuint32_t box[2] = {cuuint32_t(BK), cuuint32_t(box_rows)};
  const cuuint32_t estr[2] = {1, 1};
  return enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace

// Used by lgk_linear_fwd (linear.cu).  *done = false when the shape / alignment is not eligible (the caller then takes
// another kernel).
int launch_linear_tma(const float* x, const float* w, const float* bias, const float* rowscale, float* out, int64_t n,
                      int64_t k, int64_t m, int act, int accumulate, cudaStream_t stream, bool* done, const GatAlpha* alpha) {
  *done = false;
  if (alpha && (alpha->chan < 16 || alpha->chan > BN || BN % alpha->chan != 0 || int64_t(alpha->heads) * alpha->chan != m))
    return LGK_OK;
  const bool aligned = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(out) |
                         reinterpret_cast<uintptr_t>(bias)) & 15) == 0;
  if (!aligned || (k & 3) || n >= (int64_t(1) << 31) || k >= (int64_t(1) << 31)) return LGK_OK;
  // the attribute is per device: one flag per device ordinal
  static bool configured_dev[64] = {};
  int dev_ord = 0;
  cudaGetDevice(&dev_ord);
  bool& configured = configured_dev[dev_ord & 63];
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(linear_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
    if (e != cudaSuccess) {
      set_error("linear_tma_kernel: cannot reserve %d bytes of shared memory: %s", kSmemBytes, cudaGetErrorString(e));
      return LGK_ERR_LAUNCH;
    }
    configured = true;
  }
  CUtensorMap map_x, map_w;
  if (!make_map(&map_x, x, n, k, BM) || !make_map(&map_w, w, m, k, BN)) return LGK_OK;
  TcArgs g{bias, rowscale, out, n, k, m, act, accumulate, 0, 0, 0, nullptr, nullptr, nullptr, nullptr, 0, 0};
  if (alpha) {
    g.att_src = alpha->att_src; g.att_dst = alpha->att_dst; g.a_src = alpha->a_src; g.a_dst = alpha->a_dst;
    g.heads = alpha->heads; g.chan = alpha->chan;
  }
  static const int debug = [] { const char* v = getenv("LGK_TC_DEBUG"); return (v && *v) ? atoi(v) : 0; }();
  g.debug = debug;
  g.tiles_n = ceil_div(m, BN);
  g.tiles_total = ceil_div(n, BM) * g.tiles_n;
  const unsigned grid = unsigned(g.tiles_total < kNumSMs ? g.tiles_total : kNumSMs);
  linear_tma_kernel<<<grid, kThreadsTc, kSmemBytes, stream>>>(g, map_x, map_w);
  LGK_CHECK_LAUNCH("linear_tma_kernel");
  *done = true;
  return LGK_OK;
}

}  // namespace lgk
