// (2)+(4) windowed, row-blocked aggregation: the sum / mean SpMM of large banded graphs at HBM speed.
//
// Why a second aggregation kernel (measurements: profiles/README.md r01 v1, r02): spmm_vec_kernel lands every
// gathered neighbour row in registers through the L1 data path -- d x F x 4 bytes per output row against
// 2 x F x 4 compulsory HBM bytes -- and each batch of gathers waits on an L2 round trip, so at d = 16 it sits
// at 0.3-0.4 of the HBM roofline.  Two changes remove both costs:
//
//   (i)  quad-blocked CSR (lgk_quad_build, built once per batch next to the CSR / CSC): four consecutive
//        destination rows share ONE merged, sorted neighbour list with a 4-bit membership mask per entry.
//        Neighbouring rows of a lesion / superpixel graph share most of their neighbours (raster order), so a
//        gathered row feeds up to four accumulators: 2.1x fewer gathers on the config-4 stress graph.
//   (ii) the source rows a chunk of 256 destination rows needs form a contiguous WINDOW of node ids (graphs of
//        a collated batch are block diagonal, and inside a graph the node order follows the image raster).
//        A producer warp streams that window into a shared-memory ring with TMA (tensor-map boxes of 32 rows x
//        128 bytes, mbarrier-completed, several chunks ahead); the consumers gather from shared memory only:
//        no register-landing global gathers, no L2 round trips on the critical path, every source row is
//        read from L2/HBM once per CTA.
//
// A row is cut into 128-byte slices (32 fp32): a CTA owns one slice of a contiguous span of chunks, 8 lanes
// cover a slice (LDS.128), a warp reduces four quads (16 rows) at once; the accumulation is FFMA2 / FADD2
// (packed fp32x2, sm_100).  A chunk whose window does not fit the ring (wide band, random node order) is
// reduced by the same code with global gathers, so the kernel is correct for every graph.
// Atomic-free, fixed summation order (ascending source id): deterministic.
#include "common.cuh"
#include "umma.cuh"
#include <cuda.h>  // CUtensorMap (types only: the encoder is fetched through cudaGetDriverEntryPoint)
#include <limits.h>
#include <stdlib.h>

namespace lgk {
namespace {

constexpr int kQuadsPerChunk = 16;               // 64 destination rows per chunk (granularity of cmin / cmax)
constexpr int kBlkBytes = 4096;                  // bytes of a ring block = one TMA box (32 rows x 128 B or 16 rows x 256 B)
constexpr int kStages = 8;                       // chunks in flight between producer and consumers
constexpr int kMaxConsumerWarps = 24;
constexpr int kBatch = 32;                       // merged entries staged per group and batch (four per lane at 8 lanes)
constexpr int kSmemMax = 232448;                 // 227 KB
constexpr unsigned kColMask = 0x0FFFFFFFu;
constexpr int kStageBytes = (kBatch / 2) * 64;   // staging area of a warp: [pair][4 groups][16 B]
// GAT mode: an entry carries one soft-max weight per row of the quad: [entry][group]{address | mask, w0, w1, w2} + [entry][group]{w3}
constexpr int kStageBytesGat = kBatch * 4 * 20;
constexpr float kLog2e = 1.4426950408889634f;
// A chunk whose source window (largest - smallest source id + 1) exceeds this many rows does not leave the 256-byte-slice
// ring (~830 rows) room to run two chunks ahead; lgk_quad_build counts them (qmeta[1]) and the launch picks the slice
// width from that count on the device.
constexpr int kWide256Rows = 700;

// ---- quad-blocked CSR builder -------------------------------------------------------------------------------
// thread per quad (a warp per chunk): four-way merge of the quad's sorted rows.  Entry = col | mask << 28, written in
// place of the quad's first row (ptr[4g] + t): the merged list is never longer than the four rows together.
// qmeta[0] = widest chunk window in rows, qmeta[1] = chunks wider than kWide256Rows (both accumulated atomically:
// integer max / count, order independent).
__global__ void __launch_bounds__(128) quad_build_kernel(const int* __restrict__ ptr, const int* __restrict__ idx,
                                                         int64_t n_rows, int64_t n_quads, int* __restrict__ qlen,
                                                         unsigned* __restrict__ qent, int* __restrict__ cmin,
                                                         int* __restrict__ cmax, int* __restrict__ qmeta) {
  const int64_t chunk = (int64_t(blockIdx.x) * 128 + threadIdx.x) / kQuadsPerChunk;
  const int64_t g = int64_t(blockIdx.x) * 128 + threadIdx.x;
  int lo = INT_MAX, hi = -1;
  if (g < n_quads) {
    int pos[4], end[4], c[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int64_t r = min(g * 4 + k, n_rows);
      pos[k] = __ldg(ptr + r);
      end[k] = __ldg(ptr + min(r + 1, n_rows));
      c[k] = pos[k] < end[k] ? __ldg(idx + pos[k]) : INT_MAX;
    }
    const int start = pos[0];
    int out = start;
    unsigned selfc = 0;  // stored (i, i) entries of the four rows, 8 bits each (saturating)
    while (true) {
      const int m = min(min(c[0], c[1]), min(c[2], c[3]));
      if (m == INT_MAX) break;
      unsigned mask = 0;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (c[k] == m) {  // a duplicate edge of the same row stays at the head and is emitted again
          mask |= 1u << k;
          if (int64_t(m) == g * 4 + k && ((selfc >> (8 * k)) & 0xffu) < 255u) selfc += 1u << (8 * k);
          ++pos[k];
          c[k] = pos[k] < end[k] ? __ldg(idx + pos[k]) : INT_MAX;
        }
      }
      qent[out++] = unsigned(m) | (mask << 28);
      lo = min(lo, m);
      hi = max(hi, m);
    }
    qlen[2 * g] = out - start;
    qlen[2 * g + 1] = int(selfc);
  }
#pragma unroll
  for (int o = kQuadsPerChunk / 2; o; o >>= 1) {  // reduce over the chunk's lanes (a chunk never straddles a warp)
    lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  if ((threadIdx.x & (kQuadsPerChunk - 1)) == 0 && chunk * kQuadsPerChunk < n_quads) {
    cmin[chunk] = lo;
    cmax[chunk] = hi;
    if (hi >= lo) {
      const int width = hi - lo + 1;
      atomicMax(qmeta, width);
      if (width > kWide256Rows) atomicAdd(qmeta + 1, 1);
    }
  }
}

// ---- PTX helpers ----------------------------------------------------------------------------------------------
__device__ __forceinline__ void tma_box_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
      ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(umma::smem_u32(bar))
      : "memory");
}
// Every wait of this kernel is bounded: a protocol error (or a TMA fault) must surface as a launch failure, not as a
// hung GPU.  ~2^22 polls x >= 200 ns is seconds, orders of magnitude beyond any legitimate wait.
__device__ __forceinline__ void win_wait(uint64_t* bar, uint32_t parity) {
  if (umma::mbar_try_wait(bar, parity)) return;
  for (unsigned n = 0; !umma::mbar_try_wait(bar, parity); ++n) {
    __nanosleep(200);
    if (n > (1u << 22)) __trap();
  }
}
__device__ __forceinline__ float4 lds_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint4 lds_u4(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_u4(uint32_t addr, unsigned a, unsigned b, unsigned c, unsigned d) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void sts_u2(uint32_t addr, unsigned a, unsigned b) {
  asm volatile("st.shared.v2.u32 [%0], {%1,%2};" ::"r"(addr), "r"(a), "r"(b) : "memory");
}
// exp(d) for d <= 0 as one MUFU: ex2.approx(log2e * d); ex2.approx(-inf) = +0, ex2.approx(0) = 1 exactly, 2 ulp
__device__ __forceinline__ float exp_neg(float d) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d * 1.4426950408889634f));
  return r;
}
__device__ __forceinline__ float ex2_fast(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float rcp_fast(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
// packed fp32x2 arithmetic (FFMA2 / FADD2): two lanes of a float4 per instruction, one issue slot
__device__ __forceinline__ void fma2(float2& acc, const float2 v, const float2 w) {
  unsigned long long a = *reinterpret_cast<unsigned long long*>(&acc);
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(a) : "l"(*reinterpret_cast<const unsigned long long*>(&v)),
      "l"(*reinterpret_cast<const unsigned long long*>(&w)));
  acc = *reinterpret_cast<float2*>(&a);
}
__device__ __forceinline__ void add2(float2& acc, const float2 v) {
  unsigned long long a = *reinterpret_cast<unsigned long long*>(&acc);
  asm("add.rn.f32x2 %0, %0, %1;" : "+l"(a) : "l"(*reinterpret_cast<const unsigned long long*>(&v)));
  acc = *reinterpret_cast<float2*>(&a);
}

struct WinArgs {
  const int* ptr;        // [n_rows + 1] CSR / CSC row pointer (quad g's merged list starts at ptr[4g])
  const int* qlen;       // [n_quads][2]: {merged entries, stored self loops of the four rows (8 bits each)}
  const unsigned* qent;  // merged entries
  const int* cmin;       // [n_chunks] smallest / largest source id of a chunk (INT_MAX / -1 when it has no entries)
  const int* cmax;
  const int* qmeta;      // {widest chunk window, chunks wider than kWide256Rows}: device-side choice of the slice width
  const float* pre;
  const float* post;
  const float* selfw;
  const float* bias;
  const float* x;
  float* out;
  float self_fill;
  int self_mode;
  int mean;
  int select;            // 1: run only if qmeta selects this slice width; 0: run unconditionally
  int vec_rows;          // ptr, post and selfw are 16-byte aligned: a quad's four values are one vector load
  int64_t n_rows, n_quads, n_chunks;
  int feat, slices, nb;  // nb: ring blocks
  // GAT mode (edge soft-max + aggregate; pre = a_src TRANSPOSED [heads][pre_stride], one head per 128-byte slice group)
  const float* a_dst;    // transposed like pre: [heads][pre_stride]
  float* rowmax;         // [n_rows, heads] soft-max statistics for the backward (may be NULL)
  float* rowsum;
  float* act_out;        // ELU(out) as a second output (may be NULL)
  int heads, chan;
  int64_t pre_stride;
  float slope;
  int add_self;
};

struct WinShared {
  uint64_t full[kStages], empty[kStages];
  int direct[kStages];   // chunk is reduced with global gathers (its window does not fit the ring)
  int lo_blk[kStages];   // first ring block the chunk reads (INT_MAX: none), for the eviction rule
  int hi_blk[kStages];   // one past the last ring block of the chunk's window
};

// the 256-byte-slice kernel runs when at most 1 chunk in 32 is too wide for its ring, the 128-byte one otherwise
__device__ __forceinline__ bool select_wide_slices(const WinArgs& a) {
  return int64_t(__ldg(a.qmeta + 1)) * 32 <= a.n_chunks;
}

// Accumulate one staged entry into the quad's four accumulators.  `o` carries the membership mask in bits 28..31.
template <bool HAS_PRE>
__device__ __forceinline__ void accumulate(float2 (&acc)[4][2], const float4 v, const unsigned o, const float w) {
  const float2 lo = make_float2(v.x, v.y), hi = make_float2(v.z, v.w);
  const float2 ww = make_float2(w, w);
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    if (o & (1u << (28 + r))) {
      if (HAS_PRE) {
        fma2(acc[r][0], lo, ww);
        fma2(acc[r][1], hi, ww);
      } else {
        add2(acc[r][0], lo);
        add2(acc[r][1], hi);
      }
    }
  }
}

// LPG: lanes per group = 16-byte pieces of a slice (8: 128-byte slices, 4 quads per warp, 2 sets of 8 warps working on
// alternate chunks; 16: 256-byte slices, 2 quads per warp, 1 set).
// NW: consumer warps (+ 1 producer warp); 24 need the kernel to fit 80 registers.
// SELF (LGK_SELF_*) and SCALE (mean or post-scale present) are compile-time: per quad the kernel spends as many
// instructions on bookkeeping as on a third of its gathers (ncu, profiles/r02), so nothing optional stays in that path.
// GAT: edge soft-max + aggregation (lgk_gat_fwd_blocked): the staged factor of a source row is its attention logit half
// a_src[j, head], every (entry, row) pair gets the weight exp(e - running max) -- online soft-max at batch granularity --
// and the epilogue normalises, adds the GATConv self loop and writes the soft-max statistics.
template <bool HAS_PRE, int LPG, int NW, int SELF, bool SCALE, bool GAT = false>
__global__ void __launch_bounds__(32 * (NW + 1), 1) spmm_win_kernel(const WinArgs a, const __grid_constant__ CUtensorMap map_x) {
  static_assert(!GAT || (HAS_PRE && LPG == 8 && SELF == LGK_SELF_NONE && !SCALE), "GAT mode: 128-byte slices, logit halves staged");
  constexpr int DBG = 0;
  constexpr int kConsumerWarps = NW;
  constexpr int kSliceBytes = LPG * 16;
  constexpr int kBlkRows = kBlkBytes / kSliceBytes;
  constexpr int GPW = 32 / LPG;                          // groups (quads in flight) per warp
  // work unit of the pipeline = `kSuper` consecutive chunks of the quad structure: one for the 256-byte slices (their
  // ring only holds a 64-row chunk's window plus look-ahead), four for the 128-byte slices (fewer, larger hand-offs)
  constexpr int kSuper = LPG == 16 ? 1 : NW / 4;  // 4 chunks (256 rows) at 16 warps, 6 (384 rows) at 24
  constexpr int kUnitQuads = kQuadsPerChunk * kSuper;
  constexpr int kSets = kConsumerWarps * GPW / kUnitQuads >= 1 ? kConsumerWarps * GPW / kUnitQuads : 1;
  constexpr int kWarpsPerSet = kConsumerWarps / kSets;
  constexpr int EPL = kBatch / LPG;                      // entries a lane stages per batch
  static_assert(kWarpsPerSet * GPW == kUnitQuads && kSets * kWarpsPerSet == kConsumerWarps, "one quad per group and unit");
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  if (a.select && select_wide_slices(a) != (LPG == 16)) return;
  // layout: ring [nb][4096 B] | ring_pre [nb][kBlkRows] floats | staging [warps][8 pairs][GPW][16 B] | WinShared
  const uint32_t ring = umma::smem_u32(smem_raw);
  float* ring_pre = reinterpret_cast<float*>(smem_raw + size_t(a.nb) * kBlkBytes);
  unsigned char* stage_raw = reinterpret_cast<unsigned char*>(ring_pre + (HAS_PRE ? a.nb * 32 : 0));
  constexpr int kStageW = GAT ? kStageBytesGat : kStageBytes;
  WinShared* sh = reinterpret_cast<WinShared*>(stage_raw + kConsumerWarps * kStageW);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int slice = blockIdx.x % a.slices;
  // GAT: a slice lies inside one head (chan % 32 == 0) and the staged factors are that head's row of the transposed logits
  const int head = GAT ? (slice * (kSliceBytes / 4)) / a.chan : 0;
  const float* pre_g = GAT ? a.pre + int64_t(head) * a.pre_stride : a.pre;
  const float* adst_g = GAT ? a.a_dst + int64_t(head) * a.pre_stride : nullptr;  // GAT: a_dst transposed like a_src
  const int span = blockIdx.x / a.slices, spans = gridDim.x / a.slices;
  const int64_t n_units = (a.n_chunks + kSuper - 1) / kSuper;
  const int64_t c_begin = n_units * span / spans, c_end = n_units * (span + 1) / spans;  // in units
  const int n_my = int(c_end - c_begin);
  const int nb = a.nb;

  if (threadIdx.x == 0) {
    if (ring & (kSliceBytes - 1)) __trap();  // the gather address is formed as ring-row address | lane offset
    for (int s = 0; s < kStages; ++s) {
      umma::mbar_init(&sh->full[s], 1);
      umma::mbar_init(&sh->empty[s], kWarpsPerSet);
    }
    umma::mbar_init_fence();
  }
  __syncthreads();

  if (warp == kConsumerWarps) {
    // ===================== producer: keep the ring ahead of the consumers =====================
    // full[s] of chunk i completes when its own boxes have landed AND the producer has arrived on it -- which it does
    // one iteration later, after it has seen full of chunk i - 1 complete.  So "full[i] complete" implies that the
    // boxes of every earlier chunk have landed too: a consumer set that skipped those chunks (they belong to the other
    // sets) still finds the older part of its window in the ring.
    int base_blk = 0, top_blk = 0;  // resident source blocks [base_blk, top_blk); block b lives in slot b % nb
    int oldest = 0;                 // oldest chunk (local index) whose completion has not been waited for
    // chunk windows: lane l holds chunk 32 k + l of the current and of the next group of 32 (loaded a group ahead, so the
    // loop never waits on them)
    auto load_range = [&](int first, int& lo_r, int& hi_r) {
      lo_r = INT_MAX;
      hi_r = -1;
      if (first + lane < n_my) {
#pragma unroll
        for (int u = 0; u < kSuper; ++u) {
          const int64_t c = (c_begin + first + lane) * kSuper + u;
          if (c < a.n_chunks) {
            lo_r = min(lo_r, __ldg(a.cmin + c));
            hi_r = max(hi_r, __ldg(a.cmax + c));
          }
        }
      }
    };
    int cur_lo, cur_hi, nxt_lo, nxt_hi;
    load_range(0, cur_lo, cur_hi);
    load_range(32, nxt_lo, nxt_hi);
    for (int i = 0; i < n_my; ++i) {
      const int s = i % kStages;
      if ((i & 31) == 0 && i > 0) {
        cur_lo = nxt_lo;
        cur_hi = nxt_hi;
        load_range(i + 32, nxt_lo, nxt_hi);
      }
      const int lo = __shfl_sync(0xffffffffu, cur_lo, i & 31), hi = __shfl_sync(0xffffffffu, cur_hi, i & 31);
      // deferred arrival on the previous chunk's barrier (see above) -- FIRST, so that every chunk this iteration may
      // have to drain below can actually start
      if (lane == 0 && i >= 1) {
        if (i >= 2) win_wait(&sh->full[(i - 2) % kStages], ((i - 2) / kStages) & 1);
        umma::mbar_arrive(&sh->full[(i - 1) % kStages]);
      }
      __syncwarp();
      // stage s still belongs to chunk i - kStages until every consumer warp of its set has released it
      while (oldest + kStages <= i) {
        win_wait(&sh->empty[oldest % kStages], (oldest / kStages) & 1);
        ++oldest;
      }
      int direct = 0, lo_b = INT_MAX, load_from = 0, load_to = 0;
      if (hi >= lo) {
        lo_b = lo / kBlkRows;
        const int hi_b = hi / kBlkRows + 1;
        if (hi_b - lo_b > nb) {
          direct = 1;
          lo_b = INT_MAX;
        } else {
          if (lo_b < base_blk) {  // backward jump: drain everything in flight, restart the window
            while (oldest < i) {
              win_wait(&sh->empty[oldest % kStages], (oldest / kStages) & 1);
              ++oldest;
            }
            base_blk = top_blk = lo_b;
          } else if (lo_b > top_blk) {  // gap (next graph starts further on): nothing below is needed again
            base_blk = top_blk = lo_b;
          }
          const int new_top = max(top_blk, hi_b);
          // a block may only replace another one of its slot when no chunk in flight still reads from that slot:
          // every in-flight window [lo'', lo'' + nb) must contain the new blocks
          int need = oldest;
          for (int j = oldest; j < i; ++j)
            if (sh->lo_blk[j % kStages] != INT_MAX && sh->lo_blk[j % kStages] + nb < new_top) need = j + 1;
          while (oldest < need) {
            win_wait(&sh->empty[oldest % kStages], (oldest / kStages) & 1);
            ++oldest;
          }
          load_from = top_blk;
          load_to = new_top;
          top_blk = new_top;
          base_blk = max(base_blk, top_blk - nb);
        }
      }
      // the pre-scale factors of the new rows travel like the rows themselves: 1-D bulk copies (at most two: the ring
      // wraps) completing on the same barrier; only a ragged tail of < 4 rows at the very end of the array is stored by hand
      uint32_t pre_bytes[2] = {0, 0};
      int64_t pre_row[2] = {0, 0};
      if (HAS_PRE && load_to > load_from) {
        const int64_t R = int64_t(nb) * kBlkRows;
        const int64_t r_first = int64_t(load_from) * kBlkRows;
        const int64_t v_last = min(int64_t(load_to) * kBlkRows, a.n_rows);
        const int64_t p0 = r_first % R;
        const int64_t n1 = max(int64_t(0), min(v_last - r_first, R - p0)), n2 = max(int64_t(0), v_last - r_first - n1);
        pre_row[0] = r_first;
        pre_row[1] = r_first + n1;
        pre_bytes[0] = uint32_t(n1 * 4) & ~15u;
        pre_bytes[1] = uint32_t(n2 * 4) & ~15u;
        const int64_t tail0 = r_first + pre_bytes[0] / 4, tail1 = pre_row[1] + pre_bytes[1] / 4;
        if (tail0 + lane < pre_row[1]) ring_pre[(tail0 + lane) % R] = __ldg(pre_g + tail0 + lane);
        if (tail1 + lane < v_last) ring_pre[(tail1 + lane) % R] = __ldg(pre_g + tail1 + lane);
      }
      if (lane == 0) {
        sh->direct[s] = direct;
        sh->lo_blk[s] = lo_b;
        sh->hi_blk[s] = lo_b == INT_MAX ? 0 : hi / kBlkRows + 1;
      }
      __syncwarp();
      if (lane == 0) {
        const int nload = (DBG & 4) ? 0 : load_to - load_from;
        const uint32_t tx = uint32_t(nload) * kBlkBytes + pre_bytes[0] + pre_bytes[1];
        if (tx > 0) umma::mbar_expect_tx(&sh->full[s], tx);
        if (HAS_PRE) {
          const int64_t R = int64_t(nb) * kBlkRows;
#pragma unroll
          for (int k = 0; k < 2; ++k)
            if (pre_bytes[k]) umma::bulk_g2s(ring_pre + pre_row[k] % R, pre_g + pre_row[k], pre_bytes[k], &sh->full[s]);
        }
        for (int b = load_from; b < load_from + nload; ++b)
          tma_box_2d(ring + uint32_t(b % nb) * kBlkBytes, &map_x, slice * (kSliceBytes / 4), b * kBlkRows, &sh->full[s]);
      }
      __syncwarp();
    }
    if (lane == 0 && n_my >= 1) {
      if (n_my >= 2) win_wait(&sh->full[(n_my - 2) % kStages], ((n_my - 2) / kStages) & 1);
      umma::mbar_arrive(&sh->full[(n_my - 1) % kStages]);
    }
    return;
  }

  // ===================== consumers: GPW groups of LPG lanes per warp, one quad per group and unit =====================
  // all index arithmetic is 32-bit: the launcher checks n_rows * feat / 4 < 2^31
  const int gl = lane & (LPG - 1), grp = lane / LPG;
  const int set = warp / kWarpsPerSet;
  const int group = (warp % kWarpsPerSet) * GPW + grp;  // quad of the unit this group reduces
  const unsigned F4 = unsigned(a.feat) >> 2;
  const unsigned vec = unsigned(slice) * LPG + gl;  // this lane's float4 column
  const bool lane_on = vec < F4;
  const uint32_t stage_base = umma::smem_u32(stage_raw) + warp * kStageW + grp * 16;  // + pair * (GPW * 16)
  const uint32_t stage_w3 = umma::smem_u32(stage_raw) + warp * kStageW + kBatch * GPW * 16 + grp * 4;  // GAT: + entry * (GPW * 4)
  const float4* xv = reinterpret_cast<const float4*>(a.x) + (lane_on ? vec : 0u);
  float4* ov = reinterpret_cast<float4*>(a.out) + vec;
  const uint32_t lane_off = uint32_t(gl) * 16;
  const unsigned nb_magic = unsigned((uint64_t(1) << 32) / unsigned(nb)) + 1u;  // blk % nb without a divide (blk < 2^24)
  const int* qent_i = reinterpret_cast<const int*>(a.qent);
  const int n_rows = int(a.n_rows), n_quads = int(a.n_quads);
  const int unit0 = int(c_begin);
  const bool has_mean = a.mean != 0, has_post = a.post != nullptr, has_selfw = a.selfw != nullptr;
  float4 bias4 = f4_zero();
  if (a.bias && lane_on) bias4 = __ldg(reinterpret_cast<const float4*>(a.bias) + vec);

  // software pipeline over the quads of this group, two deep: while quad q is reduced, the first two batches of merged
  // entries of quad q + 1 and the header (start, len) of quad q + 2 are in flight -- the entry stream comes from HBM
  // (~1 us), one quad of work per warp is less than that
  auto load_header = [&](int i, int& st, int& ln, unsigned& sc) {
    const int g = (unit0 + i) * kUnitQuads + group;
    st = ln = 0;
    sc = 0;
    if (i < n_my && g < n_quads) {
      st = __ldg(a.ptr + 4 * g);
      const int2 h = __ldg(reinterpret_cast<const int2*>(a.qlen) + g);
      ln = h.x;
      sc = unsigned(h.y);
    }
  };
  auto load_batch = [&](int st, int ln, int b0, unsigned (&wd)[EPL]) {
#pragma unroll
    for (int h = 0; h < EPL; ++h) {
      const int e = b0 + gl + LPG * h;
      wd[h] = 0;
      if (e < ln) wd[h] = unsigned(ld_stream_i32(qent_i + st + e));
    }
  };
  int start_c, len_c, start_1, len_1;
  unsigned self_c, self_1;
  unsigned wc0[EPL], wc1[EPL];
  load_header(set, start_c, len_c, self_c);
  load_header(set + kSets, start_1, len_1, self_1);
  load_batch(start_c, len_c, 0, wc0);
  load_batch(start_c, len_c, kBatch, wc1);

  for (int i = set; i < n_my; i += kSets) {
    const int s = i % kStages;
    const int g = (unit0 + i) * kUnitQuads + group;
    const int row0 = 4 * g;
    const int start = start_c, len = len_c;
    const unsigned selfcnt = self_c;  // stored (i, i) entries per row, 8 bits each
    unsigned w[EPL], w2[EPL];
#pragma unroll
    for (int h = 0; h < EPL; ++h) { w[h] = wc0[h]; w2[h] = wc1[h]; }
    // next quad's entries and the header after it go out before anything of this quad is touched
    load_batch(start_1, len_1, 0, wc0);
    load_batch(start_1, len_1, kBatch, wc1);
    start_c = start_1;
    len_c = len_1;
    self_c = self_1;
    load_header(i + 2 * kSets, start_1, len_1, self_1);
    // what the epilogue reads goes out now and is only touched there (rows past the end of the array read row n_rows - 1
    // and are not stored)
    float postv[4], cself[4];
    int deg[4];
    if (SCALE || SELF != LGK_SELF_NONE) {
      const bool whole = a.vec_rows && row0 + 3 < n_rows;  // the quad's four rows as one 16-byte load per array
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        postv[r] = 1.f;
        deg[r] = 1;
        cself[r] = a.self_fill;
      }
      if (whole) {
        if (SCALE && has_post) {
          const float4 p4 = __ldg(reinterpret_cast<const float4*>(a.post + row0));
          postv[0] = p4.x; postv[1] = p4.y; postv[2] = p4.z; postv[3] = p4.w;
        }
        if (SELF != LGK_SELF_NONE && has_selfw) {
          const float4 s4 = __ldg(reinterpret_cast<const float4*>(a.selfw + row0));
          cself[0] = s4.x; cself[1] = s4.y; cself[2] = s4.z; cself[3] = s4.w;
        }
        if (SCALE && has_mean) {
          const int4 q4 = __ldg(reinterpret_cast<const int4*>(a.ptr + row0));
          const int q5 = __ldg(a.ptr + row0 + 4);
          deg[0] = q4.y - q4.x; deg[1] = q4.z - q4.y; deg[2] = q4.w - q4.z; deg[3] = q5 - q4.w;
        }
      } else {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          const int row = min(row0 + r, n_rows - 1);
          if (SCALE && has_mean) deg[r] = __ldg(a.ptr + row + 1) - __ldg(a.ptr + row);
          if (SCALE && has_post) postv[r] = __ldg(a.post + row);
          if (SELF != LGK_SELF_NONE && has_selfw) cself[r] = __ldg(a.selfw + row);
        }
      }
    }
    // GAT: destination-side logit halves of the quad's rows, and the source-side ones of the rows themselves (self loop)
    float adst[4], aself[4], mrow[4], lsum[4];
    if (GAT) {
      // both logit halves are stored transposed [head][row] (row stride a multiple of 4): one 16-byte load each per quad
      const float4 d4 = (i < n_my && g < n_quads) ? __ldg(reinterpret_cast<const float4*>(adst_g + row0)) : f4_zero();
      const float4 s4 = (a.add_self && i < n_my && g < n_quads) ? __ldg(reinterpret_cast<const float4*>(pre_g + row0)) : f4_zero();
      adst[0] = d4.x; adst[1] = d4.y; adst[2] = d4.z; adst[3] = d4.w;
      aself[0] = s4.x; aself[1] = s4.y; aself[2] = s4.z; aself[3] = s4.w;
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        mrow[r] = -INFINITY;
        lsum[r] = 0.f;
      }
    }
    win_wait(&sh->full[s], (i / kStages) & 1);
    const int direct = sh->direct[s];
    const int win_lo = sh->lo_blk[s] == INT_MAX ? INT_MAX : sh->lo_blk[s] * kBlkRows, win_hi = sh->hi_blk[s] * kBlkRows;
    const int maxlen = __reduce_max_sync(0xffffffffu, len);
    float2 acc[4][2];
#pragma unroll
    for (int r = 0; r < 4; ++r) acc[r][0] = acc[r][1] = make_float2(0.f, 0.f);

    for (int b0 = 0; b0 < maxlen; b0 += kBatch) {
      unsigned word[EPL];
#pragma unroll
      for (int h = 0; h < EPL; ++h) { word[h] = w[h]; w[h] = w2[h]; }
      load_batch(start, b0 + 2 * kBatch < maxlen ? len : 0, b0 + 2 * kBatch, w2);  // two batches ahead
      if constexpr (GAT) {
        // Groups of a warp work independently here (their merged lists differ in length: a warp-wide batch count made every
        // group pay the staging arithmetic of the longest one): group-masked shuffles and barriers, per-group entry count.
        const unsigned gmask = (LPG == 32 ? 0xffffffffu : ((1u << LPG) - 1u)) << (grp * LPG);
        const int cnt = min(kBatch, len - b0);
        if (cnt > 0) {
          // ---- GAT stage: logits of every (entry, member row), batch maximum per row, rescale, weights.  Logits and maxima
          // are formed exactly as lgk_gat_fwd forms them (the stored row maximum is bit-identical); exp(d) = exp2(log2e * d).
          float ev[EPL][4];
          unsigned oo[EPL];
          float bm[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
          for (int h = 0; h < EPL; ++h) {
            const unsigned col = word[h] & kColMask;
            unsigned mask = word[h] >> 28;
            if (a.add_self) {  // GATConv: stored self loops are removed (one is added in the epilogue)
              const unsigned d = col - unsigned(row0);
              if (d < 4u) mask &= ~(1u << d);
            }
            unsigned o;
            float pj;
            if (!direct) {
              const unsigned blk = col / kBlkRows, slot = blk - __umulhi(blk, nb_magic) * unsigned(nb);
              const unsigned rib = slot * kBlkRows + (col & (kBlkRows - 1));
              o = ring + rib * kSliceBytes;
              pj = ring_pre[rib];
            } else {
              o = col;
              pj = __ldg(pre_g + col);
            }
            oo[h] = o | (mask << 28);
#pragma unroll
            for (int r = 0; r < 4; ++r) {
              const float z = pj + adst[r];
              ev[h][r] = (mask & (1u << r)) ? fmaxf(z, a.slope * z) : -INFINITY;  // leaky-relu, slope in [0, 1]
              bm[r] = fmaxf(bm[r], ev[h][r]);
            }
          }
          float msafe[4];
#pragma unroll
          for (int r = 0; r < 4; ++r) {
#pragma unroll
            for (int o = 1; o < LPG; o <<= 1) bm[r] = fmaxf(bm[r], __shfl_xor_sync(gmask, bm[r], o));
            // branch-free rescale: factor 1 when the maximum stands (exp_neg(0) = 1), 0 on the first members of the row
            // (mrow = -inf: exp_neg(-inf) = 0); a row without any member so far (both -inf) keeps its zeros
            const float m_new = fmaxf(mrow[r], bm[r]);
            msafe[r] = m_new == -INFINITY ? 0.f : m_new;
            const float sc = exp_neg(mrow[r] - msafe[r]);
            acc[r][0].x *= sc; acc[r][0].y *= sc; acc[r][1].x *= sc; acc[r][1].y *= sc;
            lsum[r] *= sc;
            mrow[r] = m_new;
          }
#pragma unroll
          for (int r = 0; r < 4; ++r) msafe[r] *= -kLog2e;
#pragma unroll
          for (int h = 0; h < EPL; ++h) {
            float wv[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
              // exp(e - m) = ex2(e log2e - m log2e), one FFMA + one MUFU; non-members hold -inf: weight +0.  The rounding of
              // m log2e is a factor common to the row's weights: it cancels in out = acc / sum and moves the stored sum by
              // |m| 6e-8 relative
              wv[r] = ex2_fast(fmaf(ev[h][r], kLog2e, msafe[r]));
              lsum[r] += wv[r];
            }
            const int ent = gl + LPG * h;
            sts_u4(stage_base + ent * (GPW * 16), oo[h], __float_as_uint(wv[0]), __float_as_uint(wv[1]), __float_as_uint(wv[2]));
            umma::sts_f32(stage_w3 + ent * (GPW * 4), wv[3]);
          }
          __syncwarp(gmask);
          auto gat_entry = [&](const uint4 st, const float w3, const float4 v) {
            const float2 lo = make_float2(v.x, v.y), hi = make_float2(v.z, v.w);
            const float wr[4] = {__uint_as_float(st.y), __uint_as_float(st.z), __uint_as_float(st.w), w3};
#pragma unroll
            for (int r = 0; r < 4; ++r) {
              if (st.x & (1u << (28 + r))) {
                const float2 ww = make_float2(wr[r], wr[r]);
                fma2(acc[r][0], lo, ww);
                fma2(acc[r][1], hi, ww);
              }
            }
          };
          if (!direct) {
#pragma unroll 4
            for (int t = 0; t < cnt; ++t) {
              const uint4 st = lds_u4(stage_base + t * (GPW * 16));
              gat_entry(st, umma::lds_f32(stage_w3 + t * (GPW * 4)), lds_f4((st.x & kColMask) | lane_off));
            }
          } else {
#pragma unroll 2
            for (int t = 0; t < cnt; ++t) {
              const uint4 st = lds_u4(stage_base + t * (GPW * 16));
              gat_entry(st, umma::lds_f32(stage_w3 + t * (GPW * 4)), __ldg(xv + (st.x & kColMask) * F4));
            }
          }
          __syncwarp(gmask);
        }
        continue;
      }
      // ---- stage: every lane resolves its entries (ring address | mask, weight) ----
#pragma unroll
      for (int h = 0; h < EPL; ++h) {
        const unsigned col = word[h] & kColMask, mask = word[h] >> 28;
        unsigned o;
        float wt = 1.f;
        if (!direct) {
          const unsigned blk = col / kBlkRows, slot = blk - __umulhi(blk, nb_magic) * unsigned(nb);
          const unsigned rib = slot * kBlkRows + (col & (kBlkRows - 1));
          o = ring + rib * kSliceBytes;
          if (HAS_PRE) wt = ring_pre[rib];
        } else {
          o = col;
          if (HAS_PRE) wt = __ldg(pre_g + col);
        }
        // pair p = entry / 2, half = entry & 1: [pair][group][half] -> one 16-byte read serves two entries
        const int ent = gl + LPG * h;
        // with weights: [pair][group][{word, weight} x 2]; without: [four entries][group][word x 4] -- one 16-byte read
        // serves two / four entries
        if (HAS_PRE) sts_u2(stage_base + (ent >> 1) * (GPW * 16) + (ent & 1) * 8, o | (mask << 28), __float_as_uint(wt));
        else umma::sts_f32(stage_base + (ent >> 2) * (GPW * 16) + (ent & 3) * 4, __uint_as_float(o | (mask << 28)));
      }
      __syncwarp();
      const int cnt = min(kBatch, maxlen - b0);
      if (!direct && !HAS_PRE) {
#pragma unroll 2
        for (int t = 0; t < cnt; t += 4) {  // entries past cnt are staged with an empty mask
          const uint4 st = lds_u4(stage_base + (t >> 2) * (GPW * 16));
          const float4 v0 = lds_f4((st.x & kColMask) | lane_off);
          const float4 v1 = lds_f4((st.y & kColMask) | lane_off);
          const float4 v2 = lds_f4((st.z & kColMask) | lane_off);
          const float4 v3 = lds_f4((st.w & kColMask) | lane_off);
          accumulate<false>(acc, v0, st.x, 1.f);
          accumulate<false>(acc, v1, st.y, 1.f);
          accumulate<false>(acc, v2, st.z, 1.f);
          accumulate<false>(acc, v3, st.w, 1.f);
        }
      } else if (!direct) {
#pragma unroll 4
        for (int t = 0; t < cnt; t += 2) {
          const uint4 st = lds_u4(stage_base + (t >> 1) * (GPW * 16));
          const float4 v0 = lds_f4((st.x & kColMask) | lane_off);
          const float4 v1 = lds_f4((st.z & kColMask) | lane_off);
          accumulate<HAS_PRE>(acc, v0, st.x, __uint_as_float(st.y));
          accumulate<HAS_PRE>(acc, v1, st.z, __uint_as_float(st.w));
        }
      } else {
#pragma unroll 2
        for (int t = 0; t < cnt; t += 2) {
          unsigned o0, o1;
          float w0 = 1.f, w1 = 1.f;
          if (HAS_PRE) {
            const uint4 st = lds_u4(stage_base + (t >> 1) * (GPW * 16));
            o0 = st.x; o1 = st.z;
            w0 = __uint_as_float(st.y); w1 = __uint_as_float(st.w);
          } else {
            o0 = __float_as_uint(umma::lds_f32(stage_base + (t >> 2) * (GPW * 16) + (t & 3) * 4));
            o1 = __float_as_uint(umma::lds_f32(stage_base + (t >> 2) * (GPW * 16) + (t & 3) * 4 + 4));
          }
          const float4 v0 = __ldg(xv + (o0 & kColMask) * F4);
          const float4 v1 = __ldg(xv + (o1 & kColMask) * F4);
          accumulate<HAS_PRE>(acc, v0, o0, w0);
          accumulate<HAS_PRE>(acc, v1, o1, w1);
        }
      }
      __syncwarp();
    }

    // ---- epilogue: self term, mean / post scale, bias, store ----
    const int rows_here = n_rows - row0;  // >= 4 everywhere but in the last quad of the array (<= 0: no quad)
    if constexpr (GAT) {
      const bool stats_lane = gl == 0 && (unsigned(slice) * (kSliceBytes / 4)) % unsigned(a.chan) == 0 && a.rowmax != nullptr;
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        float ssum = lsum[r];
#pragma unroll
        for (int o = 1; o < LPG; o <<= 1) ssum += __shfl_xor_sync(0xffffffffu, ssum, o);
        float4 res = make_float4(acc[r][0].x, acc[r][0].y, acc[r][1].x, acc[r][1].y);
        float m = mrow[r];
        const unsigned row = unsigned(row0 + r);
        const unsigned ro = row * F4;
        if (a.add_self) {  // exactly one (i, i) entry
          const float z = aself[r] + adst[r];
          const float es = fmaxf(z, a.slope * z);
          const float m_new = fmaxf(m, es);  // finite
          const float sc = exp_neg(m - m_new);  // m = -inf (no other entry): 0
          const float ws = exp_neg(es - m_new);
          float4 xs = f4_zero();
          if (int(row) >= win_lo && int(row) < win_hi) {
            const unsigned blk = row / kBlkRows, slot = blk - __umulhi(blk, nb_magic) * unsigned(nb);
            xs = lds_f4((ring + (slot * kBlkRows + (row & (kBlkRows - 1))) * kSliceBytes) | lane_off);
          } else if (r < rows_here) {
            xs = __ldg(xv + ro);
          }
          res.x = fmaf(res.x, sc, ws * xs.x); res.y = fmaf(res.y, sc, ws * xs.y);
          res.z = fmaf(res.z, sc, ws * xs.z); res.w = fmaf(res.w, sc, ws * xs.w);
          ssum = fmaf(ssum, sc, ws);
          m = m_new;
        }
        const float inv = rcp_fast(ssum + 1e-16f);
        res.x = fmaf(res.x, inv, bias4.x); res.y = fmaf(res.y, inv, bias4.y);
        res.z = fmaf(res.z, inv, bias4.z); res.w = fmaf(res.w, inv, bias4.w);
        if (lane_on && r < rows_here) {
          st_stream_f4(ov + ro, res);
          if (a.act_out) {
            res.x = elu1(res.x); res.y = elu1(res.y); res.z = elu1(res.z); res.w = elu1(res.w);
            st_stream_f4(reinterpret_cast<float4*>(a.act_out) + vec + ro, res);
          }
          // the first slice of a head records the soft-max statistics (all slices of a head compute the same ones)
          if (stats_lane) {
            a.rowmax[row * unsigned(a.heads) + unsigned(head)] = m;
            a.rowsum[row * unsigned(a.heads) + unsigned(head)] = ssum;
          }
        }
      }
      __syncwarp();
      if (lane == 0) umma::mbar_arrive(&sh->empty[s]);
      continue;
    }
    // self term: with SELF_REPLACE on a graph that stores its loops (kNN with loop=True) every coefficient is 0 -- one
    // warp-uniform vote skips the whole block instead of four predicated copies of it
    bool self_work = false;
    if (SELF != LGK_SELF_NONE) {
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        if (SELF == LGK_SELF_REPLACE) cself[r] -= float((selfcnt >> (8 * r)) & 0xffu);
        self_work |= cself[r] != 0.f;
      }
      self_work = __any_sync(0xffffffffu, self_work);
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      float4 res = make_float4(acc[r][0].x, acc[r][0].y, acc[r][1].x, acc[r][1].y);
      const unsigned ro = unsigned(row0 + r) * F4;
      if (SELF != LGK_SELF_NONE && self_work) {
        float c = cself[r];
        if (c != 0.f && r < rows_here) {
          const int row = row0 + r;
          if (HAS_PRE) c *= __ldg(pre_g + row);
          float4 xs;
          if (row >= win_lo && row < win_hi) {  // the row itself is part of the unit's window: it is in the ring
            const unsigned blk = unsigned(row) / kBlkRows, slot = blk - __umulhi(blk, nb_magic) * unsigned(nb);
            xs = lds_f4((ring + (slot * kBlkRows + (unsigned(row) & (kBlkRows - 1))) * kSliceBytes) | lane_off);
          } else {
            xs = __ldg(xv + ro);
          }
          f4_fma(res, c, xs);
        }
      }
      if (SCALE) {
        const float sc = has_mean ? postv[r] * __frcp_rn(float(max(deg[r], 1))) : postv[r];
        res.x = fmaf(res.x, sc, bias4.x);
        res.y = fmaf(res.y, sc, bias4.y);
        res.z = fmaf(res.z, sc, bias4.z);
        res.w = fmaf(res.w, sc, bias4.w);
      } else {
        res.x += bias4.x; res.y += bias4.y; res.z += bias4.z; res.w += bias4.w;
      }
      if (lane_on && r < rows_here) st_stream_f4(ov + ro, res);
    }
    __syncwarp();
    if (lane == 0) umma::mbar_arrive(&sh->empty[s]);
  }
}

using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_win() {
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

// x [rows, feat] fp32 row-major as a 2-D tensor; boxes of {slice floats, 4096 / slice bytes rows}, no swizzle, zero fill
bool make_row_map(CUtensorMap* map, const float* base, int64_t rows, int64_t feat, int slice_bytes) {
  EncodeTiledFn enc = encode_tiled_win();
  if (!enc) return false;
  const cuuint64_t dims[2] = {cuuint64_t(feat), cuuint64_t(rows)};
  const cuuint64_t strides[1] = {cuuint64_t(4 * feat)};
  const cuuint32_t box[2] = {cuuint32_t(slice_bytes / 4), cuuint32_t(kBlkBytes / slice_bytes)};
  const cuuint32_t estr[2] = {1, 1};
  return enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

int env_int_win(const char* name, int dflt) {
  const char* v = getenv(name);
  return (v && *v) ? atoi(v) : dflt;
}

using WinKernel = void (*)(const WinArgs, const CUtensorMap);

template <int LPG, int NW>
WinKernel pick_kernel(bool has_pre, int self_mode, bool scale) {
#define LGK_WIN_ROW(P, S)                                                                    \
  if (has_pre == P && self_mode == S) return scale ? spmm_win_kernel<P, LPG, NW, S, true> : spmm_win_kernel<P, LPG, NW, S, false>;
  LGK_WIN_ROW(false, LGK_SELF_NONE) LGK_WIN_ROW(false, LGK_SELF_REPLACE) LGK_WIN_ROW(false, LGK_SELF_ADD)
  LGK_WIN_ROW(true, LGK_SELF_NONE) LGK_WIN_ROW(true, LGK_SELF_REPLACE) LGK_WIN_ROW(true, LGK_SELF_ADD)
#undef LGK_WIN_ROW
  return nullptr;
}

// one launch of the slice width LPG * 16 bytes
template <int LPG>
int launch_win(WinArgs a, int select, int dbg, cudaStream_t stream) {
  // consumer warps (tuning knob)
  static const int warps = env_int_win("LGK_WIN_WARPS", 16) == 24 ? 24 : 16;
  constexpr int slice_bytes = LPG * 16;
  a.select = select;
  a.slices = int(ceil_div(int64_t(a.feat) * 4, slice_bytes));
  const int fixed = warps * kStageBytes + int(sizeof(WinShared)) + 128;
  const int per_blk = kBlkBytes + (a.pre ? 32 * 4 : 0);
  static const int nb_cap = env_int_win("LGK_WIN_BLOCKS", 1 << 30);  // tuning / test knob: fewer ring blocks
  a.nb = std::min((kSmemMax - fixed) / per_blk, nb_cap);
  LGK_REQUIRE(a.nb >= 2, LGK_ERR_UNSUPPORTED, "lgk_spmm_blocked: no room for the ring");
  const int smem = a.nb * per_blk + fixed;
  CUtensorMap map_x;
  LGK_REQUIRE(make_row_map(&map_x, a.x, a.n_rows, a.feat, slice_bytes), LGK_ERR_UNSUPPORTED,
              "lgk_spmm_blocked: cannot encode the tensor map");
  // one CTA per SM; a group of `slices` CTAs shares a span of chunks (more slices than SMs: the CTAs queue)
  const int spans = int(std::max<int64_t>(1, std::min<int64_t>(kNumSMs / std::max(1, std::min(a.slices, kNumSMs)), a.n_chunks)));
  const unsigned grid = unsigned(spans * a.slices);
  const bool scale = a.mean || a.post != nullptr;
  WinKernel kern = warps == 24 ? pick_kernel<LPG, 24>(a.pre != nullptr, a.self_mode, scale) : pick_kernel<LPG, 16>(a.pre != nullptr, a.self_mode, scale);
  LGK_REQUIRE(kern, LGK_ERR_BAD_ARG, "lgk_spmm_blocked: no kernel for this mode");
  (void)dbg;
  int dev = 0;
  cudaGetDevice(&dev);
  static bool configured[64][2][3][2] = {};
  bool& done = configured[dev & 63][a.pre ? 1 : 0][a.self_mode][scale ? 1 : 0];
  if (!done) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemMax);
    if (e != cudaSuccess) {
      set_error("spmm_win_kernel: cannot reserve %d bytes of shared memory: %s", kSmemMax, cudaGetErrorString(e));
      return LGK_ERR_LAUNCH;
    }
    done = true;
  }
  kern<<<grid, 32 * (warps + 1), smem, stream>>>(a, map_x);
  LGK_CHECK_LAUNCH("spmm_win_kernel");
  return LGK_OK;
}

// a_src [n, heads] -> [heads][stride] (stride = n rounded up to 4: every head's row starts 16-byte aligned for the bulk copies)
__global__ void __launch_bounds__(256) transpose_heads_kernel(const float* __restrict__ in, const float* __restrict__ in2,
                                                              float* __restrict__ out, float* __restrict__ out2, int64_t n, int heads,
                                                              int64_t stride) {
  const int64_t i = int64_t(blockIdx.x) * 256 + threadIdx.x;
  if (i >= stride) return;
  for (int h = 0; h < heads; ++h) {  // the padding rows [n, stride) read as zeros (they belong to no stored row)
    out[int64_t(h) * stride + i] = i < n ? __ldg(in + i * heads + h) : 0.f;
    out2[int64_t(h) * stride + i] = i < n ? __ldg(in2 + i * heads + h) : 0.f;
  }
}

int launch_win_gat(WinArgs a, cudaStream_t stream) {
  static const int warps = env_int_win("LGK_WIN_WARPS", 16) == 24 ? 24 : 16;  // consumer warps (tuning knob)
  a.select = 0;
  a.slices = int(ceil_div(int64_t(a.feat) * 4, 128));
  const int fixed = warps * kStageBytesGat + int(sizeof(WinShared)) + 128;
  const int per_blk = kBlkBytes + 32 * 4;
  static const int nb_cap = env_int_win("LGK_WIN_BLOCKS", 1 << 30);
  a.nb = std::min((kSmemMax - fixed) / per_blk, nb_cap);
  LGK_REQUIRE(a.nb >= 2, LGK_ERR_UNSUPPORTED, "lgk_gat_fwd_blocked: no room for the ring");
  const int smem = a.nb * per_blk + fixed;
  CUtensorMap map_x;
  LGK_REQUIRE(make_row_map(&map_x, a.x, a.n_rows, a.feat, 128), LGK_ERR_UNSUPPORTED, "lgk_gat_fwd_blocked: cannot encode the tensor map");
  const int spans = int(std::max<int64_t>(1, std::min<int64_t>(kNumSMs / std::max(1, std::min(a.slices, kNumSMs)), a.n_chunks)));
  const unsigned grid = unsigned(spans * a.slices);
  WinKernel kern = warps == 24 ? spmm_win_kernel<true, 8, 24, LGK_SELF_NONE, false, true> : spmm_win_kernel<true, 8, 16, LGK_SELF_NONE, false, true>;
  int dev = 0;
  cudaGetDevice(&dev);
  static bool configured[64] = {};
  if (!configured[dev & 63]) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemMax);
    if (e != cudaSuccess) {
      set_error("spmm_win_kernel(gat): cannot reserve %d bytes of shared memory: %s", kSmemMax, cudaGetErrorString(e));
      return LGK_ERR_LAUNCH;
    }
    configured[dev & 63] = true;
  }
  kern<<<grid, 32 * (warps + 1), smem, stream>>>(a, map_x);
  LGK_CHECK_LAUNCH("spmm_win_kernel(gat)");
  return LGK_OK;
}

}  // namespace
}  // namespace lgk

using namespace lgk;

extern "C" int64_t lgk_quad_chunks(int64_t n_rows) {
  const int64_t quads = ceil_div(n_rows, 4);
  return ceil_div(quads, kQuadsPerChunk);
}

extern "C" int lgk_quad_build(const int32_t* ptr, const int32_t* idx, int64_t n_rows, int32_t* qlen, uint32_t* qent,
                              int32_t* cmin, int32_t* cmax, int32_t* qmeta, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(n_rows >= 0, LGK_ERR_BAD_ARG, "lgk_quad_build: negative size");
  LGK_REQUIRE(qmeta, LGK_ERR_BAD_ARG, "lgk_quad_build: null pointer");
  cudaMemsetAsync(qmeta, 0, 4 * sizeof(int32_t), stream);
  if (n_rows == 0) return LGK_OK;
  LGK_REQUIRE(ptr && qlen && cmin && cmax, LGK_ERR_BAD_ARG, "lgk_quad_build: null pointer");
  LGK_REQUIRE(n_rows <= int64_t(kColMask), LGK_ERR_UNSUPPORTED, "lgk_quad_build: more than 2^28 - 1 rows");
  const int64_t quads = ceil_div(n_rows, 4);
  const int64_t chunks = ceil_div(quads, kQuadsPerChunk);
  quad_build_kernel<<<unsigned(ceil_div(quads, 128)), 128, 0, stream>>>(ptr, idx, n_rows, quads, qlen, qent, cmin, cmax, qmeta);
  LGK_CHECK_LAUNCH("quad_build_kernel");
  return LGK_OK;
}

extern "C" int lgk_spmm_blocked(const int32_t* ptr, const int32_t* qlen, const uint32_t* qent, const int32_t* cmin,
                                const int32_t* cmax, const int32_t* qmeta, const float* pre, const float* post,
                                const float* selfw, float self_fill, int self_mode, const void* x, const float* bias,
                                void* out, int64_t n_rows, int64_t feat, int reduce, int dtype, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(n_rows >= 0 && feat >= 0, LGK_ERR_BAD_ARG, "lgk_spmm_blocked: negative size");
  if (n_rows == 0 || feat == 0) return LGK_OK;
  LGK_REQUIRE(ptr && qlen && qent && cmin && cmax && qmeta && x && out, LGK_ERR_BAD_ARG, "lgk_spmm_blocked: null pointer");
  LGK_REQUIRE(dtype == LGK_F32, LGK_ERR_UNSUPPORTED, "lgk_spmm_blocked: fp32 features only");
  LGK_REQUIRE(reduce == LGK_SUM || reduce == LGK_MEAN, LGK_ERR_UNSUPPORTED, "lgk_spmm_blocked: sum or mean only");
  LGK_REQUIRE(reduce == LGK_SUM || self_mode == LGK_SELF_NONE, LGK_ERR_BAD_ARG,
              "lgk_spmm_blocked: self-loop handling is defined for the sum reduction only");
  LGK_REQUIRE(self_mode >= LGK_SELF_NONE && self_mode <= LGK_SELF_ADD, LGK_ERR_BAD_ARG, "lgk_spmm_blocked: bad self_mode");
  LGK_REQUIRE(feat % 4 == 0 && feat < (1 << 20), LGK_ERR_UNSUPPORTED, "lgk_spmm_blocked: feat must be a multiple of 4");
  LGK_REQUIRE(n_rows <= int64_t(kColMask), LGK_ERR_UNSUPPORTED, "lgk_spmm_blocked: more than 2^28 - 1 rows");
  LGK_REQUIRE(n_rows * (feat / 4) < (int64_t(1) << 31), LGK_ERR_UNSUPPORTED, "lgk_spmm_blocked: n_rows * feat / 4 must be below 2^31");
  LGK_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(bias)) & 15) == 0,
              LGK_ERR_BAD_ARG, "lgk_spmm_blocked: x, out and bias must be 16-byte aligned");
  LGK_REQUIRE(x != out, LGK_ERR_BAD_ARG, "lgk_spmm_blocked: in-place aggregation is not possible");
  LGK_REQUIRE((reinterpret_cast<uintptr_t>(pre) & 15) == 0, LGK_ERR_BAD_ARG, "lgk_spmm_blocked: pre must be 16-byte aligned");

  WinArgs a{};
  a.ptr = ptr; a.qlen = qlen; a.qent = qent; a.cmin = cmin; a.cmax = cmax; a.qmeta = qmeta;
  a.pre = pre; a.post = post; a.selfw = selfw; a.bias = bias;
  a.x = static_cast<const float*>(x);
  a.out = static_cast<float*>(out);
  a.self_fill = self_fill; a.self_mode = self_mode; a.mean = reduce == LGK_MEAN;
  a.vec_rows = ((reinterpret_cast<uintptr_t>(ptr) | reinterpret_cast<uintptr_t>(post) | reinterpret_cast<uintptr_t>(selfw)) & 15) == 0;
  a.n_rows = n_rows;
  a.n_quads = ceil_div(n_rows, 4);
  a.n_chunks = ceil_div(a.n_quads, kQuadsPerChunk);
  a.feat = int(feat);
  static const int dbg = env_int_win("LGK_WIN_DEBUG", 0);
  // 128-byte slices (ring of ~1700 source rows, units of 256 destination rows) are the default: measured faster than the
  // 256-byte variant even where the latter's ring suffices (profiles/README.md r02).  LGK_WIN_SLICE=256 forces that variant,
  // LGK_WIN_SLICE=auto launches both and lets the window statistic of lgk_quad_build choose on the device.
  static const int force = [] {
    const char* v = getenv("LGK_WIN_SLICE");
    if (!v || !*v) return 128;
    return v[0] == 'a' ? 0 : atoi(v);
  }();
  if (force == 256 && feat * 4 > 128) return launch_win<16>(a, 0, dbg, stream);
  if (force == 0 && feat * 4 > 128) {
    if (int rc = launch_win<16>(a, 1, dbg, stream)) return rc;
    return launch_win<8>(a, 1, dbg, stream);
  }
  return launch_win<8>(a, 0, dbg, stream);
}

extern "C" int64_t lgk_gat_blocked_workspace_bytes(int64_t n_rows, int heads) {
  if (n_rows <= 0 || heads <= 0) return 0;
  return 2 * int64_t(heads) * ((n_rows + 3) / 4 * 4) * 4;  // a_src and a_dst, transposed
}

extern "C" int lgk_gat_fwd_blocked(const int32_t* ptr, const int32_t* qlen, const uint32_t* qent, const int32_t* cmin,
                                   const int32_t* cmax, const int32_t* qmeta, const void* h, const float* a_src, const float* a_dst,
                                   const float* bias, void* out, void* out_elu, float* rowmax, float* rowsum, int64_t n_rows,
                                   int heads, int chan, float negative_slope, int add_self_loops, void* workspace,
                                   int64_t workspace_bytes, int dtype, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(n_rows >= 0 && heads > 0 && chan > 0, LGK_ERR_BAD_ARG, "lgk_gat_fwd_blocked: bad sizes");
  if (n_rows == 0) return LGK_OK;
  LGK_REQUIRE(ptr && qlen && qent && cmin && cmax && qmeta && h && a_src && a_dst && out, LGK_ERR_BAD_ARG,
              "lgk_gat_fwd_blocked: null pointer");
  LGK_REQUIRE(dtype == LGK_F32, LGK_ERR_UNSUPPORTED, "lgk_gat_fwd_blocked: fp32 features only");
  LGK_REQUIRE(chan % 32 == 0, LGK_ERR_UNSUPPORTED, "lgk_gat_fwd_blocked: channels per head must be a multiple of 32 (a 128-byte slice lies in one head)");
  const int64_t feat = int64_t(heads) * chan;
  LGK_REQUIRE(feat < (1 << 20), LGK_ERR_UNSUPPORTED, "lgk_gat_fwd_blocked: heads * chan too large");
  LGK_REQUIRE(n_rows <= int64_t(kColMask), LGK_ERR_UNSUPPORTED, "lgk_gat_fwd_blocked: more than 2^28 - 1 rows");
  LGK_REQUIRE(n_rows * (feat / 4) < (int64_t(1) << 31) && n_rows * heads < (int64_t(1) << 31), LGK_ERR_UNSUPPORTED,
              "lgk_gat_fwd_blocked: n_rows * feat / 4 must be below 2^31");
  LGK_REQUIRE(((reinterpret_cast<uintptr_t>(h) | reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(bias) |
                reinterpret_cast<uintptr_t>(out_elu) | reinterpret_cast<uintptr_t>(workspace)) & 15) == 0,
              LGK_ERR_BAD_ARG, "lgk_gat_fwd_blocked: h, out, out_elu, bias and workspace must be 16-byte aligned");
  LGK_REQUIRE(h != out, LGK_ERR_BAD_ARG, "lgk_gat_fwd_blocked: in-place aggregation is not possible");
  const int64_t need = lgk_gat_blocked_workspace_bytes(n_rows, heads);
  LGK_REQUIRE(workspace && workspace_bytes >= need, LGK_ERR_WORKSPACE, "lgk_gat_fwd_blocked: workspace %lld < %lld bytes",
              (long long)workspace_bytes, (long long)need);
  const int64_t stride = (n_rows + 3) / 4 * 4;
  float* a_src_t = static_cast<float*>(workspace);
  float* a_dst_t = a_src_t + int64_t(heads) * stride;
  transpose_heads_kernel<<<unsigned(ceil_div(stride, 256)), 256, 0, stream>>>(a_src, a_dst, a_src_t, a_dst_t, n_rows, heads, stride);
  LGK_CHECK_LAUNCH("transpose_heads_kernel");
  WinArgs a{};
  a.ptr = ptr; a.qlen = qlen; a.qent = qent; a.cmin = cmin; a.cmax = cmax; a.qmeta = qmeta;
  a.pre = a_src_t; a.pre_stride = stride; a.a_dst = a_dst_t; a.bias = bias;
  a.x = static_cast<const float*>(h);
  a.out = static_cast<float*>(out);
  a.act_out = static_cast<float*>(out_elu);
  a.rowmax = rowmax; a.rowsum = rowsum;
  a.heads = heads; a.chan = chan; a.slope = negative_slope; a.add_self = add_self_loops;
  a.self_mode = LGK_SELF_NONE;
  a.n_rows = n_rows;
  a.n_quads = ceil_div(n_rows, 4);
  a.n_chunks = ceil_div(a.n_quads, kQuadsPerChunk);
  a.feat = int(feat);
  return launch_win_gat(a, stream);
}
