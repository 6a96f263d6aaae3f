// (1) CSR / CSC / segment builder.
//
// Replaces torch_sparse.SparseTensor(row=dst, col=src) (reference datasets/datamodule.py:45) and adds the
// transpose index for the atomic-free backward.  Integer, HBM-bound work: the design is a bucket sort keyed
// on the node id (histogram -> exclusive scan -> bucket fill) followed by a sort INSIDE each bucket on a
// unique 64-bit key, which makes the result independent of the order the atomics happened to run in:
//
//   CSR bucket = dst, key = (src << 32 | original edge id)   -> same order as a stable sort on dst*N+src
//   CSC bucket = src, key = (CSR slot << 32 | dst)           -> out-edges of a node in CSR-slot order
//
// Buckets with <= kSmallRow entries are ranked by counting from a shared-memory window of 128 rows
// (thread per entry, coalesced window load); larger buckets get one CTA each and a bitonic network
// (shared memory up to 4096 entries, in place in global memory beyond).
#include "common.cuh"
#include <stdlib.h>

namespace lgk {
namespace {

constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;  // 2048

constexpr int kSortRows = 128;     // rows per CTA window
constexpr int kSortThreads = 256;
constexpr int kSortCap = 4096;     // keys staged in shared memory per CTA (32 KB)
constexpr int kSmallRow = 64;      // rank-by-counting up to this bucket size

// ---- histogram of dst and src in one pass over edge_index ---------------------------------------
__global__ void __launch_bounds__(256) hist_kernel(const int64_t* __restrict__ src, const int64_t* __restrict__ dst,
                                                   int64_t E, int64_t N, int* __restrict__ cnt_dst,
                                                   int* __restrict__ cnt_src) {
  int64_t stride = int64_t(gridDim.x) * blockDim.x;
  for (int64_t e = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; e < E; e += stride) {
    int64_t s = __ldg(src + e), d = __ldg(dst + e);
    if (uint64_t(s) < uint64_t(N) && uint64_t(d) < uint64_t(N)) {
      atomicAdd(cnt_dst + d, 1);
      if (cnt_src) atomicAdd(cnt_src + s, 1);
    }
  }
}

// ---- exclusive scan: ptr[0..n] from cnt[0..n), three small kernels, gridDim.y selects the array --
struct ScanArgs {
  int* in[2];
  int* out[2];
  int* sums[2];
};

__device__ __forceinline__ int block_exclusive_scan(int v, int* total, int* s_warp) {
  // 256 threads -> exclusive prefix of v across the block
  int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    int w = lane < (kScanThreads / 32) ? s_warp[lane] : 0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int t = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += t;
    }
    if (lane < (kScanThreads / 32)) s_warp[lane] = w;  // inclusive warp totals
  }
  __syncthreads();
  int warp_off = warp ? s_warp[warp - 1] : 0;
  if (total) *total = s_warp[kScanThreads / 32 - 1];
  int r = warp_off + inc - v;
  __syncthreads();
  return r;
}

__global__ void __launch_bounds__(kScanThreads) scan_reduce_kernel(ScanArgs a, int64_t n) {
  __shared__ int s_warp[32];
  const int* in = a.in[blockIdx.y];
  int64_t base = int64_t(blockIdx.x) * kScanTile;
  int sum = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    int64_t idx = base + i * kScanThreads + threadIdx.x;
    if (idx < n) sum += in[idx];
  }
  int total;
  block_exclusive_scan(sum, &total, s_warp);
  if (threadIdx.x == 0) a.sums[blockIdx.y][blockIdx.x] = total;
}

__global__ void __launch_bounds__(kScanThreads) scan_sums_kernel(ScanArgs a, int64_t nblocks) {
  __shared__ int s_warp[32];
  int* sums = a.sums[blockIdx.y];
  int carry = 0;
  for (int64_t base = 0; base < nblocks; base += kScanThreads) {
    int64_t idx = base + threadIdx.x;
    int v = idx < nblocks ? sums[idx] : 0;
    int total;
    int ex = block_exclusive_scan(v, &total, s_warp);
    if (idx < nblocks) sums[idx] = carry + ex;
    carry += total;
  }
}

__global__ void __launch_bounds__(kScanThreads) scan_apply_kernel(ScanArgs a, int64_t n) {
  __shared__ int s_warp[32];
  int* in = a.in[blockIdx.y];
  int* out = a.out[blockIdx.y];
  int64_t base = int64_t(blockIdx.x) * kScanTile + int64_t(threadIdx.x) * kScanItems;
  int v[kScanItems];
  int sum = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    v[i] = (base + i < n) ? in[base + i] : 0;
    if (base + i < n) in[base + i] = 0;  // the counts become the cursors of the bucket fill
    sum += v[i];
  }
  int total;
  int ex = block_exclusive_scan(sum, &total, s_warp) + a.sums[blockIdx.y][blockIdx.x];
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    if (base + i < n) out[base + i] = ex;
    ex += v[i];
  }
  // out[n] = grand total, written by the thread that owns element n-1
  if (base <= n - 1 && n - 1 < base + kScanItems) out[n] = ex;
}

// the whole scan in one CTA of 1024 threads per array (a batch of lesion graphs has ~10^4 nodes: three dependent
// launches of a few CTAs each cost more in launch latency than the scan itself).  Tiles of 4096 counts: thread t owns
// the 4 consecutive counts [tile + 4t, tile + 4t + 4) -- one coalesced 16-byte load, clear and store per thread and tile --
// and a running carry links the tiles.  (r01's version gave every thread one contiguous range of n / 1024 counts: 48-byte
// strides, every warp access split into 32 sectors on the single SM that runs the CTA -- 18 us for 11 k counts.)
constexpr int kSingleThreads = 1024, kSingleItems = 32;
__global__ void __launch_bounds__(kSingleThreads) scan_single_kernel(ScanArgs a, int64_t n) {
  __shared__ int s_warp[32];
  int* in = a.in[blockIdx.y];
  int* out = a.out[blockIdx.y];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int carry = 0;
  for (int64_t tile = 0; tile < n; tile += 4 * kSingleThreads) {
    const int64_t i = tile + 4 * int64_t(threadIdx.x);
    int4 v = make_int4(0, 0, 0, 0);
    const bool full = i + 3 < n;
    if (full) {
      v = *reinterpret_cast<const int4*>(in + i);
      *reinterpret_cast<int4*>(in + i) = make_int4(0, 0, 0, 0);  // the counts become the cursors of the bucket fill
    } else if (i < n) {
      v.x = in[i]; in[i] = 0;
      if (i + 1 < n) { v.y = in[i + 1]; in[i + 1] = 0; }
      if (i + 2 < n) { v.z = in[i + 2]; in[i + 2] = 0; }
    }
    const int sum = v.x + v.y + v.z + v.w;
    int inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      int w = s_warp[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += t;
      }
      s_warp[lane] = w;  // inclusive warp totals
    }
    __syncthreads();
    const int ex = carry + (warp ? s_warp[warp - 1] : 0) + inc - sum;
    carry += s_warp[31];
    if (full) {
      *reinterpret_cast<int4*>(out + i) = make_int4(ex, ex + v.x, ex + v.x + v.y, ex + v.x + v.y + v.z);
    } else if (i < n) {
      out[i] = ex;
      if (i + 1 < n) out[i + 1] = ex + v.x;
      if (i + 2 < n) out[i + 2] = ex + v.x + v.y;
    }
    __syncthreads();  // s_warp is rewritten by the next tile
  }
  if (threadIdx.x == 0) out[n] = carry;  // grand total
}

// ---- bucket fill: CSR keys ------------------------------------------------------------------------
__global__ void __launch_bounds__(256) fill_csr_kernel(const int64_t* __restrict__ src, const int64_t* __restrict__ dst,
                                                       int64_t E, int64_t N, const int* __restrict__ rowptr,
                                                       int* __restrict__ cursor, uint64_t* __restrict__ keys) {
  int64_t stride = int64_t(gridDim.x) * blockDim.x;
  for (int64_t e = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; e < E; e += stride) {
    int64_t s = __ldg(src + e), d = __ldg(dst + e);
    if (uint64_t(s) < uint64_t(N) && uint64_t(d) < uint64_t(N)) {
      int slot = rowptr[d] + atomicAdd(cursor + d, 1);
      keys[slot] = (uint64_t(uint32_t(s)) << 32) | uint64_t(uint32_t(e));
    }
  }
}

// Emit policy: what to do with an entry whose final slot inside its bucket is known.
//   CSR pass: col[f] = hi, perm[f] = lo, and (optionally) drop the entry into its CSC bucket
//   CSC pass: csr2csc[f] = hi, row[f] = lo
struct EmitArgs {
  int* out_hi;
  int* out_lo;
  // CSC bucket fill (CSR pass only; null otherwise)
  const int* colptr;
  int* cursor2;
  uint64_t* keys2;
};

__device__ __forceinline__ void emit(const EmitArgs& ea, int f, int bucket, uint64_t key) {
  int hi = int(uint32_t(key >> 32)), lo = int(uint32_t(key));
  ea.out_hi[f] = hi;
  ea.out_lo[f] = lo;
  if (ea.keys2) {
    int slot2 = ea.colptr[hi] + atomicAdd(ea.cursor2 + hi, 1);
    ea.keys2[slot2] = (uint64_t(uint32_t(f)) << 32) | uint64_t(uint32_t(bucket));
  }
}

// ---- big buckets: the whole CTA sorts one bucket with a bitonic network (ascending comparators only, so the
//      virtual +inf padding above `d` never moves) ------------------------------------------------------
template <typename KeyPtr>
__device__ __forceinline__ void bitonic_sort(KeyPtr k, int d) {
  int n2 = 1;
  while (n2 < d) n2 <<= 1;
  for (int size = 2; size <= n2; size <<= 1) {
    int half = size >> 1;
    for (int i = threadIdx.x; i < (n2 >> 1); i += blockDim.x) {  // flip
      int blk = i / half, off = i - blk * half;
      int a = blk * size + off, b = blk * size + size - 1 - off;
      if (b < d) {
        uint64_t ka = k[a], kb = k[b];
        if (ka > kb) { k[a] = kb; k[b] = ka; }
      }
    }
    __syncthreads();
    for (int j = half >> 1; j >= 1; j >>= 1) {  // disperse
      for (int i = threadIdx.x; i < (n2 >> 1); i += blockDim.x) {
        int blk = i / j, off = i - blk * j;
        int a = blk * 2 * j + off, b = a + j;
        if (b < d) {
          uint64_t ka = k[a], kb = k[b];
          if (ka > kb) { k[a] = kb; k[b] = ka; }
        }
      }
      __syncthreads();
    }
  }
}

// ---- small buckets: rank by counting, thread per entry ----------------------------------------------
// Buckets above kSmallRow entries (rare: hubs of a non-kNN graph) are sorted afterwards by the same CTA, one at a
// time -- a separate launch for them cost ~4 us per build whether or not a big bucket existed.
__global__ void __launch_bounds__(kSortThreads) sort_window_kernel(const int* __restrict__ ptr, uint64_t* __restrict__ keys,
                                                                   int64_t N, EmitArgs ea) {
  __shared__ int s_ptr[kSortRows + 1];
  __shared__ uint64_t s_keys[kSortCap];
  int64_t r0 = int64_t(blockIdx.x) * kSortRows;
  int nrows = int(min(int64_t(kSortRows), N - r0));
  for (int t = threadIdx.x; t <= nrows; t += blockDim.x) s_ptr[t] = ptr[r0 + t];
  __syncthreads();
  int w0 = s_ptr[0], wn = s_ptr[nrows] - w0;
  bool staged = wn <= kSortCap;
  if (staged) {
    for (int p = threadIdx.x; p < wn; p += blockDim.x) s_keys[p] = keys[w0 + p];
  }
  __syncthreads();
  bool saw_big = false;
  for (int p = threadIdx.x; p < wn; p += blockDim.x) {
    // row of slot p: last rr with s_ptr[rr] <= w0 + p
    int lo = 0, hi = nrows;  // invariant: s_ptr[lo] <= g < s_ptr[hi]
    int g = w0 + p;
    while (hi - lo > 1) {
      int mid = (lo + hi) >> 1;
      if (s_ptr[mid] <= g) lo = mid; else hi = mid;
    }
    int a = s_ptr[lo] - w0, b = s_ptr[lo + 1] - w0;
    if (b - a > kSmallRow) { saw_big = true; continue; }
    int rank = 0;
    uint64_t mine;
    if (staged) {
      mine = s_keys[p];
      for (int q = a; q < b; ++q) rank += (s_keys[q] < mine);
    } else {
      mine = keys[w0 + p];
      for (int q = a; q < b; ++q) rank += (keys[w0 + q] < mine);
    }
    emit(ea, w0 + a + rank, int(r0) + lo, mine);
  }
  if (!__syncthreads_or(saw_big)) return;
  for (int lo = 0; lo < nrows; ++lo) {  // CTA-uniform walk over the window's rows
    const int a = s_ptr[lo], d = s_ptr[lo + 1] - a;
    if (d <= kSmallRow) continue;
    if (d <= kSortCap) {
      for (int i = threadIdx.x; i < d; i += blockDim.x) s_keys[i] = keys[a + i];
      __syncthreads();
      bitonic_sort(s_keys, d);
      for (int i = threadIdx.x; i < d; i += blockDim.x) emit(ea, a + i, int(r0) + lo, s_keys[i]);
    } else {
      bitonic_sort(keys + a, d);
      for (int i = threadIdx.x; i < d; i += blockDim.x) emit(ea, a + i, int(r0) + lo, keys[a + i]);
    }
    __syncthreads();
  }
}

// ---- fast path for edges already grouped by destination (dst non-decreasing: what lgk_knn_graph, lgk_radius_graph and
//      the reference's per-sample kNN + collate emit).  No histogram / scan / bucket fill for the CSR side at all:
//        bounds   rowptr[r] = first edge with dst >= r (boundary detection on the sorted dst row) + histogram of src
//        rows     per window of 128 rows: src ids staged as int32, STABLE rank by counting on the 32-bit id (entries of a
//                 row are in edge-id order already, so (src, edge id) order = stable order on src alone: half the
//                 shared memory and a third of the instructions of the 64-bit key compare), col / perm written, and the
//                 entry dropped into its CSC bucket as two 4-byte values (CSR slot, dst)
//        cols     per window of 128 buckets: rank by the (unique) CSR slot, csr2csc / row written
//      Same outputs, bit for bit, as the general path on the same edges.
constexpr int kFastSmall = 256;   // rank-by-counting per entry up to this row length; longer rows: the CTA together

__global__ void __launch_bounds__(256) sorted_bounds_kernel(const int64_t* __restrict__ src, const int64_t* __restrict__ dst,
                                                            int64_t E, int64_t N, int* __restrict__ rowptr,
                                                            int* __restrict__ cnt_src, int* __restrict__ flags) {
  const int64_t stride = int64_t(gridDim.x) * blockDim.x;
  const int64_t t0 = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  for (int64_t e = t0; e < E; e += stride) {
    const int64_t d = __ldg(dst + e), dp = e > 0 ? __ldg(dst + e - 1) : -1;
    if (d < dp || uint64_t(d) >= uint64_t(N)) { flags[0] = 1; continue; }  // contract violated: reported, never out of bounds
    for (int64_t r = dp + 1; r <= d; ++r) rowptr[r] = int(e);
    if (cnt_src) {
      const int64_t s = __ldg(src + e);
      if (uint64_t(s) < uint64_t(N)) atomicAdd(cnt_src + s, 1); else flags[0] = 1;
    }
  }
  const int64_t last = __ldg(dst + E - 1);
  if (uint64_t(last) < uint64_t(N))
    for (int64_t r = last + 1 + t0; r <= N; r += stride) rowptr[r] = int(E);
}

struct FastArgs {
  const int64_t* src;   // edge_index row 0 (rows pass) / unused
  const int* tmp_slot;  // cols pass: unsorted CSR slots / dst per CSC bucket
  const int* tmp_dst;
  int* out_a;           // rows: col       cols: csr2csc
  int* out_b;           // rows: perm      cols: row
  const int* colptr;    // rows pass, CSC wanted: bucket fill
  int* cursor2;
  int* drop_slot;
  int* drop_dst;
  const int* flags;     // flags[0] != 0: the bounds pass found the contract violated -- rowptr is not to be trusted
};

// ROWS = true: CSR rows from the dst-sorted edge list; false: CSC buckets from the dropped (slot, dst) pairs
template <bool ROWS>
__global__ void __launch_bounds__(kSortThreads) sorted_window_kernel(const int* __restrict__ ptr, int64_t N, FastArgs fa) {
  __shared__ int s_ptr[kSortRows + 1];
  __shared__ int s_key[kSortCap];
  __shared__ int s_pay[ROWS ? 1 : kSortCap];
  __shared__ unsigned char s_row[kSortCap];
  if (*reinterpret_cast<const volatile int*>(fa.flags) != 0) return;
  const int64_t r0 = int64_t(blockIdx.x) * kSortRows;
  const int nrows = int(min(int64_t(kSortRows), N - r0));
  for (int t = threadIdx.x; t <= nrows; t += blockDim.x) s_ptr[t] = ptr[r0 + t];
  __syncthreads();
  const int w0 = s_ptr[0], wn = s_ptr[nrows] - w0;
  const bool staged = wn <= kSortCap;
  auto key_at = [&](int p) -> int { return ROWS ? int(__ldg(fa.src + w0 + p)) : __ldg(fa.tmp_slot + w0 + p); };
  if (staged) {
    for (int p = threadIdx.x; p < wn; p += blockDim.x) {
      s_key[p] = key_at(p);
      if (!ROWS) s_pay[p] = __ldg(fa.tmp_dst + w0 + p);
    }
    for (int rr = threadIdx.x; rr < nrows; rr += blockDim.x)
      for (int p = s_ptr[rr] - w0, pe = s_ptr[rr + 1] - w0; p < pe; ++p) s_row[p] = static_cast<unsigned char>(rr);
  }
  __syncthreads();
  auto emit = [&](int f, int bucket, int key, int p) {
    fa.out_a[f] = key;
    if (ROWS) {
      fa.out_b[f] = w0 + p;  // the edge id: rowptr indexes the (sorted) input directly
      if (fa.drop_slot) {
        const int slot2 = __ldg(fa.colptr + key) + atomicAdd(fa.cursor2 + key, 1);
        fa.drop_slot[slot2] = f;
        fa.drop_dst[slot2] = bucket;
      }
    } else {
      fa.out_b[f] = staged ? s_pay[p] : __ldg(fa.tmp_dst + w0 + p);
    }
  };
  bool saw_big = false;
  for (int p = threadIdx.x; p < wn; p += blockDim.x) {
    int rr;
    if (staged) {
      rr = s_row[p];
    } else {  // row of slot p: last rr with s_ptr[rr] <= w0 + p
      int lo = 0, hi = nrows;
      const int g = w0 + p;
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (s_ptr[mid] <= g) lo = mid; else hi = mid;
      }
      rr = lo;
    }
    const int a = s_ptr[rr] - w0, b = s_ptr[rr + 1] - w0;
    if (b - a > kFastSmall) { saw_big = true; continue; }
    int rank = 0, mine;
    if (staged) {
      mine = s_key[p];
      for (int q = a; q < p; ++q) rank += (s_key[q] <= mine);      // stable: equal ids keep their (edge id) order
      for (int q = p + 1; q < b; ++q) rank += (s_key[q] < mine);
    } else {
      mine = key_at(p);
      for (int q = a; q < p; ++q) rank += (key_at(q) <= mine);
      for (int q = p + 1; q < b; ++q) rank += (key_at(q) < mine);
    }
    emit(w0 + a + rank, int(r0) + rr, mine, p);
  }
  if (!__syncthreads_or(saw_big)) return;
  for (int rr = 0; rr < nrows; ++rr) {  // hubs: the CTA ranks one long row together, keys read through L1 / L2
    const int a = s_ptr[rr] - w0, b = s_ptr[rr + 1] - w0;
    if (b - a <= kFastSmall) continue;
    for (int p = a + threadIdx.x; p < b; p += blockDim.x) {
      const int mine = key_at(p);
      int rank = 0;
      for (int q = a; q < p; ++q) rank += (key_at(q) <= mine);
      for (int q = p + 1; q < b; ++q) rank += (key_at(q) < mine);
      emit(w0 + a + rank, int(r0) + rr, mine, p);
    }
  }
}

__global__ void __launch_bounds__(256) batch_ptr_kernel(const int64_t* __restrict__ batch, int64_t N, int64_t B,
                                                        int* __restrict__ ptr) {
  int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (N == 0) {
    if (i <= B) ptr[i] = 0;
    return;
  }
  if (i >= N) return;
  int64_t b = min(__ldg(batch + i), B - 1);
  int64_t prev = i > 0 ? min(__ldg(batch + i - 1), B - 1) : -1;
  for (int64_t g = prev + 1; g <= b; ++g) ptr[g] = int(i);
  if (i == N - 1)
    for (int64_t g = b + 1; g <= B; ++g) ptr[g] = int(N);
}

__global__ void __launch_bounds__(256) gather_f32_kernel(const float* __restrict__ in, const int* __restrict__ index,
                                                         float* __restrict__ out, int64_t n) {
  int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) out[i] = __ldg(in + ld_stream_i32(index + i));
}

struct Workspace {
  uint64_t* keys;
  int* cnt_dst;   // N+1
  int* cnt_src;   // N+1
  int* sums_a;    // nblocks
  int* sums_b;    // nblocks
  int* flags;     // 2
  int64_t bytes;
};

Workspace carve(void* base, int64_t N, int64_t E) {
  Workspace w;
  int64_t nblocks = ceil_div(N + 1, kScanTile);
  char* p = static_cast<char*>(base);
  int64_t off = 0;
  auto take = [&](int64_t bytes) {
    char* r = p ? p + off : nullptr;
    off += align_up(bytes, 256);
    return r;
  };
  w.keys = reinterpret_cast<uint64_t*>(take(16 * (E > 0 ? E : 1)));  // CSR keys | CSC keys
  w.cnt_dst = reinterpret_cast<int*>(take(4 * (N + 1)));
  w.cnt_src = reinterpret_cast<int*>(take(4 * (N + 1)));
  w.sums_a = reinterpret_cast<int*>(take(4 * nblocks));
  w.sums_b = reinterpret_cast<int*>(take(4 * nblocks));
  w.flags = reinterpret_cast<int*>(take(16));
  w.bytes = off;
  return w;
}

int grid_for(int64_t n, int threads, int max_ctas_per_sm = 8) {
  int64_t g = ceil_div(n, threads);
  int64_t cap = int64_t(kNumSMs) * max_ctas_per_sm;
  return int(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace
}  // namespace lgk

using namespace lgk;

extern "C" int64_t lgk_csr_workspace_bytes(int64_t num_nodes, int64_t num_edges) {
  if (num_nodes < 0 || num_edges < 0) return 0;
  return carve(nullptr, num_nodes, num_edges).bytes;
}

extern "C" int lgk_csr_build(const int64_t* edge_index, int64_t E, int64_t N, int32_t* rowptr, int32_t* col,
                             int32_t* perm, int32_t* colptr, int32_t* row, int32_t* csr2csc, void* workspace,
                             int64_t workspace_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(E >= 0 && N >= 0, LGK_ERR_BAD_ARG, "lgk_csr_build: negative size");
  LGK_REQUIRE(N < (int64_t(1) << 31) - 1 && E < (int64_t(1) << 31) - 1, LGK_ERR_UNSUPPORTED,
              "lgk_csr_build: N and E must be < 2^31 - 1");
  LGK_REQUIRE(rowptr && (E == 0 || (edge_index && col && perm)), LGK_ERR_BAD_ARG, "lgk_csr_build: null pointer");
  bool want_csc = colptr != nullptr;
  LGK_REQUIRE(!want_csc || (E == 0 || (row && csr2csc)), LGK_ERR_BAD_ARG, "lgk_csr_build: partial CSC outputs");
  Workspace w = carve(workspace, N, E);
  LGK_REQUIRE(workspace && workspace_bytes >= w.bytes, LGK_ERR_WORKSPACE,
              "lgk_csr_build: workspace %lld < %lld bytes", (long long)workspace_bytes, (long long)w.bytes);

  if (E == 0 || N == 0) {
    cudaMemsetAsync(rowptr, 0, 4 * (N + 1), stream);
    if (want_csc) cudaMemsetAsync(colptr, 0, 4 * (N + 1), stream);
    LGK_CHECK_LAUNCH("lgk_csr_build(memset)");
    return LGK_OK;
  }
  const int64_t* src = edge_index;
  const int64_t* dst = edge_index + E;

  // counters: cnt_dst | cnt_src | block sums are one contiguous region of the workspace: one memset
  cudaMemsetAsync(w.cnt_dst, 0, size_t(reinterpret_cast<char*>(w.flags) + 16 - reinterpret_cast<char*>(w.cnt_dst)), stream);

  hist_kernel<<<grid_for(E, 256), 256, 0, stream>>>(src, dst, E, N, w.cnt_dst, want_csc ? w.cnt_src : nullptr);
  LGK_CHECK_LAUNCH("hist_kernel");

  ScanArgs sa;
  sa.in[0] = w.cnt_dst; sa.out[0] = rowptr; sa.sums[0] = w.sums_a;
  sa.in[1] = w.cnt_src; sa.out[1] = colptr; sa.sums[1] = w.sums_b;
  int nblocks = int(ceil_div(N, kScanTile));
  dim3 sgrid(nblocks, want_csc ? 2 : 1);
  // LGK_CSR_SINGLE_SCAN=0: always the three-kernel scan (profiling knob)
  static const bool single_scan = [] { const char* v = getenv("LGK_CSR_SINGLE_SCAN"); return !(v && *v == '0'); }();
  const bool ptr16 = ((reinterpret_cast<uintptr_t>(rowptr) | reinterpret_cast<uintptr_t>(colptr)) & 15) == 0;  // 16-byte stores
  if (single_scan && ptr16 && N <= int64_t(kSingleThreads) * kSingleItems) {
    scan_single_kernel<<<dim3(1, want_csc ? 2 : 1), kSingleThreads, 0, stream>>>(sa, N);
    LGK_CHECK_LAUNCH("scan_single_kernel");
  } else {
    scan_reduce_kernel<<<sgrid, kScanThreads, 0, stream>>>(sa, N);
    LGK_CHECK_LAUNCH("scan_reduce_kernel");
    scan_sums_kernel<<<dim3(1, want_csc ? 2 : 1), kScanThreads, 0, stream>>>(sa, nblocks);
    LGK_CHECK_LAUNCH("scan_sums_kernel");
    scan_apply_kernel<<<sgrid, kScanThreads, 0, stream>>>(sa, N);
    LGK_CHECK_LAUNCH("scan_apply_kernel");
  }

  // cursors: the count arrays again -- the scan zeroed every count it read
  fill_csr_kernel<<<grid_for(E, 256), 256, 0, stream>>>(src, dst, E, N, rowptr, w.cnt_dst, w.keys);
  LGK_CHECK_LAUNCH("fill_csr_kernel");

  // CSR sort; every entry is dropped into its CSC bucket the moment its final CSR slot is known, so the
  // transpose costs no extra pass over edge_index.  CSC keys live in the second half of the key region.
  uint64_t* keys2 = want_csc ? w.keys + E : nullptr;
  EmitArgs ea1{col, perm, colptr, w.cnt_src, keys2};
  int sgrid1 = int(ceil_div(N, kSortRows));
  sort_window_kernel<<<sgrid1, kSortThreads, 0, stream>>>(rowptr, w.keys, N, ea1);
  LGK_CHECK_LAUNCH("sort_window_kernel(csr)");

  if (want_csc) {
    EmitArgs ea2{csr2csc, row, nullptr, nullptr, nullptr};
    sort_window_kernel<<<sgrid1, kSortThreads, 0, stream>>>(colptr, keys2, N, ea2);
    LGK_CHECK_LAUNCH("sort_window_kernel(csc)");
  }
  return LGK_OK;
}

extern "C" int lgk_csr_build_sorted(const int64_t* edge_index, int64_t E, int64_t N, int32_t* rowptr, int32_t* col, int32_t* perm,
                                    int32_t* colptr, int32_t* row, int32_t* csr2csc, int32_t* violation, void* workspace,
                                    int64_t workspace_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(E >= 0 && N >= 0, LGK_ERR_BAD_ARG, "lgk_csr_build_sorted: negative size");
  LGK_REQUIRE(N < (int64_t(1) << 31) - 1 && E < (int64_t(1) << 31) - 1, LGK_ERR_UNSUPPORTED,
              "lgk_csr_build_sorted: N and E must be < 2^31 - 1");
  LGK_REQUIRE(rowptr && (E == 0 || (edge_index && col && perm)), LGK_ERR_BAD_ARG, "lgk_csr_build_sorted: null pointer");
  const bool want_csc = colptr != nullptr;
  LGK_REQUIRE(!want_csc || (E == 0 || (row && csr2csc)), LGK_ERR_BAD_ARG, "lgk_csr_build_sorted: partial CSC outputs");
  Workspace w = carve(workspace, N, E);
  LGK_REQUIRE(workspace && workspace_bytes >= w.bytes, LGK_ERR_WORKSPACE, "lgk_csr_build_sorted: workspace %lld < %lld bytes",
              (long long)workspace_bytes, (long long)w.bytes);
  if (E == 0 || N == 0) {
    cudaMemsetAsync(rowptr, 0, 4 * (N + 1), stream);
    if (want_csc) cudaMemsetAsync(colptr, 0, 4 * (N + 1), stream);
    if (violation) cudaMemsetAsync(violation, 0, 4, stream);
    LGK_CHECK_LAUNCH("lgk_csr_build_sorted(memset)");
    return LGK_OK;
  }
  const int64_t* src = edge_index;
  const int64_t* dst = edge_index + E;
  // cnt_src | block sums | flags: one contiguous region, one memset (cnt_dst is not used on this path)
  cudaMemsetAsync(w.cnt_src, 0, size_t(reinterpret_cast<char*>(w.flags) + 16 - reinterpret_cast<char*>(w.cnt_src)), stream);
  int* flag = violation ? violation : w.flags;
  if (violation) cudaMemsetAsync(violation, 0, 4, stream);
  sorted_bounds_kernel<<<grid_for(E, 256), 256, 0, stream>>>(src, dst, E, N, rowptr, want_csc ? w.cnt_src : nullptr, flag);
  LGK_CHECK_LAUNCH("sorted_bounds_kernel");
  if (want_csc) {
    ScanArgs sa;
    sa.in[0] = w.cnt_src; sa.out[0] = colptr; sa.sums[0] = w.sums_b;
    sa.in[1] = nullptr; sa.out[1] = nullptr; sa.sums[1] = nullptr;
    const int nblocks = int(ceil_div(N, kScanTile));
    if ((reinterpret_cast<uintptr_t>(colptr) & 15) == 0 && N <= int64_t(kSingleThreads) * kSingleItems) {
      scan_single_kernel<<<dim3(1, 1), kSingleThreads, 0, stream>>>(sa, N);
      LGK_CHECK_LAUNCH("scan_single_kernel");
    } else {
      scan_reduce_kernel<<<dim3(nblocks, 1), kScanThreads, 0, stream>>>(sa, N);
      LGK_CHECK_LAUNCH("scan_reduce_kernel");
      scan_sums_kernel<<<dim3(1, 1), kScanThreads, 0, stream>>>(sa, nblocks);
      LGK_CHECK_LAUNCH("scan_sums_kernel");
      scan_apply_kernel<<<dim3(nblocks, 1), kScanThreads, 0, stream>>>(sa, N);
      LGK_CHECK_LAUNCH("scan_apply_kernel");
    }
  }
  int* tmp_slot = reinterpret_cast<int*>(w.keys);
  int* tmp_dst = tmp_slot + E;
  const int windows = int(ceil_div(N, kSortRows));
  FastArgs fr{src, nullptr, nullptr, col, perm, colptr, w.cnt_src, want_csc ? tmp_slot : nullptr, tmp_dst, flag};
  sorted_window_kernel<true><<<windows, kSortThreads, 0, stream>>>(rowptr, N, fr);
  LGK_CHECK_LAUNCH("sorted_window_kernel(rows)");
  if (want_csc) {
    FastArgs fc{nullptr, tmp_slot, tmp_dst, csr2csc, row, nullptr, nullptr, nullptr, nullptr, flag};
    sorted_window_kernel<false><<<windows, kSortThreads, 0, stream>>>(colptr, N, fc);
    LGK_CHECK_LAUNCH("sorted_window_kernel(cols)");
  }
  return LGK_OK;
}

extern "C" int lgk_batch_ptr(const int64_t* batch, int64_t N, int64_t B, int32_t* ptr, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(N >= 0 && B >= 0 && ptr && (N == 0 || batch), LGK_ERR_BAD_ARG, "lgk_batch_ptr: bad argument");
  LGK_REQUIRE(N == 0 || B > 0, LGK_ERR_BAD_ARG, "lgk_batch_ptr: nodes without graphs");
  int64_t n = N > 0 ? N : B + 1;
  batch_ptr_kernel<<<int(ceil_div(n, 256)), 256, 0, stream>>>(batch, N, B, ptr);
  LGK_CHECK_LAUNCH("batch_ptr_kernel");
  return LGK_OK;
}

extern "C" int lgk_gather_f32(const float* in, const int32_t* index, float* out, int64_t n, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  LGK_REQUIRE(n >= 0 && (n == 0 || (in && index && out)), LGK_ERR_BAD_ARG, "lgk_gather_f32: bad argument");
  if (n == 0) return LGK_OK;
  gather_f32_kernel<<<int(ceil_div(n, 256)), 256, 0, stream>>>(in, index, out, n);
  LGK_CHECK_LAUNCH("gather_f32_kernel");
  return LGK_OK;
}
