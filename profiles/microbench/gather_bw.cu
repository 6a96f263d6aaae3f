// Microbenchmark: how fast can one B200 SM pull gathered feature rows into registers?
// Every aggregation kernel of liblgk (spmm, gat fwd/bwd) moves E x F x s bytes through this path, so its
// throughput -- not HBM -- bounds them once the graph has more than a couple of edges per node.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather_bw gather_bw.cu && ./gather_bw
//
// Each warp gathers pseudo-random 512-byte rows (128 fp32) from a table and sums them; 8 independent loads are
// kept in flight.  Table sizes: 64 KB (L1 resident), 32 MB (L2 resident), 2 GB (HBM).  Load shapes:
//   v4x1   one LDG.128 per lane, a warp covers the 512-byte row          (what spmm_vec_kernel does, F = 128)
//   v2x2   two LDG.64 per lane, each instruction covers 256 contiguous bytes
//   v1x4   four LDG.32 per lane, each instruction covers one 128-byte line
//   v8h    one LDG.256 per lane, 16 lanes per row, two different rows per instruction
//   v4h    one LDG.128 per lane, 16 lanes per 256-byte row (F = 64), two rows per instruction
//   lds4   LDS.128 from a shared-memory copy of the 64 KB table
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdint.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

constexpr int UNR = 8;

__device__ __forceinline__ uint32_t next_row(uint32_t& s, uint32_t mask) {
  s = s * 1664525u + 1013904223u;
  return (s >> 8) & mask;
}

template <int MODE>
__global__ void __launch_bounds__(512) gather_kernel(const float* __restrict__ tab, uint32_t row_mask, int iters, float* __restrict__ out) {
  extern __shared__ float4 smem[];
  const int lane = threadIdx.x & 31;
  const uint32_t warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  uint32_t s = warp_global * 2654435761u + 12345u;
  if (MODE == 5) {
    for (int i = threadIdx.x; i < int(row_mask + 1) * 32; i += blockDim.x) smem[i] = reinterpret_cast<const float4*>(tab)[i];
    __syncthreads();
  }
  float4 acc = make_float4(0, 0, 0, 0);
  for (int it = 0; it < iters; ++it) {
    uint32_t r[UNR];
#pragma unroll
    for (int u = 0; u < UNR; ++u) r[u] = next_row(s, row_mask);  // warp-uniform row ids
    if (MODE == 0) {
      float4 v[UNR];
#pragma unroll
      for (int u = 0; u < UNR; ++u) v[u] = __ldg(reinterpret_cast<const float4*>(tab + size_t(r[u]) * 128) + lane);
#pragma unroll
      for (int u = 0; u < UNR; ++u) { acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w; }
    } else if (MODE == 1) {
      float2 v[UNR][2];
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        v[u][0] = __ldg(reinterpret_cast<const float2*>(tab + size_t(r[u]) * 128) + lane);
        v[u][1] = __ldg(reinterpret_cast<const float2*>(tab + size_t(r[u]) * 128 + 64) + lane);
      }
#pragma unroll
      for (int u = 0; u < UNR; ++u) { acc.x += v[u][0].x; acc.y += v[u][0].y; acc.z += v[u][1].x; acc.w += v[u][1].y; }
    } else if (MODE == 2) {
      float v[UNR][4];
#pragma unroll
      for (int u = 0; u < UNR; ++u)
#pragma unroll
        for (int c = 0; c < 4; ++c) v[u][c] = __ldg(tab + size_t(r[u]) * 128 + c * 32 + lane);
#pragma unroll
      for (int u = 0; u < UNR; ++u) { acc.x += v[u][0]; acc.y += v[u][1]; acc.z += v[u][2]; acc.w += v[u][3]; }
    } else if (MODE == 3) {
      // two rows per instruction: lanes 0-15 row r, lanes 16-31 row r^1 ... 32 bytes per lane
      float v[UNR / 2][8];
#pragma unroll
      for (int u = 0; u < UNR / 2; ++u) {
        const uint32_t rr = (lane < 16) ? r[2 * u] : r[2 * u + 1];
        const float* p = tab + size_t(rr) * 128 + (lane & 15) * 8;
        asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=f"(v[u][0]), "=f"(v[u][1]), "=f"(v[u][2]), "=f"(v[u][3]), "=f"(v[u][4]), "=f"(v[u][5]), "=f"(v[u][6]), "=f"(v[u][7])
                     : "l"(p));
      }
#pragma unroll
      for (int u = 0; u < UNR / 2; ++u) { acc.x += v[u][0] + v[u][4]; acc.y += v[u][1] + v[u][5]; acc.z += v[u][2] + v[u][6]; acc.w += v[u][3] + v[u][7]; }
    } else if (MODE == 4) {
      // F = 64: 256-byte rows, 16 lanes each, two rows per LDG.128 (counts UNR rows of 256 B per iteration)
      float4 v[UNR / 2];
#pragma unroll
      for (int u = 0; u < UNR / 2; ++u) {
        const uint32_t rr = (lane < 16) ? r[2 * u] : r[2 * u + 1];
        v[u] = __ldg(reinterpret_cast<const float4*>(tab + size_t(rr) * 128) + (lane & 15));
      }
#pragma unroll
      for (int u = 0; u < UNR / 2; ++u) { acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w; }
    } else if (MODE == 5) {
      float4 v[UNR];
#pragma unroll
      for (int u = 0; u < UNR; ++u) v[u] = smem[r[u] * 32 + lane];
#pragma unroll
      for (int u = 0; u < UNR; ++u) { acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w; }
    }
  }
  if (acc.x + acc.y + acc.z + acc.w == 1234.5678f) out[0] = acc.x;  // keep the loads alive
}

template <int MODE>
void run(const char* name, const float* tab, uint32_t rows, int threads, int ctas_per_sm, float* out, int bytes_per_unit, int units_per_iter) {
  const int sms = 148;
  const int iters = rows > (1u << 20) ? 200 : 2000;
  size_t smem = MODE == 5 ? size_t(rows) * 512 : 0;
  if (smem) CK(cudaFuncSetAttribute(gather_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  gather_kernel<MODE><<<sms * ctas_per_sm, threads, smem>>>(tab, rows - 1, 10, out);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  gather_kernel<MODE><<<sms * ctas_per_sm, threads, smem>>>(tab, rows - 1, iters, out);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  float ms;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  const double warps = double(sms) * ctas_per_sm * threads / 32;
  const double units = warps * iters * units_per_iter;  // row gathers
  const double bytes = units * bytes_per_unit;
  int clk_khz = 0;
  cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
  const double cyc_per_unit_sm = (ms * 1e-3 * clk_khz * 1e3) / (units / sms);
  printf("{\"shape\": \"%s\", \"table_rows\": %u, \"threads\": %d, \"ctas_per_sm\": %d, \"ms\": %.3f, \"TBps\": %.2f, \"bytes_per_row\": %d, \"cycles_per_row_per_sm_at_max_clock\": %.2f}\n",
         name, rows, threads, ctas_per_sm, ms, bytes / (ms * 1e-3) / 1e12, bytes_per_unit, cyc_per_unit_sm);
}

int main() {
  const size_t big_rows = size_t(1) << 22;  // 2 GB
  float *tab, *out;
  CK(cudaMalloc(&tab, big_rows * 512));
  CK(cudaMemset(tab, 0, big_rows * 512));
  CK(cudaMalloc(&out, 4));
  const uint32_t sizes[3] = {128, 1u << 16, uint32_t(big_rows)};
  for (int occ = 0; occ < 2; ++occ) {
    const int threads = 512, ctas = occ ? 4 : 2;
    for (int si = 0; si < 3; ++si) {
      run<0>("v4x1", tab, sizes[si], threads, ctas, out, 512, UNR);
      run<1>("v2x2", tab, sizes[si], threads, ctas, out, 512, UNR);
      run<2>("v1x4", tab, sizes[si], threads, ctas, out, 512, UNR);
      run<3>("v8h", tab, sizes[si], threads, ctas, out, 512, UNR);
      run<4>("v4h_256B_rows", tab, sizes[si], threads, ctas, out, 256, UNR);
    }
    if (!occ) run<5>("lds4", tab, 128, threads, ctas, out, 512, UNR);
  }
  return 0;
}
