// How much of a kernel-to-kernel boundary inside a captured CUDA graph does programmatic dependent launch hide?
// A chain of dependent small kernels (y = f(x), ping-pong buffers), captured once and replayed, with and without the
// programmatic-stream-serialization launch attribute (+ griddepcontrol.wait / launch_dependents in the kernel).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pdl_chain pdl_chain.cu ; run: ./pdl_chain
#include <cuda_runtime.h>
#include <cstdio>
#include <vector>

__global__ void __launch_bounds__(256) work(const float4* __restrict__ x, float4* __restrict__ y, int n4, int pdl) {
  if (pdl) {
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
  }
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += gridDim.x * blockDim.x) {
    float4 v = x[i];
    v.x = v.x * 1.0001f + 1.f; v.y = v.y * 1.0001f + 1.f; v.z = v.z * 1.0001f + 1.f; v.w = v.w * 1.0001f + 1.f;
    y[i] = v;
  }
}

static float run(int n4, int chain, bool pdl, int grid) {
  float4 *a, *b;
  cudaMalloc(&a, size_t(n4) * 16); cudaMalloc(&b, size_t(n4) * 16);
  cudaMemset(a, 0, size_t(n4) * 16);
  cudaStream_t s; cudaStreamCreate(&s);
  cudaGraph_t g; cudaGraphExec_t ge;
  cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
  for (int k = 0; k < chain; ++k) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(256); cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    const float4* x = (k & 1) ? b : a; float4* y = (k & 1) ? a : b;
    cudaLaunchKernelEx(&cfg, work, x, y, n4, int(pdl));
  }
  cudaStreamEndCapture(s, &g);
  cudaGraphInstantiate(&ge, g, 0);
  for (int i = 0; i < 5; ++i) cudaGraphLaunch(ge, s);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0, s);
  const int reps = 50;
  for (int i = 0; i < reps; ++i) cudaGraphLaunch(ge, s);
  cudaEventRecord(e1, s);
  cudaStreamSynchronize(s);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  std::vector<float> h(4);
  cudaMemcpy(h.data(), (chain & 1) ? b : a, 16, cudaMemcpyDeviceToHost);
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) printf("error: %s\n", cudaGetErrorString(err));
  cudaFree(a); cudaFree(b);
  return ms * 1000.f / (reps * chain);
}

int main() {
  const int chain = 40;
  for (int n4 : {1 << 12, 1 << 16, 1 << 18, 1 << 20}) {
    for (int grid : {148, 592}) {
      float t0 = run(n4, chain, false, grid), t1 = run(n4, chain, true, grid);
      printf("{\"bytes_per_kernel\": %d, \"grid\": %d, \"us_per_kernel_plain\": %.2f, \"us_per_kernel_pdl\": %.2f}\n", n4 * 32, grid, t0, t1);
    }
  }
  return 0;
}
