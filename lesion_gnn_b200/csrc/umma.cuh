// Minimal sm_100a tensor-core plumbing for liblgk: mbarrier, tcgen05 (alloc / mma / commit / ld) and the shared
// memory matrix descriptor for K-major, 128-byte-swizzled operand tiles written by ordinary threads (no TMA:
// the 1025-float rows of the reference's lesion features are not 16-byte aligned, so tiles are staged by hand).
// Bit layouts follow the PTX ISA "tcgen05 matrix descriptor" / "instruction descriptor" tables.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace lgk {
namespace umma {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

// ---- mbarrier ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_init_fence() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ bool mbar_test_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// latency-critical waits (the per-K-block hand-offs): non-blocking test in a spin loop -- try_wait may suspend the
// warp for a system-dependent time, which measured ~0.4 us per hand-off
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_test_wait(bar, parity)) {
  }
}
// for waits that are expected to be long (a whole tile): back off so the polling does not eat issue slots
__device__ __forceinline__ void mbar_wait_relaxed(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) __nanosleep(200);
}
// explicit shared-window accesses with 32-bit addresses (keeps ptxas away from generic 64-bit pointer arithmetic)
__device__ __forceinline__ float lds_f32(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_f32(uint32_t addr, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory"); }

// bulk asynchronous copy global -> shared (the TMA engine without a tensor map): `bytes` and both addresses are
// multiples of 16; completion is signalled on `bar` as `bytes` of transaction count
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// generic-proxy shared-memory writes -> visible to the async proxy (tensor core operand reads)
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- tensor memory ------------------------------------------------------------------------------------------
// one full warp; writes the TMEM base address of `cols` columns (power of two >= 32) to *slot (shared memory)
__device__ __forceinline__ void tmem_alloc(uint32_t* slot, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// ---- descriptors --------------------------------------------------------------------------------------------
// K-major operand tile, SWIZZLE_128B: rows of 128 bytes (32 tf32 / 64 bf16 along K), 8-row groups 1024 bytes
// apart (SBO), 16-byte chunk index XOR (row & 7).  The tile base must be 1024-byte aligned; K steps inside the
// 128-byte row advance the start address (+32 bytes per tf32 MMA of K = 8).
__device__ __forceinline__ uint64_t smem_desc_k_sw128(uint32_t saddr) {
  return uint64_t((saddr & 0x3FFFFu) >> 4)   // start address, bits [0,14)
         | (uint64_t(1) << 16)               // leading byte offset (unused for swizzled K-major), bits [16,30)
         | (uint64_t(1024 >> 4) << 32)       // stride byte offset, bits [32,46)
         | (uint64_t(1) << 46)               // descriptor version 1 (sm_100), bits [46,48)
         | (uint64_t(2) << 61);              // layout type SWIZZLE_128B, bits [61,64)
}
// byte offset of element (row, k) of fp32/tf32 inside such a tile (k < 32)
__device__ __forceinline__ uint32_t sw128_offset_f32(uint32_t row, uint32_t k) {
  return row * 128u + ((((k >> 2) ^ row) & 7u) << 4) + ((k & 3u) << 2);
}

// instruction descriptor, kind::tf32, fp32 accumulate, A and B K-major, dense
__host__ __device__ constexpr uint32_t idesc_tf32(int m, int n) {
  return (1u << 4)                  // D format F32, bits [4,6)
         | (2u << 7)                // A format TF32, bits [7,10)
         | (2u << 10)               // B format TF32, bits [10,13)
         | (uint32_t(n >> 3) << 17) // N / 8, bits [17,23)
         | (uint32_t(m >> 4) << 24);// M / 16, bits [24,29)
}

// D[tmem] (+)= A[smem] * B[smem]^T, one elected thread
__device__ __forceinline__ void mma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, bool accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(uint32_t(accumulate))
      : "memory");
}
// arrive on `bar` when every MMA issued so far by this thread has completed (implies fence::before_thread_sync)
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// warp-wide: 16 consecutive fp32 columns of this thread's TMEM lane (lane = 32 * (warp % 4) + laneid)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n\t"
      "tcgen05.wait::ld.sync.aligned;"  // same statement: no consumer of r[] can be scheduled before the wait
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// x = hi + lo with hi exactly representable in TF32 (round to nearest, ties away) and lo = x - hi exact in fp32
__device__ __forceinline__ void split_tf32(float x, float& hi, float& lo) {
  uint32_t h;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(x));
  hi = __uint_as_float(h);
  lo = x - hi;
}

}  // namespace umma
}  // namespace lgk
