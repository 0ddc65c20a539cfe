// tcgen05 / TMEM / mbarrier primitives and the operand-tile layout of the fp32 sweep path (accumulate_tc.cuh).
//
// The contraction of a chunk, G = SUM_b w_b V_b V_b^T (V_b: D x k1, SSVI_TF_d.py:47-56), runs on the 5th-generation tensor
// cores.  tcgen05.mma has no FP64 kind, so this path is the fp32 option of the contract (BASELINE.json: 1e-4): V is formed
// in fp32, split into two fp16 pieces (hi = rn(a V), lo = rn(a V - hi): 22 significand bits, a = a power of two that
// centres the fp16 range) and  hi hi^T + hi lo^T + lo hi^T  is accumulated in fp32 in tensor memory.
//
// Operand tile (shared memory, written from registers): K-major, NO swizzle -- the canonical "interleaved" layout of the
// UMMA shared-memory descriptor: core matrices of 8 rows x 16 bytes stored as 128 contiguous bytes, consecutive core
// matrices along K LBO bytes apart, consecutive 8-row groups SBO bytes apart.  With k1 = 64 fp16 elements per row:
//     byte(r, k) = (r >> 3) * 1024 + (k >> 3) * 128 + (r & 7) * 16 + (k & 7) * 2        LBO = 128, SBO = 1024
// One tcgen05.mma of kind::f16 consumes K = 16 (two core matrices): K-step s starts 256 * s bytes into the tile.
#pragma once
#include <cuda_fp16.h>
#include <stdint.h>
#include "tc_layout.cuh"

#define TC_LBO 128
#define TC_SBO 1024
#define TC_TILE_ROWS 128                               // four warps x 32 rows
#define TC_TILE_BYTES (TC_TILE_ROWS / 8 * TC_SBO)      // one piece (hi or lo) of one buffer: 16 KB

__device__ __forceinline__ uint32_t tc_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- mbarrier ----------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void tc_mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(tc_smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void tc_mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(tc_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool tc_mbar_try(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n"
      "selp.u32 %0, 1, 0, P1;\n"
      "}\n"
      : "=r"(ok)
      : "r"(tc_smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// A phase that never completes is a protocol bug: trap (the launch fails loudly) instead of hanging the GPU.
__device__ __forceinline__ void tc_mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!tc_mbar_try(bar, parity))
    if (++spins > (1u << 22)) __trap();
}
// Counting arrival with release + acquire semantics in the CTA scope: returns the count before this arrival.  The warp
// that finds itself the last of a group both sees the others' writes and may act for the group.
__device__ __forceinline__ uint32_t tc_arrive_count(uint32_t* cnt) {
  uint32_t old;
  asm volatile("atom.acq_rel.cta.shared::cta.add.u32 %0, [%1], 1;\n" : "=r"(old) : "r"(tc_smem_u32(cnt)) : "memory");
  return old;
}
// generic-proxy writes (st.shared of the operand tiles) -> visible to the async proxy (the tensor core's reads)
__device__ __forceinline__ void tc_fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }

// ---- tensor memory -----------------------------------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }
// whole warp; the base address (lane 0, first column) lands in *holder (shared memory)
__device__ __forceinline__ void tc_tmem_alloc(uint32_t* holder, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(tc_smem_u32(holder)), "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void tc_tmem_dealloc(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(taddr), "r"(cols) : "memory");
}
// 32 lanes x 32 consecutive columns: thread t of the warp gets TMEM lane (lane field of taddr) + t, columns c .. c + 31
__device__ __forceinline__ void tc_tmem_ld32(uint32_t taddr, float (&d)[32]) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) d[i] = __uint_as_float(r[i]);
}

__device__ __forceinline__ void tc_tmem_ld16(uint32_t taddr, float (&d)[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) d[i] = __uint_as_float(r[i]);
}

// ---- descriptors -------------------------------------------------------------------------------------------------------
// shared-memory matrix descriptor (cute::UMMA::SmemDescriptor): start >> 4 [0,14) | LBO >> 4 [16,30) | SBO >> 4 [32,46) |
// version 1 [46,48) | base offset 0 [49,52) | layout type [61,64) = 0 (no swizzle)
__device__ __forceinline__ uint64_t tc_smem_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr >> 4) & 0x3fffu) | ((uint64_t)((lbo >> 4) & 0x3fffu) << 16) |
         ((uint64_t)((sbo >> 4) & 0x3fffu) << 32) | (1ull << 46);
}
// instruction descriptor (cute::UMMA::InstrDescriptor), kind::f16: D fp32 [4,6) = 1, A / B fp16 (0), both K-major (0),
// N >> 3 at [17,23), M >> 4 at [24,29)
__host__ __device__ constexpr uint32_t tc_idesc_f16(int M, int N) {
  return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// D[tmem] (+)= A[smem] B[smem]^T, issued by ONE thread
__device__ __forceinline__ void tc_mma_f16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on an mbarrier once every tcgen05.mma issued so far by this thread has completed (implies fence::before_thread_sync)
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(tc_smem_u32(bar)) : "memory");
}

// ---- operand tile ------------------------------------------------------------------------------------------------------
// 8 consecutive-k values of row r (k = 8 * k8 .. 8 * k8 + 7), split into fp16 hi / lo and stored as one 16-byte vector
// each (the caller has scaled them into the fp16 window).  The 8 lanes of a quarter-warp write 8 consecutive rows of one core matrix: 128 contiguous bytes.
__device__ __forceinline__ void tc_store_split8(unsigned char* tile_hi, unsigned char* tile_lo, int r, int k8,
                                                const float (&x)[8]) {
  uint32_t hi[4], lo[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float a = x[2 * i], b = x[2 * i + 1];
    const __half2 h = __floats2half2_rn(a, b);
    const float2 hf = __half22float2(h);
    const __half2 l = __floats2half2_rn(a - hf.x, b - hf.y);
    hi[i] = *reinterpret_cast<const uint32_t*>(&h);
    lo[i] = *reinterpret_cast<const uint32_t*>(&l);
  }
  const int off = (r >> 3) * TC_SBO + k8 * TC_LBO + (r & 7) * 16;
  *reinterpret_cast<uint4*>(tile_hi + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
  *reinterpret_cast<uint4*>(tile_lo + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
}
