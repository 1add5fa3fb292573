// sm_100a tensor-core plumbing for libpaged_b200: TMA tensor-map loads, UMMA (tcgen05.mma)
// shared-memory / instruction descriptors, tensor-memory (TMEM) allocation and access.
// Inline PTX only; descriptor bit layouts follow the PTX ISA "tcgen05" matrix-descriptor and
// instruction-descriptor tables.
#pragma once

#include <cuda.h>  // CUtensorMap (types only; the encode entry point is resolved at run time)

#include "pa_common.cuh"

namespace pa {
namespace tc05 {

#ifdef __CUDACC__

// ---------------------------------------------------------------------------
// shared-memory matrix descriptor (64 bit)
//   [ 0,14) start address >> 4        [16,30) leading byte offset >> 4
//   [32,46) stride byte offset >> 4   [46,48) version = 1 (sm_100)
//   [49,52) base offset = 0           [61,64) swizzle: 0 none, 2 128 B, 4 64 B, 6 32 B
// 128-byte swizzle, operand rows of 128 bytes (64 x 16-bit), 8-row groups 1024 bytes apart:
//   K-major  (rows = M or N, 128 B of K per row): SBO = 1024 B between 8-row groups; LBO unused (1)
//   MN-major (rows = K, 128 B of M/N per row):    SBO = 1024 B between 8-row K groups,
//                                                 LBO = distance between 64-element M/N chunks
// ---------------------------------------------------------------------------
constexpr uint64_t kDescSw128 = (uint64_t{1} << 46) | (uint64_t{2} << 61);

__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t smem_byte_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return kDescSw128 | static_cast<uint64_t>((smem_byte_addr & 0x3ffffu) >> 4) |
         (static_cast<uint64_t>((lbo_bytes >> 4) & 0x3fffu) << 16) |
         (static_cast<uint64_t>((sbo_bytes >> 4) & 0x3fffu) << 32);
}

// Any canonical layout: swizzle code 0 (none: 8-row x 16-byte core matrices of 128 contiguous bytes), 2 (128 B),
// 4 (64 B), 6 (32 B).  K-major operands (cute's canonical forms, in 16-byte units):
//   none  ((8,n),2):((1,SBO),LBO)   rows 16 B apart inside a core matrix, 8-row groups SBO apart, the two
//                                   16-byte K chunks of one MMA k-step LBO apart
//   32 B  ((8,n),2):((2,SBO),1)     rows 32 B apart (one k-step per row), 8-row groups SBO (= 256 B dense) apart
//   64 B  ((8,n),2):((4,SBO),1)     rows 64 B apart (two k-steps per row, +32 B start address for the second)
__device__ __forceinline__ uint64_t smem_desc(uint32_t smem_byte_addr, uint32_t lbo_bytes, uint32_t sbo_bytes,
                                              uint32_t swizzle_code) {
  return (uint64_t{1} << 46) | (static_cast<uint64_t>(swizzle_code) << 61) |
         static_cast<uint64_t>((smem_byte_addr & 0x3ffffu) >> 4) |
         (static_cast<uint64_t>((lbo_bytes >> 4) & 0x3fffu) << 16) |
         (static_cast<uint64_t>((sbo_bytes >> 4) & 0x3fffu) << 32);
}

// ---------------------------------------------------------------------------
// instruction descriptor of tcgen05.mma.kind::f16 (32 bit, passed in the upper half of nothing:
// the PTX operand is the 32-bit value itself)
//   [4,6) D format: 1 = f32      [7,10) A format: 0 f16, 1 bf16     [10,13) B format
//   [15] A major: 0 K, 1 MN      [16] B major                      [17,23) N >> 3    [24,29) M >> 4
// ---------------------------------------------------------------------------
__host__ __device__ constexpr uint32_t instr_desc_f16(bool bf16, int m, int n, bool a_mn_major, bool b_mn_major) {
  return (1u << 4) | ((bf16 ? 1u : 0u) << 7) | ((bf16 ? 1u : 0u) << 10) | ((a_mn_major ? 1u : 0u) << 15) |
         ((b_mn_major ? 1u : 0u) << 16) | (static_cast<uint32_t>(n >> 3) << 17) | (static_cast<uint32_t>(m >> 4) << 24);
}

// D[tmem] (+)= A[smem] . B[smem]        (one thread issues for the CTA)
__device__ __forceinline__ void mma_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, bool accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      :: "r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(static_cast<uint32_t>(accumulate)) : "memory");
}
// D[tmem] (+)= A[tmem] . B[smem]        (A: lane = row, 32-bit column c = K elements 2c, 2c+1)
__device__ __forceinline__ void mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, bool accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      :: "r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(static_cast<uint32_t>(accumulate)) : "memory");
}
// The same two for an issuer WARP that stays converged: every lane executes the call with the same operands,
// the instruction itself is predicated on `leader` (one lane, chosen once with elect_one()).  A divergent
// `if (lane == 0)` around tcgen05.mma makes ptxas wrap every single instruction in a loop that moves its operands
// into uniform registers (R2UR / ELECT / BRA.U.ANY, ~15 instructions, ~70 clk per MMA: slower than the 64 clk the
// tensor core needs for it); with converged control flow the operands live in uniform registers to begin with.
__device__ __forceinline__ bool elect_one() {
  uint32_t e;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(e));
  return e != 0;
}
__device__ __forceinline__ void mma_ss_w(bool leader, uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                         bool accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, e;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "setp.ne.b32 e, %5, 0;\n\t"
      "@e tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      :: "r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(static_cast<uint32_t>(accumulate)),
         "r"(static_cast<uint32_t>(leader)) : "memory");
}
__device__ __forceinline__ void mma_ts_w(bool leader, uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                         bool accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, e;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "setp.ne.b32 e, %5, 0;\n\t"
      "@e tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      :: "r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(static_cast<uint32_t>(accumulate)),
         "r"(static_cast<uint32_t>(leader)) : "memory");
}
__device__ __forceinline__ void mma_commit_w(bool leader, uint32_t bar) {
  asm volatile(
      "{\n\t.reg .pred e;\n\t"
      "setp.ne.b32 e, %1, 0;\n\t"
      "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}"
      :: "r"(bar), "r"(static_cast<uint32_t>(leader)) : "memory");
}
// mbarrier arrive once every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void mma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(bar) : "memory");
}
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// ---------------------------------------------------------------------------
// tensor memory: 128 lanes x 512 columns x 32 bit; address = lane << 16 | column
// ---------------------------------------------------------------------------
template <int kCols>
__device__ __forceinline__ void tmem_alloc(uint32_t smem_slot) {  // whole warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_slot), "n"(kCols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int kCols>
__device__ __forceinline__ void tmem_free(uint32_t taddr) {       // whole warp (the allocating one)
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(taddr), "n"(kCols) : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// warp-wide: thread l of warp w reads lane 32 (w % 4) + l, 32 consecutive columns from taddr
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%32], "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31};"
      :: "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
         "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]),
         "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]),
         "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31]),
         "r"(taddr) : "memory");
}

// ---------------------------------------------------------------------------
// TMA: 3-D tiled tensor-map load into this CTA's shared memory, completion on an mbarrier
// ---------------------------------------------------------------------------
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, int c0, int c1, int c2, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      :: "r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, int c0, int c1, int c2, int c3, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      :: "r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
// Cluster launch control: a running CTA asks the hardware scheduler for the next CTA of the grid that has not
// started yet.  16-byte answer in shared memory, completion (16 bytes) counted on `bar`.  Canceled = that CTA will
// never be launched and its blockIdx is ours to work on; not canceled = the grid has no pending CTAs left.
__device__ __forceinline__ void clc_try_cancel(uint32_t result, uint32_t bar) {
  asm volatile("clusterlaunchcontrol.try_cancel.async.shared::cta.mbarrier::complete_tx::bytes.b128 [%0], [%1];"
               :: "r"(result), "r"(bar) : "memory");
}
__device__ __forceinline__ bool clc_read(uint32_t result, int& x, int& y) {
  uint32_t ok, bx = 0, by = 0, bz = 0;
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b128 r;\n\t"
      "ld.shared.b128 r, [%4];\n\t"
      "clusterlaunchcontrol.query_cancel.is_canceled.pred.b128 p, r;\n\t"
      "selp.u32 %3, 1, 0, p;\n\t"
      "@p clusterlaunchcontrol.query_cancel.get_first_ctaid.v4.b32.b128 {%0, %1, %2, _}, r;\n\t}"
      : "+r"(bx), "+r"(by), "+r"(bz), "=r"(ok) : "r"(result) : "memory");
  x = static_cast<int>(bx);
  y = static_cast<int>(by);
  return ok != 0;
}
// this CTA's shared memory -> global through a tensor map (bulk async-group completion); the writer must have
// made its shared-memory stores visible to the async proxy (fence_async_smem) before the issue
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* map, uint32_t src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
               :: "l"(map), "r"(src), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// the stores issued so far have READ their shared memory (it may be reused, the CTA may exit)
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(map)) : "memory");
}
// generic-proxy writes to shared memory -> visible to the async proxy (TMA / UMMA reads)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

#endif  // __CUDACC__

}  // namespace tc05
}  // namespace pa
