// Shared device helpers for libpaged_b200 (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "paged_attention_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libpaged_b200 is written for sm_100a (B200) only"
#endif

namespace pa {

// ---------------------------------------------------------------------------
// host side: error state + launch accounting (defined in pa_api.cu)
// ---------------------------------------------------------------------------
int set_error(int code, const char* fmt, ...);
int check_cuda(cudaError_t e, const char* what);
void count_launch(uint64_t n = 1);

#define PA_CUDA_TRY(expr)                                      \
  do {                                                         \
    int _rc = ::pa::check_cuda((expr), #expr);                 \
    if (_rc != PA_OK) return _rc;                              \
  } while (0)

constexpr int kWarp = 32;

inline int dtype_size(int dtype) { return dtype == PA_F32 ? 4 : 2; }

// ---------------------------------------------------------------------------
// element-type tags.  16-bit types are carried as raw bit patterns; all
// arithmetic on them goes through the sm_100a mixed-precision FMA below.
// ---------------------------------------------------------------------------
struct F16 {
  static constexpr int kSize = 2;
  static constexpr int kId = PA_F16;
};
struct BF16 {
  static constexpr int kSize = 2;
  static constexpr int kId = PA_BF16;
};
struct F32 {
  static constexpr int kSize = 4;
  static constexpr int kId = PA_F32;
};

#ifdef __CUDACC__

// d = a.lo * b.lo + c with a, b the low halves of packed 16-bit pairs.
// sm_100a FHFMA: binary16/bfloat16 operands, fp32 accumulate, exact product,
// one rounding -- no separate conversion instruction.
template <typename T>
__device__ __forceinline__ float mixfma_lo(uint32_t a, uint32_t b, float c);
template <typename T>
__device__ __forceinline__ float mixfma_hi(uint32_t a, uint32_t b, float c);

template <>
__device__ __forceinline__ float mixfma_lo<F16>(uint32_t a, uint32_t b, float c) {
  float d;
  asm("{.reg .b16 al, ah, bl, bh;\n\t"
      "mov.b32 {al, ah}, %1; mov.b32 {bl, bh}, %2;\n\t"
      "fma.rn.f32.f16 %0, al, bl, %3;}"
      : "=f"(d) : "r"(a), "r"(b), "f"(c));
  return d;
}
template <>
__device__ __forceinline__ float mixfma_hi<F16>(uint32_t a, uint32_t b, float c) {
  float d;
  asm("{.reg .b16 al, ah, bl, bh;\n\t"
      "mov.b32 {al, ah}, %1; mov.b32 {bl, bh}, %2;\n\t"
      "fma.rn.f32.f16 %0, ah, bh, %3;}"
      : "=f"(d) : "r"(a), "r"(b), "f"(c));
  return d;
}
template <>
__device__ __forceinline__ float mixfma_lo<BF16>(uint32_t a, uint32_t b, float c) {
  float d;
  asm("{.reg .b16 al, ah, bl, bh;\n\t"
      "mov.b32 {al, ah}, %1; mov.b32 {bl, bh}, %2;\n\t"
      "fma.rn.f32.bf16 %0, al, bl, %3;}"
      : "=f"(d) : "r"(a), "r"(b), "f"(c));
  return d;
}
template <>
__device__ __forceinline__ float mixfma_hi<BF16>(uint32_t a, uint32_t b, float c) {
  float d;
  asm("{.reg .b16 al, ah, bl, bh;\n\t"
      "mov.b32 {al, ah}, %1; mov.b32 {bl, bh}, %2;\n\t"
      "fma.rn.f32.bf16 %0, ah, bh, %3;}"
      : "=f"(d) : "r"(a), "r"(b), "f"(c));
  return d;
}

// fp32 pair -> packed 16-bit pair, round to nearest even (lo in bits 0-15).
template <typename T>
__device__ __forceinline__ uint32_t pack2(float lo, float hi);
template <>
__device__ __forceinline__ uint32_t pack2<F16>(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
template <>
__device__ __forceinline__ uint32_t pack2<BF16>(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

// low / high 16-bit element of a packed pair -> fp32 (exact).
template <typename T>
__device__ __forceinline__ float unpack_lo(uint32_t v);
template <typename T>
__device__ __forceinline__ float unpack_hi(uint32_t v);
template <>
__device__ __forceinline__ float unpack_lo<F16>(uint32_t v) {
  float f;
  asm("{.reg .b16 l, h; mov.b32 {l, h}, %1; cvt.f32.f16 %0, l;}" : "=f"(f) : "r"(v));
  return f;
}
template <>
__device__ __forceinline__ float unpack_hi<F16>(uint32_t v) {
  float f;
  asm("{.reg .b16 l, h; mov.b32 {l, h}, %1; cvt.f32.f16 %0, h;}" : "=f"(f) : "r"(v));
  return f;
}
template <>
__device__ __forceinline__ float unpack_lo<BF16>(uint32_t v) {
  return __uint_as_float(v << 16);
}
template <>
__device__ __forceinline__ float unpack_hi<BF16>(uint32_t v) {
  return __uint_as_float(v & 0xffff0000u);
}

// ---------------------------------------------------------------------------
// warp-level matrix pieces for the sibling-group path (16-bit types only):
// D[16x8] += A[16x16] * B[16x8], fp32 accumulate (legacy tensor-core path, HMMA: the tile a group of
// <= 8 sibling queries forms is 8 wide, far below the 64-row minimum of tcgen05)
// ---------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1);
template <>
__device__ __forceinline__ void mma16816<F16>(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <>
__device__ __forceinline__ void mma16816<BF16>(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <>
__device__ __forceinline__ void mma16816<F32>(float (&)[4], const uint32_t (&)[4], uint32_t, uint32_t) {}
// four 8x8 b16 matrices from shared memory; lane l supplies the address of row l%8 of matrix l/8
__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr) : "memory");
}
// transpose of an 8x8 b16 matrix held one b16 pair per lane (row lane/4, columns 2*(lane%4), +1)
__device__ __forceinline__ uint32_t movmatrix_trans(uint32_t v) {
  uint32_t r;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(r) : "r"(v));
  return r;
}

// ---------------------------------------------------------------------------
// warp collectives
// ---------------------------------------------------------------------------
__device__ __forceinline__ float warp_max(float v) {
  float m;  // sm_100a: single-instruction fp32 warp max (CREDUX.MAX.F32)
  asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(m) : "f"(v));
  return m;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---------------------------------------------------------------------------
// streaming global access: 16-byte, no L1 allocation (every KV byte is read
// exactly once per call)
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint4 ldg_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void stg_stream(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1, %2, %3, %4};"
               :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// ---------------------------------------------------------------------------
// mbarrier + bulk async copy (TMA engine, no tensor map: pages are contiguous)
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_addr(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
               :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}"
      :: "r"(bar), "r"(parity) : "memory");
}
// global -> this CTA's shared memory, completion counted on `bar`.
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes,
                                         uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      :: "r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// L2 residency: every KV byte is read exactly once per call (evict_first), the block tables, context lengths
// and q rows are read by every call and sit on the critical path of each CTA's prologue (evict_last).
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void bulk_g2s_hint(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar,
                                              uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
      :: "r"(dst), "l"(src), "r"(bytes), "r"(bar), "l"(policy) : "memory");
}
__device__ __forceinline__ int32_t ldg_i32_hint(const int32_t* p, uint64_t policy) {
  int32_t v;
  asm volatile("ld.global.nc.L2::cache_hint.b32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(policy));
  return v;
}
__device__ __forceinline__ uint4 ldg_v4_hint(const void* p, uint64_t policy) {
  uint4 r;
  asm volatile("ld.global.nc.L2::cache_hint.v4.u32 {%0, %1, %2, %3}, [%4], %5;"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p), "l"(policy));
  return r;
}

// Programmatic dependent launch.  pdl_wait(): everything the preceding kernel on the stream wrote is visible
// after it (a no-op for a launch without the attribute).  pdl_launch_dependents(): the next kernel on the
// stream, if it was launched with the attribute, may start its CTAs as resources free up; it must pdl_wait()
// before it reads anything this kernel writes.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

#endif  // __CUDACC__

}  // namespace pa
