// Microbenchmark of the prompt kernel's exponential section: one warp per SM sub-partition (4 warps per CTA, one
// CTA per SM) runs 128 ex2 per thread with the surrounding packed-fp32 work in several orders, clocked with %clock.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ubench_exp tools/ubench_exp.cu && /tmp/ubench_exp
// Result on B200 is recorded in profiles/r2_prefill_timeline.txt.
#include <cstdio>
#include <cuda_runtime.h>
#include <cstdint>

__device__ __forceinline__ float ex2v(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float ex2n(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ void ffma2(float& d0, float& d1, float a0, float a1, float b, float c) {
  asm("{.reg .b64 ra, rb, rc, rd;\n\tmov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %4};\n\tmov.b64 rc, {%5, %5};\n\t"
      "fma.rn.f32x2 rd, ra, rb, rc;\n\tmov.b64 {%0, %1}, rd;}" : "=f"(d0), "=f"(d1) : "f"(a0), "f"(a1), "f"(b), "f"(c));
}
__device__ __forceinline__ void fadd2(float& a0, float& a1, float b0, float b1) {
  asm("{.reg .b64 ra, rb;\n\tmov.b64 ra, {%0, %1};\n\tmov.b64 rb, {%2, %3};\n\tadd.rn.f32x2 ra, ra, rb;\n\t"
      "mov.b64 {%0, %1}, ra;}" : "+f"(a0), "+f"(a1) : "f"(b0), "f"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  uint32_t r; asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo)); return r;
}

// integer-pipe bf16 packs: round half up = (bits + 0x8000) >> 16, both halves picked by one PRMT
__device__ __forceinline__ uint32_t pack_bf16_int(float lo, float hi) {
  uint32_t a = __float_as_uint(lo) + 0x8000u, b = __float_as_uint(hi) + 0x8000u, r;
  asm("prmt.b32 %0, %1, %2, 0x7632;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ uint32_t pack_bf16_int64(float lo, float hi) {   // one 64-bit add: no carry can cross (p < 2^9)
  uint64_t v; uint32_t a, b, r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(lo), "f"(hi));
  v += 0x0000800000008000ull;
  asm("mov.b64 {%0, %1}, %2;" : "=r"(a), "=r"(b) : "l"(v));
  asm("prmt.b32 %0, %1, %2, 0x7632;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ uint32_t pack_bf16_trunc(float lo, float hi) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, 0x7632;" : "=r"(r) : "r"(__float_as_uint(lo)), "r"(__float_as_uint(hi)));
  return r;
}

// packed 16-bit exponentials: one MUFU lane-op gives two results, already in the P operand's format
__device__ __forceinline__ uint32_t ex2_bf16x2(uint32_t x) { uint32_t y; asm volatile("ex2.approx.ftz.bf16x2 %0, %1;" : "=r"(y) : "r"(x)); return y; }
__device__ __forceinline__ uint32_t ex2_f16x2(uint32_t x) { uint32_t y; asm volatile("ex2.approx.f16x2 %0, %1;" : "=r"(y) : "r"(x)); return y; }
__device__ __forceinline__ uint32_t pack_f16(float lo, float hi) {
  uint32_t r; asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo)); return r;
}
__device__ __forceinline__ float sum_bf16_lo(uint32_t a, float c) {   // c + lo(a) * 1.0, exact (FHFMA)
  float d; asm("{.reg .b16 al, ah, o;\n\tmov.b32 {al, ah}, %1; mov.b16 o, 0x3f80;\n\tfma.rn.f32.bf16 %0, al, o, %2;}" : "=f"(d) : "r"(a), "f"(c)); return d;
}
__device__ __forceinline__ float sum_bf16_hi(uint32_t a, float c) {
  float d; asm("{.reg .b16 al, ah, o;\n\tmov.b32 {al, ah}, %1; mov.b16 o, 0x3f80;\n\tfma.rn.f32.bf16 %0, ah, o, %2;}" : "=f"(d) : "r"(a), "f"(c)); return d;
}
__device__ __forceinline__ float sum_f16_lo(uint32_t a, float c) {
  float d; asm("{.reg .b16 al, ah, o;\n\tmov.b32 {al, ah}, %1; mov.b16 o, 0x3c00;\n\tfma.rn.f32.f16 %0, al, o, %2;}" : "=f"(d) : "r"(a), "f"(c)); return d;
}
__device__ __forceinline__ float sum_f16_hi(uint32_t a, float c) {
  float d; asm("{.reg .b16 al, ah, o;\n\tmov.b32 {al, ah}, %1; mov.b16 o, 0x3c00;\n\tfma.rn.f32.f16 %0, ah, o, %2;}" : "=f"(d) : "r"(a), "f"(c)); return d;
}
// mode 8 / 9: ffma2 -> pack the two arguments -> ONE packed ex2 (bf16x2 / f16x2) = P as stored -> two mixed-precision adds
// mode 10: as 8, sums by shift / mask + fadd2
// mode 11: 64 x ex2.bf16x2 alone
// mode 5 / 6 / 7: the kernel's section with the pack on the integer pipe (2 adds + prmt / 64-bit add + prmt / prmt only)
// mode 0: 128 ex2 only (sum by plain adds at the end)
// mode 1: the kernel's section: ffma2 -> 2 x ex2 (volatile) -> fadd2 + pack, per pair
// mode 2: as 1 with non-volatile ex2 (ptxas free to order)
// mode 3: ffma2 + ex2 + pack, sums with scalar FADD instead of FADD2
// mode 4: as 1 but sums taken from the PACKED values later (no fadd in the section): ex2 + pack only
template <int MODE>
__global__ void __launch_bounds__(128, 1) k(const float* __restrict__ in, float* __restrict__ out, uint32_t* clk, float sc, float nm) {
  float s[128];
#pragma unroll
  for (int i = 0; i < 128; ++i) s[i] = in[(i * 128 + threadIdx.x)];
  float chk = 0.f;                                  // the loads have landed before the clock is read
#pragma unroll
  for (int i = 0; i < 128; ++i) chk += s[i];
  if (chk == 12345.f) out[0] = chk;
  __syncthreads();
  uint32_t t0, t1;
  asm volatile("mov.u32 %0, %%clock;" : "=r"(t0) : "f"(chk));
  float sum[4] = {0.f, 0.f, 0.f, 0.f};
  uint32_t pk[64];
  if (MODE == 0) {
#pragma unroll
    for (int i = 0; i < 128; ++i) s[i] = ex2v(s[i]);
  } else if (MODE == 11) {
#pragma unroll
    for (int i = 0; i < 64; ++i) pk[i] = ex2_bf16x2(__float_as_uint(s[i]));
  } else if (MODE >= 8) {
#pragma unroll
    for (int i = 0; i < 128; i += 2) {
      float a0, a1;
      ffma2(a0, a1, s[i], s[i + 1], sc, nm);
      const uint32_t pp = MODE == 9 ? ex2_f16x2(pack_f16(a0, a1)) : ex2_bf16x2(pack_bf16(a0, a1));
      if (MODE == 8) { sum[(i >> 1) & 1] = sum_bf16_lo(pp, sum[(i >> 1) & 1]); sum[2 + ((i >> 1) & 1)] = sum_bf16_hi(pp, sum[2 + ((i >> 1) & 1)]); }
      if (MODE == 9) { sum[(i >> 1) & 1] = sum_f16_lo(pp, sum[(i >> 1) & 1]); sum[2 + ((i >> 1) & 1)] = sum_f16_hi(pp, sum[2 + ((i >> 1) & 1)]); }
      if (MODE == 10) fadd2(sum[(i >> 1) & 1 ? 2 : 0], sum[(i >> 1) & 1 ? 3 : 1], __uint_as_float(pp << 16), __uint_as_float(pp & 0xffff0000u));
      pk[i >> 1] = pp;
    }
  } else {
#pragma unroll
    for (int i = 0; i < 128; i += 2) {
      float a0, a1;
      ffma2(a0, a1, s[i], s[i + 1], sc, nm);
      const float p0 = MODE == 2 ? ex2n(a0) : ex2v(a0), p1 = MODE == 2 ? ex2n(a1) : ex2v(a1);
      if (MODE == 1 || MODE == 2 || MODE >= 5) fadd2(sum[(i >> 1) & 1 ? 2 : 0], sum[(i >> 1) & 1 ? 3 : 1], p0, p1);
      if (MODE == 3) { sum[(i >> 1) & 3] += p0; sum[((i >> 1) + 1) & 3] += p1; }
      pk[i >> 1] = MODE == 5 ? pack_bf16_int(p0, p1) : MODE == 6 ? pack_bf16_int64(p0, p1)
                   : MODE == 7 ? pack_bf16_trunc(p0, p1) : pack_bf16(p0, p1);
    }
  }
  asm volatile("mov.u32 %0, %%clock;" : "=r"(t1));
  float acc = sum[0] + sum[1] + sum[2] + sum[3];
  if (MODE == 0) {
#pragma unroll
    for (int i = 0; i < 128; ++i) acc += s[i];
  } else {
    if (MODE == 11) acc += s[127];
    uint32_t x = 0;
#pragma unroll
    for (int i = 0; i < 64; ++i) x ^= pk[i];
    acc += __uint_as_float(x);
  }
  out[blockIdx.x * 128 + threadIdx.x] = acc;
  if (threadIdx.x % 32 == 0) clk[blockIdx.x * 4 + threadIdx.x / 32] = t1 - t0;
}

template <int MODE>
static void run(const char* name, const float* in, float* out, uint32_t* clk) {
  uint32_t h[148 * 4];
  for (int rep = 0; rep < 3; ++rep) k<MODE><<<148, 128>>>(in, out, clk, 0.125f, -3.f);
  cudaDeviceSynchronize();
  cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
  double avg = 0; uint32_t mn = ~0u, mx = 0;
  for (int i = 0; i < 148 * 4; ++i) { avg += h[i]; mn = h[i] < mn ? h[i] : mn; mx = h[i] > mx ? h[i] : mx; }
  printf("%-58s clk per 128 ex2: avg %.0f min %u max %u  (%s)\n", name, avg / (148 * 4), mn, mx, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  float *in, *out; uint32_t* clk;
  cudaMalloc(&in, 128 * 128 * 4); cudaMalloc(&out, 148 * 128 * 4); cudaMalloc(&clk, 148 * 4 * 4);
  cudaMemset(in, 0, 128 * 128 * 4);
  run<0>("0: 128 x ex2 alone", in, out, clk);
  run<1>("1: ffma2, 2 ex2 (pinned), fadd2, pack  [the kernel]", in, out, clk);
  run<2>("2: the same, ex2 free to move", in, out, clk);
  run<3>("3: scalar FADD sums", in, out, clk);
  run<4>("4: no sums in the section (ffma2, ex2, pack)", in, out, clk);
  run<5>("5: kernel section, pack = 2 x IADD + PRMT", in, out, clk);
  run<6>("6: kernel section, pack = 64-bit add + PRMT", in, out, clk);
  run<7>("7: kernel section, pack = PRMT (truncation)", in, out, clk);
  run<11>("11: 64 x ex2.bf16x2 alone (= 128 exponentials)", in, out, clk);
  run<8>("8: ffma2, pack args, ex2.bf16x2, 2 x fma.f32.bf16 sums", in, out, clk);
  run<9>("9: ffma2, pack args, ex2.f16x2, 2 x fma.f32.f16 sums", in, out, clk);
  run<10>("10: as 8, sums by shl / and + fadd2", in, out, clk);
  return 0;
}
