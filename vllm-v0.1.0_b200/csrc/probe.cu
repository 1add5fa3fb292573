// Read-bandwidth probe: streams a buffer through shared memory with the same TMA bulk copies the
// decode kernel uses (4 warps per CTA, a 2-stage ring of 8 KiB per warp) and does no arithmetic on
// it.  It answers "what can this access pattern read from HBM at most?", the practical ceiling the
// decode kernel's achieved GB/s is compared with (bench.py reports both).  Not an operator.
#include "pa_common.cuh"

namespace pa {

constexpr int kProbeWarps = 4;
constexpr int kProbeStages = 2;
constexpr int kProbeChunk = 8192;

__global__ void __launch_bounds__(kProbeWarps * kWarp) stream_read_probe_kernel(const char* src, int64_t chunks,
                                                                                 uint32_t* sink) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) uint64_t bars_mem[kProbeWarps * kProbeStages];
  const int warp = threadIdx.x / kWarp, lane = threadIdx.x % kWarp;
  const uint32_t bars = smem_addr(bars_mem) + warp * kProbeStages * 8;
  const uint32_t ring = smem_addr(smem) + warp * kProbeStages * kProbeChunk;
  if (lane == 0) {
    for (int s = 0; s < kProbeStages; ++s) mbar_init(bars + s * 8, 1);
    mbar_init_fence();
  }
  __syncwarp();
  // chunk index of this warp's n-th chunk: grid-strided so neighbouring warps read far apart, like pages
  const int64_t gw = static_cast<int64_t>(blockIdx.x) * kProbeWarps + warp;
  const int64_t nw = static_cast<int64_t>(gridDim.x) * kProbeWarps;
  const int64_t n_my = chunks > gw ? (chunks - gw + nw - 1) / nw : 0;
  auto issue = [&](int64_t n, int stage) {
    const uint32_t bar = bars + stage * 8;
    mbar_expect_tx(bar, kProbeChunk);
    bulk_g2s(ring + stage * kProbeChunk, src + (gw + n * nw) * kProbeChunk, kProbeChunk, bar);
  };
  if (lane == 0)
    for (int n = 0; n < kProbeStages && n < n_my; ++n) issue(n, static_cast<int>(n));
  int stage = 0;
  uint32_t parity = 0, acc = 0;
  for (int64_t n = 0; n < n_my; ++n) {
    mbar_wait(bars + stage * 8, parity);
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(ring + stage * kProbeChunk + lane * 4));
    acc ^= v;
    __syncwarp();
    if (lane == 0 && n + kProbeStages < n_my) issue(n + kProbeStages, stage);
    if (++stage == kProbeStages) { stage = 0; parity ^= 1u; }
  }
  if (acc == 0x9e3779b9u) *sink = acc;   // keeps the loads alive
}

}  // namespace pa

extern "C" int pa_stream_read_probe(const void* src, int64_t bytes, int32_t ctas_per_sm, void* sink, void* stream) {
  if (!src || !sink || bytes < pa::kProbeChunk) return pa::set_error(PA_ERR_INVALID_ARGUMENT, "probe: bad arguments");
  if (reinterpret_cast<uintptr_t>(src) & 15u) return pa::set_error(PA_ERR_INVALID_ARGUMENT, "probe: unaligned buffer");
  int dev = 0, sms = 148;
  PA_CUDA_TRY(cudaGetDevice(&dev));
  PA_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int smem = pa::kProbeWarps * pa::kProbeStages * pa::kProbeChunk;
  PA_CUDA_TRY(cudaFuncSetAttribute(pa::stream_read_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  if (ctas_per_sm <= 0) ctas_per_sm = 3;
  pa::stream_read_probe_kernel<<<sms * ctas_per_sm, pa::kProbeWarps * pa::kWarp, smem, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const char*>(src), bytes / pa::kProbeChunk, static_cast<uint32_t*>(sink));
  pa::count_launch();
  return pa::check_cuda(cudaGetLastError(), "stream_read_probe");
}
