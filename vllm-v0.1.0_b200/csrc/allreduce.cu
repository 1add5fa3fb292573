// In-place all-reduce (sum) of the row-parallel output that follows attention, over NVLink peer memory.
//
// Replaces: GraphAllReduce.launch (vllm/model_executor/parallel_utils/parallel_state.py:506-556), the
// NCCL all-reduce the reference replays from a CUDA graph after every RowParallelLinear
// (tensor_parallel/layers.py:419-426).  At decode sizes (256 tokens x 5120 x 2 B = 2.6 MB) the collective is
// latency-bound, so it is ONE kernel per rank working directly on the peers' buffers:
//
//   barrier (flags in every rank's memory, system-scope CAS) -- every rank's addend is in place
//   rank r owns slice r of the vector:
//     multimem path:  multimem.ld_reduce (the NVSwitch adds all ranks' copies, fp32 accumulate) then
//                     multimem.st (the switch broadcasts the sum to every rank's buffer)
//     peer path:      16-byte loads of slice r from every peer, fp32 sum in rank order, 16-byte stores of the
//                     sum into slice r of every peer's buffer
//   barrier -- every slice of this rank's buffer has been written
//
// Every rank computes each element once and all ranks receive the same bits.  Buffers and flag words live in
// symmetric memory the caller maps (torch.distributed._symmetric_memory in the Python mirror); the C ABI only
// sees raw pointers.  `emulate` runs all ranks of a world inside one launch on one device (blockIdx.y = rank):
// that is how the peer path is tested on a single GPU.
#include "pa_common.cuh"

namespace pa {

constexpr int kArMaxWorld = 8;
constexpr int kArMaxBlocks = 128;

struct ArParams {
  void* bufs[kArMaxWorld];        // every rank's buffer (same offset applied already), peer-mapped
  uint32_t* flags[kArMaxWorld];   // every rank's flag words: [kArMaxBlocks][world], zero when idle
  void* mc;                       // multicast address of the buffer (multimem path) or null
  int64_t nvec;                   // 16-byte vectors to reduce
  int32_t rank;
  int32_t world;
};

__device__ __forceinline__ uint32_t cas_release_sys(uint32_t* addr, uint32_t expect, uint32_t val) {
  uint32_t old;
  asm volatile("atom.global.release.sys.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(addr), "r"(expect), "r"(val) : "memory");
  return old;
}
__device__ __forceinline__ uint32_t cas_acquire_sys(uint32_t* addr, uint32_t expect, uint32_t val) {
  uint32_t old;
  asm volatile("atom.global.acquire.sys.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(addr), "r"(expect), "r"(val) : "memory");
  return old;
}

// Same-index blocks of all ranks meet.  Thread t < world raises this rank's flag in rank t's memory
// (0 -> 1, spinning while the previous use has not been consumed) and consumes rank t's flag here (1 -> 0).
__device__ __forceinline__ void ar_barrier(const ArParams& p, int rank) {
  __syncthreads();
  if (threadIdx.x < p.world) {
    const int peer = threadIdx.x;
    uint32_t* theirs = p.flags[peer] + blockIdx.x * p.world + rank;
    while (cas_release_sys(theirs, 0u, 1u) != 0u) {}
    uint32_t* mine = p.flags[rank] + blockIdx.x * p.world + peer;
    while (cas_acquire_sys(mine, 1u, 0u) != 1u) {}
  }
  __syncthreads();
}

template <typename T>
__device__ __forceinline__ void accumulate(float (&acc)[8], const uint4& v) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
  if constexpr (T::kSize == 2) {
#pragma unroll
    for (int i = 0; i < 4; ++i) { acc[2 * i] += unpack_lo<T>(w[i]); acc[2 * i + 1] += unpack_hi<T>(w[i]); }
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[i] += __uint_as_float(w[i]);
  }
}
template <typename T>
__device__ __forceinline__ uint4 finish(const float (&acc)[8]) {
  if constexpr (T::kSize == 2)
    return make_uint4(pack2<T>(acc[0], acc[1]), pack2<T>(acc[2], acc[3]), pack2<T>(acc[4], acc[5]), pack2<T>(acc[6], acc[7]));
  else
    return make_uint4(__float_as_uint(acc[0]), __float_as_uint(acc[1]), __float_as_uint(acc[2]), __float_as_uint(acc[3]));
}

template <typename T> __device__ __forceinline__ uint4 multimem_ld_reduce(const void* mc);
template <> __device__ __forceinline__ uint4 multimem_ld_reduce<F16>(const void* mc) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(mc) : "memory");
  return r;
}
template <> __device__ __forceinline__ uint4 multimem_ld_reduce<BF16>(const void* mc) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(mc) : "memory");
  return r;
}
template <> __device__ __forceinline__ uint4 multimem_ld_reduce<F32>(const void* mc) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(mc) : "memory");
  return r;
}
__device__ __forceinline__ void multimem_st(void* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
               :: "l"(mc), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld_peer(const void* p) {
  uint4 r;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}

template <typename T, bool kMultimem>
__global__ void __launch_bounds__(512)
allreduce_kernel(const ArParams p) {
  const int rank = gridDim.y > 1 ? static_cast<int>(blockIdx.y) : p.rank;   // gridDim.y > 1: emulated world
  ar_barrier(p, rank);
  const int64_t per = (p.nvec + p.world - 1) / p.world;
  const int64_t begin = per * rank;
  const int64_t end = begin + per < p.nvec ? begin + per : p.nvec;
  const int64_t step = static_cast<int64_t>(gridDim.x) * blockDim.x;
  constexpr int kU = 4;   // independent 16-byte vectors in flight per thread
  for (int64_t v0 = begin + static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; v0 < end; v0 += kU * step) {
    if constexpr (kMultimem) {
      uint4 val[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u)
        if (v0 + u * step < end) val[u] = multimem_ld_reduce<T>(static_cast<char*>(p.mc) + (v0 + u * step) * 16);
#pragma unroll
      for (int u = 0; u < kU; ++u)
        if (v0 + u * step < end) multimem_st(static_cast<char*>(p.mc) + (v0 + u * step) * 16, val[u]);
    } else {
#pragma unroll
      for (int u0 = 0; u0 < kU; u0 += 2) {          // two vectors x world loads in flight
        uint4 in[2][kArMaxWorld];
#pragma unroll
        for (int u = 0; u < 2; ++u)
#pragma unroll
          for (int r = 0; r < kArMaxWorld; ++r)
            if (r < p.world && v0 + (u0 + u) * step < end)
              in[u][r] = ld_peer(static_cast<const char*>(p.bufs[r]) + (v0 + (u0 + u) * step) * 16);
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int64_t v = v0 + (u0 + u) * step;
          if (v < end) {
            float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int r = 0; r < kArMaxWorld; ++r)
              if (r < p.world) accumulate<T>(acc, in[u][r]);
            const uint4 out = finish<T>(acc);
#pragma unroll
            for (int r = 0; r < kArMaxWorld; ++r)
              if (r < p.world) *reinterpret_cast<uint4*>(static_cast<char*>(p.bufs[r]) + v * 16) = out;
          }
        }
      }
    }
  }
  ar_barrier(p, rank);
}

template <typename T>
static int launch(const ArParams& p, bool multimem, int blocks, int emulate_world, cudaStream_t stream) {
  const dim3 grid(blocks, emulate_world > 1 ? emulate_world : 1);
  if (multimem) allreduce_kernel<T, true><<<grid, 512, 0, stream>>>(p);
  else allreduce_kernel<T, false><<<grid, 512, 0, stream>>>(p);
  count_launch();
  return check_cuda(cudaGetLastError(), "allreduce_kernel launch");
}

static int allreduce(void* const* bufs, void* const* flags, void* multicast, int64_t offset_bytes, int64_t numel,
                     int32_t rank, int32_t world, int32_t dtype, int32_t emulate, cudaStream_t stream) {
  if (dtype != PA_F16 && dtype != PA_BF16 && dtype != PA_F32)
    return set_error(PA_ERR_UNSUPPORTED_DTYPE, "unsupported data type: %d", dtype);
  if (world < 1 || world > kArMaxWorld || rank < 0 || rank >= world)
    return set_error(PA_ERR_INVALID_ARGUMENT, "bad rank %d / world %d (max %d)", rank, world, kArMaxWorld);
  if (numel < 0 || !bufs || !flags) return set_error(PA_ERR_INVALID_ARGUMENT, "null pointer or negative size");
  if (world == 1 || numel == 0) return PA_OK;
  const int64_t bytes = numel * dtype_size(dtype);
  if (bytes % 16 || offset_bytes % 16)
    return set_error(PA_ERR_INVALID_ARGUMENT, "all-reduce size and offset must be multiples of 16 bytes");
  if (emulate && multicast) return set_error(PA_ERR_INVALID_ARGUMENT, "the emulated world has no multicast path");
  ArParams p;
  for (int r = 0; r < kArMaxWorld; ++r) { p.bufs[r] = nullptr; p.flags[r] = nullptr; }
  for (int r = 0; r < world; ++r) {
    if (!bufs[r] || !flags[r]) return set_error(PA_ERR_INVALID_ARGUMENT, "null peer pointer for rank %d", r);
    if (reinterpret_cast<uintptr_t>(bufs[r]) % 16) return set_error(PA_ERR_INVALID_ARGUMENT, "peer buffer is not 16-byte aligned");
    p.bufs[r] = static_cast<char*>(bufs[r]) + offset_bytes;
    p.flags[r] = static_cast<uint32_t*>(flags[r]);
  }
  p.mc = multicast ? static_cast<char*>(multicast) + offset_bytes : nullptr;
  p.nvec = bytes / 16;
  p.rank = rank;
  p.world = world;
  // one 16-byte vector per thread and pass; every block must be resident at once (they meet at the barriers)
  const int64_t per = (p.nvec + world - 1) / world;
  int blocks = static_cast<int>((per + 2047) / 2048);   // 4 vectors per thread and pass
  const int cap = emulate ? 16 : kArMaxBlocks;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  const int ew = emulate ? world : 1;
  switch (dtype) {
    case PA_F16:  return launch<F16>(p, multicast != nullptr, blocks, ew, stream);
    case PA_BF16: return launch<BF16>(p, multicast != nullptr, blocks, ew, stream);
    default:      return launch<F32>(p, multicast != nullptr, blocks, ew, stream);
  }
}

}  // namespace pa

extern "C" {

int pa_allreduce_flag_bytes(void) { return pa::kArMaxBlocks * pa::kArMaxWorld * static_cast<int>(sizeof(uint32_t)); }

int pa_allreduce(void* const* peer_buffers, void* const* peer_flags, void* multicast_ptr, int64_t offset_bytes,
                 int64_t numel, int32_t rank, int32_t world, int32_t dtype, void* stream) {
  return pa::allreduce(peer_buffers, peer_flags, multicast_ptr, offset_bytes, numel, rank, world, dtype, 0,
                       static_cast<cudaStream_t>(stream));
}

int pa_allreduce_emulated(void* const* buffers, void* const* flags, int64_t offset_bytes, int64_t numel,
                          int32_t world, int32_t dtype, void* stream) {
  return pa::allreduce(buffers, flags, nullptr, offset_bytes, numel, 0, world, dtype, 1,
                       static_cast<cudaStream_t>(stream));
}

}  // extern "C"
