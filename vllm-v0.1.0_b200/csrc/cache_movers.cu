// KV-cache movers for sm_100a: reshape_and_cache, gather_cached_kv, copy_blocks,
// swap_blocks.  Pure data movement, bit-exact.  Reference behaviour:
// csrc/cache_kernels.cu:9-47 (swap), :52-138 (copy), :142-224 (reshape_and_cache),
// :229-386 (gather) of the reference tree.  All offsets are 64-bit.
#include "pa_common.cuh"

namespace pa {

// ---------------------------------------------------------------------------
// reshape_and_cache / gather_cached_kv
//
// One thread moves one 16-byte vector (x = 16/sizeof(T) consecutive head
// elements) of one token: in the key cache that vector is contiguous
// ([.., d/x, off, 0..x)), so it is a single 16-byte access; in the value cache
// its x elements sit block_size apart ([.., d, off]) and move one by one.
// ---------------------------------------------------------------------------
struct ScatterArgs {
  const void* key;      // gather: destination
  const void* value;
  void* key_cache;      // gather: source
  void* value_cache;
  const int32_t* slot_mapping;
  int64_t key_stride;   // elements
  int64_t value_stride;
  int32_t num_heads;
  int32_t head_size;
  int32_t block_size;
  int32_t vec_ok;       // key/value rows are 16-byte aligned
};

template <typename E, bool kGather>  // E = uint16_t or uint32_t
__global__ void __launch_bounds__(256) kv_scatter_gather_kernel(const ScatterArgs a) {
  constexpr int X = 16 / sizeof(E);
  const int token = blockIdx.x;
  const int slot = a.slot_mapping[token];
  if (slot < 0) return;  // padding token (not produced by the v0.1.0 scheduler)
  const int64_t blk = slot / a.block_size;
  const int off = slot % a.block_size;
  const int D = a.head_size, B = a.block_size, H = a.num_heads;
  const int nch = D / X;
  const int nvec = H * nch;

  E* kc = reinterpret_cast<E*>(a.key_cache);
  E* vc = reinterpret_cast<E*>(a.value_cache);
  E* krow = reinterpret_cast<E*>(const_cast<void*>(a.key)) + static_cast<int64_t>(token) * a.key_stride;
  E* vrow = reinterpret_cast<E*>(const_cast<void*>(a.value)) + static_cast<int64_t>(token) * a.value_stride;

  for (int v = blockIdx.y * blockDim.x + threadIdx.x; v < nvec; v += gridDim.y * blockDim.x) {
    const int h = v / nch;
    const int c = v - h * nch;
    // key:   [blk][h][c][off][0..X)
    E* kdst = kc + (((blk * H + h) * nch + c) * B + off) * X;
    // value: [blk][h][c*X + j][off]
    E* vdst = vc + ((blk * H + h) * D + static_cast<int64_t>(c) * X) * B + off;
    E* ksrc = krow + static_cast<int64_t>(v) * X;
    E* vsrc = vrow + static_cast<int64_t>(v) * X;
    if (!kGather) {
      union { uint4 q; E e[X]; } kk, vv;
      if (a.vec_ok) {
        kk.q = *reinterpret_cast<const uint4*>(ksrc);
        vv.q = *reinterpret_cast<const uint4*>(vsrc);
      } else {
#pragma unroll
        for (int j = 0; j < X; ++j) { kk.e[j] = ksrc[j]; vv.e[j] = vsrc[j]; }
      }
      *reinterpret_cast<uint4*>(kdst) = kk.q;
#pragma unroll
      for (int j = 0; j < X; ++j) vdst[static_cast<int64_t>(j) * B] = vv.e[j];
    } else {
      union { uint4 q; E e[X]; } kk, vv;
      kk.q = *reinterpret_cast<const uint4*>(kdst);
#pragma unroll
      for (int j = 0; j < X; ++j) vv.e[j] = vdst[static_cast<int64_t>(j) * B];
      if (a.vec_ok) {
        *reinterpret_cast<uint4*>(ksrc) = kk.q;
        *reinterpret_cast<uint4*>(vsrc) = vv.q;
      } else {
#pragma unroll
        for (int j = 0; j < X; ++j) { ksrc[j] = kk.e[j]; vsrc[j] = vv.e[j]; }
      }
    }
  }
}

template <bool kGather>
static int scatter_gather(const void* key, const void* value, void* key_cache, void* value_cache,
                          const int32_t* slot_mapping, int32_t num_tokens, int32_t num_heads,
                          int32_t head_size, int32_t block_size, int64_t key_stride,
                          int64_t value_stride, int32_t dtype, cudaStream_t stream) {
  const char* op = kGather ? "gather_cached_kv" : "reshape_and_cache";
  if (dtype != PA_F16 && dtype != PA_BF16 && dtype != PA_F32)
    return set_error(PA_ERR_UNSUPPORTED_DTYPE, "%s: unsupported dtype %d", op, dtype);
  if (num_tokens < 0 || num_heads <= 0 || head_size <= 0 || block_size <= 0)
    return set_error(PA_ERR_INVALID_ARGUMENT, "%s: bad sizes (tokens %d, heads %d, head_size %d, block_size %d)",
                     op, num_tokens, num_heads, head_size, block_size);
  if (num_tokens == 0) return PA_OK;
  if (!key || !value || !key_cache || !value_cache || !slot_mapping)
    return set_error(PA_ERR_INVALID_ARGUMENT, "%s: null pointer", op);
  const int es = dtype_size(dtype);
  const int x = 16 / es;
  if (head_size % x != 0)
    return set_error(PA_ERR_INVALID_ARGUMENT, "%s: head_size %d is not a multiple of x=%d", op, head_size, x);
  if ((reinterpret_cast<uintptr_t>(key_cache) | reinterpret_cast<uintptr_t>(value_cache)) & 15u)
    return set_error(PA_ERR_INVALID_ARGUMENT, "%s: caches must be 16-byte aligned", op);

  ScatterArgs a;
  a.key = key; a.value = value; a.key_cache = key_cache; a.value_cache = value_cache;
  a.slot_mapping = slot_mapping; a.key_stride = key_stride; a.value_stride = value_stride;
  a.num_heads = num_heads; a.head_size = head_size; a.block_size = block_size;
  a.vec_ok = ((reinterpret_cast<uintptr_t>(key) | reinterpret_cast<uintptr_t>(value)) % 16 == 0) &&
             ((key_stride * es) % 16 == 0) && ((value_stride * es) % 16 == 0);

  const int nvec = num_heads * (head_size / x);
  const int threads = nvec >= 256 ? 256 : ((nvec + 31) / 32) * 32;
  // few tokens (decode): spread one token over several CTAs so the grid still fills the GPU
  int ysplit = (nvec + threads - 1) / threads;
  if (static_cast<int64_t>(num_tokens) * ysplit > 148 * 32) ysplit = 1;
  const dim3 grid(num_tokens, ysplit);
  if (es == 2) kv_scatter_gather_kernel<uint16_t, kGather><<<grid, threads, 0, stream>>>(a);
  else kv_scatter_gather_kernel<uint32_t, kGather><<<grid, threads, 0, stream>>>(a);
  count_launch();
  return check_cuda(cudaGetLastError(), op);
}

// ---------------------------------------------------------------------------
// copy_blocks: pointer and pair tables ride in the kernel parameters
// ---------------------------------------------------------------------------
constexpr int kCopyMaxLayers = 64;
constexpr int kCopyMaxPairs = 160;
constexpr int kCopyThreads = 256;
constexpr int kCopyUnroll = 4;
constexpr int kCopySegBytes = kCopyThreads * 16 * kCopyUnroll;  // 16 KiB per CTA

struct CopyTable {
  char* kptr[kCopyMaxLayers];
  char* vptr[kCopyMaxLayers];
  int64_t src[kCopyMaxPairs];
  int64_t dst[kCopyMaxPairs];
  int64_t block_bytes;
};
static_assert(sizeof(CopyTable) <= 4000, "kernel parameter space");

// grid: (segments per block, pairs * 2 [key | value], layers)
__global__ void __launch_bounds__(kCopyThreads) copy_blocks_kernel(const __grid_constant__ CopyTable t) {
  const int pair = blockIdx.y >> 1;
  char* base = (blockIdx.y & 1) ? t.vptr[blockIdx.z] : t.kptr[blockIdx.z];
  const char* src = base + t.src[pair] * t.block_bytes;
  char* dst = base + t.dst[pair] * t.block_bytes;
  const int64_t seg = static_cast<int64_t>(blockIdx.x) * kCopySegBytes;
  uint4 v[kCopyUnroll];
  int64_t o[kCopyUnroll];
#pragma unroll
  for (int j = 0; j < kCopyUnroll; ++j) {
    o[j] = seg + (static_cast<int64_t>(j) * kCopyThreads + threadIdx.x) * 16;
    if (o[j] < t.block_bytes) v[j] = ldg_stream(src + o[j]);
  }
#pragma unroll
  for (int j = 0; j < kCopyUnroll; ++j) {
    if (o[j] < t.block_bytes) stg_stream(dst + o[j], v[j]);
  }
}

static int copy_blocks(void* const* key_cache_ptrs, void* const* value_cache_ptrs, int32_t num_layers,
                       const int64_t* pairs, int64_t num_pairs, int64_t block_bytes,
                       cudaStream_t stream) {
  if (num_layers < 0 || num_pairs < 0 || block_bytes < 0)
    return set_error(PA_ERR_INVALID_ARGUMENT, "copy_blocks: negative size");
  if (num_layers == 0 || num_pairs == 0 || block_bytes == 0) return PA_OK;  // reference returns early too (:88-90)
  if (!key_cache_ptrs || !value_cache_ptrs || !pairs)
    return set_error(PA_ERR_INVALID_ARGUMENT, "copy_blocks: null table");
  if (block_bytes % 16 != 0)
    return set_error(PA_ERR_INVALID_ARGUMENT, "copy_blocks: block_bytes %lld is not a multiple of 16",
                     static_cast<long long>(block_bytes));
  for (int l = 0; l < num_layers; ++l) {
    if (!key_cache_ptrs[l] || !value_cache_ptrs[l])
      return set_error(PA_ERR_INVALID_ARGUMENT, "copy_blocks: null cache pointer for layer %d", l);
    if ((reinterpret_cast<uintptr_t>(key_cache_ptrs[l]) | reinterpret_cast<uintptr_t>(value_cache_ptrs[l])) & 15u)
      return set_error(PA_ERR_INVALID_ARGUMENT, "copy_blocks: layer %d cache is not 16-byte aligned", l);
  }
  for (int64_t i = 0; i < 2 * num_pairs; ++i)
    if (pairs[i] < 0) return set_error(PA_ERR_INVALID_ARGUMENT, "copy_blocks: negative block number");

  const unsigned segs = static_cast<unsigned>((block_bytes + kCopySegBytes - 1) / kCopySegBytes);
  for (int l0 = 0; l0 < num_layers; l0 += kCopyMaxLayers) {
    const int nl = num_layers - l0 < kCopyMaxLayers ? num_layers - l0 : kCopyMaxLayers;
    for (int64_t p0 = 0; p0 < num_pairs; p0 += kCopyMaxPairs) {
      const int np = static_cast<int>(num_pairs - p0 < kCopyMaxPairs ? num_pairs - p0 : kCopyMaxPairs);
      CopyTable t;
      for (int l = 0; l < nl; ++l) {
        t.kptr[l] = static_cast<char*>(key_cache_ptrs[l0 + l]);
        t.vptr[l] = static_cast<char*>(value_cache_ptrs[l0 + l]);
      }
      for (int p = 0; p < np; ++p) {
        t.src[p] = pairs[2 * (p0 + p)];
        t.dst[p] = pairs[2 * (p0 + p) + 1];
      }
      t.block_bytes = block_bytes;
      const dim3 grid(segs, 2 * np, nl);
      copy_blocks_kernel<<<grid, kCopyThreads, 0, stream>>>(t);
      count_launch();
      const int rc = check_cuda(cudaGetLastError(), "copy_blocks");
      if (rc != PA_OK) return rc;
    }
  }
  return PA_OK;
}

// ---------------------------------------------------------------------------
// swap_blocks: DMA (copy engines stay off the SMs, so a swap on a side stream
// overlaps the model as in vllm/worker/cache_engine.py:101-118); consecutive
// block runs are merged so a swap of a whole sequence is a handful of copies
// instead of one per block (reference: one cudaMemcpyAsync per block, :35-46).
// ---------------------------------------------------------------------------
static int swap_blocks(const void* src, void* dst, int64_t block_bytes, const int64_t* src_blocks,
                       const int64_t* dst_blocks, int64_t num_blocks, int32_t kind,
                       cudaStream_t stream) {
  cudaMemcpyKind k;
  switch (kind) {
    case PA_COPY_D2D: k = cudaMemcpyDeviceToDevice; break;
    case PA_COPY_D2H: k = cudaMemcpyDeviceToHost; break;
    case PA_COPY_H2D: k = cudaMemcpyHostToDevice; break;
    default: return set_error(PA_ERR_INVALID_ARGUMENT, "swap_blocks: invalid device combination (kind %d)", kind);
  }
  if (num_blocks < 0 || block_bytes < 0) return set_error(PA_ERR_INVALID_ARGUMENT, "swap_blocks: negative size");
  if (num_blocks == 0 || block_bytes == 0) return PA_OK;
  if (!src || !dst || !src_blocks || !dst_blocks)
    return set_error(PA_ERR_INVALID_ARGUMENT, "swap_blocks: null pointer");
  for (int64_t i = 0; i < num_blocks; ++i)
    if (src_blocks[i] < 0 || dst_blocks[i] < 0)
      return set_error(PA_ERR_INVALID_ARGUMENT, "swap_blocks: negative block number");
  int64_t i = 0;
  while (i < num_blocks) {
    int64_t j = i + 1;
    while (j < num_blocks && src_blocks[j] == src_blocks[j - 1] + 1 && dst_blocks[j] == dst_blocks[j - 1] + 1) ++j;
    const char* s = static_cast<const char*>(src) + src_blocks[i] * block_bytes;
    char* d = static_cast<char*>(dst) + dst_blocks[i] * block_bytes;
    PA_CUDA_TRY(cudaMemcpyAsync(d, s, static_cast<size_t>((j - i) * block_bytes), k, stream));
    i = j;
  }
  return PA_OK;
}

}  // namespace pa

// ---------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------
extern "C" {

int pa_reshape_and_cache(const void* key, const void* value, void* key_cache, void* value_cache,
                         const int32_t* slot_mapping, int32_t num_tokens, int32_t num_heads,
                         int32_t head_size, int32_t block_size, int64_t key_stride,
                         int64_t value_stride, int32_t dtype, void* stream) {
  return pa::scatter_gather<false>(key, value, key_cache, value_cache, slot_mapping, num_tokens, num_heads,
                                   head_size, block_size, key_stride, value_stride, dtype,
                                   static_cast<cudaStream_t>(stream));
}

int pa_gather_cached_kv(void* key, void* value, const void* key_cache, const void* value_cache,
                        const int32_t* slot_mapping, int32_t num_tokens, int32_t num_heads,
                        int32_t head_size, int32_t block_size, int64_t key_stride,
                        int64_t value_stride, int32_t dtype, void* stream) {
  return pa::scatter_gather<true>(key, value, const_cast<void*>(key_cache), const_cast<void*>(value_cache),
                                  slot_mapping, num_tokens, num_heads, head_size, block_size, key_stride,
                                  value_stride, dtype, static_cast<cudaStream_t>(stream));
}

int pa_copy_blocks(void* const* key_cache_ptrs, void* const* value_cache_ptrs, int32_t num_layers,
                   const int64_t* pairs, int64_t num_pairs, int64_t block_bytes, void* stream) {
  return pa::copy_blocks(key_cache_ptrs, value_cache_ptrs, num_layers, pairs, num_pairs, block_bytes,
                         static_cast<cudaStream_t>(stream));
}

int pa_swap_blocks(const void* src, void* dst, int64_t block_bytes, const int64_t* src_blocks,
                   const int64_t* dst_blocks, int64_t num_blocks, int32_t kind, void* stream) {
  return pa::swap_blocks(src, dst, block_bytes, src_blocks, dst_blocks, num_blocks, kind,
                         static_cast<cudaStream_t>(stream));
}

}  // extern "C"
