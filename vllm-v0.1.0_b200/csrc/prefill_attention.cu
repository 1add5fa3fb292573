// Prompt (prefill) attention for sm_100a: causal self-attention over a batch of variable-length
// prompts packed along the token dimension, no KV cache involved.
//
// Replaces: PagedAttention.multi_query_kv_attention (vllm/model_executor/layers/attention.py:50-70),
// i.e. xformers.ops.memory_efficient_attention_forward with
// BlockDiagonalCausalMask.from_seqlens(prompt_lens) (vllm/model_executor/input_metadata.py:33);
// parity reference: ref_multi_query_kv_attention (tests/kernels/test_attention.py:69-101).
//
// This half of PagedAttention.forward is compute-bound, so it runs on the 5th-generation tensor
// cores:
//   * one CTA = 128 query rows of one (prompt, head); K/V tiles of 128 tokens stream through a
//     two-stage shared-memory ring filled by TMA (3-D tensor maps over the strided [token][head][d]
//     views of the fused qkv tensor, 128-byte swizzle, out-of-bounds rows and head-size padding
//     zero-filled by the TMA unit);
//   * S = Q.K^T: tcgen05.mma (M 128, N 128, K 16 per instruction), both operands K-major from shared
//     memory, fp32 accumulator in tensor memory, double-buffered so the tensor core computes tile j+1
//     while the CTA's threads run the softmax of tile j;
//   * softmax: thread r owns row r (TMEM lane r): tcgen05.ld of its 128 logits, running max / sum
//     in the exp2 domain, probabilities rounded to T and written back to TMEM (two per 32-bit column);
//   * O += P.V: tcgen05.mma with P as the A operand straight from TMEM and the V tile as an MN-major
//     B operand, O resident in TMEM for the whole CTA; it is rescaled in place only when some row's
//     maximum moved;
//   * only the diagonal tile applies the causal / length mask, tiles above it are never visited.
// fp32 inputs take an exact SIMT path (fp32 prefill is a test-matrix case, not a serving case).
#include <algorithm>
#include <cstring>
#include <mutex>
#include <vector>

#include <math_constants.h>

#include "prefill_tc.cuh"

namespace pa {

// ---------------------------------------------------------------------------
// fp32: exact SIMT path.  CTA = 64 query rows x 2 threads (each owns half of the head dimension);
// K/V tiles of 32 tokens staged in shared memory; online softmax per row.
// ---------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(128)
prefill_f32_kernel(const float* __restrict__ q, const float* __restrict__ k, const float* __restrict__ v,
                   float* __restrict__ out, int64_t q_stride, int64_t k_stride, int64_t v_stride, int64_t out_stride,
                   const __grid_constant__ PrefillTiles tiles, float scale_log2) {
  constexpr int HD = D / 2;          // elements per thread
  constexpr int KT = 32;             // tokens per staged tile
  __shared__ float sk[KT][D];
  __shared__ float sv[KT][D];
  const int head = blockIdx.x;
  const int qtok0 = tiles.qtok0[blockIdx.y], qlen = tiles.qlen[blockIdx.y], q0 = tiles.q0[blockIdx.y];
  const int ktok0 = tiles.ktok0[blockIdx.y], klen = tiles.klen[blockIdx.y];
  const int koff = klen - qlen;                     // cached keys in front of the first query row
  const int row = q0 + (threadIdx.x >> 1);          // query row inside the sequence
  const int half = threadIdx.x & 1;
  const bool row_ok = row < qlen;
  float qr[HD], acc[HD];
  {
    const float* qp = q + (static_cast<int64_t>(qtok0) + (row_ok ? row : q0)) * q_stride + head * D + half * HD;
#pragma unroll
    for (int i = 0; i < HD; ++i) { qr[i] = qp[i]; acc[i] = 0.f; }
  }
  float m = -CUDART_INF_F, l = 0.f;
  const int kv_end = min(q0 + 64 + koff, klen);
  for (int t0 = 0; t0 < kv_end; t0 += KT) {
    __syncthreads();
    for (int i = threadIdx.x; i < KT * D; i += blockDim.x) {
      const int t = i / D, d = i - t * D;
      const bool ok = t0 + t < kv_end;
      sk[t][d] = ok ? k[(static_cast<int64_t>(ktok0) + t0 + t) * k_stride + head * D + d] : 0.f;
      sv[t][d] = ok ? v[(static_cast<int64_t>(ktok0) + t0 + t) * v_stride + head * D + d] : 0.f;
    }
    __syncthreads();
    const int tmax = min(KT, kv_end - t0);
    for (int t = 0; t < tmax; ++t) {
      float dot = 0.f;
#pragma unroll
      for (int i = 0; i < HD; ++i) dot = fmaf(qr[i], sk[t][half * HD + i], dot);
      dot += __shfl_xor_sync(0xffffffffu, dot, 1);
      if (t0 + t <= row + koff) {                   // causal
        const float sc = dot * scale_log2;
        const float mn = fmaxf(m, sc);
        const float alpha = ex2(m - mn), pr = ex2(sc - mn);
        l = l * alpha + pr;
        m = mn;
#pragma unroll
        for (int i = 0; i < HD; ++i) acc[i] = fmaf(pr, sv[t][half * HD + i], acc[i] * alpha);
      }
    }
  }
  if (row_ok) {
    float* op = out + (static_cast<int64_t>(qtok0) + row) * out_stride + head * D + half * HD;
    const float inv = 1.f / l;
#pragma unroll
    for (int i = 0; i < HD; ++i) op[i] = acc[i] * inv;
  }
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  });
  return fn;
}

// [token][head][d] view with a token stride: dims (d, head, token), box 64 x 1 x 128, 128-byte swizzle
static int make_qkv_map(CUtensorMap* map, const void* base, int dtype, int head_size, int num_heads,
                        int64_t num_tokens, int64_t row_stride) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return set_error(PA_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
  const cuuint64_t dims[3] = {static_cast<cuuint64_t>(head_size), static_cast<cuuint64_t>(num_heads),
                              static_cast<cuuint64_t>(num_tokens)};
  const cuuint64_t strides[2] = {static_cast<cuuint64_t>(head_size) * 2, static_cast<cuuint64_t>(row_stride) * 2};
  const cuuint32_t box[3] = {64, 1, 128};
  const cuuint32_t estr[3] = {1, 1, 1};
  const CUresult r = fn(map, dtype == PA_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3,
                        const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(PA_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", static_cast<int>(r));
  return PA_OK;
}

// Tensor maps over the paged caches for the prompt kernel's page producer (prefill_tc.cuh).  The number of blocks
// is not part of the operator's interface, so the outermost extent is a bound on (block, head) rows, not the
// allocation: TMA only ever touches the boxes it is asked for.
//   K [block][head][D/8][B][8]: (64 elements = 8 tokens x 8 = one 128-byte core matrix, B/8 token groups, D/8 chunks,
//     block * heads + head), box 64 x 1 x D/8 x 1, no swizzle
//   V [block][head][D][B]: (B tokens, D rows, block * heads + head), box B x D x 1, swizzle = the row size
//     (B = 8: 16-byte rows, none; 16: 32 B; 32: 64 B)
static int make_page_maps(CUtensorMap* mk, CUtensorMap* mv, const void* key_cache, const void* value_cache, int dtype,
                          int head_size, int block_size) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return set_error(PA_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
  const CUtensorMapDataType dt = dtype == PA_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  const cuuint64_t rows = cuuint64_t{1} << 30;
  const cuuint64_t page_bytes = static_cast<cuuint64_t>(head_size) * block_size * 2;
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  {
    const cuuint64_t dims[4] = {64, static_cast<cuuint64_t>(block_size / 8), static_cast<cuuint64_t>(head_size / 8), rows};
    const cuuint64_t strides[3] = {128, static_cast<cuuint64_t>(block_size) * 16, page_bytes};
    const cuuint32_t box[4] = {64, 1, static_cast<cuuint32_t>(head_size / 8), 1};
    const CUresult r = fn(mk, dt, 4, const_cast<void*>(key_cache), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(PA_ERR_CUDA, "cuTensorMapEncodeTiled (key pages) failed with CUresult %d", static_cast<int>(r));
  }
  {
    const cuuint64_t dims[3] = {static_cast<cuuint64_t>(block_size), static_cast<cuuint64_t>(head_size), rows};
    const cuuint64_t strides[2] = {static_cast<cuuint64_t>(block_size) * 2, page_bytes};
    const cuuint32_t box[3] = {static_cast<cuuint32_t>(block_size), static_cast<cuuint32_t>(head_size), 1};
    const CUtensorMapSwizzle sw = block_size == 8 ? CU_TENSOR_MAP_SWIZZLE_NONE
                                  : block_size == 16 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_64B;
    const CUresult r = fn(mv, dt, 3, const_cast<void*>(value_cache), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(PA_ERR_CUDA, "cuTensorMapEncodeTiled (value pages) failed with CUresult %d", static_cast<int>(r));
  }
  return PA_OK;
}

template <typename T, int DP, bool kPaged>
static int launch_tc(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, const CUtensorMap& mo,
                     const PrefillTiles& tiles,
                     int ntiles, const PrefillParams& p, cudaStream_t stream) {
  constexpr int smem = 7 * (DP / 64) * 128 * 128 + 1024 + 256;   // Q x 2, K x 2, V x 2, output staging; alignment slack; barriers
  auto kern = prefill_tc_kernel<T, DP, kPaged>;
  PA_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  kern<<<p.tile_major ? dim3(ntiles, p.num_heads) : dim3(p.num_heads, ntiles), kPfThreads, smem, stream>>>(mq, mk, mv, mo, tiles, p);
  count_launch();
  return check_cuda(cudaGetLastError(), "prefill_tc_kernel launch");
}

// keys / values in the paged caches instead of packed rows (16-bit types)
struct PagedKV {
  const void* key_cache;
  const void* value_cache;
  const int32_t* block_tables;
  int32_t max_blocks;
  int32_t block_size;
};

template <int D>
static int launch_f32(const void* q, const void* k, const void* v, void* out, int64_t qs, int64_t ks, int64_t vs,
                      int64_t os, const PrefillTiles& tiles, int ntiles, int num_heads, float scale_log2,
                      cudaStream_t stream) {
  prefill_f32_kernel<D><<<dim3(num_heads, ntiles), 128, 0, stream>>>(
      static_cast<const float*>(q), static_cast<const float*>(k), static_cast<const float*>(v),
      static_cast<float*>(out), qs, ks, vs, os, tiles, scale_log2);
  count_launch();
  return check_cuda(cudaGetLastError(), "prefill_f32_kernel launch");
}

// key_lens == nullptr: plain prompt attention (every sequence's keys are its own query tokens' keys).
// Otherwise sequence i has seq_lens[i] query rows behind key_lens[i] - seq_lens[i] earlier keys.
static int prefill(void* out, const void* query, const void* key, const void* value, const int32_t* seq_lens,
                   const int32_t* key_lens, int32_t num_seqs, int32_t num_heads, int32_t head_size, int64_t q_stride,
                   int64_t k_stride, int64_t v_stride, int64_t out_stride, float scale, int32_t dtype, float* dbg,
                   cudaStream_t stream, const PagedKV* paged = nullptr) {
  if (dtype != PA_F16 && dtype != PA_BF16 && dtype != PA_F32)
    return set_error(PA_ERR_UNSUPPORTED_DTYPE, "unsupported data type: %d", dtype);
  if (head_size != 64 && head_size != 80 && head_size != 96 && head_size != 128)
    return set_error(PA_ERR_UNSUPPORTED_HEAD_SIZE, "unsupported head size: %d", head_size);
  if (num_seqs < 0 || num_heads <= 0 || num_heads > 65535) return set_error(PA_ERR_INVALID_ARGUMENT, "bad sizes");
  if (num_seqs == 0) return PA_OK;
  if (paged) { key = paged->key_cache; value = paged->value_cache; k_stride = v_stride = 8; }
  if (!out || !query || !key || !value || !seq_lens) return set_error(PA_ERR_INVALID_ARGUMENT, "null pointer");
  const int es = dtype_size(dtype);
  const uintptr_t ptrs = reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(query) |
                         reinterpret_cast<uintptr_t>(key) | reinterpret_cast<uintptr_t>(value);
  if ((ptrs & 15u) || ((q_stride | k_stride | v_stride | out_stride) * es) % 16)
    return set_error(PA_ERR_INVALID_ARGUMENT, "query/key/value/out rows must be 16-byte aligned");

  // tiles: one per 256 (16-bit) or 64 (fp32) query rows of a sequence; longest K/V span first
  const int rows = dtype == PA_F32 ? 64 : kPfRows;
  struct Tile { int32_t qtok0, qlen, q0, ktok0, klen, span; };
  std::vector<Tile> all;
  int64_t tok = 0, ktok = 0;
  for (int i = 0; i < num_seqs; ++i) {
    const int32_t len = seq_lens[i];
    const int32_t klen = key_lens ? key_lens[i] : len;
    if (len < 0) return set_error(PA_ERR_INVALID_ARGUMENT, "negative prompt length");
    if (klen < len) return set_error(PA_ERR_INVALID_ARGUMENT, "sequence %d: %d keys for %d query tokens", i, klen, len);
    if (tok + len > 0x7fffffff || ktok + klen > 0x7fffffff) return set_error(PA_ERR_INVALID_ARGUMENT, "too many tokens");
    for (int32_t q0 = 0; q0 < len; q0 += rows)
      all.push_back({static_cast<int32_t>(tok), len, q0, paged ? i : static_cast<int32_t>(ktok), klen,
                     std::min(q0 + rows + (klen - len), klen)});
    tok += len;
    ktok += klen;
  }
  if (all.empty()) return PA_OK;
  std::stable_sort(all.begin(), all.end(), [](const Tile& a, const Tile& b) { return a.span > b.span; });

  CUtensorMap mq, mk, mv, mo;
  const int dp = head_size <= 64 ? 64 : 128;
  if (dtype != PA_F32) {
    int rc = make_qkv_map(&mq, query, dtype, head_size, num_heads, tok, q_stride);
    if (rc == PA_OK && paged)
      rc = make_page_maps(&mk, &mv, paged->key_cache, paged->value_cache, dtype, head_size, paged->block_size);
    if (rc == PA_OK && !paged) rc = make_qkv_map(&mk, key, dtype, head_size, num_heads, ktok, k_stride);
    if (rc == PA_OK && !paged) rc = make_qkv_map(&mv, value, dtype, head_size, num_heads, ktok, v_stride);
    if (rc == PA_OK) rc = make_qkv_map(&mo, out, dtype, head_size, num_heads, tok, out_stride);
    if (rc != PA_OK) return rc;
  }
  PrefillParams p;
  p.out = out; p.out_stride = out_stride; p.num_heads = num_heads; p.head_size = head_size;
  p.scale_log2 = scale * 1.4426950408889634f; p.dbg = dbg;
  // K/V bytes of the batch against the L2 (126 MB on B200): beyond it, query tiles of one head run together
  p.tile_major = ktok * num_heads * head_size * 4 > (int64_t{96} << 20);
  p.key_cache = nullptr; p.value_cache = nullptr; p.block_tables = nullptr; p.max_blocks = 0; p.block_shift = 0;
  if (paged) {
    p.key_cache = paged->key_cache; p.value_cache = paged->value_cache; p.block_tables = paged->block_tables;
    p.max_blocks = paged->max_blocks;
    while ((1 << p.block_shift) < paged->block_size) ++p.block_shift;
  }

  static thread_local PrefillTiles tiles;
  for (size_t begin = 0; begin < all.size(); begin += kPfMaxTiles) {
    const int n = static_cast<int>(std::min<size_t>(kPfMaxTiles, all.size() - begin));
    for (int i = 0; i < n; ++i) {
      const Tile& t = all[begin + i];
      tiles.qtok0[i] = t.qtok0; tiles.qlen[i] = t.qlen; tiles.q0[i] = t.q0; tiles.ktok0[i] = t.ktok0; tiles.klen[i] = t.klen;
    }
    int rc;
    if (dtype == PA_F32) {
      switch (head_size) {
        case 64:  rc = launch_f32<64>(query, key, value, out, q_stride, k_stride, v_stride, out_stride, tiles, n, num_heads, p.scale_log2, stream); break;
        case 80:  rc = launch_f32<80>(query, key, value, out, q_stride, k_stride, v_stride, out_stride, tiles, n, num_heads, p.scale_log2, stream); break;
        case 96:  rc = launch_f32<96>(query, key, value, out, q_stride, k_stride, v_stride, out_stride, tiles, n, num_heads, p.scale_log2, stream); break;
        default:  rc = launch_f32<128>(query, key, value, out, q_stride, k_stride, v_stride, out_stride, tiles, n, num_heads, p.scale_log2, stream); break;
      }
    } else if (paged) {
      if (dtype == PA_F16) rc = dp == 64 ? launch_tc<F16, 64, true>(mq, mk, mv, mo, tiles, n, p, stream) : launch_tc<F16, 128, true>(mq, mk, mv, mo, tiles, n, p, stream);
      else rc = dp == 64 ? launch_tc<BF16, 64, true>(mq, mk, mv, mo, tiles, n, p, stream) : launch_tc<BF16, 128, true>(mq, mk, mv, mo, tiles, n, p, stream);
    } else if (dtype == PA_F16) {
      rc = dp == 64 ? launch_tc<F16, 64, false>(mq, mk, mv, mo, tiles, n, p, stream) : launch_tc<F16, 128, false>(mq, mk, mv, mo, tiles, n, p, stream);
    } else {
      rc = dp == 64 ? launch_tc<BF16, 64, false>(mq, mk, mv, mo, tiles, n, p, stream) : launch_tc<BF16, 128, false>(mq, mk, mv, mo, tiles, n, p, stream);
    }
    if (rc != PA_OK) return rc;
    p.dbg = nullptr;
  }
  return PA_OK;
}

// ---------------------------------------------------------------------------
// prompt chunks over the paged cache (multi_query_cached_kv_attention)
//
// Parity reference: ref_multi_query_cached_kv_attention (tests/kernels/test_attention.py:104-156): sequence i
// has query_lens[i] new tokens whose keys / values are already in the cache together with the
// context_lens[i] - query_lens[i] tokens before them; query row r sees keys 0 .. r + (context - query_len).
// The reference tree keeps only that test reference (no kernel).  16-bit types: the tcgen05 prompt kernel with
// the diagonal shifted by context - query_len and a page producer (prefill_tc.cuh, kPaged) -- one launch, no
// scratch.  fp32: the pages of each context are gathered once into packed [token][head][d] rows (the tile mover
// of gather_cached_kv, slots generated on the device from the block tables) for the exact SIMT kernel.
// ---------------------------------------------------------------------------
constexpr int kSlotSeqs = 1024;
struct SlotSeqs {
  int32_t ktok0[kSlotSeqs];  // first packed token of the sequence's context
  int32_t klen[kSlotSeqs];   // context length
};

__global__ void __launch_bounds__(256)
page_slots_kernel(const __grid_constant__ SlotSeqs t, const int32_t* __restrict__ block_tables, int32_t seq0,
                  int32_t max_blocks, int32_t block_size, int32_t* __restrict__ slots) {
  const int i = blockIdx.y;
  const int klen = t.klen[i];
  const int32_t* row = block_tables + static_cast<int64_t>(seq0 + i) * max_blocks;
  for (int u = blockIdx.x * blockDim.x + threadIdx.x; u < klen; u += gridDim.x * blockDim.x)
    slots[t.ktok0[i] + u] = row[u / block_size] * block_size + u % block_size;
}

static int64_t cached_ws_rows_bytes(int64_t tokens, int32_t num_heads, int32_t head_size, int32_t es) {
  return (tokens * num_heads * head_size * es + 255) / 256 * 256;
}

static int cached_prefill(void* out, const void* query, const void* key_cache, const void* value_cache, float scale,
                          const int32_t* block_tables, const int32_t* query_lens, const int32_t* context_lens,
                          int32_t num_seqs, int32_t num_heads, int32_t head_size, int32_t block_size,
                          int32_t max_blocks, int64_t q_stride, int64_t out_stride, int32_t dtype, void* workspace,
                          int64_t workspace_bytes, cudaStream_t stream) {
  if (dtype != PA_F16 && dtype != PA_BF16 && dtype != PA_F32)
    return set_error(PA_ERR_UNSUPPORTED_DTYPE, "unsupported data type: %d", dtype);
  if (head_size != 64 && head_size != 80 && head_size != 96 && head_size != 128)
    return set_error(PA_ERR_UNSUPPORTED_HEAD_SIZE, "unsupported head size: %d", head_size);
  if (block_size != 8 && block_size != 16 && block_size != 32)
    return set_error(PA_ERR_UNSUPPORTED_BLOCK_SIZE, "unsupported block size: %d", block_size);
  if (num_seqs < 0 || num_heads <= 0 || max_blocks < 0) return set_error(PA_ERR_INVALID_ARGUMENT, "bad sizes");
  if (num_seqs == 0) return PA_OK;
  if (!key_cache || !value_cache || !block_tables || !query_lens || !context_lens)
    return set_error(PA_ERR_INVALID_ARGUMENT, "null pointer");
  int64_t total = 0, max_ctx = 0, total_q = 0;
  for (int i = 0; i < num_seqs; ++i) {
    const int64_t c = context_lens[i], q = query_lens[i];
    if (q < 0 || c < q) return set_error(PA_ERR_INVALID_ARGUMENT, "sequence %d: context %lld < query length %lld", i,
                                         static_cast<long long>(c), static_cast<long long>(q));
    if (c > static_cast<int64_t>(max_blocks) * block_size)
      return set_error(PA_ERR_INVALID_ARGUMENT, "sequence %d: context %lld exceeds its block table (%d blocks of %d)", i,
                       static_cast<long long>(c), max_blocks, block_size);
    total += c; total_q += q; max_ctx = std::max(max_ctx, c);
  }
  if (total_q == 0) return PA_OK;
  if (total > 0x7fffffff) return set_error(PA_ERR_INVALID_ARGUMENT, "too many context tokens");
  if (dtype != PA_F32) {
    // 16-bit: one launch; the prompt kernel's producer warps read the pages themselves (no gather, no scratch)
    if ((reinterpret_cast<uintptr_t>(key_cache) | reinterpret_cast<uintptr_t>(value_cache)) & 15u)
      return set_error(PA_ERR_INVALID_ARGUMENT, "caches must be 16-byte aligned");
    const PagedKV paged = {key_cache, value_cache, block_tables, max_blocks, block_size};
    return prefill(out, query, nullptr, nullptr, query_lens, context_lens, num_seqs, num_heads, head_size, q_stride, 0, 0,
                   out_stride, scale, dtype, nullptr, stream, &paged);
  }
  const int es = dtype_size(dtype);
  const int64_t rows_bytes = cached_ws_rows_bytes(total, num_heads, head_size, es);
  const int64_t need = 2 * rows_bytes + total * 4;
  if (!workspace || workspace_bytes < need || (reinterpret_cast<uintptr_t>(workspace) & 255u))
    return set_error(PA_ERR_INVALID_ARGUMENT, "workspace of %lld bytes (256-byte aligned) needed, %lld given",
                     static_cast<long long>(need), static_cast<long long>(workspace_bytes));
  char* kws = static_cast<char*>(workspace);
  char* vws = kws + rows_bytes;
  int32_t* slots = reinterpret_cast<int32_t*>(vws + rows_bytes);

  static thread_local SlotSeqs seqs;
  int64_t ktok = 0;
  for (int begin = 0; begin < num_seqs; begin += kSlotSeqs) {
    const int n = std::min(kSlotSeqs, num_seqs - begin);
    int32_t chunk_max = 1;
    for (int i = 0; i < n; ++i) {
      seqs.ktok0[i] = static_cast<int32_t>(ktok); seqs.klen[i] = context_lens[begin + i];
      ktok += context_lens[begin + i];
      chunk_max = std::max(chunk_max, context_lens[begin + i]);
    }
    page_slots_kernel<<<dim3((chunk_max + 255) / 256, n), 256, 0, stream>>>(seqs, block_tables, begin, max_blocks,
                                                                           block_size, slots);
    count_launch();
    const int rc = check_cuda(cudaGetLastError(), "page_slots_kernel launch");
    if (rc != PA_OK) return rc;
  }
  const int64_t row = static_cast<int64_t>(num_heads) * head_size;
  int rc = pa_gather_cached_kv(kws, vws, key_cache, value_cache, slots, static_cast<int32_t>(total), num_heads,
                               head_size, block_size, row, row, dtype, stream);
  if (rc != PA_OK) return rc;
  return prefill(out, query, kws, vws, query_lens, context_lens, num_seqs, num_heads, head_size, q_stride, row, row,
                 out_stride, scale, dtype, nullptr, stream);
}

}  // namespace pa

extern "C" {

int64_t pa_multi_query_cached_kv_workspace_bytes(int64_t total_context_tokens, int32_t num_heads, int32_t head_size,
                                                 int32_t dtype) {
  if (total_context_tokens < 0 || num_heads <= 0 || head_size <= 0 ||
      (dtype != PA_F16 && dtype != PA_BF16 && dtype != PA_F32))
    return pa::set_error(PA_ERR_INVALID_ARGUMENT, "bad workspace query");
  if (dtype != PA_F32) return 0;                   // 16-bit types read the pages directly
  return 2 * pa::cached_ws_rows_bytes(total_context_tokens, num_heads, head_size, pa::dtype_size(dtype)) +
         total_context_tokens * 4;
}

int pa_multi_query_cached_kv_attention(void* out, const void* query, const void* key_cache, const void* value_cache,
                                       float scale, const int32_t* block_tables, const int32_t* query_lens,
                                       const int32_t* context_lens, int32_t num_seqs, int32_t num_heads,
                                       int32_t head_size, int32_t block_size, int32_t max_num_blocks_per_seq,
                                       int64_t q_stride, int64_t out_stride, int32_t dtype, void* workspace,
                                       int64_t workspace_bytes, void* stream) {
  return pa::cached_prefill(out, query, key_cache, value_cache, scale, block_tables, query_lens, context_lens, num_seqs,
                            num_heads, head_size, block_size, max_num_blocks_per_seq, q_stride, out_stride, dtype,
                            workspace, workspace_bytes, static_cast<cudaStream_t>(stream));
}


int pa_prefill_attention(void* out, const void* query, const void* key, const void* value, const int32_t* seq_lens,
                         int32_t num_seqs, int32_t num_heads, int32_t head_size, int64_t q_stride, int64_t k_stride,
                         int64_t v_stride, int64_t out_stride, float scale, int32_t dtype, void* stream) {
  return pa::prefill(out, query, key, value, seq_lens, nullptr, num_seqs, num_heads, head_size, q_stride, k_stride,
                     v_stride, out_stride, scale, dtype, nullptr, static_cast<cudaStream_t>(stream));
}

int pa_prefill_attention_debug(void* out, const void* query, const void* key, const void* value,
                               const int32_t* seq_lens, int32_t num_seqs, int32_t num_heads, int32_t head_size,
                               int64_t q_stride, int64_t k_stride, int64_t v_stride, int64_t out_stride, float scale,
                               int32_t dtype, float* debug, void* stream) {
  return pa::prefill(out, query, key, value, seq_lens, nullptr, num_seqs, num_heads, head_size, q_stride, k_stride,
                     v_stride, out_stride, scale, dtype, debug, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
