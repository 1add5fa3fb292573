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
  const int tok0 = tiles.tok0[blockIdx.y], len = tiles.len[blockIdx.y], q0 = tiles.q0[blockIdx.y];
  const int row = q0 + (threadIdx.x >> 1);          // query row inside the prompt
  const int half = threadIdx.x & 1;
  const bool row_ok = row < len;
  float qr[HD], acc[HD];
  {
    const float* qp = q + (static_cast<int64_t>(tok0) + (row_ok ? row : q0)) * q_stride + head * D + half * HD;
#pragma unroll
    for (int i = 0; i < HD; ++i) { qr[i] = qp[i]; acc[i] = 0.f; }
  }
  float m = -CUDART_INF_F, l = 0.f;
  const int kv_end = min(q0 + 64, len);
  for (int t0 = 0; t0 < kv_end; t0 += KT) {
    __syncthreads();
    for (int i = threadIdx.x; i < KT * D; i += blockDim.x) {
      const int t = i / D, d = i - t * D;
      const bool ok = t0 + t < kv_end;
      sk[t][d] = ok ? k[(static_cast<int64_t>(tok0) + t0 + t) * k_stride + head * D + d] : 0.f;
      sv[t][d] = ok ? v[(static_cast<int64_t>(tok0) + t0 + t) * v_stride + head * D + d] : 0.f;
    }
    __syncthreads();
    const int tmax = min(KT, kv_end - t0);
    for (int t = 0; t < tmax; ++t) {
      float dot = 0.f;
#pragma unroll
      for (int i = 0; i < HD; ++i) dot = fmaf(qr[i], sk[t][half * HD + i], dot);
      dot += __shfl_xor_sync(0xffffffffu, dot, 1);
      if (t0 + t <= row) {                          // causal
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
    float* op = out + (static_cast<int64_t>(tok0) + row) * out_stride + head * D + half * HD;
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

template <typename T, int DP>
static int launch_tc(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, const PrefillTiles& tiles,
                     int ntiles, const PrefillParams& p, cudaStream_t stream) {
  constexpr int smem = 6 * (DP / 64) * 128 * 128 + 1024 + 128;
  auto kern = prefill_tc_kernel<T, DP>;
  PA_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  kern<<<dim3(p.num_heads, ntiles), kPfThreads, smem, stream>>>(mq, mk, mv, tiles, p);
  count_launch();
  return check_cuda(cudaGetLastError(), "prefill_tc_kernel launch");
}

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

static int prefill(void* out, const void* query, const void* key, const void* value, const int32_t* seq_lens,
                   int32_t num_seqs, int32_t num_heads, int32_t head_size, int64_t q_stride, int64_t k_stride,
                   int64_t v_stride, int64_t out_stride, float scale, int32_t dtype, float* dbg, cudaStream_t stream) {
  if (dtype != PA_F16 && dtype != PA_BF16 && dtype != PA_F32)
    return set_error(PA_ERR_UNSUPPORTED_DTYPE, "unsupported data type: %d", dtype);
  if (head_size != 64 && head_size != 80 && head_size != 96 && head_size != 128)
    return set_error(PA_ERR_UNSUPPORTED_HEAD_SIZE, "unsupported head size: %d", head_size);
  if (num_seqs < 0 || num_heads <= 0 || num_heads > 65535) return set_error(PA_ERR_INVALID_ARGUMENT, "bad sizes");
  if (num_seqs == 0) return PA_OK;
  if (!out || !query || !key || !value || !seq_lens) return set_error(PA_ERR_INVALID_ARGUMENT, "null pointer");
  const int es = dtype_size(dtype);
  const uintptr_t ptrs = reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(query) |
                         reinterpret_cast<uintptr_t>(key) | reinterpret_cast<uintptr_t>(value);
  if ((ptrs & 15u) || ((q_stride | k_stride | v_stride | out_stride) * es) % 16)
    return set_error(PA_ERR_INVALID_ARGUMENT, "query/key/value/out rows must be 16-byte aligned");

  // tiles: one per 128 (16-bit) or 64 (fp32) query rows of a prompt; longest K/V span first
  const int rows = dtype == PA_F32 ? 64 : kPfRows;
  struct Tile { int32_t tok0, len, q0; };
  std::vector<Tile> all;
  int64_t tok = 0;
  for (int i = 0; i < num_seqs; ++i) {
    const int32_t len = seq_lens[i];
    if (len < 0) return set_error(PA_ERR_INVALID_ARGUMENT, "negative prompt length");
    if (tok + len > 0x7fffffff) return set_error(PA_ERR_INVALID_ARGUMENT, "too many tokens");
    for (int32_t q0 = 0; q0 < len; q0 += rows) all.push_back({static_cast<int32_t>(tok), len, q0});
    tok += len;
  }
  if (all.empty()) return PA_OK;
  std::stable_sort(all.begin(), all.end(), [](const Tile& a, const Tile& b) { return a.q0 > b.q0; });

  CUtensorMap mq, mk, mv;
  const int dp = head_size <= 64 ? 64 : 128;
  if (dtype != PA_F32) {
    int rc = make_qkv_map(&mq, query, dtype, head_size, num_heads, tok, q_stride);
    if (rc == PA_OK) rc = make_qkv_map(&mk, key, dtype, head_size, num_heads, tok, k_stride);
    if (rc == PA_OK) rc = make_qkv_map(&mv, value, dtype, head_size, num_heads, tok, v_stride);
    if (rc != PA_OK) return rc;
  }
  PrefillParams p;
  p.out = out; p.out_stride = out_stride; p.num_heads = num_heads; p.head_size = head_size;
  p.scale_log2 = scale * 1.4426950408889634f; p.dbg = dbg;

  static thread_local PrefillTiles tiles;
  for (size_t begin = 0; begin < all.size(); begin += kPfMaxTiles) {
    const int n = static_cast<int>(std::min<size_t>(kPfMaxTiles, all.size() - begin));
    for (int i = 0; i < n; ++i) {
      tiles.tok0[i] = all[begin + i].tok0; tiles.len[i] = all[begin + i].len; tiles.q0[i] = all[begin + i].q0;
    }
    int rc;
    if (dtype == PA_F32) {
      switch (head_size) {
        case 64:  rc = launch_f32<64>(query, key, value, out, q_stride, k_stride, v_stride, out_stride, tiles, n, num_heads, p.scale_log2, stream); break;
        case 80:  rc = launch_f32<80>(query, key, value, out, q_stride, k_stride, v_stride, out_stride, tiles, n, num_heads, p.scale_log2, stream); break;
        case 96:  rc = launch_f32<96>(query, key, value, out, q_stride, k_stride, v_stride, out_stride, tiles, n, num_heads, p.scale_log2, stream); break;
        default:  rc = launch_f32<128>(query, key, value, out, q_stride, k_stride, v_stride, out_stride, tiles, n, num_heads, p.scale_log2, stream); break;
      }
    } else if (dtype == PA_F16) {
      rc = dp == 64 ? launch_tc<F16, 64>(mq, mk, mv, tiles, n, p, stream) : launch_tc<F16, 128>(mq, mk, mv, tiles, n, p, stream);
    } else {
      rc = dp == 64 ? launch_tc<BF16, 64>(mq, mk, mv, tiles, n, p, stream) : launch_tc<BF16, 128>(mq, mk, mv, tiles, n, p, stream);
    }
    if (rc != PA_OK) return rc;
    p.dbg = nullptr;
  }
  return PA_OK;
}

}  // namespace pa

extern "C" {

int pa_prefill_attention(void* out, const void* query, const void* key, const void* value, const int32_t* seq_lens,
                         int32_t num_seqs, int32_t num_heads, int32_t head_size, int64_t q_stride, int64_t k_stride,
                         int64_t v_stride, int64_t out_stride, float scale, int32_t dtype, void* stream) {
  return pa::prefill(out, query, key, value, seq_lens, num_seqs, num_heads, head_size, q_stride, k_stride, v_stride,
                     out_stride, scale, dtype, nullptr, static_cast<cudaStream_t>(stream));
}

int pa_prefill_attention_debug(void* out, const void* query, const void* key, const void* value,
                               const int32_t* seq_lens, int32_t num_seqs, int32_t num_heads, int32_t head_size,
                               int64_t q_stride, int64_t k_stride, int64_t v_stride, int64_t out_stride, float scale,
                               int32_t dtype, float* debug, void* stream) {
  return pa::prefill(out, query, key, value, seq_lens, num_seqs, num_heads, head_size, q_stride, k_stride, v_stride,
                     out_stride, scale, dtype, debug, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
