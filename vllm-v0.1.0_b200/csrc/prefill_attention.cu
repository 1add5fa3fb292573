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

#include "tc05.cuh"

namespace pa {

constexpr int kPfRows = 128;       // query rows per CTA = TMEM lanes
constexpr int kPfCols = 128;       // K/V tokens per tile
constexpr int kPfMaxTiles = 1024;  // tiles per launch (the table travels as a kernel parameter)

struct PrefillTiles {
  int32_t tok0[kPfMaxTiles];  // first token of the prompt in the packed batch
  int32_t len[kPfMaxTiles];   // prompt length
  int32_t q0[kPfMaxTiles];    // first query row of this tile inside the prompt (multiple of 128)
};

struct PrefillParams {
  void* out;
  int64_t out_stride;  // elements between consecutive tokens of out
  int32_t num_heads;
  int32_t head_size;
  float scale_log2;
  float* dbg;          // tests: raw S of the first tile, raw O, row sums of CTA (0, 0)
};

__device__ __forceinline__ float ex2(float x) {   // exp2(-inf) == 0
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// tensor-memory columns
constexpr uint32_t kTmS0 = 0, kTmS1 = 128, kTmO = 256, kTmP = 384, kTmCols = 512;

template <typename T, int DP>
__global__ void __launch_bounds__(128, 1)
prefill_tc_kernel(const __grid_constant__ CUtensorMap tmq, const __grid_constant__ CUtensorMap tmk,
                  const __grid_constant__ CUtensorMap tmv, const __grid_constant__ PrefillTiles tiles,
                  const PrefillParams p) {
  constexpr int kAtoms = DP / 64;                 // 64-element (128-byte) column groups of a tile
  constexpr uint32_t kAtomBytes = 128 * 128;      // 128 rows x 128 bytes
  constexpr uint32_t kTileBytes = kAtoms * kAtomBytes;
  constexpr bool kBf16 = T::kId == PA_BF16;
  constexpr uint32_t kIdescQK = tc05::instr_desc_f16(kBf16, 128, 128, false, false);
  constexpr uint32_t kIdescPV = tc05::instr_desc_f16(kBf16, 128, DP, false, true);

  extern __shared__ unsigned char smem_raw[];
  const uint32_t base = (smem_addr(smem_raw) + 1023u) & ~1023u;
  const uint32_t sQ = base;
  const uint32_t sK = sQ + kTileBytes;            // two stages
  const uint32_t sV = sK + 2 * kTileBytes;        // two stages
  const uint32_t sBar = sV + 2 * kTileBytes;
  const uint32_t bar_q = sBar, bar_k = sBar + 8, bar_v = sBar + 24, bar_s = sBar + 40, bar_o = sBar + 56;
  const uint32_t tm_slot = sBar + 64;

  const int tid = threadIdx.x, warp = tid >> 5;
  const int head = blockIdx.x;
  const int tok0 = tiles.tok0[blockIdx.y], len = tiles.len[blockIdx.y], q0 = tiles.q0[blockIdx.y];
  const int kv_end = min(q0 + kPfRows, len);      // causal: nothing right of the diagonal tile
  const int ntiles = (kv_end + kPfCols - 1) / kPfCols;

  if (tid == 0) {
    tc05::tma_prefetch_desc(&tmq); tc05::tma_prefetch_desc(&tmk); tc05::tma_prefetch_desc(&tmv);
    mbar_init(bar_q, 1);
    for (int s = 0; s < 2; ++s) { mbar_init(bar_k + 8 * s, 1); mbar_init(bar_v + 8 * s, 1); mbar_init(bar_s + 8 * s, 1); }
    mbar_init(bar_o, 1);
    mbar_init_fence();
  }
  if (warp == 0) { __syncwarp(); tc05::tmem_alloc<kTmCols>(tm_slot); }
  tc05::fence_before_sync();
  __syncthreads();
  tc05::fence_after_sync();
  uint32_t tmem;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(tm_slot) : "memory");
  const uint32_t lane_base = tmem + (static_cast<uint32_t>(warp * 32) << 16);   // this warp's 32 TMEM lanes

  auto load_tile = [&](const CUtensorMap* map, uint32_t dst, int row0, uint32_t bar) {
    mbar_expect_tx(bar, kTileBytes);
#pragma unroll
    for (int a = 0; a < kAtoms; ++a) tc05::tma_load_3d(dst + a * kAtomBytes, map, a * 64, head, row0, bar);
  };
  auto issue_qk = [&](int j) {                     // S[j & 1] = Q . K_j^T
    const uint32_t kb = sK + (j & 1) * kTileBytes;
#pragma unroll
    for (int kk = 0; kk < DP / 16; ++kk) {
      const uint32_t off = (kk >> 2) * kAtomBytes + (kk & 3) * 32;
      tc05::mma_ss(tmem + ((j & 1) ? kTmS1 : kTmS0), tc05::smem_desc_sw128(sQ + off, 16, 1024),
                   tc05::smem_desc_sw128(kb + off, 16, 1024), kIdescQK, kk > 0);
    }
    tc05::mma_commit(bar_s + 8 * (j & 1));
  };
  auto issue_pv = [&](int j) {                     // O (+)= P . V_j
    const uint32_t vb = sV + (j & 1) * kTileBytes;
#pragma unroll
    for (int kk = 0; kk < kPfCols / 16; ++kk)
      tc05::mma_ts(tmem + kTmO, tmem + kTmP + kk * 8, tc05::smem_desc_sw128(vb + kk * 2048, kAtomBytes, 1024),
                   kIdescPV, j > 0 || kk > 0);
    tc05::mma_commit(bar_o);
  };

  if (tid == 0) {
    load_tile(&tmq, sQ, tok0 + q0, bar_q);
    for (int j = 0; j < 2 && j < ntiles; ++j) {
      load_tile(&tmk, sK + j * kTileBytes, tok0 + j * kPfCols, bar_k + 8 * j);
      load_tile(&tmv, sV + j * kTileBytes, tok0 + j * kPfCols, bar_v + 8 * j);
    }
    mbar_wait(bar_q, 0);
    mbar_wait(bar_k, 0);
    tc05::fence_after_sync();
    issue_qk(0);
  }
  __syncwarp();

  float m_run = -CUDART_INF_F, l_run = 0.f;
  const int r = tid;                                // row of this thread inside the tile
  for (int j = 0; j < ntiles; ++j) {
    if (tid == 0 && j + 1 < ntiles) {               // keep the tensor core busy during this tile's softmax
      mbar_wait(bar_k + 8 * ((j + 1) & 1), ((j + 1) >> 1) & 1);
      tc05::fence_after_sync();
      issue_qk(j + 1);
    }
    __syncwarp();
    mbar_wait(bar_s + 8 * (j & 1), (j >> 1) & 1);
    tc05::fence_after_sync();
    if (tid == 0 && j + 2 < ntiles)                 // K stage j & 1 is free again
      load_tile(&tmk, sK + (j & 1) * kTileBytes, tok0 + (j + 2) * kPfCols, bar_k + 8 * (j & 1));
    __syncwarp();

    // ---- logits of this row -------------------------------------------------------------
    uint32_t s[4][32];
#pragma unroll
    for (int c = 0; c < 4; ++c) tc05::tmem_ld32(lane_base + ((j & 1) ? kTmS1 : kTmS0) + c * 32, s[c]);
    tc05::tmem_wait_ld();
    if (p.dbg && blockIdx.x == 0 && blockIdx.y == 0 && j == 0) {
#pragma unroll
      for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int i = 0; i < 32; ++i) p.dbg[r * 128 + c * 32 + i] = __uint_as_float(s[c][i]);
    }
    const bool diag = j == ntiles - 1;              // the only tile that needs masking
    const int lim = min(r, len - 1 - q0);           // last valid column of this row in the diagonal tile
    float mx = -CUDART_INF_F;
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        float v = __uint_as_float(s[c][i]) * p.scale_log2;
        if (diag && c * 32 + i > lim) v = -CUDART_INF_F;
        s[c][i] = __float_as_uint(v);
        mx = fmaxf(mx, v);
      }
    const float m_new = fmaxf(m_run, mx);
    const float m_use = m_new == -CUDART_INF_F ? 0.f : m_new;   // row with nothing valid yet (beyond len)
    const float alpha = ex2(m_run - m_use);
    float sum = 0.f;
    uint32_t pk[2][32];
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
      for (int i = 0; i < 32; i += 2) {
        const float p0 = ex2(__uint_as_float(s[c][i]) - m_use);
        const float p1 = ex2(__uint_as_float(s[c][i + 1]) - m_use);
        sum += p0 + p1;
        pk[c >> 1][(c & 1) * 16 + (i >> 1)] = pack2<T>(p0, p1);
      }
    l_run = l_run * alpha + sum;
    m_run = m_new;

    // ---- previous P.V must be done before P and O are touched ------------------------------
    if (j > 0) {
      mbar_wait(bar_o, (j - 1) & 1);
      tc05::fence_after_sync();
      if (tid == 0 && j + 1 < ntiles)               // V stage (j + 1) & 1 is free again
        load_tile(&tmv, sV + ((j + 1) & 1) * kTileBytes, tok0 + (j + 1) * kPfCols, bar_v + 8 * ((j + 1) & 1));
      __syncwarp();
      if (__any_sync(0xffffffffu, alpha != 1.f)) {
#pragma unroll
        for (int c = 0; c < DP / 32; ++c) {
          uint32_t o[32];
          tc05::tmem_ld32(lane_base + kTmO + c * 32, o);
          tc05::tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
          tc05::tmem_st32(lane_base + kTmO + c * 32, o);
        }
      }
    }
    tc05::tmem_st32(lane_base + kTmP, pk[0]);
    tc05::tmem_st32(lane_base + kTmP + 32, pk[1]);
    tc05::tmem_wait_st();
    tc05::fence_before_sync();
    __syncthreads();
    if (tid == 0) {
      tc05::fence_after_sync();
      mbar_wait(bar_v + 8 * (j & 1), (j >> 1) & 1);
      issue_pv(j);
    }
    __syncwarp();
  }

  // ---- epilogue: O / l -> out ---------------------------------------------------------------
  mbar_wait(bar_o, (ntiles - 1) & 1);
  tc05::fence_after_sync();
  const int D = p.head_size;
  const bool row_ok = q0 + r < len;
  const float inv = 1.f / l_run;
  char* orow = reinterpret_cast<char*>(p.out) +
               ((static_cast<int64_t>(tok0) + q0 + r) * p.out_stride + static_cast<int64_t>(head) * D) * 2;
#pragma unroll
  for (int c = 0; c < DP / 32; ++c) {
    uint32_t o[32];
    tc05::tmem_ld32(lane_base + kTmO + c * 32, o);
    tc05::tmem_wait_ld();
    if (p.dbg && blockIdx.x == 0 && blockIdx.y == 0) {
#pragma unroll
      for (int i = 0; i < 32; ++i) p.dbg[128 * 128 + r * DP + c * 32 + i] = __uint_as_float(o[i]);
      if (c == 0) { p.dbg[128 * 128 + 128 * DP + r] = l_run; p.dbg[128 * 128 + 128 * DP + 128 + r] = m_run; }
    }
    if (row_ok) {
#pragma unroll
      for (int v = 0; v < 4; ++v) {                 // 8 elements = 16 bytes per store
        if (c * 32 + v * 8 < D) {
          uint4 w;
          w.x = pack2<T>(__uint_as_float(o[v * 8 + 0]) * inv, __uint_as_float(o[v * 8 + 1]) * inv);
          w.y = pack2<T>(__uint_as_float(o[v * 8 + 2]) * inv, __uint_as_float(o[v * 8 + 3]) * inv);
          w.z = pack2<T>(__uint_as_float(o[v * 8 + 4]) * inv, __uint_as_float(o[v * 8 + 5]) * inv);
          w.w = pack2<T>(__uint_as_float(o[v * 8 + 6]) * inv, __uint_as_float(o[v * 8 + 7]) * inv);
          *reinterpret_cast<uint4*>(orow + (c * 32 + v * 8) * 2) = w;
        }
      }
    }
  }
  tc05::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc05::tmem_free<kTmCols>(tmem);
}

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
  constexpr int smem = 5 * (DP / 64) * 128 * 128 + 1024 + 128;
  auto kern = prefill_tc_kernel<T, DP>;
  PA_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  kern<<<dim3(p.num_heads, ntiles), 128, smem, stream>>>(mq, mk, mv, tiles, p);
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
