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
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <map>
#include <mutex>
#include <queue>
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

template <typename T, int DP, bool kPaged, int kTiles>
static int launch_tc_n(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, const CUtensorMap& mo,
                       const PrefillTilesT<kTiles>& tiles, int ntiles, const PrefillParams& p, cudaStream_t stream) {
  constexpr int smem = 7 * (DP / 64) * 128 * 128 + 1024 + 288;   // Q x 2, K x 2, V x 2, output staging; alignment slack; barriers
  auto kern = prefill_tc_kernel<T, DP, kPaged, kTiles>;
  PA_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = p.tile_major ? dim3(ntiles, p.num_heads) : dim3(p.num_heads, ntiles);
  cfg.blockDim = dim3(kPfThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  // programmatic dependent launch: barrier set-up, tensor-memory allocation and descriptor prefetch of the CTAs
  // overlap the tail of the preceding kernel on the stream; the kernel touches no global memory before
  // griddepcontrol.wait (PA_PREFILL_PDL=0 switches it off)
  static const bool pdl = [] { const char* e = getenv("PA_PREFILL_PDL"); return !(e && e[0] == '0'); }();
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, mq, mk, mv, mo, tiles, p);
  count_launch();
  return check_cuda(e, "prefill_tc_kernel launch");
}

// up to kPfFewTiles entries travel in the small table (see PrefillTilesT)
template <typename T, int DP, bool kPaged>
static int launch_tc(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, const CUtensorMap& mo,
                     const PrefillTiles& tiles, int ntiles, const PrefillParams& p, cudaStream_t stream) {
  if (ntiles > kPfFewTiles) return launch_tc_n<T, DP, kPaged, kPfMaxTiles>(mq, mk, mv, mo, tiles, ntiles, p, stream);
  PrefillTilesT<kPfFewTiles> few;
  for (int i = 0; i < ntiles; ++i) {
    few.qtok0[i] = tiles.qtok0[i]; few.qlen[i] = tiles.qlen[i]; few.q0[i] = tiles.q0[i]; few.ktok0[i] = tiles.ktok0[i];
    few.klen[i] = tiles.klen[i]; few.krange[i] = tiles.krange[i]; few.slot[i] = tiles.slot[i];
  }
  return launch_tc_n<T, DP, kPaged, kPfFewTiles>(mq, mk, mv, mo, few, ntiles, p, stream);
}

// ---------------------------------------------------------------------------
// Context splits.  One long item on a GPU that has nothing else left to do (a short chunk behind a long cached
// context; the last, incomplete wave of a batch of equal items) is cut over its K/V tiles into parts that run as
// separate work items and leave (O, m, l) in a workspace; this kernel combines the parts of one 128-row group:
// out = sum_p 2^(m_p - M) O_p / sum_p 2^(m_p - M) l_p  (the same log-sum-exp merge the decode kernel uses).
// ---------------------------------------------------------------------------
constexpr int kPfMaxMerge = 256;
struct MergeTable {
  int32_t qtok0[kPfMaxMerge];   // first query token of the sequence in the packed batch
  int32_t qlen[kPfMaxMerge];
  int32_t q0[kPfMaxMerge];      // first query row of the 256-row entry
  int32_t slot0[kPfMaxMerge];   // workspace slot of (head 0, part 0); slot = slot0 + head * nparts + part
  int32_t nparts[kPfMaxMerge];
};

// grid (head, entry, group * chunks + chunk): one CTA merges 8 columns of one 128-row group, thread = row
template <typename T>
__global__ void __launch_bounds__(128)
prefill_merge_kernel(const __grid_constant__ MergeTable t, const float* __restrict__ ws, void* __restrict__ out,
                     int64_t out_stride, int head_size, int dp) {
  const int chunks = head_size / 8;
  const int head = blockIdx.x, e = blockIdx.y, grp = blockIdx.z / chunks, c0 = (blockIdx.z % chunks) * 8;
  const int r = threadIdx.x;
  const int row = t.q0[e] + 128 * grp + r;          // row of the sequence this thread owns
  if (t.q0[e] + 128 * grp >= t.qlen[e]) return;     // the entry has no second group
  const int np = t.nparts[e];
  const int64_t pf = pf_part_floats(dp);
  const float* base = ws + (static_cast<int64_t>(t.slot0[e] + head * np) * 2 + grp) * pf;
  float gm = -CUDART_INF_F;
  for (int p = 0; p < np; ++p) gm = fmaxf(gm, __ldg(base + p * 2 * pf + 128 * dp + r));
  float l = 0.f;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int p = 0; p < np; ++p) {
    const float* pb = base + p * 2 * pf;
    const float m = __ldg(pb + 128 * dp + r);
    const float w = m == -CUDART_INF_F ? 0.f : ex2(m - gm);
    l = fmaf(w, __ldg(pb + 128 * dp + 128 + r), l);
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = fmaf(w, __ldg(pb + (c0 + i) * 128 + r), acc[i]);   // 128 contiguous bytes per warp
  }
  if (row < t.qlen[e]) {
    const float inv = 1.f / l;
    uint4 w4;
    w4.x = pack2<T>(acc[0] * inv, acc[1] * inv); w4.y = pack2<T>(acc[2] * inv, acc[3] * inv);
    w4.z = pack2<T>(acc[4] * inv, acc[5] * inv); w4.w = pack2<T>(acc[6] * inv, acc[7] * inv);
    char* orow = static_cast<char*>(out) + ((static_cast<int64_t>(t.qtok0[e]) + row) * out_stride +
                                            static_cast<int64_t>(head) * head_size) * 2;
    *reinterpret_cast<uint4*>(orow + c0 * 2) = w4;
  }
}

// scratch for the parts, per (device, stream) like the decode kernel's; a request that would have to grow while
// the stream is being captured is refused (the caller then simply does not split)
struct SplitScratch { void* ptr = nullptr; size_t bytes = 0; bool in_graph = false; };
static std::mutex g_split_mu;
static std::map<std::pair<int, cudaStream_t>, SplitScratch> g_split_ws;
static std::vector<std::pair<int, void*>> g_split_retired;   // outgrown buffers a captured graph may still replay into

static float* split_scratch(cudaStream_t stream, size_t bytes) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
  std::lock_guard<std::mutex> lock(g_split_mu);
  SplitScratch& w = g_split_ws[std::make_pair(dev, stream)];
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(stream, &cap) != cudaSuccess) return nullptr;
  if (w.bytes < bytes) {
    if (cap != cudaStreamCaptureStatusNone) return nullptr;
    if (w.ptr) {
      if (cudaStreamSynchronize(stream) != cudaSuccess) return nullptr;
      if (w.in_graph) g_split_retired.emplace_back(dev, w.ptr);
      else cudaFree(w.ptr);
      w = SplitScratch();
    }
    const size_t want = std::max<size_t>(bytes, size_t{16} << 20);
    if (cudaMalloc(&w.ptr, want) != cudaSuccess) { cudaGetLastError(); w = SplitScratch(); return nullptr; }
    w.bytes = want;
  }
  if (cap != cudaStreamCaptureStatusNone) w.in_graph = true;   // a graph will hold this pointer
  return static_cast<float*>(w.ptr);
}

void release_prefill_scratch(int dev) {             // pa_release_workspaces
  std::lock_guard<std::mutex> lock(g_split_mu);
  for (auto it = g_split_retired.begin(); it != g_split_retired.end();) {
    if (it->first == dev) {
      cudaFree(it->second);
      it = g_split_retired.erase(it);
    } else {
      ++it;
    }
  }
  for (auto it = g_split_ws.begin(); it != g_split_ws.end();) {
    if (it->first.first == dev) {
      if (it->second.ptr) cudaFree(it->second.ptr);
      it = g_split_ws.erase(it);
    } else {
      ++it;
    }
  }
}

static int device_sms() {
  static int sms = [] {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
      n = 148;
    return n > 0 ? n : 148;
  }();
  return sms;
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

// One entry of the work list: 256 (16-bit) or 64 (fp32) query rows of a sequence against K/V tiles [kt0, kt1)
struct Tile { int32_t qtok0, qlen, q0, ktok0, klen, span, kt0, kt1, slot; };
struct MergeEntry { int32_t qtok0, qlen, q0, slot0, nparts; };

// Which entries to split over their context, decided on a model of the launch: the items (entry x head) go, in
// the grid's own order, to whichever SM is free first; an item costs its K/V tiles plus ~2 tile periods of fixed
// work (the softmax chain filling, masked tiles, epilogue -- profiles/r2_prefill_timeline.txt; measured on 480
// one-group items of 11 tiles: 1.84 us per tile, ~3 us per item), a part half a period more for its partial
// result, and a launch that splits anything pays the merge kernel (~6 periods).  One knob, the same for every
// entry: the longest part allowed, g tiles; an entry longer than g is cut into equal parts of at most g.  The g with
// the shortest modelled makespan wins if it beats the unsplit launch by 6 %.  Large launches balance themselves
// and are left alone.  Returns the workspace (nullptr: nothing is split); entries of *all are replaced by their
// parts, re-sorted.
static float* plan_splits(std::vector<Tile>* all, std::vector<MergeEntry>* merges, int num_heads, int dp,
                          bool tile_major, cudaStream_t stream) {
  // read per call (tests switch it): PA_PREFILL_SPLIT=0 keeps every entry whole
  const char* env = getenv("PA_PREFILL_SPLIT");
  const bool off = env && env[0] == '0';
  static const bool verbose = [] { const char* e = getenv("PA_PREFILL_SPLIT_DEBUG"); return e && e[0] == '1'; }();
  const int sms = device_sms();
  const size_t n = all->size();
  if (off || n * num_heads > static_cast<size_t>(8 * sms) || n > 512) return nullptr;
  constexpr double kFixed = 2.0, kPartExtra = 0.5, kMerge = 6.0;   // in K/V tile periods
  constexpr int kMinPart = 3;                       // tiles
  struct Ent { int nfull, first_diag, parts; };
  std::vector<Ent> ents(n);
  int longest = 0;
  for (size_t i = 0; i < n; ++i) {
    const Tile& t = (*all)[i];
    const int koff = t.klen - t.qlen;
    const int last_row = std::min(t.q0 + kPfRows, t.qlen) - 1;
    ents[i].nfull = std::min(last_row + koff, t.klen - 1) / kPfCols + 1;
    ents[i].first_diag = (t.q0 + koff) / kPfCols;   // the last part must start at or before this tile
    ents[i].parts = 1;
    longest = std::max(longest, ents[i].nfull);
  }
  // part boundaries of an entry cut into k parts: the last part starts at `last`, the others share [0, last) evenly
  auto last_start = [&](const Ent& e, int k) { return std::min(e.first_diag, (k - 1) * e.nfull / k); };
  auto can_split = [&](const Ent& e, int k) {
    if (k > 64) return false;
    const int last = last_start(e, k);
    return last / (k - 1) >= kMinPart && e.nfull - last >= kMinPart;
  };
  auto part_begin = [&](const Ent& e, int k, int i) { return k == 1 ? 0 : (i == k - 1 ? last_start(e, k) : i * last_start(e, k) / (k - 1)); };
  if (longest < 2 * kMinPart) return nullptr;       // nothing to decide for batches of short prompts
  // Modelled makespan of the current ents[].parts, items dispatched in the grid's order.  Costs are counted in half
  // periods; the SMs' free times are kept as a short ascending list of (time, SMs) and every run of equal-cost
  // items is handed out as a batch, so one evaluation costs a few hundred list operations, not one heap operation
  // per item.
  std::vector<int> costs;
  std::vector<std::pair<int, int>> runs, sm;       // (cost, items), (free at, SMs)
  auto simulate = [&]() {
    costs.clear();
    bool any = false;
    for (size_t i = 0; i < n; ++i)
      for (int k = 0; k < ents[i].parts; ++k) {
        const int b = part_begin(ents[i], ents[i].parts, k);
        const int e = k + 1 < ents[i].parts ? part_begin(ents[i], ents[i].parts, k + 1) : ents[i].nfull;
        costs.push_back(2 * (e - b) + static_cast<int>(2 * kFixed) + (ents[i].parts > 1 ? static_cast<int>(2 * kPartExtra) : 0));
        any |= ents[i].parts > 1;
      }
    std::sort(costs.begin(), costs.end(), std::greater<int>());   // the launch sorts by span
    runs.clear();
    for (int c : costs) {
      if (!runs.empty() && runs.back().first == c) ++runs.back().second;
      else runs.emplace_back(c, 1);
    }
    sm.assign(1, std::make_pair(0, sms));
    int makespan = 0;
    auto put = [&](int c, int m) {
      while (m > 0) {
        const int t = sm.front().first, x = std::min(sm.front().second, m);
        if (x == sm.front().second) sm.erase(sm.begin());
        else sm.front().second -= x;
        const int nt = t + c;
        makespan = std::max(makespan, nt);
        auto it = sm.end();
        while (it != sm.begin() && (it - 1)->first > nt) --it;
        if (it != sm.begin() && (it - 1)->first == nt) (it - 1)->second += x;
        else sm.insert(it, std::make_pair(nt, x));
        m -= x;
      }
    };
    if (tile_major) {
      for (int h = 0; h < num_heads; ++h)
        for (const auto& r : runs) put(r.first, r.second);
    } else {
      for (const auto& r : runs) put(r.first, r.second * num_heads);
    }
    return 0.5 * makespan + (any ? kMerge : 0.0);
  };
  const double unsplit = simulate();
  double best = unsplit;
  int best_g = 0;
  static const int kCuts[] = {2, 3, 4, 6, 8, 12, 16, 24};
  int prev_g = 0, worse = 0;
  for (int cut : kCuts) {
    const int g = std::max((longest + cut - 1) / cut, kMinPart);
    if (g == prev_g) continue;
    prev_g = g;
    size_t pieces = 0;
    for (size_t i = 0; i < n; ++i) {
      int k = (ents[i].nfull + g - 1) / g;
      while (k > 1 && !can_split(ents[i], k)) --k;
      ents[i].parts = k;
      pieces += k;
    }
    if (pieces * num_heads > static_cast<size_t>(24 * sms) || pieces > kPfMaxTiles) break;
    const double m = simulate();
    if (m < best) { best = m; best_g = g; worse = 0; }
    else if (++worse == 2) break;                  // finer parts only add fixed costs from here on
  }
  if (best_g == 0 || best > 0.94 * unsplit) return nullptr;
  int64_t slots = 0;
  for (size_t i = 0; i < n; ++i) {
    int k = (ents[i].nfull + best_g - 1) / best_g;
    while (k > 1 && !can_split(ents[i], k)) --k;
    ents[i].parts = k;
    if (k > 1) slots += static_cast<int64_t>(k) * num_heads;
  }
  const int64_t bytes = slots * 2 * pf_part_floats(dp) * 4;
  if (slots == 0 || slots >= (1 << 23) || bytes > (int64_t{1} << 30)) return nullptr;
  float* ws = split_scratch(stream, static_cast<size_t>(bytes));
  if (!ws && !verbose) return nullptr;
  std::vector<Tile> out;
  int32_t slot0 = 0;
  for (size_t i = 0; i < n; ++i) {
    const Tile& t = (*all)[i];
    const int k = ents[i].parts;
    if (k == 1) { out.push_back(t); continue; }
    for (int j = 0; j < k; ++j) {
      Tile pt = t;
      pt.kt0 = part_begin(ents[i], k, j);
      pt.kt1 = j + 1 < k ? part_begin(ents[i], k, j + 1) : 0xffff;
      pt.span = (std::min(pt.kt1, ents[i].nfull) - pt.kt0) * kPfCols;
      pt.slot = ((slot0 + j) << 8) | k;
      out.push_back(pt);
    }
    merges->push_back({t.qtok0, t.qlen, t.q0, slot0, k});
    slot0 += k * num_heads;
  }
  std::stable_sort(out.begin(), out.end(), [](const Tile& a, const Tile& b) { return a.span > b.span; });
  if (verbose) {
    fprintf(stderr, "[plan_splits] %zu entries x %d heads -> %zu (parts of at most %d tiles), modelled makespan %.1f vs %.1f unsplit:",
            n, num_heads, out.size(), best_g, best, unsplit);
    for (const Tile& t : out) fprintf(stderr, " q0=%d[%d,%d)%s", t.q0, t.kt0, t.kt1 == 0xffff ? -1 : t.kt1, t.slot < 0 ? "" : "*");
    fprintf(stderr, "\n");
  }
  if (!ws) { merges->clear(); return nullptr; }     // planner debugging without a GPU
  all->swap(out);
  return ws;
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
                     std::min(q0 + rows + (klen - len), klen), 0, 0xffff, -1});
    tok += len;
    ktok += klen;
  }
  if (all.empty()) return PA_OK;
  std::stable_sort(all.begin(), all.end(), [](const Tile& a, const Tile& b) { return a.span > b.span; });
  float* split_ws = nullptr;
  std::vector<MergeEntry> merges;
  // K/V bytes of the batch against the L2 (126 MB on B200): beyond it, query tiles of one head run together
  const bool tile_major = ktok * num_heads * head_size * 4 > (int64_t{96} << 20);
  if (dtype != PA_F32) split_ws = plan_splits(&all, &merges, num_heads, head_size <= 64 ? 64 : 128, tile_major, stream);

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
  p.tile_major = tile_major;
  p.split_ws = split_ws;
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
      tiles.krange[i] = t.kt0 | (t.kt1 << 16); tiles.slot[i] = t.slot;
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
  // the parts of split entries -> out
  static thread_local MergeTable mt;
  for (size_t begin = 0; begin < merges.size(); begin += kPfMaxMerge) {
    const int n = static_cast<int>(std::min<size_t>(kPfMaxMerge, merges.size() - begin));
    bool second_group = false;
    for (int i = 0; i < n; ++i) {
      const MergeEntry& m = merges[begin + i];
      mt.qtok0[i] = m.qtok0; mt.qlen[i] = m.qlen; mt.q0[i] = m.q0; mt.slot0[i] = m.slot0; mt.nparts[i] = m.nparts;
      second_group |= m.q0 + 128 < m.qlen;
    }
    const dim3 grid(num_heads, n, (second_group ? 2 : 1) * (head_size / 8));
    if (dtype == PA_F16) prefill_merge_kernel<F16><<<grid, 128, 0, stream>>>(mt, split_ws, out, out_stride, head_size, dp);
    else prefill_merge_kernel<BF16><<<grid, 128, 0, stream>>>(mt, split_ws, out, out_stride, head_size, dp);
    count_launch();
    const int rc = check_cuda(cudaGetLastError(), "prefill_merge_kernel launch");
    if (rc != PA_OK) return rc;
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
