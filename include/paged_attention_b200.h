/*
 * paged_attention_b200.h -- C ABI of libpaged_b200.so
 *
 * B200 (sm_100a) implementation of vLLM v0.1.0's PagedAttention decode path:
 * the decode-attention operator and the KV-cache movers.  Every entry point
 * below replaces one pybind11-bound function of the reference; the citation
 * after "replaces:" is the reference interface it is a drop-in for
 * (paths relative to the reference tree).
 *
 * Conventions
 *   - plain pointers and sizes only; no torch / ATen types;
 *   - all device pointers belong to the CUDA device that is current when the
 *     call is made; work is enqueued on `stream` (a cudaStream_t passed as
 *     void*; NULL = the legacy default stream) and the call returns without
 *     synchronising the host;
 *   - every function returns PA_OK (0) or a negative PA_ERR_* code;
 *     pa_last_error() then returns a thread-local, NUL-terminated message;
 *   - the caller owns every buffer.  The only state kept by the library is
 *     per-(device, stream) scratch for split-context partial results of the
 *     decode and the prompt kernels (see pa_release_workspaces).
 *
 * Data layout (reference: vllm/worker/cache_engine.py:48-63,
 * csrc/cache_kernels.cu:165-178), with x = 16 / sizeof(T):
 *   key_cache   [num_blocks][num_heads][head_size / x][block_size][x]
 *   value_cache [num_blocks][num_heads][head_size][block_size]
 * A "page" is the contiguous (block, head) slice of either cache:
 * head_size * block_size * sizeof(T) bytes.
 */
#ifndef PAGED_ATTENTION_B200_H_
#define PAGED_ATTENTION_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* 2: pa_decode_attention_reserve, pa_decode_attention_static, pa_reshape_and_cache_ex; the prompt kernels keep
 * library-owned scratch.  Bindings compare pa_abi_version() with the value they were built against. */
#define PA_ABI_VERSION 2

#if defined(__GNUC__) || defined(__clang__)
#define PA_API __attribute__((visibility("default")))
#else
#define PA_API
#endif

/* element type of q / k / v / out and both caches */
enum pa_dtype {
  PA_F16 = 0,  /* IEEE binary16  (torch.half)     */
  PA_BF16 = 1, /* bfloat16       (torch.bfloat16) */
  PA_F32 = 2   /* IEEE binary32  (torch.float)    */
};

/* direction of pa_swap_blocks */
enum pa_copy_kind {
  PA_COPY_D2D = 0, /* both caches on the current device */
  PA_COPY_D2H = 1, /* swap-out: GPU cache -> host cache */
  PA_COPY_H2D = 2  /* swap-in : host cache -> GPU cache */
};

enum pa_status {
  PA_OK = 0,
  PA_ERR_UNSUPPORTED_HEAD_SIZE = -1,  /* head_size not in {64, 80, 96, 128}  */
  PA_ERR_UNSUPPORTED_BLOCK_SIZE = -2, /* block_size not in {8, 16, 32}       */
  PA_ERR_UNSUPPORTED_DTYPE = -3,
  PA_ERR_INVALID_ARGUMENT = -4, /* NULL pointer, negative size, misalignment */
  PA_ERR_CUDA = -5,             /* a CUDA runtime call failed                */
  PA_ERR_NEED_TABLE = -6        /* pa_step_prepare_inputs: a sequence needs its full block table */
};

/* ------------------------------------------------------------------------ */
/* library                                                                  */
/* ------------------------------------------------------------------------ */

PA_API int pa_abi_version(void);

/* Message of the last failed call made by the calling thread ("" if none). */
PA_API const char* pa_last_error(void);

/* Compute capability the kernels were compiled for (100 -> sm_100a). */
PA_API int pa_compiled_sm(void);

/* Number of kernels launched by this library since load (all threads).
 * bench.py reads it before/after the timed region ("gpu_launches"). */
PA_API uint64_t pa_launch_count(void);

/* Frees every split-context scratch buffer held for the current device.
 * Synchronises the device.  Safe to call at any time. */
PA_API int pa_release_workspaces(void);

/* Measurement aid, not an operator: streams `bytes` of device memory through
 * shared memory with the decode kernel's TMA bulk-copy pattern (8 KiB chunks,
 * 4 warps x 2 stages per CTA, ctas_per_sm CTAs per SM; <= 0 -> 3) and no
 * arithmetic.  Timing it gives the read-bandwidth ceiling of that pattern.
 * `sink` is a 4-byte device scratch word. */
PA_API int pa_stream_read_probe(const void* src, int64_t bytes, int32_t ctas_per_sm,
                         void* sink, void* stream);

/* ------------------------------------------------------------------------ */
/* decode attention                                                         */
/* ------------------------------------------------------------------------ */

/*
 * replaces: attention_ops.single_query_cached_kv_attention
 *           csrc/attention.cpp:3-19, csrc/attention/attention_kernels.cu:450-469
 *           (launcher :332-401, kernel :67-314)
 *
 * For every sequence i < num_seqs and head h < num_heads:
 *   out[i][h][:] = softmax(scale * q[i][h][:] . K[i][h]^T) . V[i][h]
 * where K/V[i][h] are the first context_lens[i] tokens reached through
 * block_tables[i][token / block_size].
 *
 *   out            [num_seqs][num_heads][head_size]          T, contiguous
 *   query          [num_seqs][num_heads][head_size]          T, row stride
 *                  q_stride elements (query.stride(0)), inner dims contiguous
 *   key_cache      layout above                              T
 *   value_cache    layout above                              T
 *   block_tables   [num_seqs][max_num_blocks_per_seq]        int32
 *   context_lens   [num_seqs]                                int32
 *   max_context_len upper bound of context_lens (sizes the split-context
 *                  schedule; the reference used it to size shared memory).
 *                  REQUIRED: max_context_len >= max(context_lens), as
 *                  worker.py:191 computes it.  The page walk is capped by the
 *                  schedule derived from it, so a sequence longer than
 *                  max_context_len would lose its tail tokens without an error
 *                  (context_lens lives on the device; the host cannot check --
 *                  unless PA_DEBUG_CHECKS=1 is set in the environment, which reads
 *                  context_lens back on every call, synchronising the stream, and
 *                  returns PA_ERR_INVALID_ARGUMENT instead).
 *
 * head_size in {64, 80, 96, 128}, block_size in {8, 16, 32}, any pa_dtype --
 * the reference's dispatch set (attention_kernels.cu:370-400, 416-448,
 * 460-468); anything else returns the matching PA_ERR_UNSUPPORTED_* code
 * (the reference raises through TORCH_CHECK).
 *
 * Differences from the reference kernel, all deliberate:
 *   - 64-bit element offsets (num_blocks*num_heads*head_size*block_size may
 *     exceed 2^31; the reference overflows there);
 *   - tokens >= context_len in the last block are masked, never multiplied,
 *     so non-finite garbage in unwritten cache slots cannot reach the output;
 *   - no limit on context length from shared memory;
 *   - context_len == 0 yields an all-zero output row.
 * key_cache, value_cache, query rows and out must be 16-byte aligned.
 */
PA_API int pa_decode_attention(void* out, const void* query, const void* key_cache,
                        const void* value_cache, float scale,
                        const int32_t* block_tables,
                        const int32_t* context_lens, int32_t num_seqs,
                        int32_t num_heads, int32_t head_size,
                        int32_t block_size, int32_t max_num_blocks_per_seq,
                        int32_t max_context_len, int64_t q_stride,
                        int32_t dtype, void* stream);

/* Launch-shape overrides for tuning and tests; 0 / negative = automatic. */
typedef struct pa_decode_tuning {
  int32_t variant;    /* 0 auto, 1 direct global loads, 2 bulk-copy staged (sibling groups
                         served together), 3 bulk-copy staged without sibling groups */
  int32_t num_warps;  /* warps working together on one (seq, head, split)   */
  int32_t num_stages; /* pages in flight per warp (variants 2, 3)            */
  int32_t num_splits; /* context partitions per (seq, head)                  */
} pa_decode_tuning;

/* Same operator with explicit launch shape (tuning == NULL -> automatic). */
PA_API int pa_decode_attention_ex(void* out, const void* query, const void* key_cache,
                           const void* value_cache, float scale,
                           const int32_t* block_tables,
                           const int32_t* context_lens, int32_t num_seqs,
                           int32_t num_heads, int32_t head_size,
                           int32_t block_size, int32_t max_num_blocks_per_seq,
                           int32_t max_context_len, int64_t q_stride,
                           int32_t dtype, const pa_decode_tuning* tuning,
                           void* stream);

/* Sizes the per-(device, stream) split-context scratch for this shape without launching anything, so a
 * later pa_decode_attention call of that shape never allocates -- required before capturing the call into
 * a CUDA graph (growing the scratch synchronises the stream, which capture forbids; the call then fails
 * with PA_ERR_INVALID_ARGUMENT instead). */
PA_API int pa_decode_attention_reserve(int32_t num_seqs, int32_t num_heads,
                                int32_t head_size, int32_t block_size,
                                int32_t max_num_blocks_per_seq,
                                int32_t max_context_len, int32_t dtype,
                                const pa_decode_tuning* tuning, void* stream);

/* The launch shape pa_decode_attention would pick for these sizes (plan_out->variant selects the
 * kernel family on entry; 0 = automatic). */
PA_API int pa_decode_attention_plan(int32_t num_seqs, int32_t num_heads,
                             int32_t head_size, int32_t block_size,
                             int32_t max_context_len, int32_t dtype,
                             pa_decode_tuning* plan_out);

/*
 * No counterpart in the reference: this entry point exists for hosts that are
 * not allowed to allocate device memory.
 *
 * pa_decode_attention keeps the partial results of a context that is split
 * over several CTAs in a buffer it cudaMalloc's on first use and grows later.
 * Some callers cannot live with that: an engine that hands the whole GPU to
 * its own pool at start-up and leaves nothing for a driver allocation to
 * succeed from, a plug-in running under a host whose rules forbid cudaMalloc
 * after initialisation, a latency-critical loop that cannot take the implicit
 * device synchronisation an allocation or a free may bring.  For them this
 * variant computes the same operator (same arguments and results as
 * pa_decode_attention_ex) but keeps those partial results in a fixed array
 * compiled into the library.  It never calls cudaMalloc, cudaFree or any
 * routine that allocates; apart from the kernel launch it issues one
 * cudaMemsetAsync on `stream`.
 *
 * Limits.  The array is PA_STATIC_SCRATCH_BYTES long and every GPU that loads
 * the library holds one copy of it until the process exits.  When the launch
 * shape splits the context (num_splits > 1 in the plan) the call needs
 *   num_seqs * num_heads                                  <= PA_STATIC_SCRATCH_MAX_ITEMS
 *   num_seqs * num_heads * num_splits * (head_size+2) * 4 <= PA_STATIC_SCRATCH_BYTES
 *                                                            - 4 * PA_STATIC_SCRATCH_MAX_ITEMS
 * and returns PA_ERR_INVALID_ARGUMENT, with the sizes in pa_last_error(), when
 * it does not fit; nothing is launched then.  Automatic plans stay below
 * 148 * 32 (item, split) pairs and always fit; a forced num_splits may not --
 * lower it or use pa_decode_attention.  Shapes that do not split use no scratch
 * and have no limit.
 *
 * Sharing.  There is ONE array per device, whatever the stream.  The call
 * clears the part that has to start from zero, so whatever an earlier call
 * left behind does no harm, but two calls that can be on the GPU at the same
 * time (different streams, or one stream from two host threads without
 * ordering) overwrite each other's partial results: keep calls to this entry
 * point ordered one after the other on each device.  pa_decode_attention and
 * the other entry points use their own buffers and may overlap with it.
 *
 * First call.  The first call on each device looks up the array's device
 * address (with that device current) and remembers it.  Make that first call
 * ahead of any CUDA stream capture on the device.
 */
#define PA_STATIC_SCRATCH_BYTES (4 * 1024 * 1024)
#define PA_STATIC_SCRATCH_MAX_ITEMS 16384
PA_API int pa_decode_attention_static(void* out, const void* query, const void* key_cache,
                               const void* value_cache, float scale,
                               const int32_t* block_tables,
                               const int32_t* context_lens, int32_t num_seqs,
                               int32_t num_heads, int32_t head_size,
                               int32_t block_size, int32_t max_num_blocks_per_seq,
                               int32_t max_context_len, int64_t q_stride,
                               int32_t dtype, const pa_decode_tuning* tuning,
                               void* stream);

/* ------------------------------------------------------------------------ */
/* prompt (prefill) attention                                               */
/* ------------------------------------------------------------------------ */

/*
 * replaces: PagedAttention.multi_query_kv_attention
 *           vllm/model_executor/layers/attention.py:50-70, i.e. xformers
 *           memory_efficient_attention_forward with
 *           BlockDiagonalCausalMask.from_seqlens(prompt_lens)
 *           (vllm/model_executor/input_metadata.py:33); parity reference
 *           tests/kernels/test_attention.py:69-101 (ref_multi_query_kv_attention)
 *
 * Causal self-attention over num_seqs prompts packed along the token dimension
 * (prompt i occupies tokens [sum(seq_lens[:i]), sum(seq_lens[:i+1]))):
 *   out[t][h][:] = softmax_{u <= t, same prompt}(scale * q[t][h] . k[u][h]) . v[u][h]
 * query / key / value / out are [num_tokens][num_heads][head_size] views with
 * token strides *_stride (elements; 3*num_heads*head_size when sliced from the
 * fused qkv tensor), inner dimensions contiguous, rows 16-byte aligned.
 * seq_lens is a HOST array (the reference holds prompt_lens as a Python list).
 * head_size in {64, 80, 96, 128}.  PA_F16 / PA_BF16 run on the tcgen05 tensor
 * cores (TMA-fed, accumulators in tensor memory); PA_F32 takes an exact fp32
 * SIMT path.  Tokens past sum(seq_lens) (the padding the model adds) are not
 * read or written.
 */
PA_API int pa_prefill_attention(void* out, const void* query, const void* key,
                         const void* value, const int32_t* seq_lens,
                         int32_t num_seqs, int32_t num_heads, int32_t head_size,
                         int64_t q_stride, int64_t k_stride, int64_t v_stride,
                         int64_t out_stride, float scale, int32_t dtype,
                         void* stream);

/* Same, and (16-bit types) CTA (head 0, first tile) also dumps to `debug` (device, floats):
 * [128][128] raw q.k of its first K/V tile, [128][64 or 128] unnormalised output rows,
 * [128] row sums, [128] row maxima (log2 domain).  Tests only. */
PA_API int pa_prefill_attention_debug(void* out, const void* query, const void* key,
                               const void* value, const int32_t* seq_lens,
                               int32_t num_seqs, int32_t num_heads, int32_t head_size,
                               int64_t q_stride, int64_t k_stride, int64_t v_stride,
                               int64_t out_stride, float scale, int32_t dtype,
                               float* debug, void* stream);

/*
 * Prompt chunks over the paged cache ("multi-query cached-KV attention"): sequence i
 * has query_lens[i] new tokens whose keys / values already sit in the cache behind
 * context_lens[i] - query_lens[i] earlier tokens (a cached prefix, or the earlier
 * chunks of a long prompt); query row r of the sequence attends to cached tokens
 * 0 .. r + (context_lens[i] - query_lens[i]).
 *
 * parity reference: ref_multi_query_cached_kv_attention,
 *           tests/kernels/test_attention.py:104-156 (the reference tree holds this
 *           test reference but no kernel for it; the operator name and argument order
 *           of the Python shim follow attention_ops.single_query_cached_kv_attention,
 *           csrc/attention.cpp:3-19)
 *
 * query / out: [sum(query_lens)][num_heads][head_size] with token strides q_stride /
 * out_stride (elements), rows 16-byte aligned.  key_cache / value_cache / block_tables
 * (DEVICE, int32 [num_seqs][max_num_blocks_per_seq]) as for pa_decode_attention.
 * query_lens and context_lens are HOST arrays (the reference keeps cu_query_lens as a
 * Python list).  head_size in {64, 80, 96, 128}, block_size in {8, 16, 32}.
 * PA_F16 / PA_BF16: the tcgen05 prompt kernel reads the pages itself -- a producer warp issues
 * TMA loads through tensor maps over the caches (a K page is a set of no-swizzle core matrices,
 * a V page a K-major operand under 32- / 64-byte swizzle); tokens past the context are zeroed in
 * shared memory.  No caller scratch: workspace may be NULL and
 * pa_multi_query_cached_kv_workspace_bytes returns 0.  One launch, or two when long items of a
 * small batch are split over their context (parts + a merge kernel; the partial results live in
 * the library's own per-(device, stream) scratch, see pa_release_workspaces; PA_PREFILL_SPLIT=0
 * in the environment, read per call, keeps every item whole and the summation order fixed).
 * PA_F32 (a test-matrix case): slot generation + page gather into `workspace` (DEVICE scratch,
 * 256-byte aligned, at least pa_multi_query_cached_kv_workspace_bytes(sum(context_lens), ...)
 * bytes, owned by the caller, free for reuse once the stream has passed this call) + the exact
 * fp32 SIMT prompt kernel.
 */
PA_API int64_t pa_multi_query_cached_kv_workspace_bytes(int64_t total_context_tokens,
                                                 int32_t num_heads, int32_t head_size,
                                                 int32_t dtype);

PA_API int pa_multi_query_cached_kv_attention(void* out, const void* query,
                                       const void* key_cache, const void* value_cache,
                                       float scale, const int32_t* block_tables,
                                       const int32_t* query_lens,
                                       const int32_t* context_lens, int32_t num_seqs,
                                       int32_t num_heads, int32_t head_size,
                                       int32_t block_size, int32_t max_num_blocks_per_seq,
                                       int64_t q_stride, int64_t out_stride, int32_t dtype,
                                       void* workspace, int64_t workspace_bytes,
                                       void* stream);

/* ------------------------------------------------------------------------ */
/* KV-cache movers (bit-exact)                                              */
/* ------------------------------------------------------------------------ */

/*
 * replaces: cache_ops.reshape_and_cache
 *           csrc/cache.cpp:17-22, csrc/cache_kernels.cu:186-224 (kernel :142-182)
 *
 * Scatters the new tokens' key/value rows into the paged caches:
 *   slot = slot_mapping[t]; blk = slot / block_size; off = slot % block_size
 *   key_cache  [blk][h][d / x][off][d % x] = key  [t][h][d]
 *   value_cache[blk][h][d][off]            = value[t][h][d]
 * key/value are [num_tokens][num_heads][head_size] with row strides
 * key_stride / value_stride (elements).  A negative slot skips the token
 * (the reference has no such case and would write out of bounds).
 */
PA_API int pa_reshape_and_cache(const void* key, const void* value, void* key_cache,
                         void* value_cache, const int32_t* slot_mapping,
                         int32_t num_tokens, int32_t num_heads,
                         int32_t head_size, int32_t block_size,
                         int64_t key_stride, int64_t value_stride,
                         int32_t dtype, void* stream);

/* What the caller knows about slot_mapping (it lives on the device, the library cannot look): a performance hint
 * only, the result is the same whatever is passed. */
enum pa_slot_pattern {
  PA_SLOTS_AUTO = 0,      /* nothing known */
  PA_SLOTS_SCATTERED = 1, /* every token in a block of its own: a decode step (worker.py:170-199)    */
  PA_SLOTS_RUNS = 2       /* consecutive slots: prompt tokens (worker.py:138-167), or a mixed batch  */
};

/* pa_reshape_and_cache with that hint.  A decode step of ~100..512 sequences is served by a
 * one-CTA-per-token element sweep when the caller declares PA_SLOTS_SCATTERED (about 15 % faster than the tile
 * mover there); everything else, and every call without a hint, goes through the tile mover, which merges
 * the tokens of a block into whole sectors.
 * replaces: the same operator, cache_ops.reshape_and_cache (csrc/cache.cpp:17-22); the hint stands for what the
 *           call site knows from InputMetadata.num_prompt_tokens (vllm/model_executor/layers/attention.py:129-140). */
PA_API int pa_reshape_and_cache_ex(const void* key, const void* value, void* key_cache,
                            void* value_cache, const int32_t* slot_mapping,
                            int32_t num_tokens, int32_t num_heads,
                            int32_t head_size, int32_t block_size,
                            int64_t key_stride, int64_t value_stride,
                            int32_t dtype, int32_t slot_pattern, void* stream);

/*
 * replaces: cache_ops.gather_cached_kv
 *           csrc/cache.cpp:24-29, csrc/cache_kernels.cu:348-386 (kernels :229-344)
 * Inverse of pa_reshape_and_cache: key/value rows <- cache at slot_mapping.
 */
PA_API int pa_gather_cached_kv(void* key, void* value, const void* key_cache,
                        const void* value_cache, const int32_t* slot_mapping,
                        int32_t num_tokens, int32_t num_heads,
                        int32_t head_size, int32_t block_size,
                        int64_t key_stride, int64_t value_stride,
                        int32_t dtype, void* stream);

/*
 * replaces: pos_encoding_ops.rotary_embedding_neox
 *           csrc/pos_encoding.cpp:3-15, csrc/pos_encoding_kernels.cu:5-83
 *           (the operator that runs immediately before this path:
 *           vllm/model_executor/layers/attention.py:200-206)
 *
 * GPT-NeoX rotary embedding in place on query and key
 * ([num_tokens][num_heads][head_size], row strides in elements): for i < rot_dim/2
 *   (x, y) = (row[i], row[rot_dim/2 + i]);  c = cache[pos][i];  s = cache[pos][rot_dim/2 + i]
 *   row[i] = x*c - y*s;   row[rot_dim/2 + i] = y*c + x*s
 * positions is int64 [num_tokens]; cos_sin_cache is [max_position][rot_dim] of T.
 * Every product and sum is rounded to T as the reference does, so results are
 * bit-identical to the reference's PyTorch implementation of the same formula.
 */
PA_API int pa_rotary_embedding_neox(const int64_t* positions, void* query, void* key,
                             const void* cos_sin_cache, int32_t num_tokens,
                             int32_t num_heads, int32_t head_size, int32_t rot_dim,
                             int64_t query_stride, int64_t key_stride,
                             int32_t dtype, void* stream);

/*
 * Fusion of the two operators either side of the KV write (one launch, one pass
 * over k and v): pa_rotary_embedding_neox on tokens [0, num_rope_tokens) followed by
 * pa_reshape_and_cache of tokens [0, num_cache_tokens) (num_cache_tokens <=
 * num_rope_tokens: the model pads its token batch, the slot mapping is not
 * padded).  query and key are left rotated in place exactly as the unfused
 * sequence leaves them; the caches receive the rotated key and the value.
 * replaces: vllm/model_executor/layers/attention.py:200-206 + :134-140.
 */
PA_API int pa_rope_reshape_and_cache(const int64_t* positions, void* query, void* key,
                              const void* value, void* key_cache, void* value_cache,
                              const int32_t* slot_mapping, const void* cos_sin_cache,
                              int32_t num_rope_tokens, int32_t num_cache_tokens,
                              int32_t num_heads, int32_t head_size,
                              int32_t block_size, int32_t rot_dim,
                              int64_t query_stride, int64_t key_stride,
                              int64_t value_stride, int32_t dtype, void* stream);

/*
 * replaces: cache_ops.copy_blocks
 *           csrc/cache.cpp:11-15, csrc/cache_kernels.cu:82-138 (kernel :52-78)
 *
 * Copy-on-write: for every layer and every pair p, copies block
 * pairs[2p] -> block pairs[2p+1] inside that layer's key cache and inside
 * its value cache.  key_cache_ptrs / value_cache_ptrs are HOST arrays of
 * num_layers device base pointers; pairs is a HOST array of 2*num_pairs
 * block numbers; block_bytes = num_heads*head_size*block_size*sizeof(T).
 * Tables travel as kernel parameters: no staging allocation, no host
 * synchronisation (the reference does three blocking pageable H2D copies).
 * As in the reference, a destination must not be another pair's source.
 */
PA_API int pa_copy_blocks(void* const* key_cache_ptrs, void* const* value_cache_ptrs,
                   int32_t num_layers, const int64_t* pairs, int64_t num_pairs,
                   int64_t block_bytes, void* stream);

/*
 * replaces: cache_ops.swap_blocks
 *           csrc/cache.cpp:6-9, csrc/cache_kernels.cu:9-47
 *
 * dst[dst_blocks[i]] = src[src_blocks[i]] for i < num_blocks, whole blocks of
 * block_bytes.  kind selects device/host sides; host memory may be pageable
 * or pinned (pinned is what vllm/worker/cache_engine.py:88-97 allocates).
 * src_blocks / dst_blocks are HOST arrays.
 * Few contiguous runs of (src, dst) blocks go through the copy engines, one DMA
 * per run; many scattered blocks with pinned host memory (or GPU->GPU) go
 * through one gather/scatter kernel per 240 blocks that accesses the host
 * buffer directly over PCIe.  Both are stream-ordered and asynchronous.
 */
PA_API int pa_swap_blocks(const void* src, void* dst, int64_t block_bytes,
                   const int64_t* src_blocks, const int64_t* dst_blocks,
                   int64_t num_blocks, int32_t kind, void* stream);

/* Same with the engine forced: 0 automatic, 1 copy engines, 2 kernel (falls
 * back to the copy engines when the host side is pageable).  Tests and tuning. */
PA_API int pa_swap_blocks_ex(const void* src, void* dst, int64_t block_bytes,
                      const int64_t* src_blocks, const int64_t* dst_blocks,
                      int64_t num_blocks, int32_t kind, int32_t engine,
                      void* stream);

/* ------------------------------------------------------------------------ */
/* all-reduce of the row-parallel output that follows attention             */
/* ------------------------------------------------------------------------ */

/*
 * replaces: GraphAllReduce.launch
 *           vllm/model_executor/parallel_utils/parallel_state.py:506-556 (NCCL
 *           all_reduce replayed from a CUDA graph), called after every
 *           RowParallelLinear (tensor_parallel/layers.py:419-426) -- the one
 *           collective next to the attention path (head-sharded attention itself
 *           needs none).
 *
 * In-place sum over `world` ranks of numel elements starting offset_bytes into
 * every rank's buffer; all ranks end up with identical bits.  One kernel per
 * rank over NVLink peer memory: rank r reduces slice r (multimem.ld_reduce in
 * the NVSwitch when multicast_ptr is given, else 16-byte peer loads summed in
 * fp32 in rank order) and writes the sum to every rank (multimem.st / peer
 * stores), between two flag barriers.  Stream-ordered, no host synchronisation,
 * capturable in a CUDA graph.
 *   peer_buffers  HOST array [world]: this buffer's address on every rank, mapped
 *                 into the calling process (symmetric memory); [rank] is local
 *   peer_flags    HOST array [world]: pa_allreduce_flag_bytes() bytes on every
 *                 rank, zeroed once before first use (they return to zero)
 *   multicast_ptr multicast mapping of the buffer (NVLS) or NULL
 * All ranks must call with the same offset / numel / dtype.  numel * sizeof(T)
 * and offset_bytes must be multiples of 16; world <= 8.
 */
PA_API int pa_allreduce(void* const* peer_buffers, void* const* peer_flags,
                 void* multicast_ptr, int64_t offset_bytes, int64_t numel,
                 int32_t rank, int32_t world, int32_t dtype, void* stream);

PA_API int pa_allreduce_flag_bytes(void);

/* The peer path with all `world` ranks inside ONE launch on ONE device
 * (buffers[r] / flags[r] are ordinary device allocations): how the kernel is
 * tested where only one GPU is present.  Tests only. */
PA_API int pa_allreduce_emulated(void* const* buffers, void* const* flags,
                          int64_t offset_bytes, int64_t numel, int32_t world,
                          int32_t dtype, void* stream);

/* ------------------------------------------------------------------------ */
/* per-step inputs: device-resident block tables                            */
/* ------------------------------------------------------------------------ */

/*
 * replaces: Worker._prepare_inputs, vllm/worker/worker.py:128-237 (with
 *           _pad_to_alignment / _pad_to_max, :304-309) -- the producer of
 *           InputMetadata's tensors (vllm/model_executor/input_metadata.py:10-40).
 *
 * The reference rebuilds every block table as nested Python lists each step and
 * uploads tokens, positions, slot mapping, context lengths and the padded
 * [S, MB] table with five blocking pageable copies.  Here the block tables LIVE
 * on the device in a pa_step_state; a step sends only what changed (the block a
 * sequence appended or copy-on-wrote, full tables only for sequences that were
 * not in the previous step) in ONE pinned, stream-ordered copy, and ONE kernel
 * patches the resident tables and writes all five tensors exactly as the
 * reference lays them out:
 *   tokens / positions  int64 [pad8(num_prompt_tokens + S)]: prompt tokens of all
 *                       prompts, then one token per generation sequence, then
 *                       zeros up to a multiple of 8 (worker.py:201-204)
 *   slot_mapping        int32 [num_prompt_tokens + S]: table[pos / B] * B + pos % B
 *   context_lens        int32 [S]
 *   block_tables        int32 [S, width], rows zero-padded (worker.py:211-214)
 * Bit-identical to the reference's tensors (tests/golden/step_inputs_trace.json.gz).
 *
 * A state belongs to one worker thread; calls on it must not overlap.  Work is
 * enqueued on `stream`; the host arrays may be reused as soon as the call
 * returns (they are packed into the state's pinned staging ring).
 */
typedef struct pa_step_state pa_step_state;

PA_API int pa_step_state_create(int32_t max_seqs, int32_t max_blocks_per_seq,
                         int32_t block_size, pa_step_state** state_out);
PA_API int pa_step_state_destroy(pa_step_state* state);
/* Forgets every sequence (the next step must bring full tables). */
PA_API int pa_step_state_reset(pa_step_state* state);

typedef struct pa_step_batch {
  /* generation sequences, in batch order (worker.py:170-199) */
  int32_t num_seqs;
  const int64_t* seq_ids;       /* [S] scheduler sequence ids (unique in a step)      */
  const int64_t* last_tokens;   /* [S] seq_data.get_last_token_id()                   */
  const int32_t* context_lens;  /* [S] seq_data.get_len()                             */
  const int32_t* num_blocks;    /* [S] len(block_table)                               */
  const int32_t* last_block;    /* [S] block_table[-1]                                */
  /* Full tables (CSR), for the sequences that bring one: table_offsets[i+1] -
   * table_offsets[i] is 0 or num_blocks[i].  NULL = nobody brings one.  Required
   * for a sequence that was not in the previous step; a sequence that was may
   * omit it if its table only grew by one block or had its last block replaced
   * (BlockSpaceManager.append_slot, core/block_manager.py:111-135) -- anything
   * else (beam-search forks, block_manager.py:137-143) must bring the table,
   * which is diffed against the resident copy and only the changed span is sent. */
  const int64_t* table_offsets; /* [S + 1] or NULL */
  const int32_t* table_blocks;
  /* prompts, in batch order (worker.py:138-167) */
  int32_t num_prompts;
  const int64_t* prompt_offsets;        /* [P + 1] into prompt_tokens                 */
  const int64_t* prompt_tokens;
  const int64_t* prompt_table_offsets;  /* [P + 1] into prompt_table_blocks; an empty */
  const int32_t* prompt_table_blocks;   /* row = "block tables not initialised yet":  */
                                        /* dummy slot 0 (worker.py:160-164)           */
} pa_step_batch;

typedef struct pa_step_outputs {  /* DEVICE pointers, caller-allocated */
  int64_t* tokens;        /* [pad8(num_prompt_tokens + S)] */
  int64_t* positions;     /* same length */
  int32_t* slot_mapping;  /* [num_prompt_tokens + S] */
  int32_t* context_lens;  /* [S] */
  int32_t* block_tables;  /* [S, block_table_width] */
  int32_t block_table_width; /* >= max(num_blocks); the reference uses == */
} pa_step_outputs;

typedef struct pa_step_info {
  int32_t max_context_len;  /* worker.py:191 */
  int32_t max_num_blocks;   /* max(num_blocks), worker.py:192-193 */
  int32_t num_tokens;       /* num_prompt_tokens + S (before padding) */
  int32_t num_incremental;  /* sequences updated without a full table */
  int32_t num_full;         /* sequences whose table was sent whole */
  int32_t num_diffed;       /* sequences that brought a table of which only a span was sent */
  int64_t h2d_bytes;        /* bytes of the one host->device copy of this step */
} pa_step_info;

/* On PA_ERR_NEED_TABLE the state is unchanged and need_table[i] (optional, [S])
 * is 1 for every sequence that must bring its full table; call again. */
PA_API int pa_step_prepare_inputs(pa_step_state* state, const pa_step_batch* batch,
                           const pa_step_outputs* out, pa_step_info* info,
                           uint8_t* need_table, void* stream);

#ifdef __cplusplus
} /* extern "C" */
#endif

#endif /* PAGED_ATTENTION_B200_H_ */
