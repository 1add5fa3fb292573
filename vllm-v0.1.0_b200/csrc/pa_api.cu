// Library plumbing and the decode-attention entry points of libpaged_b200:
// error state, launch accounting, the split-context scratch buffers and the
// launch-shape heuristic.  See include/paged_attention_b200.h for the ABI.
#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <vector>
#include <utility>

#include "decode_attention.cuh"

namespace pa {

// ---------------------------------------------------------------------------
// error state (thread-local) and launch counter
// ---------------------------------------------------------------------------
static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return PA_OK;
  return set_error(PA_ERR_CUDA, "CUDA error in %s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
}

void count_launch(uint64_t n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// ---------------------------------------------------------------------------
// per-(device, stream) scratch for split-context partials.
// Layout: [tickets: items x uint32, padded to 256 B][partials: floats].
// Tickets are zeroed when the buffer is created and re-armed by the kernels; after a
// launch that returned an error they are zeroed again before the next use (an aborted launch may have
// left them counting).  The buffer only ever grows; growing needs a stream synchronisation, so it is
// refused while the stream is being captured into a CUDA graph -- size it first with
// pa_decode_attention_reserve (or one eager call of the same shape).
// ---------------------------------------------------------------------------
struct Workspace {
  void* ptr = nullptr;
  size_t bytes = 0;
  size_t ticket_bytes = 0;
  bool dirty = false;
  bool in_graph = false;   // handed out while the stream was being captured: a CUDA graph holds the pointer
};
static std::mutex g_ws_mu;
static std::map<std::pair<int, cudaStream_t>, Workspace> g_ws;
// buffers that a captured graph may still replay into; kept until pa_release_workspaces instead of being freed when a
// later, larger call outgrows them
static std::vector<std::pair<int, void*>> g_ws_retired;

static int get_workspace(cudaStream_t stream, size_t ticket_bytes, size_t part_bytes, uint32_t** tickets,
                         float** parts) {
  int dev = 0;
  PA_CUDA_TRY(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lock(g_ws_mu);
  Workspace& w = g_ws[std::make_pair(dev, stream)];
  ticket_bytes = (ticket_bytes + 255) / 256 * 256;
  part_bytes = (part_bytes + 255) / 256 * 256;
  // the capture status is asked for at most once per call: when the buffer must grow, or while no capture has used it yet
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  const bool must_grow = w.ticket_bytes < ticket_bytes || w.bytes - w.ticket_bytes < part_bytes;
  if (must_grow || !w.in_graph) PA_CUDA_TRY(cudaStreamIsCapturing(stream, &cap));
  if (must_grow) {
    if (cap != cudaStreamCaptureStatusNone)
      return set_error(PA_ERR_INVALID_ARGUMENT,
                       "decode scratch must grow (%zu B) while the stream is being captured: call "
                       "pa_decode_attention_reserve for this shape before capturing", ticket_bytes + part_bytes);
    if (w.ptr) {
      PA_CUDA_TRY(cudaStreamSynchronize(stream));
      if (w.in_graph) g_ws_retired.emplace_back(dev, w.ptr);
      else PA_CUDA_TRY(cudaFree(w.ptr));
      w = Workspace();
    }
    const size_t tb = ticket_bytes * 2;            // head-room: item count may grow
    const size_t total = tb + part_bytes * 2;
    PA_CUDA_TRY(cudaMalloc(&w.ptr, total));
    w.bytes = total;
    w.ticket_bytes = tb;
    w.dirty = true;
  }
  if (w.dirty) {
    PA_CUDA_TRY(cudaMemsetAsync(w.ptr, 0, w.ticket_bytes, stream));
    w.dirty = false;
  }
  if (cap != cudaStreamCaptureStatusNone) w.in_graph = true;
  *tickets = static_cast<uint32_t*>(w.ptr);
  *parts = reinterpret_cast<float*>(static_cast<char*>(w.ptr) + w.ticket_bytes);
  return PA_OK;
}

// a launch on `stream` failed: its control words / tickets may be left counting
static void poison_workspace(cudaStream_t stream) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return;
  std::lock_guard<std::mutex> lock(g_ws_mu);
  auto it = g_ws.find(std::make_pair(dev, stream));
  if (it != g_ws.end()) it->second.dirty = true;
}

// ---------------------------------------------------------------------------
// launch-shape heuristic
// ---------------------------------------------------------------------------
static int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

// sibling-group detection needs one CTA warp per sequence of a group and room for the group's table slices
// Sibling groups are served together by a second instantiation of the kernel (tensor-core tiles for the
// shared pages: 16-bit types, pages of whole 16-token tiles, one CTA warp per sequence of a group).  It
// needs 128 registers per thread against 80, so it is used only where shared memory, not registers,
// already bounds the resident CTAs -- elsewhere it would cost the plain path occupancy.
constexpr int kMaxSmem = 227 * 1024;
constexpr int kMaxSmemPerSm = 228 * 1024;
constexpr int kGroupKernelRegs = 128;
static int group_rows_for(bool staged, int variant, int warps, int num_seqs, int dtype, int block_size, int smem_bytes) {
  if (!(staged && variant != 3 && warps == kGroup && num_seqs >= 2 && dtype != PA_F32 && block_size % 16 == 0)) return 1;
  const int by_smem = kMaxSmemPerSm / (smem_bytes + 1024);
  const int by_regs = 65536 / (kGroupKernelRegs * warps * kWarp);
  return by_regs >= by_smem ? kGroup : 1;
}

int decode_smem_bytes(int dtype, int head_size, int block_size, int pages_per_split, int warps, int stages,
                      bool staged) {
  const int page = head_size * block_size * dtype_size(dtype);
  const int pbuf = ((2 * block_size * dtype_size(dtype) + 15) / 16) * 16;
  int o = ((pages_per_split * 4 + 15) / 16) * 16;
  o += warps * (head_size + 2) * 4;
  o = ((o + 15) / 16) * 16;
  o += warps * pbuf;
  o += staged ? warps * stages * 8 : 0;
  o = ((o + 127) / 128) * 128;
  o += staged ? warps * stages * 2 * page : 0;
  return o;
}


// CTAs of this launch shape that fit one SM (shared memory, threads)
static int resident_ctas(int dtype, int head_size, int block_size, int warps, int stages, bool staged) {
  const int smem1 = decode_smem_bytes(dtype, head_size, block_size, 64, warps, stages, staged);
  int resident = kMaxSmem / (smem1 + 1024);
  const int by_threads = 2048 / (warps * kWarp);
  if (resident > by_threads) resident = by_threads;
  if (resident > 32) resident = 32;
  return resident < 1 ? 1 : resident;
}

constexpr int kMaxPagesPerSplit = 8192;  // block-table slice kept in shared memory (32 KiB)

// Fills the zero fields of `t`.  Returns pages_per_split through the pointer.
//
// The rules come from sweeps on B200 (profiles/r1_sweep_*.json):
//   * two ring stages per warp, deeper rings only cost resident CTAs -- except for the page sizes listed at
//     num_stages below (round 3: profiles/r3_dispatch_tune.json);
//   * warps per CTA follow the pages one (sequence, head) has: a warp wants >= 4 pages,
//     and short contexts run best as many small CTAs (W = 1: 14 resident CTAs per SM);
//   * the context is split only when the (sequence, head) items cannot occupy the
//     resident CTA slots once, and never below 8 pages per warp: every extra CTA
//     costs ~2.7 us of prologue/epilogue during which its slot loads nothing.
static int make_plan(int num_seqs, int num_heads, int head_size, int block_size, int max_context_len,
                     int dtype, pa_decode_tuning* t, int* pages_per_split) {
  const int page = head_size * block_size * dtype_size(dtype);
  const int stage = 2 * page;
  const int max_pages = max_context_len > 0 ? (max_context_len + block_size - 1) / block_size : 1;

  if (t->variant <= 0) t->variant = 2;
  const bool staged = t->variant >= 2;
  const bool auto_warps = t->num_warps <= 0;
  if (t->num_warps <= 0) {
    t->num_warps = max_pages <= 4 ? 1 : (max_pages <= 8 ? 2 : 4);
    while (staged && t->num_warps > 1 && t->num_warps * 2 * stage > 160 * 1024) t->num_warps >>= 1;
    // 16 KiB pages (fp32, block 32, head 128): two warps' ring already takes more than half an SM, so one CTA is
    // resident either way -- a third warp raises the bytes in flight from 128 to 192 KiB (5.4 -> 7.1 TB/s)
    if (staged && t->num_warps == 2 && 2 * 2 * stage > 113 * 1024 && 3 * 2 * stage <= 200 * 1024) t->num_warps = 3;
    // 8 KiB pages of 32 tokens, fp32 (head 64): 4 warps x 2 stages is one CTA of 128 KiB per SM; CTAs of two warps fit
    // three times (192 KiB in flight): 6.96 -> 7.46 TB/s.  The 16-bit twin (head 128) keeps its four warps, which the
    // sibling-group path needs, and gets a third stage instead (below).
    if (staged && t->num_warps == 4 && block_size == 32 && page == 8192 && dtype == PA_F32) t->num_warps = 2;
  }
  if (t->num_warps > 8) t->num_warps = 8;
  if (t->num_stages <= 0) {
    // All rules measured over the reference's dispatch set at 256 x 32 x 1024 (profiles/r3_dispatch_tune.json).
    // Two stages per warp, except:
    //  * pages under 2 KiB (block 8 with head 64 / 80 / 96 in 16 bits): a warp is through such a page long before the
    //    next copy lands, and two stages leave it waiting -- 8 KiB in flight per warp: 4.6 -> 7.1 TB/s (head 64, four
    //    stages), 5.9 -> 7.3 (head 96, three);
    //  * 8 KiB pages of 32 tokens in 16 bits: one CTA per SM, a third stage brings 192 KiB in flight: 7.19 -> 7.39.
    t->num_stages = 2;
    if (staged && stage < 4096) t->num_stages = (8192 + stage - 1) / stage;
    if (staged && auto_warps && t->num_warps == 4 && block_size == 32 && page == 8192 && dtype != PA_F32) t->num_stages = 3;
  }
  if (!staged) t->num_stages = 1;
  if (t->num_stages > 32) t->num_stages = 32;
  // Just over one wave of short items (444 < items <= 1.25 x 444 at 4 warps, <= 64 pages): the second wave
  // would run nearly empty behind a full first one and pay a second prologue; CTAs of two warps fit twice as
  // many per SM, so the launch becomes a single wave with every page in flight from the start (measured,
  // profiles/r1_short_ctx_vs_reference_kernel.json S16_L512: 28.8 -> 25.0 us; profiles/r2_wave_probe_h5.json
  // shows the step at the first wave boundary, +6..9 us, and none at the later ones).
  if (auto_warps && staged && t->num_warps == 4 && max_pages <= 64) {
    const int64_t items = static_cast<int64_t>(num_seqs) * num_heads;
    const int64_t slots4 = static_cast<int64_t>(sm_count()) *
                           resident_ctas(dtype, head_size, block_size, 4, t->num_stages, staged);
    if (items > slots4 && items * 4 <= slots4 * 5) t->num_warps = 2;
  }

  if (t->num_splits <= 0) {
    const int resident = resident_ctas(dtype, head_size, block_size, t->num_warps, t->num_stages, staged);
    const int64_t items = static_cast<int64_t>(num_seqs) * num_heads;
    const int64_t slots = static_cast<int64_t>(sm_count()) * resident;
    int64_t ns = items > 0 ? slots / items : 1;                  // fill the slots once, no more
    const int64_t cap = max_pages / (8 * t->num_warps);          // keep >= 8 pages per warp
    if (ns > cap) ns = cap;
    if (ns < 1) ns = 1;
    t->num_splits = static_cast<int>(ns);
  }
  if (t->num_splits > max_pages) t->num_splits = max_pages;
  if (t->num_splits < 1) t->num_splits = 1;
  // pages per split, a multiple of the warp count so warps stay balanced
  int pps = (max_pages + t->num_splits - 1) / t->num_splits;
  pps = (pps + t->num_warps - 1) / t->num_warps * t->num_warps;
  if (pps > kMaxPagesPerSplit) pps = kMaxPagesPerSplit / t->num_warps * t->num_warps;
  t->num_splits = (max_pages + pps - 1) / pps;
  if (t->num_splits > 65535) return set_error(PA_ERR_INVALID_ARGUMENT, "context too long: %d splits", t->num_splits);
  *pages_per_split = pps;

  // shrink the ring until the CTA fits
  while (staged && t->num_stages > 1 &&
         decode_smem_bytes(dtype, head_size, block_size, pps, t->num_warps, t->num_stages, true) > kMaxSmem)
    --t->num_stages;
  if (decode_smem_bytes(dtype, head_size, block_size, pps, t->num_warps, t->num_stages, staged) > kMaxSmem)
    return set_error(PA_ERR_INVALID_ARGUMENT, "launch shape does not fit shared memory (warps %d, stages %d)",
                     t->num_warps, t->num_stages);
  return PA_OK;
}

static int check_shape(int head_size, int block_size, int dtype) {
  // the reference's dispatch set: attention_kernels.cu:370-400 (head), :416-448 (block), :460-468 (dtype)
  if (dtype != PA_F16 && dtype != PA_BF16 && dtype != PA_F32)
    return set_error(PA_ERR_UNSUPPORTED_DTYPE, "unsupported data type: %d", dtype);
  if (head_size != 64 && head_size != 80 && head_size != 96 && head_size != 128)
    return set_error(PA_ERR_UNSUPPORTED_HEAD_SIZE, "unsupported head size: %d", head_size);
  if (block_size != 8 && block_size != 16 && block_size != 32)
    return set_error(PA_ERR_UNSUPPORTED_BLOCK_SIZE, "unsupported block size: %d", block_size);
  return PA_OK;
}

static int decode(void* out, const void* query, const void* key_cache, const void* value_cache, float scale,
                  const int32_t* block_tables, const int32_t* context_lens, int32_t num_seqs,
                  int32_t num_heads, int32_t head_size, int32_t block_size, int32_t max_num_blocks_per_seq,
                  int32_t max_context_len, int64_t q_stride, int32_t dtype, const pa_decode_tuning* tuning,
                  cudaStream_t stream, bool reserve_only = false) {
  int rc = check_shape(head_size, block_size, dtype);
  if (rc != PA_OK) return rc;
  if (num_seqs < 0 || num_heads <= 0 || max_num_blocks_per_seq < 0 || max_context_len < 0)
    return set_error(PA_ERR_INVALID_ARGUMENT, "bad sizes (seqs %d, heads %d, max blocks %d, max context %d)",
                     num_seqs, num_heads, max_num_blocks_per_seq, max_context_len);
  if (num_seqs == 0) return PA_OK;
  if (num_seqs > 65535) return set_error(PA_ERR_INVALID_ARGUMENT, "too many sequences: %d (max 65535)", num_seqs);
  if (!reserve_only && (!out || !query || !key_cache || !value_cache || !block_tables || !context_lens))
    return set_error(PA_ERR_INVALID_ARGUMENT, "null pointer");
  if ((reinterpret_cast<uintptr_t>(key_cache) | reinterpret_cast<uintptr_t>(value_cache)) & 15u)
    return set_error(PA_ERR_INVALID_ARGUMENT, "key/value cache must be 16-byte aligned");
  const int es = dtype_size(dtype);
  if (reinterpret_cast<uintptr_t>(query) % es || reinterpret_cast<uintptr_t>(out) % es)
    return set_error(PA_ERR_INVALID_ARGUMENT, "query/out are not element-aligned");

  pa_decode_tuning t = {0, 0, 0, 0};
  if (tuning) t = *tuning;
  if (t.variant < 0 || t.variant > 3) return set_error(PA_ERR_INVALID_ARGUMENT, "invalid variant %d", t.variant);
  int pps = 0;
  rc = make_plan(num_seqs, num_heads, head_size, block_size, max_context_len, dtype, &t, &pps);
  if (rc != PA_OK) return rc;

  DecodeLaunch l;
  std::memset(&l, 0, sizeof(l));
  DecodeParams& p = l.params;
  p.out = out; p.q = query; p.k_cache = key_cache; p.v_cache = value_cache;
  p.block_tables = block_tables; p.context_lens = context_lens;
  p.q_stride = q_stride;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.num_heads = num_heads;
  p.max_blocks_per_seq = max_num_blocks_per_seq;
  p.pages_per_split = pps;
  p.num_splits = t.num_splits;
  p.num_stages = t.num_stages;
  p.q_vec_ok = (reinterpret_cast<uintptr_t>(query) % 16 == 0) && ((q_stride * es) % 16 == 0);
  p.num_seqs = num_seqs;
  p.group_rows = group_rows_for(t.variant >= 2, t.variant, t.num_warps, num_seqs, dtype, block_size,
                                decode_smem_bytes(dtype, head_size, block_size, pps, t.num_warps, t.num_stages, t.variant >= 2));
  if (t.num_splits > 1) {
    const size_t items = static_cast<size_t>(num_seqs) * num_heads;
    rc = get_workspace(stream, items * sizeof(uint32_t), items * t.num_splits * (head_size + 2) * sizeof(float),
                       &p.ws_ticket, &p.ws_part);
    if (rc != PA_OK) return rc;
  }
  if (reserve_only) return PA_OK;
  // PA_DEBUG_CHECKS=1 (read once per process): the one contract the host cannot normally verify, max_context_len >=
  // max(context_lens) -- a longer sequence would silently lose its tail tokens -- is checked by reading context_lens
  // back.  This synchronises the stream on every call: a debugging aid, skipped while the stream is being captured.
  static const bool debug_checks = [] { const char* e = getenv("PA_DEBUG_CHECKS"); return e && e[0] == '1'; }();
  if (debug_checks) {
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    PA_CUDA_TRY(cudaStreamIsCapturing(stream, &cap));
    if (cap == cudaStreamCaptureStatusNone) {
      std::vector<int32_t> lens(num_seqs);
      PA_CUDA_TRY(cudaMemcpyAsync(lens.data(), context_lens, sizeof(int32_t) * num_seqs, cudaMemcpyDeviceToHost, stream));
      PA_CUDA_TRY(cudaStreamSynchronize(stream));
      for (int i = 0; i < num_seqs; ++i)
        if (lens[i] > max_context_len || lens[i] < 0)
          return set_error(PA_ERR_INVALID_ARGUMENT, "context_lens[%d] = %d is outside [0, max_context_len = %d]: the "
                           "page walk is sized from max_context_len (worker.py:191 passes the maximum)", i, lens[i],
                           max_context_len);
    }
  }
  l.num_seqs = num_seqs; l.head_size = head_size; l.block_size = block_size;
  l.num_warps = t.num_warps; l.staged = (t.variant >= 2); l.stream = stream;
  static const bool pdl = [] { const char* e = getenv("PA_DECODE_PDL"); return !(e && e[0] == '0'); }();
  l.pdl = pdl;
  switch (dtype) {
    case PA_F16:  rc = launch_decode_f16(l); break;
    case PA_BF16: rc = launch_decode_bf16(l); break;
    default:      rc = launch_decode_f32(l); break;
  }
  if (rc != PA_OK) poison_workspace(stream);
  return rc;
}


// ---------------------------------------------------------------------------
// decode attention for callers that must not allocate (pa_decode_attention_static)
//
// The split-context partials and tickets live in an array that is part of the library image instead of
// a cudaMalloc'ed buffer: [0, kStaticTicketBytes) tickets, the rest partial results.  Every device that
// loads the library carries its own copy until the process exits.
// ---------------------------------------------------------------------------
constexpr size_t kStaticScratchBytes = PA_STATIC_SCRATCH_BYTES;
constexpr size_t kStaticTicketBytes = PA_STATIC_SCRATCH_MAX_ITEMS * sizeof(uint32_t);
static_assert(kStaticTicketBytes % 256 == 0 && kStaticTicketBytes < kStaticScratchBytes, "ticket region");
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

// Device address of g_static_scratch on the current device.  In host code the array's name is only a handle
// for the runtime -- it is not an address a kernel can use -- so the address is asked for once per device,
// with that device current, and remembered.
static int static_scratch_address(unsigned char** base) {
  static std::mutex mu;
  static unsigned char* cached[64] = {nullptr};
  int dev = 0;
  PA_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return set_error(PA_ERR_INVALID_ARGUMENT, "device ordinal %d out of range (max 63)", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!cached[dev]) {
    void* ptr = nullptr;
    PA_CUDA_TRY(cudaGetSymbolAddress(&ptr, g_static_scratch));
    if (!ptr || (reinterpret_cast<uintptr_t>(ptr) & 255u))
      return set_error(PA_ERR_CUDA, "cudaGetSymbolAddress returned an unusable address for the static scratch");
    cached[dev] = static_cast<unsigned char*>(ptr);
  }
  *base = cached[dev];
  return PA_OK;
}

static int decode_static(void* out, const void* query, const void* key_cache, const void* value_cache, float scale,
                         const int32_t* block_tables, const int32_t* context_lens, int32_t num_seqs,
                         int32_t num_heads, int32_t head_size, int32_t block_size, int32_t max_num_blocks_per_seq,
                         int32_t max_context_len, int64_t q_stride, int32_t dtype, const pa_decode_tuning* tuning,
                         cudaStream_t stream) {
  int rc = check_shape(head_size, block_size, dtype);
  if (rc != PA_OK) return rc;
  if (num_seqs < 0 || num_heads <= 0 || max_num_blocks_per_seq < 0 || max_context_len < 0)
    return set_error(PA_ERR_INVALID_ARGUMENT, "bad sizes (seqs %d, heads %d, max blocks %d, max context %d)",
                     num_seqs, num_heads, max_num_blocks_per_seq, max_context_len);
  if (num_seqs == 0) return PA_OK;
  if (num_seqs > 65535) return set_error(PA_ERR_INVALID_ARGUMENT, "too many sequences: %d (max 65535)", num_seqs);
  if (!out || !query || !key_cache || !value_cache || !block_tables || !context_lens)
    return set_error(PA_ERR_INVALID_ARGUMENT, "null pointer");
  if ((reinterpret_cast<uintptr_t>(key_cache) | reinterpret_cast<uintptr_t>(value_cache)) & 15u)
    return set_error(PA_ERR_INVALID_ARGUMENT, "key/value cache must be 16-byte aligned");
  const int es = dtype_size(dtype);
  if (reinterpret_cast<uintptr_t>(query) % es || reinterpret_cast<uintptr_t>(out) % es)
    return set_error(PA_ERR_INVALID_ARGUMENT, "query/out are not element-aligned");

  pa_decode_tuning t = {0, 0, 0, 0};
  if (tuning) t = *tuning;
  if (t.variant < 0 || t.variant > 3) return set_error(PA_ERR_INVALID_ARGUMENT, "invalid variant %d", t.variant);
  int pps = 0;
  rc = make_plan(num_seqs, num_heads, head_size, block_size, max_context_len, dtype, &t, &pps);
  if (rc != PA_OK) return rc;

  DecodeLaunch l;
  std::memset(&l, 0, sizeof(l));
  DecodeParams& p = l.params;
  p.out = out; p.q = query; p.k_cache = key_cache; p.v_cache = value_cache;
  p.block_tables = block_tables; p.context_lens = context_lens;
  p.q_stride = q_stride;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.num_heads = num_heads;
  p.max_blocks_per_seq = max_num_blocks_per_seq;
  p.pages_per_split = pps;
  p.num_splits = t.num_splits;
  p.num_stages = t.num_stages;
  p.q_vec_ok = (reinterpret_cast<uintptr_t>(query) % 16 == 0) && ((q_stride * es) % 16 == 0);
  p.num_seqs = num_seqs;
  p.group_rows = group_rows_for(t.variant >= 2, t.variant, t.num_warps, num_seqs, dtype, block_size,
                                decode_smem_bytes(dtype, head_size, block_size, pps, t.num_warps, t.num_stages, t.variant >= 2));
  if (t.num_splits > 1) {
    const size_t items = static_cast<size_t>(num_seqs) * num_heads;
    const size_t ticket_bytes = items * sizeof(uint32_t);
    const size_t part_bytes = items * t.num_splits * (head_size + 2) * sizeof(float);
    if (ticket_bytes > kStaticTicketBytes || part_bytes > kStaticScratchBytes - kStaticTicketBytes)
      return set_error(PA_ERR_INVALID_ARGUMENT,
                       "pa_decode_attention_static: %zu (sequence, head) items in %d context splits need %zu B of tickets and "
                       "%zu B of partial results; the built-in scratch holds %zu B and %zu B (PA_STATIC_SCRATCH_MAX_ITEMS "
                       "items, PA_STATIC_SCRATCH_BYTES in all).  Pass a tuning with fewer num_splits, or use "
                       "pa_decode_attention", items, t.num_splits, ticket_bytes, part_bytes, kStaticTicketBytes,
                       kStaticScratchBytes - kStaticTicketBytes);
    unsigned char* base = nullptr;
    rc = static_scratch_address(&base);
    if (rc != PA_OK) return rc;
    // the array keeps whatever the previous call (or a failed launch) left in it; tickets must count from zero
    PA_CUDA_TRY(cudaMemsetAsync(base, 0, (ticket_bytes + 255) / 256 * 256, stream));
    p.ws_ticket = reinterpret_cast<uint32_t*>(base);
    p.ws_part = reinterpret_cast<float*>(base + kStaticTicketBytes);
  }
  l.num_seqs = num_seqs; l.head_size = head_size; l.block_size = block_size;
  l.num_warps = t.num_warps; l.staged = (t.variant >= 2); l.stream = stream;
  l.pdl = false;   // the launch follows a memset node, not a kernel: nothing to overlap with
  switch (dtype) {
    case PA_F16:  return launch_decode_f16(l);
    case PA_BF16: return launch_decode_bf16(l);
    default:      return launch_decode_f32(l);
  }
}

}  // namespace pa

namespace pa { void release_prefill_scratch(int dev); }   // prefill_attention.cu

extern "C" {

int pa_abi_version(void) { return PA_ABI_VERSION; }
const char* pa_last_error(void) { return pa::g_err; }
int pa_compiled_sm(void) { return 100; }
uint64_t pa_launch_count(void) { return pa::g_launches.load(std::memory_order_relaxed); }

int pa_release_workspaces(void) {
  int dev = 0;
  PA_CUDA_TRY(cudaGetDevice(&dev));
  PA_CUDA_TRY(cudaDeviceSynchronize());
  pa::release_prefill_scratch(dev);
  std::lock_guard<std::mutex> lock(pa::g_ws_mu);
  for (auto it = pa::g_ws.begin(); it != pa::g_ws.end();) {
    if (it->first.first == dev) {
      if (it->second.ptr) cudaFree(it->second.ptr);
      it = pa::g_ws.erase(it);
    } else {
      ++it;
    }
  }
  for (auto it = pa::g_ws_retired.begin(); it != pa::g_ws_retired.end();) {
    if (it->first == dev) {
      cudaFree(it->second);
      it = pa::g_ws_retired.erase(it);
    } else {
      ++it;
    }
  }
  return PA_OK;
}

int pa_decode_attention(void* out, const void* query, const void* key_cache, const void* value_cache,
                        float scale, const int32_t* block_tables, const int32_t* context_lens,
                        int32_t num_seqs, int32_t num_heads, int32_t head_size, int32_t block_size,
                        int32_t max_num_blocks_per_seq, int32_t max_context_len, int64_t q_stride,
                        int32_t dtype, void* stream) {
  return pa::decode(out, query, key_cache, value_cache, scale, block_tables, context_lens, num_seqs, num_heads,
                    head_size, block_size, max_num_blocks_per_seq, max_context_len, q_stride, dtype, nullptr,
                    static_cast<cudaStream_t>(stream));
}

int pa_decode_attention_ex(void* out, const void* query, const void* key_cache, const void* value_cache,
                           float scale, const int32_t* block_tables, const int32_t* context_lens,
                           int32_t num_seqs, int32_t num_heads, int32_t head_size, int32_t block_size,
                           int32_t max_num_blocks_per_seq, int32_t max_context_len, int64_t q_stride,
                           int32_t dtype, const pa_decode_tuning* tuning, void* stream) {
  return pa::decode(out, query, key_cache, value_cache, scale, block_tables, context_lens, num_seqs, num_heads,
                    head_size, block_size, max_num_blocks_per_seq, max_context_len, q_stride, dtype, tuning,
                    static_cast<cudaStream_t>(stream));
}

int pa_decode_attention_reserve(int32_t num_seqs, int32_t num_heads, int32_t head_size, int32_t block_size,
                                int32_t max_num_blocks_per_seq, int32_t max_context_len, int32_t dtype,
                                const pa_decode_tuning* tuning, void* stream) {
  return pa::decode(nullptr, nullptr, nullptr, nullptr, 1.f, nullptr, nullptr, num_seqs, num_heads, head_size,
                    block_size, max_num_blocks_per_seq, max_context_len, head_size, dtype, tuning,
                    static_cast<cudaStream_t>(stream), true);
}

int pa_decode_attention_static(void* out, const void* query, const void* key_cache, const void* value_cache,
                               float scale, const int32_t* block_tables, const int32_t* context_lens,
                               int32_t num_seqs, int32_t num_heads, int32_t head_size, int32_t block_size,
                               int32_t max_num_blocks_per_seq, int32_t max_context_len, int64_t q_stride,
                               int32_t dtype, const pa_decode_tuning* tuning, void* stream) {
  return pa::decode_static(out, query, key_cache, value_cache, scale, block_tables, context_lens, num_seqs,
                           num_heads, head_size, block_size, max_num_blocks_per_seq, max_context_len, q_stride,
                           dtype, tuning, static_cast<cudaStream_t>(stream));
}

int pa_decode_attention_plan(int32_t num_seqs, int32_t num_heads, int32_t head_size, int32_t block_size,
                             int32_t max_context_len, int32_t dtype, pa_decode_tuning* plan_out) {
  if (!plan_out) return pa::set_error(PA_ERR_INVALID_ARGUMENT, "plan_out is null");
  int rc = pa::check_shape(head_size, block_size, dtype);
  if (rc != PA_OK) return rc;
  int pps = 0;
  return pa::make_plan(num_seqs, num_heads, head_size, block_size, max_context_len, dtype, plan_out, &pps);
}

}  // extern "C"
