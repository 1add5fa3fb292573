// Per-step model inputs from device-resident block tables.
//
// Replaces: Worker._prepare_inputs (vllm/worker/worker.py:128-237).  The reference turns the scheduler's
// per-sequence Python lists into five tensors every step (nested-list -> torch.cuda.IntTensor, blocking
// pageable copies, ~137 KB at 256 sequences x 2048 tokens).  Here the tables stay in HBM:
//
//   host (this file)   seq id -> row of a resident [max_seqs][max_blocks] int32 table, with a shadow copy;
//                      per step it packs ONE command buffer in pinned memory: per sequence
//                      {row, token, context length, block count, changed span} + the span's block numbers
//                      (one block every `block_size` steps in steady decode; a whole table only for a
//                      sequence that was not in the previous step), plus the prompts' tokens and tables;
//   device             one cudaMemcpyAsync of that buffer, one launch of step_inputs_kernel:
//                      a warp per generation sequence patches its resident row and writes the padded
//                      block-table row, context length, slot, position and token; a thread per prompt
//                      token writes token / position / slot.
// Outputs are bit-identical to the reference's tensors.
#include <algorithm>
#include <cstring>
#include <unordered_map>
#include <vector>

#include "pa_common.cuh"

namespace pa {

// Command buffer (int32 words).  Header, then seven arrays of S words, then the prompt arrays, then payload.
enum StepHdr : int {
  kHdrNumSeqs = 0,        // S
  kHdrNumPrompts,         // P
  kHdrPromptTokens,       // tokens of all prompts
  kHdrPaddedTokens,       // pad8(prompt tokens + S)
  kHdrWidth,              // columns of the block-table output
  kHdrGenBase,            // word offset of the generation arrays
  kHdrPromptBase,         // word offset of the prompt arrays
  kHdrPayloadBase,        // word offset of the block-number payload
  kHdrWords
};
enum GenField : int { kGenRow = 0, kGenToken, kGenLen, kGenBlocks, kGenSpanStart, kGenSpanCount, kGenSpanOff, kGenFields };

constexpr int kStepThreads = 128;
constexpr int kStepWarps = kStepThreads / kWarp;

struct StepKernelArgs {
  const int32_t* cmd;        // device copy of the command buffer
  int32_t* tables;           // resident [max_seqs][stride]
  int32_t stride;            // max_blocks_per_seq
  int32_t block_size;
  int32_t gen_ctas;          // CTAs [0, gen_ctas) serve generation sequences, the rest prompt tokens
  int64_t* tokens;
  int64_t* positions;
  int32_t* slot_mapping;
  int32_t* context_lens;
  int32_t* block_tables;
};

__global__ void __launch_bounds__(kStepThreads) step_inputs_kernel(const StepKernelArgs a) {
  const int32_t* __restrict__ cmd = a.cmd;
  const int S = cmd[kHdrNumSeqs], P = cmd[kHdrNumPrompts], npt = cmd[kHdrPromptTokens];
  const int padded = cmd[kHdrPaddedTokens], width = cmd[kHdrWidth];
  const int32_t* __restrict__ payload = cmd + cmd[kHdrPayloadBase];
  const int B = a.block_size;

  if (blockIdx.x == 0) {   // zero padding up to a multiple of 8 (worker.py:201-204)
    for (int i = npt + S + threadIdx.x; i < padded; i += kStepThreads) {
      a.tokens[i] = 0;
      a.positions[i] = 0;
    }
  }

  if (static_cast<int>(blockIdx.x) < a.gen_ctas) {
    // ---- generation sequences: one warp each (worker.py:170-199) ----
    const int i = blockIdx.x * kStepWarps + (threadIdx.x >> 5);
    if (i >= S) return;
    const int lane = threadIdx.x & 31;
    const int32_t* __restrict__ g = cmd + cmd[kHdrGenBase];
    const int row = g[kGenRow * S + i], len = g[kGenLen * S + i], nb = g[kGenBlocks * S + i];
    const int s0 = g[kGenSpanStart * S + i], sn = g[kGenSpanCount * S + i];
    const int32_t* __restrict__ span = payload + g[kGenSpanOff * S + i];
    int32_t* __restrict__ resident = a.tables + static_cast<int64_t>(row) * a.stride;
    int32_t* __restrict__ out_row = a.block_tables + static_cast<int64_t>(i) * width;
    // patch the resident row with the changed span, and emit the padded output row; an entry inside the
    // span is taken from the payload, so nothing is read back after being written
    for (int c = lane; c < sn; c += kWarp) resident[s0 + c] = span[c];
    for (int c = lane; c < width; c += kWarp) {
      int32_t v = 0;                                   // _pad_to_max pads with block 0 (worker.py:308-309)
      if (c < nb) v = (c >= s0 && c < s0 + sn) ? span[c - s0] : resident[c];
      out_row[c] = v;
    }
    if (lane == 0) {
      const int pos = len - 1, bi = pos / B;
      const int32_t blk = (bi >= s0 && bi < s0 + sn) ? span[bi - s0] : resident[bi];
      a.context_lens[i] = len;
      a.slot_mapping[npt + i] = blk * B + (pos - bi * B);
      a.positions[npt + i] = pos;
      a.tokens[npt + i] = g[kGenToken * S + i];
    }
    return;
  }

  // ---- prompt tokens: one thread each (worker.py:138-167) ----
  const int t = (blockIdx.x - a.gen_ctas) * kStepThreads + threadIdx.x;
  if (t >= npt) return;
  const int32_t* __restrict__ pr = cmd + cmd[kHdrPromptBase];
  const int32_t* __restrict__ tok_off = pr;               // [P + 1]
  const int32_t* __restrict__ tab_off = pr + (P + 1);     // [P + 1] into payload, -1 = no table
  const int32_t* __restrict__ ptok = pr + 2 * (P + 1);    // [npt]
  int lo = 0, hi = P - 1;                                 // last p with tok_off[p] <= t
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (tok_off[mid] <= t) lo = mid; else hi = mid - 1;
  }
  const int i = t - tok_off[lo];
  a.tokens[t] = ptok[t];
  a.positions[t] = i;
  const int off = tab_off[lo];
  const int bi = i / B;
  a.slot_mapping[t] = off < 0 ? 0 : payload[off + bi] * B + (i - bi * B);
}

}  // namespace pa

// ---------------------------------------------------------------------------
// host state
// ---------------------------------------------------------------------------
struct pa_step_state {
  static constexpr int kSlots = 4;          // staging ring: steps may run ahead of the device by 3
  int32_t max_seqs = 0, stride = 0, block_size = 0;
  int device = 0;
  int32_t* tables = nullptr;                // device, [max_seqs][stride]
  struct Slot {
    int32_t* host = nullptr;                // pinned
    int32_t* dev = nullptr;
    size_t words = 0;
    cudaEvent_t done = nullptr;
    bool in_flight = false;
  } slots[kSlots];
  int next_slot = 0;

  struct Row {
    int64_t seq_id = -1;
    uint64_t seen = 0;
    std::vector<int32_t> blocks;            // shadow of the resident row
  };
  std::vector<Row> rows;
  std::vector<int32_t> free_rows;
  std::vector<int32_t> live;                // rows of the previous step
  std::unordered_map<int64_t, int32_t> row_of;
  uint64_t step_no = 0;
  // scratch reused across steps
  std::vector<int32_t> row_scratch, span_start, span_count;
};

namespace pa {

static int ensure_slot(pa_step_state* st, pa_step_state::Slot& s, size_t words) {
  if (s.in_flight) {
    PA_CUDA_TRY(cudaEventSynchronize(s.done));
    s.in_flight = false;
  }
  if (!s.done) PA_CUDA_TRY(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
  if (s.words >= words) return PA_OK;
  size_t cap = std::max<size_t>(words + words / 2, 4096);
  if (s.host) PA_CUDA_TRY(cudaFreeHost(s.host));
  if (s.dev) PA_CUDA_TRY(cudaFree(s.dev));
  s.host = nullptr; s.dev = nullptr; s.words = 0;
  PA_CUDA_TRY(cudaMallocHost(reinterpret_cast<void**>(&s.host), cap * sizeof(int32_t)));
  PA_CUDA_TRY(cudaMalloc(reinterpret_cast<void**>(&s.dev), cap * sizeof(int32_t)));
  s.words = cap;
  (void)st;
  return PA_OK;
}

static void forget_all(pa_step_state* st) {
  st->row_of.clear();
  st->live.clear();
  st->free_rows.clear();
  for (int32_t r = st->max_seqs - 1; r >= 0; --r) {
    st->rows[r].seq_id = -1;
    st->rows[r].seen = 0;
    st->rows[r].blocks.clear();
    st->free_rows.push_back(r);
  }
}

}  // namespace pa

extern "C" {

int pa_step_state_create(int32_t max_seqs, int32_t max_blocks_per_seq, int32_t block_size, pa_step_state** out) {
  if (!out) return pa::set_error(PA_ERR_INVALID_ARGUMENT, "state_out is null");
  *out = nullptr;
  if (max_seqs <= 0 || max_blocks_per_seq <= 0 || block_size <= 0)
    return pa::set_error(PA_ERR_INVALID_ARGUMENT, "bad sizes (max_seqs %d, max_blocks_per_seq %d, block_size %d)",
                         max_seqs, max_blocks_per_seq, block_size);
  pa_step_state* st = new pa_step_state();
  st->max_seqs = max_seqs;
  st->stride = max_blocks_per_seq;
  st->block_size = block_size;
  st->rows.resize(max_seqs);
  pa::forget_all(st);
  cudaError_t e = cudaGetDevice(&st->device);
  if (e == cudaSuccess)
    e = cudaMalloc(reinterpret_cast<void**>(&st->tables), sizeof(int32_t) * static_cast<size_t>(max_seqs) * max_blocks_per_seq);
  if (e != cudaSuccess) {
    delete st;
    return pa::check_cuda(e, "pa_step_state_create");
  }
  *out = st;
  return PA_OK;
}

int pa_step_state_destroy(pa_step_state* st) {
  if (!st) return PA_OK;
  for (auto& s : st->slots) {
    if (s.in_flight) cudaEventSynchronize(s.done);
    if (s.done) cudaEventDestroy(s.done);
    if (s.host) cudaFreeHost(s.host);
    if (s.dev) cudaFree(s.dev);
  }
  if (st->tables) cudaFree(st->tables);
  delete st;
  return PA_OK;
}

int pa_step_state_reset(pa_step_state* st) {
  if (!st) return pa::set_error(PA_ERR_INVALID_ARGUMENT, "state is null");
  pa::forget_all(st);
  return PA_OK;
}

}  // extern "C"

extern "C" int pa_step_prepare_inputs(pa_step_state* st, const pa_step_batch* b, const pa_step_outputs* out,
                                      pa_step_info* info, uint8_t* need_table, void* stream_) {
  using namespace pa;
  if (!st || !b || !out) return set_error(PA_ERR_INVALID_ARGUMENT, "null state / batch / outputs");
  const int S = b->num_seqs, P = b->num_prompts, B = st->block_size;
  if (S < 0 || P < 0) return set_error(PA_ERR_INVALID_ARGUMENT, "negative counts (seqs %d, prompts %d)", S, P);
  if (S > 0 && (!b->seq_ids || !b->last_tokens || !b->context_lens || !b->num_blocks || !b->last_block))
    return set_error(PA_ERR_INVALID_ARGUMENT, "generation arrays missing");
  if (P > 0 && (!b->prompt_offsets || !b->prompt_tokens))
    return set_error(PA_ERR_INVALID_ARGUMENT, "prompt arrays missing");
  if (S > st->max_seqs) return set_error(PA_ERR_INVALID_ARGUMENT, "%d sequences > max_seqs %d", S, st->max_seqs);
  const int64_t npt64 = P > 0 ? b->prompt_offsets[P] - b->prompt_offsets[0] : 0;
  if (npt64 < 0 || npt64 + S > (1 << 30)) return set_error(PA_ERR_INVALID_ARGUMENT, "bad prompt offsets");
  const int npt = static_cast<int>(npt64);
  const int total = npt + S, padded = (total + 7) / 8 * 8;
  if (total == 0) return set_error(PA_ERR_INVALID_ARGUMENT, "empty step");
  if (!out->tokens || !out->positions || !out->slot_mapping || (S > 0 && (!out->context_lens || !out->block_tables)))
    return set_error(PA_ERR_INVALID_ARGUMENT, "null output pointer");

  // ---- pass 1: validate and decide every sequence's changed span; the state is not touched ----
  struct Plan { int32_t row, s0, sn; const int32_t* src; };
  std::vector<Plan> plan(S);
  std::vector<uint8_t> present(st->max_seqs, 0);
  std::vector<int64_t> fresh;
  int max_len = 0, max_nb = 0, n_incr = 0, n_full = 0, n_diff = 0, need = 0;
  int64_t span_words = 0;
  if (need_table && S > 0) std::memset(need_table, 0, S);
  for (int i = 0; i < S; ++i) {
    const int nb = b->num_blocks[i], len = b->context_lens[i];
    if (nb < 1 || nb > st->stride)
      return set_error(PA_ERR_INVALID_ARGUMENT, "sequence %lld: %d blocks (max_blocks_per_seq %d)",
                       static_cast<long long>(b->seq_ids[i]), nb, st->stride);
    if (len < 1 || (len - 1) / B >= nb)
      return set_error(PA_ERR_INVALID_ARGUMENT, "sequence %lld: context length %d does not fit %d blocks of %d",
                       static_cast<long long>(b->seq_ids[i]), len, nb, B);
    if (b->last_tokens[i] < INT32_MIN || b->last_tokens[i] > INT32_MAX)
      return set_error(PA_ERR_INVALID_ARGUMENT, "token id out of the 32-bit range");
    max_len = std::max(max_len, len);
    max_nb = std::max(max_nb, nb);
    const int32_t* table = nullptr;
    if (b->table_offsets) {
      const int64_t n = b->table_offsets[i + 1] - b->table_offsets[i];
      if (n != 0 && n != nb)
        return set_error(PA_ERR_INVALID_ARGUMENT, "sequence %lld brings %lld table entries, num_blocks says %d",
                         static_cast<long long>(b->seq_ids[i]), static_cast<long long>(n), nb);
      if (n) {
        if (!b->table_blocks) return set_error(PA_ERR_INVALID_ARGUMENT, "table_blocks is null");
        table = b->table_blocks + b->table_offsets[i];
        if (table[nb - 1] != b->last_block[i])
          return set_error(PA_ERR_INVALID_ARGUMENT, "sequence %lld: last_block disagrees with its table",
                           static_cast<long long>(b->seq_ids[i]));
      }
    }
    Plan& p = plan[i];
    auto it = st->row_of.find(b->seq_ids[i]);
    if (it != st->row_of.end()) {
      p.row = it->second;
      if (present[p.row])
        return set_error(PA_ERR_INVALID_ARGUMENT, "sequence %lld appears twice", static_cast<long long>(b->seq_ids[i]));
      present[p.row] = 1;
      const std::vector<int32_t>& have = st->rows[p.row].blocks;
      const int hn = static_cast<int>(have.size());
      if (table) {                        // diff against the resident copy, send only [first, last] changed
        int first = 0;
        const int common = std::min(hn, nb);
        while (first < common && have[first] == table[first]) ++first;
        int last = nb - 1;
        if (nb <= hn) while (last >= first && have[last] == table[last]) --last;
        p.s0 = first;
        p.sn = std::max(0, last - first + 1);
        p.src = table + first;
        ++n_diff;
      } else if (hn == nb) {              // in place, or last block copied on write (block_manager.py:124-135)
        p.s0 = nb - 1;
        p.sn = have[nb - 1] != b->last_block[i] ? 1 : 0;
        p.src = b->last_block + i;
        ++n_incr;
      } else if (hn == nb - 1) {          // a new block was appended (block_manager.py:117-122)
        p.s0 = nb - 1;
        p.sn = 1;
        p.src = b->last_block + i;
        ++n_incr;
      } else {
        ++need;
        if (need_table) need_table[i] = 1;
        continue;
      }
    } else {
      if (!table) {
        ++need;
        if (need_table) need_table[i] = 1;
        continue;
      }
      p.row = -1;
      p.s0 = 0;
      p.sn = nb;
      p.src = table;
      fresh.push_back(b->seq_ids[i]);
      ++n_full;
    }
    span_words += p.sn;
  }
  if (need)
    return set_error(PA_ERR_NEED_TABLE, "%d sequence(s) were not in the previous step (or changed by more than "
                     "one block) and brought no block table", need);
  if (!fresh.empty()) {
    std::vector<int64_t> sorted(fresh);
    std::sort(sorted.begin(), sorted.end());
    if (std::adjacent_find(sorted.begin(), sorted.end()) != sorted.end())
      return set_error(PA_ERR_INVALID_ARGUMENT, "a new sequence id appears twice");
  }
  if (S > 0 && out->block_table_width < max_nb)
    return set_error(PA_ERR_INVALID_ARGUMENT, "block_table_width %d < widest table %d", out->block_table_width, max_nb);
  int64_t prompt_table_words = 0;
  for (int p = 0; p < P; ++p) {
    const int64_t n = b->prompt_offsets[p + 1] - b->prompt_offsets[p];
    if (n < 0) return set_error(PA_ERR_INVALID_ARGUMENT, "prompt offsets must not decrease");
    for (int64_t k = b->prompt_offsets[p]; k < b->prompt_offsets[p + 1]; ++k)
      if (b->prompt_tokens[k] < INT32_MIN || b->prompt_tokens[k] > INT32_MAX)
        return set_error(PA_ERR_INVALID_ARGUMENT, "token id out of the 32-bit range");
    const int64_t tn = b->prompt_table_offsets ? b->prompt_table_offsets[p + 1] - b->prompt_table_offsets[p] : 0;
    if (tn != 0 && tn * B < n)
      return set_error(PA_ERR_INVALID_ARGUMENT, "prompt %d: %lld tokens do not fit %lld blocks", p,
                       static_cast<long long>(n), static_cast<long long>(tn));
    if (tn && !b->prompt_table_blocks) return set_error(PA_ERR_INVALID_ARGUMENT, "prompt_table_blocks is null");
    prompt_table_words += tn;
  }
  {   // rows: those of sequences that left are reusable right away (nobody reads them this step)
    size_t leaving = 0;
    for (int32_t r : st->live) leaving += present[r] ? 0 : 1;
    if (fresh.size() > st->free_rows.size() + leaving)
      return set_error(PA_ERR_INVALID_ARGUMENT, "more than max_seqs (%d) sequences alive", st->max_seqs);
  }

  // ---- pass 2: commit to the host state and pack the command buffer ----
  const size_t words = kHdrWords + static_cast<size_t>(kGenFields) * S + 2 * static_cast<size_t>(P + 1) + npt +
                       static_cast<size_t>(span_words) + static_cast<size_t>(prompt_table_words);
  pa_step_state::Slot& slot = st->slots[st->next_slot];
  {
    const int rc = ensure_slot(st, slot, words);
    if (rc != PA_OK) return rc;
  }
  st->next_slot = (st->next_slot + 1) % pa_step_state::kSlots;
  ++st->step_no;
  for (int32_t r : st->live) {
    if (present[r]) continue;
    st->row_of.erase(st->rows[r].seq_id);
    st->rows[r].seq_id = -1;
    st->rows[r].blocks.clear();
    st->free_rows.push_back(r);
  }
  st->live.clear();

  int32_t* cmd = slot.host;
  const int gen_base = kHdrWords, prompt_base = gen_base + kGenFields * S;
  const int payload_base = prompt_base + 2 * (P + 1) + npt;
  cmd[kHdrNumSeqs] = S;
  cmd[kHdrNumPrompts] = P;
  cmd[kHdrPromptTokens] = npt;
  cmd[kHdrPaddedTokens] = padded;
  cmd[kHdrWidth] = out->block_table_width;
  cmd[kHdrGenBase] = gen_base;
  cmd[kHdrPromptBase] = prompt_base;
  cmd[kHdrPayloadBase] = payload_base;
  int32_t* g = cmd + gen_base;
  int32_t* payload = cmd + payload_base;
  int32_t cursor = 0;
  for (int i = 0; i < S; ++i) {
    Plan& p = plan[i];
    const int nb = b->num_blocks[i];
    if (p.row < 0) {
      p.row = st->free_rows.back();
      st->free_rows.pop_back();
      st->rows[p.row].seq_id = b->seq_ids[i];
      st->row_of.emplace(b->seq_ids[i], p.row);
    }
    pa_step_state::Row& row = st->rows[p.row];
    row.seen = st->step_no;
    row.blocks.resize(nb);
    std::copy(p.src, p.src + p.sn, row.blocks.begin() + p.s0);
    st->live.push_back(p.row);
    g[kGenRow * S + i] = p.row;
    g[kGenToken * S + i] = static_cast<int32_t>(b->last_tokens[i]);
    g[kGenLen * S + i] = b->context_lens[i];
    g[kGenBlocks * S + i] = nb;
    g[kGenSpanStart * S + i] = p.s0;
    g[kGenSpanCount * S + i] = p.sn;
    g[kGenSpanOff * S + i] = cursor;
    std::copy(p.src, p.src + p.sn, payload + cursor);
    cursor += p.sn;
  }
  int32_t* pr = cmd + prompt_base;
  int32_t* tok_off = pr;
  int32_t* tab_off = pr + (P + 1);
  int32_t* ptok = pr + 2 * (P + 1);
  for (int p = 0; p < P; ++p) {
    const int64_t t0 = b->prompt_offsets[p], n = b->prompt_offsets[p + 1] - t0;
    tok_off[p] = static_cast<int32_t>(t0 - b->prompt_offsets[0]);
    for (int64_t k = 0; k < n; ++k) {
      ptok[tok_off[p] + k] = static_cast<int32_t>(b->prompt_tokens[t0 + k]);   // range checked in pass 1
    }
    const int64_t tn = b->prompt_table_offsets ? b->prompt_table_offsets[p + 1] - b->prompt_table_offsets[p] : 0;
    if (tn) {
      tab_off[p] = cursor;
      std::copy(b->prompt_table_blocks + b->prompt_table_offsets[p],
                b->prompt_table_blocks + b->prompt_table_offsets[p + 1], payload + cursor);
      cursor += static_cast<int32_t>(tn);
    } else {
      tab_off[p] = -1;
    }
  }
  tok_off[P] = npt;
  tab_off[P] = -1;

  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  PA_CUDA_TRY(cudaMemcpyAsync(slot.dev, slot.host, words * sizeof(int32_t), cudaMemcpyHostToDevice, stream));
  StepKernelArgs a;
  a.cmd = slot.dev;
  a.tables = st->tables;
  a.stride = st->stride;
  a.block_size = B;
  a.gen_ctas = (S + kStepWarps - 1) / kStepWarps;
  a.tokens = out->tokens;
  a.positions = out->positions;
  a.slot_mapping = out->slot_mapping;
  a.context_lens = out->context_lens;
  a.block_tables = out->block_tables;
  const int ctas = a.gen_ctas + (npt + kStepThreads - 1) / kStepThreads;
  step_inputs_kernel<<<ctas, kStepThreads, 0, stream>>>(a);
  PA_CUDA_TRY(cudaGetLastError());
  count_launch();
  PA_CUDA_TRY(cudaEventRecord(slot.done, stream));
  slot.in_flight = true;
  if (info) {
    info->max_context_len = max_len;
    info->max_num_blocks = max_nb;
    info->num_tokens = total;
    info->num_incremental = n_incr;
    info->num_full = n_full;
    info->num_diffed = n_diff;
    info->h2d_bytes = static_cast<int64_t>(words * sizeof(int32_t));
  }
  return PA_OK;
}
