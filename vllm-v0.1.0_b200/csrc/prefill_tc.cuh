// tcgen05 prompt-attention kernel (16-bit types).  See prefill_attention.cu for the operator.
//
// One CTA = 256 query rows of one (prompt, head), as two 128-row tiles A and B that share every K/V
// tile.  Twelve warps, four roles (warp 11 only donates its registers):
//   warps 0-3 / 4-7  softmax group of tile A / B: thread r owns row r (TMEM lane r % 128)
//   warp 8           TMA producer: Q tiles once, then K_j / V_j through two-stage rings (packed rows, or the
//                    pages of the paged cache as they lie)
//   warp 9           MMA issuer: one thread issues every tcgen05.mma of the CTA
//   warp 10          output store: TMA stores of finished 128-row tiles from the staging buffer
// Tensor memory (512 columns): S_A | S_B | O_A | O_B, 128 fp32 columns each; the probabilities
// P_i overwrite the first 64 columns of S_i (two 16-bit values per column) and are consumed
// from there as the A operand of P.V.  Per tile the dependency chain is
//   Q.K_j^T -> S_i --softmax group i--> P_i -> O_i += P_i.V_j, then Q.K_{j+1}^T into S_i again;
// the issuer interleaves the chains of A and B, so the tensor core works on one tile's matrices
// while the other tile's threads are in their softmax.  tcgen05.mma executes in issue order, which
// is what makes the S_i / P_i aliasing safe without an extra barrier.
//
// CTAs are persistent by work stealing: the grid still has one CTA per (tile, head) in the host's order, but a
// running CTA cancels the next pending one (cluster launch control) and takes over its blockIdx, so every role
// loops over work items.  Measured before (tools/prefill_timeline.py): 5400 clk from CTA entry to the first MMA
// (barrier init, TMEM allocation, Q / K_0 from HBM) and 3000-4400 clk from the last P.V to exit, with the
// tensor core idle in both; per item now the producer has Q and K_0 / V_0 of the NEXT item in shared memory
// before the current one ends (the rings simply continue), the issuer runs the next item's first Q.K^T behind
// the last P.V, and the output tile leaves through its own staging buffer.
#pragma once

#include <math_constants.h>

#include "tc05.cuh"

namespace pa {

#ifdef PA_PF_WATCHDOG
// debugging aid (compile prefill_attention.cu with -DPA_PF_WATCHDOG and <cstdio>): every mbarrier wait of the prompt
// kernel that is stuck for ~2^28 clocks says once where it is, and traps later -- the wait-for graph of a deadlock
__device__ __noinline__ void mbar_wait_dbg(uint32_t bar, uint32_t parity, int line) {
  const long long t0 = clock64();
  bool said = false;
  for (;;) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    if (ok) return;
    const long long dt = clock64() - t0;
    if (dt > (1ll << 28) && !said) {
      said = true;
      if ((threadIdx.x & 31) == 0 && blockIdx.y == 0 && blockIdx.x < 12)
        printf("[watchdog] cta (%d,%d) warp %d line %d bar+%u parity %u\n", blockIdx.x, blockIdx.y, threadIdx.x >> 5,
               line, bar & 1023u, parity);
    }
    if (dt > (3ll << 28)) asm volatile("trap;");
  }
}
#define mbar_wait(b, p) mbar_wait_dbg(b, p, __LINE__)
#endif

constexpr int kPfRows = 256;       // query rows per CTA (two tiles of 128 = TMEM lanes)
constexpr int kPfCols = 128;       // K/V tokens per tile
constexpr int kPfMaxTiles = 1024;  // CTAs per launch (the table travels as a kernel parameter)
constexpr int kPfThreads = 384;   // 8 softmax warps + producer + issuer + output store + 1 warp that only donates registers

// A sequence has qlen query rows (packed from token qtok0 of q / out) and klen >= qlen keys (packed from
// token ktok0 of k / v).  Query row i sees keys 0 .. i + (klen - qlen): plain prompt attention is
// klen == qlen; klen > qlen is a prompt chunk behind klen - qlen tokens that are already cached.
// The table travels as a kernel parameter, and parameters cost launch time (measured: 5 x 512 tokens, eager
// back-to-back launches, 36.7 us per call with a 20 KB table, 41.5 us with 28 KB, the kernel itself 32-33 us), so
// launches of up to kPfFewTiles entries use the small instantiation (1.8 KB).
constexpr int kPfFewTiles = 64;
template <int N>
struct PrefillTilesT {
  int32_t qtok0[N];            // first query token of the sequence in the packed batch
  int32_t qlen[N];             // query rows of the sequence
  int32_t q0[N];               // first query row of this CTA inside the sequence (multiple of its row count)
  int32_t ktok0[N];            // first key token of the sequence in the packed keys / values
  int32_t klen[N];             // keys of the sequence
  // Context splits: an entry may cover only K/V tiles [kt0, kt1) of its query tile (kt0 | kt1 << 16; 0 | 0xffff =
  // everything).  Such an entry leaves un-normalised partial results in the workspace slot (slot >> 8) +
  // head * (slot & 255) instead of writing out (slot < 0: a whole item); prefill_merge_kernel combines them.
  // Every part starts at or before the tile that holds the last key of the entry's FIRST row, so each of its rows
  // sees at least one key in the part's first tile (the masked tiles all belong to the last part).
  int32_t krange[N];
  int32_t slot[N];
};
using PrefillTiles = PrefillTilesT<kPfMaxTiles>;

// floats per (slot, 128-row group) of the split workspace: O transposed [DP][128] (a warp's stores of one column
// are 128 contiguous bytes), then the rows' running maxima (exp2 domain) and sums
__host__ __device__ constexpr int64_t pf_part_floats(int dp) { return static_cast<int64_t>(128) * dp + 256; }

struct PrefillParams {
  void* out;
  int64_t out_stride;  // elements between consecutive tokens of out
  int32_t num_heads;
  int32_t head_size;
  float scale_log2;
  float* dbg;          // tests: raw S of the first tile, raw O, row sums / maxima of tile A of CTA (0, 0)
  // kPaged: keys / values come straight from the paged caches (tiles.ktok0 then holds the sequence index)
  const void* key_cache;        // [block][head][head_size / 8][block_size][8]
  const void* value_cache;      // [block][head][head_size][block_size]
  const int32_t* block_tables;  // [sequence][max_blocks]
  int32_t max_blocks;
  int32_t block_shift;          // log2(block_size): 3, 4 or 5
  // 0: grid (head, tile) -- all heads' longest tiles first, the best balance when K/V fits L2 anyway;
  // 1: grid (tile, head) -- for batches whose K/V exceeds L2.  Measured on B200: 8 x 4096 tokens 948 -> 1064
  // TFLOP/s with 1, 5 x 512 tokens 317 -> 304 and a ragged 2560-token batch 418 -> 360 with it.
  int32_t tile_major;
  float* split_ws;              // partial results of split entries (see PrefillTiles::slot)
};

// tests / tools: SM-clock stamps of the hand-offs of CTA (head 0, tile 0) into the debug buffer, after the raw
// tiles: slot ((who * kTlIters + j) * 8 + event), who = 0 / 1 softmax group A / B, 2 issuer (tools/prefill_timeline.py)
constexpr int kTlBase = 40000, kTlIters = 40;
__device__ __forceinline__ void stamp(float* dbg, bool on, int who, int j, int ev) {
  if (on && j < kTlIters) {
    uint32_t c;
    asm volatile("mov.u32 %0, %%clock;" : "=r"(c));
    reinterpret_cast<uint32_t*>(dbg)[kTlBase + (who * kTlIters + j) * 8 + ev] = c;
  }
}

__device__ __forceinline__ float ex2(float x) {   // exp2(-inf) == 0
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// the same, pinned between the named barriers of the ping-pong section (volatile asm keeps its order)
__device__ __forceinline__ float ex2_pinned(float x) {
  float y;
  asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ float max3(float a, float b, float c) {   // FMNMX3
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
// (d0, d1) = (a0, a1) * b + c on the packed fp32 pipe (FFMA2)
__device__ __forceinline__ void ffma2(float& d0, float& d1, float a0, float a1, float b, float c) {
  asm("{.reg .b64 ra, rb, rc, rd;\n\tmov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %4};\n\tmov.b64 rc, {%5, %5};\n\t"
      "fma.rn.f32x2 rd, ra, rb, rc;\n\tmov.b64 {%0, %1}, rd;}"
      : "=f"(d0), "=f"(d1) : "f"(a0), "f"(a1), "f"(b), "f"(c));
}
// (a0, a1) += (b0, b1) (FADD2)
__device__ __forceinline__ void fadd2(float& a0, float& a1, float b0, float b1) {
  asm("{.reg .b64 ra, rb;\n\tmov.b64 ra, {%0, %1};\n\tmov.b64 rb, {%2, %3};\n\tadd.rn.f32x2 ra, ra, rb;\n\t"
      "mov.b64 {%0, %1}, ra;}"
      : "+f"(a0), "+f"(a1) : "f"(b0), "f"(b1));
}

// Named barriers 1 and 2 hand the special-function unit back and forth between the two softmax groups
// (bar.sync by the 128 threads that wait, bar.arrive by the 128 that release).
__device__ __forceinline__ void named_sync(int id) { asm volatile("bar.sync %0, 256;" :: "r"(id) : "memory"); }
__device__ __forceinline__ void named_sync_n(int id, int n) { asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void named_arrive(int id) { asm volatile("bar.arrive %0, 256;" :: "r"(id) : "memory"); }

template <typename T, int DP, bool kPaged = false, int kTiles = kPfMaxTiles>
__global__ void __launch_bounds__(kPfThreads, 1)
prefill_tc_kernel(const __grid_constant__ CUtensorMap tmq, const __grid_constant__ CUtensorMap tmk,
                  const __grid_constant__ CUtensorMap tmv, const __grid_constant__ CUtensorMap tmo,
                  const __grid_constant__ PrefillTilesT<kTiles> tiles,
                  const PrefillParams p) {
  constexpr int kAtoms = DP / 64;                 // 64-element (128-byte) column groups of a tile
  constexpr uint32_t kAtomBytes = 128 * 128;      // 128 rows x 128 bytes
  constexpr uint32_t kTileBytes = kAtoms * kAtomBytes;
  constexpr bool kBf16 = T::kId == PA_BF16;
  constexpr uint32_t kIdescQK = tc05::instr_desc_f16(kBf16, 128, 128, false, false);
  // head sizes 80 and 96 ride in the 128-wide tiles (TMA zero-fills the padding) but only their own
  // k-steps of Q.K^T and their own N columns of P.V are issued
  // V is the B operand of P.V: packed rows give an MN-major tile ([token][d]); cache pages give a K-major one
  // ([d][token], the page's own order)
  const uint32_t idesc_pv = tc05::instr_desc_f16(kBf16, 128, p.head_size, false, !kPaged);
  const int qk_steps = p.head_size / 16;
  constexpr uint32_t kTmS = 0, kTmO = 256, kTmCols = 512;   // S_i at 128 i, O_i at 256 + 128 i

  extern __shared__ unsigned char smem_raw[];
  const uint32_t base = (smem_addr(smem_raw) + 1023u) & ~1023u;
  const uint32_t sQ = base;                        // two tiles
  const uint32_t sK = sQ + 2 * kTileBytes;         // two stages
  const uint32_t sV = sK + 2 * kTileBytes;         // two stages
  const uint32_t sStg = sV + 2 * kTileBytes;       // output tile on its way out (one group at a time)
  const uint32_t sBar = sStg + kTileBytes;
  const uint32_t bar_q = sBar, bar_q1 = sBar + 224;   // Q_A / Q_B landed
  const uint32_t bar_kf = sBar + 8, bar_ke = sBar + 24, bar_vf = sBar + 40, bar_ve = sBar + 56;   // [2] each
  const uint32_t bar_s = sBar + 72, bar_p = sBar + 88, bar_o = sBar + 104;                        // [2] each
  const uint32_t bar_ph = sBar + 128;              // [2]: the first half of P_i (keys 0-63 of the tile) is in tensor memory
  const uint32_t bar_qe = sBar + 144, bar_qe1 = sBar + 232;   // tile A's / B's last Q.K^T of the item has run: Q_i may be overwritten
  const uint32_t bar_stg = sBar + 152;             // the staging buffer's previous user is done with it
  const uint32_t bar_cf = sBar + 160, bar_ce = sBar + 176;   // [2] each: next work item landed / read by everybody
  const uint32_t sClc = sBar + 192;                // [2] x 16 bytes: the scheduler's answers
  const uint32_t bar_vl = sBar + 240;              // kPaged: a V tile whose tail is zeroed lands here first
  const uint32_t bar_sf = sBar + 248;              // [2]: group i has filled the staging buffer with a whole output tile
  const uint32_t bar_sd = sBar + 264;              // [2]: ... and the store warp has sent it off and released the buffer
  const uint32_t bar_bl = sBar + 280;              // group B has taken its last turn of the item (see the ping-pong below)
  const uint32_t tm_slot = sBar + 120;
  const uint32_t sc_slot = sBar + 124;             // scale_log2, re-read after the ping-pong barrier (see below)

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  // ---- work items ------------------------------------------------------------------------------------
  // grid order (host's choice, see PrefillParams::tile_major): tiles fastest keeps the CTAs that run together on
  // few heads, so a head's K/V is read from HBM once and from L2 by its other query tiles
  struct Item {
    int head, tile, qtok0, qlen, q0, ktok0, klen, koff, nA, nB, nmax, kt0, slot;
  };
  auto decode = [&](int bx, int by) {
    Item it;
    it.head = p.tile_major ? by : bx;
    it.tile = p.tile_major ? bx : by;
    it.qtok0 = tiles.qtok0[it.tile]; it.qlen = tiles.qlen[it.tile]; it.q0 = tiles.q0[it.tile];
    it.ktok0 = tiles.ktok0[it.tile]; it.klen = tiles.klen[it.tile];
    it.koff = it.klen - it.qlen;                  // keys in front of the first query row
    // K/V tiles each query tile visits: everything up to and including the tile that holds the last key its
    // last row sees (plain prompts: its diagonal tile).  nB is nA or nA + 1.
    const uint32_t kr = static_cast<uint32_t>(tiles.krange[it.tile]);
    const int kt1 = static_cast<int>(kr >> 16);
    it.kt0 = static_cast<int>(kr & 0xffffu);
    it.nA = min(min(it.q0 + 127 + it.koff, it.klen - 1) / kPfCols + 1, kt1) - it.kt0;
    it.nB = (it.q0 + 128 < it.qlen) ? min(min(it.q0 + 255 + it.koff, it.klen - 1) / kPfCols + 1, kt1) - it.kt0 : 0;
    it.nmax = it.nB > it.nA ? it.nB : it.nA;
    const int si = tiles.slot[it.tile];
    it.slot = si < 0 ? -1 : (si >> 8) + it.head * (si & 255);
    return it;
  };
  // Item n + 1 of this CTA is the answer to the n-th request (slot n & 1), issued by the scheduler thread when
  // it starts item n.  Every thread that works on items calls next_item() once per item, in order: it waits for
  // the answer, reads it, releases the slot, and learns whether to go on.
  constexpr uint32_t kClcReaders = 256 + 1 + 1 + 1;   // softmax threads, issuer lane, producer lane, store lane
  auto request_next = [&](int n) {                // scheduler thread only
    const uint32_t slot = n & 1;
    mbar_wait(bar_ce + 8 * slot, ((n >> 1) & 1) ^ 1);   // first use of a slot passes immediately
    mbar_expect_tx(bar_cf + 8 * slot, 16);
    tc05::clc_try_cancel(sClc + 16 * slot, bar_cf + 8 * slot);
  };
  // whole_warp: all 32 lanes call together and lane 0 alone arrives for them -- it may do so only once every lane
  // has read the slot, or the scheduler could overwrite it under a lane that is still reading.
  auto next_item = [&](int n, Item& it, bool arrive, bool whole_warp = false) {
    const uint32_t slot = n & 1;
    mbar_wait(bar_cf + 8 * slot, (n >> 1) & 1);
    int bx, by;
    const bool ok = tc05::clc_read(sClc + 16 * slot, bx, by);
    tc05::fence_async_smem();                     // my read is ordered before the slot's next asynchronous write
    if (whole_warp) __syncwarp();
    if (arrive) mbar_arrive(bar_ce + 8 * slot);
    if (ok) it = decode(bx, by);
    return ok;
  };
  auto peek_item = [&](int n, Item& nx) {          // the same answer, early and without releasing the slot
    const uint32_t slot = n & 1;
    mbar_wait(bar_cf + 8 * slot, (n >> 1) & 1);
    int bx, by;
    const bool ok = tc05::clc_read(sClc + 16 * slot, bx, by);
    if (ok) nx = decode(bx, by);
    return ok;
  };
  Item it = decode(blockIdx.x, blockIdx.y);

  stamp(p.dbg, p.dbg && it.head == 0 && it.tile == 0 && tid == 0, 2, kTlIters - 2, 7);   // kernel entry
  if (tid == 0) {
    tc05::tma_prefetch_desc(&tmq); tc05::tma_prefetch_desc(&tmk); tc05::tma_prefetch_desc(&tmv);
    tc05::tma_prefetch_desc(&tmo);
    mbar_init(bar_q, 1); mbar_init(bar_q1, 1);
    mbar_init(bar_qe, 1); mbar_init(bar_qe1, 1);
    mbar_init(bar_stg, 1);
    mbar_init(bar_vl, 1);
    mbar_init(bar_bl, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_kf + 8 * s, 1); mbar_init(bar_ke + 8 * s, 1);
      mbar_init(bar_vf + 8 * s, 1); mbar_init(bar_ve + 8 * s, 1);
      mbar_init(bar_s + 8 * s, 1); mbar_init(bar_p + 8 * s, 128); mbar_init(bar_o + 8 * s, 1);
      mbar_init(bar_ph + 8 * s, 128);
      mbar_init(bar_cf + 8 * s, 1); mbar_init(bar_ce + 8 * s, kClcReaders);
      mbar_init(bar_sf + 8 * s, 128); mbar_init(bar_sd + 8 * s, 1);
    }
    asm volatile("st.shared.f32 [%0], %1;" :: "r"(sc_slot), "f"(p.scale_log2) : "memory");
    mbar_init_fence();
  }
  if (warp == 0) { __syncwarp(); tc05::tmem_alloc<kTmCols>(tm_slot); }
  tc05::fence_before_sync();
  __syncthreads();
  tc05::fence_after_sync();
  // Programmatic dependent launch: everything above touched only parameters and this CTA's own shared / tensor
  // memory, so it may run while the preceding kernel on the stream is still draining; from here on global memory
  // is read and written.
  pdl_launch_dependents();
  pdl_wait();
  uint32_t tmem;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(tm_slot) : "memory");

  if (kPaged && warp == 8) {
    // ===================== page producer (kPaged) =====================
    // One warp, no per-thread copies: a cache page already IS an operand tile of tcgen05.mma, so whole pages go
    // from HBM to shared memory by TMA.
    //   K page [d / 8][token][8]: for one 8-token group the D / 8 chunks are 128 contiguous bytes each = the
    //     8-row x 16-byte core matrices of the no-swizzle K-major layout.  Tensor map (64 elements = one core
    //     matrix, token group, chunk, block * heads + head), box 64 x 1 x D/8 x 1: group G of the tile lands at
    //     G * D * 16 bytes as [chunk][8 tokens][16 B] -> LBO (next K chunk) 128 B, SBO (next token group) D * 16.
    //   V page [d][token]: rows of B tokens = 16 / 32 / 64 bytes, K-major with N = d.  Tensor map (token, d,
    //     block * heads + head), box B x D x 1 with no / 32-byte / 64-byte swizzle: a page of 16 tokens is one MMA
    //     k-step (SBO 256 B), of 32 tokens two (start address + 32 B), two pages of 8 tokens make one (LBO = page).
    // Lane G issues K group G, lane l V page l; block numbers are fetched one tile ahead (one coalesced load) and
    // passed by shuffle.  Pages beyond the context are not loaded (their k-steps of P.V are not issued, their
    // logits are masked by selection); in the context's last k-step the tokens past the end are zeroed in shared
    // memory behind the TMA (never-written slots may hold NaN; 0 x NaN must not reach O).
    asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
    const int shB = p.block_shift, B = 1 << shB;
    const int D = p.head_size, H = p.num_heads;
    const int npg = kPfCols >> shB;                 // pages per 128-token tile
    const uint32_t kgrp_bytes = D * 16;             // one 8-token group of K: D / 8 core matrices
    const uint32_t vpage_bytes = D * B * 2;
    const int32_t* table = nullptr;
    int klen = 0;
    auto tile_blocks = [&](int t) {                 // lane l: block number of page l of tile t
      const int ptok = t * kPfCols + (lane << shB);
      return (lane < npg && ptok < klen) ? __ldg(table + (ptok >> shB)) : 0;
    };
    int g = 0;                                      // K/V tiles loaded so far: the rings run on across items
    int cB = 0;                                     // items with a tile B so far (phases of bar_q1 / bar_qe1)
    int ctail = 0;                                  // tiles with a zeroed tail so far (phase of bar_vl)
    for (int n = 0;; ++n) {
      if (lane == 0) {
        request_next(n);
        if (n > 0) mbar_wait(bar_qe, (n - 1) & 1);  // the previous item's last Q_A.K^T has read Q_A
        mbar_expect_tx(bar_q, kTileBytes);
#pragma unroll
        for (int a = 0; a < kAtoms; ++a)
          tc05::tma_load_3d(sQ + a * kAtomBytes, &tmq, a * 64, it.head, it.qtok0 + it.q0, bar_q);
      }
      __syncwarp();
      table = p.block_tables + static_cast<int64_t>(it.ktok0) * p.max_blocks;   // ktok0 = sequence index
      klen = it.klen;
      int next_blocks = tile_blocks(it.kt0);
      for (int j = 0; j < it.nmax; ++j, ++g) {
        const int s = g & 1;
        const uint32_t ph = ((g >> 1) & 1) ^ 1;     // first use of a stage passes immediately
        const int blocks = next_blocks;
        if (j + 1 < it.nmax) next_blocks = tile_blocks(it.kt0 + j + 1);
        const int nvalid = min(klen - (it.kt0 + j) * kPfCols, kPfCols);   // tokens of this tile inside the context (>= 1)
        const int ngrp = (nvalid + 7) >> 3, npages = (nvalid + B - 1) >> shB;
        const int row = blocks * H + it.head;       // lane l: (block, head) row of page l in both caches
        mbar_wait(bar_ke + 8 * s, ph);
        if (lane == 0) mbar_expect_tx(bar_kf + 8 * s, ngrp * kgrp_bytes);
        __syncwarp();
        {
          const int krow = __shfl_sync(0xffffffffu, row, (lane << 3) >> shB);   // lane G: the page of group G
          if (lane < ngrp)
            tc05::tma_load_4d(sK + s * kTileBytes + lane * kgrp_bytes, &tmk, 0, lane & ((B >> 3) - 1), 0, krow,
                              bar_kf + 8 * s);
        }
        if (j == 0 && it.nB && lane == 0) {         // Q_B behind K_0: it is free one iteration later than Q_A
          if (cB > 0) mbar_wait(bar_qe1, (cB - 1) & 1);
          mbar_expect_tx(bar_q1, kTileBytes);
#pragma unroll
          for (int a = 0; a < kAtoms; ++a)
            tc05::tma_load_3d(sQ + kTileBytes + a * kAtomBytes, &tmq, a * 64, it.head, it.qtok0 + it.q0 + 128, bar_q1);
        }
        if (j == 0 && it.nB) ++cB;
        __syncwarp();
        mbar_wait(bar_ve + 8 * s, ph);
        const bool tail = (nvalid & 15) != 0;       // the last k-step holds tokens past the context
        const uint32_t vbar = tail ? bar_vl : bar_vf + 8 * s;
        if (lane == 0) mbar_expect_tx(vbar, npages * vpage_bytes);
        __syncwarp();
        const uint32_t vdst = sV + s * kTileBytes;
        if (lane < npages) tc05::tma_load_3d(vdst + lane * vpage_bytes, &tmv, 0, 0, row, vbar);
        if (tail) {
          mbar_wait(bar_vl, ctail & 1);
          ++ctail;
          const int ks = (nvalid - 1) >> 4;         // the k-step to repair; vt of its 16 tokens are valid
          const int vt = nvalid - ks * 16;
          for (int idx = lane; idx < 2 * D; idx += 32) {
            const int d = idx >> 1, h = idx & 1;    // 16-byte chunk h (8 tokens) of row d of this k-step
            const int nv = min(max(vt - 8 * h, 0), 8);
            if (nv == 8) continue;
            uint32_t a;
            if (shB == 4) a = ks * vpage_bytes + d * 32 + ((h ^ ((d >> 2) & 1)) << 4);
            else if (shB == 5) a = (ks >> 1) * vpage_bytes + d * 64 + ((((ks & 1) * 2 + h) ^ ((d >> 1) & 3)) << 4);
            else a = (2 * ks + h) * vpage_bytes + d * 16;
            a += vdst;
            uint32_t w[4] = {0u, 0u, 0u, 0u};
            if (nv > 0) {
              asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "r"(a) : "memory");
#pragma unroll
              for (int e = 0; e < 4; ++e)
                w[e] = 2 * e >= nv ? 0u : (2 * e + 1 >= nv ? w[e] & 0xffffu : w[e]);
            }
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" :: "r"(a), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]) : "memory");
          }
          tc05::fence_async_smem();                 // generic-proxy stores -> visible to the MMA's reads
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_vf + 8 * s);
        }
      }
      if (!next_item(n, it, lane == 0, true)) break;
    }
  } else if (warp == 8) {
    // ===================== TMA producer =====================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
    if (lane == 0) {
      auto load_tile = [&](const CUtensorMap* map, uint32_t dst, int row0, uint32_t bar) {
#pragma unroll
        for (int a = 0; a < kAtoms; ++a) tc05::tma_load_3d(dst + a * kAtomBytes, map, a * 64, it.head, row0, bar);
      };
      int g = 0;                                    // K/V tiles loaded so far: the rings run on across items
      int cB = 0;                                   // items with a tile B so far (phases of bar_q1 / bar_qe1)
      for (int n = 0;; ++n) {
        request_next(n);
        // Q_A is free one iteration before the item ends (A's last Q.K^T), Q_B one later: loaded in that order
        // with K_0 in between, the next item's first Q_A.K_0^T can run right behind this item's last P_A.V
        if (n > 0) mbar_wait(bar_qe, (n - 1) & 1);
        mbar_expect_tx(bar_q, kTileBytes);
        load_tile(&tmq, sQ, it.qtok0 + it.q0, bar_q);
        for (int j = 0; j < it.nmax; ++j, ++g) {
          const int s = g & 1;
          const uint32_t ph = ((g >> 1) & 1) ^ 1;  // first use of a stage passes immediately
          mbar_wait(bar_ke + 8 * s, ph);
          mbar_expect_tx(bar_kf + 8 * s, kTileBytes);
          load_tile(&tmk, sK + s * kTileBytes, it.ktok0 + (it.kt0 + j) * kPfCols, bar_kf + 8 * s);
          if (j == 0 && it.nB) {
            if (cB > 0) mbar_wait(bar_qe1, (cB - 1) & 1);
            mbar_expect_tx(bar_q1, kTileBytes);
            load_tile(&tmq, sQ + kTileBytes, it.qtok0 + it.q0 + 128, bar_q1);
            ++cB;
          }
          mbar_wait(bar_ve + 8 * s, ph);
          mbar_expect_tx(bar_vf + 8 * s, kTileBytes);
          load_tile(&tmv, sV + s * kTileBytes, it.ktok0 + (it.kt0 + j) * kPfCols, bar_vf + 8 * s);
        }
        if (!next_item(n, it, true)) break;
      }
    }
  } else if (warp == 9) {
    // ===================== MMA issuer =====================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
    // The whole warp runs this loop converged (all lanes poll the barriers); one pre-elected lane issues.
    // Descriptors: one 32-bit add on the low word of a per-tile base (the start-address field, bytes >> 4).
    const bool leader = tc05::elect_one();
    const uint32_t tm = __shfl_sync(0xffffffffu, tmem, 0);
    auto desc = [](uint64_t base, uint32_t byte_off) { return base + (byte_off >> 4); };
    // kPaged: the operand tiles are the cache pages as the producer's TMA boxes left them (see there)
    const int shB = p.block_shift;
    const uint32_t vstep = p.head_size * 32;        // bytes of V per 16-token k-step (all d rows)
    const uint64_t dq = tc05::smem_desc_sw128(sQ, 16, 1024);
    const uint64_t dk = kPaged ? tc05::smem_desc(sK, 128, p.head_size * 16, 0) : tc05::smem_desc_sw128(sK, 16, 1024);
    const uint64_t dv = !kPaged ? tc05::smem_desc_sw128(sV, kAtomBytes, 1024)
                        : shB == 4 ? tc05::smem_desc(sV, 16, 256, 6)
                        : shB == 5 ? tc05::smem_desc(sV, 16, 512, 4)
                                   : tc05::smem_desc(sV, p.head_size * 16, 128, 0);
    auto issue_qk = [&](int i, int g) {             // S_i = Q_i . K^T, K in ring slot g
      const uint64_t qb = desc(dq, i * kTileBytes), kb = desc(dk, (g & 1) * kTileBytes);
#pragma unroll
      for (int kk = 0; kk < DP / 16; ++kk) {
        if (kk >= qk_steps) break;
        const uint32_t off = (kk >> 2) * kAtomBytes + (kk & 3) * 32;
        tc05::mma_ss_w(leader, tm + kTmS + 128 * i, desc(qb, off), desc(kb, kPaged ? kk * 256 : off), kIdescQK, kk > 0);
      }
    };
    // O_i (+)= P_i . V, keys 64 half .. 64 half + 63 of the tile; kmax: k-steps that hold keys of the context
    auto issue_pv = [&](int i, int g, bool first, int half, int kmax) {
      const uint64_t vb = desc(dv, (g & 1) * kTileBytes);
#pragma unroll
      for (int kk = half * (kPfCols / 32); kk < (half + 1) * (kPfCols / 32); ++kk) {
        if (kPaged && kk >= kmax) break;
        const uint32_t off = !kPaged ? kk * 2048 : shB == 5 ? (kk >> 1) * 2 * vstep + (kk & 1) * 32 : kk * vstep;
        tc05::mma_ts_w(leader, tm + kTmO + 128 * i, tm + kTmS + 128 * i + kk * 8, desc(vb, off), idesc_pv, !first || kk > 0);
      }
    };
    int g0 = 0;                                     // ring slot of this item's K_0 / V_0
    int cs[2] = {0, 0};                             // S tiles of A / B so far: the phases of bar_s, bar_p, bar_ph
    int cB = 0;                                     // items with a tile B so far (phase of bar_q1)
    int started = 0;                                // bit i: S_i of this item's tile 0 was issued during the previous item
    bool k0_released = false;                       // ... and K_0's stage already handed back behind both of them
    for (int n = 0;; ++n) {
      const bool tl = p.dbg && blockIdx.x == 0 && blockIdx.y == 0 && lane == 0;   // stamps follow CTA 0 through its items
      const int nA = it.nA, nB = it.nB, nmax = it.nmax;
      // first Q.K^T of the item, unless the previous item's loop has issued it already (see below)
      auto first_qk = [&](int i, int item, int cb, int gk) {
        mbar_wait(i ? bar_q1 : bar_q, (i ? cb : item) & 1);
        mbar_wait(bar_kf + 8 * (gk & 1), (gk >> 1) & 1);
        tc05::fence_after_sync();
        issue_qk(i, gk);
        tc05::mma_commit_w(leader, bar_s + 8 * i);
      };
      if (!(started & 1)) first_qk(0, n, cB, g0);
      stamp(p.dbg, tl, 2, g0, 6);
      if (nB && !(started & 2)) first_qk(1, n, cB, g0);
      if (!k0_released) tc05::mma_commit_w(leader, bar_ke + 8 * (g0 & 1));   // K_0 is free once these have run
      if (nA == 1) tc05::mma_commit_w(leader, bar_qe);
      if (nB == 1) tc05::mma_commit_w(leader, bar_qe1);
      const int cB_next = cB + (nB ? 1 : 0);
      started = 0;
      k0_released = false;
      Item nx;
      bool peeked = false, have_next = false;
      for (int j = 0; j < nmax; ++j) {
        const int g = g0 + j, s = g & 1;
        const bool more_k = j + 1 < nmax;
        mbar_wait(bar_vf + 8 * s, (g >> 1) & 1);
        const int kmax = kPaged ? (min(it.klen - (it.kt0 + j) * kPfCols, kPfCols) + 15) >> 4 : kPfCols / 16;
        stamp(p.dbg, tl, 2, g, 4);
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const int ni = i ? nB : nA;
          if (j < ni) {
            // P_i of tile j arrives in two halves: the first half's product runs while the softmax group is
            // still in the exponentials of the second
            const uint32_t php = (cs[i] + j) & 1;
            mbar_wait(bar_ph + 8 * i, php);
            tc05::fence_after_sync();
            stamp(p.dbg, tl, 2, g, 2 * i);
            issue_pv(i, g, j == 0, 0, kmax);
            mbar_wait(bar_p + 8 * i, php);
            tc05::fence_after_sync();
            issue_pv(i, g, false, 1, kmax);
            if (j + 1 < ni) {
              if (i == 0 || j + 1 >= nA) {          // K_{j+1} is needed from here on, not before the P.V above
                mbar_wait(bar_kf + 8 * (s ^ 1), ((g + 1) >> 1) & 1);
              }
              issue_qk(i, g + 1);
              tc05::mma_commit_w(leader, bar_s + 8 * i);
              if (j + 2 == ni) tc05::mma_commit_w(leader, i ? bar_qe1 : bar_qe);   // tile i's last Q.K^T: Q_i may go
            } else {
              tc05::mma_commit_w(leader, bar_o + 8 * i);
              // Tile i of this item is done on the tensor core.  Its group still has the output to write, but S_i
              // is free (the last P_i has been consumed, in order): the NEXT item's first Q_i.K_0^T goes in right
              // here, so that group i finds its S ready when it comes out of the epilogue instead of waiting for
              // the other tile's tail (B's diagonal tile and its softmax) to pass through the issuer first.
              if (!peeked) { have_next = peek_item(n, nx); peeked = true; }
              if (have_next && (i == 0 || nx.nB)) {
                first_qk(i, n + 1, cB_next, g0 + nmax);
                started |= 1 << i;
                stamp(p.dbg, tl, 2, g, 5);
              }
            }
            stamp(p.dbg, tl, 2, g, 2 * i + 1);
          }
        }
        tc05::mma_commit_w(leader, bar_ve + 8 * s);   // V_j and K_{j+1} are consumed by everything issued so far
        if (more_k) tc05::mma_commit_w(leader, bar_ke + 8 * (s ^ 1));
      }
      if (have_next && started == (nx.nB ? 3 : 1)) {  // both first products of the next item are in the queue
        tc05::mma_commit_w(leader, bar_ke + 8 * ((g0 + nmax) & 1));
        k0_released = true;
      }
      g0 += nmax;
      cs[0] += nA;
      cs[1] += nB;
      cB = cB_next;
      if (!next_item(n, it, lane == 0, true)) break;
    }
    // the last commits must have landed in this CTA's shared memory before the CTA may exit
    mbar_wait(bar_ve + 8 * ((g0 - 1) & 1), ((g0 - 1) >> 1) & 1);
  } else if (warp == 10) {
    // ===================== output store =====================
    // A whole 128-row output tile leaves as TMA stores from the staging buffer.  The softmax group only fills the
    // buffer (tensor memory -> registers -> swizzled shared memory), fences and arrives; this warp issues the
    // stores, waits until they have read the buffer and hands it on -- 1000 clk per tile that used to sit in the
    // group's own chain (the storing thread waited for the TMA read, and its warp with it).  Partial tiles and the
    // parts of split entries are written by the groups themselves, which then release the buffer.
    asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
    int cf[2] = {0, 0};                             // whole tiles handed over by group A / B so far (phase of bar_sf)
    for (int n = 0;; ++n) {
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        const int ni = g ? it.nB : it.nA;
        const int qi0 = it.q0 + 128 * g;
        if (ni > 0 && it.slot < 0 && it.qlen - qi0 >= 128) {
          mbar_wait(bar_sf + 8 * g, cf[g] & 1);
          ++cf[g];
          if (lane == 0) {
#pragma unroll
            for (int a = 0; a < kAtoms; ++a)
              if (a * 64 < p.head_size) tc05::tma_store_3d(&tmo, sStg + a * kAtomBytes, a * 64, it.head, it.qtok0 + qi0);
            tc05::tma_store_commit();
            tc05::tma_store_wait_read();
            mbar_arrive(bar_stg);
            mbar_arrive(bar_sd + 8 * g);
          }
          __syncwarp();
        }
      }
      if (!next_item(n, it, lane == 0, true)) break;
    }
  } else if (warp >= 10) {
    // register donors.  384 threads x 168 registers is the CTA's pool = 128 x 56 + 256 x 224; setmaxnreg must be
    // the same instruction in all four warps of a warpgroup (warps 8-11), whatever branch they sit in
    asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
  } else {
    // ===================== softmax groups =====================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 224;");
    const int grp = warp >> 2;                      // 0: tile A, 1: tile B
    const int r = tid & 127;                        // row inside the tile
    const uint32_t lane_base = tmem + (static_cast<uint32_t>((warp & 3) * 32) << 16);
    const uint32_t tS = lane_base + kTmS + 128 * grp, tO = lane_base + kTmO + 128 * grp;
    const int bar_mine = 1 + grp, bar_other = 2 - grp;
    int cs = 0, co = 0;                             // S tiles / finished tiles of this group so far (barrier phases)
    int ch = 0, chw = 0;                            // tiles handed to the store warp / of those, seen off (phase of bar_sd)
    for (int n = 0;; ++n) {
      const int nA = it.nA, nB = it.nB;
      const int ni = grp ? nB : nA;
      const int head = it.head, qtok0 = it.qtok0, qlen = it.qlen, klen = it.klen, koff = it.koff, kt0 = it.kt0;
      const int slot = it.slot;
      const int qi0 = it.q0 + 128 * grp;            // first row of this tile inside the prompt
      const bool dump = p.dbg && head == 0 && it.tile == 0 && grp == 0;
      const bool tl = p.dbg && blockIdx.x == 0 && blockIdx.y == 0 && r == 0;
      float m_run = -CUDART_INF_F, l_run = 0.f;
      // The exponentials of a tile keep the SM's special-function unit busy for as long as the tensor core
      // needs for that tile's two matrix products.  Left alone, the two groups drift into phase -- both in
      // their exponentials (sharing the unit), then both waiting for the tensor core -- and each unit idles
      // half of the time.  So the exponential section is a critical section that alternates A, B, A, B, ...:
      // while one group is in it the tensor core runs the other tile's P.V and next Q.K^T.
      // A's first tile of an item takes no turn (it may share the unit with the tail of B's previous item; waiting
      // for B to get there cost 2000 clk per item), from then on the two alternate.
      const bool pingpong = nB > 0;
      if (!pingpong && grp == 1 && r == 0) mbar_arrive(bar_bl);   // nothing of A's to consume in this item (see below)
      for (int j = 0; j < ni; ++j) {
        stamp(p.dbg, tl, grp, cs + j, 0);
        mbar_wait(bar_s + 8 * grp, (cs + j) & 1);
        tc05::fence_after_sync();
        stamp(p.dbg, tl, grp, cs + j, 1);
        uint32_t s[4][32];
#pragma unroll
        for (int c = 0; c < 4; ++c) tc05::tmem_ld32(tS + c * 32, s[c]);
        tc05::tmem_wait_ld();
        if (dump && j == 0) {
#pragma unroll
          for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int i = 0; i < 32; ++i) p.dbg[r * 128 + c * 32 + i] = __uint_as_float(s[c][i]);
        }
        // only tiles reaching past the last key of the tile's FIRST row need masking: the diagonal tile of a
        // plain prompt, at most the last two tiles when cached keys shift the diagonal off the tile grid.
        // A row wholly masked in such a tile (never tile 0) keeps its running maximum: all its p are 0.
        if ((kt0 + j + 1) * kPfCols - 1 > min(qi0 + koff, klen - 1)) {
          const int lim = min(qi0 + r + koff, klen - 1) - (kt0 + j) * kPfCols;   // last valid column of this row
          // the row's valid columns are a prefix: one 32-bit mask per chunk, tested bit by bit (ptxas moves the bits
          // into predicates seven at a time, R2P, instead of one compare per element)
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const int m = lim - 32 * c;             // last valid column inside this chunk
            const uint32_t bits = m >= 31 ? 0xffffffffu : (m < 0 ? 0u : (2u << m) - 1u);
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (!((bits >> i) & 1u)) s[c][i] = 0xff800000u;   // -inf
          }
        }
        // row maximum of the raw logits (scale > 0), four independent chains of 3-input max
        float mx[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          mx[c] = max3(__uint_as_float(s[c][0]), __uint_as_float(s[c][1]), __uint_as_float(s[c][2]));
#pragma unroll
          for (int i = 3; i + 1 < 32; i += 2) mx[c] = max3(mx[c], __uint_as_float(s[c][i]), __uint_as_float(s[c][i + 1]));
          mx[c] = fmaxf(mx[c], __uint_as_float(s[c][31]));
        }
        const float m_tile = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])) * p.scale_log2;
        // Lazy rescaling: the reference maximum only moves when the row maximum grew by more than 2^8; until then
        // probabilities may exceed 1 (by at most 2^8), which fp32 sums and 16-bit P hold without loss, and
        // O / l is unchanged because both carry the same reference.  O_i is quiescent here: S_j ready implies
        // the previous P.V of this tile completed (in-order tensor pipe).
        const bool move = m_tile > m_run + 8.f;     // always true on the first tile (m_run = -inf)
        if (__any_sync(0xffffffffu, move)) {
          const float alpha = move ? ex2(m_run - m_tile) : 1.f;
          if (move) { m_run = m_tile; l_run *= alpha; }
          if (j > 0) {
#pragma unroll
            for (int c = 0; c < DP / 32; ++c) {
              uint32_t o[32];
              tc05::tmem_ld32(tO + c * 32, o);
              tc05::tmem_wait_ld();
#pragma unroll
              for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
              tc05::tmem_st32(tO + c * 32, o);
            }
          }
        }
        const float neg_m = -m_run;
        float sum[4] = {0.f, 0.f, 0.f, 0.f};
        stamp(p.dbg, tl, grp, cs + j, 2);
        if (pingpong && j < nA && (grp == 1 || j > 0)) named_sync(bar_mine);   // B's extra (diagonal) tile runs after A's last: no turn to wait for
        // B has now consumed every arrival A made in this item.  A's first arrival of the NEXT item must not come
        // before that (two arrivals with no sync between them would complete the named barrier with nobody
        // waiting): B says so here, once per item, A checks in its epilogue before it leaves the item.  It used to be
        // a margin of a few thousand clocks -- A's epilogue and first tile against B's lag of half a period -- not a
        // guarantee.
        if (pingpong && grp == 1 && j == nA - 1 && r == 0) mbar_arrive(bar_bl);
        // ptxas moves arithmetic (and MUFU) across BAR.SYNC freely; what it cannot move is a value loaded
        // after the barrier, so the section's multiplier is re-read from shared memory here
        float sc;
        asm volatile("ld.volatile.shared.f32 %0, [%1];" : "=f"(sc) : "r"(sc_slot) : "memory");
        stamp(p.dbg, tl, grp, cs + j, 3);
#pragma unroll
        for (int c2 = 0; c2 < 2; ++c2) {
          uint32_t pk[32];
#pragma unroll
          for (int h = 0; h < 2; ++h)
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
              float a0, a1;
              ffma2(a0, a1, __uint_as_float(s[c2 * 2 + h][i]), __uint_as_float(s[c2 * 2 + h][i + 1]), sc, neg_m);
              const float p0 = ex2_pinned(a0), p1 = ex2_pinned(a1);
              fadd2(sum[(i >> 1) & 1 ? 2 : 0], sum[(i >> 1) & 1 ? 3 : 1], p0, p1);
              pk[h * 16 + (i >> 1)] = pack2<T>(p0, p1);
            }
          tc05::tmem_st32(tS + c2 * 32, pk);
          if (c2 == 0) {                              // keys 0-63 of P: the issuer starts P.V on them
            tc05::tmem_wait_st();
            tc05::fence_before_sync();
            mbar_arrive(bar_ph + 8 * grp);
            stamp(p.dbg, tl, grp, cs + j, 7);
          }
        }
        stamp(p.dbg, tl, grp, cs + j, 4);
        // hand over: B after its tile j lets A into tile j + 1 (if A has one); A after tile j lets B into tile j.
        // Every arrive is matched by exactly one later sync of the other group before the next arrive on the same
        // barrier can happen (two back-to-back arrives would complete the barrier with nobody waiting); the
        // counts balance within an item (A syncs nA - 1 times, B arrives nA - 1 times; B syncs nA times).
        if (pingpong && (grp == 0 || j + 1 < nA)) named_arrive(bar_other);
        l_run += (sum[0] + sum[1]) + (sum[2] + sum[3]);
        tc05::tmem_wait_st();
        tc05::fence_before_sync();
        mbar_arrive(bar_p + 8 * grp);
        stamp(p.dbg, tl, grp, cs + j, 5);
      }
      cs += ni;
      // ---- epilogue: O / l -> out ---------------------------------------------------------
      // Every thread holds one row of O; stored from here, a warp's 16-byte stores would land in 32 different
      // rows (32 wavefronts per instruction: measured 4400 clk for the tile).  So the tile is staged in shared
      // memory in the 128-byte-swizzled layout and leaves as one TMA store per 64-column group; a prompt's last,
      // partial tile (rows past the prompt belong to the next one) leaves as row-contiguous 16-byte stores of
      // its valid rows instead.  The staging buffer is used in turns, A and B of item 0, A and B of item 1, ...
      // (use u = 2 n + group waits for the u-th release); a group without a tile only passes the turn on.
      bool handed = false;                            // this item's tile went to the store warp, which releases the buffer
      // The turn barrier tells phases apart by parity only.  A release that the store warp makes on this group's
      // behalf comes whenever the TMA read is over -- possibly after this group has run through a short next item --
      // and a wait that starts before it would take the phase before for the one it wants.  So a handed tile is
      // first seen off (its own barrier, one phase per handed tile, consumed in order), then the turn is awaited.
      if (r == 0 && ch > chw) { mbar_wait(bar_sd + 8 * grp, chw & 1); ++chw; }
      if (r == 0) mbar_wait(bar_stg, (grp & 1) ^ 1);   // u & 1 == group
      if (ni > 0) {
        mbar_wait(bar_o + 8 * grp, co & 1);
        ++co;
        tc05::fence_after_sync();
        if (grp == 0) mbar_wait(bar_bl, n & 1);       // B has taken its last turn of this item (one phase per item)
        stamp(p.dbg, tl, grp, cs - 1, 6);             // O ready (the item's last P.V has run)
        if (slot >= 0) {
          // one part of a split context: O as it stands (relative to m_run, not normalised), m_run and l_run go to
          // the workspace; column-major, so that the 32 rows of a warp store 128 contiguous bytes per column
          float* ws = p.split_ws + (static_cast<int64_t>(slot) * 2 + grp) * pf_part_floats(DP);
#pragma unroll
          for (int c = 0; c < DP / 32; ++c) {
            uint32_t o[32];
            tc05::tmem_ld32(tO + c * 32, o);
            tc05::tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) ws[(c * 32 + i) * 128 + r] = __uint_as_float(o[i]);
          }
          ws[128 * DP + r] = m_run;
          ws[128 * DP + 128 + r] = l_run;
          // every warp of the group is past its bar_bl wait before row 0's thread passes the turn on: bar_bl tells
          // its phases apart by parity only, and B's arrival for the NEXT item must not come while a warp of A still
          // waits for this one (the whole-item path below has the same barrier ahead of its stores)
          named_sync_n(3 + grp, 128);
        } else {
        const int D = p.head_size;
        const float inv = 1.f / l_run;
        named_sync_n(3 + grp, 128);                   // row 0's thread has seen the buffer released
#pragma unroll
        for (int c = 0; c < DP / 32; ++c) {
          uint32_t o[32];
          tc05::tmem_ld32(tO + c * 32, o);
          tc05::tmem_wait_ld();
          if (dump) {
#pragma unroll
            for (int i = 0; i < 32; ++i) p.dbg[128 * 128 + r * DP + c * 32 + i] = __uint_as_float(o[i]);
            if (c == 0) { p.dbg[128 * 128 + 128 * DP + r] = l_run; p.dbg[128 * 128 + 128 * DP + 128 + r] = m_run; }
          }
#pragma unroll
          for (int v = 0; v < 4; ++v) {               // 8 elements = one 16-byte chunk
            const uint32_t w0 = pack2<T>(__uint_as_float(o[v * 8 + 0]) * inv, __uint_as_float(o[v * 8 + 1]) * inv);
            const uint32_t w1 = pack2<T>(__uint_as_float(o[v * 8 + 2]) * inv, __uint_as_float(o[v * 8 + 3]) * inv);
            const uint32_t w2 = pack2<T>(__uint_as_float(o[v * 8 + 4]) * inv, __uint_as_float(o[v * 8 + 5]) * inv);
            const uint32_t w3 = pack2<T>(__uint_as_float(o[v * 8 + 6]) * inv, __uint_as_float(o[v * 8 + 7]) * inv);
            const uint32_t chunk = (c & 1) * 4 + v;
            const uint32_t dst = sStg + (c >> 1) * kAtomBytes + r * 128 + ((chunk ^ (r & 7)) << 4);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" :: "r"(dst), "r"(w0), "r"(w1), "r"(w2), "r"(w3) : "memory");
          }
        }
        const int valid = min(128, qlen - qi0);       // rows of this tile inside the prompt
        if (valid == 128) {
          tc05::fence_async_smem();                   // my stores -> visible to the TMA unit
          mbar_arrive(bar_sf + 8 * grp);              // the store warp takes it from here and releases the buffer
          handed = true;
          ++ch;
        } else {
          named_sync_n(3 + grp, 128);
          const int cpr = D / 8;                      // 16-byte chunks per row: 8, 10, 12 or 16
          char* obase = reinterpret_cast<char*>(p.out) +
                        ((static_cast<int64_t>(qtok0) + qi0) * p.out_stride + static_cast<int64_t>(head) * D) * 2;
          for (int idx = r; idx < valid * cpr; idx += 128) {
            const int row = idx / cpr, chunk = idx - row * cpr;
            const uint32_t src = sStg + (chunk >> 3) * kAtomBytes + row * 128 + (((chunk & 7) ^ (row & 7)) << 4);
            uint4 w;
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(w.x), "=r"(w.y), "=r"(w.z), "=r"(w.w) : "r"(src) : "memory");
            *reinterpret_cast<uint4*>(obase + static_cast<int64_t>(row) * p.out_stride * 2 + chunk * 16) = w;
          }
          named_sync_n(3 + grp, 128);                 // everybody has read its chunks
        }
        }
      }
      if (r == 0 && !handed) mbar_arrive(bar_stg);
      if (ni > 0) stamp(p.dbg, tl, 3, cs - 1, grp);     // output written or handed to the store warp (row 3 of the stamp table)
      if (!next_item(n, it, true)) break;
    }
  }
  tc05::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc05::tmem_free<kTmCols>(tmem);
}

}  // namespace pa
