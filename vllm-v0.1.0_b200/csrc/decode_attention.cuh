// Paged decode attention for sm_100a.
//
// Replaces the reference kernel single_query_cached_kv_attention_kernel
// (csrc/attention/attention_kernels.cu:67-314 of the reference) behind the
// same operator contract; the design is not derived from it:
//
//   * one CTA per (head, sequence, context split); its W warps take whole
//     pages round-robin and never synchronise with each other until the end;
//   * every warp runs a single-pass (online) softmax over its pages, so no
//     logits row lives in shared memory and context length is unbounded;
//   * pages -- the contiguous D*B*sizeof(T) bytes of one (block, head) in the
//     key cache and in the value cache -- are fetched by the TMA engine with
//     cp.async.bulk into a per-warp ring of stages, completion signalled on an
//     mbarrier; the block-table row is staged in shared memory first so the
//     indirection never sits on the critical path;
//   * 16-bit types never go through a convert: products use the sm_100a
//     mixed-precision FMA (FHFMA: fp16/bf16 operands, exact product, fp32
//     accumulate); the warp max is one CREDUX.MAX.F32;
//   * tokens past context_len are masked by selection, never by multiplying,
//     so garbage (NaN/Inf) in unwritten slots of the last block is harmless;
//   * every offset is 64-bit.
//
// A second load path (PageSrc = GmemPage) reads pages straight from global
// memory with streaming 16-byte loads; it shares all of the arithmetic.
#pragma once

#include <math_constants.h>

#include "pa_common.cuh"

namespace pa {

// ---------------------------------------------------------------------------
// compile-time geometry of one (T, head_size D, block_size B) combination
// ---------------------------------------------------------------------------
template <typename T, int D, int B>
struct Geom {
  static constexpr int kS = T::kSize;
  static constexpr int kX = 16 / kS;                   // elements per 16-byte vector
  // key page  [D/X][B] vectors: vector (c, t) = elements c*X..c*X+X-1 of token t
  static constexpr int kNch = D / kX;                  // chunks per token
  static constexpr int kTl = B < 16 ? B : 16;          // lanes along tokens
  static constexpr int kG = kWarp / kTl;               // lane groups along chunks
  static constexpr int kTpl = B / kTl;                 // tokens per lane
  static constexpr int kCpl = (kNch + kG - 1) / kG;    // chunks per lane
  static constexpr bool kKExact = (kNch % kG) == 0;
  // value page [D][B/X] vectors: vector (d, tb) = tokens tb*X..tb*X+X-1 of row d
  static constexpr int kVpr = B / kX;                  // vectors per row
  static constexpr int kRpi = kWarp / kVpr;            // rows per warp iteration
  static constexpr int kNvi = (D + kRpi - 1) / kRpi;   // iterations = accumulators per lane
  static constexpr bool kVExact = (D % kRpi) == 0;
  static constexpr int kPageBytes = D * B * kS;
  static constexpr int kPbufBytes = B * kS;            // one page of probabilities
  static_assert(D % kX == 0, "head_size must be a multiple of 16 bytes");
  static_assert(B % kX == 0, "block_size must be a multiple of 16 bytes");
};

// page handles ---------------------------------------------------------------
struct SmemPage {
  uint32_t addr;
  __device__ __forceinline__ uint4 ld(int vec) const {
    uint4 r;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "r"(addr + static_cast<uint32_t>(vec) * 16u));
    return r;
  }
};
struct GmemPage {
  const uint4* ptr;
  __device__ __forceinline__ uint4 ld(int vec) const { return ldg_stream(ptr + vec); }
};

// 16-byte dot product step ----------------------------------------------------
template <typename T>
__device__ __forceinline__ void dot16(const uint4& a, const uint4& b, float& lo, float& hi) {
  lo = mixfma_lo<T>(a.x, b.x, lo);
  hi = mixfma_hi<T>(a.x, b.x, hi);
  lo = mixfma_lo<T>(a.y, b.y, lo);
  hi = mixfma_hi<T>(a.y, b.y, hi);
  lo = mixfma_lo<T>(a.z, b.z, lo);
  hi = mixfma_hi<T>(a.z, b.z, hi);
  lo = mixfma_lo<T>(a.w, b.w, lo);
  hi = mixfma_hi<T>(a.w, b.w, hi);
}
template <>
__device__ __forceinline__ void dot16<F32>(const uint4& a, const uint4& b, float& lo, float& hi) {
  lo = fmaf(__uint_as_float(a.x), __uint_as_float(b.x), lo);
  hi = fmaf(__uint_as_float(a.y), __uint_as_float(b.y), hi);
  lo = fmaf(__uint_as_float(a.z), __uint_as_float(b.z), lo);
  hi = fmaf(__uint_as_float(a.w), __uint_as_float(b.w), hi);
}

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Probabilities are kept relative to the running maximum (<= 1).  For fp16 they
// are stored scaled by 2^14 so that the 16-bit copy used in the P.V products
// keeps tokens down to 2^-38 of the maximum; the row sum carries the same
// factor, so it cancels in the final division.
template <typename T> struct ProbScale { static constexpr float kLog2 = 0.f; };
template <> struct ProbScale<F16> { static constexpr float kLog2 = 14.f; };

// running state of one warp ----------------------------------------------------
template <typename T, int D, int B>
struct WarpState {
  using Gm = Geom<T, D, B>;
  float m;              // running maximum of scaled logits (log2 domain)
  float l;              // this lane's share of the running sum of probabilities
  float acc[Gm::kNvi];  // this lane's rows of the running P.V
};

// One page: logits for its B tokens, online-softmax update, P.V accumulate.
// nvalid = number of tokens of this page inside the context (B unless last).
template <typename T, int D, int B, typename Page>
__device__ __forceinline__ void process_page(const Page& kp, const Page& vp,
                                             const uint4 (&q)[Geom<T, D, B>::kCpl],
                                             WarpState<T, D, B>& st, float scale_log2,
                                             int nvalid, uint32_t pbuf, int lane) {
  using Gm = Geom<T, D, B>;
  const int tl = lane % Gm::kTl;
  const int g = lane / Gm::kTl;

  // ---- q . k -------------------------------------------------------------
  float lo[Gm::kTpl], hi[Gm::kTpl];
#pragma unroll
  for (int tt = 0; tt < Gm::kTpl; ++tt) { lo[tt] = 0.f; hi[tt] = 0.f; }
#pragma unroll
  for (int cc = 0; cc < Gm::kCpl; ++cc) {
    const int c = cc * Gm::kG + g;
    if (Gm::kKExact || c < Gm::kNch) {
#pragma unroll
      for (int tt = 0; tt < Gm::kTpl; ++tt) {
        const uint4 kv = kp.ld(c * B + tt * Gm::kTl + tl);
        dot16<T>(q[cc], kv, lo[tt], hi[tt]);
      }
    }
  }
  float s[Gm::kTpl];
  float pm = -CUDART_INF_F;
#pragma unroll
  for (int tt = 0; tt < Gm::kTpl; ++tt) {
    float v = lo[tt] + hi[tt];
#pragma unroll
    for (int o = Gm::kTl; o < kWarp; o <<= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    v *= scale_log2;
    // masked tokens are replaced, not multiplied: their logit may be NaN/Inf
    if (tt * Gm::kTl + tl >= nvalid) v = -CUDART_INF_F;
    s[tt] = v;
    pm = fmaxf(pm, v);
  }

  // ---- online softmax -----------------------------------------------------
  pm = warp_max(pm);
  if (pm > st.m) {  // warp-uniform
    const float alpha = fast_exp2(st.m - pm);
    st.m = pm;
    st.l *= alpha;
#pragma unroll
    for (int i = 0; i < Gm::kNvi; ++i) st.acc[i] *= alpha;
  }
  const float off = ProbScale<T>::kLog2 - st.m;
#pragma unroll
  for (int tt = 0; tt < Gm::kTpl; ++tt) {
    const float p = fast_exp2(s[tt] + off);  // exp2(-inf) == 0 for masked tokens
    const int t = tt * Gm::kTl + tl;
    if constexpr (T::kSize == 2) {
      const uint32_t pk = pack2<T>(p, 0.f);
      if (g == 0) {
        st.l += unpack_lo<T>(pk);  // sum what the products will actually use
        asm volatile("st.shared.u16 [%0], %1;"
                     :: "r"(pbuf + t * 2), "h"(static_cast<uint16_t>(pk)) : "memory");
      }
    } else {
      if (g == 0) {
        st.l += p;
        asm volatile("st.shared.f32 [%0], %1;" :: "r"(pbuf + t * 4), "f"(p) : "memory");
      }
    }
  }
  __syncwarp();

  // ---- p . v --------------------------------------------------------------
  const int tb = lane % Gm::kVpr;
  uint4 pv;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(pv.x), "=r"(pv.y), "=r"(pv.z), "=r"(pv.w)
               : "r"(pbuf + tb * 16) : "memory");
  if (nvalid >= B) {
#pragma unroll
    for (int i = 0; i < Gm::kNvi; ++i) {
      if (Gm::kVExact || i * Gm::kRpi + lane / Gm::kVpr < D) {
        const uint4 vv = vp.ld(i * kWarp + lane);
        float hi2 = 0.f;
        dot16<T>(pv, vv, st.acc[i], hi2);
        st.acc[i] += hi2;
      }
    }
  } else {
    // last block of the context: zero the value elements of masked tokens
    uint32_t msk[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if constexpr (T::kSize == 2) {
        const int t0 = tb * 8 + 2 * j;
        msk[j] = (t0 < nvalid ? 0x0000ffffu : 0u) | (t0 + 1 < nvalid ? 0xffff0000u : 0u);
      } else {
        msk[j] = (tb * 4 + j < nvalid) ? 0xffffffffu : 0u;
      }
    }
#pragma unroll
    for (int i = 0; i < Gm::kNvi; ++i) {
      if (Gm::kVExact || i * Gm::kRpi + lane / Gm::kVpr < D) {
        uint4 vv = vp.ld(i * kWarp + lane);
        vv.x &= msk[0]; vv.y &= msk[1]; vv.z &= msk[2]; vv.w &= msk[3];
        float hi2 = 0.f;
        dot16<T>(pv, vv, st.acc[i], hi2);
        st.acc[i] += hi2;
      }
    }
  }
}

// ---------------------------------------------------------------------------
// kernel
// ---------------------------------------------------------------------------
struct DecodeParams {
  void* out;                  // [S][H][D] T
  const void* q;              // [S][H][D] T, row stride q_stride elements
  const void* k_cache;
  const void* v_cache;
  const int32_t* block_tables;  // [S][MB]
  const int32_t* context_lens;  // [S]
  float* ws_part;             // [S][H][nsplit][D + 2]   (nsplit > 1)
  uint32_t* ws_ticket;        // [S][H], zero between calls (nsplit > 1)
  int64_t q_stride;
  float scale_log2;           // scale * log2(e)
  int32_t num_heads;
  int32_t max_blocks_per_seq;
  int32_t pages_per_split;
  int32_t num_splits;
  int32_t num_stages;
  int32_t q_vec_ok;           // q rows are 16-byte aligned
  int32_t num_seqs;
  int32_t group_rows;         // kGroup: sibling detection over aligned groups of 4 sequences; 1: off
};

constexpr int kGroup = 4;          // sequences per sibling group (= warps per CTA in group mode)
constexpr int kGroupMinShared = 4; // shared full pages needed before a group is processed together

// dynamic shared memory map (bytes), W = warps per CTA:
//   [0, tbl_bytes)                       int32 block numbers of this split
//   + W * (D + 2) floats                 end-of-CTA merge area
//   + W * 2 * kPbufBytes(aligned 16)     probability pages (double-buffered)
//   + W * S * 8                          mbarriers            (staged only)
//   + W * S * 2 * kPageBytes             stage ring, 128-byte aligned (staged only)
template <typename T, int D, int B>
struct SmemMap {
  using Gm = Geom<T, D, B>;
  static constexpr int kPbufStride = ((2 * Gm::kPbufBytes + 15) / 16) * 16;
  int tbl, red, pbuf, bars, ring, total;
  __host__ __device__ SmemMap(int pages_per_split, int warps, int stages, bool staged) {
    int o = 0;
    tbl = o;  o += ((pages_per_split * 4 + 15) / 16) * 16;
    red = o;  o += warps * (D + 2) * 4;  o = ((o + 15) / 16) * 16;
    pbuf = o; o += warps * kPbufStride;
    bars = o; o += staged ? warps * stages * 8 : 0;
    o = ((o + 127) / 128) * 128;
    ring = o; o += staged ? warps * stages * 2 * Gm::kPageBytes : 0;
    total = o;
  }
};

template <typename T, int D, int B, bool kStaged, bool kGroups = false>
__global__ void __launch_bounds__(256)
paged_decode_kernel(const DecodeParams p) {
  using Gm = Geom<T, D, B>;
  extern __shared__ __align__(128) unsigned char smem[];

  const int head = blockIdx.x;
  const int seq = blockIdx.y;
  const int split = blockIdx.z;
  const int warp = threadIdx.x / kWarp;
  const int lane = threadIdx.x % kWarp;
  const int nwarps = blockDim.x / kWarp;
  const int H = p.num_heads;

  const SmemMap<T, D, B> map(p.pages_per_split, nwarps, p.num_stages, kStaged);
  int32_t* tbl = reinterpret_cast<int32_t*>(smem + map.tbl);
  float* red = reinterpret_cast<float*>(smem + map.red);

  const int first_page = split * p.pages_per_split;
  // ---- barriers first: nothing here depends on memory, so under programmatic dependent launch this part
  //      (and the CTA's launch itself) overlaps the tail of the preceding kernel on the stream
  const uint32_t bars = smem_addr(smem + map.bars) + warp * p.num_stages * 8;
  const uint32_t ring = smem_addr(smem + map.ring) + warp * p.num_stages * 2 * Gm::kPageBytes;
  if constexpr (kStaged) {
    if (lane == 0) {
      for (int s = 0; s < p.num_stages; ++s) mbar_init(bars + s * 8, 1);
      mbar_init_fence();
    }
  }
  pdl_launch_dependents();
  pdl_wait();
  const uint64_t keep = l2_policy_evict_last();      // tables, lengths, q: re-read by every call
  const uint64_t once = l2_policy_evict_first();     // KV pages: read exactly once per call
  // ---- stage the block-table slice (does not wait for the context length: the row is
  //      max_blocks_per_seq entries long whatever the context, so both loads fly together)
  int avail = p.max_blocks_per_seq - first_page;
  avail = avail < 0 ? 0 : (avail > p.pages_per_split ? p.pages_per_split : avail);
  {
    const int32_t* row = p.block_tables + static_cast<int64_t>(seq) * p.max_blocks_per_seq + first_page;
    for (int i = threadIdx.x; i < avail; i += blockDim.x) tbl[i] = ldg_i32_hint(row + i, keep);
  }
  // sibling detection (kGroups): first page and context length of every sequence of this aligned group
  constexpr bool kGroupPath = kGroups && kStaged && T::kSize == 2 && B % 16 == 0;
  const bool grouping = kGroupPath && p.group_rows > 1;
  const int r0 = grouping ? seq - seq % kGroup : seq;                       // first sequence of the group
  const int nrows = grouping ? min(kGroup, p.num_seqs - r0) : 1;
  __shared__ int s_first[kGroup];
  __shared__ int s_ctx[kGroup];
  if (grouping && threadIdx.x < nrows) {
    s_first[threadIdx.x] = avail > 0 ? ldg_i32_hint(p.block_tables + static_cast<int64_t>(r0 + threadIdx.x) * p.max_blocks_per_seq + first_page, keep) : -1 - static_cast<int>(threadIdx.x);
    s_ctx[threadIdx.x] = ldg_i32_hint(p.context_lens + r0 + threadIdx.x, keep);
  }
  const int ctx = ldg_i32_hint(p.context_lens + seq, keep);
  // never walk past the block-table row or the split schedule, whatever context_lens says
  const int pages_true = ctx > 0 ? (ctx + B - 1) / B : 0;
  const int pages_cap = min(p.max_blocks_per_seq, p.pages_per_split * p.num_splits);
  const int total_pages = min(pages_true, pages_cap);
  int npages = total_pages - first_page;
  npages = npages < 0 ? 0 : (npages > p.pages_per_split ? p.pages_per_split : npages);

  if (p.num_splits == 1 && npages == 0) {
    // empty context: the reference leaves an all-zero row (attention_kernels.cu:142,213)
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
      const int64_t o = (static_cast<int64_t>(seq) * H + head) * D + d;
      if constexpr (T::kSize == 2) reinterpret_cast<uint16_t*>(p.out)[o] = 0;
      else reinterpret_cast<float*>(p.out)[o] = 0.f;
    }
    return;
  }

  // ---- q row -> registers ------------------------------------------------------
  uint4 q[Gm::kCpl];
  auto load_q = [&](int row_seq) {
    const int g = lane / Gm::kTl;
    const char* qrow = reinterpret_cast<const char*>(p.q) +
                       (static_cast<int64_t>(row_seq) * p.q_stride + static_cast<int64_t>(head) * D) * T::kSize;
#pragma unroll
    for (int cc = 0; cc < Gm::kCpl; ++cc) {
      const int c = cc * Gm::kG + g;
      q[cc] = make_uint4(0u, 0u, 0u, 0u);
      if (Gm::kKExact || c < Gm::kNch) {
        if (p.q_vec_ok) {
          q[cc] = ldg_v4_hint(qrow + c * 16, keep);
        } else {
          const uint16_t* h16 = reinterpret_cast<const uint16_t*>(qrow + c * 16);
          uint32_t w[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) w[j] = static_cast<uint32_t>(h16[2 * j]) | (static_cast<uint32_t>(h16[2 * j + 1]) << 16);
          q[cc] = make_uint4(w[0], w[1], w[2], w[3]);
        }
      }
    }
  };
  load_q(seq);
  __syncthreads();  // tbl visible to every warp; barrier inits visible

  WarpState<T, D, B> st;
  st.m = -CUDART_INF_F;
  st.l = 0.f;
#pragma unroll
  for (int i = 0; i < Gm::kNvi; ++i) st.acc[i] = 0.f;

  const uint32_t pbuf0 = smem_addr(smem + map.pbuf) + warp * SmemMap<T, D, B>::kPbufStride;
  const char* kbase = reinterpret_cast<const char*>(p.k_cache) + static_cast<int64_t>(head) * Gm::kPageBytes;
  const char* vbase = reinterpret_cast<const char*>(p.v_cache) + static_cast<int64_t>(head) * Gm::kPageBytes;
  const int64_t block_bytes = static_cast<int64_t>(H) * Gm::kPageBytes;
  const float eps = 1e-6f * exp2f(ProbScale<T>::kLog2);  // reference: 1/(sum + 1e-6), attention_kernels.cu:213

  // ---- sibling groups ----------------------------------------------------------------
  // Sequences of one request (parallel sampling, beam search) sit next to each other in the batch and
  // share the physical pages of their common prefix.  The CTAs of an aligned group of kGroup sequences
  // all evaluate the same test on the same staged tables: how many leading pages of this split are the
  // same block in every sequence and lie wholly inside every context.  If enough are, the first CTA of
  // the group serves the whole group -- one warp per sequence, every shared page fetched from HBM once
  // and read by all warps from shared memory -- and the others leave.
  if constexpr (kGroupPath) {
    bool maybe = grouping && nrows > 1;                                              // cheap reject: first pages differ
    for (int r = 1; maybe && r < nrows; ++r) maybe = s_first[0] == s_first[r];
    if (maybe) {
      __shared__ int s_shared;
      if (threadIdx.x == 0) s_shared = 0x7fffffff;
      __syncthreads();
      int limit = avail;
      for (int r = 0; r < nrows; ++r) {
        int full = s_ctx[r] / B - first_page;                                      // pages wholly inside the context
        full = full < 0 ? 0 : full;
        limit = limit < full ? limit : full;
      }
      int first_diff = limit;
      const int32_t* rows0 = p.block_tables + static_cast<int64_t>(r0) * p.max_blocks_per_seq + first_page;
      for (int i = threadIdx.x; i < limit && i < first_diff; i += blockDim.x) {
        const int32_t b0 = tbl[i];
        bool same = true;
        for (int r = 0; r < nrows; ++r)
          if (r0 + r != seq) same = same && (__ldg(rows0 + static_cast<int64_t>(r) * p.max_blocks_per_seq + i) == b0);
        if (!same) first_diff = i;
      }
      atomicMin(&s_shared, first_diff);
      __syncthreads();
      const int P = s_shared;                                                      // shared pages of this split
      if (P >= kGroupMinShared) {
        if (seq != r0) return;                                                     // the group's first CTA does it all
        // -------- group mode ------------------------------------------------------------------
        // Phase 1, shared pages: the warps take them round-robin, each page once for ALL sequences of
        // the group on the tensor cores: S^T[token][seq] = K[token][:] . Q^T (K straight from its x-packed
        // page as the A operand), column-wise online softmax, O^T[d][seq] += V^T[d][token] . P^T.
        // Phase 2, the pages that are a sequence's own: warp w <-> sequence r0 + w, as in the plain path.
        constexpr int KD = D / 16;            // k-steps of Q.K^T = 16-row tiles of O^T
        constexpr int TT = B / 16;            // 16-token tiles per page
        const int S = p.num_stages;
        const uint32_t full0 = smem_addr(smem + map.bars);                         // stage s of warp w: full0 + 8 (w S + s)
        const uint32_t ring0 = smem_addr(smem + map.ring);
        const bool active = warp < nrows;
        const int my_seq = r0 + (active ? warp : 0);
        const int my_ctx = s_ctx[active ? warp : 0];
        const int32_t* my_row = p.block_tables + static_cast<int64_t>(my_seq) * p.max_blocks_per_seq + first_page;
        auto issue_stage = [&](int64_t blk, int stage) {
          const uint32_t bar = full0 + stage * 8;
          const uint32_t dst = ring0 + stage * 2 * Gm::kPageBytes;
          mbar_expect_tx(bar, 2 * Gm::kPageBytes);
          bulk_g2s_hint(dst, kbase + blk * block_bytes, Gm::kPageBytes, bar, once);
          bulk_g2s_hint(dst + Gm::kPageBytes, vbase + blk * block_bytes, Gm::kPageBytes, bar, once);
        };
        const int n1 = P > warp ? (P - warp + nwarps - 1) / nwarps : 0;            // shared pages of this warp
        if (lane == 0)
          for (int n = 0; n < S && n < n1; ++n) issue_stage(tbl[warp + n * nwarps], warp * S + n);

        // B fragments of Q^T: lane holds seq n = lane / 4, elements d = 16 kk + 2 (lane % 4) + {0, 1} (+ 8)
        uint32_t qb[KD][2];
        {
          const int n = lane >> 2;
          const uint16_t* qn = reinterpret_cast<const uint16_t*>(p.q) +
                               (static_cast<int64_t>(r0 + (n < nrows ? n : 0)) * p.q_stride + static_cast<int64_t>(head) * D);
#pragma unroll
          for (int kk = 0; kk < KD; ++kk) {
            const int d = 16 * kk + 2 * (lane & 3);
            qb[kk][0] = n < nrows ? (static_cast<uint32_t>(qn[d]) | (static_cast<uint32_t>(qn[d + 1]) << 16)) : 0u;
            qb[kk][1] = n < nrows ? (static_cast<uint32_t>(qn[d + 8]) | (static_cast<uint32_t>(qn[d + 9]) << 16)) : 0u;
          }
        }
        float om[KD][4];                     // O^T: rows d = 16 mt + lane/4 (+8), columns seq 2 (lane%4), +1
#pragma unroll
        for (int mt = 0; mt < KD; ++mt) { om[mt][0] = om[mt][1] = om[mt][2] = om[mt][3] = 0.f; }
        float gm0 = -CUDART_INF_F, gm1 = -CUDART_INF_F, gl0 = 0.f, gl1 = 0.f;      // per column; gl: this lane's share
        {
          const int mi = lane >> 3, rr = lane & 7;                                 // ldmatrix: matrix, row
          int stage2 = 0;
          uint32_t parity = 0;
          for (int n = 0; n < n1; ++n) {
            const int stage = warp * S + stage2;
            mbar_wait(full0 + stage * 8, parity);
            const uint32_t kpage = ring0 + stage * 2 * Gm::kPageBytes;
            const uint32_t vpage = kpage + Gm::kPageBytes;
#pragma unroll
            for (int tt = 0; tt < TT; ++tt) {
              float sc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
              for (int kk = 0; kk < KD; ++kk) {
                uint32_t a[4];   // matrices: (chunk 2kk, tokens 0-7), (2kk, 8-15), (2kk+1, 0-7), (2kk+1, 8-15)
                ldmatrix_x4(a, kpage + (((2 * kk + (mi >> 1)) * B + tt * 16 + (mi & 1) * 8 + rr) * 16));
                mma16816<T>(sc, a, qb[kk][0], qb[kk][1]);
              }
#pragma unroll
              for (int j = 0; j < 4; ++j) sc[j] *= p.scale_log2;
              float mx0 = fmaxf(sc[0], sc[2]), mx1 = fmaxf(sc[1], sc[3]);
#pragma unroll
              for (int o = 4; o < kWarp; o <<= 1) {
                mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, o));
                mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, o));
              }
              const float nm0 = fmaxf(gm0, mx0), nm1 = fmaxf(gm1, mx1);
              if (__any_sync(0xffffffffu, nm0 > gm0 || nm1 > gm1)) {
                const float al0 = fast_exp2(gm0 - nm0), al1 = fast_exp2(gm1 - nm1);
                gl0 *= al0; gl1 *= al1;
#pragma unroll
                for (int mt = 0; mt < KD; ++mt) { om[mt][0] *= al0; om[mt][2] *= al0; om[mt][1] *= al1; om[mt][3] *= al1; }
                gm0 = nm0; gm1 = nm1;
              }
              const float off0 = ProbScale<T>::kLog2 - gm0, off1 = ProbScale<T>::kLog2 - gm1;
              const uint32_t plo = pack2<T>(fast_exp2(sc[0] + off0), fast_exp2(sc[1] + off1));   // token lane/4
              const uint32_t phi = pack2<T>(fast_exp2(sc[2] + off0), fast_exp2(sc[3] + off1));   // token lane/4 + 8
              gl0 += unpack_lo<T>(plo) + unpack_lo<T>(phi);    // sum what the products will actually use
              gl1 += unpack_hi<T>(plo) + unpack_hi<T>(phi);
              const uint32_t b0 = movmatrix_trans(plo), b1 = movmatrix_trans(phi);               // P^T as the B operand
#pragma unroll
              for (int mt = 0; mt < KD; ++mt) {
                uint32_t a[4];   // matrices: (rows 0-7, tokens 0-7), (rows 8-15, 0-7), (rows 0-7, 8-15), (rows 8-15, 8-15)
                ldmatrix_x4(a, vpage + ((16 * mt + (mi & 1) * 8 + rr) * B + tt * 16 + (mi >> 1) * 8) * 2);
                mma16816<T>(om[mt], a, b0, b1);
              }
            }
            __syncwarp();
            if (lane == 0 && n + S < n1) issue_stage(tbl[warp + (n + S) * nwarps], stage);
            if (++stage2 == S) { stage2 = 0; parity ^= 1u; }
          }
        }
#pragma unroll
        for (int o = 4; o < kWarp; o <<= 1) {
          gl0 += __shfl_xor_sync(0xffffffffu, gl0, o);
          gl1 += __shfl_xor_sync(0xffffffffu, gl1, o);
        }
        __syncthreads();   // every warp is done with its ring: it is reused to hand the partials over
        // partials of warp w for sequence n: xfer[(w * kGroup + n) * (D + 2) + d], then m, l
        float* xfer = reinterpret_cast<float*>(smem + map.ring);
        {
          const int n0 = 2 * (lane & 3);
          if (n0 < kGroup) {
            float* x0 = xfer + (warp * kGroup + n0) * (D + 2);
            float* x1 = x0 + (D + 2);
#pragma unroll
            for (int mt = 0; mt < KD; ++mt) {
              const int d = 16 * mt + (lane >> 2);
              x0[d] = om[mt][0]; x1[d] = om[mt][1]; x0[d + 8] = om[mt][2]; x1[d + 8] = om[mt][3];
            }
            if ((lane >> 2) == 0) { x0[D] = gm0; x0[D + 1] = gl0; x1[D] = gm1; x1[D + 1] = gl1; }
          }
        }
        __syncthreads();
        if (active) {
          // warp w gathers sequence w: the log-sum-exp merge of the four warps' shared-page partials
          // becomes its running state (value and sum parked in the lanes the final reduction reads)
          float mm = -CUDART_INF_F;
          for (int w = 0; w < nwarps; ++w) mm = fmaxf(mm, xfer[(w * kGroup + warp) * (D + 2) + D]);
          float ll = 0.f;
          float wgt[8];
#pragma unroll
          for (int w = 0; w < 8; ++w) {
            wgt[w] = 0.f;
            if (w < nwarps) {
              const float mw = xfer[(w * kGroup + warp) * (D + 2) + D];
              wgt[w] = mw > -CUDART_INF_F ? fast_exp2(mw - mm) : 0.f;
              ll += xfer[(w * kGroup + warp) * (D + 2) + D + 1] * wgt[w];
            }
          }
          st.m = mm;
          st.l = lane == 0 ? ll : 0.f;
#pragma unroll
          for (int i = 0; i < Gm::kNvi; ++i) {
            const int d = i * Gm::kRpi + lane / Gm::kVpr;
            float a = 0.f;
            if (lane % Gm::kVpr == 0 && (Gm::kVExact || d < D)) {
#pragma unroll
              for (int w = 0; w < 8; ++w)
                if (w < nwarps) a += xfer[(w * kGroup + warp) * (D + 2) + d] * wgt[w];
            }
            st.acc[i] = a;
          }
          load_q(my_seq);   // not held across phase 1: its registers carry the accumulator tiles there
        }
        // the ring was written (xfer) and read through the generic proxy; the bulk copies of phase 2 write it
        // through the async proxy
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();   // partials consumed: the ring carries pages again
        // Phase 2: the pages that are this sequence's own, through the warp's own S stages
        if (active) {
          const int pages_true_w = my_ctx > 0 ? (my_ctx + B - 1) / B : 0;
          int npages_w = min(pages_true_w, pages_cap) - first_page;
          npages_w = npages_w < 0 ? 0 : (npages_w > p.pages_per_split ? p.pages_per_split : npages_w);
          const int last_local_w = pages_true_w <= pages_cap ? pages_true_w - 1 - first_page : -1;
          const int tail_w = my_ctx - (pages_true_w - 1) * B;
          const int n_priv = npages_w - P;
          uint32_t parbits = 0;                                                    // parity of the warp's stages
          for (int s2 = 0; s2 < S; ++s2) {
            const int used = n1 > s2 ? (n1 - s2 + S - 1) / S : 0;                  // phases completed in phase 1
            parbits |= static_cast<uint32_t>(used & 1) << s2;
          }
          // the sequence's own page numbers come straight from its block-table row, 32 at a time, one per lane
          auto load_ids = [&](int batch) {
            const int i = P + batch * kWarp + lane;
            return i < npages_w ? __ldg(my_row + i) : 0;
          };
          int ids_cur = load_ids(0), ids_next = load_ids(1);
          int batch = 0;
          auto page_id = [&](int k) {            // k in batch or batch + 1; executed by the whole warp
            const int a = __shfl_sync(0xffffffffu, ids_cur, k & 31), b2 = __shfl_sync(0xffffffffu, ids_next, k & 31);
            return (k >> 5) == batch ? a : b2;
          };
          for (int n = 0; n < S && n < n_priv; ++n) {
            const int blk = page_id(n);
            if (lane == 0) issue_stage(blk, warp * S + n);
          }
          for (int n = 0; n < n_priv; ++n) {
            if ((n >> 5) != batch) { batch = n >> 5; ids_cur = ids_next; ids_next = load_ids(batch + 1); }
            const int s2 = n % S, stage = warp * S + s2;
            mbar_wait(full0 + stage * 8, (parbits >> s2) & 1u);
            parbits ^= 1u << s2;
            const int local = P + n;
            const int nvalid = (local == last_local_w) ? tail_w : B;
            const SmemPage kp{ring0 + stage * 2 * Gm::kPageBytes};
            const SmemPage vp{kp.addr + Gm::kPageBytes};
            process_page<T, D, B>(kp, vp, q, st, p.scale_log2, nvalid, pbuf0 + (n & 1) * Gm::kPbufBytes, lane);
            __syncwarp();
            if (n + S < n_priv) {
              const int blk = page_id(n + S);
              if (lane == 0) issue_stage(blk, stage);
            }
          }
          // this warp's result is the whole (sequence, head, split)
#pragma unroll
          for (int i = 0; i < Gm::kNvi; ++i) {
#pragma unroll
            for (int o = 1; o < Gm::kVpr; o <<= 1) st.acc[i] += __shfl_xor_sync(0xffffffffu, st.acc[i], o);
          }
          st.l = warp_sum(st.l);
          const int64_t item = static_cast<int64_t>(my_seq) * H + head;
          const bool writer = lane % Gm::kVpr == 0;
          if (p.num_splits == 1) {
            const float inv = 1.f / (st.l + eps);
#pragma unroll
            for (int i = 0; i < Gm::kNvi; ++i) {
              const int d = i * Gm::kRpi + lane / Gm::kVpr;
              if (writer && (Gm::kVExact || d < D)) {
                const float r = st.acc[i] * inv;
                if constexpr (T::kSize == 2) reinterpret_cast<uint16_t*>(p.out)[item * D + d] = static_cast<uint16_t>(pack2<T>(r, 0.f));
                else reinterpret_cast<float*>(p.out)[item * D + d] = r;
              }
            }
          } else {
            float* part = p.ws_part + (item * p.num_splits + split) * (D + 2);
#pragma unroll
            for (int i = 0; i < Gm::kNvi; ++i) {
              const int d = i * Gm::kRpi + lane / Gm::kVpr;
              if (writer && (Gm::kVExact || d < D)) __stcg(part + d, st.acc[i]);
            }
            if (lane == 0) { __stcg(part + D, st.m); __stcg(part + D + 1, st.l); }
            __threadfence();
            __syncwarp();
            uint32_t prev = 0;
            if (lane == 0) prev = atomicAdd(p.ws_ticket + item, 1u);
            prev = __shfl_sync(0xffffffffu, prev, 0);
            if (prev == static_cast<uint32_t>(p.num_splits) - 1u) {                // last split of this (sequence, head)
              if (lane == 0) p.ws_ticket[item] = 0u;
              __threadfence();
              const float* parts = p.ws_part + item * p.num_splits * (D + 2);
              float gm = -CUDART_INF_F;
              for (int s2 = 0; s2 < p.num_splits; ++s2) gm = fmaxf(gm, __ldcg(parts + s2 * (D + 2) + D));
              float gl = 0.f;
              for (int s2 = 0; s2 < p.num_splits; ++s2) {
                const float ms = __ldcg(parts + s2 * (D + 2) + D);
                if (ms > -CUDART_INF_F) gl += __ldcg(parts + s2 * (D + 2) + D + 1) * fast_exp2(ms - gm);
              }
              for (int d = lane; d < D; d += kWarp) {
                float a = 0.f;
                for (int s2 = 0; s2 < p.num_splits; ++s2) {
                  const float ms = __ldcg(parts + s2 * (D + 2) + D);
                  if (ms > -CUDART_INF_F) a += __ldcg(parts + s2 * (D + 2) + d) * fast_exp2(ms - gm);
                }
                const float r = a / (gl + eps);
                if constexpr (T::kSize == 2) reinterpret_cast<uint16_t*>(p.out)[item * D + d] = static_cast<uint16_t>(pack2<T>(r, 0.f));
                else reinterpret_cast<float*>(p.out)[item * D + d] = r;
              }
            }
          }
        }
        return;
      }
    }
  }

  // pages of this warp: local index warp + n * nwarps, n < n_my
  const int n_my = npages > warp ? (npages - warp + nwarps - 1) / nwarps : 0;
  // local index of the context's last page and the number of tokens in it
  const int last_local = pages_true <= pages_cap ? pages_true - 1 - first_page : -1;
  const int tail = ctx - (pages_true - 1) * B;

  if constexpr (kStaged) {
    const int S = p.num_stages;
    auto issue = [&](int n, int stage) {
      const int64_t blk = tbl[warp + n * nwarps];
      const uint32_t bar = bars + stage * 8;
      const uint32_t dst = ring + stage * 2 * Gm::kPageBytes;
      mbar_expect_tx(bar, 2 * Gm::kPageBytes);
      bulk_g2s_hint(dst, kbase + blk * block_bytes, Gm::kPageBytes, bar, once);
      bulk_g2s_hint(dst + Gm::kPageBytes, vbase + blk * block_bytes, Gm::kPageBytes, bar, once);
    };
    if (lane == 0) {
      const int pre = n_my < S ? n_my : S;
      for (int n = 0; n < pre; ++n) issue(n, n);
    }
    int stage = 0;
    uint32_t parity = 0;
    for (int n = 0; n < n_my; ++n) {
      mbar_wait(bars + stage * 8, parity);
      const int local = warp + n * nwarps;
      const int nvalid = (local == last_local) ? tail : B;
      const SmemPage kp{ring + stage * 2 * Gm::kPageBytes};
      const SmemPage vp{kp.addr + Gm::kPageBytes};
      process_page<T, D, B>(kp, vp, q, st, p.scale_log2, nvalid, pbuf0 + (n & 1) * Gm::kPbufBytes, lane);
      __syncwarp();  // every lane is done reading this stage
      if (lane == 0 && n + S < n_my) issue(n + S, stage);
      if (++stage == S) { stage = 0; parity ^= 1u; }
    }
  } else {
    for (int n = 0; n < n_my; ++n) {
      const int local = warp + n * nwarps;
      const int64_t blk = tbl[local];
      const int nvalid = (local == last_local) ? tail : B;
      const GmemPage kp{reinterpret_cast<const uint4*>(kbase + blk * block_bytes)};
      const GmemPage vp{reinterpret_cast<const uint4*>(vbase + blk * block_bytes)};
      process_page<T, D, B>(kp, vp, q, st, p.scale_log2, nvalid, pbuf0 + (n & 1) * Gm::kPbufBytes, lane);
    }
  }

  // ---- merge the warps of this CTA ------------------------------------------------
#pragma unroll
  for (int i = 0; i < Gm::kNvi; ++i) {
#pragma unroll
    for (int o = 1; o < Gm::kVpr; o <<= 1) st.acc[i] += __shfl_xor_sync(0xffffffffu, st.acc[i], o);
  }
  st.l = warp_sum(st.l);
  float* my = red + warp * (D + 2);
  if (lane == 0) { my[D] = st.m; my[D + 1] = st.l; }
  if (lane % Gm::kVpr == 0) {
#pragma unroll
    for (int i = 0; i < Gm::kNvi; ++i) {
      const int d = i * Gm::kRpi + lane / Gm::kVpr;
      if (Gm::kVExact || d < D) my[d] = st.acc[i];
    }
  }
  __syncthreads();

  float cta_m = -CUDART_INF_F;
  for (int w = 0; w < nwarps; ++w) cta_m = fmaxf(cta_m, red[w * (D + 2) + D]);
  float cta_l = 0.f;
  for (int w = 0; w < nwarps; ++w) {
    const float mw = red[w * (D + 2) + D];
    if (mw > -CUDART_INF_F) cta_l += red[w * (D + 2) + D + 1] * fast_exp2(mw - cta_m);
  }

  if (p.num_splits == 1) {
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
      float a = 0.f;
      for (int w = 0; w < nwarps; ++w) {
        const float mw = red[w * (D + 2) + D];
        if (mw > -CUDART_INF_F) a += red[w * (D + 2) + d] * fast_exp2(mw - cta_m);
      }
      const float r = a / (cta_l + eps);
      const int64_t o = (static_cast<int64_t>(seq) * H + head) * D + d;
      if constexpr (T::kSize == 2) reinterpret_cast<uint16_t*>(p.out)[o] = static_cast<uint16_t>(pack2<T>(r, 0.f));
      else reinterpret_cast<float*>(p.out)[o] = r;
    }
    return;
  }

  // ---- split context: publish the partial, the last CTA to arrive combines -----
  const int64_t item = static_cast<int64_t>(seq) * H + head;
  float* part = p.ws_part + (item * p.num_splits + split) * (D + 2);
  for (int d = threadIdx.x; d < D; d += blockDim.x) {
    float a = 0.f;
    for (int w = 0; w < nwarps; ++w) {
      const float mw = red[w * (D + 2) + D];
      if (mw > -CUDART_INF_F) a += red[w * (D + 2) + d] * fast_exp2(mw - cta_m);
    }
    __stcg(part + d, a);
  }
  if (threadIdx.x == 0) { __stcg(part + D, cta_m); __stcg(part + D + 1, cta_l); }
  __threadfence();
  __syncthreads();
  __shared__ uint32_t s_last;
  if (threadIdx.x == 0) {
    const uint32_t prev = atomicAdd(p.ws_ticket + item, 1u);
    s_last = (prev == static_cast<uint32_t>(p.num_splits) - 1u) ? 1u : 0u;
    if (s_last) p.ws_ticket[item] = 0u;  // re-arm for the next call on this stream
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const float* parts = p.ws_part + item * p.num_splits * (D + 2);
  float gm = -CUDART_INF_F;
  for (int s = 0; s < p.num_splits; ++s) gm = fmaxf(gm, __ldcg(parts + s * (D + 2) + D));
  float gl = 0.f;
  for (int s = 0; s < p.num_splits; ++s) {
    const float ms = __ldcg(parts + s * (D + 2) + D);
    if (ms > -CUDART_INF_F) gl += __ldcg(parts + s * (D + 2) + D + 1) * fast_exp2(ms - gm);
  }
  for (int d = threadIdx.x; d < D; d += blockDim.x) {
    float a = 0.f;
    for (int s = 0; s < p.num_splits; ++s) {
      const float ms = __ldcg(parts + s * (D + 2) + D);
      if (ms > -CUDART_INF_F) a += __ldcg(parts + s * (D + 2) + d) * fast_exp2(ms - gm);
    }
    const float r = a / (gl + eps);  // all splits empty (context_len == 0): 0 / eps == 0
    const int64_t o = item * D + d;
    if constexpr (T::kSize == 2) reinterpret_cast<uint16_t*>(p.out)[o] = static_cast<uint16_t>(pack2<T>(r, 0.f));
    else reinterpret_cast<float*>(p.out)[o] = r;
  }
}

// ---------------------------------------------------------------------------
// host-side launch of one dtype (defined per dtype in decode_*.cu)
// ---------------------------------------------------------------------------
struct DecodeLaunch {
  DecodeParams params;
  int32_t num_seqs;
  int32_t head_size;
  int32_t block_size;
  int32_t num_warps;
  bool staged;
  bool pdl;                   // launch with programmatic stream serialization (PA_DECODE_PDL=0 switches it off)
  cudaStream_t stream;
};

// bytes of dynamic shared memory a launch needs (same formula as the kernel)
int decode_smem_bytes(int dtype, int head_size, int block_size, int pages_per_split,
                      int warps, int stages, bool staged);

int launch_decode_f16(const DecodeLaunch& l);
int launch_decode_bf16(const DecodeLaunch& l);
int launch_decode_f32(const DecodeLaunch& l);

#ifdef PA_DECODE_IMPL
template <typename T, int D, int B, bool kStaged, bool kGroups = false>
static int launch_one(const DecodeLaunch& l) {
  if constexpr (kStaged && !kGroups && T::kSize == 2 && B % 16 == 0) {
    if (l.params.group_rows > 1) return launch_one<T, D, B, true, true>(l);
  }
  const SmemMap<T, D, B> map(l.params.pages_per_split, l.num_warps, l.params.num_stages, kStaged);
  auto kern = paged_decode_kernel<T, D, B, kStaged, kGroups>;
  if (map.total > 48 * 1024) {
    PA_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, map.total));
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(l.params.num_heads, l.num_seqs, l.params.num_splits);
  cfg.blockDim = dim3(l.num_warps * kWarp);
  cfg.dynamicSmemBytes = map.total;
  cfg.stream = l.stream;
  // programmatic dependent launch: the CTAs may be placed (and set their barriers up) while the preceding
  // kernel on the stream drains; the kernel reads no memory before griddepcontrol.wait
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = l.pdl ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, l.params);
  count_launch();
  return check_cuda(e, "paged_decode_kernel launch");
}

template <typename T, int D>
static int launch_block(const DecodeLaunch& l) {
  switch (l.block_size) {
    case 8:  return l.staged ? launch_one<T, D, 8, true>(l) : launch_one<T, D, 8, false>(l);
    case 16: return l.staged ? launch_one<T, D, 16, true>(l) : launch_one<T, D, 16, false>(l);
    case 32: return l.staged ? launch_one<T, D, 32, true>(l) : launch_one<T, D, 32, false>(l);
    default: return set_error(PA_ERR_UNSUPPORTED_BLOCK_SIZE, "unsupported block size: %d", l.block_size);
  }
}

template <typename T>
static int launch_dtype(const DecodeLaunch& l) {
  switch (l.head_size) {
    case 64:  return launch_block<T, 64>(l);
    case 80:  return launch_block<T, 80>(l);
    case 96:  return launch_block<T, 96>(l);
    case 128: return launch_block<T, 128>(l);
    default:  return set_error(PA_ERR_UNSUPPORTED_HEAD_SIZE, "unsupported head size: %d", l.head_size);
  }
}
#endif  // PA_DECODE_IMPL

}  // namespace pa
