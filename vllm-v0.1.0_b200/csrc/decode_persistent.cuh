// Persistent work-queue form of the paged decode kernel (sm_100a).
//
// The grid kernel in decode_attention.cuh gives every (head, sequence, split) its own CTA.  When the
// per-GPU problem is a few waves of equally long CTAs (head-sharded shapes: 5 heads per GPU at TP = 8; small
// batches) every CTA pays its prologue -- block-table row, q row, barrier init, first page fetches, three
// dependent memory latencies -- with an empty ring, all CTAs of a wave pay it at the same time, and the last
// partial wave leaves SMs idle.  Here the unit of scheduling is a WARP and the unit of work a chunk of pages:
//
//   * one CTA per SM, W independent warps, each with its own ring of S (K page + V page) stages fed by
//     cp.async.bulk -- the same page pipeline as the grid kernel, but the ring never drains: when the pages
//     of the unit a warp is consuming are all issued, the warp goes on issuing the first pages (and the q
//     row) of its NEXT unit, so a unit boundary costs no memory bubble;
//   * work units = (sequence, head, chunk of <= C pages), enumerated exactly from the context lengths: every
//     CTA builds the prefix sum of chunks per sequence in shared memory (one pass over context_lens) and a
//     unit index is turned into (sequence, chunk, head) by a binary search there;
//   * units are handed out by an atomic counter in the workspace (the first two per warp are static), fetched
//     one unit ahead so the counter's latency is never waited for; the tail of the launch is at most one
//     chunk per warp;
//   * a (sequence, head) that spans several chunks is combined through the workspace exactly like the grid
//     kernel's context splits (partials + ticket, the last arriver merges); a single-chunk item is written
//     directly;
//   * warps never synchronise with each other after the prologue.
//
// Arithmetic (process_page) is shared with the grid kernel, so results are bit-identical to it whenever the
// chunking matches its splits.
#pragma once

#include "decode_attention.cuh"

namespace pa {

constexpr int kPersistMaxSeqs = 4096;   // context lengths + prefix sums live in shared memory

// dynamic shared memory map (bytes)
template <typename T, int D, int B>
struct PersistMap {
  using Gm = Geom<T, D, B>;
  static constexpr int kQBytes = D * Gm::kS;                                   // one q row
  static constexpr int kPbufStride = ((2 * Gm::kPbufBytes + 15) / 16) * 16;   // probabilities, double-buffered
  static constexpr int kWarpMisc = 2 * kQBytes + kPbufStride;                  // q rows (double-buffered) + pbuf
  int ctx, prefix, misc, bars, ring, total;
  __host__ __device__ PersistMap(int num_seqs, int warps, int stages) {
    int o = 0;
    ctx = o;    o += ((num_seqs * 4 + 15) / 16) * 16;
    prefix = o; o += (((num_seqs + 1) * 4 + 15) / 16) * 16;
    misc = o;   o += warps * kWarpMisc;
    bars = o;   o += warps * stages * 8;
    o = ((o + 127) / 128) * 128;
    ring = o;   o += warps * stages * 2 * Gm::kPageBytes;
    total = o;
  }
};

struct Unit {
  int idx;     // linear unit index, -1: no more work
  int seq, head;
  int first;   // first page (logical block) of the chunk
  int n;       // pages in the chunk (0 only for an empty context)
  int nparts;  // chunks of this (sequence, head)
  int tail;    // valid tokens in the chunk's last page
};

template <typename T, int D, int B>
__global__ void __launch_bounds__(512, 1)
paged_decode_persistent_kernel(const DecodeParams p) {
  using Gm = Geom<T, D, B>;
  using Map = PersistMap<T, D, B>;
  extern __shared__ __align__(128) unsigned char smem[];

  const int warp = threadIdx.x / kWarp;
  const int lane = threadIdx.x % kWarp;
  const int W = blockDim.x / kWarp;
  const int S = p.num_stages;
  const int H = p.num_heads;
  const int C = p.pages_per_split;     // pages per chunk
  const int NS = p.num_seqs;
  const Map map(NS, W, S);
  int* s_ctx = reinterpret_cast<int*>(smem + map.ctx);
  int* s_pre = reinterpret_cast<int*>(smem + map.prefix);

  const uint32_t bars = smem_addr(smem + map.bars) + warp * S * 8;
  const uint32_t ring = smem_addr(smem + map.ring) + warp * S * 2 * Gm::kPageBytes;
  const uint32_t qbuf = smem_addr(smem + map.misc) + warp * Map::kWarpMisc;
  const uint32_t pbuf0 = qbuf + 2 * Map::kQBytes;
  if (lane == 0) {
    for (int s = 0; s < S; ++s) mbar_init(bars + s * 8, 1);
    mbar_init_fence();
  }

  // ---- plan: pages and chunks per sequence, exclusive prefix sum of chunks ---------------------------
  // never walk past the block-table row or the host's partial-slot budget, whatever context_lens says
  const int pages_cap = min(p.max_blocks_per_seq, C * p.num_splits);
  auto pages_of = [&](int ctx) {
    const int pg = ctx > 0 ? (ctx + B - 1) / B : 0;
    return pg < pages_cap ? pg : pages_cap;
  };
  for (int s = threadIdx.x; s < NS; s += blockDim.x) s_ctx[s] = __ldg(p.context_lens + s);
  __syncthreads();
  if (warp == 0) {
    int carry = 0;
    for (int base = 0; base < NS; base += kWarp) {
      const int s = base + lane;
      int v = 0;
      if (s < NS) {
        const int pg = pages_of(s_ctx[s]);
        v = pg > 0 ? (pg + C - 1) / C : 1;      // an empty context still owns one unit: it writes the zero row
      }
      int incl = v;
#pragma unroll
      for (int o = 1; o < kWarp; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
      }
      if (s < NS) s_pre[s] = carry + incl - v;
      carry += __shfl_sync(0xffffffffu, incl, kWarp - 1);
    }
    if (lane == 0) s_pre[NS] = carry;
  }
  __syncthreads();

  const int total_warps = gridDim.x * W;
  const int gw = blockIdx.x * W + warp;
  const int total_units = s_pre[NS] * H;
  uint32_t* ctrl = p.ws_ctrl;            // [0] next dynamic unit, [1] warps that have left

  auto resolve = [&](int idx) {
    Unit u;
    u.idx = -1; u.seq = 0; u.head = 0; u.first = 0; u.n = 0; u.nparts = 1; u.tail = B;
    if (idx < 0 || idx >= total_units) return u;
    const int rc = idx / H;
    int lo = 0, hi = NS;                 // s_pre[lo] <= rc < s_pre[hi]
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (s_pre[mid] <= rc) lo = mid; else hi = mid;
    }
    const int ctx = s_ctx[lo];
    const int pg = pages_of(ctx);
    const int chunk = rc - s_pre[lo];
    u.idx = idx;
    u.seq = lo;
    u.head = idx - rc * H;
    u.first = chunk * C;
    u.n = min(C, pg - u.first);
    u.nparts = s_pre[lo + 1] - s_pre[lo];
    const int pages_true = ctx > 0 ? (ctx + B - 1) / B : 0;
    u.tail = (u.first + u.n == pages_true) ? ctx - (pages_true - 1) * B : B;
    return u;
  };
  auto row_of = [&](const Unit& u) {
    return p.block_tables + static_cast<int64_t>(u.seq) * p.max_blocks_per_seq + u.first;
  };
  auto load_ids = [&](const Unit& u, int batch) {
    const int i = batch * kWarp + lane;
    return (u.idx >= 0 && i < u.n) ? __ldg(row_of(u) + i) : 0;
  };
  // lane 0's return value is the index; the others hold garbage until broadcast
  auto fetch_async = [&]() -> int {
    int v = 0;
    if (lane == 0) v = static_cast<int>(atomicAdd(ctrl, 1u)) + 2 * total_warps;
    return v;
  };

  const char* kcache = reinterpret_cast<const char*>(p.k_cache);
  const char* vcache = reinterpret_cast<const char*>(p.v_cache);
  const int64_t block_bytes = static_cast<int64_t>(H) * Gm::kPageBytes;
  const float eps = 1e-6f * exp2f(ProbScale<T>::kLog2);   // reference: 1/(sum + 1e-6), attention_kernels.cu:213
  const bool q_bulk = p.q_vec_ok != 0;

  int i_stage = 0;
  // one page of unit `u` (block `blk`) into the next ring stage; its first page brings the q row along
  auto issue_page = [&](const Unit& u, int64_t blk, bool first, int qpar) {
    if (lane == 0) {
      const uint32_t bar = bars + i_stage * 8;
      const uint32_t dst = ring + i_stage * 2 * Gm::kPageBytes;
      const int64_t off = blk * block_bytes + static_cast<int64_t>(u.head) * Gm::kPageBytes;
      const bool with_q = first && q_bulk;
      mbar_expect_tx(bar, 2 * Gm::kPageBytes + (with_q ? Map::kQBytes : 0));
      bulk_g2s(dst, kcache + off, Gm::kPageBytes, bar);
      bulk_g2s(dst + Gm::kPageBytes, vcache + off, Gm::kPageBytes, bar);
      if (with_q) {
        const char* qrow = reinterpret_cast<const char*>(p.q) +
                           (static_cast<int64_t>(u.seq) * p.q_stride + static_cast<int64_t>(u.head) * D) * T::kSize;
        bulk_g2s(qbuf + qpar * Map::kQBytes, qrow, Map::kQBytes, bar);
      }
    }
    i_stage = (i_stage + 1 == S) ? 0 : i_stage + 1;
  };

  Unit cur = resolve(gw);
  Unit nxt = resolve(gw + total_warps);
  int pend = nxt.idx >= 0 ? fetch_async() : -1;
  int ids_a = load_ids(cur, 0), ids_b = load_ids(cur, 1), ids_n = load_ids(nxt, 0);
  int ia = 0;                 // batch of `cur` held in ids_a (ids_b: the next one)
  int cur_issued = 0, nxt_issued = 0;
  int c_stage = 0, upar = 0;
  uint32_t parbits = 0, pcount = 0;

  while (cur.idx >= 0) {
    // keep S pages in flight: the rest of this unit first, then the head of the next one
    auto pump = [&](int consumed) {
      while ((cur_issued - consumed) + nxt_issued < S) {
        if (cur_issued < cur.n) {
          if ((cur_issued >> 5) != ia) { ids_a = ids_b; ++ia; ids_b = load_ids(cur, ia + 1); }
          const int blk = __shfl_sync(0xffffffffu, ids_a, cur_issued & 31);
          issue_page(cur, blk, cur_issued == 0, upar);
          ++cur_issued;
        } else if (nxt.idx >= 0 && nxt_issued < nxt.n && nxt_issued < kWarp) {
          const int blk = __shfl_sync(0xffffffffu, ids_n, nxt_issued);
          issue_page(nxt, blk, nxt_issued == 0, upar ^ 1);
          ++nxt_issued;
        } else {
          break;
        }
      }
    };
    pump(0);

    WarpState<T, D, B> st;
    st.m = -CUDART_INF_F;
    st.l = 0.f;
#pragma unroll
    for (int i = 0; i < Gm::kNvi; ++i) st.acc[i] = 0.f;
    uint4 q[Gm::kCpl];

    for (int n = 0; n < cur.n; ++n) {
      mbar_wait(bars + c_stage * 8, (parbits >> c_stage) & 1u);
      parbits ^= 1u << c_stage;
      if (n == 0) {
        const int g = lane / Gm::kTl;
        if (q_bulk) {
#pragma unroll
          for (int cc = 0; cc < Gm::kCpl; ++cc) {
            const int c = cc * Gm::kG + g;
            q[cc] = make_uint4(0u, 0u, 0u, 0u);
            if (Gm::kKExact || c < Gm::kNch) q[cc] = SmemPage{qbuf + upar * Map::kQBytes}.ld(c);
          }
        } else {
          const char* qrow = reinterpret_cast<const char*>(p.q) +
                             (static_cast<int64_t>(cur.seq) * p.q_stride + static_cast<int64_t>(cur.head) * D) * T::kSize;
#pragma unroll
          for (int cc = 0; cc < Gm::kCpl; ++cc) {
            const int c = cc * Gm::kG + g;
            q[cc] = make_uint4(0u, 0u, 0u, 0u);
            if (Gm::kKExact || c < Gm::kNch) {
              if constexpr (T::kSize == 2) {
                const uint16_t* h16 = reinterpret_cast<const uint16_t*>(qrow + c * 16);
                uint32_t w[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) w[j] = static_cast<uint32_t>(h16[2 * j]) | (static_cast<uint32_t>(h16[2 * j + 1]) << 16);
                q[cc] = make_uint4(w[0], w[1], w[2], w[3]);
              } else {
                const uint32_t* w32 = reinterpret_cast<const uint32_t*>(qrow + c * 16);
                q[cc] = make_uint4(w32[0], w32[1], w32[2], w32[3]);
              }
            }
          }
        }
      }
      const int nvalid = (n == cur.n - 1) ? cur.tail : B;
      const SmemPage kp{ring + c_stage * 2 * Gm::kPageBytes};
      const SmemPage vp{kp.addr + Gm::kPageBytes};
      process_page<T, D, B>(kp, vp, q, st, p.scale_log2, nvalid, pbuf0 + (pcount & 1u) * Gm::kPbufBytes, lane);
      ++pcount;
      __syncwarp();   // every lane is done reading this stage
      c_stage = (c_stage + 1 == S) ? 0 : c_stage + 1;
      pump(n + 1);
    }

    // ---- this warp's result is the whole chunk ------------------------------------------------------
#pragma unroll
    for (int i = 0; i < Gm::kNvi; ++i) {
#pragma unroll
      for (int o = 1; o < Gm::kVpr; o <<= 1) st.acc[i] += __shfl_xor_sync(0xffffffffu, st.acc[i], o);
    }
    st.l = warp_sum(st.l);
    const int64_t item = static_cast<int64_t>(cur.seq) * H + cur.head;
    const bool writer = lane % Gm::kVpr == 0;
    if (cur.nparts == 1) {
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
      float* part = p.ws_part + static_cast<int64_t>(cur.idx) * (D + 2);
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
      if (prev == static_cast<uint32_t>(cur.nparts) - 1u) {          // last chunk of this (sequence, head) to finish
        if (lane == 0) p.ws_ticket[item] = 0u;                        // re-arm for the next call on this stream
        __threadfence();
        const int64_t u0 = static_cast<int64_t>(s_pre[cur.seq]) * H + cur.head;   // chunk c lives at unit u0 + c H
        const int64_t ustride = static_cast<int64_t>(H) * (D + 2);
        const float* parts = p.ws_part + u0 * (D + 2);
        float gm = -CUDART_INF_F;
        for (int c = 0; c < cur.nparts; ++c) gm = fmaxf(gm, __ldcg(parts + c * ustride + D));
        float gl = 0.f;
        for (int c = 0; c < cur.nparts; ++c) {
          const float ms = __ldcg(parts + c * ustride + D);
          if (ms > -CUDART_INF_F) gl += __ldcg(parts + c * ustride + D + 1) * fast_exp2(ms - gm);
        }
        for (int d = lane; d < D; d += kWarp) {
          float a = 0.f;
          for (int c = 0; c < cur.nparts; ++c) {
            const float ms = __ldcg(parts + c * ustride + D);
            if (ms > -CUDART_INF_F) a += __ldcg(parts + c * ustride + d) * fast_exp2(ms - gm);
          }
          const float r = a / (gl + eps);
          if constexpr (T::kSize == 2) reinterpret_cast<uint16_t*>(p.out)[item * D + d] = static_cast<uint16_t>(pack2<T>(r, 0.f));
          else reinterpret_cast<float*>(p.out)[item * D + d] = r;
        }
      }
    }

    // ---- next unit becomes current; resolve the one after it (its index was asked for a unit ago) ----
    cur = nxt;
    cur_issued = nxt_issued;
    nxt_issued = 0;
    upar ^= 1;
    ids_a = ids_n;
    ia = 0;
    ids_b = load_ids(cur, 1);
    if (cur.idx >= 0) {
      nxt = resolve(__shfl_sync(0xffffffffu, pend, 0));
      pend = nxt.idx >= 0 ? fetch_async() : -1;
    } else {
      nxt.idx = -1;
    }
    ids_n = load_ids(nxt, 0);
  }

  // ---- leave: the last warp out re-arms the counters for the next launch on this stream -----------------
  __syncwarp();
  if (lane == 0) {
    __threadfence();
    const uint32_t prev = atomicAdd(ctrl + 1, 1u);
    if (prev == static_cast<uint32_t>(total_warps) - 1u) {
      ctrl[0] = 0u;
      ctrl[1] = 0u;
      __threadfence();
    }
  }
}

#ifdef PA_DECODE_IMPL
template <typename T, int D, int B>
static int launch_persistent_one(const DecodeLaunch& l) {
  const PersistMap<T, D, B> map(l.num_seqs, l.num_warps, l.params.num_stages);
  auto kern = paged_decode_persistent_kernel<T, D, B>;
  if (map.total > 48 * 1024) {
    PA_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, map.total));
  }
  kern<<<l.grid_ctas, l.num_warps * kWarp, map.total, l.stream>>>(l.params);
  count_launch();
  return check_cuda(cudaGetLastError(), "paged_decode_persistent_kernel launch");
}

template <typename T, int D>
static int launch_persistent_block(const DecodeLaunch& l) {
  switch (l.block_size) {
    case 8:  return launch_persistent_one<T, D, 8>(l);
    case 16: return launch_persistent_one<T, D, 16>(l);
    case 32: return launch_persistent_one<T, D, 32>(l);
    default: return set_error(PA_ERR_UNSUPPORTED_BLOCK_SIZE, "unsupported block size: %d", l.block_size);
  }
}

template <typename T>
static int launch_persistent_dtype(const DecodeLaunch& l) {
  switch (l.head_size) {
    case 64:  return launch_persistent_block<T, 64>(l);
    case 80:  return launch_persistent_block<T, 80>(l);
    case 96:  return launch_persistent_block<T, 96>(l);
    case 128: return launch_persistent_block<T, 128>(l);
    default:  return set_error(PA_ERR_UNSUPPORTED_HEAD_SIZE, "unsupported head size: %d", l.head_size);
  }
}
#endif  // PA_DECODE_IMPL

}  // namespace pa
