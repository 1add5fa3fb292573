// KV-cache movers for sm_100a: reshape_and_cache, gather_cached_kv, copy_blocks,
// swap_blocks.  Pure data movement, bit-exact.  Reference behaviour:
// csrc/cache_kernels.cu:9-47 (swap), :52-138 (copy), :142-224 (reshape_and_cache),
// :229-386 (gather) of the reference tree.  All offsets are 64-bit.
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <type_traits>

#include "pa_common.cuh"

namespace pa {

// ---------------------------------------------------------------------------
// reshape_and_cache / gather_cached_kv
//
// One CTA moves a tile of kTile consecutive tokens x one head through shared
// memory.  Rows (key[t][h][:], value[t][h][:]) are contiguous D*sizeof(E) bytes
// and move as 16-byte vectors.  The cache side is where the layout bites:
//   key   [blk][h][d/x][off][x]   one 16-byte chunk per (token, d/x); chunks of
//                                 consecutive tokens of a block are adjacent
//   value [blk][h][d][off]        one element per (token, d); the B tokens of a
//                                 block share a 2*B-byte row
// so the lane order of the cache accesses is chosen per tile: when the tile's
// slots run consecutively (prefill: a prompt fills its blocks in order) lanes
// walk tokens first and a half-warp touches one whole 32-byte sector / 256-byte
// key row; when they are scattered (decode: every token in its own block) lanes
// walk the head dimension first, which touches 4x fewer 128-byte lines per
// instruction than any other order.
// ---------------------------------------------------------------------------
struct ScatterArgs {
  const void* key;      // gather: destination
  const void* value;
  void* key_cache;      // gather: source
  void* value_cache;
  const int32_t* slot_mapping;
  int64_t key_stride;   // elements
  int64_t value_stride;
  int32_t num_tokens;
  int32_t num_heads;
  int32_t head_size;
  int32_t block_size;
  int32_t vec_ok;       // key/value rows are 16-byte aligned
  int32_t chunk_shift;  // log2 of the lanes laid along the 16-byte chunks of a row (>= head_size / x)
  // fused rotary embedding (kRope != 0): q and k rows of tokens [0, num_rope_tokens) are rotated in place
  const int64_t* positions;   // [num_rope_tokens]
  void* query;                // [num_rope_tokens][H][D], row stride query_stride
  const void* cos_sin;        // [max_position][rot_dim]: cos | sin halves
  int64_t query_stride;
  int32_t rot_dim;
  int32_t num_rope_tokens;
  int32_t rot_shift;          // log2 of the lanes laid along rot_dim / 2
  int32_t rot_vec_shift;      // >= 0: 16-byte path, log2 of the lanes laid along (rot_dim / 2) / x; -1: element-wise
};

// ---------------------------------------------------------------------------
// GPT-NeoX rotary embedding of one (x, y) pair, rounded exactly as the reference computes it
// (csrc/pos_encoding_kernels.cu:36-45 on c10::Half / c10::BFloat16 / float: every product and
// every sum is formed in fp32 and rounded to T before the next operation, no fused
// multiply-add), so the result is bit-identical to the reference's PyTorch implementation
// (tests/kernels/test_pos_encoding.py:12-27, q * cos + rotate_half(q) * sin in dtype T).
// ---------------------------------------------------------------------------
template <typename T> struct RopeMath;
template <> struct RopeMath<F16> {
  using E = uint16_t;
  static __device__ __forceinline__ float up(E v) { return __half2float(__ushort_as_half(v)); }
  static __device__ __forceinline__ E down(float f) { return __half_as_ushort(__float2half_rn(f)); }
};
template <> struct RopeMath<BF16> {
  using E = uint16_t;
  static __device__ __forceinline__ float up(E v) { return __uint_as_float(static_cast<uint32_t>(v) << 16); }
  static __device__ __forceinline__ E down(float f) { return __bfloat16_as_ushort(__float2bfloat16_rn(f)); }
};
template <> struct RopeMath<F32> {
  using E = uint32_t;
  static __device__ __forceinline__ float up(E v) { return __uint_as_float(v); }
  static __device__ __forceinline__ E down(float f) { return __float_as_uint(f); }
};
template <typename T>
__device__ __forceinline__ void rope_pair(typename RopeMath<T>::E& x, typename RopeMath<T>::E& y,
                                          typename RopeMath<T>::E c, typename RopeMath<T>::E s) {
  using M = RopeMath<T>;
  const float fx = M::up(x), fy = M::up(y), fc = M::up(c), fs = M::up(s);
  const float xc = M::up(M::down(__fmul_rn(fx, fc))), ys = M::up(M::down(__fmul_rn(fy, fs)));
  const float yc = M::up(M::down(__fmul_rn(fy, fc))), xs = M::up(M::down(__fmul_rn(fx, fs)));
  x = M::down(__fsub_rn(xc, ys));
  y = M::down(__fadd_rn(yc, xs));
}
// the same on a 16-byte vector of pairs: x and y hold X consecutive elements of the two halves
template <typename T>
__device__ __forceinline__ void rope_vec(uint4& x, uint4& y, const uint4& c, const uint4& s) {
  using E = typename RopeMath<T>::E;
  constexpr int X = 16 / sizeof(E);
  union V { uint4 q; E e[X]; };
  V vx, vy, vc, vs;
  vx.q = x; vy.q = y; vc.q = c; vs.q = s;
#pragma unroll
  for (int i = 0; i < X; ++i) rope_pair<T>(vx.e[i], vy.e[i], vc.e[i], vs.e[i]);
  x = vx.q; y = vy.q;
}
struct NoRope { using E = void; };

constexpr int kTile = 16;          // tokens per CTA
constexpr int kTileThreads = 128;

template <typename E, bool kGather, typename R = NoRope>  // E = uint16_t or uint32_t; R = NoRope or F16/BF16/F32
__global__ void __launch_bounds__(kTileThreads) kv_tile_kernel(const ScatterArgs a) {
  constexpr bool kRope = !std::is_same<R, NoRope>::value;
  constexpr int X = 16 / sizeof(E);
  extern __shared__ __align__(16) unsigned char tile_smem[];
  __shared__ int64_t s_kbase[kTile];   // element offset of (blk, h, chunk 0, off, 0) in the key cache
  __shared__ int64_t s_vbase[kTile];   // element offset of (blk, h, row 0, off) in the value cache
  __shared__ int s_valid[kTile];       // token is written to / read from the caches
  __shared__ int s_tok[kTile];         // token exists for the rotary embedding (kRope)
  __shared__ int64_t s_pos[kTile];
  __shared__ int s_runs;

  const int D = a.head_size, B = a.block_size, H = a.num_heads;
  const int nch = D / X;
  const int h = blockIdx.y;
  const int t0 = blockIdx.x * kTile;
  const int nt = min(kTile, (kRope ? max(a.num_tokens, a.num_rope_tokens) : a.num_tokens) - t0);
  const int row_bytes = D * static_cast<int>(sizeof(E)) + 16;   // +16: spreads the rows over the banks
  unsigned char* ktile = tile_smem;
  unsigned char* vtile = tile_smem + kTile * row_bytes;

  if (threadIdx.x < kTile) {
    const int j = threadIdx.x;
    int slot = -1;
    if (j < nt && t0 + j < a.num_tokens) slot = a.slot_mapping[t0 + j];
    if constexpr (kRope) {
      s_tok[j] = j < nt && t0 + j < a.num_rope_tokens;
      s_pos[j] = s_tok[j] ? a.positions[t0 + j] : 0;
    }
    s_valid[j] = slot >= 0;                      // negative slot: padding token (not produced by the v0.1.0 scheduler)
    const int64_t blk = slot >= 0 ? slot / B : 0;
    const int off = slot >= 0 ? slot % B : 0;
    s_kbase[j] = ((blk * H + h) * nch * B + off) * X;
    s_vbase[j] = (blk * H + h) * static_cast<int64_t>(D) * B + off;
    // does this token continue the previous token's block?
    const int prev = __shfl_up_sync(0x0000ffffu, slot, 1);
    const bool cont = j > 0 && j < nt && slot >= 0 && prev >= 0 && slot == prev + 1 && (slot % B) != 0;
    const unsigned m = __ballot_sync(0x0000ffffu, cont);
    if (j == 0) s_runs = __popc(m);
  }
  __syncthreads();
  const bool token_major = s_runs * 2 >= nt;     // at least half of the tile continues a run

  E* kc = reinterpret_cast<E*>(a.key_cache);
  E* vc = reinterpret_cast<E*>(a.value_cache);
  E* krows = reinterpret_cast<E*>(const_cast<void*>(a.key)) + static_cast<int64_t>(h) * D;
  E* vrows = reinterpret_cast<E*>(const_cast<void*>(a.value)) + static_cast<int64_t>(h) * D;

  // Thread -> (token, chunk/row) maps are built from shifts and masks only (no integer division
  // in the loops): cshift = log2(lanes along the 16-byte chunks of a row), dshift likewise for D.
  const int cshift = a.chunk_shift;
  const int clanes = 1 << cshift;

  // rows <-> tile (16-byte vectors; element-wise when the rows are not 16-byte aligned)
  auto rows_pass = [&](bool to_tile) {
    const int c = threadIdx.x & (clanes - 1);
    if (c >= nch) return;
    for (int j = threadIdx.x >> cshift; j < 2 * nt; j += kTileThreads >> cshift) {
      const int which = j >= nt;                           // 0 key, 1 value
      const int jj = which ? j - nt : j;
      if (kRope && !which ? !(s_tok[jj] || s_valid[jj]) : !s_valid[jj]) continue;
      E* row = (which ? vrows + (t0 + jj) * a.value_stride : krows + (t0 + jj) * a.key_stride) + c * X;
      uint4* cell = reinterpret_cast<uint4*>((which ? vtile : ktile) + jj * row_bytes + c * 16);
      if (to_tile) {
        if (a.vec_ok) {
          *cell = *reinterpret_cast<const uint4*>(row);
        } else {
          E* ce = reinterpret_cast<E*>(cell);
#pragma unroll
          for (int e = 0; e < X; ++e) ce[e] = row[e];
        }
      } else {
        if (a.vec_ok) {
          *reinterpret_cast<uint4*>(row) = *cell;
        } else {
          const E* ce = reinterpret_cast<const E*>(cell);
#pragma unroll
          for (int e = 0; e < X; ++e) row[e] = ce[e];
        }
      }
    }
  };

  // tile <-> caches
  auto cache_pass = [&]() {
    if (token_major) {
      // lanes walk the tile's tokens first: 16 lanes cover one 256-byte key row / one value sector
      const int j = threadIdx.x & (kTile - 1);
      if (j < nt && s_valid[j]) {
        const int64_t kb = s_kbase[j], vb = s_vbase[j];
        for (int c = threadIdx.x / kTile; c < nch; c += kTileThreads / kTile) {
          uint4* cell = reinterpret_cast<uint4*>(ktile + j * row_bytes + c * 16);
          uint4* g = reinterpret_cast<uint4*>(kc + kb + static_cast<int64_t>(c) * B * X);
          if (kGather) *cell = *g; else *g = *cell;
        }
        E* vrow = reinterpret_cast<E*>(vtile + j * row_bytes);
#pragma unroll 4
        for (int d = threadIdx.x / kTile; d < D; d += kTileThreads / kTile) {
          E* g = vc + vb + static_cast<int64_t>(d) * B;
          if (kGather) vrow[d] = *g; else *g = vrow[d];
        }
      }
    } else {
      // lanes walk the head dimension first (scattered tokens: fewest 128-byte lines per instruction)
      const int c = threadIdx.x & (clanes - 1);
      if (c < nch) {
        for (int j = threadIdx.x >> cshift; j < nt; j += kTileThreads >> cshift) {
          if (!s_valid[j]) continue;
          uint4* cell = reinterpret_cast<uint4*>(ktile + j * row_bytes + c * 16);
          uint4* g = reinterpret_cast<uint4*>(kc + s_kbase[j] + static_cast<int64_t>(c) * B * X);
          if (kGather) *cell = *g; else *g = *cell;
        }
      }
      for (int j = 0; j < nt; ++j) {
        if (!s_valid[j]) continue;
        E* vrow = reinterpret_cast<E*>(vtile + j * row_bytes);
        const int64_t vb = s_vbase[j];
        for (int d = threadIdx.x; d < D; d += kTileThreads) {
          E* g = vc + vb + static_cast<int64_t>(d) * B;
          if (kGather) vrow[d] = *g; else *g = vrow[d];
        }
      }
    }
  };

  if (!kRope && !token_major && a.vec_ok && H >= nt) {
    // Scattered tokens (a decode step: every token is the newest one of its own sequence).  Nothing to
    // gain from staging, and DRAM locality is per token: one token's K/V destinations are a contiguous
    // [heads][...] region of its block, 2 bytes (V) or 16 bytes (K) written per 32-byte sector, so the
    // sector fills dominate and want to stream through one DRAM page.  The tile's H CTAs therefore
    // regroup: CTA h takes token h % nt and a contiguous share of the heads.
    const int j = h % nt;
    if (!s_valid[j]) return;
    const int m = (H - j + nt - 1) / nt;           // CTAs of this tile that landed on token j
    const int r = h / nt;
    const int lo = r * H / m, hi = (r + 1) * H / m;
    const int64_t hk = static_cast<int64_t>(h) * nch * B * X, hv = static_cast<int64_t>(h) * D * B;
    E* kdst = kc + (s_kbase[j] - hk);              // head 0 of the token's block
    E* vdst = vc + (s_vbase[j] - hv);
    E* krow = krows - static_cast<int64_t>(h) * D + (t0 + j) * a.key_stride;
    E* vrow = vrows - static_cast<int64_t>(h) * D + (t0 + j) * a.value_stride;
#pragma unroll 2
    for (int ci = lo * nch + threadIdx.x; ci < hi * nch; ci += kTileThreads) {
      uint4* rp = reinterpret_cast<uint4*>(krow + static_cast<int64_t>(ci) * X);
      uint4* g = reinterpret_cast<uint4*>(kdst + static_cast<int64_t>(ci) * B * X);
      if (kGather) *rp = *g; else *g = __ldg(rp);
    }
#pragma unroll 8
    for (int i = lo * D + threadIdx.x; i < hi * D; i += kTileThreads) {
      E* g = vdst + static_cast<int64_t>(i) * B;
      if (kGather) vrow[i] = *g; else *g = __ldg(vrow + i);
    }
    return;
  }
  if constexpr (kRope) {
    // rows -> tile, rotate q (in global memory) and k (in the tile, and back to its row), tile -> caches
    using M = RopeMath<R>;
    static_assert(std::is_same<typename M::E, E>::value, "element width");
    rows_pass(true);
    __syncthreads();
    const int embed = a.rot_dim / 2;
    const E* cs = reinterpret_cast<const E*>(a.cos_sin);
    E* qrows = reinterpret_cast<E*>(a.query) + static_cast<int64_t>(h) * D;
    if (a.rot_vec_shift >= 0) {
      // 16 bytes of each half per thread: cos/sin, k (tile) and q (global) all move as vectors
      const int ech = embed / X;
      const int c = threadIdx.x & ((1 << a.rot_vec_shift) - 1);
      if (c < ech) {
        for (int j = threadIdx.x >> a.rot_vec_shift; j < nt; j += kTileThreads >> a.rot_vec_shift) {
          if (!s_tok[j]) continue;
          const E* csr = cs + s_pos[j] * a.rot_dim;
          const uint4 cv = __ldg(reinterpret_cast<const uint4*>(csr) + c);
          const uint4 sv = __ldg(reinterpret_cast<const uint4*>(csr + embed) + c);
          uint4* qx = reinterpret_cast<uint4*>(qrows + (t0 + j) * a.query_stride) + c;
          uint4* qy = qx + ech;
          uint4 x = *qx, y = *qy;
          uint4* kx = reinterpret_cast<uint4*>(ktile + j * row_bytes) + c;
          uint4* ky = kx + ech;
          uint4 u = *kx, v = *ky;
          rope_vec<R>(x, y, cv, sv);
          rope_vec<R>(u, v, cv, sv);
          *qx = x; *qy = y;
          *kx = u; *ky = v;
          uint4* gx = reinterpret_cast<uint4*>(krows + (t0 + j) * a.key_stride) + c;
          gx[0] = u; gx[ech] = v;
        }
      }
    } else {
      const int i = threadIdx.x & ((1 << a.rot_shift) - 1);
      if (i < embed) {
        for (int j = threadIdx.x >> a.rot_shift; j < nt; j += kTileThreads >> a.rot_shift) {
          if (!s_tok[j]) continue;
          const E c = __ldg(cs + s_pos[j] * a.rot_dim + i), sn = __ldg(cs + s_pos[j] * a.rot_dim + embed + i);
          E* kt = reinterpret_cast<E*>(ktile + j * row_bytes);
          E kx = kt[i], ky = kt[embed + i];
          rope_pair<R>(kx, ky, c, sn);
          kt[i] = kx; kt[embed + i] = ky;
          E* krow = krows + (t0 + j) * a.key_stride;
          krow[i] = kx; krow[embed + i] = ky;
          E* qrow = qrows + (t0 + j) * a.query_stride;
          E qx = qrow[i], qy = qrow[embed + i];
          rope_pair<R>(qx, qy, c, sn);
          qrow[i] = qx; qrow[embed + i] = qy;
        }
      }
    }
    __syncthreads();
    cache_pass();
  } else if (!kGather) {
    rows_pass(true);
    __syncthreads();
    cache_pass();
  } else {
    cache_pass();
    __syncthreads();
    rows_pass(false);
  }
}

// reshape_and_cache for a handful of tokens (a decode step: one new token per sequence, every token in its own
// block).  Nothing can be staged or merged there -- each 2-byte V element and each 16-byte K chunk lands in a
// sector of its own -- so the launch is bound by the requests the L2 sees: one CTA per token sweeps the token's
// [heads][head_size] row; lanes walk consecutive elements of V, so one warp store covers whole 128-byte lines
// of the cache (4 sectors each: 8 requests per warp store; a thread per 16-byte chunk would make 32), and the
// K chunk of 8 elements goes out as one 16-byte store.  All loads of a thread are issued before its stores.
template <typename E, int kThreads>
__global__ void __launch_bounds__(kThreads) kv_scatter_direct_kernel(const ScatterArgs a) {
  constexpr int X = 16 / sizeof(E);
  constexpr int kU = 4;
  const int t = blockIdx.x;
  const int slot = __ldg(a.slot_mapping + t);
  if (slot < 0) return;                                      // padding token
  const int B = a.block_size, H = a.num_heads, D = a.head_size;
  const int nch = D / X;
  const int64_t blk = slot / B;
  const int off = slot % B;
  const E* krow = reinterpret_cast<const E*>(a.key) + t * a.key_stride;
  const E* vrow = reinterpret_cast<const E*>(a.value) + t * a.value_stride;
  // value: ((blk H + h) D + d) B + off with i = h D + d
  E* vc = reinterpret_cast<E*>(a.value_cache) + blk * H * D * B + off;
  const int n = H * D;
  for (int i0 = threadIdx.x; i0 < n; i0 += kThreads * kU) {
    E v[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int i = i0 + u * kThreads;
      if (i < n) v[u] = __ldg(vrow + i);
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int i = i0 + u * kThreads;
      if (i < n) vc[static_cast<int64_t>(i) * B] = v[u];
    }
  }
  // key: ((blk H + h) nch + c) B X + off X with ci = h nch + c
  E* kc = reinterpret_cast<E*>(a.key_cache) + (blk * H * nch * B + off) * X;
  for (int ci = threadIdx.x; ci < H * nch; ci += kThreads) {
    const uint4 kv = __ldg(reinterpret_cast<const uint4*>(krow) + ci);
    *reinterpret_cast<uint4*>(kc + static_cast<int64_t>(ci) * B * X) = kv;
  }
}
constexpr int kDirectMinTokens = 96;
constexpr int kDirectMaxTokens = 512;

template <bool kGather>
static int scatter_gather(const void* key, const void* value, void* key_cache, void* value_cache,
                          const int32_t* slot_mapping, int32_t num_tokens, int32_t num_heads,
                          int32_t head_size, int32_t block_size, int64_t key_stride,
                          int64_t value_stride, int32_t dtype, cudaStream_t stream,
                          int32_t slot_pattern = PA_SLOTS_AUTO) {
  const char* op = kGather ? "gather_cached_kv" : "reshape_and_cache";
  if (slot_pattern < PA_SLOTS_AUTO || slot_pattern > PA_SLOTS_RUNS)
    return set_error(PA_ERR_INVALID_ARGUMENT, "%s: invalid slot pattern %d", op, slot_pattern);
  if (dtype != PA_F16 && dtype != PA_BF16 && dtype != PA_F32)
    return set_error(PA_ERR_UNSUPPORTED_DTYPE, "%s: unsupported dtype %d", op, dtype);
  if (num_tokens < 0 || num_heads <= 0 || head_size <= 0 || block_size <= 0)
    return set_error(PA_ERR_INVALID_ARGUMENT, "%s: bad sizes (tokens %d, heads %d, head_size %d, block_size %d)",
                     op, num_tokens, num_heads, head_size, block_size);
  if (num_tokens == 0) return PA_OK;
  if (!key || !value || !key_cache || !value_cache || !slot_mapping)
    return set_error(PA_ERR_INVALID_ARGUMENT, "%s: null pointer", op);
  const int es = dtype_size(dtype);
  const int x = 16 / es;
  if (head_size % x != 0)
    return set_error(PA_ERR_INVALID_ARGUMENT, "%s: head_size %d is not a multiple of x=%d", op, head_size, x);
  if ((reinterpret_cast<uintptr_t>(key_cache) | reinterpret_cast<uintptr_t>(value_cache)) & 15u)
    return set_error(PA_ERR_INVALID_ARGUMENT, "%s: caches must be 16-byte aligned", op);
  if (num_heads > 65535) return set_error(PA_ERR_INVALID_ARGUMENT, "%s: too many heads (%d)", op, num_heads);

  ScatterArgs a;
  a.key = key; a.value = value; a.key_cache = key_cache; a.value_cache = value_cache;
  a.slot_mapping = slot_mapping; a.key_stride = key_stride; a.value_stride = value_stride;
  a.num_tokens = num_tokens; a.num_heads = num_heads; a.head_size = head_size; a.block_size = block_size;
  a.vec_ok = ((reinterpret_cast<uintptr_t>(key) | reinterpret_cast<uintptr_t>(value)) % 16 == 0) &&
             ((key_stride * es) % 16 == 0) && ((value_stride * es) % 16 == 0) && ((head_size * es) % 16 == 0);

  a.chunk_shift = 0;
  while ((1 << a.chunk_shift) < head_size / x) ++a.chunk_shift;
  if ((1 << a.chunk_shift) > kTileThreads) return set_error(PA_ERR_INVALID_ARGUMENT, "%s: head_size %d too large", op, head_size);
  a.positions = nullptr; a.query = nullptr; a.cos_sin = nullptr; a.query_stride = 0;
  a.rot_dim = 0; a.num_rope_tokens = 0; a.rot_shift = 0; a.rot_vec_shift = -1;

  // Measured on B200 against the reference's kernel (same one-CTA-per-token sweep, element stores): 256 tokens
  // 12.31 vs 12.32 us, 512 tokens 34.7 vs 36.7 us -- the time is the L2's: 1.3 M partial-sector stores and the
  // DRAM fills behind them, whatever issues them.  Below ~100 tokens the tile mover's head-parallel grid wins
  // (8 tokens: 4.1 vs 7.2 us, 64 tokens: 7.9 vs 8.2 us): one CTA per token cannot hide its own latency chain.
  // That holds for scattered slots only (a decode step).  The same token counts are also a short prompt, whose slots
  // are consecutive: there the tile mover merges a block's tokens into whole sectors and the element sweep does not.
  // The slot mapping lives on the device, so the host cannot tell the two apart: the caller says which it is
  // (pa_reshape_and_cache_ex; PagedAttention knows from num_prompt_tokens), and without a hint the tile mover, which
  // is never far off, is taken.
  if (!kGather && a.vec_ok && slot_pattern == PA_SLOTS_SCATTERED && num_tokens > kDirectMinTokens &&
      num_tokens <= kDirectMaxTokens) {
    if (es == 2) kv_scatter_direct_kernel<uint16_t, 512><<<num_tokens, 512, 0, stream>>>(a);
    else kv_scatter_direct_kernel<uint32_t, 512><<<num_tokens, 512, 0, stream>>>(a);
    count_launch();
    return check_cuda(cudaGetLastError(), op);
  }
  const int smem = 2 * kTile * (head_size * es + 16);
  if (smem > 200 * 1024) return set_error(PA_ERR_INVALID_ARGUMENT, "%s: head_size %d too large", op, head_size);
  const dim3 grid((num_tokens + kTile - 1) / kTile, num_heads);
  auto launch = [&](auto kern) -> int {
    if (smem > 48 * 1024)
      PA_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    kern<<<grid, kTileThreads, smem, stream>>>(a);
    return PA_OK;
  };
  int rc = es == 2 ? launch(kv_tile_kernel<uint16_t, kGather>) : launch(kv_tile_kernel<uint32_t, kGather>);
  if (rc != PA_OK) return rc;
  count_launch();
  return check_cuda(cudaGetLastError(), op);
}

// ---------------------------------------------------------------------------
// rotary_embedding_neox (stand-alone) and rotary embedding fused with reshape_and_cache
// ---------------------------------------------------------------------------
struct RopeArgs {
  const int64_t* positions;
  void* query;
  void* key;
  const void* cos_sin;
  int64_t query_stride, key_stride;
  int32_t num_heads, head_size, rot_dim, rot_shift, rot_vec_shift;
};

// grid: one CTA per token; lanes along rot_dim / 2, thread rows along heads.
template <typename T>
__global__ void __launch_bounds__(256) rope_kernel(const RopeArgs a) {
  using M = RopeMath<T>;
  using E = typename M::E;
  const int64_t token = blockIdx.x;
  const int64_t pos = a.positions[token];
  const int embed = a.rot_dim / 2;
  if (a.rot_vec_shift >= 0) {
    constexpr int X = 16 / sizeof(E);
    const int ech = embed / X;
    const int c = threadIdx.x & ((1 << a.rot_vec_shift) - 1);
    if (c >= ech) return;
    const E* csr = reinterpret_cast<const E*>(a.cos_sin) + pos * a.rot_dim;
    const uint4 cv = __ldg(reinterpret_cast<const uint4*>(csr) + c);
    const uint4 sv = __ldg(reinterpret_cast<const uint4*>(csr + embed) + c);
    E* q = reinterpret_cast<E*>(a.query) + token * a.query_stride;
    E* k = reinterpret_cast<E*>(a.key) + token * a.key_stride;
    for (int hh = threadIdx.x >> a.rot_vec_shift; hh < a.num_heads; hh += blockDim.x >> a.rot_vec_shift) {
      uint4* qx = reinterpret_cast<uint4*>(q + static_cast<int64_t>(hh) * a.head_size) + c;
      uint4* kx = reinterpret_cast<uint4*>(k + static_cast<int64_t>(hh) * a.head_size) + c;
      uint4 x = qx[0], y = qx[ech], u = kx[0], v = kx[ech];
      rope_vec<T>(x, y, cv, sv);
      rope_vec<T>(u, v, cv, sv);
      qx[0] = x; qx[ech] = y;
      kx[0] = u; kx[ech] = v;
    }
    return;
  }
  const int i = threadIdx.x & ((1 << a.rot_shift) - 1);
  if (i >= embed) return;
  const E* cs = reinterpret_cast<const E*>(a.cos_sin) + pos * a.rot_dim;
  const E c = __ldg(cs + i), sn = __ldg(cs + embed + i);
  E* q = reinterpret_cast<E*>(a.query) + token * a.query_stride;
  E* k = reinterpret_cast<E*>(a.key) + token * a.key_stride;
  for (int hh = threadIdx.x >> a.rot_shift; hh < a.num_heads; hh += blockDim.x >> a.rot_shift) {
    E* qh = q + static_cast<int64_t>(hh) * a.head_size;
    E* kh = k + static_cast<int64_t>(hh) * a.head_size;
    E qx = qh[i], qy = qh[embed + i], kx = kh[i], ky = kh[embed + i];
    rope_pair<T>(qx, qy, c, sn);
    rope_pair<T>(kx, ky, c, sn);
    qh[i] = qx; qh[embed + i] = qy;
    kh[i] = kx; kh[embed + i] = ky;
  }
}

// 16-byte rotary path: both halves of the rotated part, the cos/sin rows and the q/k rows split into whole,
// aligned 16-byte vectors.  Returns log2 of the lanes along (rot_dim / 2) / x, or -1 for the element-wise path.
static int rope_vec_shift(int32_t head_size, int32_t rot_dim, int32_t dtype, const void* query, const void* key,
                          const void* cos_sin, int64_t query_stride, int64_t key_stride) {
  const int es = dtype_size(dtype), x = 16 / es;
  if ((rot_dim / 2) % x != 0 || head_size % x != 0) return -1;
  if ((reinterpret_cast<uintptr_t>(query) | reinterpret_cast<uintptr_t>(key) | reinterpret_cast<uintptr_t>(cos_sin)) & 15u)
    return -1;
  if ((query_stride * es) % 16 != 0 || (key_stride * es) % 16 != 0) return -1;
  int sh = 0;
  while ((1 << sh) < (rot_dim / 2) / x) ++sh;
  return sh;
}

static int rope_checks(const char* op, int32_t num_heads, int32_t head_size, int32_t rot_dim, int32_t dtype, int* shift) {
  if (dtype != PA_F16 && dtype != PA_BF16 && dtype != PA_F32)
    return set_error(PA_ERR_UNSUPPORTED_DTYPE, "%s: unsupported dtype %d", op, dtype);
  if (num_heads <= 0 || head_size <= 0 || rot_dim <= 0 || rot_dim % 2 != 0 || rot_dim > head_size)
    return set_error(PA_ERR_INVALID_ARGUMENT, "%s: bad sizes (heads %d, head_size %d, rot_dim %d)", op, num_heads,
                     head_size, rot_dim);
  int sh = 0;
  while ((1 << sh) < rot_dim / 2) ++sh;
  if ((1 << sh) > kTileThreads) return set_error(PA_ERR_INVALID_ARGUMENT, "%s: rot_dim %d too large", op, rot_dim);
  *shift = sh;
  return PA_OK;
}

static int rotary_embedding_neox(const int64_t* positions, void* query, void* key, const void* cos_sin,
                                 int32_t num_tokens, int32_t num_heads, int32_t head_size, int32_t rot_dim,
                                 int64_t query_stride, int64_t key_stride, int32_t dtype, cudaStream_t stream) {
  int shift = 0;
  int rc = rope_checks("rotary_embedding_neox", num_heads, head_size, rot_dim, dtype, &shift);
  if (rc != PA_OK) return rc;
  if (num_tokens < 0) return set_error(PA_ERR_INVALID_ARGUMENT, "rotary_embedding_neox: negative token count");
  if (num_tokens == 0) return PA_OK;
  if (!positions || !query || !key || !cos_sin) return set_error(PA_ERR_INVALID_ARGUMENT, "rotary_embedding_neox: null pointer");
  RopeArgs a;
  a.positions = positions; a.query = query; a.key = key; a.cos_sin = cos_sin;
  a.query_stride = query_stride; a.key_stride = key_stride;
  a.num_heads = num_heads; a.head_size = head_size; a.rot_dim = rot_dim; a.rot_shift = shift;
  a.rot_vec_shift = rope_vec_shift(head_size, rot_dim, dtype, query, key, cos_sin, query_stride, key_stride);
  int threads = num_heads << (a.rot_vec_shift >= 0 ? a.rot_vec_shift : shift);
  threads = threads > 256 ? 256 : (threads < 32 ? 32 : threads);
  if (dtype == PA_F16) rope_kernel<F16><<<num_tokens, threads, 0, stream>>>(a);
  else if (dtype == PA_BF16) rope_kernel<BF16><<<num_tokens, threads, 0, stream>>>(a);
  else rope_kernel<F32><<<num_tokens, threads, 0, stream>>>(a);
  count_launch();
  return check_cuda(cudaGetLastError(), "rotary_embedding_neox");
}

static int rope_reshape_and_cache(const int64_t* positions, void* query, void* key, const void* value, void* key_cache,
                                  void* value_cache, const int32_t* slot_mapping, const void* cos_sin,
                                  int32_t num_rope_tokens, int32_t num_cache_tokens, int32_t num_heads,
                                  int32_t head_size, int32_t block_size, int32_t rot_dim, int64_t query_stride,
                                  int64_t key_stride, int64_t value_stride, int32_t dtype, cudaStream_t stream) {
  const char* op = "rope_reshape_and_cache";
  int shift = 0;
  int rc = rope_checks(op, num_heads, head_size, rot_dim, dtype, &shift);
  if (rc != PA_OK) return rc;
  if (num_rope_tokens < 0 || num_cache_tokens < 0 || num_cache_tokens > num_rope_tokens || block_size <= 0)
    return set_error(PA_ERR_INVALID_ARGUMENT, "%s: bad token counts (%d rotated, %d cached) or block size %d", op,
                     num_rope_tokens, num_cache_tokens, block_size);
  if (num_rope_tokens == 0) return PA_OK;
  if (!positions || !query || !key || !cos_sin || (num_cache_tokens > 0 && (!value || !key_cache || !value_cache || !slot_mapping)))
    return set_error(PA_ERR_INVALID_ARGUMENT, "%s: null pointer", op);
  const int es = dtype_size(dtype);
  const int x = 16 / es;
  if (head_size % x != 0)
    return set_error(PA_ERR_INVALID_ARGUMENT, "%s: head_size %d is not a multiple of x=%d", op, head_size, x);
  if ((reinterpret_cast<uintptr_t>(key_cache) | reinterpret_cast<uintptr_t>(value_cache)) & 15u)
    return set_error(PA_ERR_INVALID_ARGUMENT, "%s: caches must be 16-byte aligned", op);
  if (num_heads > 65535) return set_error(PA_ERR_INVALID_ARGUMENT, "%s: too many heads (%d)", op, num_heads);

  ScatterArgs a;
  a.key = key; a.value = value; a.key_cache = key_cache; a.value_cache = value_cache;
  a.slot_mapping = slot_mapping; a.key_stride = key_stride; a.value_stride = value_stride;
  a.num_tokens = num_cache_tokens; a.num_heads = num_heads; a.head_size = head_size; a.block_size = block_size;
  a.vec_ok = ((reinterpret_cast<uintptr_t>(key) | reinterpret_cast<uintptr_t>(value)) % 16 == 0) &&
             ((key_stride * es) % 16 == 0) && ((value_stride * es) % 16 == 0);
  a.chunk_shift = 0;
  while ((1 << a.chunk_shift) < head_size / x) ++a.chunk_shift;
  if ((1 << a.chunk_shift) > kTileThreads) return set_error(PA_ERR_INVALID_ARGUMENT, "%s: head_size %d too large", op, head_size);
  a.positions = positions; a.query = query; a.cos_sin = cos_sin; a.query_stride = query_stride;
  a.rot_dim = rot_dim; a.num_rope_tokens = num_rope_tokens; a.rot_shift = shift;
  a.rot_vec_shift = a.vec_ok ? rope_vec_shift(head_size, rot_dim, dtype, query, key, cos_sin, query_stride, key_stride) : -1;

  const int smem = 2 * kTile * (head_size * es + 16);
  if (smem > 200 * 1024) return set_error(PA_ERR_INVALID_ARGUMENT, "%s: head_size %d too large", op, head_size);
  const dim3 grid((num_rope_tokens + kTile - 1) / kTile, num_heads);
  auto launch = [&](auto kern) -> int {
    if (smem > 48 * 1024)
      PA_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    kern<<<grid, kTileThreads, smem, stream>>>(a);
    return PA_OK;
  };
  if (dtype == PA_F16) rc = launch(kv_tile_kernel<uint16_t, false, F16>);
  else if (dtype == PA_BF16) rc = launch(kv_tile_kernel<uint16_t, false, BF16>);
  else rc = launch(kv_tile_kernel<uint32_t, false, F32>);
  if (rc != PA_OK) return rc;
  count_launch();
  return check_cuda(cudaGetLastError(), op);
}

// ---------------------------------------------------------------------------
// copy_blocks: pointer and pair tables ride in the kernel parameters
// ---------------------------------------------------------------------------
constexpr int kCopyMaxLayers = 64;
constexpr int kCopyMaxPairs = 160;
constexpr int kCopyThreads = 256;
constexpr int kCopyUnroll = 4;
constexpr int kCopySegBytes = kCopyThreads * 16 * kCopyUnroll;  // 16 KiB per CTA

struct CopyTable {
  char* kptr[kCopyMaxLayers];
  char* vptr[kCopyMaxLayers];
  int64_t src[kCopyMaxPairs];
  int64_t dst[kCopyMaxPairs];
  int64_t block_bytes;
};
static_assert(sizeof(CopyTable) <= 4000, "kernel parameter space");

// grid: (segments per block, pairs * 2 [key | value], layers)
__global__ void __launch_bounds__(kCopyThreads) copy_blocks_kernel(const __grid_constant__ CopyTable t) {
  const int pair = blockIdx.y >> 1;
  char* base = (blockIdx.y & 1) ? t.vptr[blockIdx.z] : t.kptr[blockIdx.z];
  const char* src = base + t.src[pair] * t.block_bytes;
  char* dst = base + t.dst[pair] * t.block_bytes;
  const int64_t seg = static_cast<int64_t>(blockIdx.x) * kCopySegBytes;
  uint4 v[kCopyUnroll];
  int64_t o[kCopyUnroll];
#pragma unroll
  for (int j = 0; j < kCopyUnroll; ++j) {
    o[j] = seg + (static_cast<int64_t>(j) * kCopyThreads + threadIdx.x) * 16;
    if (o[j] < t.block_bytes) v[j] = ldg_stream(src + o[j]);
  }
#pragma unroll
  for (int j = 0; j < kCopyUnroll; ++j) {
    if (o[j] < t.block_bytes) stg_stream(dst + o[j], v[j]);
  }
}

static int copy_blocks(void* const* key_cache_ptrs, void* const* value_cache_ptrs, int32_t num_layers,
                       const int64_t* pairs, int64_t num_pairs, int64_t block_bytes,
                       cudaStream_t stream) {
  if (num_layers < 0 || num_pairs < 0 || block_bytes < 0)
    return set_error(PA_ERR_INVALID_ARGUMENT, "copy_blocks: negative size");
  if (num_layers == 0 || num_pairs == 0 || block_bytes == 0) return PA_OK;  // reference returns early too (:88-90)
  if (!key_cache_ptrs || !value_cache_ptrs || !pairs)
    return set_error(PA_ERR_INVALID_ARGUMENT, "copy_blocks: null table");
  if (block_bytes % 16 != 0)
    return set_error(PA_ERR_INVALID_ARGUMENT, "copy_blocks: block_bytes %lld is not a multiple of 16",
                     static_cast<long long>(block_bytes));
  for (int l = 0; l < num_layers; ++l) {
    if (!key_cache_ptrs[l] || !value_cache_ptrs[l])
      return set_error(PA_ERR_INVALID_ARGUMENT, "copy_blocks: null cache pointer for layer %d", l);
    if ((reinterpret_cast<uintptr_t>(key_cache_ptrs[l]) | reinterpret_cast<uintptr_t>(value_cache_ptrs[l])) & 15u)
      return set_error(PA_ERR_INVALID_ARGUMENT, "copy_blocks: layer %d cache is not 16-byte aligned", l);
  }
  for (int64_t i = 0; i < 2 * num_pairs; ++i)
    if (pairs[i] < 0) return set_error(PA_ERR_INVALID_ARGUMENT, "copy_blocks: negative block number");

  const unsigned segs = static_cast<unsigned>((block_bytes + kCopySegBytes - 1) / kCopySegBytes);
  for (int l0 = 0; l0 < num_layers; l0 += kCopyMaxLayers) {
    const int nl = num_layers - l0 < kCopyMaxLayers ? num_layers - l0 : kCopyMaxLayers;
    for (int64_t p0 = 0; p0 < num_pairs; p0 += kCopyMaxPairs) {
      const int np = static_cast<int>(num_pairs - p0 < kCopyMaxPairs ? num_pairs - p0 : kCopyMaxPairs);
      CopyTable t;
      for (int l = 0; l < nl; ++l) {
        t.kptr[l] = static_cast<char*>(key_cache_ptrs[l0 + l]);
        t.vptr[l] = static_cast<char*>(value_cache_ptrs[l0 + l]);
      }
      for (int p = 0; p < np; ++p) {
        t.src[p] = pairs[2 * (p0 + p)];
        t.dst[p] = pairs[2 * (p0 + p) + 1];
      }
      t.block_bytes = block_bytes;
      const dim3 grid(segs, 2 * np, nl);
      copy_blocks_kernel<<<grid, kCopyThreads, 0, stream>>>(t);
      count_launch();
      const int rc = check_cuda(cudaGetLastError(), "copy_blocks");
      if (rc != PA_OK) return rc;
    }
  }
  return PA_OK;
}

// ---------------------------------------------------------------------------
// swap_blocks.  Two engines, chosen per call:
//   * few contiguous runs (or pageable host memory): the copy engines, one
//     cudaMemcpyAsync per run of consecutive (src, dst) blocks -- no SM is used
//     and a whole-sequence swap is a handful of DMAs (the reference issues one
//     cudaMemcpyAsync per block, cache_kernels.cu:35-46);
//   * many scattered blocks with pinned host memory (what
//     vllm/worker/cache_engine.py:88-97 allocates) or GPU->GPU: one gather/scatter
//     kernel per <= kSwapMaxPairs blocks that reads/writes the host buffer
//     directly over PCIe (zero-copy through the unified address space); the
//     block table rides in the kernel parameters.  Per-block DMA setup
//     (~7 us for a 160 KiB block) is what limits the first engine here.
// Either way the work is stream-ordered on the caller's (side) stream, so a swap
// still overlaps the model as in cache_engine.py:101-118.
// ---------------------------------------------------------------------------
constexpr int kSwapMaxPairs = 240;
constexpr int kSwapMinRunsForKernel = 6;

struct SwapTable {
  int64_t src_off[kSwapMaxPairs];   // byte offsets
  int64_t dst_off[kSwapMaxPairs];
  const char* src;
  char* dst;
  int64_t block_bytes;
};
static_assert(sizeof(SwapTable) <= 4000, "kernel parameter space");

// grid: (segments per block, blocks)
__global__ void __launch_bounds__(kCopyThreads) swap_blocks_kernel(const __grid_constant__ SwapTable t) {
  const char* src = t.src + t.src_off[blockIdx.y];
  char* dst = t.dst + t.dst_off[blockIdx.y];
  const int64_t seg = static_cast<int64_t>(blockIdx.x) * kCopySegBytes;
  uint4 v[kCopyUnroll];
  int64_t o[kCopyUnroll];
#pragma unroll
  for (int j = 0; j < kCopyUnroll; ++j) {
    o[j] = seg + (static_cast<int64_t>(j) * kCopyThreads + threadIdx.x) * 16;
    if (o[j] < t.block_bytes) v[j] = *reinterpret_cast<const uint4*>(src + o[j]);
  }
#pragma unroll
  for (int j = 0; j < kCopyUnroll; ++j) {
    if (o[j] < t.block_bytes) *reinterpret_cast<uint4*>(dst + o[j]) = v[j];
  }
}

// device-visible alias of a host pointer if it is pinned (mapped), else nullptr
static void* mapped_host_pointer(const void* p) {
  cudaPointerAttributes attr;
  if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  if (attr.type == cudaMemoryTypeHost && attr.devicePointer) return attr.devicePointer;
  return nullptr;
}

static int swap_blocks(const void* src, void* dst, int64_t block_bytes, const int64_t* src_blocks,
                       const int64_t* dst_blocks, int64_t num_blocks, int32_t kind, int32_t engine,
                       cudaStream_t stream) {
  cudaMemcpyKind k;
  switch (kind) {
    case PA_COPY_D2D: k = cudaMemcpyDeviceToDevice; break;
    case PA_COPY_D2H: k = cudaMemcpyDeviceToHost; break;
    case PA_COPY_H2D: k = cudaMemcpyHostToDevice; break;
    default: return set_error(PA_ERR_INVALID_ARGUMENT, "swap_blocks: invalid device combination (kind %d)", kind);
  }
  if (num_blocks < 0 || block_bytes < 0) return set_error(PA_ERR_INVALID_ARGUMENT, "swap_blocks: negative size");
  if (num_blocks == 0 || block_bytes == 0) return PA_OK;
  if (!src || !dst || !src_blocks || !dst_blocks)
    return set_error(PA_ERR_INVALID_ARGUMENT, "swap_blocks: null pointer");
  int64_t runs = 1;
  for (int64_t i = 0; i < num_blocks; ++i) {
    if (src_blocks[i] < 0 || dst_blocks[i] < 0)
      return set_error(PA_ERR_INVALID_ARGUMENT, "swap_blocks: negative block number");
    if (i > 0 && !(src_blocks[i] == src_blocks[i - 1] + 1 && dst_blocks[i] == dst_blocks[i - 1] + 1)) ++runs;
  }

  // engine: 0 automatic, 1 copy engines, 2 kernel (falls back to 1 when the host side is not pinned)
  bool use_kernel = engine == 2 || (engine == 0 && runs >= kSwapMinRunsForKernel);
  const char* ksrc = static_cast<const char*>(src);
  char* kdst = static_cast<char*>(dst);
  if (use_kernel && block_bytes % 16 != 0) use_kernel = false;
  if (use_kernel && kind == PA_COPY_D2H) {
    kdst = static_cast<char*>(mapped_host_pointer(dst));
    use_kernel = kdst != nullptr;
  }
  if (use_kernel && kind == PA_COPY_H2D) {
    ksrc = static_cast<const char*>(mapped_host_pointer(src));
    use_kernel = ksrc != nullptr;
  }
  if (use_kernel && ((reinterpret_cast<uintptr_t>(ksrc) | reinterpret_cast<uintptr_t>(kdst)) & 15u)) use_kernel = false;

  if (use_kernel) {
    const unsigned segs = static_cast<unsigned>((block_bytes + kCopySegBytes - 1) / kCopySegBytes);
    for (int64_t p0 = 0; p0 < num_blocks; p0 += kSwapMaxPairs) {
      const int np = static_cast<int>(num_blocks - p0 < kSwapMaxPairs ? num_blocks - p0 : kSwapMaxPairs);
      SwapTable t;
      for (int p = 0; p < np; ++p) {
        t.src_off[p] = src_blocks[p0 + p] * block_bytes;
        t.dst_off[p] = dst_blocks[p0 + p] * block_bytes;
      }
      t.src = ksrc; t.dst = kdst; t.block_bytes = block_bytes;
      swap_blocks_kernel<<<dim3(segs, np), kCopyThreads, 0, stream>>>(t);
      count_launch();
      const int rc = check_cuda(cudaGetLastError(), "swap_blocks");
      if (rc != PA_OK) return rc;
    }
    return PA_OK;
  }

  int64_t i = 0;
  while (i < num_blocks) {
    int64_t j = i + 1;
    while (j < num_blocks && src_blocks[j] == src_blocks[j - 1] + 1 && dst_blocks[j] == dst_blocks[j - 1] + 1) ++j;
    const char* s = static_cast<const char*>(src) + src_blocks[i] * block_bytes;
    char* d = static_cast<char*>(dst) + dst_blocks[i] * block_bytes;
    PA_CUDA_TRY(cudaMemcpyAsync(d, s, static_cast<size_t>((j - i) * block_bytes), k, stream));
    i = j;
  }
  return PA_OK;
}

}  // namespace pa

// ---------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------
extern "C" {

int pa_reshape_and_cache(const void* key, const void* value, void* key_cache, void* value_cache,
                         const int32_t* slot_mapping, int32_t num_tokens, int32_t num_heads,
                         int32_t head_size, int32_t block_size, int64_t key_stride,
                         int64_t value_stride, int32_t dtype, void* stream) {
  return pa::scatter_gather<false>(key, value, key_cache, value_cache, slot_mapping, num_tokens, num_heads,
                                   head_size, block_size, key_stride, value_stride, dtype,
                                   static_cast<cudaStream_t>(stream));
}

int pa_reshape_and_cache_ex(const void* key, const void* value, void* key_cache, void* value_cache,
                            const int32_t* slot_mapping, int32_t num_tokens, int32_t num_heads,
                            int32_t head_size, int32_t block_size, int64_t key_stride,
                            int64_t value_stride, int32_t dtype, int32_t slot_pattern, void* stream) {
  return pa::scatter_gather<false>(key, value, key_cache, value_cache, slot_mapping, num_tokens, num_heads,
                                   head_size, block_size, key_stride, value_stride, dtype,
                                   static_cast<cudaStream_t>(stream), slot_pattern);
}

int pa_gather_cached_kv(void* key, void* value, const void* key_cache, const void* value_cache,
                        const int32_t* slot_mapping, int32_t num_tokens, int32_t num_heads,
                        int32_t head_size, int32_t block_size, int64_t key_stride,
                        int64_t value_stride, int32_t dtype, void* stream) {
  return pa::scatter_gather<true>(key, value, const_cast<void*>(key_cache), const_cast<void*>(value_cache),
                                  slot_mapping, num_tokens, num_heads, head_size, block_size, key_stride,
                                  value_stride, dtype, static_cast<cudaStream_t>(stream));
}

int pa_rotary_embedding_neox(const int64_t* positions, void* query, void* key, const void* cos_sin_cache,
                             int32_t num_tokens, int32_t num_heads, int32_t head_size, int32_t rot_dim,
                             int64_t query_stride, int64_t key_stride, int32_t dtype, void* stream) {
  return pa::rotary_embedding_neox(positions, query, key, cos_sin_cache, num_tokens, num_heads, head_size, rot_dim,
                                   query_stride, key_stride, dtype, static_cast<cudaStream_t>(stream));
}

int pa_rope_reshape_and_cache(const int64_t* positions, void* query, void* key, const void* value, void* key_cache,
                              void* value_cache, const int32_t* slot_mapping, const void* cos_sin_cache,
                              int32_t num_rope_tokens, int32_t num_cache_tokens, int32_t num_heads,
                              int32_t head_size, int32_t block_size, int32_t rot_dim, int64_t query_stride,
                              int64_t key_stride, int64_t value_stride, int32_t dtype, void* stream) {
  return pa::rope_reshape_and_cache(positions, query, key, value, key_cache, value_cache, slot_mapping, cos_sin_cache,
                                    num_rope_tokens, num_cache_tokens, num_heads, head_size, block_size, rot_dim,
                                    query_stride, key_stride, value_stride, dtype, static_cast<cudaStream_t>(stream));
}

int pa_copy_blocks(void* const* key_cache_ptrs, void* const* value_cache_ptrs, int32_t num_layers,
                   const int64_t* pairs, int64_t num_pairs, int64_t block_bytes, void* stream) {
  return pa::copy_blocks(key_cache_ptrs, value_cache_ptrs, num_layers, pairs, num_pairs, block_bytes,
                         static_cast<cudaStream_t>(stream));
}

int pa_swap_blocks(const void* src, void* dst, int64_t block_bytes, const int64_t* src_blocks,
                   const int64_t* dst_blocks, int64_t num_blocks, int32_t kind, void* stream) {
  return pa::swap_blocks(src, dst, block_bytes, src_blocks, dst_blocks, num_blocks, kind, 0,
                         static_cast<cudaStream_t>(stream));
}

int pa_swap_blocks_ex(const void* src, void* dst, int64_t block_bytes, const int64_t* src_blocks,
                      const int64_t* dst_blocks, int64_t num_blocks, int32_t kind, int32_t engine, void* stream) {
  return pa::swap_blocks(src, dst, block_bytes, src_blocks, dst_blocks, num_blocks, kind, engine,
                         static_cast<cudaStream_t>(stream));
}

}  // extern "C"
