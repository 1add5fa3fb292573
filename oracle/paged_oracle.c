/*
 * paged_oracle.c -- CPU restatement of the reference's PagedAttention decode
 * path in plain C.  TEST INFRASTRUCTURE ONLY: loaded by tests/, smoke() and
 * bench.py's CPU-baseline legs through oracle/c_oracle.py; never by the product.
 *
 * It follows the reference's *CUDA* algorithm (the Python test reference is
 * restated separately in paged_oracle.py), one function per kernel:
 *
 *   po_decode_attention   csrc/attention/attention_kernels.cu:67-314
 *                         two passes per (sequence, head): logits = scale * q.k
 *                         through the block table (:142-178), maximum (:184-201),
 *                         exp(logit - max) and their sum (:204-211),
 *                         1 / (sum + 1e-6) (:213), then P.V (:236-253).
 *                         Arithmetic here is done in double precision so the
 *                         result is a truth value for tolerance checks, not a
 *                         bit-level model of the GPU's fp16 products.
 *   po_reshape_and_cache  csrc/cache_kernels.cu:142-182   (index arithmetic :165-178)
 *   po_gather_cached_kv   csrc/cache_kernels.cu:229-270
 *   po_copy_blocks        csrc/cache_kernels.cu:52-78
 *   po_swap_blocks        csrc/cache_kernels.cu:32-46
 *
 * Parity pinning: checked against the golden files produced from the
 * reference's own Python test references (tests/golden/, tests/test_oracle_golden.py).
 *
 * dtype codes match include/paged_attention_b200.h: 0 = fp16, 1 = bf16, 2 = fp32.
 * Layouts: key_cache [NB][H][D/x][B][x], value_cache [NB][H][D][B], x = 16/sizeof(T).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

static float half_to_float(uint16_t h) {
  const uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
  uint32_t exp = (h >> 10) & 0x1fu;
  uint32_t man = h & 0x3ffu;
  uint32_t bits;
  if (exp == 0) {
    if (man == 0) {
      bits = sign;
    } else { /* subnormal: renormalise */
      int e = -1;
      do { man <<= 1; ++e; } while (!(man & 0x400u));
      bits = sign | ((uint32_t)(127 - 15 - e) << 23) | ((man & 0x3ffu) << 13);
    }
  } else if (exp == 31) {
    bits = sign | 0x7f800000u | (man << 13);
  } else {
    bits = sign | ((exp + 112u) << 23) | (man << 13);
  }
  float f;
  memcpy(&f, &bits, 4);
  return f;
}

static float bf16_to_float(uint16_t b) {
  const uint32_t bits = (uint32_t)b << 16;
  float f;
  memcpy(&f, &bits, 4);
  return f;
}

/* round-to-nearest-even, as cvt.rn / torch */
static uint16_t float_to_half(float f) {
  uint32_t x;
  memcpy(&x, &f, 4);
  const uint32_t sign = (x >> 16) & 0x8000u;
  x &= 0x7fffffffu;
  if (x >= 0x7f800000u) return (uint16_t)(sign | 0x7c00u | ((x > 0x7f800000u) ? 0x200u : 0u));
  if (x >= 0x477ff000u) return (uint16_t)(sign | 0x7c00u); /* rounds to inf */
  if (x < 0x33000001u) return (uint16_t)sign;              /* rounds to zero */
  int exp = (int)(x >> 23) - 127;
  uint32_t man = (x & 0x7fffffu) | 0x800000u;
  int shift;
  uint32_t base;
  if (exp < -14) { shift = 13 + (-14 - exp); base = 0; }
  else { shift = 13; base = (uint32_t)(exp + 15) << 10; man &= 0x7fffffu; }
  uint32_t q = man >> shift;
  const uint32_t rem = man & ((1u << shift) - 1u);
  const uint32_t half = 1u << (shift - 1);
  if (rem > half || (rem == half && (q & 1u))) ++q;
  return (uint16_t)(sign | (base + q));
}

static uint16_t float_to_bf16(float f) {
  uint32_t x;
  memcpy(&x, &f, 4);
  if ((x & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((x >> 16) | 0x40u);
  x += 0x7fffu + ((x >> 16) & 1u);
  return (uint16_t)(x >> 16);
}

static inline float load_elem(const void* base, int64_t idx, int dtype) {
  if (dtype == 2) return ((const float*)base)[idx];
  const uint16_t v = ((const uint16_t*)base)[idx];
  return dtype == 0 ? half_to_float(v) : bf16_to_float(v);
}

static inline void store_elem(void* base, int64_t idx, int dtype, float v) {
  if (dtype == 2) ((float*)base)[idx] = v;
  else ((uint16_t*)base)[idx] = dtype == 0 ? float_to_half(v) : float_to_bf16(v);
}

int po_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* out [S][H][D]; q [S][H][D] with row stride q_stride (elements). */
int po_decode_attention(void* out, const void* q, const void* key_cache, const void* value_cache, float scale,
                        const int32_t* block_tables, const int32_t* context_lens, int32_t num_seqs,
                        int32_t num_heads, int32_t head_size, int32_t block_size, int32_t max_blocks_per_seq,
                        int64_t q_stride, int32_t dtype) {
  if (dtype < 0 || dtype > 2) return -3;
  const int esize = dtype == 2 ? 4 : 2;
  const int x = 16 / esize;
  const int64_t H = num_heads, D = head_size, B = block_size;
  int max_ctx = 0;
  for (int i = 0; i < num_seqs; ++i) if (context_lens[i] > max_ctx) max_ctx = context_lens[i];
  int rc = 0;
#pragma omp parallel
  {
    double* logits = (double*)malloc(sizeof(double) * (size_t)(max_ctx > 0 ? max_ctx : 1));
    double* acc = (double*)malloc(sizeof(double) * (size_t)D);
    float* qf = (float*)malloc(sizeof(float) * (size_t)D);
    if (!logits || !acc || !qf) {
#pragma omp critical
      rc = -4;
    } else {
#pragma omp for collapse(2) schedule(dynamic, 1)
      for (int s = 0; s < num_seqs; ++s) {
        for (int h = 0; h < num_heads; ++h) {
          const int ctx = context_lens[s];
          const int32_t* table = block_tables + (int64_t)s * max_blocks_per_seq;
          for (int d = 0; d < D; ++d) qf[d] = load_elem(q, (int64_t)s * q_stride + h * D + d, dtype);
          /* pass 1: logits and their maximum */
          double mx = -INFINITY;
          for (int t = 0; t < ctx; ++t) {
            const int64_t blk = table[t / B];
            const int off = (int)(t % B);
            const int64_t kbase = (blk * H + h) * D * B; /* [blk][h] page */
            double dot = 0.0;
            for (int d = 0; d < D; ++d) {
              /* [d / x][off][d % x] inside the page */
              const int64_t ki = kbase + (int64_t)(d / x) * B * x + (int64_t)off * x + (d % x);
              dot += (double)qf[d] * (double)load_elem(key_cache, ki, dtype);
            }
            logits[t] = (double)scale * dot;
            if (logits[t] > mx) mx = logits[t];
          }
          /* pass 2: softmax weights */
          double sum = 0.0;
          for (int t = 0; t < ctx; ++t) {
            logits[t] = exp(logits[t] - mx);
            sum += logits[t];
          }
          const double inv = 1.0 / (sum + 1e-6);
          /* pass 3: weighted sum of values */
          for (int d = 0; d < D; ++d) acc[d] = 0.0;
          for (int t = 0; t < ctx; ++t) {
            const int64_t blk = table[t / B];
            const int off = (int)(t % B);
            const int64_t vbase = (blk * H + h) * D * B;
            const double p = logits[t] * inv;
            for (int d = 0; d < D; ++d) acc[d] += p * (double)load_elem(value_cache, vbase + (int64_t)d * B + off, dtype);
          }
          for (int d = 0; d < D; ++d) store_elem(out, ((int64_t)s * H + h) * D + d, dtype, (float)acc[d]);
        }
      }
    }
    free(logits);
    free(acc);
    free(qf);
  }
  return rc;
}

/* Byte-exact movers.  esize = sizeof(T); x = 16 / esize. */
static void scatter_gather(char* key, char* value, char* key_cache, char* value_cache, const int32_t* slot_mapping,
                           int32_t num_tokens, int32_t num_heads, int32_t head_size, int32_t block_size,
                           int64_t key_stride, int64_t value_stride, int esize, int gather) {
  const int x = 16 / esize;
  const int64_t H = num_heads, D = head_size, B = block_size;
#pragma omp parallel for schedule(static)
  for (int t = 0; t < num_tokens; ++t) {
    const int32_t slot = slot_mapping[t];
    if (slot < 0) continue;
    const int64_t blk = slot / block_size;
    const int64_t off = slot % block_size;
    for (int64_t i = 0; i < H * D; ++i) {
      const int64_t h = i / D, d = i % D;
      const int64_t kc = blk * H * (D / x) * B * x + h * (D / x) * B * x + (d / x) * B * x + off * x + (d % x);
      const int64_t vc = blk * H * D * B + h * D * B + d * B + off;
      char* kp = key + ((int64_t)t * key_stride + i) * esize;
      char* vp = value + ((int64_t)t * value_stride + i) * esize;
      if (gather) {
        memcpy(kp, key_cache + kc * esize, (size_t)esize);
        memcpy(vp, value_cache + vc * esize, (size_t)esize);
      } else {
        memcpy(key_cache + kc * esize, kp, (size_t)esize);
        memcpy(value_cache + vc * esize, vp, (size_t)esize);
      }
    }
  }
}

void po_reshape_and_cache(const void* key, const void* value, void* key_cache, void* value_cache,
                          const int32_t* slot_mapping, int32_t num_tokens, int32_t num_heads, int32_t head_size,
                          int32_t block_size, int64_t key_stride, int64_t value_stride, int32_t esize) {
  scatter_gather((char*)key, (char*)value, (char*)key_cache, (char*)value_cache, slot_mapping, num_tokens, num_heads,
                 head_size, block_size, key_stride, value_stride, esize, 0);
}

void po_gather_cached_kv(void* key, void* value, const void* key_cache, const void* value_cache,
                         const int32_t* slot_mapping, int32_t num_tokens, int32_t num_heads, int32_t head_size,
                         int32_t block_size, int64_t key_stride, int64_t value_stride, int32_t esize) {
  scatter_gather((char*)key, (char*)value, (char*)key_cache, (char*)value_cache, slot_mapping, num_tokens, num_heads,
                 head_size, block_size, key_stride, value_stride, esize, 1);
}

/* pairs = {src0, dst0, src1, dst1, ...}; applied in order, layer by layer. */
void po_copy_blocks(void* const* key_caches, void* const* value_caches, int32_t num_layers, const int64_t* pairs,
                    int64_t num_pairs, int64_t block_bytes) {
  for (int l = 0; l < num_layers; ++l) {
    for (int64_t p = 0; p < num_pairs; ++p) {
      const int64_t s = pairs[2 * p], d = pairs[2 * p + 1];
      memmove((char*)key_caches[l] + d * block_bytes, (char*)key_caches[l] + s * block_bytes, (size_t)block_bytes);
      memmove((char*)value_caches[l] + d * block_bytes, (char*)value_caches[l] + s * block_bytes, (size_t)block_bytes);
    }
  }
}

void po_swap_blocks(const void* src, void* dst, int64_t block_bytes, const int64_t* src_blocks,
                    const int64_t* dst_blocks, int64_t num_blocks) {
  for (int64_t i = 0; i < num_blocks; ++i)
    memcpy((char*)dst + dst_blocks[i] * block_bytes, (const char*)src + src_blocks[i] * block_bytes, (size_t)block_bytes);
}

/* GPT-NeoX rotary embedding in place (csrc/pos_encoding_kernels.cu:5-47).  As in the reference's
 * c10::Half / c10::BFloat16 / float arithmetic every product and every sum is rounded to T before
 * it is used again; compiled with -ffp-contract=off so no multiply-add is fused. */
static inline float round_to(int dtype, float v) {
  if (dtype == 2) return v;
  return dtype == 0 ? half_to_float(float_to_half(v)) : bf16_to_float(float_to_bf16(v));
}

void po_rotary_embedding_neox(const int64_t* positions, void* query, void* key, const void* cos_sin_cache,
                              int32_t num_tokens, int32_t num_heads, int32_t head_size, int32_t rot_dim,
                              int64_t query_stride, int64_t key_stride, int32_t dtype) {
  const int embed = rot_dim / 2;
#pragma omp parallel for schedule(static)
  for (int t = 0; t < num_tokens; ++t) {
    const int64_t pos = positions[t];
    for (int w = 0; w < 2; ++w) {
      void* base = w == 0 ? query : key;
      const int64_t row = (int64_t)t * (w == 0 ? query_stride : key_stride);
      for (int h = 0; h < num_heads; ++h) {
        for (int i = 0; i < embed; ++i) {
          const float c = load_elem(cos_sin_cache, pos * rot_dim + i, dtype);
          const float s = load_elem(cos_sin_cache, pos * rot_dim + embed + i, dtype);
          const int64_t ix = row + (int64_t)h * head_size + i, iy = ix + embed;
          const float x = load_elem(base, ix, dtype), y = load_elem(base, iy, dtype);
          const float xc = round_to(dtype, x * c), ys = round_to(dtype, y * s);
          const float yc = round_to(dtype, y * c), xs = round_to(dtype, x * s);
          store_elem(base, ix, dtype, xc - ys);
          store_elem(base, iy, dtype, yc + xs);
        }
      }
    }
  }
}
