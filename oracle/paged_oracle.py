"""CPU oracle for the PagedAttention decode path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import
this module; nothing under vllm-v0.1.0_b200/ does.  It restates, in plain
PyTorch on CPU tensors, the reference's own *test references* for the path
(/root/reference is not available on the GPU box, so the restatement travels
instead):

  decode_attention_loop   tests/kernels/test_attention.py:30-66  (ref_single_query_cached_kv_attention)
                          + :14-27 (ref_masked_attention)
  prefill_attention_loop  tests/kernels/test_attention.py:69-101 (ref_multi_query_kv_attention)
  cached_prefill_attention_loop  tests/kernels/test_attention.py:104-156 (ref_multi_query_cached_kv_attention)
  reshape_and_cache       tests/kernels/test_cache.py:95-100
  gather_cached_kv        tests/kernels/test_cache.py:138-143
  copy_blocks             tests/kernels/test_cache.py:51-56
  swap_blocks             csrc/cache_kernels.cu:32-46 (the reference has no test for it)
  rotary_embedding_neox   tests/kernels/test_pos_encoding.py:12-77 (RefRotaryEmbeddingNeox)

Parity pinning: tests/golden/make_golden.py runs the reference's functions
themselves (extracted from /root/reference with `ast`) on seeded inputs and
stores their outputs; tests/test_oracle_golden.py checks this module against
those files.  swap_blocks is "parity unpinned" by the reference (no test, no
golden vector exists there); it is pinned here against plain index assignment.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import torch


# ---------------------------------------------------------------------------
# layout (vllm/worker/cache_engine.py:48-63 of the reference)
# ---------------------------------------------------------------------------
def pack_factor(dtype: torch.dtype) -> int:
    """x: elements per 16-byte group along head_size in the key cache."""
    return 16 // torch.empty((), dtype=dtype).element_size()


def kv_cache_shapes(num_blocks: int, num_heads: int, head_size: int, block_size: int, dtype: torch.dtype):
    x = pack_factor(dtype)
    return ((num_blocks, num_heads, head_size // x, block_size, x),
            (num_blocks, num_heads, head_size, block_size))


# ---------------------------------------------------------------------------
# decode attention
# ---------------------------------------------------------------------------
def _masked_attention(query: torch.Tensor, key: torch.Tensor, value: torch.Tensor, scale: float) -> torch.Tensor:
    """test_attention.py:14-27 without the (unused on this path) mask: scores and
    softmax stay in the input dtype, exactly as the reference computes them."""
    scores = torch.einsum("qhd,khd->hqk", query * scale, key)
    probs = torch.softmax(scores, dim=-1)
    return torch.einsum("hqk,khd->qhd", probs, value)


@torch.inference_mode()
def decode_attention_loop(output: torch.Tensor, query: torch.Tensor, key_cache: torch.Tensor,
                          value_cache: torch.Tensor, block_tables: torch.Tensor, context_lens: torch.Tensor,
                          scale: Optional[float] = None) -> None:
    """Token-by-token gather then dense attention, per sequence (test_attention.py:30-66).

    `scale` defaults to 1/sqrt(head_size), which the reference hard-codes (:63).
    Sequences with context_len == 0 get zeros (the reference oracle cannot
    express them; its CUDA kernel writes zeros, attention_kernels.cu:142,213).
    """
    num_heads, head_size, block_size = value_cache.shape[1:4]
    if scale is None:
        scale = 1.0 / (head_size ** 0.5)
    tables = block_tables.tolist()
    lens = [int(v) for v in context_lens.tolist()]
    for i, ctx in enumerate(lens):
        if ctx == 0:
            output[i].zero_()
            continue
        keys, values = [], []
        for tok in range(ctx):
            blk = tables[i][tok // block_size]
            off = tok % block_size
            keys.append(key_cache[blk, :, :, off, :].reshape(num_heads, head_size))
            values.append(value_cache[blk, :, :, off])
        out = _masked_attention(query[i].unsqueeze(0), torch.stack(keys), torch.stack(values), scale)
        output[i].copy_(out.view(num_heads, head_size))


@torch.inference_mode()
def decode_attention_fp32(query: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                          block_tables: torch.Tensor, context_lens: torch.Tensor, scale: float) -> torch.Tensor:
    """Same mathematics in fp32 with a vectorised gather; returns fp32 [S, H, D].

    Used where the loop oracle is too slow and as the high-precision truth for
    unit-scale inputs.  Proven equal to decode_attention_loop (fp32) in
    tests/test_oracle_golden.py.
    """
    num_blocks, num_heads, head_size, block_size = value_cache.shape
    out = torch.zeros(query.shape[0], num_heads, head_size, dtype=torch.float32)
    kc = key_cache.permute(0, 3, 1, 2, 4).reshape(num_blocks, block_size, num_heads, head_size)  # [blk, off, h, d]
    vc = value_cache.permute(0, 3, 1, 2)                                                        # [blk, off, h, d]
    for i, ctx in enumerate(int(v) for v in context_lens.tolist()):
        if ctx == 0:
            continue
        nblk = (ctx + block_size - 1) // block_size
        blks = block_tables[i, :nblk].long()
        k = kc[blks].reshape(nblk * block_size, num_heads, head_size)[:ctx].float()
        v = vc[blks].reshape(nblk * block_size, num_heads, head_size)[:ctx].float()
        q = query[i].float() * scale
        scores = torch.einsum("hd,khd->hk", q, k)
        probs = torch.softmax(scores, dim=-1)
        out[i] = torch.einsum("hk,khd->hd", probs, v)
    return out


# ---------------------------------------------------------------------------
# prompt (prefill) attention
# ---------------------------------------------------------------------------
@torch.inference_mode()
def prefill_attention_loop(seq_lens: Sequence[int], query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                           scale: Optional[float] = None) -> torch.Tensor:
    """Per-prompt dense causal attention (test_attention.py:69-101): an additive upper-triangular
    mask of finfo(dtype).min, scores / softmax in the input dtype (:14-27).  query/key/value are
    [num_tokens, num_heads, head_size]; prompts are packed along the token dimension."""
    dtype = query.dtype
    if scale is None:
        scale = 1.0 / (query.shape[-1] ** 0.5)
    outs, start = [], 0
    for n in seq_lens:
        end = start + int(n)
        mask = torch.triu(torch.ones(n, n, dtype=dtype), diagonal=1) * torch.finfo(dtype).min
        scores = torch.einsum("qhd,khd->hqk", query[start:end] * scale, key[start:end]) + mask
        probs = torch.softmax(scores, dim=-1)
        outs.append(torch.einsum("hqk,khd->qhd", probs, value[start:end]))
        start = end
    return torch.cat(outs, dim=0) if outs else query[:0].clone()


@torch.inference_mode()
def prefill_attention_fp32(seq_lens: Sequence[int], query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                           scale: Optional[float] = None) -> torch.Tensor:
    """Same attention with every intermediate in fp32 on the (already rounded) inputs: the yardstick
    for unit-scale inputs, where the reference's in-dtype softmax is itself only ~1e-3 accurate."""
    if scale is None:
        scale = 1.0 / (query.shape[-1] ** 0.5)
    q, k, v = query.float(), key.float(), value.float()
    outs, start = [], 0
    for n in seq_lens:
        end = start + int(n)
        scores = torch.einsum("qhd,khd->hqk", q[start:end], k[start:end]) * scale
        scores = scores.masked_fill(torch.triu(torch.ones(n, n, dtype=torch.bool), diagonal=1), float("-inf"))
        outs.append(torch.einsum("hqk,khd->qhd", torch.softmax(scores, dim=-1), v[start:end]))
        start = end
    return torch.cat(outs, dim=0) if outs else q[:0].clone()


# ---------------------------------------------------------------------------
# prompt chunks over the paged cache
# ---------------------------------------------------------------------------
def _context_rows(key_cache: torch.Tensor, value_cache: torch.Tensor, block_table: torch.Tensor, context_len: int):
    """Keys / values of the first context_len cached tokens as [context_len, num_heads, head_size]
    (the token-by-token gather of test_attention.py:131-143)."""
    num_heads, head_size, block_size = value_cache.shape[1], value_cache.shape[2], value_cache.shape[3]
    pos = torch.arange(context_len)
    blocks = block_table.long()[pos // block_size]
    offs = pos % block_size
    keys = key_cache[blocks, :, :, offs, :].reshape(context_len, num_heads, head_size)
    values = value_cache[blocks, :, :, offs]
    return keys, values


@torch.inference_mode()
def cached_prefill_attention_loop(cu_query_lens: Sequence[int], query: torch.Tensor, key_cache: torch.Tensor,
                                  value_cache: torch.Tensor, block_tables: torch.Tensor, context_lens,
                                  scale: Optional[float] = None) -> torch.Tensor:
    """Per sequence: gather the cached context, add a mask of -1e5 above the diagonal shifted by
    context_len - query_len, dense attention with scores / softmax in the input dtype
    (test_attention.py:104-156 with :14-27)."""
    dtype = query.dtype
    if scale is None:
        scale = 1.0 / (query.shape[-1] ** 0.5)
    outs = []
    for i in range(len(cu_query_lens) - 1):
        start, end = int(cu_query_lens[i]), int(cu_query_lens[i + 1])
        qlen, ctx = end - start, int(context_lens[i])
        keys, values = _context_rows(key_cache, value_cache, block_tables[i], ctx)
        mask = (torch.triu(torch.ones(qlen, ctx), diagonal=ctx - qlen + 1) * -1e5).to(dtype)
        scores = torch.einsum("qhd,khd->hqk", query[start:end] * scale, keys) + mask
        probs = torch.softmax(scores, dim=-1)
        outs.append(torch.einsum("hqk,khd->qhd", probs, values))
    return torch.cat(outs, dim=0) if outs else query[:0].clone()


@torch.inference_mode()
def cached_prefill_attention_fp32(cu_query_lens: Sequence[int], query: torch.Tensor, key_cache: torch.Tensor,
                                  value_cache: torch.Tensor, block_tables: torch.Tensor, context_lens,
                                  scale: Optional[float] = None) -> torch.Tensor:
    """The same with every intermediate in fp32 on the (already rounded) inputs and a true -inf mask."""
    if scale is None:
        scale = 1.0 / (query.shape[-1] ** 0.5)
    outs = []
    for i in range(len(cu_query_lens) - 1):
        start, end = int(cu_query_lens[i]), int(cu_query_lens[i + 1])
        qlen, ctx = end - start, int(context_lens[i])
        keys, values = _context_rows(key_cache, value_cache, block_tables[i], ctx)
        scores = torch.einsum("qhd,khd->hqk", query[start:end].float(), keys.float()) * scale
        hidden = torch.triu(torch.ones(qlen, ctx, dtype=torch.bool), diagonal=ctx - qlen + 1)
        scores = scores.masked_fill(hidden, float("-inf"))
        outs.append(torch.einsum("hqk,khd->qhd", torch.softmax(scores, dim=-1), values.float()))
    return torch.cat(outs, dim=0) if outs else query[:0].float().clone()


# ---------------------------------------------------------------------------
# cache movers (bit-exact)
# ---------------------------------------------------------------------------
@torch.inference_mode()
def reshape_and_cache(key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                      slot_mapping: torch.Tensor) -> None:
    """test_cache.py:95-100: key viewed as [N, H, D/x, x] lands at [blk, :, :, off, :]."""
    num_tokens, num_heads, head_size = key.shape
    block_size, x = key_cache.shape[3], key_cache.shape[4]
    packed = key.reshape(num_tokens, num_heads, head_size // x, x)
    for i, slot in enumerate(slot_mapping.tolist()[:num_tokens]):
        if slot < 0:
            continue
        blk, off = slot // block_size, slot % block_size
        key_cache[blk, :, :, off, :] = packed[i]
        value_cache[blk, :, :, off] = value[i]


@torch.inference_mode()
def gather_cached_kv(key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                     slot_mapping: torch.Tensor) -> None:
    """test_cache.py:138-143: the inverse assignment."""
    num_tokens, num_heads, head_size = key.shape
    block_size = key_cache.shape[3]
    for i, slot in enumerate(slot_mapping.tolist()[:num_tokens]):
        if slot < 0:
            continue
        blk, off = slot // block_size, slot % block_size
        key[i] = key_cache[blk, :, :, off, :].reshape(num_heads, head_size)
        value[i] = value_cache[blk, :, :, off]


@torch.inference_mode()
def copy_blocks(key_caches: Sequence[torch.Tensor], value_caches: Sequence[torch.Tensor],
                block_mapping: Dict[int, List[int]]) -> None:
    """test_cache.py:51-56: every layer, key and value cache, dst block <- src block."""
    for src in sorted(block_mapping):
        for dst in block_mapping[src]:
            for cache in list(key_caches) + list(value_caches):
                cache[dst] = cache[src]


@torch.inference_mode()
def swap_blocks(src: torch.Tensor, dst: torch.Tensor, block_mapping: Dict[int, int]) -> None:
    """csrc/cache_kernels.cu:32-46: dst[d] = src[s], whole blocks, in src order."""
    for s in sorted(block_mapping):
        dst[block_mapping[s]] = src[s].to(dst.device)


# ---------------------------------------------------------------------------
# rotary embedding (the operator next to the path)
# ---------------------------------------------------------------------------
def _rotate_half(x: torch.Tensor) -> torch.Tensor:
    """tests/kernels/test_pos_encoding.py:12-16."""
    half = x.shape[-1] // 2
    return torch.cat((-x[..., half:], x[..., :half]), dim=-1)


@torch.inference_mode()
def rotary_embedding_neox(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor, head_size: int,
                          cos_sin_cache: torch.Tensor) -> None:
    """GPT-NeoX rotary embedding in place on query / key ([N, H*D] or [N, H, D], possibly row-strided), in the
    tensors' own dtype: q*cos + rotate_half(q)*sin on the first rot_dim elements of every head
    (tests/kernels/test_pos_encoding.py:19-77 RefRotaryEmbeddingNeox; cos/sin rows come from the
    [max_position, rot_dim] cache the operator is given, csrc/pos_encoding_kernels.cu:19-36)."""
    n = query.shape[0]
    rot = cos_sin_cache.shape[1]
    pos = positions.long()
    cos = cos_sin_cache[pos, : rot // 2]
    sin = cos_sin_cache[pos, rot // 2:]
    cos = torch.cat((cos, cos), dim=-1).unsqueeze(1)          # [N, 1, rot]
    sin = torch.cat((sin, sin), dim=-1).unsqueeze(1)
    for t in (query, key):
        v = t.view(n, -1, head_size) if t.dim() == 2 else t
        part = v[..., :rot]
        part.copy_((part * cos) + (_rotate_half(part) * sin))


def rope_cos_sin_cache(rot_dim: int, max_position: int, dtype: torch.dtype, base: int = 10000) -> torch.Tensor:
    """[max_position, rot_dim] table as vllm/model_executor/layers/attention.py:172-186 builds it."""
    inv_freq = 1.0 / (base ** (torch.arange(0, rot_dim, 2) / rot_dim))
    t = torch.arange(max_position).float()
    freqs = torch.einsum("i,j -> ij", t, inv_freq.float())
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1).to(dtype)


# ---------------------------------------------------------------------------
# accounting shared by bench.py and the docs
# ---------------------------------------------------------------------------
def decode_algorithmic_bytes(context_lens: Sequence[int], num_heads: int, head_size: int, block_size: int,
                             elem_size: int) -> int:
    """Bytes a perfect decode-attention kernel must move: K and V of exactly
    context_len tokens per head, q read + out write per head, the used part of the
    block-table row and the context length once per sequence."""
    total = 0
    for ctx in context_lens:
        total += num_heads * (2 * ctx * head_size * elem_size + 2 * head_size * elem_size)
        total += ((ctx + block_size - 1) // block_size) * 4 + 4
    return total
