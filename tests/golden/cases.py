"""Seeded input generators shared by make_golden.py (run where /root/reference
exists) and the tests (run anywhere).  Inputs come from numpy's frozen
RandomState streams, so they are identical on every machine; the golden files
store a SHA-256 of the inputs next to the reference outputs to prove it.
"""
from __future__ import annotations

import hashlib
from typing import Dict, List

import numpy as np
import torch

DTYPES = {"f16": torch.float16, "bf16": torch.bfloat16, "f32": torch.float32}
HEAD_SIZES = (64, 80, 96, 128)
BLOCK_SIZES = (8, 16, 32)


def _tensor(rs: np.random.RandomState, shape, dtype: torch.dtype, dist: str) -> torch.Tensor:
    if dist == "ref":      # the reference tests' distribution (test_attention.py:170,177,181)
        a = rs.uniform(-1e-3, 1e-3, size=shape)
    elif dist == "unit":   # unit-scale values: a non-vacuous check at the stated tolerance
        a = rs.standard_normal(size=shape)
    else:
        raise ValueError(dist)
    return torch.from_numpy(a.astype(np.float32)).to(dtype)


def attention_case_ids() -> List[str]:
    ids = []
    for dt in DTYPES:
        for b in BLOCK_SIZES:
            for d in HEAD_SIZES:
                ids.append(f"{dt}-b{b}-d{d}-ref")
    # unit-scale inputs on the shapes the benchmarks use
    for dt in DTYPES:
        ids.append(f"{dt}-b16-d128-unit")
        ids.append(f"{dt}-b8-d80-unit")
        ids.append(f"{dt}-b32-d96-unit")
    return ids


def attention_case(case_id: str) -> Dict[str, object]:
    """Small decode-attention problem: 5 sequences, 3 heads, 24 blocks, context <= 180,
    block tables with repeated blocks, query sliced out of a fused qkv tensor."""
    dt, b, d, dist = case_id.split("-")
    dtype, block_size, head_size = DTYPES[dt], int(b[1:]), int(d[1:])
    seed = int(hashlib.sha256(case_id.encode()).hexdigest()[:8], 16)
    rs = np.random.RandomState(seed)
    num_seqs, num_heads, num_blocks, max_ctx = 5, 3, 24, 180
    x = 16 // torch.empty((), dtype=dtype).element_size()
    qkv = _tensor(rs, (num_seqs, 3, num_heads, head_size), dtype, dist)
    query = qkv[:, 0]                                      # row stride 3*H*D, as in the reference test (:168-171)
    key_cache = _tensor(rs, (num_blocks, num_heads, head_size // x, block_size, x), dtype, dist)
    value_cache = _tensor(rs, (num_blocks, num_heads, head_size, block_size), dtype, dist)
    ctx = [int(v) for v in rs.randint(1, max_ctx + 1, size=num_seqs)]
    ctx[0] = block_size * 3            # ends exactly on a block boundary
    ctx[1] = 1                         # single token
    max_blocks = (max(ctx) + block_size - 1) // block_size + 2     # rows longer than needed (zero padded)
    tables = np.zeros((num_seqs, max_blocks), dtype=np.int32)
    for i, c in enumerate(ctx):
        n = (c + block_size - 1) // block_size
        tables[i, :n] = rs.randint(0, num_blocks, size=n)
    return {
        "dtype": dtype, "block_size": block_size, "head_size": head_size, "num_heads": num_heads,
        "scale": float(1.0 / (head_size ** 0.5)),
        "query": query, "key_cache": key_cache, "value_cache": value_cache,
        "block_tables": torch.from_numpy(tables), "context_lens": torch.tensor(ctx, dtype=torch.int32),
        "max_context_len": max(ctx),
    }


def attention_large_case_ids() -> List[str]:
    """BASELINE.json configs[0] exactly as the reference test builds it, and one long-context case per dtype
    and input distribution (the small cases above stop at 180 tokens)."""
    return ["cfg1-f32-ref"] + [f"{dt}-long-{dist}" for dt in DTYPES for dist in ("ref", "unit")]


def attention_large_case(case_id: str) -> Dict[str, object]:
    """cfg1-f32-ref: 7 sequences x 40 heads x 128, block 16, 1024 blocks, fp32, values U(-1e-3, 1e-3), context
    lengths and block tables drawn from Python's `random` seeded with 0 in the order the reference test draws
    them (tests/kernels/test_attention.py:183-195: all lengths in [1, 4096] first, then one row of
    ceil(max/16) block numbers in [0, 1023] per sequence -- rows are longer than most contexts need and blocks
    repeat within and across sequences).
    {dt}-long-{dist}: 3 sequences x 2 heads x 128, block 16, contexts 4096 / 4090 / 3000 over 320 blocks."""
    import random
    seed = int(hashlib.sha256(("large-" + case_id).encode()).hexdigest()[:8], 16)
    rs = np.random.RandomState(seed)
    if case_id == "cfg1-f32-ref":
        dtype, dist = torch.float32, "ref"
        num_seqs, num_heads, head_size, block_size, num_blocks = 7, 40, 128, 16, 1024
        rnd = random.Random(0)
        ctx = [rnd.randint(1, 4096) for _ in range(num_seqs)]
        max_blocks = (max(ctx) + block_size - 1) // block_size
        tables = np.array([[rnd.randint(0, num_blocks - 1) for _ in range(max_blocks)] for _ in range(num_seqs)],
                          dtype=np.int32)
    else:
        dt, _, dist = case_id.split("-")
        dtype = DTYPES[dt]
        num_seqs, num_heads, head_size, block_size, num_blocks = 3, 2, 128, 16, 320
        ctx = [4096, 4090, 3000]
        max_blocks = (max(ctx) + block_size - 1) // block_size
        tables = rs.randint(0, num_blocks, size=(num_seqs, max_blocks)).astype(np.int32)
    x = 16 // torch.empty((), dtype=dtype).element_size()
    qkv = _tensor(rs, (num_seqs, 3, num_heads, head_size), dtype, dist)
    key_cache = _tensor(rs, (num_blocks, num_heads, head_size // x, block_size, x), dtype, dist)
    value_cache = _tensor(rs, (num_blocks, num_heads, head_size, block_size), dtype, dist)
    return {
        "dtype": dtype, "block_size": block_size, "head_size": head_size, "num_heads": num_heads,
        "scale": float(1.0 / (head_size ** 0.5)),
        "query": qkv[:, 0], "key_cache": key_cache, "value_cache": value_cache,
        "block_tables": torch.from_numpy(tables), "context_lens": torch.tensor(ctx, dtype=torch.int32),
        "max_context_len": max(ctx),
    }


def cache_case_ids() -> List[str]:
    return [f"{dt}-b{b}-d{d}" for dt in DTYPES for (b, d) in ((8, 16), (16, 64), (32, 128), (16, 80))]


def cache_case(case_id: str) -> Dict[str, object]:
    """Inputs for reshape_and_cache / gather_cached_kv / copy_blocks / swap_blocks."""
    dt, b, d = case_id.split("-")
    dtype, block_size, head_size = DTYPES[dt], int(b[1:]), int(d[1:])
    seed = int(hashlib.sha256(("cache" + case_id).encode()).hexdigest()[:8], 16)
    rs = np.random.RandomState(seed)
    num_tokens, num_heads, num_blocks, num_layers = 7, 3, 10, 3
    x = 16 // torch.empty((), dtype=dtype).element_size()
    qkv = _tensor(rs, (num_tokens, 3, num_heads, head_size), dtype, "unit")
    slots = rs.choice(num_blocks * block_size, size=num_tokens, replace=False).astype(np.int32)
    kshape = (num_blocks, num_heads, head_size // x, block_size, x)
    vshape = (num_blocks, num_heads, head_size, block_size)
    key_caches = [_tensor(rs, kshape, dtype, "unit") for _ in range(num_layers)]
    value_caches = [_tensor(rs, vshape, dtype, "unit") for _ in range(num_layers)]
    perm = rs.permutation(num_blocks)
    srcs, dsts = perm[:3], perm[3:8]
    copy_map = {int(srcs[0]): [int(dsts[0])], int(srcs[1]): [int(dsts[1]), int(dsts[2])], int(srcs[2]): [int(dsts[3])]}
    swap_map = {int(s): int(t) for s, t in zip(rs.permutation(num_blocks)[:4], rs.permutation(num_blocks)[:4])}
    return {
        "dtype": dtype, "block_size": block_size, "head_size": head_size, "num_heads": num_heads,
        "key": qkv[:, 1], "value": qkv[:, 2], "slot_mapping": torch.from_numpy(slots),
        "key_caches": key_caches, "value_caches": value_caches, "copy_map": copy_map, "swap_map": swap_map,
    }


def rope_case_ids() -> List[str]:
    # dtype x (head_size, rotary_dim): the reference test's head sizes (test_pos_encoding.py:138) with full
    # rotation, plus partial rotation as GPT-NeoX uses it (rotary_pct 0.25: 96 -> 24)
    shapes = [(32, 32), (64, 64), (80, 80), (96, 96), (128, 128), (160, 160), (192, 192), (256, 256), (96, 24), (128, 64)]
    return [f"{dt}-d{d}-r{r}" for dt in DTYPES for (d, r) in shapes]


def rope_case(case_id: str) -> Dict[str, object]:
    """Inputs of rotary_embedding_neox: 53 tokens, 5 heads, positions < 2048, q/k sliced out of a fused qkv."""
    dt, d, r = case_id.split("-")
    dtype, head_size, rot_dim = DTYPES[dt], int(d[1:]), int(r[1:])
    seed = int(hashlib.sha256(("rope" + case_id).encode()).hexdigest()[:8], 16)
    rs = np.random.RandomState(seed)
    num_tokens, num_heads, max_position = 53, 5, 2048
    qkv = _tensor(rs, (num_tokens, 3, num_heads * head_size), dtype, "unit")
    positions = torch.from_numpy(rs.randint(0, max_position, size=num_tokens).astype(np.int64))
    return {"dtype": dtype, "head_size": head_size, "rot_dim": rot_dim, "num_heads": num_heads,
            "max_position": max_position, "qkv": qkv, "positions": positions}


def prefill_case_ids() -> List[str]:
    ids = [f"{dt}-d{d}-ref" for dt in DTYPES for d in HEAD_SIZES]
    ids += [f"{dt}-d{d}-unit" for dt in DTYPES for d in (128, 80)]
    return ids


def prefill_case(case_id: str) -> Dict[str, object]:
    """Prompt attention: 4 prompts packed along the token dimension (lengths 1, 130, 70 and 128: a
    single token, one that crosses a 128-row tile, a short one, and exactly one tile), 3 heads,
    q/k/v sliced out of a fused qkv tensor as the model does (test_attention.py:236-239)."""
    dt, d, dist = case_id.split("-")
    dtype, head_size = DTYPES[dt], int(d[1:])
    seed = int(hashlib.sha256(("prefill-" + case_id).encode()).hexdigest()[:8], 16)
    rs = np.random.RandomState(seed)
    seq_lens = [1, 130, 70, 128]
    num_heads = 3
    qkv = _tensor(rs, (sum(seq_lens), 3, num_heads, head_size), dtype, dist)
    return {"dtype": dtype, "head_size": head_size, "num_heads": num_heads, "seq_lens": seq_lens,
            "scale": float(1.0 / (head_size ** 0.5)), "qkv": qkv,
            "query": qkv[:, 0], "key": qkv[:, 1], "value": qkv[:, 2]}


def cached_prefill_case_ids() -> List[str]:
    ids = [f"{dt}-b16-d{d}-ref" for dt in DTYPES for d in HEAD_SIZES]
    ids += [f"{dt}-b{b}-d{d}-unit" for dt in DTYPES for (b, d) in ((16, 128), (8, 80), (32, 96))]
    return ids


def cached_prefill_case(case_id: str) -> Dict[str, object]:
    """Prompt chunks over the paged cache (ref_multi_query_cached_kv_attention, test_attention.py:104-156):
    5 sequences, 3 heads, (query_len, context_len) = (1, 1) a lone first token, (70, 70) no cached prefix,
    (130, 300) a chunk crossing a 128-row tile behind 170 cached tokens, (1, 257) a decode-like row,
    (128, 133) a diagonal shifted by 5; block tables with repeated blocks and zero padding; the query
    sliced out of a fused qkv tensor."""
    dt, b, d, dist = case_id.split("-")
    dtype, block_size, head_size = DTYPES[dt], int(b[1:]), int(d[1:])
    seed = int(hashlib.sha256(("cached-prefill-" + case_id).encode()).hexdigest()[:8], 16)
    rs = np.random.RandomState(seed)
    num_heads, num_blocks = 3, 40
    pairs = [(1, 1), (70, 70), (130, 300), (1, 257), (128, 133)]
    qlens = [p[0] for p in pairs]
    ctx = [p[1] for p in pairs]
    cu = [0]
    for n in qlens:
        cu.append(cu[-1] + n)
    x = 16 // torch.empty((), dtype=dtype).element_size()
    qkv = _tensor(rs, (cu[-1], 3, num_heads, head_size), dtype, dist)
    key_cache = _tensor(rs, (num_blocks, num_heads, head_size // x, block_size, x), dtype, dist)
    value_cache = _tensor(rs, (num_blocks, num_heads, head_size, block_size), dtype, dist)
    max_blocks = (max(ctx) + block_size - 1) // block_size + 1
    tables = np.zeros((len(pairs), max_blocks), dtype=np.int32)
    for i, c in enumerate(ctx):
        n = (c + block_size - 1) // block_size
        tables[i, :n] = rs.randint(0, num_blocks, size=n)
    return {"dtype": dtype, "block_size": block_size, "head_size": head_size, "num_heads": num_heads,
            "scale": float(1.0 / (head_size ** 0.5)), "qkv": qkv, "query": qkv[:, 0], "cu_query_lens": cu,
            "key_cache": key_cache, "value_cache": value_cache, "block_tables": torch.from_numpy(tables),
            "context_lens": torch.tensor(ctx, dtype=torch.int32)}


def to_numpy(t: torch.Tensor) -> np.ndarray:
    """Bit pattern of a tensor as a numpy array (bf16/f16 -> uint16)."""
    t = t.contiguous()
    if t.dtype in (torch.float16, torch.bfloat16):
        return t.view(torch.int16).numpy().view(np.uint16).copy()
    return t.numpy().copy()


def from_numpy(a: np.ndarray, dtype: torch.dtype) -> torch.Tensor:
    if dtype in (torch.float16, torch.bfloat16):
        return torch.from_numpy(a.view(np.int16).copy()).view(dtype)
    return torch.from_numpy(a.copy())


def digest(*tensors: torch.Tensor) -> str:
    h = hashlib.sha256()
    for t in tensors:
        h.update(to_numpy(t).tobytes())
    return h.hexdigest()
