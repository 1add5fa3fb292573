"""Helpers shared by the GPU parity tests, smoke() and the tuning sweep: seeded
problem generators (host tensors first, so the oracle and the GPU see the same
bits) and the comparison against the CPU oracle."""
from __future__ import annotations

import random
from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np
import torch

from oracle import c_oracle

# north-star tolerances (BASELINE.json): fp16 1e-3, bf16 1e-2; fp32 held to 1e-4
TOL = {torch.float16: 1e-3, torch.bfloat16: 1e-2, torch.float32: 1e-4}


def pack_factor(dtype: torch.dtype) -> int:
    return 16 // torch.empty((), dtype=dtype).element_size()


@dataclass
class DecodeProblem:
    query: torch.Tensor          # [S, H, D] view into a fused [S, 3, H, D] tensor (row stride 3*H*D)
    key_cache: torch.Tensor
    value_cache: torch.Tensor
    block_tables: torch.Tensor   # int32 [S, MB]
    context_lens: torch.Tensor   # int32 [S]
    scale: float
    block_size: int
    max_context_len: int

    def to(self, device) -> "DecodeProblem":
        qkv = self.query._base if self.query._base is not None else self.query
        if self.query._base is not None:
            q = qkv.to(device)[:, 0]
        else:
            q = self.query.to(device)
        return DecodeProblem(q, self.key_cache.to(device), self.value_cache.to(device),
                             self.block_tables.to(device), self.context_lens.to(device), self.scale,
                             self.block_size, self.max_context_len)


def make_decode_problem(num_seqs: int, num_heads: int, head_size: int, block_size: int, num_blocks: int,
                        dtype: torch.dtype, context_lens: Sequence[int], *, dist: str = "unit", seed: int = 0,
                        unique_blocks: bool = False, extra_table_cols: int = 0) -> DecodeProblem:
    """Host-side decode problem.  dist: 'ref' = U(-1e-3, 1e-3) as the reference tests,
    'unit' = N(0, 1).  Block tables draw blocks with repetition unless unique_blocks."""
    g = torch.Generator().manual_seed(seed)
    rnd = random.Random(seed)
    x = pack_factor(dtype)

    def fill(shape):
        if dist == "ref":
            return (torch.rand(shape, generator=g) * 2e-3 - 1e-3).to(dtype)
        return torch.randn(shape, generator=g).to(dtype)

    qkv = fill((num_seqs, 3, num_heads, head_size))
    key_cache = fill((num_blocks, num_heads, head_size // x, block_size, x))
    value_cache = fill((num_blocks, num_heads, head_size, block_size))
    ctx = [int(c) for c in context_lens]
    max_ctx = max(ctx) if ctx else 0
    max_blocks = (max_ctx + block_size - 1) // block_size + extra_table_cols
    tables = np.zeros((num_seqs, max(max_blocks, 1)), dtype=np.int32)
    free = list(range(num_blocks))
    rnd.shuffle(free)
    for i, c in enumerate(ctx):
        n = (c + block_size - 1) // block_size
        if unique_blocks:
            assert len(free) >= n, "not enough blocks for unique tables"
            tables[i, :n] = [free.pop() for _ in range(n)]
        else:
            tables[i, :n] = [rnd.randint(0, num_blocks - 1) for _ in range(n)]
    return DecodeProblem(qkv[:, 0], key_cache, value_cache, torch.from_numpy(tables),
                         torch.tensor(ctx, dtype=torch.int32), float(1.0 / (head_size ** 0.5)), block_size, max_ctx)


def oracle_decode(p: DecodeProblem, rows: Optional[Sequence[int]] = None) -> torch.Tensor:
    """C-oracle output (in the problem's dtype) for all sequences or the given rows."""
    assert p.query.device.type == "cpu"
    q, bt, cl = p.query, p.block_tables, p.context_lens
    if rows is not None:
        idx = torch.tensor(list(rows), dtype=torch.long)
        q, bt, cl = q[idx].contiguous(), bt[idx].contiguous(), cl[idx].contiguous()
    out = torch.empty(q.shape, dtype=q.dtype)
    c_oracle.decode_attention(out, q if q.stride(1) == q.shape[2] else q.contiguous(), p.key_cache, p.value_cache,
                              p.scale, bt, cl)
    return out


def assert_decode_close(got: torch.Tensor, want: torch.Tensor, dtype: torch.dtype, what: str = "") -> None:
    """|got - want| <= tol * max|want| + tol * |want| with the north-star tolerance of the dtype."""
    tol = TOL[dtype]
    g, w = got.detach().float().cpu(), want.detach().float().cpu()
    assert torch.isfinite(g).all(), f"{what}: non-finite output"
    scale = float(w.abs().max()) if w.numel() else 0.0
    sub = 1.2e-7 if dtype == torch.float16 else 0.0     # two fp16 subnormal steps: U(+-1e-3) inputs over long contexts
    torch.testing.assert_close(g, w, atol=tol * max(scale, 1e-30) + sub, rtol=tol, msg=lambda m: f"{what}: {m}")


def assert_flat_north_star(got: torch.Tensor, want: torch.Tensor, dtype: torch.dtype, what: str = "") -> None:
    """The north-star tolerance by the letter: allclose(got, want, atol = rtol = 1e-3 fp16 / 1e-2 bf16), NOT
    scaled by the output magnitude (fp32 is held to the reference test's own atol = 1e-3 with rtol = 1e-4).
    Used beside assert_decode_close, which is the stricter one whenever max|want| < 1."""
    tol = {torch.float16: 1e-3, torch.bfloat16: 1e-2, torch.float32: 1e-3}[dtype]
    rtol = 1e-4 if dtype == torch.float32 else tol
    g, w = got.detach().float().cpu(), want.detach().float().cpu()
    torch.testing.assert_close(g, w, atol=tol, rtol=rtol, msg=lambda m: f"{what} (flat north-star tolerance): {m}")


def run_decode(p_dev: DecodeProblem, *, tuning=None) -> torch.Tensor:
    """Calls the operator exactly as vllm/model_executor/layers/attention.py:81-91 does."""
    from vllm_b200 import attention_ops
    out = torch.empty(p_dev.query.shape, dtype=p_dev.query.dtype, device=p_dev.query.device)
    attention_ops.single_query_cached_kv_attention(out, p_dev.query, p_dev.key_cache, p_dev.value_cache, p_dev.scale,
                                                   p_dev.block_tables, p_dev.context_lens, p_dev.block_size,
                                                   p_dev.max_context_len, tuning=tuning)
    return out


def bits(t: torch.Tensor) -> torch.Tensor:
    """Integer view for bit-exact comparison."""
    t = t.detach().cpu().contiguous()
    return t.view(torch.int16) if t.element_size() == 2 else t.view(torch.int32)


def tuning(variant: int = 0, warps: int = 0, stages: int = 0, splits: int = 0):
    from vllm_b200 import _cabi
    return _cabi.DecodeTuning(variant, warps, stages, splits)
