"""Drop-in for the reference's `vllm.attention_ops` extension module.

Reference interface: csrc/attention.cpp:3-19 (pybind11) and
csrc/attention/attention_kernels.cu:450-469 (host entry); call site
vllm/model_executor/layers/attention.py:81-91.  Same name, positional
arguments and in-place semantics; errors surface as RuntimeError.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _cabi

_DTYPES = {torch.float16: _cabi.PA_F16, torch.bfloat16: _cabi.PA_BF16, torch.float32: _cabi.PA_F32}


def _dtype_code(t: torch.Tensor) -> int:
    try:
        return _DTYPES[t.dtype]
    except KeyError:
        raise RuntimeError(f"Unsupported data type: {t.dtype}") from None


def _require_cuda(name: str, t: torch.Tensor, device: torch.device) -> None:
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (there is no CPU path)")
    if t.device != device:
        raise RuntimeError(f"{name} is on {t.device}, expected {device}")


def single_query_cached_kv_attention(
    out: torch.Tensor,            # [num_seqs, num_heads, head_size]
    query: torch.Tensor,          # [num_seqs, num_heads, head_size], row-strided
    key_cache: torch.Tensor,      # [num_blocks, num_heads, head_size/x, block_size, x]
    value_cache: torch.Tensor,    # [num_blocks, num_heads, head_size, block_size]
    scale: float,
    block_tables: torch.Tensor,   # [num_seqs, max_num_blocks_per_seq] int32
    context_lens: torch.Tensor,   # [num_seqs] int32
    block_size: int,
    max_context_len: int,
    *,
    tuning: Optional[_cabi.DecodeTuning] = None,
) -> None:
    """out[i, h] = softmax(scale * q[i, h] . K^T) . V over the first context_lens[i]
    cached tokens of sequence i, reached through block_tables[i].

    `tuning` (keyword-only, not part of the reference signature) forces a launch
    shape; tests and the tuning sweep use it.
    """
    ops = _cabi.torch_ops()
    if ops is not None:
        if tuning is None:
            ops.single_query_cached_kv_attention(out, query, key_cache, value_cache, scale, block_tables, context_lens,
                                                 block_size, max_context_len)
        else:
            ops.single_query_cached_kv_attention(out, query, key_cache, value_cache, scale, block_tables, context_lens,
                                                 block_size, max_context_len, tuning.variant, tuning.num_warps,
                                                 tuning.num_stages, tuning.num_splits)
        return
    dev = query.device
    for name, t in (("out", out), ("query", query), ("key_cache", key_cache), ("value_cache", value_cache),
                    ("block_tables", block_tables), ("context_lens", context_lens)):
        _require_cuda(name, t, dev)
    code = _dtype_code(query)
    if out.dtype != query.dtype or key_cache.dtype != query.dtype or value_cache.dtype != query.dtype:
        raise RuntimeError("out, query, key_cache and value_cache must share one dtype")
    if block_tables.dtype != torch.int32 or context_lens.dtype != torch.int32:
        # the reference's data_ptr<int>() throws on anything else (attention_kernels.cu:358-359)
        raise RuntimeError("block_tables and context_lens must be int32")
    if query.dim() != 3 or out.dim() != 3 or key_cache.dim() != 5 or value_cache.dim() != 4 or block_tables.dim() != 2:
        raise RuntimeError("unexpected tensor rank")
    num_seqs, num_heads, head_size = query.shape
    if tuple(out.shape) != (num_seqs, num_heads, head_size) or not out.is_contiguous():
        raise RuntimeError("out must be a contiguous [num_seqs, num_heads, head_size] tensor")
    if query.stride(2) != 1 or query.stride(1) != head_size:
        raise RuntimeError("query must be contiguous in its last two dimensions")
    if not key_cache.is_contiguous() or not value_cache.is_contiguous() or not block_tables.is_contiguous() \
            or not context_lens.is_contiguous():
        raise RuntimeError("caches, block_tables and context_lens must be contiguous")
    if value_cache.shape[1] != num_heads or value_cache.shape[2] != head_size or value_cache.shape[3] != block_size:
        raise RuntimeError("value_cache shape does not match [num_blocks, num_heads, head_size, block_size]")
    if block_tables.shape[0] != num_seqs or context_lens.numel() != num_seqs:
        raise RuntimeError("block_tables / context_lens row count does not match query")

    lib = _cabi.lib()
    stream = torch.cuda.current_stream(dev).cuda_stream
    with torch.cuda.device(dev):
        rc = lib.pa_decode_attention_ex(
            out.data_ptr(), query.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(), float(scale),
            block_tables.data_ptr(), context_lens.data_ptr(), num_seqs, num_heads, head_size, int(block_size),
            block_tables.shape[1], int(max_context_len), query.stride(0), code,
            C.byref(tuning) if tuning is not None else None, stream)
    _cabi.check(rc)


def decode_plan(num_seqs: int, num_heads: int, head_size: int, block_size: int, max_context_len: int,
                dtype: torch.dtype) -> _cabi.DecodeTuning:
    """The launch shape the library picks for these sizes (needs no GPU)."""
    plan = _cabi.DecodeTuning(0, 0, 0, 0)
    _cabi.check(_cabi.lib().pa_decode_attention_plan(num_seqs, num_heads, head_size, block_size, max_context_len,
                                                     _DTYPES[dtype], C.byref(plan)))
    return plan


def multi_query_kv_attention(
    output: torch.Tensor,         # [num_prompt_tokens, num_heads, head_size]
    query: torch.Tensor,          # [num_prompt_tokens, num_heads, head_size], row-strided
    key: torch.Tensor,            # same
    value: torch.Tensor,          # same
    prompt_lens,                  # List[int]: the prompts packed along the token dimension
    scale: float,
    *,
    debug: Optional[torch.Tensor] = None,
) -> None:
    """Causal self-attention of every prompt over its own tokens, written into `output`.

    Replaces PagedAttention.multi_query_kv_attention (vllm/model_executor/layers/attention.py:50-70:
    xformers memory_efficient_attention_forward under BlockDiagonalCausalMask.from_seqlens(prompt_lens));
    the mask is passed as the list of prompt lengths it was built from.  fp16 / bf16 run on the
    tcgen05 tensor cores, fp32 on an exact SIMT path.  `debug` (tests): float32 device buffer the
    first CTA dumps its raw tiles into.
    """
    dev = query.device
    for name, t in (("output", output), ("query", query), ("key", key), ("value", value)):
        _require_cuda(name, t, dev)
    code = _dtype_code(query)
    if not (output.dtype == query.dtype == key.dtype == value.dtype):
        raise RuntimeError("output, query, key and value must share one dtype")
    if query.dim() != 3 or key.shape != query.shape or value.shape != query.shape or output.shape != query.shape:
        raise RuntimeError("output, query, key and value must be [num_tokens, num_heads, head_size] of one shape")
    num_tokens, num_heads, head_size = query.shape
    for name, t in (("output", output), ("query", query), ("key", key), ("value", value)):
        if t.stride(2) != 1 or t.stride(1) != head_size:
            raise RuntimeError(f"{name} must be contiguous in its last two dimensions")
    lens = [int(n) for n in prompt_lens]
    if sum(lens) > num_tokens:
        raise RuntimeError("prompt_lens add up to more tokens than query holds")
    arr = (C.c_int32 * max(1, len(lens)))(*lens)
    lib = _cabi.lib()
    stream = torch.cuda.current_stream(dev).cuda_stream
    with torch.cuda.device(dev):
        if debug is None:
            rc = lib.pa_prefill_attention(output.data_ptr(), query.data_ptr(), key.data_ptr(), value.data_ptr(), arr,
                                          len(lens), num_heads, head_size, query.stride(0), key.stride(0),
                                          value.stride(0), output.stride(0), float(scale), code, stream)
        else:
            rc = lib.pa_prefill_attention_debug(output.data_ptr(), query.data_ptr(), key.data_ptr(), value.data_ptr(),
                                                arr, len(lens), num_heads, head_size, query.stride(0), key.stride(0),
                                                value.stride(0), output.stride(0), float(scale), code,
                                                debug.data_ptr(), stream)
    _cabi.check(rc)


_workspaces = {}   # (device index, stream) -> uint8 tensor, grown on demand


def _workspace(dev: torch.device, stream: int, nbytes: int) -> torch.Tensor:
    key = (dev.index, stream)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        # the old buffer may still be in use by kernels queued on this stream; torch's caching allocator
        # keeps a freed block stream-ordered, so dropping the reference is safe
        ws = torch.empty(max(nbytes, 1 << 20) + 256, dtype=torch.uint8, device=dev)
        _workspaces[key] = ws
    return ws


def multi_query_cached_kv_attention(
    cu_query_lens,                # List[int] (or CPU tensor) [num_seqs + 1]: query tokens of sequence i are [cu[i], cu[i+1])
    output: torch.Tensor,         # [num_tokens, num_heads, head_size]
    query: torch.Tensor,          # [num_tokens, num_heads, head_size], row-strided
    key_cache: torch.Tensor,      # [num_blocks, num_heads, head_size/x, block_size, x]
    value_cache: torch.Tensor,    # [num_blocks, num_heads, head_size, block_size]
    scale: float,
    block_tables: torch.Tensor,   # [num_seqs, max_num_blocks_per_seq] int32, CUDA
    context_lens,                 # [num_seqs]: List[int] or tensor (a CUDA tensor is read back: one sync)
    block_size: int,
) -> None:
    """Prompt chunks over the paged cache: the new tokens of sequence i (whose K/V are already written
    to the cache) attend to the context_lens[i] cached tokens, causally: row r sees cached tokens
    0 .. r + (context_lens[i] - query_len_i).

    Parity reference: ref_multi_query_cached_kv_attention (tests/kernels/test_attention.py:104-156), which
    the reference tree keeps without a kernel; argument order follows single_query_cached_kv_attention
    (csrc/attention.cpp:3-19) with the cumulative query lengths in front, as that reference takes them.
    """
    dev = query.device
    for name, t in (("output", output), ("query", query), ("key_cache", key_cache), ("value_cache", value_cache),
                    ("block_tables", block_tables)):
        _require_cuda(name, t, dev)
    code = _dtype_code(query)
    if not (output.dtype == query.dtype == key_cache.dtype == value_cache.dtype):
        raise RuntimeError("output, query, key_cache and value_cache must share one dtype")
    if block_tables.dtype != torch.int32 or block_tables.dim() != 2 or not block_tables.is_contiguous():
        raise RuntimeError("block_tables must be a contiguous int32 [num_seqs, max_num_blocks_per_seq] tensor")
    if query.dim() != 3 or output.shape != query.shape or key_cache.dim() != 5 or value_cache.dim() != 4:
        raise RuntimeError("unexpected tensor rank or shape")
    num_tokens, num_heads, head_size = query.shape
    for name, t in (("output", output), ("query", query)):
        if t.stride(2) != 1 or t.stride(1) != head_size:
            raise RuntimeError(f"{name} must be contiguous in its last two dimensions")
    if not key_cache.is_contiguous() or not value_cache.is_contiguous():
        raise RuntimeError("caches must be contiguous")
    if value_cache.shape[1] != num_heads or value_cache.shape[2] != head_size or value_cache.shape[3] != block_size:
        raise RuntimeError("value_cache shape does not match [num_blocks, num_heads, head_size, block_size]")
    cu = [int(v) for v in (cu_query_lens.tolist() if isinstance(cu_query_lens, torch.Tensor) else cu_query_lens)]
    ctx = [int(v) for v in (context_lens.tolist() if isinstance(context_lens, torch.Tensor) else context_lens)]
    num_seqs = len(cu) - 1
    if num_seqs < 0 or len(ctx) != num_seqs or block_tables.shape[0] != num_seqs:
        raise RuntimeError("cu_query_lens, context_lens and block_tables disagree on the number of sequences")
    if num_seqs == 0:
        return
    qlens = [cu[i + 1] - cu[i] for i in range(num_seqs)]
    if cu[0] != 0 or min(qlens) < 0 or cu[-1] > num_tokens:
        raise RuntimeError("cu_query_lens must start at 0, be non-decreasing and end within query")
    lib = _cabi.lib()
    need = lib.pa_multi_query_cached_kv_workspace_bytes(sum(ctx), num_heads, head_size, code)
    if need < 0:
        _cabi.check(int(need))
    stream = torch.cuda.current_stream(dev).cuda_stream
    if need:
        ws = _workspace(dev, stream, int(need))
        base = (ws.data_ptr() + 255) // 256 * 256
        ws_bytes = ws.numel() - (base - ws.data_ptr())
    else:                                           # 16-bit types: the kernel reads the pages itself
        base, ws_bytes = None, 0
    qarr = (C.c_int32 * num_seqs)(*qlens)
    carr = (C.c_int32 * num_seqs)(*ctx)
    with torch.cuda.device(dev):
        rc = lib.pa_multi_query_cached_kv_attention(
            output.data_ptr(), query.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(), float(scale),
            block_tables.data_ptr(), qarr, carr, num_seqs, num_heads, head_size, int(block_size),
            block_tables.shape[1], query.stride(0), output.stride(0), code, base, ws_bytes, stream)
    _cabi.check(rc)
