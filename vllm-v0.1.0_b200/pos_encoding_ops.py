"""Drop-in for the reference's `vllm.pos_encoding_ops` extension module, plus the fusion of the
rotary embedding with the KV-cache write that follows it on the decode path.

Reference interface: csrc/pos_encoding.cpp:3-15 (pybind11), csrc/pos_encoding_kernels.cu:51-83;
call site vllm/model_executor/layers/attention.py:200-206 (PagedAttentionWithRoPE.forward), which is
followed by cache_ops.reshape_and_cache at :134-140.
"""
from __future__ import annotations

import torch

from . import _cabi
from .attention_ops import _dtype_code, _require_cuda


def _rope_args(op: str, positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor, head_size: int,
               cos_sin_cache: torch.Tensor):
    dev = query.device
    for name, t in (("positions", positions), ("query", query), ("key", key), ("cos_sin_cache", cos_sin_cache)):
        _require_cuda(name, t, dev)
    code = _dtype_code(query)
    if key.dtype != query.dtype or cos_sin_cache.dtype != query.dtype:
        raise RuntimeError(f"{op}: query, key and cos_sin_cache must share one dtype")
    if positions.dtype != torch.int64:            # the reference's data_ptr<int64_t>() throws otherwise
        raise RuntimeError(f"{op}: positions must be int64")
    if query.dim() != 2 or key.dim() != 2 or cos_sin_cache.dim() != 2 or query.shape != key.shape:
        raise RuntimeError(f"{op}: query and key must be [num_tokens, num_heads * head_size]")
    if query.stride(1) != 1 or key.stride(1) != 1 or not cos_sin_cache.is_contiguous() or not positions.is_contiguous():
        raise RuntimeError(f"{op}: rows must be contiguous")
    if head_size <= 0 or query.shape[1] % head_size:
        raise RuntimeError(f"{op}: hidden size {query.shape[1]} is not a multiple of head_size {head_size}")
    if positions.numel() < query.shape[0]:
        raise RuntimeError(f"{op}: positions is shorter than the token count")
    return dev, code, query.shape[0], query.shape[1] // head_size, cos_sin_cache.shape[1]


def rotary_embedding_neox(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor, head_size: int,
                          cos_sin_cache: torch.Tensor) -> None:
    """GPT-NeoX rotary embedding in place on query and key ([num_tokens, num_heads * head_size])."""
    ops = _cabi.torch_ops()
    if ops is not None:
        return ops.rotary_embedding_neox(positions, query, key, head_size, cos_sin_cache)
    dev, code, n, heads, rot = _rope_args("rotary_embedding_neox", positions, query, key, head_size, cos_sin_cache)
    with torch.cuda.device(dev):
        rc = _cabi.lib().pa_rotary_embedding_neox(positions.data_ptr(), query.data_ptr(), key.data_ptr(),
                                                  cos_sin_cache.data_ptr(), n, heads, head_size, rot, query.stride(0),
                                                  key.stride(0), code, torch.cuda.current_stream(dev).cuda_stream)
    _cabi.check(rc)


def rotary_embedding_neox_reshape_and_cache(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor,
                                            value: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                                            slot_mapping: torch.Tensor, head_size: int,
                                            cos_sin_cache: torch.Tensor) -> None:
    """One launch for rotary_embedding_neox(positions, query, key, head_size, cos_sin_cache) followed by
    cache_ops.reshape_and_cache(key[:n], value[:n], key_cache, value_cache, slot_mapping) with
    n = slot_mapping.numel() (the model pads its token batch; the slot mapping is not padded,
    vllm/model_executor/layers/attention.py:129-140).  query and key end up rotated in place exactly
    as the two separate operators leave them."""
    ops = _cabi.torch_ops()
    if ops is not None:
        return ops.rotary_embedding_neox_reshape_and_cache(positions, query, key, value, key_cache, value_cache,
                                                           slot_mapping, head_size, cos_sin_cache)
    op = "rotary_embedding_neox_reshape_and_cache"
    dev, code, n, heads, rot = _rope_args(op, positions, query, key, head_size, cos_sin_cache)
    for name, t in (("value", value), ("key_cache", key_cache), ("value_cache", value_cache),
                    ("slot_mapping", slot_mapping)):
        _require_cuda(name, t, dev)
    if value.dtype != query.dtype or key_cache.dtype != query.dtype or value_cache.dtype != query.dtype:
        raise RuntimeError(f"{op}: value and both caches must have the dtype of query")
    if slot_mapping.dtype != torch.int32 or not slot_mapping.is_contiguous():
        raise RuntimeError(f"{op}: slot_mapping must be a contiguous int32 tensor")
    if value.dim() != 2 or value.shape != key.shape or value.stride(1) != 1:
        raise RuntimeError(f"{op}: value must be [num_tokens, num_heads * head_size] with contiguous rows")
    if key_cache.dim() != 5 or value_cache.dim() != 4 or not key_cache.is_contiguous() or not value_cache.is_contiguous():
        raise RuntimeError(f"{op}: unexpected cache rank or layout")
    block_size, x = key_cache.shape[3], key_cache.shape[4]
    if x != 16 // key.element_size() or key_cache.shape[1] != heads or key_cache.shape[2] * x != head_size \
            or tuple(value_cache.shape[1:]) != (heads, head_size, block_size):
        raise RuntimeError(f"{op}: cache shapes do not match [blocks, heads, head_size/x, block_size, x] / "
                           "[blocks, heads, head_size, block_size]")
    n_cache = slot_mapping.numel()
    if n_cache > n:
        raise RuntimeError(f"{op}: slot_mapping is longer than the token batch")
    with torch.cuda.device(dev):
        rc = _cabi.lib().pa_rope_reshape_and_cache(
            positions.data_ptr(), query.data_ptr(), key.data_ptr(), value.data_ptr(), key_cache.data_ptr(),
            value_cache.data_ptr(), slot_mapping.data_ptr(), cos_sin_cache.data_ptr(), n, n_cache, heads, head_size,
            block_size, rot, query.stride(0), key.stride(0), value.stride(0), code,
            torch.cuda.current_stream(dev).cuda_stream)
    _cabi.check(rc)
