"""Drop-in for the reference's `vllm.cache_ops` extension module.

Reference interface: csrc/cache.cpp:6-47 (pybind11), csrc/cache_kernels.cu;
call sites vllm/model_executor/layers/attention.py:134-140 and
vllm/worker/cache_engine.py:112-116,130.  Same names, positional arguments
and in-place semantics; errors surface as RuntimeError.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional

import torch

from . import _cabi
from .attention_ops import _dtype_code, _require_cuda


def _kv_rows(op: str, key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
             slot_mapping: torch.Tensor):
    dev = key_cache.device
    for name, t in (("key", key), ("value", value), ("key_cache", key_cache), ("value_cache", value_cache),
                    ("slot_mapping", slot_mapping)):
        _require_cuda(name, t, dev)
    code = _dtype_code(key)
    if value.dtype != key.dtype or key_cache.dtype != key.dtype or value_cache.dtype != key.dtype:
        raise RuntimeError(f"{op}: key, value and both caches must share one dtype")
    if slot_mapping.dtype != torch.int32:
        raise RuntimeError(f"{op}: slot_mapping must be int32")
    if key.dim() != 3 or value.shape != key.shape or key_cache.dim() != 5 or value_cache.dim() != 4:
        raise RuntimeError(f"{op}: unexpected tensor rank or shape")
    num_tokens, num_heads, head_size = key.shape
    for name, t in (("key", key), ("value", value)):
        if t.stride(2) != 1 or t.stride(1) != head_size:
            raise RuntimeError(f"{op}: {name} must be contiguous in its last two dimensions")
    if not key_cache.is_contiguous() or not value_cache.is_contiguous() or not slot_mapping.is_contiguous():
        raise RuntimeError(f"{op}: caches and slot_mapping must be contiguous")
    block_size, x = key_cache.shape[3], key_cache.shape[4]   # as the reference: cache_kernels.cu:196-197
    if x != 16 // key.element_size() or key_cache.shape[1] != num_heads or key_cache.shape[2] * x != head_size:
        raise RuntimeError(f"{op}: key_cache shape does not match [blocks, heads, head_size/x, block_size, x]")
    if tuple(value_cache.shape[1:]) != (num_heads, head_size, block_size):
        raise RuntimeError(f"{op}: value_cache shape does not match [blocks, heads, head_size, block_size]")
    if slot_mapping.numel() < num_tokens:
        raise RuntimeError(f"{op}: slot_mapping is shorter than the token count")
    return dev, code, num_tokens, num_heads, head_size, block_size


_SLOT_PATTERNS = {None: _cabi.PA_SLOTS_AUTO, "scattered": _cabi.PA_SLOTS_SCATTERED, "runs": _cabi.PA_SLOTS_RUNS}


def reshape_and_cache(key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                      slot_mapping: torch.Tensor, *, slots: Optional[str] = None) -> None:
    """Scatter the new tokens' key/value rows into the paged caches at slot_mapping.

    `slots` (keyword-only, not part of the reference signature) tells the library what the caller knows about
    slot_mapping: "scattered" = a decode step, one token per sequence, each in a block of its own; "runs" =
    prompt tokens in consecutive slots (or a mixed batch); None = unknown.  A hint only: the result is the same."""
    try:
        pattern = _SLOT_PATTERNS[slots]
    except KeyError:
        raise RuntimeError(f"reshape_and_cache: slots must be None, 'scattered' or 'runs', not {slots!r}") from None
    ops = _cabi.torch_ops()
    if ops is not None:
        return ops.reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, pattern)
    dev, code, n, h, d, b = _kv_rows("reshape_and_cache", key, value, key_cache, value_cache, slot_mapping)
    with torch.cuda.device(dev):
        rc = _cabi.lib().pa_reshape_and_cache_ex(
            key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(), slot_mapping.data_ptr(),
            n, h, d, b, key.stride(0), value.stride(0), code, pattern, torch.cuda.current_stream(dev).cuda_stream)
    _cabi.check(rc)


def gather_cached_kv(key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                     slot_mapping: torch.Tensor) -> None:
    """Inverse of reshape_and_cache: key/value rows <- caches at slot_mapping."""
    ops = _cabi.torch_ops()
    if ops is not None:
        return ops.gather_cached_kv(key, value, key_cache, value_cache, slot_mapping)
    dev, code, n, h, d, b = _kv_rows("gather_cached_kv", key, value, key_cache, value_cache, slot_mapping)
    with torch.cuda.device(dev):
        rc = _cabi.lib().pa_gather_cached_kv(
            key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(), slot_mapping.data_ptr(),
            n, h, d, b, key.stride(0), value.stride(0), code, torch.cuda.current_stream(dev).cuda_stream)
    _cabi.check(rc)


def copy_blocks(key_caches: List[torch.Tensor], value_caches: List[torch.Tensor],
                block_mapping: Dict[int, List[int]]) -> None:
    """Copy-on-write: block src -> each dst of block_mapping[src], in every layer's key and value cache."""
    ops = _cabi.torch_ops()
    if ops is not None:
        return ops.copy_blocks(key_caches, value_caches, block_mapping)
    num_layers = len(key_caches)
    if num_layers != len(value_caches):
        raise RuntimeError("copy_blocks: key_caches and value_caches differ in length")
    if num_layers == 0:
        return
    dev = key_caches[0].device
    if not key_caches[0].is_cuda:
        raise RuntimeError("copy_blocks: caches must be CUDA tensors")
    block_bytes = key_caches[0][0].numel() * key_caches[0].element_size()
    for kc, vc in zip(key_caches, value_caches):
        if kc.device != dev or vc.device != dev or not kc.is_contiguous() or not vc.is_contiguous():
            raise RuntimeError("copy_blocks: every cache must be contiguous and on one GPU")
        if kc[0].numel() * kc.element_size() != block_bytes or vc[0].numel() * vc.element_size() != block_bytes:
            raise RuntimeError("copy_blocks: caches differ in block size")
    flat: List[int] = []
    for src in sorted(block_mapping):            # std::map order of the reference (cache_kernels.cu:105)
        for dst in block_mapping[src]:
            flat += (int(src), int(dst))
    num_pairs = len(flat) // 2
    if num_pairs == 0:
        return
    kptrs = (C.c_void_p * num_layers)(*[t.data_ptr() for t in key_caches])
    vptrs = (C.c_void_p * num_layers)(*[t.data_ptr() for t in value_caches])
    pairs = (C.c_int64 * len(flat))(*flat)
    with torch.cuda.device(dev):
        rc = _cabi.lib().pa_copy_blocks(kptrs, vptrs, num_layers, pairs, num_pairs, block_bytes,
                                        torch.cuda.current_stream(dev).cuda_stream)
    _cabi.check(rc)


def swap_blocks(src: torch.Tensor, dst: torch.Tensor, block_mapping: Dict[int, int], *, engine: int = 0) -> None:
    """dst[block_mapping[s]] = src[s] for whole cache blocks; GPU<->GPU (same device), GPU->CPU or CPU->GPU.

    `engine` (keyword-only, not part of the reference signature): 0 automatic, 1 copy engines, 2 kernel."""
    ops = _cabi.torch_ops()
    if ops is not None:
        return ops.swap_blocks(src, dst, block_mapping, engine)
    if src.is_cuda and dst.is_cuda:
        if src.device != dst.device:
            raise RuntimeError("src and dst must be on the same GPU")
        kind, dev = _cabi.PA_COPY_D2D, src.device
    elif src.is_cuda and not dst.is_cuda:
        kind, dev = _cabi.PA_COPY_D2H, src.device
    elif not src.is_cuda and dst.is_cuda:
        kind, dev = _cabi.PA_COPY_H2D, dst.device
    else:
        raise RuntimeError("Invalid device combination")
    if not src.is_contiguous() or not dst.is_contiguous():
        raise RuntimeError("swap_blocks: caches must be contiguous")
    block_bytes = src.element_size() * src[0].numel()           # cache_kernels.cu:32
    if dst.element_size() * dst[0].numel() != block_bytes:
        raise RuntimeError("swap_blocks: src and dst differ in block size")
    n = len(block_mapping)
    if n == 0:
        return
    keys = sorted(block_mapping)                                  # std::map order
    srcs = (C.c_int64 * n)(*[int(k) for k in keys])
    dsts = (C.c_int64 * n)(*[int(block_mapping[k]) for k in keys])
    with torch.cuda.device(dev):
        rc = _cabi.lib().pa_swap_blocks_ex(src.data_ptr(), dst.data_ptr(), block_bytes, srcs, dsts, n, kind, engine,
                                           torch.cuda.current_stream(dev).cuda_stream)
    _cabi.check(rc)
