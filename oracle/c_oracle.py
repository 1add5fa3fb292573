"""ctypes access to libpaged_oracle.so (oracle/paged_oracle.c) on CPU torch tensors.

TEST INFRASTRUCTURE ONLY -- see paged_oracle.c for what each function restates.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Dict, List, Sequence

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpaged_oracle.so")
_DT = {torch.float16: 0, torch.bfloat16: 1, torch.float32: 2}
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "paged_oracle.c")
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE, "-B", "libpaged_oracle.so"], check=True, capture_output=True)
    return LIB_PATH


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        _lib = C.CDLL(LIB_PATH)
        _lib.po_num_threads.restype = C.c_int
        _lib.po_decode_attention.restype = C.c_int
        vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
        _lib.po_decode_attention.argtypes = [vp, vp, vp, vp, C.c_float, vp, vp, i32, i32, i32, i32, i32, i64, i32]
        for name in ("po_reshape_and_cache", "po_gather_cached_kv"):
            fn = getattr(_lib, name)
            fn.restype = None
            fn.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, i64, i64, i32]
        _lib.po_copy_blocks.restype = None
        _lib.po_copy_blocks.argtypes = [C.POINTER(vp), C.POINTER(vp), i32, C.POINTER(i64), i64, i64]
        _lib.po_rotary_embedding_neox.restype = None
        _lib.po_rotary_embedding_neox.argtypes = [vp, vp, vp, vp, i32, i32, i32, i32, i64, i64, i32]
        _lib.po_swap_blocks.restype = None
        _lib.po_swap_blocks.argtypes = [vp, vp, i64, C.POINTER(i64), C.POINTER(i64), i64]
    return _lib


def num_threads() -> int:
    return int(lib().po_num_threads())


def _cpu(*tensors: torch.Tensor) -> None:
    for t in tensors:
        assert t.device.type == "cpu", "the C oracle works on host tensors"


def decode_attention(out: torch.Tensor, query: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                     scale: float, block_tables: torch.Tensor, context_lens: torch.Tensor) -> None:
    _cpu(out, query, key_cache, value_cache, block_tables, context_lens)
    assert out.is_contiguous() and key_cache.is_contiguous() and value_cache.is_contiguous()
    assert block_tables.dtype == torch.int32 and context_lens.dtype == torch.int32
    assert query.stride(2) == 1 and query.stride(1) == query.shape[2]
    s, h, d = query.shape
    rc = lib().po_decode_attention(out.data_ptr(), query.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
                                   float(scale), block_tables.data_ptr(), context_lens.data_ptr(), s, h, d,
                                   value_cache.shape[3], block_tables.shape[1], query.stride(0), _DT[query.dtype])
    if rc != 0:
        raise RuntimeError(f"po_decode_attention failed: {rc}")


def reshape_and_cache(key, value, key_cache, value_cache, slot_mapping) -> None:
    _cpu(key, value, key_cache, value_cache, slot_mapping)
    n, h, d = key.shape
    lib().po_reshape_and_cache(key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
                               slot_mapping.data_ptr(), n, h, d, key_cache.shape[3], key.stride(0), value.stride(0),
                               key.element_size())


def gather_cached_kv(key, value, key_cache, value_cache, slot_mapping) -> None:
    _cpu(key, value, key_cache, value_cache, slot_mapping)
    n, h, d = key.shape
    lib().po_gather_cached_kv(key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
                              slot_mapping.data_ptr(), n, h, d, key_cache.shape[3], key.stride(0), value.stride(0),
                              key.element_size())


def copy_blocks(key_caches: Sequence[torch.Tensor], value_caches: Sequence[torch.Tensor],
                block_mapping: Dict[int, List[int]]) -> None:
    _cpu(*key_caches, *value_caches)
    flat: List[int] = []
    for src in sorted(block_mapping):
        for dst in block_mapping[src]:
            flat += (src, dst)
    n = len(key_caches)
    kp = (C.c_void_p * n)(*[t.data_ptr() for t in key_caches])
    vp = (C.c_void_p * n)(*[t.data_ptr() for t in value_caches])
    pairs = (C.c_int64 * max(len(flat), 1))(*flat)
    block_bytes = key_caches[0][0].numel() * key_caches[0].element_size()
    lib().po_copy_blocks(kp, vp, n, pairs, len(flat) // 2, block_bytes)


def swap_blocks(src: torch.Tensor, dst: torch.Tensor, block_mapping: Dict[int, int]) -> None:
    _cpu(src, dst)
    keys = sorted(block_mapping)
    n = len(keys)
    s = (C.c_int64 * max(n, 1))(*keys)
    d = (C.c_int64 * max(n, 1))(*[block_mapping[k] for k in keys])
    lib().po_swap_blocks(src.data_ptr(), dst.data_ptr(), src.element_size() * src[0].numel(), s, d, n)


def rotary_embedding_neox(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor, head_size: int,
                          cos_sin_cache: torch.Tensor) -> None:
    _cpu(positions, query, key, cos_sin_cache)
    assert positions.dtype == torch.int64 and positions.is_contiguous() and cos_sin_cache.is_contiguous()
    n = query.shape[0]
    heads = query.shape[1] // head_size if query.dim() == 2 else query.shape[1]
    lib().po_rotary_embedding_neox(positions.data_ptr(), query.data_ptr(), key.data_ptr(), cos_sin_cache.data_ptr(), n,
                                   heads, head_size, cos_sin_cache.shape[1], query.stride(0), key.stride(0),
                                   _DT[query.dtype])
