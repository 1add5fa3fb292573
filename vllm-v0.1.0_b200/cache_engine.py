"""Host-side mirror of the reference's CacheEngine (vllm/worker/cache_engine.py:12-146).

Owns the paged KV layout of one GPU worker: per-layer key/value caches on the
GPU, pinned per-layer caches on the host, a side stream for swaps with one
event per layer, and copy-on-write through cache_ops.copy_blocks.  Method names
and semantics follow the reference; the constructor takes plain sizes instead of
the reference's config objects (configuration is outside this path).
"""
from __future__ import annotations

from typing import Dict, List, Tuple

import torch

from . import cache_ops

KVCache = Tuple[torch.Tensor, torch.Tensor]


def kv_cache_shapes(num_heads: int, head_size: int, block_size: int, dtype: torch.dtype):
    """(key block shape, value block shape) -- reference cache_engine.py:48-63.

    key   block: (num_heads, head_size // x, block_size, x), x = 16 // element size
    value block: (num_heads, head_size, block_size)
    """
    x = 16 // torch.empty((), dtype=dtype).element_size()
    if head_size % x:
        raise ValueError(f"head_size {head_size} is not a multiple of x={x}")
    return (num_heads, head_size // x, block_size, x), (num_heads, head_size, block_size)


class CacheEngine:
    """KV-cache owner of one worker: allocation, swap in/out, copy-on-write."""

    def __init__(self, num_layers: int, num_heads: int, head_size: int, block_size: int, dtype: torch.dtype,
                 num_gpu_blocks: int, num_cpu_blocks: int, device: str = "cuda") -> None:
        self.num_layers = num_layers
        self.num_heads = num_heads          # heads of THIS worker (total heads / tensor-parallel size)
        self.head_size = head_size
        self.block_size = block_size
        self.dtype = dtype
        self.num_gpu_blocks = num_gpu_blocks
        self.num_cpu_blocks = num_cpu_blocks
        self.device = torch.device(device)

        self.gpu_cache = self.allocate_gpu_cache()
        self.cpu_cache = self.allocate_cpu_cache()

        # swaps run beside the model on their own stream; attention waits on events[layer]
        self.cache_stream = torch.cuda.Stream(device=self.device)
        self.events = [torch.cuda.Event() for _ in range(num_layers)]

    def get_key_block_shape(self) -> Tuple[int, int, int, int]:
        return kv_cache_shapes(self.num_heads, self.head_size, self.block_size, self.dtype)[0]

    def get_value_block_shape(self) -> Tuple[int, int, int]:
        return kv_cache_shapes(self.num_heads, self.head_size, self.block_size, self.dtype)[1]

    def _allocate(self, num_blocks: int, **kw) -> List[KVCache]:
        kshape, vshape = self.get_key_block_shape(), self.get_value_block_shape()
        return [(torch.empty((num_blocks, *kshape), dtype=self.dtype, **kw),
                 torch.empty((num_blocks, *vshape), dtype=self.dtype, **kw)) for _ in range(self.num_layers)]

    def allocate_gpu_cache(self) -> List[KVCache]:
        return self._allocate(self.num_gpu_blocks, device=self.device)

    def allocate_cpu_cache(self) -> List[KVCache]:
        return self._allocate(self.num_cpu_blocks, pin_memory=True)

    def _swap(self, src: List[KVCache], dst: List[KVCache], src_to_dst: Dict[int, int]) -> None:
        with torch.cuda.stream(self.cache_stream):
            for layer in range(self.num_layers):
                cache_ops.swap_blocks(src[layer][0], dst[layer][0], src_to_dst)
                cache_ops.swap_blocks(src[layer][1], dst[layer][1], src_to_dst)
                self.events[layer].record(stream=self.cache_stream)

    def swap_in(self, src_to_dst: Dict[int, int]) -> None:
        self._swap(self.cpu_cache, self.gpu_cache, src_to_dst)

    def swap_out(self, src_to_dst: Dict[int, int]) -> None:
        self._swap(self.gpu_cache, self.cpu_cache, src_to_dst)

    def copy(self, src_to_dsts: Dict[int, List[int]]) -> None:
        cache_ops.copy_blocks([k for k, _ in self.gpu_cache], [v for _, v in self.gpu_cache], src_to_dsts)

    @staticmethod
    def get_cache_block_size(block_size: int, num_heads: int, head_size: int, num_layers: int,
                             dtype: torch.dtype) -> int:
        """Bytes of one block across all layers, key + value (reference :132-146)."""
        per = block_size * num_heads * head_size
        return torch.empty((), dtype=dtype).element_size() * num_layers * 2 * per
