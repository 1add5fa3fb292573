"""Head sharding of the decode path across the GPUs of one node (tensor parallelism of the
reference: vllm/model_executor/models/llama.py:86-89, vllm/worker/cache_engine.py:31,
vllm/config.py:54-60,77-79).  Every (sequence, head) is independent, so GPU r of g owns heads
[r*H/g, (r+1)*H/g) of q, out and both caches and the path needs no collective."""
from __future__ import annotations

from typing import Tuple


def heads_per_gpu(total_heads: int, world_size: int) -> int:
    """total_heads / world_size; the reference rejects anything that does not divide (config.py:54-60)."""
    if world_size <= 0 or total_heads % world_size != 0:
        raise ValueError(f"Total number of attention heads ({total_heads}) must be divisible by "
                         f"tensor parallel size ({world_size}).")
    return total_heads // world_size


def head_range(rank: int, world_size: int, total_heads: int) -> Tuple[int, int]:
    """[first, last) head owned by `rank`."""
    n = heads_per_gpu(total_heads, world_size)
    if not 0 <= rank < world_size:
        raise ValueError(f"rank {rank} outside world of {world_size}")
    return rank * n, (rank + 1) * n


def shard_heads(t, rank: int, world_size: int, head_dim: int = 1):
    """The slice of a [.., heads, ..] tensor (q, out, key_cache, value_cache) that `rank` owns."""
    lo, hi = head_range(rank, world_size, t.shape[head_dim])
    return t.narrow(head_dim, lo, hi - lo)
