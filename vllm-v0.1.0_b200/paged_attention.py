"""Mirror of the reference's attention layer (vllm/model_executor/layers/attention.py:16-217):
`PagedAttention` and `PagedAttentionWithRoPE` with the same constructor arguments, method names
and `forward` signatures, every step of which runs on this library's sm_100a kernels:

  1. prompt tokens      multi_query_kv_attention           -> tcgen05 prefill kernel (no xformers)
  2. cache_event.wait() (swap / copy issued by the cache engine)
  3. KV write           cache_ops.reshape_and_cache        -> tile mover
  4. generation tokens  single_query_cached_kv_attention   -> paged decode kernel
  RoPE layer: rotary embedding fused with step 3 in one launch when `fuse_rope_cache` is on.

torch.nn.Module is used only as the container the reference uses; there is no PyTorch math here.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from . import attention_ops, cache_ops, pos_encoding_ops
from .input_metadata import BlockDiagonalCausalMask, InputMetadata

_SUPPORTED_HEAD_SIZES = [64, 80, 96, 128]


class PagedAttention(nn.Module):
    """GPT-style multi-head paged attention over a flat token batch laid out as
    |prompt_0 ... prompt_N-1|generation_0 ... generation_M-1|padding| (attention.py:16-38)."""

    def __init__(self, num_heads: int, head_size: int, scale: float) -> None:
        super().__init__()
        self.num_heads = num_heads
        self.head_size = head_size
        self.scale = float(scale)
        if self.head_size not in _SUPPORTED_HEAD_SIZES:
            raise ValueError(f"head_size ({self.head_size}) is not supported. "
                             f"Supported head sizes: {_SUPPORTED_HEAD_SIZES}.")

    def multi_query_kv_attention(
        self,
        output: torch.Tensor,                   # [num_prompt_tokens, num_heads, head_size]
        query: torch.Tensor,                    # [num_prompt_tokens, num_heads, head_size]
        key: torch.Tensor,                      # [num_prompt_tokens, num_heads, head_size]
        value: torch.Tensor,                    # [num_prompt_tokens, num_heads, head_size]
        attn_bias: BlockDiagonalCausalMask,
    ) -> torch.Tensor:
        # attention.py:50-70; the result lands in `output` directly (the reference copies it there)
        attention_ops.multi_query_kv_attention(output, query, key, value, attn_bias.seqlens, self.scale)
        return output

    def single_query_cached_kv_attention(
        self,
        output: torch.Tensor,           # [num_generation_tokens, num_heads, head_size]
        query: torch.Tensor,            # [num_generation_tokens, num_heads, head_size]
        key_cache: torch.Tensor,        # [num_blocks, num_heads, head_size/x, block_size, x]
        value_cache: torch.Tensor,      # [num_blocks, num_heads, head_size, block_size]
        input_metadata: InputMetadata,
    ) -> None:
        block_size = value_cache.shape[3]
        attention_ops.single_query_cached_kv_attention(
            output, query, key_cache, value_cache, self.scale, input_metadata.block_tables,
            input_metadata.context_lens, block_size, input_metadata.max_context_len)

    def _attend(self, query, key, value, key_cache, value_cache, input_metadata, cache_event, kv_written):
        query = query.view(-1, self.num_heads, self.head_size)
        key = key.view(-1, self.num_heads, self.head_size)
        value = value.view(-1, self.num_heads, self.head_size)
        # dense output whatever the strides of the qkv slice (torch.empty_like would keep the gaps)
        output = torch.empty(query.shape, dtype=query.dtype, device=query.device)

        num_prompt_tokens = input_metadata.num_prompt_tokens
        if num_prompt_tokens > 0:
            self.multi_query_kv_attention(output[:num_prompt_tokens], query[:num_prompt_tokens],
                                          key[:num_prompt_tokens], value[:num_prompt_tokens],
                                          input_metadata.attn_bias)
        if cache_event is not None:
            cache_event.wait()
        num_valid_tokens = input_metadata.num_valid_tokens
        if not kv_written and num_valid_tokens > 0 and key_cache is not None and value_cache is not None:
            # a pure decode step writes one token per sequence, each into a block of its own; prompts fill runs of slots
            cache_ops.reshape_and_cache(key[:num_valid_tokens], value[:num_valid_tokens], key_cache, value_cache,
                                        input_metadata.slot_mapping,
                                        slots="scattered" if num_prompt_tokens == 0 else "runs")
        if input_metadata.num_generation_tokens > 0:
            assert key_cache is not None and value_cache is not None, \
                "key_cache and value_cache must be provided when generating tokens."
            self.single_query_cached_kv_attention(output[num_prompt_tokens:num_valid_tokens],
                                                  query[num_prompt_tokens:num_valid_tokens],
                                                  key_cache, value_cache, input_metadata)
        return output.view(-1, self.num_heads * self.head_size)

    def forward(
        self,
        query: torch.Tensor,                    # [num_tokens, num_heads * head_size]
        key: torch.Tensor,                      # [num_tokens, num_heads * head_size]
        value: torch.Tensor,                    # [num_tokens, num_heads * head_size]
        key_cache: Optional[torch.Tensor],      # [num_blocks, num_heads, head_size/x, block_size, x]
        value_cache: Optional[torch.Tensor],    # [num_blocks, num_heads, head_size, block_size]
        input_metadata: InputMetadata,
        cache_event: Optional[torch.cuda.Event],
    ) -> torch.Tensor:                          # [num_tokens, num_heads * head_size]
        # attention.py:93-156; query/key/value are slices of one [num_tokens, 3 * num_heads * head_size] tensor
        return self._attend(query, key, value, key_cache, value_cache, input_metadata, cache_event, False)


class PagedAttentionWithRoPE(PagedAttention):
    """PagedAttention with GPT-NeoX style rotary embedding (attention.py:159-217)."""

    def __init__(self, num_heads: int, head_size: int, scale: float, rotary_dim: int, max_position: int = 8192,
                 base: int = 10000, fuse_rope_cache: bool = True) -> None:
        super().__init__(num_heads, head_size, scale)
        self.fuse_rope_cache = fuse_rope_cache
        # cos / sin cache exactly as the reference builds it (:174-186)
        inv_freq = 1.0 / (base ** (torch.arange(0, rotary_dim, 2) / rotary_dim))
        t = torch.arange(max_position).float()
        freqs = torch.einsum("i,j -> ij", t, inv_freq.float())
        cache = torch.cat((freqs.cos(), freqs.sin()), dim=-1).to(torch.get_default_dtype())
        self.register_buffer("cos_sin_cache", cache, persistent=False)   # [max_position, rotary_dim]

    def forward(
        self,
        positions: torch.Tensor,                # [num_tokens]
        query: torch.Tensor,                    # [num_tokens, num_heads * head_size]
        key: torch.Tensor,                      # [num_tokens, num_heads * head_size]
        value: torch.Tensor,                    # [num_tokens, num_heads * head_size]
        key_cache: torch.Tensor,                # [num_blocks, num_heads, head_size/x, block_size, x]
        value_cache: torch.Tensor,              # [num_blocks, num_heads, head_size, block_size]
        input_metadata: InputMetadata,
        cache_event: Optional[torch.cuda.Event],
    ) -> torch.Tensor:                          # [num_tokens, num_heads * head_size]
        fused = (self.fuse_rope_cache and key_cache is not None and value_cache is not None
                 and input_metadata.num_valid_tokens > 0)
        if fused:
            # one launch: rotate q and k in place, write rotated k and v to their cache slots.  The cache write
            # moves ahead of the prompt attention, so the cache-op event is waited for here instead.
            if cache_event is not None:
                cache_event.wait()
            pos_encoding_ops.rotary_embedding_neox_reshape_and_cache(
                positions, query, key, value, key_cache, value_cache, input_metadata.slot_mapping, self.head_size,
                self.cos_sin_cache)
            return self._attend(query, key, value, key_cache, value_cache, input_metadata, None, True)
        pos_encoding_ops.rotary_embedding_neox(positions, query, key, self.head_size, self.cos_sin_cache)
        return self._attend(query, key, value, key_cache, value_cache, input_metadata, cache_event, False)
