"""Mirror of the reference's `InputMetadata` (vllm/model_executor/input_metadata.py:10-40) for the
attention path: same constructor arguments and attributes.  The only difference is `attn_bias`:
the reference builds an xformers `BlockDiagonalCausalMask.from_seqlens(prompt_lens)` (:33); the
prompt-attention kernel here takes the prompt lengths themselves, so `attn_bias` is a small
stand-in object that just remembers them (same `from_seqlens` constructor)."""
from __future__ import annotations

from typing import Any, Dict, List, Sequence, Tuple

import torch


class BlockDiagonalCausalMask:
    """Stand-in for xformers.ops.fmha.attn_bias.BlockDiagonalCausalMask: block-diagonal (one block
    per prompt), causal inside each block.  Holds only what defines it."""

    def __init__(self, seqlens: Sequence[int]) -> None:
        self.seqlens: List[int] = [int(n) for n in seqlens]

    @classmethod
    def from_seqlens(cls, seqlens: Sequence[int]) -> "BlockDiagonalCausalMask":
        return cls(seqlens)


class InputMetadata:

    def __init__(
        self,
        seq_groups: List[Tuple[List[int], Any]],     # (seq_ids, sampling_params) per group
        seq_data: Dict[int, Any],                    # seq_id -> SequenceData
        prompt_lens: List[int],
        slot_mapping: torch.Tensor,                  # [num_valid_tokens] int32
        context_lens: torch.Tensor,                  # [num_generation_tokens] int32
        max_context_len: int,
        block_tables: torch.Tensor,                  # [num_generation_tokens, max_num_blocks_per_seq] int32
    ) -> None:
        self.seq_groups = seq_groups
        self.seq_data = seq_data
        self.prompt_lens = prompt_lens
        self.slot_mapping = slot_mapping
        self.context_lens = context_lens
        self.max_context_len = max_context_len
        self.block_tables = block_tables

        self.attn_bias = BlockDiagonalCausalMask.from_seqlens(prompt_lens)
        self.num_prompts = len(prompt_lens)
        self.num_prompt_tokens = sum(prompt_lens)
        self.num_generation_tokens = context_lens.shape[0]
        self.num_valid_tokens = slot_mapping.shape[0]
        if block_tables.numel() > 0:
            self.max_num_blocks_per_seq = block_tables.shape[1]
        else:
            self.max_num_blocks_per_seq = 0
        assert block_tables.shape[0] == self.num_generation_tokens
        assert context_lens.shape[0] == self.num_generation_tokens

    def __repr__(self) -> str:
        return (f"InputMetadata(num_valid_tokens={self.num_valid_tokens}, "
                f"num_prompt_tokens={self.num_prompt_tokens}, num_prompts={self.num_prompts}, "
                f"prompt_lens={self.prompt_lens}, num_generation_tokens={self.num_generation_tokens}, "
                f"max_context_len={self.max_context_len}, max_num_blocks_per_seq={self.max_num_blocks_per_seq})")
