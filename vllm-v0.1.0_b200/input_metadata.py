"""Mirror of the reference's `InputMetadata` (vllm/model_executor/input_metadata.py:10-40) for the
attention path: same constructor arguments and attributes.  The only difference is `attn_bias`:
the reference builds an xformers `BlockDiagonalCausalMask.from_seqlens(prompt_lens)` (:33); the
prompt-attention kernel here takes the prompt lengths themselves, so `attn_bias` is a small
stand-in object that just remembers them (same `from_seqlens` constructor)."""
from __future__ import annotations

import dataclasses
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


@dataclasses.dataclass(repr=False, eq=False)
class InputMetadata:
    """The seven constructor fields, in the reference's order (positional and keyword construction both work as
    there); everything else is derived once after construction.  Derived counts are properties of the tensors'
    shapes, checked against each other like the reference's asserts (:36-37)."""
    seq_groups: List[Tuple[List[int], Any]]      # (seq_ids, sampling_params) per group
    seq_data: Dict[int, Any]                     # seq_id -> SequenceData
    prompt_lens: List[int]
    slot_mapping: torch.Tensor                   # [num_valid_tokens] int32
    context_lens: torch.Tensor                   # [num_generation_tokens] int32
    max_context_len: int
    block_tables: torch.Tensor                   # [num_generation_tokens, max_num_blocks_per_seq] int32

    # derived in __post_init__ (read by the attention layer and the sampler under exactly these names)
    attn_bias: BlockDiagonalCausalMask = dataclasses.field(init=False)
    num_prompts: int = dataclasses.field(init=False)
    num_prompt_tokens: int = dataclasses.field(init=False)
    num_generation_tokens: int = dataclasses.field(init=False)
    num_valid_tokens: int = dataclasses.field(init=False)
    max_num_blocks_per_seq: int = dataclasses.field(init=False)

    def __post_init__(self) -> None:
        self.attn_bias = BlockDiagonalCausalMask.from_seqlens(self.prompt_lens)
        self.num_prompts, self.num_prompt_tokens = len(self.prompt_lens), sum(self.prompt_lens)
        self.num_generation_tokens = int(self.context_lens.shape[0])
        self.num_valid_tokens = int(self.slot_mapping.shape[0])
        self.max_num_blocks_per_seq = int(self.block_tables.shape[1]) if self.block_tables.numel() > 0 else 0
        if self.block_tables.shape[0] != self.num_generation_tokens:
            raise AssertionError(f"{self.block_tables.shape[0]} block-table rows for {self.num_generation_tokens} "
                                 "generation tokens")

    def __repr__(self) -> str:
        shown = ("num_valid_tokens", "num_prompt_tokens", "num_prompts", "prompt_lens", "num_generation_tokens",
                 "max_context_len", "max_num_blocks_per_seq")
        return "InputMetadata(" + ", ".join(f"{k}={getattr(self, k)}" for k in shown) + ")"
