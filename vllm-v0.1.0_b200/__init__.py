"""B200-native PagedAttention decode path (vLLM v0.1.0 operator contract).

`attention_ops` and `cache_ops` carry the reference's operator signatures and
call libpaged_b200.so (hand-written sm_100a CUDA) through its C ABI.
Import as `vllm_b200` (the repo-root alias package; this directory's name is
not a valid Python identifier).
"""
from . import _cabi, attention_ops, cache_ops, pos_encoding_ops  # noqa: F401
from .cache_engine import CacheEngine, kv_cache_shapes  # noqa: F401
from . import sharding  # noqa: F401
from .input_metadata import BlockDiagonalCausalMask, InputMetadata  # noqa: F401
from .paged_attention import PagedAttention, PagedAttentionWithRoPE  # noqa: F401
from .step_inputs import StepInputs  # noqa: F401

__all__ = ["attention_ops", "cache_ops", "pos_encoding_ops", "CacheEngine", "kv_cache_shapes", "InputMetadata",
           "BlockDiagonalCausalMask", "PagedAttention", "PagedAttentionWithRoPE", "StepInputs"]
