"""Mirror of the reference's `GraphAllReduce` (vllm/model_executor/parallel_utils/parallel_state.py:506-556)
and of the module-level launcher accessors around it (:191-204, :474-476).

Same constructor arguments, same `buffer` attribute (RowParallelLinear writes its matmul output into a
slice of it, tensor_parallel/layers.py:419-426) and the same `launch(x)` contract: `x` must be a leading
slice of `buffer`; it is summed over the tensor-parallel group in place and returned.

The reference replays one NCCL all-reduce per token count from a CUDA graph.  Here `buffer` lives in
symmetric memory (torch.distributed._symmetric_memory: allocation + handle exchange only) and `launch`
is one launch of libpaged_b200's own NVLink kernel (pa_allreduce: multimem.ld_reduce / multimem.st through
the NVSwitch when the buffer has a multicast mapping, peer loads / stores otherwise).  A single stream-ordered
kernel needs no graph, so `disable_graph` is accepted and ignored; the launch is itself graph-capturable.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _cabi
from .attention_ops import _dtype_code

_ALL_REDUCE_LAUNCHER: Optional["GraphAllReduce"] = None


class GraphAllReduce:

    def __init__(self, max_num_tokens: int, hidden_size: int, dtype: torch.dtype, disable_graph: bool = False,
                 group=None, use_multimem: Optional[bool] = None,
                 two_rank_peer_threshold_bytes: int = 4 << 20) -> None:
        # two_rank_peer_threshold_bytes: with two ranks and use_multimem=None, messages larger than this take the peer
        # path although a multicast mapping exists (one box, two sizes: 26 MB 54 us peer vs 86 us multimem, 2.6 MB
        # equal -- the crossover has not been swept, hence a parameter); path_for() reports the choice
        import torch.distributed as dist
        self.max_num_tokens = max_num_tokens
        self.hidden_size = hidden_size
        self.disable_graph = disable_graph
        self.two_rank_peer_threshold_bytes = two_rank_peer_threshold_bytes
        self.world_size = dist.get_world_size(group) if dist.is_initialized() else 1
        if self.world_size == 1:
            return
        if self.world_size > 8:
            raise RuntimeError("GraphAllReduce: at most 8 ranks (one NVSwitch domain)")
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group if group is not None else dist.group.WORLD
        self.rank = dist.get_rank(self.group)
        dev = torch.device("cuda", torch.cuda.current_device())
        self.buffer = symm_mem.empty((max_num_tokens, hidden_size), dtype=dtype, device=dev)
        self._buf_hdl = symm_mem.rendezvous(self.buffer, self.group)
        lib = _cabi.lib()
        self._flags = symm_mem.empty((lib.pa_allreduce_flag_bytes() // 4,), dtype=torch.int32, device=dev)
        self._flags.zero_()
        self._flag_hdl = symm_mem.rendezvous(self._flags, self.group)
        torch.cuda.synchronize(dev)
        dist.barrier(group=self.group)                # every rank's flags are zero before anyone signals
        n = self.world_size
        self._bufs = (C.c_void_p * n)(*[int(p) for p in self._buf_hdl.buffer_ptrs])
        self._flag_ptrs = (C.c_void_p * n)(*[int(p) for p in self._flag_hdl.buffer_ptrs])
        mc = int(getattr(self._buf_hdl, "multicast_ptr", 0) or 0)
        self._forced_multimem = use_multimem is True
        if use_multimem is None:
            use_multimem = mc != 0
        if use_multimem and mc == 0:
            raise RuntimeError("GraphAllReduce: this buffer has no multicast mapping (no NVLS on this system)")
        self.multicast_ptr = mc if use_multimem else 0
        self._code = _dtype_code(self.buffer)

    def path_for(self, nbytes: int) -> str:
        """"multimem" (reduction inside the NVSwitch) or "peer" (16-byte peer loads summed in rank order) for a message
        of nbytes.  Every rank derives the same answer from the same arguments."""
        if self.world_size == 1 or not self.multicast_ptr:
            return "peer"
        # two ranks, large message: reading the one peer directly beats the round trip through the switch's reduction
        if self.world_size == 2 and nbytes > self.two_rank_peer_threshold_bytes and not self._forced_multimem:
            return "peer"
        return "multimem"

    def launch(self, x: torch.Tensor) -> torch.Tensor:
        # NOTE: x must be a slice of self.buffer (parallel_state.py:550)
        if self.world_size == 1:
            return x
        offset = x.data_ptr() - self.buffer.data_ptr()
        if offset < 0 or offset + x.numel() * x.element_size() > self.buffer.numel() * self.buffer.element_size() \
                or not x.is_contiguous():
            raise RuntimeError("GraphAllReduce.launch: x must be a contiguous slice of self.buffer")
        stream = torch.cuda.current_stream(x.device).cuda_stream
        mc = self.multicast_ptr if self.path_for(x.numel() * x.element_size()) == "multimem" else 0
        rc = _cabi.lib().pa_allreduce(self._bufs, self._flag_ptrs, mc or None, offset, x.numel(),
                                      self.rank, self.world_size, self._code, stream)
        _cabi.check(rc)
        return x


def initialize_all_reduce_launcher(max_num_tokens: int, hidden_size: int, dtype: torch.dtype,
                                   disable_graph: bool = False, group=None) -> None:
    global _ALL_REDUCE_LAUNCHER
    _ALL_REDUCE_LAUNCHER = GraphAllReduce(max_num_tokens=max_num_tokens, hidden_size=hidden_size, dtype=dtype,
                                          disable_graph=disable_graph, group=group)


def get_all_reduce_launcher() -> GraphAllReduce:
    assert _ALL_REDUCE_LAUNCHER is not None, "all reduce launcher is not initialized"
    return _ALL_REDUCE_LAUNCHER


def allreduce_emulated(buffers, flags, numel: int, offset_bytes: int = 0) -> None:
    """Tests: the peer path of pa_allreduce with every rank of the world inside one launch on one GPU.
    `buffers` / `flags`: one device tensor per emulated rank (flags int32, zero)."""
    n = len(buffers)
    bufs = (C.c_void_p * n)(*[b.data_ptr() for b in buffers])
    fl = (C.c_void_p * n)(*[f.data_ptr() for f in flags])
    dev = buffers[0].device
    with torch.cuda.device(dev):
        rc = _cabi.lib().pa_allreduce_emulated(bufs, fl, offset_bytes, numel, n, _dtype_code(buffers[0]),
                                               torch.cuda.current_stream(dev).cuda_stream)
    _cabi.check(rc)
