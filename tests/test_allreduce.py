"""All-reduce after attention (pa_allreduce; mirror of the reference's GraphAllReduce,
vllm/model_executor/parallel_utils/parallel_state.py:506-556).  The peer path is checked bit-exactly on one
GPU with every rank of the world inside one launch; the multimem path and the real multi-process wiring are
checked by tools/allreduce_check.py under torchrun on 2+ GPUs."""
import ctypes as C

import pytest
import torch

import cases


def _oracle_sum(bufs):
    """fp32 accumulation in rank order from zero, one rounding to the dtype: what the peer path computes."""
    acc = torch.zeros(bufs[0].shape, dtype=torch.float32)
    for b in bufs:
        acc += b.float()
    return acc.to(bufs[0].dtype)


def test_allreduce_symbols_and_argument_checks():
    from vllm_b200 import _cabi
    lib = _cabi.lib()
    assert lib.pa_allreduce_flag_bytes() >= 128 * 8 * 4
    one = (C.c_void_p * 1)(None)
    assert lib.pa_allreduce(one, one, None, 0, 16, 0, 9, _cabi.PA_F16, None) == _cabi.PA_ERR_INVALID_ARGUMENT
    assert lib.pa_allreduce(one, one, None, 0, 16, 3, 2, _cabi.PA_F16, None) == _cabi.PA_ERR_INVALID_ARGUMENT
    assert lib.pa_allreduce(one, one, None, 0, 16, 0, 2, 7, None) == _cabi.PA_ERR_UNSUPPORTED_DTYPE
    assert lib.pa_allreduce(one, one, None, 0, 16, 0, 1, _cabi.PA_F16, None) == _cabi.PA_OK     # world 1: nothing to do
    two = (C.c_void_p * 2)(None, None)
    assert lib.pa_allreduce(two, two, None, 0, 16, 0, 2, _cabi.PA_F16, None) == _cabi.PA_ERR_INVALID_ARGUMENT
    assert "null peer pointer" in _cabi.last_error()


def test_graph_all_reduce_single_rank_is_identity():
    from vllm_b200.all_reduce import GraphAllReduce
    ar = GraphAllReduce(max_num_tokens=16, hidden_size=64, dtype=torch.float16)
    x = torch.ones(4, 64)
    assert ar.launch(x) is x and ar.world_size == 1


@pytest.mark.gpu
@pytest.mark.parametrize("dt", ["f16", "bf16", "f32"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_allreduce_peer_path_emulated_world(dt, world):
    from vllm_b200 import _cabi
    from vllm_b200.all_reduce import allreduce_emulated
    dtype = cases.DTYPES[dt]
    tokens, hidden = 37, 5120                       # 37 rows: slices that do not divide evenly
    g = torch.Generator().manual_seed(world)
    host = [torch.randn((tokens, hidden), generator=g).to(dtype) for _ in range(world)]
    bufs = [h.cuda() for h in host]
    flags = [torch.zeros(_cabi.lib().pa_allreduce_flag_bytes() // 4, dtype=torch.int32, device="cuda") for _ in range(world)]
    want = _oracle_sum(host)
    allreduce_emulated(bufs, flags, tokens * hidden)
    torch.cuda.synchronize()
    for r in range(world):
        assert torch.equal(bufs[r].cpu().view(torch.uint8), want.view(torch.uint8)), f"rank {r} differs"
        assert int(flags[r].abs().sum()) == 0        # flags return to zero: reusable without a reset
    # second call on the results (flags reused), on a sub-range only: rows [8, 24)
    host2 = [want.clone() for _ in range(world)]
    want2 = want.clone()
    want2[8:24] = _oracle_sum([h[8:24] for h in host2])
    allreduce_emulated(bufs, flags, 16 * hidden, offset_bytes=8 * hidden * want.element_size())
    torch.cuda.synchronize()
    for r in range(world):
        assert torch.equal(bufs[r].cpu().view(torch.uint8), want2.view(torch.uint8)), f"rank {r} differs (2nd call)"
