"""pa_decode_attention_static: decode attention for hosts that may not allocate device memory.

Some programs that embed this library cannot let it call cudaMalloc: a serving engine that gives the whole GPU to
its own memory pool at start-up (a later driver allocation has nothing left to succeed from), a plug-in whose host
forbids allocation once initialisation is over, a step loop that cannot afford the device-wide synchronisation an
allocation or a free may cause.  The ordinary entry point allocates its split-context scratch on first use; this
variant keeps that scratch in a 4 MiB array built into the library (one copy per GPU, shared by all calls, which is
why calls must not overlap on the GPU and why sizes beyond the array are refused).

The file is named so that pytest collects it after every other test file: `pytest -x` stops at the first failure,
and a failure here must not hide the rest of the suite.
"""
import ctypes as C

import pytest
import torch

import pa_testing as pt


def _call(fn, p, out, tuning=None):
    """One call through the C ABI (ctypes, no operator shim in between) on the current stream."""
    from vllm_b200 import _cabi
    code = {torch.float16: _cabi.PA_F16, torch.bfloat16: _cabi.PA_BF16, torch.float32: _cabi.PA_F32}[p.query.dtype]
    s, h, d = p.query.shape
    return fn(out.data_ptr(), p.query.data_ptr(), p.key_cache.data_ptr(), p.value_cache.data_ptr(), float(p.scale),
              p.block_tables.data_ptr(), p.context_lens.data_ptr(), s, h, d, p.block_size, p.block_tables.shape[1],
              p.max_context_len, p.query.stride(0), code, C.byref(tuning) if tuning is not None else None,
              torch.cuda.current_stream().cuda_stream)


@pytest.mark.gpu
def test_static_scratch_decode_matches_the_allocating_entry_point_twice_in_a_row():
    """Two different split-context problems back to back: the second one starts from whatever the first left in the
    shared array (other ticket count, other partials) and must still agree with pa_decode_attention_ex."""
    from vllm_b200 import _cabi, attention_ops
    lib, dev = _cabi.lib(), "cuda"
    problems = [
        # (dtype, heads, head size, block size, context lengths): few items and long contexts, so the plan splits
        (torch.float16, 4, 128, 16, [4096, 3000, 4095]),
        (torch.bfloat16, 2, 96, 32, [6000, 1, 257, 5999, 33]),      # ragged: some splits find no pages at all
    ]
    for call_no, (dtype, heads, d, b, ctx) in enumerate(problems):
        plan = attention_ops.decode_plan(len(ctx), heads, d, b, max(ctx), dtype)
        assert plan.num_splits > 1, "the shape is meant to use the split-context scratch"
        blocks = sum((c + b - 1) // b for c in ctx) + 8
        host = pt.make_decode_problem(len(ctx), heads, d, b, blocks, dtype, ctx, dist="unit", seed=40 + call_no,
                                      unique_blocks=True)
        p = host.to(dev)
        want = torch.empty((len(ctx), heads, d), dtype=dtype, device=dev)
        _cabi.check(_call(lib.pa_decode_attention_ex, p, want))
        got = torch.full_like(want, float("nan"))
        _cabi.check(_call(lib.pa_decode_attention_static, p, got))
        torch.cuda.synchronize()
        pt.assert_decode_close(got, want, dtype, f"static scratch, call {call_no}")
        pt.assert_decode_close(got, pt.oracle_decode(host), dtype, f"static scratch vs oracle, call {call_no}")
    # the same problem once more: the array now holds this very problem's finished tickets and partials
    got2 = torch.full_like(want, float("nan"))
    _cabi.check(_call(lib.pa_decode_attention_static, p, got2))
    torch.cuda.synchronize()
    pt.assert_decode_close(got2, want, dtype, "static scratch, repeated call")


@pytest.mark.gpu
def test_static_scratch_decode_without_splits_needs_no_scratch():
    """A batch large enough to fill the GPU is not split; the entry point then has no size limit to apply."""
    from vllm_b200 import _cabi, attention_ops
    lib, dev, dtype = _cabi.lib(), "cuda", torch.float16
    ctx = [100 + 3 * i for i in range(64)]
    assert attention_ops.decode_plan(len(ctx), 32, 64, 8, max(ctx), dtype).num_splits == 1
    p = pt.make_decode_problem(len(ctx), 32, 64, 8, 512, dtype, ctx, dist="unit", seed=43).to(dev)
    want = torch.empty((len(ctx), 32, 64), dtype=dtype, device=dev)
    got = torch.full_like(want, float("nan"))
    _cabi.check(_call(lib.pa_decode_attention_ex, p, want))
    _cabi.check(_call(lib.pa_decode_attention_static, p, got))
    torch.cuda.synchronize()
    pt.assert_decode_close(got, want, dtype, "static scratch, unsplit")


def test_static_scratch_decode_refuses_what_does_not_fit():
    """Sizes are checked on the host before anything is launched, so this runs without a GPU: a forced split count
    whose partial results exceed the built-in array, and more (sequence, head) items than it has tickets for."""
    from vllm_b200 import _cabi
    lib = _cabi.lib()
    args = lambda seqs, heads, ctx, tun: (16, 16, 16, 16, 1.0, 16, 16, seqs, heads, 128, 16, (ctx + 15) // 16, ctx,
                                          128, _cabi.PA_F16, C.byref(tun), None)
    too_many_parts = _cabi.DecodeTuning(0, 0, 0, 64)          # 64 x 40 items x 64 splits x 130 floats = 85 MB
    assert lib.pa_decode_attention_static(*args(64, 40, 16384, too_many_parts)) == _cabi.PA_ERR_INVALID_ARGUMENT
    msg = _cabi.last_error()
    assert "PA_STATIC_SCRATCH_BYTES" in msg and "pa_decode_attention_static" in msg
    too_many_items = _cabi.DecodeTuning(0, 0, 0, 2)           # 512 x 40 = 20480 items > 16384 tickets
    assert lib.pa_decode_attention_static(*args(512, 40, 4096, too_many_items)) == _cabi.PA_ERR_INVALID_ARGUMENT
    assert "PA_STATIC_SCRATCH_MAX_ITEMS" in _cabi.last_error()
    # shape and argument errors are the ordinary entry point's
    assert lib.pa_decode_attention_static(16, 16, 16, 16, 1.0, 16, 16, 1, 1, 100, 16, 1, 1, 100, 0, None, None) \
        == _cabi.PA_ERR_UNSUPPORTED_HEAD_SIZE
    assert lib.pa_decode_attention_static(None, None, None, None, 1.0, None, None, 4, 2, 128, 16, 4, 64, 256, 0, None,
                                          None) == _cabi.PA_ERR_INVALID_ARGUMENT
