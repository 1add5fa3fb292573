"""CUDA-graph capture of the operators that keep scratch inside the library (split-context decode, split prompt
items): a captured call must replay to the same bits, also after a later, larger eager call has outgrown the scratch
buffer the graph holds a pointer to (the buffer is retired, not freed)."""
import pytest
import torch

import cases
import pa_testing as pt


def _capture(fn):
    s = torch.cuda.Stream()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(s):
        fn()                                   # sizes every scratch buffer before the capture
        torch.cuda.synchronize()
        with torch.cuda.graph(g, stream=s):
            fn()
    torch.cuda.synchronize()
    return g, s


@pytest.mark.gpu
def test_split_context_decode_replays_after_the_workspace_grew():
    from vllm_b200 import attention_ops
    dtype, dev = torch.float16, "cuda"
    ctx = [4096, 3000, 4095]
    p = pt.make_decode_problem(len(ctx), 4, 128, 16, 1024, dtype, ctx, dist="unit", seed=5, unique_blocks=True).to(dev)
    plan = attention_ops.decode_plan(len(ctx), 4, 128, 16, max(ctx), dtype)
    assert plan.num_splits > 1, "this shape is meant to split its contexts"
    out = torch.empty((len(ctx), 4, 128), dtype=dtype, device=dev)
    call = lambda: attention_ops.single_query_cached_kv_attention(out, p.query, p.key_cache, p.value_cache, p.scale,
                                                                  p.block_tables, p.context_lens, 16, max(ctx))
    g, s = _capture(call)
    want = out.clone()
    out.zero_()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(pt.bits(out), pt.bits(want))
    # a larger problem on the SAME stream outgrows the split workspace
    ctx2 = [8192] * 24
    p2 = pt.make_decode_problem(len(ctx2), 8, 128, 16, 24 * 512 + 8, dtype, ctx2, dist="unit", seed=6, unique_blocks=True).to(dev)
    out2 = torch.empty((len(ctx2), 8, 128), dtype=dtype, device=dev)
    with torch.cuda.stream(s):
        attention_ops.single_query_cached_kv_attention(out2, p2.query, p2.key_cache, p2.value_cache, p2.scale,
                                                       p2.block_tables, p2.context_lens, 16, max(ctx2))
    torch.cuda.synchronize()
    out.zero_()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(pt.bits(out), pt.bits(want))


@pytest.mark.gpu
def test_split_prompt_items_replay_after_the_scratch_grew():
    from vllm_b200 import _cabi, attention_ops
    import test_cached_prefill as tcp
    dtype, dev = torch.bfloat16, "cuda"
    c = tcp._random_problem(dtype, 128, 16, [(100, 3007), (256, 2500), (1, 2049)], h=8, seed=31)
    q = c["qkv"].to(dev)[:, 0].contiguous()
    kc, vc, bt = c["key_cache"].to(dev), c["value_cache"].to(dev), c["block_tables"].to(dev)
    out = torch.empty_like(q)
    ctx = [int(x) for x in c["context_lens"]]
    call = lambda: attention_ops.multi_query_cached_kv_attention(c["cu_query_lens"], out, q, kc, vc, c["scale"], bt, ctx, 16)
    before = _cabi.launch_count()
    g, s = _capture(call)
    assert _cabi.launch_count() - before == 4, "two calls of prompt kernel + merge kernel expected (the batch is split)"
    want = out.clone()
    out.zero_()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(pt.bits(out), pt.bits(want))
    # a batch with many more parts on the same stream
    c2 = tcp._random_problem(dtype, 128, 16, [(128, 8000)] * 6, h=16, seed=32)
    q2 = c2["qkv"].to(dev)[:, 0].contiguous()
    out2 = torch.empty_like(q2)
    with torch.cuda.stream(s):
        attention_ops.multi_query_cached_kv_attention(c2["cu_query_lens"], out2, q2, c2["key_cache"].to(dev),
                                                      c2["value_cache"].to(dev), c2["scale"], c2["block_tables"].to(dev),
                                                      [int(x) for x in c2["context_lens"]], 16)
    torch.cuda.synchronize()
    out.zero_()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(pt.bits(out), pt.bits(want))
