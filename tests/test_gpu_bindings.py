"""The two routes to the C ABI -- the compiled operator binding (csrc/torch_binding.cpp) and
plain ctypes (_cabi.py) -- launch the same kernels: bit-identical results, same errors."""
import pytest
import torch

import pa_testing as pt
from vllm_b200 import _cabi, attention_ops, cache_ops

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(autouse=True)
def _restore_binding():
    yield
    _cabi.use_binding("")


def _both(fn):
    res = []
    for name in ("torch", "ctypes"):
        _cabi.use_binding(name)
        assert (_cabi.torch_ops() is not None) == (name == "torch")
        res.append(fn())
    return res


def test_default_binding_is_the_compiled_one():
    _cabi.use_binding("")
    assert _cabi.torch_ops() is not None


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
def test_decode_is_bit_identical_through_both_bindings(dtype):
    p = pt.make_decode_problem(9, 5, 128, 16, 300, dtype, [1, 16, 17, 300, 1024, 2049, 64, 999, 4000], seed=4)
    want = pt.oracle_decode(p)
    pd = p.to(DEV)
    for tun in (None, pt.tuning(1, 4, 0, 1), pt.tuning(2, 4, 2, 3)):
        a, b = _both(lambda: pt.run_decode(pd, tuning=tun))
        assert torch.equal(a, b)
        pt.assert_decode_close(a, want, dtype, "binding")


def test_movers_are_bit_identical_through_both_bindings():
    n, h, d, b, nb = 100, 4, 128, 16, 32
    qkv = torch.randn((n, 3, h, d), device=DEV).half()
    slots = torch.randperm(nb * b, device=DEV)[:n].to(torch.int32)

    def movers():
        kc = torch.zeros((nb, h, d // 8, b, 8), dtype=torch.float16, device=DEV)
        vc = torch.zeros((nb, h, d, b), dtype=torch.float16, device=DEV)
        cache_ops.reshape_and_cache(qkv[:, 1], qkv[:, 2], kc, vc, slots)
        cache_ops.copy_blocks([kc], [vc], {0: [20, 21], 3: [22]})
        back = torch.zeros_like(qkv)
        cache_ops.gather_cached_kv(back[:, 1], back[:, 2], kc, vc, slots)
        host = torch.zeros((nb, h, d, b), dtype=torch.float16).pin_memory()
        cache_ops.swap_blocks(vc, host, {i: (i * 5) % nb for i in range(0, nb, 3)})
        torch.cuda.synchronize()
        return kc, vc, back, host

    a, b2 = _both(movers)
    for x, y in zip(a, b2):
        assert torch.equal(x.cpu(), y.cpu())


@pytest.mark.parametrize("name", ["torch", "ctypes"])
def test_errors_through_both_bindings(name):
    _cabi.use_binding(name)
    p = pt.make_decode_problem(2, 2, 64, 16, 8, torch.float16, [20, 20]).to(DEV)
    out = torch.empty_like(p.query.contiguous())
    with pytest.raises(RuntimeError):
        attention_ops.single_query_cached_kv_attention(out, p.query, p.key_cache, p.value_cache, p.scale,
                                                       p.block_tables.long(), p.context_lens, 16, 20)
    q112 = torch.zeros(2, 2, 112, dtype=torch.float16, device=DEV)
    kc = torch.zeros(8, 2, 14, 16, 8, dtype=torch.float16, device=DEV)
    vc = torch.zeros(8, 2, 112, 16, dtype=torch.float16, device=DEV)
    with pytest.raises(RuntimeError, match="head size"):
        attention_ops.single_query_cached_kv_attention(torch.empty_like(q112), q112, kc, vc, 0.1, p.block_tables,
                                                       p.context_lens, 16, 20)
    with pytest.raises(RuntimeError, match="same GPU|Invalid device"):
        cache_ops.swap_blocks(torch.zeros(2, 4), torch.zeros(2, 4), {0: 1})
    with pytest.raises(RuntimeError):
        cache_ops.copy_blocks([p.key_cache], [p.value_cache, p.value_cache], {0: [1]})
    with pytest.raises(RuntimeError):
        cache_ops.reshape_and_cache(p.query, p.query, p.key_cache, p.value_cache,
                                    torch.zeros(2, dtype=torch.int64, device=DEV))
