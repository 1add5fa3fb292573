"""The C-ABI boundary without a GPU: the library loads, exports exactly what the header
declares, validates arguments before touching CUDA, and the Python operator shims mirror
the reference's signatures and error behaviour.  No compute is launched here."""
import ctypes as C
import inspect
import os
import re
import subprocess

import pytest
import torch

import vllm_b200
from vllm_b200 import _cabi, attention_ops, cache_ops

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "paged_attention_b200.h")


def _declared():
    src = open(HEADER).read()
    return sorted(set(re.findall(r"PA_API\s+[\w\s\*]+?\b(pa_\w+)\s*\(", src)))


def test_header_symbols_are_exported_and_bound():
    declared = _declared()
    assert len(declared) >= 12
    lib = _cabi.lib()
    for name in declared:
        assert hasattr(lib, name), f"{name} is declared in the header but not exported"
    assert sorted(_cabi.SIGNATURES) == declared, "ctypes SIGNATURES must list every PA_API symbol"
    exported = subprocess.run(["nm", "-D", "--defined-only", _cabi.LIB_PATH], capture_output=True, text=True).stdout
    extra = [l.split()[-1] for l in exported.splitlines() if " T " in l and not l.split()[-1].startswith("pa_")]
    assert not extra, f"unexpected exported symbols: {extra}"


def test_version_and_target():
    lib = _cabi.lib()
    declared = int(re.search(r"#define PA_ABI_VERSION (\d+)", open(HEADER).read()).group(1))
    assert lib.pa_abi_version() == declared == _cabi.PA_ABI_VERSION
    assert lib.pa_compiled_sm() == 100
    assert lib.pa_launch_count() == 0 or lib.pa_launch_count() > 0


def test_library_is_sm100a_only():
    out = subprocess.run(["cuobjdump", "-lelf", _cabi.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_library_has_no_torch_dependency():
    out = subprocess.run(["ldd", _cabi.LIB_PATH], capture_output=True, text=True).stdout
    assert "torch" not in out and "c10" not in out


@pytest.mark.parametrize("head,block,dtype,code", [
    (100, 16, 0, _cabi.PA_ERR_UNSUPPORTED_HEAD_SIZE),      # reference: attention_kernels.cu:398
    (128, 4, 0, _cabi.PA_ERR_UNSUPPORTED_BLOCK_SIZE),      # :446
    (128, 16, 7, _cabi.PA_ERR_UNSUPPORTED_DTYPE),          # :467
])
def test_unsupported_shapes_are_rejected_before_cuda(head, block, dtype, code):
    lib = _cabi.lib()
    rc = lib.pa_decode_attention(1, 1, 16, 16, 1.0, 1, 1, 1, 1, head, block, 1, 1, head, dtype, None)
    assert rc == code
    assert _cabi.last_error() != ""
    with pytest.raises(RuntimeError):
        _cabi.check(rc)


def test_invalid_arguments():
    lib = _cabi.lib()
    assert lib.pa_decode_attention(None, None, None, None, 1.0, None, None, 4, 2, 128, 16, 4, 64, 256, 0, None) \
        == _cabi.PA_ERR_INVALID_ARGUMENT
    assert lib.pa_decode_attention(16, 16, 24, 16, 1.0, 16, 16, 4, 2, 128, 16, 4, 64, 256, 0, None) \
        == _cabi.PA_ERR_INVALID_ARGUMENT       # misaligned key cache
    assert lib.pa_decode_attention(None, None, None, None, 1.0, None, None, 0, 2, 128, 16, 4, 64, 256, 0, None) == 0
    assert lib.pa_swap_blocks(16, 16, 64, None, None, 0, 9, None) == _cabi.PA_ERR_INVALID_ARGUMENT   # CPU<->CPU etc.
    assert lib.pa_swap_blocks(16, 16, 64, None, None, 0, _cabi.PA_COPY_D2H, None) == 0                # empty map
    assert lib.pa_copy_blocks(None, None, 0, None, 0, 64, None) == 0                                  # no layers (:88-90)
    assert lib.pa_reshape_and_cache(None, None, None, None, None, 0, 2, 16, 8, 32, 32, 0, None) == 0
    assert lib.pa_reshape_and_cache(16, 16, 16, 16, 16, 1, 2, 12, 8, 24, 24, 0, None) == _cabi.PA_ERR_INVALID_ARGUMENT


def test_plan_is_deterministic_and_sane():
    for (s, h, d, b, ctx, dt) in [(256, 32, 128, 16, 1024, torch.float16), (256, 5, 128, 16, 2048, torch.bfloat16),
                                  (64, 8, 96, 16, 8192, torch.float16), (64, 9, 128, 32, 8192, torch.float32),
                                  (1, 1, 64, 8, 1, torch.float32), (3, 40, 80, 8, 300000, torch.float16)]:
        p = attention_ops.decode_plan(s, h, d, b, ctx, dt)
        q = attention_ops.decode_plan(s, h, d, b, ctx, dt)
        assert (p.variant, p.num_warps, p.num_stages, p.num_splits) == (q.variant, q.num_warps, q.num_stages, q.num_splits)
        assert p.variant in (1, 2) and 1 <= p.num_warps <= 8 and 1 <= p.num_stages <= 32 and 1 <= p.num_splits <= 65535
        pages = (ctx + b - 1) // b
        assert p.num_splits <= max(pages, 1)
    # enough (seq, head) items to occupy the GPU: no split; few items and a long context: split
    assert attention_ops.decode_plan(256, 32, 128, 16, 1024, torch.float16).num_splits == 1
    assert attention_ops.decode_plan(64, 8, 96, 16, 8192, torch.float16).num_splits == 1
    assert attention_ops.decode_plan(4, 8, 128, 16, 8192, torch.float16).num_splits > 1
    # short contexts run as small CTAs
    assert attention_ops.decode_plan(256, 32, 128, 16, 64, torch.float16).num_warps == 1
    assert attention_ops.decode_plan(256, 32, 128, 16, 2048, torch.float16).num_warps == 4


def test_operator_signatures_match_the_reference():
    # csrc/attention.cpp:3-12 and csrc/cache.cpp:6-29 of the reference (positional names)
    sig = lambda f: [p.name for p in inspect.signature(f).parameters.values() if p.kind is p.POSITIONAL_OR_KEYWORD]
    assert sig(attention_ops.single_query_cached_kv_attention) == [
        "out", "query", "key_cache", "value_cache", "scale", "block_tables", "context_lens", "block_size",
        "max_context_len"]
    assert sig(cache_ops.reshape_and_cache) == ["key", "value", "key_cache", "value_cache", "slot_mapping"]
    assert sig(cache_ops.gather_cached_kv) == ["key", "value", "key_cache", "value_cache", "slot_mapping"]
    assert sig(cache_ops.copy_blocks) == ["key_caches", "value_caches", "block_mapping"]
    assert sig(cache_ops.swap_blocks) == ["src", "dst", "block_mapping"]


@pytest.fixture(params=["torch", "ctypes"])
def binding(request):
    """Both routes to the C ABI: the compiled operator binding (lib/pa_torch_ops.so) and plain ctypes."""
    _cabi.use_binding(request.param)
    yield request.param
    _cabi.use_binding("")


def test_compiled_binding_is_built_and_links_the_c_abi():
    ops = _cabi.torch_ops()
    assert ops is not None, "lib/pa_torch_ops.so must be built by build()"
    for name in ("single_query_cached_kv_attention", "reshape_and_cache", "gather_cached_kv", "copy_blocks",
                 "swap_blocks"):
        assert hasattr(ops, name)
    out = subprocess.run(["ldd", _cabi.TORCH_OPS_PATH], capture_output=True, text=True).stdout
    assert "libpaged_b200.so" in out
    syms = subprocess.run(["nm", "-D", "--undefined-only", _cabi.TORCH_OPS_PATH], capture_output=True, text=True).stdout
    for name in ("pa_decode_attention_ex", "pa_reshape_and_cache", "pa_gather_cached_kv", "pa_copy_blocks",
                 "pa_swap_blocks_ex"):
        assert name in syms, f"the binding must forward to {name}"
    assert "cudaLaunchKernel" not in syms          # no kernel code in the binding


def test_no_cpu_fallback(binding):
    """CPU tensors are refused: the product path never computes on the host."""
    kshape, vshape = vllm_b200.kv_cache_shapes(2, 64, 16, torch.float16)
    kc, vc = torch.zeros(4, *kshape, dtype=torch.float16), torch.zeros(4, *vshape, dtype=torch.float16)
    q = torch.zeros(3, 2, 64, dtype=torch.float16)
    with pytest.raises(RuntimeError):
        attention_ops.single_query_cached_kv_attention(torch.empty_like(q), q, kc, vc, 0.125,
                                                       torch.zeros(3, 2, dtype=torch.int32),
                                                       torch.ones(3, dtype=torch.int32), 16, 16)
    with pytest.raises(RuntimeError):
        cache_ops.reshape_and_cache(q, q, kc, vc, torch.zeros(3, dtype=torch.int32))
    with pytest.raises(RuntimeError):
        cache_ops.gather_cached_kv(q, q, kc, vc, torch.zeros(3, dtype=torch.int32))
    with pytest.raises(RuntimeError):
        cache_ops.copy_blocks([kc], [vc], {0: [1]})
    with pytest.raises(RuntimeError, match="Invalid device combination"):
        cache_ops.swap_blocks(kc, kc.clone(), {0: 1})          # CPU<->CPU, cache_kernels.cu:26
    with pytest.raises(RuntimeError):
        cache_ops.copy_blocks([kc], [vc, vc], {0: [1]})        # layer-count mismatch, cache_kernels.cu:87
    cache_ops.copy_blocks([], [], {0: [1]})                    # zero layers returns early, :88-90


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setattr(_cabi, "_lib", None)
    monkeypatch.setattr(_cabi, "LIB_PATH", str(tmp_path / "libpaged_b200.so"))
    with pytest.raises(RuntimeError, match="no CPU or PyTorch fallback"):
        _cabi.lib()


def test_stale_library_is_refused(monkeypatch):
    """A library built from an older header must not be bound silently (symbols may be missing or changed)."""
    monkeypatch.setattr(_cabi, "_lib", None)
    monkeypatch.setattr(_cabi, "PA_ABI_VERSION", _cabi.PA_ABI_VERSION + 1)
    with pytest.raises(RuntimeError, match="C ABI version"):
        _cabi.lib()


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "vllm-v0.1.0_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text, f"{f} mentions the oracle"


def test_kv_cache_shapes_follow_cache_engine():
    # vllm/worker/cache_engine.py:48-63
    assert vllm_b200.kv_cache_shapes(40, 128, 16, torch.float16) == ((40, 16, 16, 8), (40, 128, 16))
    assert vllm_b200.kv_cache_shapes(8, 96, 32, torch.float32) == ((8, 24, 32, 4), (8, 96, 32))
    assert vllm_b200.CacheEngine.get_cache_block_size(16, 40, 128, 40, torch.float16) == 2 * 40 * 163840
