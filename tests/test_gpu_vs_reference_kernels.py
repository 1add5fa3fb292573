"""Our sm_100a operators against the REFERENCE's own CUDA operators on the same GPU tensors.

oracle/_ref/ref_attention_ops.so and ref_cache_ops.so are the reference's csrc/ sources compiled
unmodified for sm_100 by oracle/build_ref.py (test infrastructure; built in the container that
has /root/reference, shipped to the GPU box as binaries).  Skipped when they are absent.

Cache movers must agree bit for bit -- this is also what pins swap_blocks, for which the
reference has no test of its own.  Decode attention must agree within the north-star tolerance
(both kernels round differently: the reference multiplies P.V in T precision)."""
import importlib.util
import os
import random

import pytest
import torch

import pa_testing as pt

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
REF_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref")


def _load(name):
    path = os.path.join(REF_DIR, name + ".so")
    if not os.path.exists(path):
        pytest.skip(f"{path} not built (python oracle/build_ref.py where /root/reference exists)")
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="module")
def ref_attention():
    return _load("ref_attention_ops")


@pytest.fixture(scope="module")
def ref_cache():
    return _load("ref_cache_ops")


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("block_size", [8, 16, 32])
@pytest.mark.parametrize("head_size", [64, 80, 96, 128])
@pytest.mark.parametrize("dist", ["ref", "unit"])
def test_decode_attention_matches_reference_kernel(ref_attention, dtype, block_size, head_size, dist):
    """The reference test's shape (tests/kernels/test_attention.py:270-287): 37 sequences, 3 heads,
    1024 blocks, context <= 4096, strided query."""
    rnd = random.Random(block_size * 1000 + head_size)
    ctx = [rnd.randint(1, 4096) for _ in range(37)]
    p = pt.make_decode_problem(37, 3, head_size, block_size, 1024, dtype, ctx, dist=dist, seed=head_size).to(DEV)
    ours = pt.run_decode(p)
    theirs = torch.empty_like(ours)
    ref_attention.single_query_cached_kv_attention(theirs, p.query, p.key_cache, p.value_cache, p.scale,
                                                   p.block_tables, p.context_lens, block_size, p.max_context_len)
    torch.cuda.synchronize()
    if dist == "ref":
        # the reference test's own criterion (test_attention.py:224).  On these U(+-1e-3) inputs the reference
        # kernel's fp16 P.V products (p*v ~ 1e-7) fall into fp16 subnormals, so it is itself only good to a few
        # percent of the output scale there; ours is held to 1e-3 of the fp64 oracle in test_gpu_decode.py.
        assert torch.allclose(ours, theirs, atol=1e-3, rtol=1e-5)
        scale = float(theirs.float().abs().max())
        assert float((ours.float() - theirs.float()).abs().max()) <= 0.05 * scale
    else:
        pt.assert_decode_close(ours, theirs, dtype, f"vs reference kernel {dtype} B{block_size} D{head_size} {dist}")


def test_decode_attention_config2_matches_reference_kernel(ref_attention):
    """BASELINE configs[1] at full size (NB*H*D*B = 1.07e9 < 2^31, so the reference's int offsets hold)."""
    S, H, D, B, L = 256, 32, 128, 16, 1024
    g = torch.Generator(device=DEV).manual_seed(0)
    nblk = L // B
    kc = torch.empty((S * nblk, H, D // 8, B, 8), dtype=torch.float16, device=DEV).normal_(generator=g)
    vc = torch.empty((S * nblk, H, D, B), dtype=torch.float16, device=DEV).normal_(generator=g)
    q = torch.empty((S, 3, H, D), dtype=torch.float16, device=DEV).normal_(generator=g)[:, 0]
    tables = torch.randperm(S * nblk, generator=g, device=DEV).to(torch.int32).view(S, nblk).contiguous()
    lens = torch.randint(1, L + 1, (S,), generator=g, device=DEV, dtype=torch.int32)
    p = pt.DecodeProblem(q, kc, vc, tables, lens, D ** -0.5, B, L)
    ours = pt.run_decode(p)
    theirs = torch.empty_like(ours)
    ref_attention.single_query_cached_kv_attention(theirs, q, kc, vc, p.scale, tables, lens, B, L)
    torch.cuda.synchronize()
    pt.assert_decode_close(ours, theirs, torch.float16, "config 2 vs reference kernel")


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("tokens,heads,head_size,block_size,nb", [(3, 2, 16, 8, 2), (256, 40, 128, 16, 64),
                                                                  (2560, 8, 96, 32, 96), (77, 5, 80, 8, 40)])
def test_reshape_and_gather_bit_exact_vs_reference_kernel(ref_cache, dtype, tokens, heads, head_size, block_size, nb):
    from vllm_b200 import cache_ops
    x = pt.pack_factor(dtype)
    g = torch.Generator(device=DEV).manual_seed(tokens)
    qkv = torch.empty((tokens, 3, heads, head_size), dtype=dtype, device=DEV).normal_(generator=g)
    kc = torch.empty((nb, heads, head_size // x, block_size, x), dtype=dtype, device=DEV).normal_(generator=g)
    vc = torch.empty((nb, heads, head_size, block_size), dtype=dtype, device=DEV).normal_(generator=g)
    slots = torch.randperm(nb * block_size, generator=g, device=DEV)[:tokens].to(torch.int32)
    if tokens % 2 == 0:       # prefill-like: consecutive slots from a random start (runs cross block boundaries)
        slots = (torch.arange(tokens, device=DEV, dtype=torch.int32) + 5) % (nb * block_size)
    kc_r, vc_r = kc.clone(), vc.clone()
    ref_cache.reshape_and_cache(qkv[:, 1], qkv[:, 2], kc_r, vc_r, slots)
    cache_ops.reshape_and_cache(qkv[:, 1], qkv[:, 2], kc, vc, slots)
    assert torch.equal(pt.bits(kc), pt.bits(kc_r)) and torch.equal(pt.bits(vc), pt.bits(vc_r))
    a, b = torch.zeros_like(qkv), torch.zeros_like(qkv)
    ref_cache.gather_cached_kv(a[:, 1], a[:, 2], kc_r, vc_r, slots)
    cache_ops.gather_cached_kv(b[:, 1], b[:, 2], kc, vc, slots)
    assert torch.equal(pt.bits(a), pt.bits(b))


@pytest.mark.parametrize("dtype", [torch.float16, torch.float32])
def test_copy_blocks_bit_exact_vs_reference_kernel(ref_cache, dtype):
    """The reference test's shape (tests/kernels/test_cache.py:149-154) and a parallel-sampling fork."""
    from vllm_b200 import cache_ops
    for layers, heads, head_size, block_size, nb, pairs in ((7, 17, 16, 8, 1024, 23), (4, 40, 128, 16, 400, 192)):
        x = pt.pack_factor(dtype)
        kcs = [torch.randn((nb, heads, head_size // x, block_size, x), device=DEV).to(dtype) for _ in range(layers)]
        vcs = [torch.randn((nb, heads, head_size, block_size), device=DEV).to(dtype) for _ in range(layers)]
        rnd = random.Random(pairs)
        ids = rnd.sample(range(nb), pairs // 3 + 1 + pairs)
        srcs, dsts = ids[:pairs // 3 + 1], ids[pairs // 3 + 1:]
        mapping = {}
        for i, d in enumerate(dsts):
            mapping.setdefault(srcs[i % len(srcs)], []).append(d)
        kr, vr = [t.clone() for t in kcs], [t.clone() for t in vcs]
        ref_cache.copy_blocks(kr, vr, mapping)
        cache_ops.copy_blocks(kcs, vcs, mapping)
        for l in range(layers):
            assert torch.equal(pt.bits(kcs[l]), pt.bits(kr[l])) and torch.equal(pt.bits(vcs[l]), pt.bits(vr[l]))


def test_swap_blocks_bit_exact_vs_reference_implementation(ref_cache):
    """Pins swap_blocks (untested in the reference) on the reference's own implementation, all three directions."""
    from vllm_b200 import cache_ops
    nb, heads, d, b = 48, 5, 128, 16
    gpu = torch.randn((nb, heads, d, b), device=DEV).half()
    cpu = torch.randn((nb, heads, d, b)).half().pin_memory()
    out_map = {3: 7, 4: 8, 5: 9, 20: 0, 47: 46, 10: 11, 30: 2, 33: 31, 35: 21, 37: 23}
    for engine in (0, 1, 2):             # automatic, copy engines, zero-copy kernel
        cpu_r, cpu_o = cpu.clone().pin_memory(), cpu.clone().pin_memory()
        ref_cache.swap_blocks(gpu, cpu_r, out_map)
        cache_ops.swap_blocks(gpu, cpu_o, out_map, engine=engine)
        torch.cuda.synchronize()
        assert torch.equal(pt.bits(cpu_r), pt.bits(cpu_o)), f"gpu->cpu engine {engine}"
        gpu_r, gpu_o = gpu.clone(), gpu.clone()
        in_map = {7: 1, 8: 2, 9: 3, 0: 40, 46: 47, 11: 13, 21: 15, 23: 17}
        ref_cache.swap_blocks(cpu, gpu_r, in_map)
        cache_ops.swap_blocks(cpu, gpu_o, in_map, engine=engine)
        assert torch.equal(pt.bits(gpu_r), pt.bits(gpu_o)), f"cpu->gpu engine {engine}"
        d_r, d_o = torch.zeros_like(gpu), torch.zeros_like(gpu)
        ref_cache.swap_blocks(gpu, d_r, {1: 2, 2: 1, 40: 41})
        cache_ops.swap_blocks(gpu, d_o, {1: 2, 2: 1, 40: 41}, engine=engine)
        assert torch.equal(pt.bits(d_r), pt.bits(d_o)), f"gpu->gpu engine {engine}"


@pytest.fixture(scope="module")
def ref_pos_encoding():
    return _load("ref_pos_encoding_ops")


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
@pytest.mark.parametrize("head_size,rot", [(32, 32), (64, 64), (80, 80), (96, 96), (128, 128), (160, 160), (192, 192),
                                           (256, 256), (96, 24)])
def test_rotary_embedding_vs_reference_kernel(ref_pos_encoding, dtype, head_size, rot):
    """The reference test's shape (tests/kernels/test_pos_encoding.py:134-146): 2145 tokens, 5 heads,
    max_position 8192.  16-bit types: bit-identical (both round every product and sum to T).  fp32: the
    reference kernel's x*cos - y*sin is contracted into a fused multiply-add by nvcc, ours keeps the two
    roundings of the PyTorch formula, so the last bit may differ."""
    from oracle import paged_oracle
    from vllm_b200 import pos_encoding_ops
    tokens, heads = 2145, 5
    g = torch.Generator(device=DEV).manual_seed(head_size)
    qkv = torch.empty((tokens, 3, heads * head_size), dtype=dtype, device=DEV).normal_(generator=g)
    pos = torch.randint(0, 8192, (tokens,), generator=g, device=DEV)
    cache = paged_oracle.rope_cos_sin_cache(rot, 8192, dtype).to(DEV)
    a, b = qkv.clone(), qkv.clone()
    ref_pos_encoding.rotary_embedding_neox(pos, a[:, 0], a[:, 1], head_size, cache)
    pos_encoding_ops.rotary_embedding_neox(pos, b[:, 0], b[:, 1], head_size, cache)
    if dtype == torch.float32:
        torch.testing.assert_close(b, a, atol=1e-6, rtol=1e-6)
        assert torch.allclose(b, a, atol=1e-3, rtol=1e-5)          # the reference test's criterion (:130-131)
    else:
        assert torch.equal(pt.bits(a), pt.bits(b))
