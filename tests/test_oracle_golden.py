"""Pins the oracle: our restatements (oracle/paged_oracle.py, oracle/paged_oracle.c)
against outputs of the reference's own test references (tests/golden/, produced by
make_golden.py from /root/reference).  CPU only."""
import hashlib
import json
import os

import numpy as np
import pytest
import torch

import cases
from oracle import c_oracle, paged_oracle

GOLD = os.path.join(os.path.dirname(__file__), "golden")
ATT = np.load(os.path.join(GOLD, "attention_golden.npz"))
ATT_SHA = json.load(open(os.path.join(GOLD, "attention_inputs.sha256.json")))
CACHE_SHA_IN = json.load(open(os.path.join(GOLD, "cache_inputs.sha256.json")))
CACHE_SHA_OUT = json.load(open(os.path.join(GOLD, "cache_outputs.sha256.json")))
CACHE_SMALL = np.load(os.path.join(GOLD, "cache_golden_small.npz"))

# A re-run of the same in-dtype arithmetic on another CPU may round differently in the
# last place; the fp32 truth vs the in-dtype reference differs by the dtype's resolution
# (the north-star tolerances: 1e-3 fp16, 1e-2 bf16).
SAME_MATH_TOL = {torch.float16: 2e-3, torch.bfloat16: 1.6e-2, torch.float32: 1e-5}
TRUTH_TOL = {torch.float16: 1e-3, torch.bfloat16: 1e-2, torch.float32: 1e-5}


def _sha(t):
    return hashlib.sha256(cases.to_numpy(t).tobytes()).hexdigest()


@pytest.mark.parametrize("cid", cases.attention_case_ids())
def test_attention_inputs_are_the_golden_inputs(cid):
    c = cases.attention_case(cid)
    assert cases.digest(c["query"], c["key_cache"], c["value_cache"], c["block_tables"], c["context_lens"]) == ATT_SHA[cid]


@pytest.mark.parametrize("cid", cases.attention_case_ids())
def test_loop_oracle_matches_reference_output(cid):
    c = cases.attention_case(cid)
    gold = cases.from_numpy(ATT[cid], c["dtype"])
    out = torch.empty_like(gold)
    paged_oracle.decode_attention_loop(out, c["query"], c["key_cache"], c["value_cache"], c["block_tables"],
                                       c["context_lens"])
    tol = SAME_MATH_TOL[c["dtype"]]
    torch.testing.assert_close(out.float(), gold.float(), atol=tol * 1e-3, rtol=tol)


@pytest.mark.parametrize("cid", cases.attention_case_ids())
def test_fp32_and_c_oracles_match_reference_output(cid):
    c = cases.attention_case(cid)
    gold = cases.from_numpy(ATT[cid], c["dtype"]).float()
    truth = paged_oracle.decode_attention_fp32(c["query"], c["key_cache"], c["value_cache"], c["block_tables"],
                                               c["context_lens"], c["scale"])
    tol = TRUTH_TOL[c["dtype"]]
    scale = float(gold.abs().max())
    torch.testing.assert_close(truth, gold, atol=tol * scale, rtol=tol)
    cout = torch.empty(c["query"].shape, dtype=c["dtype"])
    c_oracle.decode_attention(cout, c["query"], c["key_cache"], c["value_cache"], c["scale"], c["block_tables"],
                              c["context_lens"])
    # C oracle (double accumulation, rounded once to the dtype) against the fp32 truth rounded the same way
    # (one unit in the last place of the dtype; 1.2e-7 = two fp16 subnormal steps)
    torch.testing.assert_close(cout.float(), truth.to(c["dtype"]).float(), atol=2e-6 * scale + 1.2e-7,
                               rtol={torch.float16: 2e-3, torch.bfloat16: 1.6e-2, torch.float32: 2e-5}[c["dtype"]])


ATT_LARGE = np.load(os.path.join(GOLD, "attention_golden_large.npz"))
ATT_LARGE_SHA = json.load(open(os.path.join(GOLD, "attention_large_inputs.sha256.json")))


@pytest.mark.parametrize("cid", cases.attention_large_case_ids())
def test_large_cases_config1_and_long_contexts(cid):
    """BASELINE.json configs[0] (7 x 40 x 128, fp32, contexts <= 4096, the reference test's own procedure) and
    4096-token contexts per dtype: both oracles against outputs of the reference's function at those sizes."""
    c = cases.attention_large_case(cid)
    assert cases.digest(c["query"], c["key_cache"], c["value_cache"], c["block_tables"], c["context_lens"]) == ATT_LARGE_SHA[cid]
    if cid.startswith("cfg1"):
        assert c["query"].shape == (7, 40, 128) and c["key_cache"].shape == (1024, 40, 32, 16, 4)
        assert c["dtype"] == torch.float32 and int(c["context_lens"].max()) <= 4096
    gold = cases.from_numpy(ATT_LARGE[cid], c["dtype"]).float()
    scale = float(gold.abs().max())
    tol = TRUTH_TOL[c["dtype"]]
    cout = torch.empty(c["query"].shape, dtype=c["dtype"])
    c_oracle.decode_attention(cout, c["query"].contiguous(), c["key_cache"], c["value_cache"], c["scale"],
                              c["block_tables"], c["context_lens"])
    sub = 1.2e-7 if c["dtype"] == torch.float16 else 0.0     # two fp16 subnormal steps (outputs ~4e-5 are subnormal)
    torch.testing.assert_close(cout.float(), gold, atol=tol * scale + sub, rtol=tol)
    # by the letter of the reference test (test_attention.py:224) and of the north star
    assert torch.allclose(cout.float(), gold, atol=1e-3, rtol=1e-5) or not cid.endswith("-ref")
    flat = {torch.float16: 1e-3, torch.bfloat16: 1e-2, torch.float32: 1e-3}[c["dtype"]]
    torch.testing.assert_close(cout.float(), gold, atol=flat, rtol=flat)
    truth = paged_oracle.decode_attention_fp32(c["query"], c["key_cache"], c["value_cache"], c["block_tables"],
                                               c["context_lens"], c["scale"])
    torch.testing.assert_close(truth, gold, atol=tol * scale + sub, rtol=tol)


def test_flat_north_star_tolerance_holds_for_the_oracle_on_unit_scale_goldens():
    """The unit-scale goldens (outputs of the reference function in the input dtype) against the C oracle under
    the flat tolerance the north star states -- the same check the GPU tests apply to the kernels."""
    for cid in cases.attention_case_ids():
        if not cid.endswith("-unit"):
            continue
        c = cases.attention_case(cid)
        gold = cases.from_numpy(ATT[cid], c["dtype"]).float()
        cout = torch.empty(c["query"].shape, dtype=c["dtype"])
        c_oracle.decode_attention(cout, c["query"].contiguous(), c["key_cache"], c["value_cache"], c["scale"],
                                  c["block_tables"], c["context_lens"])
        flat = {torch.float16: 1e-3, torch.bfloat16: 1e-2, torch.float32: 1e-3}[c["dtype"]]
        torch.testing.assert_close(cout.float(), gold, atol=flat, rtol=flat, msg=lambda m: f"{cid}: {m}")


def test_zero_context_is_zero_output():
    c = cases.attention_case("f32-b16-d64-ref")
    lens = c["context_lens"].clone()
    lens[2] = 0
    out = torch.full(c["query"].shape, 7.0)
    paged_oracle.decode_attention_loop(out, c["query"], c["key_cache"], c["value_cache"], c["block_tables"], lens)
    assert torch.count_nonzero(out[2]) == 0
    cout = torch.full(c["query"].shape, 7.0)
    c_oracle.decode_attention(cout, c["query"].contiguous(), c["key_cache"], c["value_cache"], c["scale"],
                              c["block_tables"], lens)
    assert torch.count_nonzero(cout[2]) == 0
    torch.testing.assert_close(cout, out, atol=1e-9, rtol=2e-5)


@pytest.mark.parametrize("cid", cases.cache_case_ids())
@pytest.mark.parametrize("impl", ["python", "c"])
def test_cache_movers_match_reference_output(cid, impl):
    c = cases.cache_case(cid)
    assert cases.digest(c["key"], c["value"], c["slot_mapping"], *c["key_caches"], *c["value_caches"]) == CACHE_SHA_IN[cid]
    mod = paged_oracle if impl == "python" else c_oracle
    got = {}
    # reshape_and_cache on layer 0
    kc, vc = c["key_caches"][0].clone(), c["value_caches"][0].clone()
    mod.reshape_and_cache(c["key"], c["value"], kc, vc, c["slot_mapping"])
    got["reshape_key_cache"], got["reshape_value_cache"] = kc, vc
    # gather_cached_kv from layer 1 into a zeroed fused qkv
    qkv = torch.zeros(c["key"].shape[0], 3, c["num_heads"], c["head_size"], dtype=c["dtype"])
    mod.gather_cached_kv(qkv[:, 1], qkv[:, 2], c["key_caches"][1], c["value_caches"][1], c["slot_mapping"])
    got["gather_key"], got["gather_value"] = qkv[:, 1], qkv[:, 2]
    # copy_blocks over all layers
    kcs, vcs = [t.clone() for t in c["key_caches"]], [t.clone() for t in c["value_caches"]]
    mod.copy_blocks(kcs, vcs, c["copy_map"])
    for l, (k, v) in enumerate(zip(kcs, vcs)):
        got[f"copy_key_cache{l}"], got[f"copy_value_cache{l}"] = k, v
    for name, t in got.items():
        assert _sha(t) == CACHE_SHA_OUT[f"{cid}/{name}"], f"{cid}/{name} differs from the reference output"
        if f"{cid}/{name}" in CACHE_SMALL.files:
            assert np.array_equal(cases.to_numpy(t), CACHE_SMALL[f"{cid}/{name}"])


@pytest.mark.parametrize("cid", cases.cache_case_ids()[:4])
def test_swap_blocks_oracles_agree(cid):
    """swap_blocks has no test in the reference ("parity unpinned"): pin both restatements on
    plain block assignment, and on each other."""
    c = cases.cache_case(cid)
    src = c["key_caches"][0]
    want = c["key_caches"][1].clone()
    for s, d in c["swap_map"].items():
        want[d] = src[s]
    a, b = c["key_caches"][1].clone(), c["key_caches"][1].clone()
    paged_oracle.swap_blocks(src, a, c["swap_map"])
    c_oracle.swap_blocks(src, b, c["swap_map"])
    assert torch.equal(a.view(torch.uint8), want.view(torch.uint8))
    assert torch.equal(b.view(torch.uint8), want.view(torch.uint8))


def test_algorithmic_bytes_match_baseline_md():
    # 4.299 GB
    n = paged_oracle.decode_algorithmic_bytes([1024] * 256, 32, 128, 16, 2)
    assert n == 4294967296 + 4194304 + 256 * (64 * 4 + 4)
