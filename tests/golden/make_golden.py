"""Generates tests/golden/*.npz by running the REFERENCE's own test references.

Run in the build container, where /root/reference exists:

    python tests/golden/make_golden.py

The reference test modules cannot be imported (top-level xformers / compiled
extension imports), so the pure-PyTorch reference functions are cut out of their
source with `ast` and executed unmodified:

  tests/kernels/test_attention.py  ref_masked_attention (:14-27),
                                   ref_single_query_cached_kv_attention (:30-66)
                                   ref_multi_query_kv_attention (:69-101), with device='cuda' of its
                                   mask (:87) mapped to the CPU while it runs
                                   ref_multi_query_cached_kv_attention (:104-156), likewise (:127)
  tests/kernels/test_cache.py      the reference loops inside run_copy_blocks (:51-56),
                                   run_reshape_and_cache (:95-100), run_gather_cached_kv (:138-143)
  tests/kernels/test_pos_encoding.py  rotate_half, apply_rotary_pos_emb, RefRotaryEmbeddingNeox (:12-77)

Nothing from the reference is copied into this repository: only the outputs of
those functions on the seeded inputs of cases.py are stored.
"""
from __future__ import annotations

import ast
import json
import os
import sys
from typing import List, Optional

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import cases  # noqa: E402

REF = os.environ.get("PA_REFERENCE_ROOT", "/root/reference")


def _functions(path: str, names) -> dict:
    tree = ast.parse(open(path).read())
    picked = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in names]
    assert len(picked) == len(names), f"missing functions in {path}"
    ns = {"torch": torch, "Optional": Optional, "List": List}
    exec(compile(ast.Module(body=picked, type_ignores=[]), path, "exec"), ns)
    return ns


def _reference_loop(path: str, func: str):
    """The first top-level `for` after the cache_ops call inside run_* -- the reference computation."""
    tree = ast.parse(open(path).read())
    fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == func)
    seen_call = False
    for stmt in fn.body:
        if isinstance(stmt, ast.Expr) and isinstance(stmt.value, ast.Call) and "cache_ops" in ast.unparse(stmt.value.func):
            seen_call = True
        elif seen_call and isinstance(stmt, ast.For):
            return compile(ast.Module(body=[stmt], type_ignores=[]), f"{path}:{func}", "exec")
    raise AssertionError(f"reference loop not found in {func}")


def make_attention() -> None:
    ns = _functions(os.path.join(REF, "tests/kernels/test_attention.py"),
                    {"ref_masked_attention", "ref_single_query_cached_kv_attention"})
    ref = ns["ref_single_query_cached_kv_attention"]
    out, meta = {}, {}
    for cid in cases.attention_case_ids():
        c = cases.attention_case(cid)
        o = torch.empty_like(c["query"])
        with torch.inference_mode():
            ref(o, c["query"], c["key_cache"], c["value_cache"], c["block_tables"], c["context_lens"])
        out[cid] = cases.to_numpy(o)
        meta[cid] = cases.digest(c["query"], c["key_cache"], c["value_cache"], c["block_tables"], c["context_lens"])
        print(cid, float(o.float().abs().max()))
    np.savez_compressed(os.path.join(HERE, "attention_golden.npz"), **out)
    json.dump(meta, open(os.path.join(HERE, "attention_inputs.sha256.json"), "w"), indent=1, sort_keys=True)


def make_attention_large() -> None:
    """BASELINE.json configs[0] (the reference test's own shape and procedure) and the long-context cases."""
    ns = _functions(os.path.join(REF, "tests/kernels/test_attention.py"),
                    {"ref_masked_attention", "ref_single_query_cached_kv_attention"})
    ref = ns["ref_single_query_cached_kv_attention"]
    out, meta = {}, {}
    for cid in cases.attention_large_case_ids():
        c = cases.attention_large_case(cid)
        o = torch.empty_like(c["query"])
        with torch.inference_mode():
            ref(o, c["query"], c["key_cache"], c["value_cache"], c["block_tables"], c["context_lens"])
        out[cid] = cases.to_numpy(o)
        meta[cid] = cases.digest(c["query"], c["key_cache"], c["value_cache"], c["block_tables"], c["context_lens"])
        print(cid, float(o.float().abs().max()))
    np.savez_compressed(os.path.join(HERE, "attention_golden_large.npz"), **out)
    json.dump(meta, open(os.path.join(HERE, "attention_large_inputs.sha256.json"), "w"), indent=1, sort_keys=True)


def make_cache() -> None:
    path = os.path.join(REF, "tests/kernels/test_cache.py")
    loop_copy = _reference_loop(path, "run_copy_blocks")
    loop_reshape = _reference_loop(path, "run_reshape_and_cache")
    loop_gather = _reference_loop(path, "run_gather_cached_kv")
    out, meta = {}, {}
    for cid in cases.cache_case_ids():
        c = cases.cache_case(cid)
        x = 16 // c["key"].element_size()
        num_tokens = c["key"].shape[0]
        # --- reshape_and_cache: reference writes into cloned_key_cache / cloned_value_cache
        ns = {"torch": torch, "num_tokens": num_tokens, "num_heads": c["num_heads"], "head_size": c["head_size"],
              "block_size": c["block_size"], "x": x, "key": c["key"], "value": c["value"],
              "slot_mapping": c["slot_mapping"], "cloned_key_cache": c["key_caches"][0].clone(),
              "cloned_value_cache": c["value_caches"][0].clone()}
        with torch.inference_mode():
            exec(loop_reshape, ns)
        out[f"{cid}/reshape_key_cache"] = cases.to_numpy(ns["cloned_key_cache"])
        out[f"{cid}/reshape_value_cache"] = cases.to_numpy(ns["cloned_value_cache"])
        # --- gather_cached_kv: reference writes into cloned_key / cloned_value
        qkv_clone = torch.zeros(num_tokens, 3, c["num_heads"], c["head_size"], dtype=c["dtype"])
        ns = {"torch": torch, "num_tokens": num_tokens, "num_heads": c["num_heads"], "head_size": c["head_size"],
              "block_size": c["block_size"], "x": x, "slot_mapping": c["slot_mapping"],
              "key_cache": c["key_caches"][1], "value_cache": c["value_caches"][1],
              "cloned_key": qkv_clone[:, 1], "cloned_value": qkv_clone[:, 2]}
        with torch.inference_mode():
            exec(loop_gather, ns)
        out[f"{cid}/gather_key"] = cases.to_numpy(ns["cloned_key"])
        out[f"{cid}/gather_value"] = cases.to_numpy(ns["cloned_value"])
        # --- copy_blocks: reference mutates cloned_key_caches / cloned_value_caches
        ns = {"torch": torch, "block_mapping": c["copy_map"],
              "key_caches": c["key_caches"], "value_caches": c["value_caches"],
              "cloned_key_caches": [t.clone() for t in c["key_caches"]],
              "cloned_value_caches": [t.clone() for t in c["value_caches"]]}
        with torch.inference_mode():
            exec(loop_copy, ns)
        for l, (k, v) in enumerate(zip(ns["cloned_key_caches"], ns["cloned_value_caches"])):
            out[f"{cid}/copy_key_cache{l}"] = cases.to_numpy(k)
            out[f"{cid}/copy_value_cache{l}"] = cases.to_numpy(v)
        meta[cid] = cases.digest(c["key"], c["value"], c["slot_mapping"], *c["key_caches"], *c["value_caches"])
        print(cid, "ok")
    # bit-exact operators: a SHA-256 per output pins every case; full arrays are kept
    # only for the smallest shape (for debugging a mismatch)
    import hashlib
    hashes = {k: hashlib.sha256(v.tobytes()).hexdigest() for k, v in out.items()}
    small = {k: v for k, v in out.items() if k.split("/")[0].endswith("-b8-d16")}
    np.savez_compressed(os.path.join(HERE, "cache_golden_small.npz"), **small)
    json.dump(hashes, open(os.path.join(HERE, "cache_outputs.sha256.json"), "w"), indent=1, sort_keys=True)
    json.dump(meta, open(os.path.join(HERE, "cache_inputs.sha256.json"), "w"), indent=1, sort_keys=True)


def make_rope() -> None:
    """RefRotaryEmbeddingNeox (tests/kernels/test_pos_encoding.py:12-77), run exactly as the reference test
    runs it (:118-130): module built in fp32, moved to the dtype, called on [N, H, D] views."""
    import torch.nn as nn
    import torch.nn.functional as F
    from typing import Tuple
    path = os.path.join(REF, "tests/kernels/test_pos_encoding.py")
    tree = ast.parse(open(path).read())
    names = {"rotate_half", "apply_rotary_pos_emb", "RefRotaryEmbeddingNeox"}
    picked = [n for n in tree.body if isinstance(n, (ast.FunctionDef, ast.ClassDef)) and n.name in names]
    assert len(picked) == 3
    ns = {"torch": torch, "nn": nn, "F": F, "Tuple": Tuple}
    exec(compile(ast.Module(body=picked, type_ignores=[]), path, "exec"), ns)
    import hashlib
    out, meta = {}, {}
    for cid in cases.rope_case_ids():
        c = cases.rope_case(cid)
        n, h, d = c["qkv"].shape[0], c["num_heads"], c["head_size"]
        ref = ns["RefRotaryEmbeddingNeox"](dim=c["rot_dim"], max_position_embeddings=c["max_position"]).to(dtype=c["dtype"])
        with torch.inference_mode():
            q, k = ref(c["positions"], c["qkv"][:, 0].reshape(n, h, d), c["qkv"][:, 1].reshape(n, h, d))
        out[f"{cid}/query"] = cases.to_numpy(q.reshape(n, h * d))
        out[f"{cid}/key"] = cases.to_numpy(k.reshape(n, h * d))
        meta[cid] = cases.digest(c["qkv"], c["positions"])
        print(cid, "ok")
    hashes = {k: hashlib.sha256(v.tobytes()).hexdigest() for k, v in out.items()}
    small = {k: v for k, v in out.items() if "-d32-" in k}
    np.savez_compressed(os.path.join(HERE, "rope_golden_small.npz"), **small)
    json.dump(hashes, open(os.path.join(HERE, "rope_outputs.sha256.json"), "w"), indent=1, sort_keys=True)
    json.dump(meta, open(os.path.join(HERE, "rope_inputs.sha256.json"), "w"), indent=1, sort_keys=True)


def make_prefill() -> None:
    ns = _functions(os.path.join(REF, "tests/kernels/test_attention.py"),
                    {"ref_masked_attention", "ref_multi_query_kv_attention"})
    ref = ns["ref_multi_query_kv_attention"]
    # the reference moves its mask to 'cuda' (:87); there is no GPU here, so .to() is told to stay on the CPU
    real_to = torch.Tensor.to

    def cpu_to(self, *args, **kwargs):
        if kwargs.get("device") == "cuda":
            kwargs["device"] = "cpu"
        return real_to(self, *args, **kwargs)

    out, meta = {}, {}
    torch.Tensor.to = cpu_to
    try:
        for cid in cases.prefill_case_ids():
            c = cases.prefill_case(cid)
            cu = [0]
            for n in c["seq_lens"]:
                cu.append(cu[-1] + n)
            with torch.inference_mode():
                o = ref(cu, c["query"], c["key"], c["value"], c["dtype"])
            out[cid] = cases.to_numpy(o)
            meta[cid] = cases.digest(c["qkv"])
            print(cid, float(o.float().abs().max()))
    finally:
        torch.Tensor.to = real_to
    np.savez_compressed(os.path.join(HERE, "prefill_golden.npz"), **out)
    json.dump(meta, open(os.path.join(HERE, "prefill_inputs.sha256.json"), "w"), indent=1, sort_keys=True)


def make_cached_prefill() -> None:
    """ref_multi_query_cached_kv_attention (tests/kernels/test_attention.py:104-156), unmodified; its
    mask's device='cuda' (:127) is mapped to the CPU while it runs."""
    ns = _functions(os.path.join(REF, "tests/kernels/test_attention.py"),
                    {"ref_masked_attention", "ref_multi_query_cached_kv_attention"})
    ref = ns["ref_multi_query_cached_kv_attention"]
    real_to = torch.Tensor.to

    def cpu_to(self, *args, **kwargs):
        if kwargs.get("device") == "cuda":
            kwargs["device"] = "cpu"
        return real_to(self, *args, **kwargs)

    out, meta = {}, {}
    torch.Tensor.to = cpu_to
    try:
        for cid in cases.cached_prefill_case_ids():
            c = cases.cached_prefill_case(cid)
            with torch.inference_mode():
                o = ref(c["cu_query_lens"], c["query"], c["key_cache"], c["value_cache"], c["block_tables"],
                        c["context_lens"], c["dtype"])
            out[cid] = cases.to_numpy(o)
            meta[cid] = cases.digest(c["qkv"], c["key_cache"], c["value_cache"], c["block_tables"], c["context_lens"])
            print(cid, float(o.float().abs().max()))
    finally:
        torch.Tensor.to = real_to
    np.savez_compressed(os.path.join(HERE, "cached_prefill_golden.npz"), **out)
    json.dump(meta, open(os.path.join(HERE, "cached_prefill_inputs.sha256.json"), "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    torch.manual_seed(0)
    which = sys.argv[1:] or ["attention", "attention_large", "cache", "rope", "prefill", "cached_prefill"]
    if "attention" in which:
        make_attention()
    if "attention_large" in which:
        make_attention_large()
    if "cache" in which:
        make_cache()
    if "rope" in which:
        make_rope()
    if "prefill" in which:
        make_prefill()
    if "cached_prefill" in which:
        make_cached_prefill()
