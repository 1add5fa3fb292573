"""PagedAttention / PagedAttentionWithRoPE layer mirrors (vllm/model_executor/layers/attention.py:16-217):
a mixed batch -- two prompts, five generation tokens, padding to a multiple of 8 -- through `forward`,
against the oracle composition of the same steps (rope -> prompt attention -> KV write -> paged decode)."""
import pytest
import torch

import pa_testing as pt
from oracle import paged_oracle


def _make_batch(dtype, num_heads, head_size, block_size, seed):
    g = torch.Generator().manual_seed(seed)
    prompt_lens = [37, 150]
    gen_ctx = [1, 16, 17, 90, 200]                # context length of each generation token AFTER its own KV write
    num_blocks = 96
    x = 16 // torch.empty((), dtype=dtype).element_size()
    n_prompt, n_gen = sum(prompt_lens), len(gen_ctx)
    n_valid = n_prompt + n_gen
    n_tok = (n_valid + 7) // 8 * 8
    qkv = torch.randn((n_tok, 3 * num_heads * head_size), generator=g).to(dtype)
    key_cache = torch.randn((num_blocks, num_heads, head_size // x, block_size, x), generator=g).to(dtype)
    value_cache = torch.randn((num_blocks, num_heads, head_size, block_size), generator=g).to(dtype)
    perm = torch.randperm(num_blocks, generator=g).tolist()
    # prompts get fresh blocks; generation sequences own distinct blocks
    slots, tables, positions = [], [], []
    for n in prompt_lens:
        blocks = [perm.pop() for _ in range((n + block_size - 1) // block_size)]
        slots += [blocks[i // block_size] * block_size + i % block_size for i in range(n)]
        positions += list(range(n))
    max_blocks = (max(gen_ctx) + block_size - 1) // block_size
    for c in gen_ctx:
        blocks = [perm.pop() for _ in range((c + block_size - 1) // block_size)]
        tables.append(blocks + [0] * (max_blocks - len(blocks)))
        slots.append(blocks[(c - 1) // block_size] * block_size + (c - 1) % block_size)
        positions.append(c - 1)
    positions += [0] * (n_tok - n_valid)
    return dict(prompt_lens=prompt_lens, gen_ctx=gen_ctx, qkv=qkv, key_cache=key_cache, value_cache=value_cache,
                slot_mapping=torch.tensor(slots, dtype=torch.int32), block_tables=torch.tensor(tables, dtype=torch.int32),
                context_lens=torch.tensor(gen_ctx, dtype=torch.int32), positions=torch.tensor(positions, dtype=torch.int64),
                n_prompt=n_prompt, n_valid=n_valid, n_tok=n_tok)


def _oracle_forward(b, num_heads, head_size, scale, cos_sin_cache):
    """The reference layer's steps on CPU copies, fp32 attention maths on the dtype-rounded tensors."""
    hid = num_heads * head_size
    qkv = b["qkv"].clone()
    q, k, v = qkv[:, :hid], qkv[:, hid:2 * hid], qkv[:, 2 * hid:]
    if cos_sin_cache is not None:
        paged_oracle.rotary_embedding_neox(b["positions"], q, k, head_size, cos_sin_cache)
    q3, k3, v3 = (t.reshape(-1, num_heads, head_size) for t in (q, k, v))
    kc, vc = b["key_cache"].clone(), b["value_cache"].clone()
    out = torch.zeros((b["n_tok"], num_heads, head_size), dtype=torch.float32)
    n_p, n_v = b["n_prompt"], b["n_valid"]
    out[:n_p] = paged_oracle.prefill_attention_fp32(b["prompt_lens"], q3[:n_p], k3[:n_p], v3[:n_p], scale)
    paged_oracle.reshape_and_cache(k3[:n_v], v3[:n_v], kc, vc, b["slot_mapping"])
    out[n_p:n_v] = paged_oracle.decode_attention_fp32(q3[n_p:n_v], kc, vc, b["block_tables"], b["context_lens"], scale)
    return out, q, k, kc, vc


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("head_size", [128, 80])
@pytest.mark.parametrize("rope,fuse", [(False, False), (True, False), (True, True)])
def test_layer_forward_mixed_batch(dtype, head_size, rope, fuse):
    import vllm_b200 as vb
    num_heads, block_size = 4, 16
    scale = head_size ** -0.5
    b = _make_batch(dtype, num_heads, head_size, block_size, seed=head_size + int(rope))
    dev = "cuda"
    prev = torch.get_default_dtype()
    torch.set_default_dtype(dtype)                       # the reference reads the cache dtype from the default
    try:
        layer = (vb.PagedAttentionWithRoPE(num_heads, head_size, scale, rotary_dim=head_size, max_position=512,
                                           fuse_rope_cache=fuse) if rope else vb.PagedAttention(num_heads, head_size, scale))
    finally:
        torch.set_default_dtype(prev)
    layer = layer.to(dev)
    meta = vb.InputMetadata(seq_groups=[], seq_data={}, prompt_lens=b["prompt_lens"],
                            slot_mapping=b["slot_mapping"].to(dev), context_lens=b["context_lens"].to(dev),
                            max_context_len=max(b["gen_ctx"]), block_tables=b["block_tables"].to(dev))
    assert meta.num_prompt_tokens == b["n_prompt"] and meta.num_valid_tokens == b["n_valid"]
    hid = num_heads * head_size
    qkv = b["qkv"].to(dev)
    q, k, v = qkv[:, :hid], qkv[:, hid:2 * hid], qkv[:, 2 * hid:]
    kc, vc = b["key_cache"].to(dev), b["value_cache"].to(dev)
    ev = torch.cuda.Event()
    ev.record()
    if rope:
        out = layer(b["positions"].to(dev), q, k, v, kc, vc, meta, ev)
    else:
        out = layer(q, k, v, kc, vc, meta, ev)
    torch.cuda.synchronize()
    want, q_ref, k_ref, kc_ref, vc_ref = _oracle_forward(b, num_heads, head_size, scale,
                                                         layer.cos_sin_cache.cpu() if rope else None)
    # movers and rope are bit-exact
    assert torch.equal(pt.bits(kc.cpu()), pt.bits(kc_ref)) and torch.equal(pt.bits(vc.cpu()), pt.bits(vc_ref))
    if rope:
        assert torch.equal(pt.bits(q.cpu().contiguous()), pt.bits(q_ref.contiguous()))
        assert torch.equal(pt.bits(k.cpu().contiguous()), pt.bits(k_ref.contiguous()))
    got = out.view(-1, num_heads, head_size)[:b["n_valid"]].cpu()
    pt.assert_decode_close(got, want[:b["n_valid"]], dtype, f"layer {dtype} d{head_size} rope={rope} fuse={fuse}")


def test_layer_rejects_unsupported_head_size():
    import vllm_b200 as vb
    with pytest.raises(ValueError):
        vb.PagedAttention(4, 48, 1.0)


def test_input_metadata_mirrors_reference_fields():
    import vllm_b200 as vb
    meta = vb.InputMetadata([], {}, [3, 5], torch.zeros(10, dtype=torch.int32), torch.zeros(2, dtype=torch.int32), 7,
                            torch.zeros((2, 4), dtype=torch.int32))
    assert (meta.num_prompts, meta.num_prompt_tokens, meta.num_generation_tokens, meta.num_valid_tokens,
            meta.max_num_blocks_per_seq) == (2, 8, 2, 10, 4)
    assert meta.attn_bias.seqlens == [3, 5]
    with pytest.raises(AssertionError):
        vb.InputMetadata([], {}, [], torch.zeros(1, dtype=torch.int32), torch.zeros(2, dtype=torch.int32), 7,
                         torch.zeros((1, 4), dtype=torch.int32))
