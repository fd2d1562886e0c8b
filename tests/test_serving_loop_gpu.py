"""Caller-side integration: what a vLLM attention backend does with this library —
prefill over a paged cache through flash_attn_varlen_func(block_table=...), then a decode loop in which
every step appends the new token's K/V to the paged cache and attends, with the decode step captured
ONCE in a CUDA graph and replayed.  Every step is checked against the fp32 oracle on the gathered cache."""
import pytest
import torch

import _util  # noqa: F401  (sys.path)
import attention_oracle as orc

pytestmark = pytest.mark.gpu


def test_prefill_then_graph_captured_decode_loop():
    import vllm_flash_attn as fa
    torch.manual_seed(0)
    dev, dt = torch.device("cuda", 0), torch.bfloat16
    h, hk, d, page, steps = 8, 2, 128, 16, 6
    prompt = [300, 57, 1000, 129]
    b = len(prompt)
    max_len = max(prompt) + steps
    blocks_per_seq = (max_len + page - 1) // page
    nblk = b * blocks_per_seq + 7
    k_cache = torch.full((nblk, page, hk, d), float("nan"), device=dev, dtype=dt)   # unused rows must never matter
    v_cache = torch.full_like(k_cache, float("nan"))
    bt = torch.randperm(nblk, device=dev)[: b * blocks_per_seq].to(torch.int32).view(b, blocks_per_seq).contiguous()

    # ---- the cache-layout producer: scatter the prompts' K/V into their pages ----
    k_seq = [torch.randn(max_len, hk, d, device=dev, dtype=dt) for _ in range(b)]   # ground truth, contiguous
    v_seq = [torch.randn(max_len, hk, d, device=dev, dtype=dt) for _ in range(b)]
    from vllm_flash_attn.cache_ops import reshape_and_cache_flash
    slots = []
    for i, n in enumerate(prompt):
        pos = torch.arange(n, device=dev)
        slots.append(bt[i, pos // page].long() * page + pos % page)
    reshape_and_cache_flash(torch.cat([k_seq[i][:n] for i, n in enumerate(prompt)]), torch.cat([v_seq[i][:n] for i, n in enumerate(prompt)]),
                            k_cache, v_cache, torch.cat(slots))
    for i, n in enumerate(prompt):             # the writer against plain indexing
        pos = torch.arange(n, device=dev)
        blk = bt[i, pos // page].long()
        assert torch.equal(k_cache[blk, pos % page], k_seq[i][:n]) and torch.equal(v_cache[blk, pos % page], v_seq[i][:n])

    # ---- prefill: packed queries against the paged cache ----
    cu = torch.tensor([0] + list(torch.tensor(prompt).cumsum(0)), dtype=torch.int32)
    q_prompt = torch.randn(sum(prompt), h, d, device=dev, dtype=dt)
    out = fa.flash_attn_varlen_func(q_prompt, k_cache, v_cache, cu.to(dev), cu.to(dev), max(prompt), max(prompt),
                                    causal=True, block_table=bt)
    k_packed = torch.cat([k_seq[i][:n] for i, n in enumerate(prompt)])
    v_packed = torch.cat([v_seq[i][:n] for i, n in enumerate(prompt)])
    ref32, _ = orc.attention_ref_varlen(q_prompt, k_packed, v_packed, cu, cu, causal=True)
    pt, _ = orc.attention_ref_varlen(q_prompt, k_packed, v_packed, cu, cu, causal=True, upcast=False, reorder_ops=True)
    err, err_pt = (out.float() - ref32.float()).abs().max().item(), (pt.float() - ref32.float()).abs().max().item()
    assert not torch.isnan(out.float()).any() and err <= 2 * err_pt + 1e-5, (err, err_pt)

    # ---- decode: one step captured in a CUDA graph, replayed with new inputs in static buffers ----
    lens = torch.tensor(prompt, dtype=torch.int32, device=dev)       # tokens already in the cache
    q_buf = torch.zeros(b, 1, h, d, device=dev, dtype=dt)
    kn_buf = torch.zeros(b, 1, hk, d, device=dev, dtype=dt)
    vn_buf = torch.zeros_like(kn_buf)
    o_buf = torch.zeros_like(q_buf)

    def step():
        o = fa.flash_attn_with_kvcache(q_buf, k_cache, v_cache, k=kn_buf, v=vn_buf, cache_seqlens=lens,
                                       block_table=bt, causal=True)
        o_buf.copy_(o)
        lens.add_(1)

    # warm up on a side stream with throw-away state, then restore and capture
    snap = (k_cache.clone(), v_cache.clone(), lens.clone())
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(2):
            step()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    k_cache.copy_(snap[0]); v_cache.copy_(snap[1]); lens.copy_(snap[2])
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        step()
    k_cache.copy_(snap[0]); v_cache.copy_(snap[1]); lens.copy_(snap[2])     # capture does not execute

    for s in range(steps):
        q_new = torch.randn(b, 1, h, d, device=dev, dtype=dt)
        q_buf.copy_(q_new)
        for i, n in enumerate(prompt):
            kn_buf[i, 0] = k_seq[i][n + s]; vn_buf[i, 0] = v_seq[i][n + s]
        graph.replay()
        torch.cuda.synchronize()
        for i, n in enumerate(prompt):
            L = n + s + 1
            r32, _ = orc.attention_ref(q_new[i:i + 1], k_seq[i][:L].unsqueeze(0), v_seq[i][:L].unsqueeze(0))
            rpt, _ = orc.attention_ref(q_new[i:i + 1], k_seq[i][:L].unsqueeze(0), v_seq[i][:L].unsqueeze(0),
                                       upcast=False, reorder_ops=True)
            e = (o_buf[i:i + 1].float() - r32.float()).abs().max().item()
            e_pt = (rpt.float() - r32.float()).abs().max().item()
            assert e <= 3 * e_pt + 1e-5, (s, i, e, e_pt)
    assert lens.tolist() == [n + steps for n in prompt]
    # the appended rows are where the producer would have put them
    for i, n in enumerate(prompt):
        pos = torch.arange(n, n + steps, device=dev)
        blk = bt[i, pos // page].long()
        assert torch.equal(k_cache[blk, pos % page], k_seq[i][n:n + steps])
        assert torch.equal(v_cache[blk, pos % page], v_seq[i][n:n + steps])


def test_reshape_and_cache_flash_strided_inputs_and_padding_slots():
    """The cache writer: K/V as strided views of one packed projection output (how a model hands them over), padding
    tokens with slot -1 left out, every other cache row untouched."""
    from vllm_flash_attn.cache_ops import reshape_and_cache_flash
    torch.manual_seed(1)
    dev, dt = "cuda", torch.float16
    n, hk, d, page, nblk = 77, 4, 96, 32, 9
    qkv = torch.randn(n, (8 + 2 * hk) * d, device=dev, dtype=dt)
    key = qkv[:, 8 * d:(8 + hk) * d].view(n, hk, d)
    value = qkv[:, (8 + hk) * d:].view(n, hk, d)
    kc = torch.randn(nblk, page, hk, d, device=dev, dtype=dt); vc = torch.randn_like(kc)
    kc0, vc0 = kc.clone(), vc.clone()
    slot = torch.randperm(nblk * page, device=dev)[:n].to(torch.int64)
    slot[5] = -1; slot[40] = -1
    reshape_and_cache_flash(key, value, kc, vc, slot)
    keep = slot >= 0
    kc0.view(-1, hk, d)[slot[keep]] = key[keep]; vc0.view(-1, hk, d)[slot[keep]] = value[keep]
    assert torch.equal(kc, kc0) and torch.equal(vc, vc0)
    with pytest.raises(RuntimeError):
        reshape_and_cache_flash(key, value, kc, vc, slot.to(torch.int32))
