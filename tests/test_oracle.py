"""CPU: the oracle restatement vs golden vectors produced by the reference's own attention_ref
(tests/golden/make_golden.py), plus the oracle's internal consistency (blockwise kernel-arithmetic
model, split + combine algebra, paged gather)."""
import glob
import os

import pytest
import torch

import attention_oracle as orc
from golden.make_golden import case_inputs

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.pt")))
CASE_FILES = [g for g in GOLD if not g.endswith("paged_layout.pt")]


@pytest.mark.parametrize("path", CASE_FILES, ids=lambda p: os.path.basename(p)[:-3])
def test_attention_ref_matches_reference_golden(path):
    blob = torch.load(path)
    spec = blob["spec"]
    inp = case_inputs(spec)
    causal, window = spec.get("causal", False), spec.get("window", (-1, -1))
    qm, km = inp.get("query_padding_mask"), inp.get("key_padding_mask")
    bias = None
    if "alibi_slopes" in inp:
        bias = orc.attn_bias_from_alibi_slopes(inp["alibi_slopes"], spec["sq"], spec["sk"], qm, km, causal=causal)
        assert torch.equal(bias.double().sum(), blob["bias_sum"])
    out, _ = orc.attention_ref(inp["q"], inp["k"], inp["v"], qm, km, bias, causal=causal, window_size=window)
    out_pt, _ = orc.attention_ref(inp["q"], inp["k"], inp["v"], qm, km, bias, causal=causal, window_size=window,
                                  upcast=False, reorder_ops=True)
    # same ops in the same order on the same seeded inputs: bit-exact
    assert torch.equal(out, blob["out"])
    assert torch.equal(out_pt, blob["out_pt"])
    mask = orc.construct_local_mask(spec["sq"], spec["sk"], (window[0], 0) if causal else window, qm, km)
    assert torch.equal((~mask).reshape(-1, mask.shape[-1]).any(-1).to(torch.uint8), blob["mask_rows_any"])


def test_paged_layout_matches_reference_generator():
    blob = torch.load(os.path.join(os.path.dirname(__file__), "golden", "paged_layout.pt"))
    torch.manual_seed(0)
    kc, vc, bt, kp, vp, nb = orc.generate_block_kvcache(100, 16, 2, 2, 32, "cpu", torch.bfloat16)
    assert nb == blob["num_blocks"]
    assert torch.equal(bt, blob["block_table"])
    assert torch.equal(kp, blob["k_paged"])
    assert torch.equal(kc, blob["k_cache"])


@pytest.mark.parametrize("causal,window", [(False, (-1, -1)), (True, (-1, -1)), (False, (7, 3)), (False, (-1, 4))])
@pytest.mark.parametrize("sq,sk", [(1, 77), (37, 50), (64, 64), (50, 20)])
def test_blockwise_model_matches_attention_ref(causal, window, sq, sk):
    torch.manual_seed(1)
    q = torch.randn(1, sq, 4, 32).bfloat16(); k = torch.randn(1, sk, 2, 32).bfloat16(); v = torch.randn(1, sk, 2, 32).bfloat16()
    o, _ = orc.attention_ref(q, k, v, causal=causal, window_size=window)
    opt, _ = orc.attention_ref(q, k, v, causal=causal, window_size=window, upcast=False, reorder_ops=True)
    ob, lse = orc.flash_blockwise_ref(q[0], k[0], v[0], causal=causal, window_size=window, block_n=16)
    lr = orc.lse_ref(q, k, causal=causal, window_size=window)[0]
    o32 = orc.attention_ref(q.float(), k.float(), v.float(), causal=causal, window_size=window)[0][0]
    assert (ob - o32).abs().max() <= 2 * (opt[0].float() - o32).abs().max() + 1e-5
    fin = torch.isfinite(lr)
    assert torch.equal(torch.isinf(lse), torch.isinf(lr))
    if fin.any():
        assert (lse[fin] - lr[fin]).abs().max() < 1e-4
    assert torch.equal(o, o)  # no NaN
    assert not torch.isnan(ob).any()


def test_split_combine_is_exact_algebra():
    torch.manual_seed(2)
    q = torch.randn(5, 4, 64).half(); k = torch.randn(300, 2, 64).half(); v = torch.randn(300, 2, 64).half()
    full, lse = orc.flash_blockwise_ref(q, k, v, causal=True, block_n=32, round_p=False)
    ranges = [(0, 64), (64, 64), (64, 200), (200, 300)]          # includes an empty split
    parts = [orc.flash_blockwise_ref(q, k, v, causal=True, block_n=32, split_range=r, round_p=False) for r in ranges]
    o_p = torch.stack([p[0].permute(1, 0, 2).reshape(-1, 64) for p in parts])
    l_p = torch.stack([p[1].reshape(-1) for p in parts])
    assert torch.isinf(l_p[1]).all() and (l_p[1] < 0).all()
    oc, lc = orc.combine_ref(o_p, l_p)
    assert (oc.view(4, 5, 64).permute(1, 0, 2) - full).abs().max() < 1e-5
    assert (lc.view(4, 5) - lse).abs().max() < 1e-5
    # all splits empty -> O = 0, lse = +inf
    oc, lc = orc.combine_ref(torch.zeros(3, 2, 8), torch.full((3, 2), float("-inf")))
    assert (oc == 0).all() and torch.isinf(lc).all() and (lc > 0).all()


def test_gather_paged_roundtrip():
    g = torch.Generator().manual_seed(3)
    kp = torch.randn(12, 16, 2, 8, generator=g)
    bt = torch.randperm(12, generator=g).to(torch.int32).view(3, 4)
    kc = orc.gather_paged(kp, bt)
    for b in range(3):
        for r in (0, 15, 16, 63):
            assert torch.equal(kc[b, r], kp[bt[b, r // 16], r % 16])
