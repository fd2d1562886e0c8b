"""Generate tests/golden/*.pt from the REFERENCE's own attention_ref.

Run in the build container only (needs /root/reference):  python tests/golden/make_golden.py

The reference test module cannot be imported on a CPU box (module-level
torch.cuda.get_device_capability, tests/test_flash_attn.py:23-26), so the function
definitions we need are lifted out of its source with `ast` and executed here, unmodified,
in a namespace that holds torch / einops / math.  Nothing from the reference is written into
this repo except the numeric outputs.
"""
import ast
import math
import os
import sys

import torch
from einops import rearrange, repeat

REF_TESTS = "/root/reference/tests/test_flash_attn.py"
HERE = os.path.dirname(os.path.abspath(__file__))
WANTED = {"attn_bias_from_alibi_slopes", "generate_random_padding_mask", "construct_local_mask",
          "attention_ref", "_generate_block_kvcache"}


def load_reference_functions():
    src = open(REF_TESTS).read()
    tree = ast.parse(src)
    ns = {"torch": torch, "math": math, "rearrange": rearrange, "repeat": repeat}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in WANTED:
            mod = ast.Module(body=[node], type_ignores=[])
            exec(compile(mod, REF_TESTS, "exec"), ns)
    missing = WANTED - set(ns)
    assert not missing, missing
    return ns


def case_inputs(spec):
    """Seeded inputs; shared with tests/test_oracle.py so only outputs need storing."""
    g = torch.Generator().manual_seed(spec["seed"])
    dt = getattr(torch, spec["dtype"])
    b, sq, sk, h, hk, d = (spec[x] for x in ("b", "sq", "sk", "h", "hk", "d"))
    q = torch.randn(b, sq, h, d, generator=g).to(dt)
    k = torch.randn(b, sk, hk, d, generator=g).to(dt)
    v = torch.randn(b, sk, hk, d, generator=g).to(dt)
    out = {"q": q, "k": k, "v": v}
    if spec.get("alibi"):
        out["alibi_slopes"] = torch.rand(b, h, generator=g) * 0.3
    if spec.get("padding"):
        lq = torch.randint(max(1, sq - 20), sq + 1, (b, 1), generator=g)
        lk = torch.randint(max(1, sk - 20), sk + 1, (b, 1), generator=g)
        out["query_padding_mask"] = torch.arange(sq).expand(b, sq) < lq
        out["key_padding_mask"] = torch.arange(sk).expand(b, sk) < lk
    return out


CASES = [
    dict(name="c1_dense_b2_s512_h8_d64", seed=0, dtype="bfloat16", b=2, sq=512, sk=512, h=8, hk=8, d=64),
    dict(name="causal_gqa_d128", seed=1, dtype="bfloat16", b=2, sq=113, sk=203, h=8, hk=2, d=128, causal=True),
    dict(name="causal_sq_gt_sk", seed=2, dtype="float16", b=1, sq=239, sk=100, h=4, hk=4, d=64, causal=True),
    dict(name="decode_q1_mqa", seed=3, dtype="bfloat16", b=3, sq=1, sk=239, h=6, hk=1, d=128, causal=True),
    dict(name="local_window", seed=4, dtype="float16", b=2, sq=128, sk=217, h=4, hk=2, d=64, window=(17, 5)),
    dict(name="alibi_causal", seed=5, dtype="bfloat16", b=2, sq=97, sk=97, h=4, hk=4, d=64, causal=True, alibi=True),
    dict(name="alibi_noncausal_padding", seed=6, dtype="float16", b=2, sq=64, sk=128, h=4, hk=4, d=32, alibi=True, padding=True),
    dict(name="padding_causal_d128", seed=7, dtype="bfloat16", b=3, sq=130, sk=130, h=4, hk=2, d=128, causal=True, padding=True),
]


def main():
    ns = load_reference_functions()
    torch.set_num_threads(8)
    for spec in CASES:
        inp = case_inputs(spec)
        causal = spec.get("causal", False)
        window = spec.get("window", (-1, -1))
        qm, km = inp.get("query_padding_mask"), inp.get("key_padding_mask")
        bias = None
        if "alibi_slopes" in inp:
            bias = ns["attn_bias_from_alibi_slopes"](inp["alibi_slopes"], spec["sq"], spec["sk"], qm, km, causal=causal)
        out, _ = ns["attention_ref"](inp["q"], inp["k"], inp["v"], qm, km, bias, 0.0, None, causal=causal, window_size=window)
        out_pt, _ = ns["attention_ref"](inp["q"], inp["k"], inp["v"], qm, km, bias, 0.0, None, causal=causal,
                                        window_size=window, upcast=False, reorder_ops=True)
        mask = ns["construct_local_mask"](spec["sq"], spec["sk"], (window[0], 0) if causal else window, qm, km)
        blob = {"spec": spec, "out": out, "out_pt": out_pt,
                "mask_rows_any": (~mask).reshape(-1, mask.shape[-1]).any(-1).to(torch.uint8)}
        if bias is not None:
            blob["bias_sum"] = bias.double().sum()
        torch.save(blob, os.path.join(HERE, spec["name"] + ".pt"))
        print(spec["name"], tuple(out.shape), float(out.float().abs().mean()))
    # paged cache generator: pin the block-table layout (random permutation, 3x pool)
    torch.manual_seed(0)
    kc, vc, bt, kp, vp, nb = ns["_generate_block_kvcache"](100, 16, 2, 2, 32, "cpu", torch.bfloat16)
    torch.save({"k_cache": kc, "block_table": bt, "k_paged": kp, "num_blocks": nb},
               os.path.join(HERE, "paged_layout.pt"))
    print("paged_layout", tuple(kc.shape), nb)


if __name__ == "__main__":
    sys.exit(main())
