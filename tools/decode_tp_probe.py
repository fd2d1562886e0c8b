"""Decode half of C4's TP shard alone (192 sequences, ctx randint(1024, 4097), 64/tp q heads, 8/tp kv heads, paged block
16) through flash_attn_with_kvcache at forced split counts, with random and with uniform context lengths: how much of
the shard's decode time is imbalance / tail and how much per-CTA overhead.  Development tool.
  python tools/decode_tp_probe.py [--tp 8]"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vllm-flash-attention_b200"))
import torch
import vllm_flash_attn as fa
from vllm_flash_attn import vllm_flash_attn_2_cuda as ext

def timeit(f, iters=50):
    for _ in range(5): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters

ap = argparse.ArgumentParser(); ap.add_argument("--tp", type=int, nargs="*", default=[8, 4])
a = ap.parse_args()
dev, dt, D, page, B = "cuda", torch.bfloat16, 128, 16, 192
for tp in a.tp:
    Hq, Hk = 64 // tp, 8 // tp
    g = torch.Generator(device="cpu").manual_seed(0)
    rnd = torch.randint(1024, 4097, (B,), generator=g)
    mb = 4096 // page
    kc = torch.randn(B * mb, page, Hk, D, device=dev, dtype=dt); vc = torch.randn_like(kc)
    bt = torch.randperm(B * mb, device=dev).to(torch.int32).view(B, mb).contiguous()
    q = torch.randn(B, 1, Hq, D, device=dev, dtype=dt)
    for name, lens in (("random", rnd), ("uniform", torch.full((B,), int(rnd.float().mean())))):
        cl = lens.to(torch.int32).to(dev)
        nbytes = 2.0 * float(lens.sum()) * Hk * D * 2
        for s in (0, 1, 2, 3, 4, 6, 8, 12):
            f = lambda: fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=cl, block_table=bt, causal=True, num_splits=s)
            f(); used = ext.last_call_info["num_splits"]
            ms = timeit(f)
            print(json.dumps({"tp": tp, "lens": name, "num_splits": s, "used": used, "ms": round(ms, 4), "gbs": round(nbytes / ms / 1e6, 1)}), flush=True)
