"""Long-context decode (C5 shape, one GPU): time vs num_splits, CUDA-graph replay (device time).
  python tools/decode_splits.py [--ctx 131072] [--batch 1]"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vllm-flash-attention_b200"))
import torch
import vllm_flash_attn as fa
from vllm_flash_attn import vllm_flash_attn_2_cuda as ext

ap = argparse.ArgumentParser(); ap.add_argument("--ctx", type=int, default=131072); ap.add_argument("--batch", type=int, default=1)
a = ap.parse_args()
dev, dt, H, HK, D, page = "cuda", torch.bfloat16, 32, 8, 128, 16
b, ctx = a.batch, a.ctx
nblk = b * ctx // page
kc = torch.randn(nblk, page, HK, D, device=dev, dtype=dt); vc = torch.randn_like(kc)
bt = torch.randperm(nblk, device=dev).to(torch.int32).view(b, -1).contiguous()
q = torch.randn(b, 1, H, D, device=dev, dtype=dt)
lens = torch.full((b,), ctx, dtype=torch.int32, device=dev)
nbytes = 2.0 * b * ctx * HK * D * 2
for ns in (0, 16, 24, 32, 37, 48, 64, 74, 96, 111, 128):
    f = lambda: fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=lens, block_table=bt, causal=True, num_splits=ns)
    side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3): f()
    torch.cuda.current_stream().wait_stream(side); torch.cuda.synchronize()
    used = ext.last_call_info["num_splits"]
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(10): f()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): g.replay()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 100
    print(json.dumps({"num_splits_arg": ns, "num_splits_used": used, "us": round(us, 2), "GBps": round(nbytes / us / 1e3, 1)}), flush=True)
