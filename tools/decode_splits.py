"""Split-count sweep of the paged decode kernel for small batches (C5 shapes), CUDA-graph timed (development tool).
  python tools/decode_splits.py [--batch 1] [--ctx 16384 32768 65536 131072] [--splits 0 16 24 32 37 48 64 74 111 128]"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vllm-flash-attention_b200"))
import torch
import vllm_flash_attn as fa
from vllm_flash_attn import vllm_flash_attn_2_cuda as ext

def timeit_graph(f, iters=100):
    for _ in range(3): f()
    torch.cuda.synchronize()
    s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s): f()
    torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g): f()
    for _ in range(5): g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): g.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, nargs="*", default=[1]); ap.add_argument("--ctx", type=int, nargs="*", default=[16384, 32768, 65536, 131072])
ap.add_argument("--splits", type=int, nargs="*", default=[0, 16, 24, 32, 37, 48, 64, 74, 111, 128])
ap.add_argument("--hk", type=int, default=8); ap.add_argument("--h", type=int, default=32); ap.add_argument("--d", type=int, default=128)
a = ap.parse_args()
dev, dt, page = "cuda", torch.bfloat16, 16
for b in a.batch:
    for ctx in a.ctx:
        nblk = b * ctx // page
        kc = torch.randn(nblk, page, a.hk, a.d, device=dev, dtype=dt); vc = torch.randn_like(kc)
        bt = torch.randperm(nblk, device=dev).to(torch.int32).view(b, -1).contiguous()
        q = torch.randn(b, 1, a.h, a.d, device=dev, dtype=dt); lens = torch.full((b,), ctx, dtype=torch.int32, device=dev)
        out = torch.empty_like(q)
        nbytes = 2.0 * b * ctx * a.hk * a.d * 2
        row = {"batch": b, "ctx": ctx, "MB": nbytes / 1e6, "ideal_us_at_6555": nbytes / 6554.9e9 * 1e6}
        for s in a.splits:
            f = lambda: fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=lens, block_table=bt, causal=True, num_splits=s, out=out)
            f(); used = ext.last_call_info["num_splits"]
            us = timeit_graph(f) * 1e3
            row[f"s{s}" + (f"->{used}" if s == 0 else "")] = round(us, 1)
        print(json.dumps(row), flush=True)
