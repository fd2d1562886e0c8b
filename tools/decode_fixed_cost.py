"""What a decode launch costs beside its stream (development tool): one CUDA-graph replay of flash_attn_with_kvcache per
context length, from 0 keys (every CTA exits at once: launch + graph floor) over 16 keys (one chunk: prologue chain +
epilogue) up to 4096, for a few batch sizes; the slope is the stream, the intercept the fixed cost.
  python tools/decode_fixed_cost.py [--batch 32 128] [--hk 8] [--h 32]"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vllm-flash-attention_b200"))
import torch
import vllm_flash_attn as fa
from vllm_flash_attn import vllm_flash_attn_2_cuda as ext


def timeit_graph(f, iters=200):
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
    return e0.elapsed_time(e1) / iters * 1e3


ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, nargs="*", default=[32, 128]); ap.add_argument("--hk", type=int, default=8)
ap.add_argument("--h", type=int, default=32); ap.add_argument("--d", type=int, default=128)
ap.add_argument("--ctx", type=int, nargs="*", default=[0, 16, 64, 256, 512, 1024, 2048, 4096])
a = ap.parse_args()
dev, dt, page = "cuda", torch.bfloat16, 16
for b in a.batch:
    cmax = max(max(a.ctx), 16)
    nblk = b * cmax // page
    kc = torch.randn(nblk, page, a.hk, a.d, device=dev, dtype=dt); vc = torch.randn_like(kc)
    bt = torch.randperm(nblk, device=dev).to(torch.int32).view(b, -1).contiguous()
    q = torch.randn(b, 1, a.h, a.d, device=dev, dtype=dt); out = torch.empty_like(q)
    row = {"batch": b, "units": b * a.hk}
    for ctx in a.ctx:
        lens = torch.full((b,), ctx, dtype=torch.int32, device=dev)
        f = lambda: fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=lens, block_table=bt, causal=True, num_splits=1, out=out)
        f()
        row[f"ctx{ctx}"] = round(timeit_graph(f), 2)
    print(json.dumps(row), flush=True)
