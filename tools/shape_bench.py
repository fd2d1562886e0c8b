"""Throughput of the shapes beside the headline configs (development / reporting tool): prefill at every head
dim the tcgen05 kernel serves (32, 64, 96, 128) plus the tile kernel's d=256, a Mistral-style sliding window,
and paged decode at every head dim of the reference build.  One JSON line per shape (CUDA events, inputs > L2
or rotated pools).
  python tools/shape_bench.py [--iters 20] > gpurun_out/shapes.jsonl
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402
import _util  # noqa: E402,F401
sys.path.insert(0, os.path.join(ROOT, "tools"))
import vllm_flash_attn as fa  # noqa: E402
from vllm_flash_attn import vllm_flash_attn_2_cuda as ext  # noqa: E402
from prefill_bench import timeit  # noqa: E402

KNAME = {1: "tcgen05_prefill", 2: "tile", 3: "decode"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    dev, dt = "cuda", torch.bfloat16
    torch.manual_seed(0)
    # ---- prefill: 16384 tokens as 4 x 4096, causal, 2048 model width (h = 2048 / d), GQA 4 ----
    for d in (32, 64, 96, 128, 256):
        h = 2048 // d if d != 96 else 24
        hk = max(h // 4, 1)
        b, s = 4, 4096
        q = torch.randn(b, s, h, d, device=dev, dtype=dt)
        k = torch.randn(b, s, hk, d, device=dev, dtype=dt); v = torch.randn_like(k)
        ms = timeit(lambda: fa.flash_attn_func(q, k, v, causal=True), a.iters)
        fl = 4.0 * b * s * s * h * d / 2 / 1e12
        print(json.dumps({"shape": f"prefill causal b{b} s{s} h{h}/{hk} d{d}", "ms": round(ms, 4), "TFLOPs": round(fl / ms * 1e3, 1),
                          "kernel": KNAME[ext.last_call_info["kernel_used"]]}), flush=True)
    # ---- sliding window (Mistral 7B: 32q/8kv d128, window 4096) over one 32k sequence ----
    s, h, hk, d, w = 32768, 32, 8, 128, 4096
    q = torch.randn(1, s, h, d, device=dev, dtype=dt)
    k = torch.randn(1, s, hk, d, device=dev, dtype=dt); v = torch.randn_like(k)
    ms = timeit(lambda: fa.flash_attn_func(q, k, v, causal=True, window_size=(w - 1, 0)), a.iters)
    vis = sum(min(i + 1, w) for i in range(s))
    print(json.dumps({"shape": f"prefill causal window {w} s{s} h{h}/{hk} d{d}", "ms": round(ms, 4),
                      "TFLOPs": round(4.0 * vis * h * d / 1e12 / ms * 1e3, 1), "kernel": KNAME[ext.last_call_info["kernel_used"]]}), flush=True)
    # ---- paged decode: batch 128, context 4096, page 16, 8 KV heads, group 4; 3 pools rotated (each > L2) ----
    for d in (32, 64, 96, 128, 160, 192, 224, 256):
        b, ctx, page, hk, g = 128, 4096, 16, 8, 4
        nblk = b * ctx // page
        pools = [(torch.randn(nblk, page, hk, d, device=dev, dtype=dt), torch.randn(nblk, page, hk, d, device=dev, dtype=dt)) for _ in range(3)]
        bt = torch.randperm(nblk, device=dev).to(torch.int32).view(b, -1).contiguous()
        lens = torch.full((b,), ctx, dtype=torch.int32, device=dev)
        q = torch.randn(b, 1, hk * g, d, device=dev, dtype=dt)
        i = [0]

        def step():
            kc, vc = pools[i[0] % 3]; i[0] += 1
            fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=lens, block_table=bt, causal=True)
        ms = timeit(step, a.iters * 3)
        byts = 2.0 * b * ctx * hk * d * 2
        print(json.dumps({"shape": f"decode paged16 b{b} ctx{ctx} h{hk * g}/{hk} d{d}", "ms": round(ms, 4), "GBps": round(byts / ms / 1e6, 1),
                          "kernel": KNAME[ext.last_call_info["kernel_used"]], "splits": ext.last_call_info["num_splits"]}), flush=True)
        del pools


if __name__ == "__main__":
    main()
