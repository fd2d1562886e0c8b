"""Paged-decode bandwidth sweep over head dim / batch / context (development tool): one JSON line per point."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch
import _util  # noqa: F401
import vllm_flash_attn as fa
from vllm_flash_attn import vllm_flash_attn_2_cuda as ext
from prefill_bench import timeit

dims = [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "64,128").split(",")]
dev, dt, page, hk, g = "cuda", torch.bfloat16, 16, 8, 4
for d in dims:
    for b, ctx in ((128, 4096), (256, 4096), (512, 1024), (64, 16384)):
        nblk = b * ctx // page
        pools = [(torch.randn(nblk, page, hk, d, device=dev, dtype=dt), torch.randn(nblk, page, hk, d, device=dev, dtype=dt)) for _ in range(3)]
        bt = torch.randperm(nblk, device=dev).to(torch.int32).view(b, -1).contiguous()
        lens = torch.full((b,), ctx, dtype=torch.int32, device=dev)
        q = torch.randn(b, 1, hk * g, d, device=dev, dtype=dt)
        i = [0]
        def step():
            kc, vc = pools[i[0] % 3]; i[0] += 1
            fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=lens, block_table=bt, causal=True)
        ms = timeit(step, 30)
        print(json.dumps({"d": d, "b": b, "ctx": ctx, "ms": round(ms, 4), "GBps": round(2.0 * b * ctx * hk * d * 2 / ms / 1e6, 1),
                          "splits": ext.last_call_info["num_splits"]}), flush=True)
        del pools
