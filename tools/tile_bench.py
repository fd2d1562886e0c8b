"""Prefill throughput of the head dims above 128 (mma.sync tile kernel), development tool."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch
import _util  # noqa: F401
import vllm_flash_attn as fa
from vllm_flash_attn import vllm_flash_attn_2_cuda as ext
from prefill_bench import timeit
dev, dt = "cuda", torch.bfloat16
for d in (160, 192, 224, 256):
    for causal in (True, False):
        h, hk, b, s = 2048 // 256 * 2, 4, 4, 4096
        q = torch.randn(b, s, h, d, device=dev, dtype=dt); k = torch.randn(b, s, hk, d, device=dev, dtype=dt); v = torch.randn_like(k)
        ms = timeit(lambda: fa.flash_attn_func(q, k, v, causal=causal), 10)
        fl = 4.0 * b * s * s * h * d / (2 if causal else 1) / 1e12
        print(json.dumps({"d": d, "causal": causal, "ms": round(ms, 3), "TFLOPs": round(fl / ms * 1e3, 1), "kernel": ext.last_call_info["kernel_used"]}), flush=True)
