"""Quick iteration tool for the tcgen05 prefill kernel: a correctness spot check against the fp32
oracle (TEST INFRASTRUCTURE) and CUDA-event timings of a few shapes.  Not part of the product.

  python tools/prefill_bench.py [--iters 20] [--quick]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402
import _util  # noqa: E402,F401
import attention_oracle as orc  # noqa: E402
import vllm_flash_attn as fa  # noqa: E402
from vllm_flash_attn import vllm_flash_attn_2_cuda as ext  # noqa: E402


def timeit(f, iters, warmup=3):
    for _ in range(warmup):
        f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        f()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def check(name, d, sq, sk, causal, h=4, hk=2, b=2, dt=torch.bfloat16, window=(-1, -1)):
    torch.manual_seed(0)
    q = torch.randn(b, sq, h, d, device="cuda", dtype=dt)
    k = torch.randn(b, sk, hk, d, device="cuda", dtype=dt); v = torch.randn_like(k)
    out, lse, _ = fa.flash_attn_func(q, k, v, causal=causal, window_size=window, return_attn_probs=True)
    ref32, _ = orc.attention_ref(q, k, v, causal=causal, window_size=window)
    pt, _ = orc.attention_ref(q, k, v, causal=causal, window_size=window, upcast=False, reorder_ops=True)
    lref = orc.lse_ref(q, k, causal=causal, window_size=window)
    e, ept = (out.float() - ref32.float()).abs().max().item(), (pt.float() - ref32.float()).abs().max().item()
    fin = torch.isfinite(lref)
    le = (lse[fin] - lref[fin]).abs().max().item()
    ok = e <= 2 * ept + 1e-5 and le < 2e-3 and not torch.isnan(out.float()).any().item()
    print(json.dumps({"check": name, "ok": ok, "err": e, "err_pt": ept, "lse_err": le, "kernel": ext.last_call_info["kernel_used"]}), flush=True)
    return ok


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--quick", action="store_true")
    a = ap.parse_args()
    dev, dt = "cuda", torch.bfloat16
    ok = True
    ok &= check("d128 causal 1000x1000", 128, 1000, 1000, True)
    ok &= check("d128 noncausal 384x777", 128, 384, 777, False)
    ok &= check("d64 causal 515x515 fp16", 64, 515, 515, True, dt=torch.float16)
    ok &= check("d128 local 700x900", 128, 700, 900, False, window=(100, 30))
    ok &= check("d128 causal sq>sk", 128, 600, 200, True)
    print(json.dumps({"all_checks_ok": bool(ok)}), flush=True)

    seqlens = [8192, 4096, 2048, 1024, 512, 256, 128, 128]
    tot = sum(seqlens)
    q = torch.randn(tot, 32, 128, device=dev, dtype=dt)
    k = torch.randn(tot, 8, 128, device=dev, dtype=dt); v = torch.randn_like(k)
    cu = torch.tensor([0] + list(torch.tensor(seqlens).cumsum(0)), dtype=torch.int32, device=dev)
    flops = 2.0 * 32 * 128 * sum(s * s for s in seqlens) / 1e12
    ms = timeit(lambda: fa.flash_attn_varlen_func(q, k, v, cu, cu, 8192, 8192, causal=True), a.iters)
    print(json.dumps({"config": "C2 varlen mixed 16k", "ms": ms, "TFLOPs": flops / (ms * 1e-3)}), flush=True)
    if a.quick:
        return
    for hq, hk, b, s in ((32, 8, 8, 2048), (64, 8, 4, 2048), (32, 8, 2, 8192)):   # GQA: head-pair work units
        for causal in (True, False):
            q = torch.randn(b, s, hq, 128, device=dev, dtype=dt); k = torch.randn(b, s, hk, 128, device=dev, dtype=dt); v = torch.randn_like(k)
            fl = 4.0 * b * s * s * hq * 128 / (2 if causal else 1) / 1e12
            ms = timeit(lambda: fa.flash_attn_func(q, k, v, causal=causal), a.iters)
            print(json.dumps({"config": f"gqa d128 {hq}q/{hk}kv b{b} s{s} causal={causal}", "ms": ms, "TFLOPs": fl / (ms * 1e-3)}), flush=True)
    for d, h in ((128, 16), (64, 32)):
        for causal in (False, True):
            for s in (2048, 8192):
                b = 16384 // s
                q = torch.randn(b, s, h, d, device=dev, dtype=dt); k = torch.randn_like(q); v = torch.randn_like(q)
                fl = 4.0 * b * s * s * h * d / (2 if causal else 1) / 1e12
                ms = timeit(lambda: fa.flash_attn_func(q, k, v, causal=causal), a.iters)
                print(json.dumps({"config": f"dense d{d} h{h} b{b} s{s} causal={causal}", "ms": ms, "TFLOPs": fl / (ms * 1e-3)}), flush=True)


if __name__ == "__main__":
    main()
