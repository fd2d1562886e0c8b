"""Time the REFERENCE's own FA2 kernels (oracle/_ref, built for sm_100 from /root/reference) and
ours on the BASELINE.json configs C1/C2/C3/C5 on one B200 — the "reference on B200" column
of BASELINE.md §3.  Development/measurement tool (TEST INFRASTRUCTURE: uses oracle/_ref).

  python tools/ref_kernel_bench.py [--iters 20] > gpurun_out/ref_kernel_bench.jsonl
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402
import _util  # noqa: E402,F401
import vllm_flash_attn as fa  # noqa: E402
from vllm_flash_attn import vllm_flash_attn_2_cuda as ext  # noqa: E402


def timeit(f, iters, warmup=5):
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    dev, dt = "cuda", torch.bfloat16
    have_ref = _util.ref_lib() is not None
    torch.manual_seed(0)
    rows = []

    def emit(name, work, unit, ms_ours, ms_ref, extra=None):
        r = {"config": name, "unit": unit, "ours_ms": ms_ours, "ours": work / (ms_ours * 1e-3),
             "ref_kernel_ms": ms_ref, "ref_kernel": work / (ms_ref * 1e-3) if ms_ref else None}
        if ms_ref:
            r["speedup"] = ms_ref / ms_ours
        r.update(extra or {})
        rows.append(r)
        print(json.dumps(r), flush=True)

    # ---- C1 dense B=2 S=512 H=8 D=64 non-causal
    q = torch.randn(2, 512, 8, 64, device=dev, dtype=dt); k = torch.randn_like(q); v = torch.randn_like(q)
    ms = timeit(lambda: fa.flash_attn_func(q, k, v), a.iters)
    msr = timeit(lambda: _util.ref_kernel_fwd(q, k, v), a.iters) if have_ref else None
    emit("C1 dense B2 S512 H8 D64", 4 * 2 * 512 * 512 * 8 * 64 / 1e12, "TFLOP/s", ms, msr, {"kernel": ext.last_call_info["kernel_used"]})

    # ---- C2 varlen causal GQA 32/8 D=128 16k tokens
    for name, seqlens in (("C2 varlen mixed 16k", [8192, 4096, 2048, 1024, 512, 256, 128, 128]), ("C2 varlen uniform 8x2048", [2048] * 8)):
        tot = sum(seqlens)
        q = torch.randn(tot, 32, 128, device=dev, dtype=dt)
        k = torch.randn(tot, 8, 128, device=dev, dtype=dt); v = torch.randn_like(k)
        cu = torch.tensor([0] + list(torch.tensor(seqlens).cumsum(0)), dtype=torch.int32, device=dev)
        mx = max(seqlens)
        flops = 2.0 * 32 * 128 * sum(s * s for s in seqlens) / 1e12
        ms = timeit(lambda: fa.flash_attn_varlen_func(q, k, v, cu, cu, mx, mx, causal=True), a.iters)
        msr = timeit(lambda: _util.ref_kernel_varlen(q, k, v, cu, cu, mx, mx, causal=True), a.iters) if have_ref else None
        emit(name, flops, "TFLOP/s", ms, msr, {"kernel": ext.last_call_info["kernel_used"]})

    # ---- dense causal/non-causal sweeps d=128 (16k tokens total)
    for causal in (False, True):
        for s in (1024, 4096, 16384):
            b = 16384 // s
            q = torch.randn(b, s, 16, 128, device=dev, dtype=dt); k = torch.randn_like(q); v = torch.randn_like(q)
            flops = 4.0 * b * s * s * 16 * 128 / (2 if causal else 1) / 1e12
            ms = timeit(lambda: fa.flash_attn_func(q, k, v, causal=causal), a.iters)
            msr = timeit(lambda: _util.ref_kernel_fwd(q, k, v, causal=causal), a.iters) if have_ref else None
            emit(f"dense d128 h16 b{b} s{s} causal={causal}", flops, "TFLOP/s", ms, msr)

    # ---- C3 paged decode B=256 ctx 4096 block 16
    for name, b, ctx in (("C3 paged decode B256 ctx4096", 256, 4096), ("C5 long decode B1 ctx131072", 1, 131072),
                         ("C5 long decode B8 ctx131072", 8, 131072)):
        page = 16
        bps = ctx // page
        nblk = b * bps * 5 // 4
        kc = torch.randn(nblk, page, 8, 128, device=dev, dtype=dt); vc = torch.randn_like(kc)
        bt = torch.randperm(nblk, device=dev)[: b * bps].to(torch.int32).view(b, -1).contiguous()
        q = torch.randn(b, 1, 32, 128, device=dev, dtype=dt)
        lens = torch.full((b,), ctx, dtype=torch.int32, device=dev)
        nbytes = 2.0 * b * ctx * 8 * 128 * 2 / 1e9
        ms = timeit(lambda: fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=lens, block_table=bt, causal=True), a.iters)
        info = dict(ext.last_call_info)
        msr = timeit(lambda: _util.ref_kernel_kvcache(q, kc, vc, lens, bt, causal=True), a.iters) if have_ref else None
        emit(name, nbytes, "GB/s", ms, msr, {"num_splits": info["num_splits"], "kernel": info["kernel_used"]})
        del kc, vc


if __name__ == "__main__":
    main()
