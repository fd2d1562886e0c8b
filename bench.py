#!/usr/bin/env python
"""bench.py — headline benchmark of the B200-native flash-attention forward library.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload prefill|decode]

A "step" = one pass of the hot path over one batch of synthetic input.
N=1 workload (BASELINE.json configs[1], "C2"): Llama-3-8B prefill — flash_attn_varlen_func, causal,
GQA 32q/8kv heads, D=128, 16 384 tokens over seqlens [8192,4096,2048,1024,512,256,128,128], bf16.
metric = attention forward TFLOP/s, FLOPs = 2*H*D*sum(s_i^2) (= 4*s^2*H*D/2 causal; reference
benchmarks/benchmark_flash_attention.py:27-30).  N>1: head-sharded replicas of the same per-GPU
work (each rank owns 32q/8kv heads of a 32N-head layer; no collective) -> weak scaling.
The JSON line also carries the paged-KV decode figure (C3: B=256, ctx 4096, block 16) under "decode".
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (os.path.join(ROOT, "vllm-flash-attention_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

SEQLENS_C2 = [8192, 4096, 2048, 1024, 512, 256, 128, 128]
H, HK, D = 32, 8, 128


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


def prefill_flops(seqlens=SEQLENS_C2, h=H, d=D):
    return 2.0 * h * d * sum(s * s for s in seqlens)     # causal: half of 4*s^2*h*d


class ClockSampler:
    """SM clock + throttle reasons sampled DURING the timed region (B200_PROFILING.md clocks line): NVML polled
    every 2 ms from a thread; falls back to `nvidia-smi -lms 100` when pynvml is unavailable."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.samples, self.proc, self.nv, self.stop_flag = index, [], None, None, False
        self.sm, self.mx, self.reasons = [], None, set()

    def start(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = int(vis.split(",")[self.index]) if vis and vis.split(",")[self.index].isdigit() else self.index
            self.h = nv.nvmlDeviceGetHandleByIndex(idx)
            self.mx = float(nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM))
            self.nv = nv
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.nv = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _poll(self):
        nv = self.nv
        bits = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown if hasattr(nv, "nvmlClocksEventReasonHwSlowdown") else 0x8,
                "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self.stop_flag:
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                r = int(get_reasons(self.h))
                for n, b in bits.items():
                    if r & b:
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.002)

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self):
        if self.nv is not None:
            self.stop_flag = True
            self.thread.join(timeout=1.0)
            sm = sorted(self.sm)
            return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_min_mhz": sm[0] if sm else None, "sm_max_mhz": self.mx,
                    "reasons": sorted(self.reasons), "samples": len(sm), "source": "nvml 2 ms poll"}
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            f = [x.strip() for x in s.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm), "source": "nvidia-smi -lms 100"}


def ncu_traffic(kernel):
    """DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` summary
    (profiles/traffic.json, written by tools/ncu_summary.py); None when no capture is committed."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(kernel)
    except Exception:
        return None


def make_prefill_inputs(dev, seed=0):
    import torch
    g = torch.Generator(device="cpu").manual_seed(seed)
    total = sum(SEQLENS_C2)
    q = torch.randn(total, H, D, generator=g, dtype=torch.float32).to(torch.bfloat16)
    k = torch.randn(total, HK, D, generator=g, dtype=torch.float32).to(torch.bfloat16)
    v = torch.randn(total, HK, D, generator=g, dtype=torch.float32).to(torch.bfloat16)
    cu = torch.tensor([0] + list(__import__("itertools").accumulate(SEQLENS_C2)), dtype=torch.int32)
    return q, k, v, cu


def cpu_attention_sample(threads, budget_s=20.0):
    """The reference's PyTorch attention_ref (our oracle port, oracle/attention_oracle.py) timed on the
    host cores on a bounded sample of the C2 workload: whole sequences from the short end of the
    batch, all 32 heads, causal, fp32 upcast.  Returns (tflops, description)."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import attention_oracle as orc
    torch.set_num_threads(threads)
    g = torch.Generator().manual_seed(0)
    done_flops, t_total, used = 0.0, 0.0, []
    for s in [128, 128, 256, 512, 1024, 2048]:
        q = torch.randn(1, s, H, D, generator=g).to(torch.bfloat16)
        k = torch.randn(1, s, HK, D, generator=g).to(torch.bfloat16)
        v = torch.randn(1, s, HK, D, generator=g).to(torch.bfloat16)
        t0 = time.perf_counter()
        orc.attention_ref(q, k, v, causal=True)
        dt = time.perf_counter() - t0
        done_flops += prefill_flops([s]); t_total += dt; used.append(s)
        if t_total > budget_s:
            break
    return done_flops / t_total / 1e12, f"attention_ref (fp32 upcast) on sequences {used} of the C2 batch, 32q/8kv heads, causal"


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path (PyTorch attention_ref,
    restated in oracle/attention_oracle.py) on the host cores; each step = a bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    threads = os.cpu_count() or 1
    for _ in range(min(args.warmup, 1)):
        cpu_attention_sample(threads, budget_s=2.0)
    vals, desc = [], ""
    t0 = time.perf_counter()
    for _ in range(args.steps):
        v, desc = cpu_attention_sample(threads, budget_s=6.0)
        vals.append(v)
    ms = (time.perf_counter() - t0) / max(args.steps, 1) * 1e3
    val = sum(vals) / len(vals)
    line = {"impl": "reference", "metric": "attn_fwd_tflops_prefill", "value": val, "unit": "TFLOP/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "C2 Llama-3-8B prefill: flash_attn_varlen_func causal GQA 32q/8kv D=128, 16384 tokens, "
                                   "seqlens " + str(SEQLENS_C2),
                       "sample": "each step = a bounded sample of that workload on the host cores: " + desc},
            "cpu_baseline": {"value": val, "unit": "TFLOP/s", "cores": threads, "kind": "port", "sample": desc},
            "e2e": {"value": val, "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def _ensure_built():
    """Build outputs are git-ignored: a fresh checkout compiles the library in-tree (rank 0 only) before importing it."""
    pkg = os.path.join(ROOT, "vllm-flash-attention_b200")
    so = os.path.join(pkg, "vllm_flash_attn", "libfa_b200.so")
    if os.path.exists(so):
        return
    if int(os.environ.get("LOCAL_RANK", "0")) == 0:
        subprocess.run(["bash", os.path.join(pkg, "build.sh")], check=True, stdout=subprocess.DEVNULL)
    else:
        last = -1
        for _ in range(600):            # wait until rank 0 has finished linking (size stable)
            size = os.path.getsize(so) if os.path.exists(so) else -1
            if size > 0 and size == last:
                break
            last = size
            time.sleep(2)


def run_ours(args):
    _ensure_built()
    import torch
    import torch.distributed as dist
    import vllm_flash_attn as fa
    from vllm_flash_attn import vllm_flash_attn_2_cuda as ext
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    pk, pk_src = peaks()

    # ------------------------------------------------------------------ prefill (headline)
    qh, kh, vh, cuh = make_prefill_inputs(dev, seed=rank)
    qh, kh, vh = qh.pin_memory(), kh.pin_memory(), vh.pin_memory()
    q, k, v, cu = qh.to(dev), kh.to(dev), vh.to(dev), cuh.to(dev)
    out = torch.empty_like(q)
    out_h = torch.empty(out.shape, dtype=out.dtype, pin_memory=True)
    max_s = max(SEQLENS_C2)
    flops = prefill_flops()

    def step():
        fa.flash_attn_varlen_func(q, k, v, cu, cu, max_s, max_s, causal=True, out=out)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    launches_per_step = ext.last_call_info["launches"]
    kernel_used = ext.last_call_info["kernel_used"]
    sampler = ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    # ---- e2e: host buffers in, host result out, through the public API.  Every step copies its own
    # inputs from pinned host memory and its result back; the three legs run on three streams with
    # double-buffered device tensors so step i's attention overlaps step i+1's H2D and step i-1's D2H
    # (PCIe is full duplex) — what a serving loop does.  All copies are inside the timed region.
    NB = 2
    cur = torch.cuda.current_stream()
    h2d_s, d2h_s = torch.cuda.Stream(), torch.cuda.Stream()
    bufs = [dict(q=torch.empty_like(q), k=torch.empty_like(k), v=torch.empty_like(v), o=torch.empty_like(out),
                 h2d=torch.cuda.Event(), comp=torch.cuda.Event(), d2h=torch.cuda.Event()) for _ in range(NB)]

    def e2e_loop(n):
        for i in range(n):
            B = bufs[i % NB]
            with torch.cuda.stream(h2d_s):
                if i >= NB:
                    h2d_s.wait_event(B["comp"])        # device inputs of step i-NB no longer in use
                B["q"].copy_(qh, non_blocking=True); B["k"].copy_(kh, non_blocking=True); B["v"].copy_(vh, non_blocking=True)
                B["h2d"].record(h2d_s)
            cur.wait_event(B["h2d"])
            if i >= NB:
                cur.wait_event(B["d2h"])               # output buffer of step i-NB has been read back
            fa.flash_attn_varlen_func(B["q"], B["k"], B["v"], cu, cu, max_s, max_s, causal=True, out=B["o"])
            B["comp"].record(cur)
            with torch.cuda.stream(d2h_s):
                d2h_s.wait_event(B["comp"])
                out_h.copy_(B["o"], non_blocking=True)
                B["d2h"].record(d2h_s)
        for B in bufs[: min(n, NB)]:
            cur.wait_event(B["d2h"])

    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2e_loop(2); barrier()
    s0.record()
    e2e_loop(args.steps)
    s1.record()
    barrier()
    ms_e2e_total = s0.elapsed_time(s1)
    clocks = sampler.stop() if rank == 0 else None

    t = torch.tensor([ms_total, ms_e2e_total], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step, ms_e2e = t[0].item() / args.steps, t[1].item() / args.steps
    value = world * flops / (ms_step * 1e-3) / 1e12
    e2e_value = world * flops / (ms_e2e * 1e-3) / 1e12
    per_gpu_tflops = flops / (ms_step * 1e-3) / 1e12

    line = {"metric": "attn_fwd_tflops_prefill", "value": value, "unit": "TFLOP/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "C2 Llama-3-8B prefill: flash_attn_varlen_func causal GQA 32q/8kv D=128, 16384 tokens, "
                                   "seqlens " + str(SEQLENS_C2) + (" per GPU (head-sharded replicas)" if world > 1 else ""),
                       "l2": "inputs+outputs 320 MiB per step > 126 MB L2", "kernel": {1: "tcgen05_prefill", 2: "tile_mma_sync", 3: "decode"}.get(kernel_used)},
            "e2e": {"value": e2e_value, "unit": "TFLOP/s", "h2d_bytes_per_step": (qh.numel() + kh.numel() + vh.numel()) * 2,
                    "d2h_bytes_per_step": out_h.numel() * 2, "ms_per_step": ms_e2e,
                    "note": "pinned host q,k,v -> device, attention, out -> pinned host every step; 3 streams, double-buffered"},
            "gpu_launches": launches_per_step * args.steps * 2,   # kernel-only region + e2e region
            "roofline": {"bound": "tensor", "achieved": per_gpu_tflops, "peak": pk["bf16_tflops"], "unit": "TFLOP/s",
                         "frac": per_gpu_tflops / pk["bf16_tflops"], "traffic": ncu_traffic("prefill_c2"), "peak_source": pk_src + " burst (cuBLAS bf16)",
                         "frac_of_sustained": per_gpu_tflops / pk["bf16_tflops_sustained"], "frac_of_nominal_2250": per_gpu_tflops / 2250.0,
                         "flops_per_launch": flops},
            "clocks": clocks}
    del bufs

    # ------------------------------------------------------------------ paged decode (C3), rank-local
    if args.decode:
        line["decode"] = bench_decode(fa, ext, dev, pk, pk_src, args)
    if rank == 0 and world == 1 and not args.no_cpu:
        threads = os.cpu_count() or 1
        val, desc = cpu_attention_sample(threads, budget_s=15.0)
        line["cpu_baseline"] = {"value": val, "unit": "TFLOP/s", "cores": threads, "kind": "port", "sample": desc}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def bench_decode(fa, ext, dev, pk, pk_src, args, b=256, ctx=4096, page=16):
    """C3: flash_attn_with_kvcache, paged KV block 16, batch 256, context 4k, GQA 32q/8kv, D=128.
    Algorithmic bytes = K+V read once = 2*b*ctx*hk*d*2 B."""
    import torch
    g = torch.Generator(device=dev).manual_seed(0)
    blocks_per_seq = ctx // page
    nblk = b * blocks_per_seq * 5 // 4
    kc = torch.randn(nblk, page, HK, D, device=dev, dtype=torch.bfloat16, generator=g)
    vc = torch.randn(nblk, page, HK, D, device=dev, dtype=torch.bfloat16, generator=g)
    bt = torch.randperm(nblk, device=dev, generator=g)[: b * blocks_per_seq].to(torch.int32).view(b, -1).contiguous()
    q = torch.randn(b, 1, H, D, device=dev, dtype=torch.bfloat16, generator=g)
    lens = torch.full((b,), ctx, dtype=torch.int32, device=dev)
    out = torch.empty_like(q)
    f = lambda: fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=lens, block_table=bt, causal=True, out=out)
    for _ in range(max(args.warmup, 3)):
        f()
    info = dict(ext.last_call_info)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        f()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    nbytes = 2.0 * b * ctx * HK * D * 2
    gbs = nbytes / (ms * 1e-3) / 1e9
    return {"metric": "paged_kv_decode_hbm_gbs", "value": gbs, "unit": "GB/s", "ms_per_step": ms,
            "config": {"workload": f"C3 Llama-3-8B decode: flash_attn_with_kvcache paged block {page}, batch {b}, ctx {ctx}, GQA 32q/8kv D=128",
                       "l2": "KV pool 5.4 GB >> L2"},
            "kernel_used": info.get("kernel_used"), "num_splits": info.get("num_splits"),
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": gbs / pk["hbm_gbs"],
                         "traffic": ncu_traffic("decode_c3"), "peak_source": pk_src + " (copy)", "frac_of_nominal_8000": gbs / 8000.0,
                         "bytes_per_launch": nbytes}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-decode", dest="decode", action="store_false")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
