#!/usr/bin/env python
"""bench.py — headline benchmark of the B200-native flash-attention forward library.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload prefill|decode]

A "step" = one pass of the hot path over one batch of synthetic input.
N=1 workload (BASELINE.json configs[1], "C2"): Llama-3-8B prefill — flash_attn_varlen_func, causal,
GQA 32q/8kv heads, D=128, 16 384 tokens over seqlens [8192,4096,2048,1024,512,256,128,128], bf16.
metric = attention forward TFLOP/s, FLOPs = 2*H*D*sum(s_i^2) (= 4*s^2*H*D/2 causal; reference
benchmarks/benchmark_flash_attention.py:27-30).  N>1: head-sharded replicas of the same per-GPU
work (each rank owns 32q/8kv heads of a 32N-head layer; no collective) -> weak scaling.
The JSON line also carries the paged-KV decode figure (C3: B=256, ctx 4096, block 16) under "decode" and, at
every N, the two shardings north_star names, strong-scaled against a 1-GPU run made in the same process:
"c4" (64q/8kv heads head-sharded, mixed prefill+decode varlen batch, no collective) and "c5" (131072-token
decode, KV sequence split over the ranks, (out,lse) exchanged by the fused peer-memory kernel or by one NCCL
all-gather, then LSE-combined), each with parity_ok against the single-GPU result.
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


def cpu_attention_sample(threads, budget_s=60.0):
    """The reference's PyTorch attention_ref (our oracle port, oracle/attention_oracle.py) timed on the
    host cores over the WHOLE C2 batch: every sequence, all 32 heads, causal, fp32 upcast.  attention_ref
    materialises (B,H,Sq,Sk) scores, so it runs per sequence and, from 2048 tokens up, per KV-head group
    (4 query heads + their KV head: 1 GiB of fp32 scores at 8192 tokens) — SURVEY 8d's chunking.  If the
    budget runs out the remaining sequences are skipped and the description says which ones ran.
    Returns (tflops, description)."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import attention_oracle as orc
    torch.set_num_threads(threads)
    g = torch.Generator().manual_seed(0)
    done_flops, t_total, used = 0.0, 0.0, []
    grp = H // HK
    for s in sorted(SEQLENS_C2):
        q = torch.randn(1, s, H, D, generator=g).to(torch.bfloat16)
        k = torch.randn(1, s, HK, D, generator=g).to(torch.bfloat16)
        v = torch.randn(1, s, HK, D, generator=g).to(torch.bfloat16)
        t0 = time.perf_counter()
        if s >= 2048:
            for kh in range(HK):
                orc.attention_ref(q[:, :, kh * grp:(kh + 1) * grp], k[:, :, kh:kh + 1], v[:, :, kh:kh + 1], causal=True)
        else:
            orc.attention_ref(q, k, v, causal=True)
        dt = time.perf_counter() - t0
        done_flops += prefill_flops([s]); t_total += dt; used.append(s)
        if t_total > budget_s:
            break
    whole = len(used) == len(SEQLENS_C2)
    desc = ("attention_ref (fp32 upcast) over " + ("the whole C2 batch" if whole else f"sequences {used} of the C2 batch") +
            ", 32q/8kv heads, causal; per sequence, and per KV-head group from 2048 tokens up")
    return done_flops / t_total / 1e12, desc


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path (PyTorch attention_ref,
    restated in oracle/attention_oracle.py) on the host cores; each step = a bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    threads = os.cpu_count() or 1
    for _ in range(min(args.warmup, 1)):
        cpu_attention_sample(threads, budget_s=1.0)
    vals, desc = [], ""
    t0 = time.perf_counter()
    for _ in range(args.steps):
        v, desc = cpu_attention_sample(threads, budget_s=60.0)
        vals.append(v)
    ms = (time.perf_counter() - t0) / max(args.steps, 1) * 1e3
    val = sum(vals) / len(vals)
    line = {"impl": "reference", "metric": "attn_fwd_tflops_prefill", "value": val, "unit": "TFLOP/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "C2 Llama-3-8B prefill: flash_attn_varlen_func causal GQA 32q/8kv D=128, 16384 tokens, "
                                   "seqlens " + str(SEQLENS_C2),
                       "sample": "each step = one pass over that workload on the host cores: " + desc},
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
                    "note": "pinned host q,k,v -> device, attention, out -> pinned host every step; 3 streams, double-buffered" +
                            ("; at N>1 every rank pushes its 335 MB per step through the host's PCIe roots at once, so this figure is "
                             "bound by host PCIe / NUMA bandwidth, not by the kernels" if world > 1 else "")},
            "gpu_launches": launches_per_step * args.steps * 2,   # kernel-only region + e2e region
            "roofline": {"bound": "tensor", "achieved": per_gpu_tflops, "peak": pk["bf16_tflops"], "unit": "TFLOP/s",
                         "frac": per_gpu_tflops / pk["bf16_tflops"], "traffic": ncu_traffic("prefill_c2"), "peak_source": pk_src + " burst (cuBLAS bf16)",
                         "frac_of_nominal_2250": per_gpu_tflops / 2250.0,
                         "flops_per_launch": flops},
            "clocks": clocks}
    del bufs
    # ------------------------------------------------------------------ paged decode (C3), rank-local
    if args.decode:
        line["decode"] = bench_decode(fa, ext, dev, pk, pk_src, args)
    # ------------------------------------------------------------------ the two shardings (C4, C5), every N
    if args.sharded:
        torch.cuda.empty_cache()
        only = set(x for x in args.only.split(",") if x)
        if not only or "c4" in only:
            line["c4"] = bench_c4(fa, ext, dev, world, rank, pk)
        torch.cuda.empty_cache()
        line["c5"] = {}
        if not only or "c5b1" in only:
            line["c5"]["b1"] = bench_c5(fa, ext, dev, world, rank, pk, batch=1)
        if not only or "c5b8" in only:
            line["c5"]["b8"] = bench_c5(fa, ext, dev, world, rank, pk, batch=8)
        torch.cuda.empty_cache()
    if rank == 0 and world == 1 and args.ref_kernel:
        torch.cuda.empty_cache()
        line["ref_kernel"] = bench_ref_kernel(fa, dev)
        torch.cuda.empty_cache()
    # ---- the headline step in the SUSTAINED regime (last leg, so that the throttled state it leaves behind touches no other number): back to back for >= 2 s (the GPU settles at its power cap after ~0.1 s;
    # the 20-step region above is the burst regime), against the sustained cuBLAS peak of MEASURED_PEAKS.json
    if world == 1 and args.sustained and pk.get("bf16_tflops_sustained"):
        n = max(int(2.2e3 / max(ms_total / args.steps, 1e-3)), 100)
        for _ in range(n // 10):
            step()
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record()
        for _ in range(n):
            step()
        s1.record(); torch.cuda.synchronize()
        ms_s = s0.elapsed_time(s1) / n
        line["sustained"] = {"value": flops / (ms_s * 1e-3) / 1e12, "unit": "TFLOP/s", "ms_per_step": ms_s, "steps": n,
                             "seconds": s0.elapsed_time(s1) / 1e3, "peak": pk["bf16_tflops_sustained"],
                             "frac": flops / (ms_s * 1e-3) / 1e12 / pk["bf16_tflops_sustained"],
                             "peak_source": pk_src + " sustained (cuBLAS bf16, seconds-long loop under the power cap)"}

    if rank == 0 and world == 1 and not args.no_cpu:
        threads = os.cpu_count() or 1
        val, desc = cpu_attention_sample(threads, budget_s=30.0)
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


def bench_ref_kernel(fa, dev, iters=20):
    """BASELINE.md section 3's "reference kernel on B200" column, driver-run: the REFERENCE's own FA2 forward kernels,
    compiled for sm_100 from /root/reference by oracle/Makefile into oracle/_ref (test infrastructure: loaded only
    inside this leg, outside the product's timed regions — like cpu_baseline), next to ours on the same box and the
    same tensors: C1 (latency), C2 (TFLOP/s), C3 (GB/s).  None when the library was not built."""
    import torch
    import _util
    if _util.ref_lib() is None:
        return None

    def t(f, n=iters):
        for _ in range(5):
            f()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            f()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n

    dt = torch.bfloat16
    g = torch.Generator(device=dev).manual_seed(0)
    res = {"what": "reference FA2 kernels (csrc/flash_attn/src, mma.sync) recompiled for sm_100 from /root/reference, same box, same tensors; "
                   "CUDA events over back-to-back calls through each library's Python entry"}
    q = torch.randn(2, 512, 8, 64, device=dev, dtype=dt, generator=g); k = torch.randn_like(q); v = torch.randn_like(q)
    ms_o, ms_r = t(lambda: fa.flash_attn_func(q, k, v), 200), t(lambda: _util.ref_kernel_fwd(q, k, v), 200)
    res["c1"] = {"workload": "C1 dense B=2 S=512 H=8 D=64 non-causal", "ours_us_per_call": ms_o * 1e3, "ref_kernel_us_per_call": ms_r * 1e3,
                 "speedup": ms_r / ms_o}
    tot = sum(SEQLENS_C2)
    q = torch.randn(tot, H, D, device=dev, dtype=dt, generator=g)
    k = torch.randn(tot, HK, D, device=dev, dtype=dt, generator=g); v = torch.randn(tot, HK, D, device=dev, dtype=dt, generator=g)
    cu = torch.tensor([0] + list(__import__("itertools").accumulate(SEQLENS_C2)), dtype=torch.int32, device=dev)
    mx = max(SEQLENS_C2)
    ms_o = t(lambda: fa.flash_attn_varlen_func(q, k, v, cu, cu, mx, mx, causal=True))
    ms_r = t(lambda: _util.ref_kernel_varlen(q, k, v, cu, cu, mx, mx, causal=True))
    fl = prefill_flops()
    res["c2"] = {"workload": "C2 varlen causal 32q/8kv D=128, 16384 tokens", "ours_tflops": fl / (ms_o * 1e-3) / 1e12,
                 "ref_kernel_tflops": fl / (ms_r * 1e-3) / 1e12, "ours_ms": ms_o, "ref_kernel_ms": ms_r, "speedup": ms_r / ms_o}
    del q, k, v
    b, ctx, page = 256, 4096, 16
    nblk = b * (ctx // page) * 5 // 4
    kc = torch.randn(nblk, page, HK, D, device=dev, dtype=dt, generator=g); vc = torch.randn(nblk, page, HK, D, device=dev, dtype=dt, generator=g)
    bt = torch.randperm(nblk, device=dev, generator=g)[: b * (ctx // page)].to(torch.int32).view(b, -1).contiguous()
    q = torch.randn(b, 1, H, D, device=dev, dtype=dt, generator=g)
    lens = torch.full((b,), ctx, dtype=torch.int32, device=dev)
    ms_o = t(lambda: fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=lens, block_table=bt, causal=True))
    ms_r = t(lambda: _util.ref_kernel_kvcache(q, kc, vc, lens, bt, causal=True))
    nb = 2.0 * b * ctx * HK * D * 2
    res["c3"] = {"workload": "C3 paged decode B=256 ctx 4096 block 16", "ours_gbs": nb / (ms_o * 1e-3) / 1e9,
                 "ref_kernel_gbs": nb / (ms_r * 1e-3) / 1e9, "ours_ms": ms_o, "ref_kernel_ms": ms_r, "speedup": ms_r / ms_o}
    return res


# ------------------------------------------------------------------------------------------------
# The two shardings north_star names, measured at every N (strong scaling: the total work is fixed).
def _timed(f, iters, world, dev, warmup=5):
    """CUDA events on the launching stream, barrier + synchronize on both sides, max over ranks -> ms per call."""
    import torch
    import torch.distributed as dist
    for _ in range(warmup):
        f()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        f()
    e1.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def _timed_graph(f, iters, world, dev):
    """The same measurement with the call captured in a CUDA graph (how a serving loop replays decode steps).
    Every rank captures, so a failure on one rank is agreed on before anyone replays (no stuck collective)."""
    import torch
    import torch.distributed as dist
    g, ok = None, 1
    try:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                f()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            f()
    except Exception as ex:  # noqa: BLE001
        print("graph capture failed:", repr(ex)[:300], file=sys.stderr, flush=True)
        ok = 0
    t = torch.tensor([ok], device=dev, dtype=torch.int32)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if t.item() == 0:
        return None
    return _timed(g.replay, iters, world, dev)


def bench_c4(fa, ext, dev, world, rank, pk, iters=20):
    """C4 (BASELINE.json configs[3]): Llama-3-70B attention, 64q/8kv heads, D=128, head-sharded over
    the ranks (no collective); ONE mixed varlen call through flash_attn_varlen_func(block_table=...): 4 prompts of
    2048 tokens + 192 single-token decodes with contexts randint(1024, 4097), paged block 16, seed 0.
    Strong scaling: t(1 GPU, all 64 heads) is measured on rank 0 in the same run."""
    import torch
    import torch.distributed as dist
    from vllm_flash_attn import sharded
    Hq, Hk, page, dt = 64, 8, 16, torch.bfloat16
    n_pre, s_pre, n_dec = 4, 2048, 192
    gc = torch.Generator(device="cpu").manual_seed(0)
    dec_ctx = torch.randint(1024, 4097, (n_dec,), generator=gc).tolist()
    q_lens, k_lens = [s_pre] * n_pre + [1] * n_dec, [s_pre] * n_pre + dec_ctx
    acc = lambda xs: [0] + list(__import__("itertools").accumulate(xs))
    cq = torch.tensor(acc(q_lens), dtype=torch.int32, device=dev)
    ck = torch.tensor(acc(k_lens), dtype=torch.int32, device=dev)
    tq, nb = sum(q_lens), len(q_lens)
    max_blocks = (max(k_lens) + page - 1) // page
    nblk = nb * max_blocks
    g = torch.Generator(device=dev).manual_seed(0)            # same Philox stream on every rank's GPU: same tensors
    q_all = torch.randn(tq, Hq, D, device=dev, dtype=dt, generator=g)
    kc_all = torch.randn(nblk, page, Hk, D, device=dev, dtype=dt, generator=g)
    vc_all = torch.randn(nblk, page, Hk, D, device=dev, dtype=dt, generator=g)
    bt = torch.randperm(nblk, device=dev, generator=g).to(torch.int32).view(nb, max_blocks).contiguous()
    qs, ks = sharded.head_shard(rank, world, Hq, Hk)
    q_r, kc_r, vc_r = q_all[:, qs].contiguous(), kc_all[:, :, ks].contiguous(), vc_all[:, :, ks].contiguous()
    mq, mk = max(q_lens), max(k_lens)
    call = lambda q_, k_, v_: fa.flash_attn_varlen_func(q_, k_, v_, cq, ck, mq, mk, causal=True, block_table=bt)
    # ---- single-GPU reference time and result (rank 0, all heads) ----
    t1, o_full = None, None
    if rank == 0:
        o_full = call(q_all, kc_all, vc_all)
    t1 = _timed(lambda: call(q_all, kc_all, vc_all), iters, 1, dev) if rank == 0 else 0.0
    t1g = (_timed_graph(lambda: call(q_all, kc_all, vc_all), iters, 1, dev) or 0.0) if rank == 0 else 0.0
    if world > 1:
        if rank != 0:
            del q_all, kc_all, vc_all
        tt = torch.tensor([t1, t1g], device=dev, dtype=torch.float64)
        dist.broadcast(tt, 0)
        t1, t1g = tt[0].item(), tt[1].item()
    # ---- sharded run ----
    o_r = call(q_r, kc_r, vc_r)
    launches = ext.last_call_info.get("launches")
    parity = None
    if world > 1:
        parts = [torch.empty_like(o_r) for _ in range(world)]
        dist.all_gather(parts, o_r)                # check only: the product path has no collective
        if rank == 0:
            parity = (torch.cat(parts, dim=1).float() - o_full.float()).abs().max().item()
    elif rank == 0:
        parity = (o_r.float() - o_full.float()).abs().max().item()
    tn_e = _timed(lambda: call(q_r, kc_r, vc_r), iters, world, dev)
    tn_g = _timed_graph(lambda: call(q_r, kc_r, vc_r), iters, world, dev)     # the mixed call (side-stream fork included) is capture-safe
    tn = min(x for x in (tn_e, tn_g) if x)
    t1_e, t1 = t1, min(x for x in (t1, t1g) if x)                             # like C5: the better of eager / CUDA graph on both sides
    # the two parts alone (same shard): prefill sequences only / decode sequences only
    cq_p, ck_p = cq[: n_pre + 1].contiguous(), ck[: n_pre + 1].contiguous()
    t_pre = _timed(lambda: fa.flash_attn_varlen_func(q_r[: n_pre * s_pre], kc_r, vc_r, cq_p, ck_p, s_pre, s_pre, causal=True,
                                                     block_table=bt[:n_pre]), iters, world, dev)
    cq_d = (cq[n_pre:] - cq[n_pre]).contiguous(); ck_d = (ck[n_pre:] - ck[n_pre]).contiguous()
    t_dec = _timed(lambda: fa.flash_attn_varlen_func(q_r[n_pre * s_pre:], kc_r, vc_r, cq_d, ck_d, 1, max(dec_ctx), causal=True,
                                                     block_table=bt[n_pre:]), iters, world, dev)
    flops = 2.0 * Hq * D * n_pre * s_pre * s_pre                   # whole job, causal
    nbytes = 2.0 * float(sum(dec_ctx)) * Hk * D * 2                # whole job, K+V of the decode sequences once
    return {"workload": "C4 Llama-3-70B attention 64q/8kv D=128 head-sharded, one flash_attn_varlen_func(block_table) call: "
                        "4x2048-token prompts + 192 decodes (ctx randint(1024,4097)), paged block 16, seed 0",
            "n_gpus": world, "heads_per_gpu": [Hq // world, Hk // world], "scaling": "strong",
            "ms": tn, "ms_eager": tn_e, "ms_cuda_graph": tn_g, "ms_1gpu_all_heads": t1, "ms_1gpu_all_heads_eager": t1_e,
            "ms_1gpu_all_heads_cuda_graph": t1g or None, "speedup_vs_1gpu": t1 / tn, "strong_scaling_efficiency": t1 / (world * tn),
            "ms_prefill_part_alone": t_pre, "ms_decode_part_alone": t_dec,
            "prefill_part_tflops_per_gpu": flops / world / (t_pre * 1e-3) / 1e12,
            "decode_part_gbs_per_gpu": nbytes / world / (t_dec * 1e-3) / 1e9,
            "prefill_part_frac_of_peak": flops / world / (t_pre * 1e-3) / 1e12 / pk["bf16_tflops"],
            "decode_part_frac_of_peak": nbytes / world / (t_dec * 1e-3) / 1e9 / pk["hbm_gbs"],
            "launches_per_call": launches, "collective": "none (heads independent)",
            # heads are independent, so the prefill rows are bit-identical to the unsharded call; the decode rows may
            # differ by one bf16 rounding (4e-3 for |x| < 1): a shard with few KV heads splits its decode half along
            # the context (fa_api.cu), the all-heads call on one GPU does not — a different fp32 summation tree
            "parity_max_abs_vs_1gpu": parity, "parity_tolerance": 4e-3,
            "parity_ok": (parity is not None and parity <= 4e-3) if rank == 0 else None}


def bench_c5(fa, ext, dev, world, rank, pk, batch, ctx=131072, iters=30):
    """C5 (BASELINE.json configs[4]): long-context decode, 32q/8kv D=128, ctx 131072, paged block 16,
    KV sequence split over the ranks (contiguous ctx/N keys per rank); per-rank (out fp32, lse) partials are
    exchanged and LSE-combined (flash_fwd_kernel.h:1151-1221 lifted to ranks) (a) inside the decode kernel itself (ONE
    launch per step: split merge + NVLink push + wait + combine), (b) by a separate peer-memory kernel
    (fa_b200_exchange_combine) or (c) by ONE NCCL all-gather + the combine kernel.
    Strong scaling: t(1 GPU, whole sequence) is measured on rank 0 in the same run, eager and as a CUDA graph."""
    import torch
    import torch.distributed as dist
    from vllm_flash_attn import sharded
    page, dt = 16, torch.bfloat16
    g = torch.Generator(device=dev).manual_seed(1234 + batch)
    q = torch.randn(batch, 1, H, D, device=dev, dtype=dt, generator=g)
    lens_h = [ctx] * batch
    if batch > 1:
        lens_h[1] = ctx // 3 + 5                   # one ragged sequence: some ranks hold none of it
    lens = torch.tensor(lens_h, dtype=torch.int32, device=dev)
    ranges = [sharded.seq_shard_range(r, world, ctx, page) for r in range(world)]
    ranges = [(s, min(e, ctx)) for s, e in ranges]

    def shard_kv(r, which):                        # keys [s, e) of every sequence, reproducible on any rank's GPU
        s, e = ranges[r]
        gg = torch.Generator(device=dev).manual_seed(100000 * which + 10 * batch + r)
        return torch.randn(batch, e - s, HK, D, device=dev, dtype=dt, generator=gg)

    s, e = ranges[rank]
    L = e - s
    nblk = batch * (L // page)
    gp = torch.Generator(device=dev).manual_seed(77 + rank)
    bt = torch.randperm(nblk, device=dev, generator=gp).to(torch.int32).view(batch, L // page).contiguous()
    kc = torch.empty(nblk, page, HK, D, device=dev, dtype=dt); vc = torch.empty_like(kc)
    kc[bt.view(-1).long()] = shard_kv(rank, 0).view(nblk, page, HK, D)
    vc[bt.view(-1).long()] = shard_kv(rank, 1).view(nblk, page, HK, D)
    nbytes = 2.0 * float(sum(lens_h)) * HK * D * 2             # whole job: K+V of every sequence once

    # ---- single-GPU reference (rank 0): the whole sequence in one paged pool, auto split-KV ----
    t1 = t1g = 0.0
    o1 = l1 = None
    splits1 = None
    if rank == 0:
        nb1 = batch * (ctx // page)
        bt1 = torch.randperm(nb1, device=dev, generator=gp).to(torch.int32).view(batch, ctx // page).contiguous()
        kf = torch.empty(nb1, page, HK, D, device=dev, dtype=dt); vf = torch.empty_like(kf)
        kf[bt1.view(-1).long()] = torch.cat([shard_kv(r, 0) for r in range(world)], dim=1).view(nb1, page, HK, D)
        vf[bt1.view(-1).long()] = torch.cat([shard_kv(r, 1) for r in range(world)], dim=1).view(nb1, page, HK, D)
        f1 = lambda: fa.flash_attn_with_kvcache(q, kf, vf, cache_seqlens=lens, block_table=bt1, causal=True, return_softmax_lse=True)
        o1, l1 = f1()
        splits1 = ext.last_call_info.get("num_splits")
        t1 = _timed(f1, iters, 1, dev)
        t1g = _timed_graph(f1, iters, 1, dev) or 0.0
        del kf, vf
    if world > 1:
        tt = torch.tensor([t1, t1g], device=dev, dtype=torch.float64)
        dist.broadcast(tt, 0)
        t1, t1g = tt[0].item(), tt[1].item()
    res = {"workload": f"C5 long-context decode: flash_attn_with_kvcache, B={batch}, ctx {ctx}, 32q/8kv D=128, paged block {page}, "
                       "KV sequence split over the ranks + (out,lse) exchange + LSE combine",
           "n_gpus": world, "batch": batch, "scaling": "strong", "bytes_whole_job": nbytes,
           "ms_1gpu": t1, "ms_1gpu_cuda_graph": t1g or None, "num_splits_1gpu": splits1,
           "gbs_1gpu": nbytes / (t1 * 1e-3) / 1e9, "gbs_1gpu_cuda_graph": (nbytes / (t1g * 1e-3) / 1e9) if t1g else None,
           "frac_of_hbm_peak_1gpu_cuda_graph": (nbytes / (t1g * 1e-3) / 1e9 / pk["hbm_gbs"]) if t1g else None}
    if world == 1:
        res.update({"ms": t1, "ms_cuda_graph": t1g or None, "parity_ok": True, "collective": "none (one GPU: in-GPU split-KV + combine)"})
        return res

    # ---- sharded: (a) fused peer-memory exchange, (b) NCCL all-gather + combine ----
    xch = sharded.PeerExchange.for_decode(batch, H, HK, D, dev)
    f_one = lambda: sharded.flash_attn_with_kvcache_seq_sharded(q, kc, vc, lens, bt, shard_start=s, shard_end=e, causal=True,
                                                                 return_softmax_lse=True, exchange=xch, one_kernel=True)
    f_fused = lambda: sharded.flash_attn_with_kvcache_seq_sharded(q, kc, vc, lens, bt, shard_start=s, shard_end=e, causal=True,
                                                                   return_softmax_lse=True, exchange=xch, one_kernel=False)
    f_nccl = lambda: sharded.flash_attn_with_kvcache_seq_sharded(q, kc, vc, lens, bt, shard_start=s, shard_end=e, causal=True,
                                                                  return_softmax_lse=True)
    oo, lo = f_one()
    launches_one = ext.last_call_info.get("launches")
    of, lf = f_fused()
    on, ln = f_nccl()
    splits_n = ext.last_call_info.get("num_splits")
    torch.cuda.synchronize()
    par = None
    if rank == 0:
        par = {"one_kernel_out": (oo.float() - o1.float()).abs().max().item(), "one_kernel_lse": (lo - l1).abs().max().item(),
               "fused_out": (of.float() - o1.float()).abs().max().item(), "fused_lse": (lf - l1).abs().max().item(),
               "nccl_out": (on.float() - o1.float()).abs().max().item(), "nccl_lse": (ln - l1).abs().max().item()}
    t_o = _timed(f_one, iters, world, dev)
    t_og = _timed_graph(f_one, iters, world, dev)
    t_f = _timed(f_fused, iters, world, dev)
    t_n = _timed(f_nccl, iters, world, dev)
    t_fg = _timed_graph(f_fused, iters, world, dev)
    t_ng = _timed_graph(f_nccl, iters, world, dev)
    lens_local = sharded.seq_shard_lengths(lens, s, e)
    t_loc = _timed_graph(lambda: sharded.kvcache_partial(q, kc, vc, lens_local, bt), iters, world, dev)
    best = min(x for x in (t_o, t_og, t_f, t_n, t_fg, t_ng) if x)
    t1_best = min(x for x in (t1, t1g) if x)
    res.update({
        "ms_one_kernel": t_o, "ms_one_kernel_cuda_graph": t_og, "launches_one_kernel": launches_one,
        "ms_fused_exchange": t_f, "ms_nccl_allgather": t_n, "ms_fused_exchange_cuda_graph": t_fg, "ms_nccl_allgather_cuda_graph": t_ng,
        "ms_local_split_decode_only_cuda_graph": t_loc, "num_splits_per_rank": splits_n,
        "ms": best, "speedup_vs_1gpu": t1_best / best, "strong_scaling_efficiency": t1_best / (world * best),
        "gbs_per_gpu": nbytes / world / (best * 1e-3) / 1e9, "frac_of_hbm_peak_per_gpu": nbytes / world / (best * 1e-3) / 1e9 / pk["hbm_gbs"],
        "gbs_per_gpu_local_kernel_only": (nbytes / world / (t_loc * 1e-3) / 1e9) if t_loc else None,
        "exchange_bytes_per_rank": sharded.partial_record_numel(batch, 1, H, D) * 4 * (world - 1),
        "collective": "one_kernel: fa_decode_kernel merges its splits, pushes to the peers over NVLink, waits and combines (1 launch) | "
                      "fused_exchange: decode + fa_b200_exchange_combine (2 launches) | nccl: decode + all_gather_into_tensor + fa_b200_combine",
        "parity_max_abs_vs_1gpu": par,
        # bf16 output of two different split trees (ranks x splits vs splits): one bf16 ulp at |o| < 4
        "parity_ok": (max(par["one_kernel_out"], par["fused_out"], par["nccl_out"]) <= 2e-2 and
                      max(par["one_kernel_lse"], par["fused_lse"], par["nccl_lse"]) <= 1e-3) if rank == 0 else None})
    del xch
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-decode", dest="decode", action="store_false")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-sustained", dest="sustained", action="store_false", help="skip the 2 s sustained-regime leg (N=1)")
    ap.add_argument("--no-ref-kernel", dest="ref_kernel", action="store_false", help="skip the reference-kernel comparison leg (N=1)")
    ap.add_argument("--no-sharded", dest="sharded", action="store_false", help="skip the C4 / C5 sub-benchmarks")
    ap.add_argument("--only", default="", help="development: comma list of c4,c5b1,c5b8 — run only those sub-benchmarks of the sharded section")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
