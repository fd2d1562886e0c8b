"""Per-call host cost of the small-shape calls (C1: B=2 S=512 H=8 D=64 dense; a 16-sequence paged decode step) through
the two host bindings — the compiled pybind11 / ATen module and the Python-over-ctypes shim — next to the device time.
Development tool.   python tools/c1_latency.py [--profile]"""
import json, sys, time, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vllm-flash-attention_b200"))
import torch
import vllm_flash_attn as fa
from vllm_flash_attn import vllm_flash_attn_2_cuda as ext

q = torch.randn(2, 512, 8, 64, device="cuda", dtype=torch.bfloat16); k = torch.randn_like(q); v = torch.randn_like(q)
page, mb, b = 16, 64, 16
kc = torch.randn(b * mb, page, 8, 128, device="cuda", dtype=torch.bfloat16); vc = torch.randn_like(kc)
bt = torch.randperm(b * mb, device="cuda").to(torch.int32).view(b, mb).contiguous()
qd = torch.randn(b, 1, 32, 128, device="cuda", dtype=torch.bfloat16); cl = torch.full((b,), 1000, dtype=torch.int32, device="cuda")
import math
kn = torch.randn(b, 1, 8, 128, device="cuda", dtype=torch.bfloat16); vn = torch.randn_like(kn)
ang = torch.rand(mb * page, 64, device="cuda") * 2 * math.pi
cos, sin = torch.cos(ang).to(torch.bfloat16), torch.sin(ang).to(torch.bfloat16)
append = lambda: fa.flash_attn_with_kvcache(qd, kc, vc, kn, vn, rotary_cos=cos, rotary_sin=sin, cache_seqlens=cl, block_table=bt, causal=True)
calls = {"c1_flash_attn_func": lambda: fa.flash_attn_func(q, k, v),
         "decode_step_flash_attn_with_kvcache": lambda: fa.flash_attn_with_kvcache(qd, kc, vc, cache_seqlens=cl, block_table=bt, causal=True)}

def host_us(f, n=3000):
    for _ in range(50): f()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n): f()
    t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
    return 1e6 * (t1 - t0) / n, 1e6 * (t2 - t0) / n

def device_us(f):
    s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3): f()
    torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(20): f()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): g.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / 400

native = ext._N
for name, f in calls.items():
    row = {"call": name, "device_us": round(device_us(f), 2)}
    for binding in ("pybind11", "ctypes"):
        if binding == "pybind11" and native is None: continue
        ext._N = native if binding == "pybind11" else None
        enq, wall = host_us(f)
        row[binding + "_host_enqueue_us"] = round(enq, 2); row[binding + "_wall_us"] = round(wall, 2)
    ext._N = native
    print(json.dumps(row), flush=True)
# decode step that also appends the step's k/v with rotary: in-kernel append (one launch) vs the stand-alone append kernel first
for two_step in (False, True):
    ext._two_step_append = two_step
    enq, wall = host_us(append)
    print(json.dumps({"call": "decode_step_with_append_and_rotary", "append": "stand-alone kernel first" if two_step else "in the decode kernel",
                      "binding": "ctypes" if two_step or native is None else "pybind11", "device_us": round(device_us(append), 2),
                      "host_enqueue_us": round(enq, 2), "wall_us": round(wall, 2)}), flush=True)
ext._two_step_append = False
if "--profile" in sys.argv:
    import cProfile, pstats
    pr = cProfile.Profile(); pr.enable()
    for _ in range(500): calls["c1_flash_attn_func"]()
    pr.disable(); torch.cuda.synchronize()
    pstats.Stats(pr).sort_stats("tottime").print_stats(12)
