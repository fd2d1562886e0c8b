"""C4's per-rank shard at TP = 1/2/4/8 on ONE GPU (heads sliced, no collective needed): whole mixed call, the two
parts alone, and the per-kernel launch list when run under ncu.  Development tool.
  python tools/c4_shard_probe.py [--tp 8] [--iters 50]"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vllm-flash-attention_b200"))
import torch
import vllm_flash_attn as fa
from vllm_flash_attn import vllm_flash_attn_2_cuda as ext

def timeit(f, iters, graph=False):
    for _ in range(5): f()
    torch.cuda.synchronize()
    if graph:
        s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            f()
        torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g): f()
        f = g.replay
        for _ in range(3): f()
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters

ap = argparse.ArgumentParser(); ap.add_argument("--tp", type=int, nargs="*", default=[1, 2, 4, 8]); ap.add_argument("--iters", type=int, default=50)
a = ap.parse_args()
dev, dt, D, page = "cuda", torch.bfloat16, 128, 16
for tp in a.tp:
    Hq, Hk = 64 // tp, 8 // tp
    n_pre, s_pre, n_dec = 4, 2048, 192
    gc = torch.Generator(device="cpu").manual_seed(0)
    dec_ctx = torch.randint(1024, 4097, (n_dec,), generator=gc).tolist()
    q_lens, k_lens = [s_pre] * n_pre + [1] * n_dec, [s_pre] * n_pre + dec_ctx
    import itertools
    cq = torch.tensor([0] + list(itertools.accumulate(q_lens)), dtype=torch.int32, device=dev)
    ck = torch.tensor([0] + list(itertools.accumulate(k_lens)), dtype=torch.int32, device=dev)
    nb = len(q_lens); mb = (max(k_lens) + page - 1) // page; nblk = nb * mb
    q = torch.randn(sum(q_lens), Hq, D, device=dev, dtype=dt)
    kc = torch.randn(nblk, page, Hk, D, device=dev, dtype=dt); vc = torch.randn_like(kc)
    bt = torch.randperm(nblk, device=dev).to(torch.int32).view(nb, mb).contiguous()
    whole = lambda: fa.flash_attn_varlen_func(q, kc, vc, cq, ck, max(q_lens), max(k_lens), causal=True, block_table=bt)
    cq_p, ck_p = cq[: n_pre + 1].contiguous(), ck[: n_pre + 1].contiguous()
    pre = lambda: fa.flash_attn_varlen_func(q[: n_pre * s_pre], kc, vc, cq_p, ck_p, s_pre, s_pre, causal=True, block_table=bt[:n_pre])
    cq_d = (cq[n_pre:] - cq[n_pre]).contiguous(); ck_d = (ck[n_pre:] - ck[n_pre]).contiguous()
    dec = lambda: fa.flash_attn_varlen_func(q[n_pre * s_pre:], kc, vc, cq_d, ck_d, 1, max(dec_ctx), causal=True, block_table=bt[n_pre:])
    r = {"tp": tp, "heads": [Hq, Hk]}
    for name, f in (("whole", whole), ("prefill_alone", pre), ("decode_alone", dec)):
        f(); r[name + "_launches"] = ext.last_call_info["launches"]
        r[name + "_ms"] = timeit(f, a.iters); r[name + "_graph_ms"] = timeit(f, a.iters, graph=True)
    flops = 2.0 * Hq * D * n_pre * s_pre * s_pre; nbytes = 2.0 * sum(dec_ctx) * Hk * D * 2
    r["ideal_ms_at_measured_peaks"] = flops / 1658.8e12 * 1e3 + nbytes / 6554.9e9 * 1e3
    print(json.dumps(r), flush=True)
