"""Same-box comparator (development / measurement tool, NOT part of the product): vLLM's CuTe-DSL FlashAttention-4
forward for sm_100 (vllm/vllm_flash_attn/cute/flash_fwd_sm100.py, library code in this image) next to ours on C2 and on
a few dense shapes.  VERDICT r1 item 6: "if a Blackwell-native kernel on the same box is faster ... that is the number
to chase".  Prints one JSON line per shape; {"unavailable": why} if the package cannot be imported or compiled here.
  python tools/fa4_bench.py [--iters 20]"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vllm-flash-attention_b200"))
import torch
import vllm_flash_attn as fa


def timeit(f, iters, warmup=3):
    for _ in range(warmup): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    ap = argparse.ArgumentParser(); ap.add_argument("--iters", type=int, default=20); a = ap.parse_args()
    try:
        from vllm.vllm_flash_attn.cute import interface as fa4
    except Exception as ex:  # noqa: BLE001
        print(json.dumps({"unavailable": "import: " + repr(ex)[:300]})); return
    dev, dt = "cuda", torch.bfloat16
    seqlens = [8192, 4096, 2048, 1024, 512, 256, 128, 128]
    tot = sum(seqlens)
    q = torch.randn(tot, 32, 128, device=dev, dtype=dt); k = torch.randn(tot, 8, 128, device=dev, dtype=dt); v = torch.randn_like(k)
    cu = torch.tensor([0] + list(torch.tensor(seqlens).cumsum(0)), dtype=torch.int32, device=dev)
    flops = 2.0 * 32 * 128 * sum(s * s for s in seqlens) / 1e12
    ours = lambda: fa.flash_attn_varlen_func(q, k, v, cu, cu, 8192, 8192, causal=True)
    theirs = lambda: fa4.flash_attn_varlen_func(q, k, v, cu_seqlens_q=cu, cu_seqlens_k=cu, max_seqlen_q=8192, max_seqlen_k=8192, causal=True)
    try:
        o4 = theirs()
        o4 = o4[0] if isinstance(o4, (tuple, list)) else o4
        diff = (o4.float() - ours().float()).abs().max().item()
        ms4, mso = timeit(theirs, a.iters), timeit(ours, a.iters)
        print(json.dumps({"config": "C2 varlen mixed 16k", "ours_tflops": flops / (mso * 1e-3), "fa4_cute_tflops": flops / (ms4 * 1e-3),
                          "ours_ms": mso, "fa4_ms": ms4, "max_abs_diff": diff}), flush=True)
    except Exception as ex:  # noqa: BLE001
        print(json.dumps({"config": "C2 varlen mixed 16k", "unavailable": repr(ex)[:400]}), flush=True)
    for causal in (False, True):
        for b, s in ((2, 8192), (8, 2048)):
            q = torch.randn(b, s, 16, 128, device=dev, dtype=dt); k = torch.randn_like(q); v = torch.randn_like(q)
            fl = 4.0 * b * s * s * 16 * 128 / (2 if causal else 1) / 1e12
            try:
                ms4 = timeit(lambda: fa4.flash_attn_func(q, k, v, causal=causal), a.iters)
                mso = timeit(lambda: fa.flash_attn_func(q, k, v, causal=causal), a.iters)
                print(json.dumps({"config": f"dense d128 h16 b{b} s{s} causal={causal}", "ours_tflops": fl / (mso * 1e-3),
                                  "fa4_cute_tflops": fl / (ms4 * 1e-3)}), flush=True)
            except Exception as ex:  # noqa: BLE001
                print(json.dumps({"config": f"dense b{b} s{s} causal={causal}", "unavailable": repr(ex)[:300]}), flush=True)


if __name__ == "__main__":
    main()
