"""Timeline of one CTA of the tcgen05 prefill kernel (debug tool; needs the -DFA_TC5_TRACE build:
  cd vllm-flash-attention_b200 && FA_BUILD_DIR=build_trace FA_OUT=vllm_flash_attn/libfa_b200_trace.so \
      FA_NVCC_EXTRA=-DFA_TC5_TRACE bash build.sh
  python tools/tc5_trace.py [--causal] [--s 8192]
Prints, per KV iteration, SM-clock deltas between the pipeline events of the two softmax
warpgroups and the MMA issuer."""
import argparse
import os
import struct
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ["FA_B200_LIB"] = os.path.join(ROOT, "vllm-flash-attention_b200", "vllm_flash_attn", "libfa_b200_trace.so")
TRACE = os.path.join(ROOT, "gpurun_out", "tc5_trace.bin")
os.environ["FA_B200_TRACE_FILE"] = TRACE
sys.path.insert(0, os.path.join(ROOT, "vllm-flash-attention_b200"))

import torch  # noqa: E402
import vllm_flash_attn as fa  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--causal", action="store_true")
    ap.add_argument("--s", type=int, default=8192)
    ap.add_argument("--h", type=int, default=16)
    ap.add_argument("--d", type=int, default=128)
    ap.add_argument("--first", type=int, default=8)
    ap.add_argument("--count", type=int, default=10)
    a = ap.parse_args()
    os.makedirs(os.path.dirname(TRACE), exist_ok=True)
    b = max(1, 16384 // a.s)
    q = torch.randn(b, a.s, a.h, a.d, device="cuda", dtype=torch.bfloat16); k = torch.randn_like(q); v = torch.randn_like(q)
    for _ in range(3):
        fa.flash_attn_func(q, k, v, causal=a.causal)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fa.flash_attn_func(q, k, v, causal=a.causal); e1.record(); torch.cuda.synchronize()
    print(f"one call (trace build, incl. trace memset + schedule kernel + sync): {e0.elapsed_time(e1) * 1e3:.1f} us")
    raw = open(TRACE, "rb").read()
    vals = struct.unpack(f"{len(raw) // 8}q", raw)
    ev = lambda role, it, e: vals[(role * 128 + it) * 16 + e]
    print("softmax WG events: 0 wait_S start, 1 S ready, 2 first-half loaded, 3 max done, 4 P_lo published, 5 P_hi published")
    print("MMA issuer events (per tile t): V_j,K_j+1 landed / P seen / PV(j)+QK(j+1) issued")
    t00 = ev(0, a.first, 0)
    for it in range(a.first, a.first + a.count):
        for wg in (0, 1):
            e = [ev(wg, it, i) - t00 for i in range(6)]
            x = [ev(wg, it, i) for i in (6, 7, 8)]
            if x[0]:
                print(f"        WG{wg} unit end after it {it}: PV(last) done @{x[0]-t00}, epilogue stores +{x[1]-x[0]}, next unit read +{x[2]-x[1]}")
            pw = [ev(wg, it, 9 + w) - t00 for w in range(4)]
            sw = [e[1]] + [ev(wg, it, 12 + w) - t00 for w in (1, 2, 3)]
            print(f"        WG{wg} per warp: S seen {sw}  P published {pw}  (softmax time {[pw[w] - sw[w] for w in range(4)]})")
            print(f"it {it:3d} WG{wg}: wait@{e[0]:7d} S_ready +{e[1]-e[0]:5d} ld1 +{e[2]-e[1]:4d} max +{e[3]-e[2]:4d} "
                  f"exp_half +{e[4]-e[3]:5d} P_pub +{e[5]-e[4]:5d}  | iter total {e[5]-e[0]:5d}  S_ready@{e[1]:7d} P_hi@{e[5]:7d}")
        m = [ev(2, it, i) - t00 for i in range(16)]
        print(f"        MMA : top@{m[0]:7d} KV +{m[1]-m[0]:4d} | t0 P@{m[2]:7d} PVissue +{m[3]-m[2]:4d} commit +{m[4]-m[3]:4d} QKissue +{m[5]-m[4]:4d} commit +{m[6]-m[5]:4d}"
              f" | t1 P@{m[10]:7d} (waited {m[10]-m[6]:5d}) PVissue +{m[11]-m[10]:4d} commit +{m[12]-m[11]:4d} QKissue +{m[13]-m[12]:4d} commit +{m[14]-m[13]:4d}")
    c0, n0, c1, n1, its = (ev(2, 127, i) for i in (12, 13, 14, 15, 11))
    if n1 > n0:
        print(f"CTA 0 lifetime: {c1 - c0} SM clk in {(n1 - n0) / 1e3:.1f} us -> {(c1 - c0) / (n1 - n0) * 1e3:.0f} MHz effective; "
              f"{its} KV iterations -> {(c1 - c0) / max(its, 1):.0f} clk per iteration overall")
    # average period
    per = (ev(0, a.first + a.count, 5) - ev(0, a.first, 5)) / a.count
    print(f"average period per KV tile (WG0 P_hi to P_hi): {per:.0f} clk; tensor work per period = 2048 clk -> {2048 / per:.1%} of peak")


if __name__ == "__main__":
    main()
