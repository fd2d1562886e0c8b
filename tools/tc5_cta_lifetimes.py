"""Per-CTA wall-clock lifetimes of the persistent tcgen05 prefill kernel on the prefill part of C4's TP shard
(4 x 2048-token prompts, 64/tp q heads, 8/tp kv heads, paged block 16) — how well the work list balances and what one
KV iteration costs when a CTA runs only two or three units.  Debug tool; needs the -DFA_TC5_TRACE build (see
tools/tc5_trace.py).
  python tools/tc5_cta_lifetimes.py [--tp 8] [--s 2048] [--n 4]"""
import argparse, os, struct, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TRACE = os.path.join(ROOT, "gpurun_out", "tc5_trace.bin")
os.environ["FA_B200_TRACE_FILE"] = TRACE
sys.path.insert(0, os.path.join(ROOT, "vllm-flash-attention_b200"))
ap = argparse.ArgumentParser()
ap.add_argument("--tp", type=int, default=8); ap.add_argument("--s", type=int, default=2048); ap.add_argument("--n", type=int, default=4)
ap.add_argument("--c2", action="store_true", help="BASELINE C2 instead: packed varlen, 32q/8kv, 16384 tokens")
ap.add_argument("--lib", default="libfa_b200_trace.so")
a = ap.parse_args()
os.environ["FA_B200_LIB"] = os.path.join(ROOT, "vllm-flash-attention_b200", "vllm_flash_attn", a.lib)
import torch  # noqa: E402
import vllm_flash_attn as fa  # noqa: E402
os.makedirs(os.path.dirname(TRACE), exist_ok=True)
dev, dt, D, page = "cuda", torch.bfloat16, 128, 16
Hq, Hk = 64 // a.tp, 8 // a.tp
cu = torch.arange(0, (a.n + 1) * a.s, a.s, dtype=torch.int32, device=dev)
mb = a.s // page
q = torch.randn(a.n * a.s, Hq, D, device=dev, dtype=dt)
kc = torch.randn(a.n * mb, page, Hk, D, device=dev, dtype=dt); vc = torch.randn_like(kc)
bt = torch.randperm(a.n * mb, device=dev).to(torch.int32).view(a.n, mb).contiguous()
f = lambda: fa.flash_attn_varlen_func(q, kc, vc, cu, cu, a.s, a.s, causal=True, block_table=bt)
if a.c2:
    lens = [8192, 4096, 2048, 1024, 512, 256, 128, 128]
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device=dev)
    q = torch.randn(sum(lens), 32, D, device=dev, dtype=dt); k = torch.randn(sum(lens), 8, D, device=dev, dtype=dt); v = torch.randn_like(k)
    f = lambda: fa.flash_attn_varlen_func(q, k, v, cu, cu, max(lens), max(lens), causal=True)
for _ in range(3): f()
torch.cuda.synchronize()
raw = open(TRACE, "rb").read()
vals = struct.unpack(f"{len(raw) // 8}q", raw)
base = 3 * 128 * 16
rec = [vals[base + 4 * c: base + 4 * c + 4] for c in range(148)]
t0 = min(r[0] for r in rec if r[0]); t1 = max(r[1] for r in rec)
print(f"kernel span (first issuer start -> last issuer end): {(t1 - t0) / 1e3:.1f} us; units total {sum(r[3] for r in rec)}, KV iterations total {sum(r[2] for r in rec)}")
life = sorted(((r[1] - r[0]) / 1e3, (r[0] - t0) / 1e3, r[2], r[3], c) for c, r in enumerate(rec))
print("CTA lifetimes us (min / median / max):", life[0][0], life[74][0], life[-1][0])
print("start skew us (max):", max(x[1] for x in life))
n = len(life); sx = sum(x[3] for x in life); sy = sum(x[2] for x in life); sl = sum(x[0] for x in life)
# least squares  lifetime = a * iterations + b * units  over the CTAs
import numpy as np
A = np.array([[x[2], x[3]] for x in life], dtype=float); y = np.array([x[0] for x in life])
coef, *_ = np.linalg.lstsq(A, y, rcond=None)
print(f"fit over CTAs: lifetime = {coef[0] * 1e3:.0f} ns x iterations + {coef[1] * 1e3:.0f} ns x units; mean lifetime {sl / n:.1f} us, iterations per CTA {sy / n:.1f}, units per CTA {sx / n:.1f}")
by_units = {}
for lt, st, its, un, c in life: by_units.setdefault((un, its), []).append(lt)
for k in sorted(by_units)[:12]: print(f"  units {k[0]} iterations {k[1]:3d}: {len(by_units[k]):3d} CTAs, lifetime {min(by_units[k]):.1f} .. {max(by_units[k]):.1f} us -> {1e3 * sum(by_units[k]) / len(by_units[k]) / max(k[1], 1):.0f} ns per iteration")
c0, n0, c1, n1 = (vals[(2 * 128 + 127) * 16 + i] for i in (12, 13, 14, 15))
if n1 > n0: print(f"CTA 0: {c1 - c0} clk in {(n1 - n0) / 1e3:.1f} us -> {(c1 - c0) / (n1 - n0) * 1e3:.0f} MHz")
ev = lambda role, it, e: vals[(role * 128 + it) * 16 + e]
its0 = rec[0][2]
print("CTA 0, WG0 per iteration: S_ready, P_hi (clk since first wait); unit-end records")
tb = ev(0, 0, 0)
for it in range(min(its0, 40)):
    x = [ev(0, it, i) for i in (6, 7, 8)]
    extra = f"   unit end: PV done @{x[0] - tb} epilogue +{x[1] - x[0]} next unit read +{x[2] - x[1]}" if x[0] else ""
    print(f"  it {it:2d}: wait@{ev(0, it, 0) - tb:7d} S_ready@{ev(0, it, 1) - tb:7d} P_hi@{ev(0, it, 5) - tb:7d} | WG1 S_ready@{ev(1, it, 1) - tb:7d} P_hi@{ev(1, it, 5) - tb:7d}{extra}")
print("epilogue stamps of CTA 0 (clk after 'PV done'): slab free / O cols 0-63 in registers / staged / fenced / TMA issued + slab read / cols 64-127 done")
for u in range(min(rec[0][3], 10)):
    for wg in (0, 1):
        e = [ev(wg, 110 + u, i) for i in range(7)]
        if e[0]: print(f"  unit {u} WG{wg}: " + " ".join(f"+{e[i] - e[0]}" for i in range(1, 7)))
