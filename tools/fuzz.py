"""Randomised parity sweep through the automatic kernel choice (development tool): random head dims, lengths, masks,
page sizes and batch mixes, each checked against the fp32 oracle with the reference's criterion.
  python tools/fuzz.py [--n 300] [--seed 0]       # prints failing specs, then a summary line"""
import argparse
import json
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "tools"))
from parity_cases import random_spec, run_case  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=300)
    ap.add_argument("--seed", type=int, default=0)
    a = ap.parse_args()
    rng = random.Random(a.seed)
    bad, used = 0, {}
    for i in range(a.n):
        spec = random_spec(rng, a.seed * 100000 + i)
        try:
            r = run_case(spec)
            mult = 3 if spec["kind"] == "kvcache" else 2
            ok = (not r["nan"]) and r["err"] <= mult * r["err_pt"] + 1e-5 and r["lse_err"] < 2e-3
            used[r["kernel_used"]] = used.get(r["kernel_used"], 0) + 1
            if not ok:
                bad += 1
                print("FAIL", json.dumps({k: spec[k] for k in ("kind", "d", "h", "hk", "b", "sq", "sk", "causal", "window", "dtype", "alibi", "page", "mixed", "splits", "seed")}),
                      json.dumps({k: r[k] for k in ("err", "err_pt", "lse_err", "nan", "kernel_used")}), flush=True)
        except Exception as e:  # noqa: BLE001
            bad += 1
            print("EXC", json.dumps({k: spec[k] for k in ("kind", "d", "h", "hk", "b", "sq", "sk", "causal", "window", "dtype", "alibi", "page", "mixed", "splits", "seed")}), repr(e)[:300], flush=True)
    print(json.dumps({"cases": a.n, "failing": bad, "kernels_used": used}))


if __name__ == "__main__":
    main()
