"""GPU bring-up probe (development tool; the case matrix and the runner live in tests/parity_cases.py).

Runs a matrix of small cases through the public API with a pinned kernel family, each in its own
subprocess with a timeout (a hung tcgen05 pipeline must not take the whole gpurun call down), and
writes one JSON line per case to gpurun_out/probe.jsonl: max-abs error of our kernel and of the
reference FA2 kernel (oracle/_ref) against the fp32 attention_ref oracle.

  python tools/gpu_probe.py                # all cases
  python tools/gpu_probe.py --case NAME    # one case, in-process
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))

from parity_cases import CASES, C, K_DEC, K_TC5, K_TILE, run_case  # noqa: E402,F401


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--case")
    ap.add_argument("--filter", default="")
    ap.add_argument("--timeout", type=int, default=150)
    a = ap.parse_args()
    if a.case:
        spec = next(c for c in CASES if c["name"] == a.case)
        print("RESULT " + json.dumps(run_case(spec)))
        return
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    outp = open(os.path.join(ROOT, "gpurun_out", "probe.jsonl"), "a")
    nfail = 0
    for spec in CASES:
        if a.filter and a.filter not in spec["name"]:
            continue
        t0 = time.time()
        try:
            pr = subprocess.run([sys.executable, __file__, "--case", spec["name"]], capture_output=True, text=True,
                                timeout=a.timeout)
            line = next((l for l in pr.stdout.splitlines() if l.startswith("RESULT ")), None)
            res = json.loads(line[7:]) if line else dict(name=spec["name"], ok=False, rc=pr.returncode,
                                                         stderr=pr.stderr[-1500:])
        except subprocess.TimeoutExpired:
            res = dict(name=spec["name"], ok=False, timeout=True)
        res["secs"] = round(time.time() - t0, 1)
        nfail += 0 if res.get("ok") else 1
        outp.write(json.dumps(res) + "\n"); outp.flush()
        print(json.dumps(res), flush=True)
    print(f"probe done: {nfail} failing")


if __name__ == "__main__":
    main()
