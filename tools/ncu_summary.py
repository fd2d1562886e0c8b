"""Summarise an .ncu-rep (raw page) into the handful of metrics the roofline discussion needs.
  python tools/ncu_summary.py gpurun_out/prof.ncu-rep [extra_regex] [--traffic KEY]
--traffic KEY records dram read+write bytes of the longest captured launch in profiles/traffic.json under KEY
(bench.py reports it as roofline.traffic)."""
import csv, io, json, os, re, subprocess, sys
argv = sys.argv[1:]
traffic_key = None
if "--traffic" in argv:
    i = argv.index("--traffic"); traffic_key = argv[i + 1]; del argv[i:i + 2]
rep = argv[0]
extra = argv[1] if len(argv) > 1 else None
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
pat = re.compile(r"(gpu__time_duration\.sum|dram__bytes_(read|write)\.sum$|dram__throughput\.avg\.pct|"
                 r"sm__pipe_tensor.*cycles_active.*pct|sm__inst_executed_pipe_(xu|tensor|fma|alu|lsu|uniform|tmem).*pct|sm__warps_active\.avg\.pct|"
                 r"launch__registers_per_thread|launch__grid_size|launch__block_size|launch__occupancy_limit|"
                 r"sm__throughput\.avg\.pct|smsp__issue_active\.avg\.pct|l1tex__throughput\.avg\.pct|lts__throughput\.avg\.pct|"
                 r"smsp__average_warps?_issue_stalled.*_per_issue_active|smsp__cycles_active\.avg|sm__cycles_elapsed\.max|"
                 r"smsp__inst_executed\.sum$|local_(load|store))" + (("|" + extra) if extra else ""))
for r in rows[2:]:
    print("== kernel:", r[hdr.index("Kernel Name")][:90], "grid", r[hdr.index("Grid Size")], "block", r[hdr.index("Block Size")])
    for h, u, v in zip(hdr, units, r):
        if pat.search(h):
            print(f"  {h} [{u}] = {v}")

if traffic_key and len(rows) > 2:
    # the longest captured launch is the benchmark's (warm-up / self-check launches of the same kernel are tiny)
    ti = hdr.index("gpu__time_duration.sum")
    r = max(rows[2:], key=lambda x: float(x[ti].replace(",", "")) * {"us": 1, "ms": 1e3, "ns": 1e-3, "s": 1e6}.get(units[ti].lower(), 1))
    def val(name):
        i = hdr.index(name)
        v, u = float(r[i].replace(",", "")), units[i].lower()
        return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
    tot = val("dram__bytes_read.sum") + val("dram__bytes_write.sum")
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "traffic.json")
    d = json.load(open(path)) if os.path.exists(path) else {}
    d[traffic_key] = tot
    json.dump(d, open(path, "w"), indent=1)
    print(f"traffic[{traffic_key}] = {tot:.4g} bytes per launch -> {path}")
