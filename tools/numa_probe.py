"""Host <-> device copy bandwidth of pinned memory on this box: each direction alone, under the NVML-recommended CPU
affinity, and both directions at once (what bounds bench.py's e2e leg).  Development tool.   python tools/numa_probe.py"""
import os, time, torch
print("affinity", sorted(os.sched_getaffinity(0)))
os.system("lscpu | grep -i -E 'numa|socket|model name|^CPU\\(s\\)'; nvidia-smi topo -m 2>/dev/null | head -6; cat /sys/bus/pci/devices/*/local_cpulist 2>/dev/null | sort | uniq -c | head")
import pynvml; pynvml.nvmlInit(); h = pynvml.nvmlDeviceGetHandleByIndex(0)
def bw(tag):
    x = torch.empty(256 << 20, dtype=torch.uint8, pin_memory=True); x.fill_(1)
    d = torch.empty_like(x, device="cuda")
    for _ in range(3): d.copy_(x, non_blocking=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): d.copy_(x, non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    h2d = 10 * x.numel() / e0.elapsed_time(e1) / 1e6
    e0.record()
    for _ in range(10): x.copy_(d, non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    print(tag, "H2D GB/s", round(h2d, 1), "D2H GB/s", round(10 * x.numel() / e0.elapsed_time(e1) / 1e6, 1))
bw("default")
try:
    pynvml.nvmlDeviceSetCpuAffinity(h); print("nvml affinity ->", sorted(os.sched_getaffinity(0))[:8], "...", len(os.sched_getaffinity(0)))
    bw("nvml-affinity")
except Exception as e: print("affinity failed", e)
# both directions at once (what the e2e leg of bench.py does): is PCIe full duplex at full rate here?
x = torch.empty(201 << 20, dtype=torch.uint8, pin_memory=True); d = torch.empty_like(x, device="cuda")
y = torch.empty(134 << 20, dtype=torch.uint8, pin_memory=True); g = torch.empty_like(y, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
torch.cuda.synchronize()
for both in (False, True):
    a0, a1, b0, b1 = (torch.cuda.Event(enable_timing=True) for _ in range(4))
    with torch.cuda.stream(s1):
        a0.record(s1)
        for _ in range(10): d.copy_(x, non_blocking=True)
        a1.record(s1)
    if both:
        with torch.cuda.stream(s2):
            b0.record(s2)
            for _ in range(15): y.copy_(g, non_blocking=True)
            b1.record(s2)
    torch.cuda.synchronize()
    print("concurrent D2H" if both else "H2D alone", "H2D GB/s", round(10 * x.numel() / a0.elapsed_time(a1) / 1e6, 1),
          ("D2H GB/s %.1f" % (15 * y.numel() / b0.elapsed_time(b1) / 1e6)) if both else "")
