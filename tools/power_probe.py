import sys, time, torch
sys.path.insert(0, "vllm-flash-attention_b200")
import vllm_flash_attn as fa
lens = [8192, 4096, 2048, 1024, 512, 256, 128, 128]
dev, dt = "cuda", torch.bfloat16
cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device=dev)
q = torch.randn(sum(lens), 32, 128, device=dev, dtype=dt); k = torch.randn(sum(lens), 8, 128, device=dev, dtype=dt); v = torch.randn_like(k)
f = lambda: fa.flash_attn_varlen_func(q, k, v, cu, cu, 8192, 8192, causal=True)
for _ in range(5): f()
torch.cuda.synchronize()
def timed(n, gap):
    ts = []
    for _ in range(n):
        if gap: time.sleep(gap)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); f(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort(); return ts[len(ts) // 2], ts[0]
print("isolated launches (100 ms idle before each): median %.1f us, best %.1f us" % timed(15, 0.1))
print("isolated launches (5 ms idle): median %.1f us, best %.1f us" % timed(30, 0.005))
print("single launches back to back (sync between): median %.1f us, best %.1f us" % timed(50, 0))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for n in (20, 200, 2000):
    e0.record()
    for _ in range(n): f()
    e1.record(); torch.cuda.synchronize()
    print(f"{n} launches back to back: {e0.elapsed_time(e1) * 1e3 / n:.1f} us per launch")
