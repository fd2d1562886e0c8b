import sys, time, torch
sys.path.insert(0, "vllm-flash-attention_b200")
import vllm_flash_attn as fa
q = torch.randn(2, 512, 8, 64, device="cuda", dtype=torch.bfloat16); k = torch.randn_like(q); v = torch.randn_like(q)
for _ in range(20): fa.flash_attn_func(q, k, v)
torch.cuda.synchronize()
n = 2000
t0 = time.perf_counter()
for _ in range(n): fa.flash_attn_func(q, k, v)
t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
print(f"C1 host enqueue {1e6*(t1-t0)/n:.1f} us/call, wall incl. drain {1e6*(t2-t0)/n:.1f} us/call")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
g = torch.cuda.CUDAGraph()
s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(3): fa.flash_attn_func(q, k, v)
torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
with torch.cuda.graph(g):
    for _ in range(20): fa.flash_attn_func(q, k, v)
g.replay(); torch.cuda.synchronize()
e0.record(); 
for _ in range(20): g.replay()
e1.record(); torch.cuda.synchronize()
print(f"C1 device time per call (CUDA graph of 20 calls): {e0.elapsed_time(e1)*1e3/400:.2f} us")
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
for _ in range(500): fa.flash_attn_func(q, k, v)
pr.disable(); torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("tottime").print_stats(12)
