"""C3 (paged decode B=256, ctx 4096, block 16, 32q/8kv D=128) a few times — the command ncu wraps for the decode capture.
  python tools/c3_once.py [--iters 5]"""
import argparse, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vllm-flash-attention_b200"))
import torch
import vllm_flash_attn as fa
ap = argparse.ArgumentParser(); ap.add_argument("--iters", type=int, default=5); a = ap.parse_args()
dev, dt, b, ctx, page = "cuda", torch.bfloat16, 256, 4096, 16
nblk = b * (ctx // page) * 5 // 4
kc = torch.randn(nblk, page, 8, 128, device=dev, dtype=dt); vc = torch.randn_like(kc)
bt = torch.randperm(nblk, device=dev)[: b * (ctx // page)].to(torch.int32).view(b, -1).contiguous()
q = torch.randn(b, 1, 32, 128, device=dev, dtype=dt); lens = torch.full((b,), ctx, dtype=torch.int32, device=dev)
out = torch.empty_like(q)
for _ in range(a.iters):
    fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=lens, block_table=bt, causal=True, out=out)
torch.cuda.synchronize()
print("ok")
