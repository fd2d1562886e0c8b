"""Multi-GPU check + timing of the two shardings (run under torchrun, one rank per GPU):

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
      --master-port 29511 tools/multi_gpu_check.py [--ctx 131072] [--batch 1]

* C5: long-context decode, KV sequence sharded over the ranks, NCCL all-gather of (out,lse) records +
  LSE combine; checked against the single-GPU kernel run on rank 0 over the whole sequence.
* C4: Llama-3-70B attention (64q/8kv heads, D=128) head-sharded, mixed prefill+decode varlen batch
  through flash_attn_varlen_func(block_table=...); no collective; checked against rank 0 running
  all heads.
Timing: CUDA events, barrier + synchronize on both sides, max over ranks.  One JSON line per case
from rank 0.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "vllm-flash-attention_b200"))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import vllm_flash_attn as fa  # noqa: E402
from vllm_flash_attn import sharded  # noqa: E402


def timed(f, iters, world, dev):
    for _ in range(5):
        f()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        f()
    e1.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def timed_graph(f, iters, world, dev):
    """Same measurement with the step captured in a CUDA graph (how a serving loop replays decode
    steps): removes the Python/ctypes launch overhead that dominates at tens of microseconds."""
    try:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                f()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            f()
        return timed(g.replay, iters, world, dev)
    except Exception as ex:  # noqa: BLE001
        print("graph capture failed:", repr(ex)[:300], flush=True)
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ctx", type=int, default=131072)
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--iters", type=int, default=50)
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    dt = torch.bfloat16
    H, HK, D, page = 32, 8, 128, 16

    # ------------------------------------------------------------------ C5: sequence-sharded decode
    b, ctx = a.batch, a.ctx
    g = torch.Generator(device="cpu").manual_seed(1234)           # same data on every rank
    q = torch.randn(b, 1, H, D, generator=g).to(dt).to(dev)
    lens = torch.full((b,), ctx, dtype=torch.int32)
    if b > 1:
        lens[1] = ctx // 3 + 5                                      # a ragged sequence; some ranks see none of it
    lens = lens.to(dev)
    s, e = sharded.seq_shard_range(rank, world, ctx, page)
    e = min(e, ctx)
    L = e - s
    # this rank's shard of the (logically contiguous) K/V, generated chunk-wise with a per-chunk seed
    def chunk(seq, r, which):
        gg = torch.Generator(device="cpu").manual_seed(100000 * which + 1000 * seq + r)
        cs, ce = sharded.seq_shard_range(r, world, ctx, page); ce = min(ce, ctx)
        return torch.randn(ce - cs, HK, D, generator=gg).to(dt)
    nblk = b * (L // page)
    perm = torch.randperm(nblk, generator=g)
    kc = torch.empty(nblk, page, HK, D, dtype=dt, device=dev); vc = torch.empty_like(kc)
    bt = perm.view(b, L // page).to(torch.int32).to(dev)
    for i in range(b):
        kc[bt[i].long()] = chunk(i, rank, 0).view(L // page, page, HK, D).to(dev)
        vc[bt[i].long()] = chunk(i, rank, 1).view(L // page, page, HK, D).to(dev)

    f = lambda: sharded.flash_attn_with_kvcache_seq_sharded(q, kc, vc, lens, bt, shard_start=s, shard_end=e,
                                                             causal=True, return_softmax_lse=True)
    out, lse = f()
    # reference: rank 0 assembles the whole sequence and runs the single-GPU kernel
    err = lerr = None
    if rank == 0:
        kfull = torch.stack([torch.cat([chunk(i, r, 0) for r in range(world)]) for i in range(b)]).to(dev)
        vfull = torch.stack([torch.cat([chunk(i, r, 1) for r in range(world)]) for i in range(b)]).to(dev)
        o1, l1 = fa.flash_attn_with_kvcache(q, kfull, vfull, cache_seqlens=lens, causal=True, return_softmax_lse=True)
        err = (out.float() - o1.float()).abs().max().item(); lerr = (lse - l1).abs().max().item()
        del kfull, vfull
    ms = timed(f, a.iters, world, dev)
    ms_local = timed(lambda: sharded.kvcache_partial(q, kc, vc, sharded.seq_shard_lengths(lens, s, e), bt), a.iters, world, dev)
    # fused exchange: peer-memory push + wait + combine in one kernel instead of NCCL all-gather + combine
    xch = sharded.PeerExchange(sharded.partial_record_numel(b, 1, H, D), dev)
    ff = lambda: sharded.flash_attn_with_kvcache_seq_sharded(q, kc, vc, lens, bt, shard_start=s, shard_end=e,
                                                              causal=True, return_softmax_lse=True, exchange=xch)
    out_f, lse_f = ff()
    torch.cuda.synchronize()
    err_f = (out_f.float() - out.float()).abs().max().item(); lerr_f = (lse_f - lse).abs().max().item()
    ms_fused = timed(ff, a.iters, world, dev)
    ms_fused_graph = timed_graph(ff, a.iters, world, dev)
    lens_shard = sharded.seq_shard_lengths(lens, s, e)
    ms_graph = timed_graph(f, a.iters, world, dev)
    ms_local_graph = timed_graph(lambda: sharded.kvcache_partial(q, kc, vc, lens_shard, bt), a.iters, world, dev)
    if rank == 0:
        nbytes = 2.0 * float(lens.sum().item()) * HK * D * 2
        print(json.dumps({"case": "C5_seq_sharded_decode", "n_gpus": world, "batch": b, "ctx": ctx, "ms": ms,
                          "ms_fused_exchange": ms_fused, "ms_fused_exchange_cuda_graph": ms_fused_graph,
                          "fused_vs_nccl_max_abs_diff": err_f, "fused_vs_nccl_lse_diff": lerr_f,
                          "ms_cuda_graph": ms_graph, "ms_local_kernel_only_cuda_graph": ms_local_graph,
                          "agg_GBps_cuda_graph": (nbytes / (ms_graph * 1e-3) / 1e9) if ms_graph else None,
                          "ms_local_kernel_only": ms_local, "agg_GBps": nbytes / (ms * 1e-3) / 1e9,
                          "max_abs_err_vs_1gpu": err, "lse_err_vs_1gpu": lerr}), flush=True)
    del kc, vc

    # ------------------------------------------------------------------ C4: head-sharded mixed batch
    Hq, Hk = 64, 8
    qs, ks = sharded.head_shard(rank, world, Hq, Hk)
    g = torch.Generator(device="cpu").manual_seed(4321)
    n_pre, s_pre, n_dec = 4, 2048, 192
    dec_ctx = torch.randint(1024, 4097, (n_dec,), generator=g)
    q_lens = [s_pre] * n_pre + [1] * n_dec
    k_lens = [s_pre] * n_pre + dec_ctx.tolist()
    cu_q = torch.tensor([0] + list(torch.tensor(q_lens).cumsum(0)), dtype=torch.int32)
    cu_k = torch.tensor([0] + list(torch.tensor(k_lens).cumsum(0)), dtype=torch.int32)
    tq = int(cu_q[-1]); nb = len(q_lens)
    max_blocks = (max(k_lens) + page - 1) // page
    nblk = nb * max_blocks
    q_all = torch.randn(tq, Hq, D, generator=g).to(dt)
    kc_all = torch.randn(nblk, page, Hk, D, generator=g).to(dt); vc_all = torch.randn(nblk, page, Hk, D, generator=g).to(dt)
    bt = torch.randperm(nblk, generator=g).view(nb, max_blocks).to(torch.int32).to(dev)
    cq, ck = cu_q.to(dev), cu_k.to(dev)
    q_r = q_all[:, qs].contiguous().to(dev); kc_r = kc_all[:, :, ks].contiguous().to(dev); vc_r = vc_all[:, :, ks].contiguous().to(dev)
    f4 = lambda: fa.flash_attn_varlen_func(q_r, kc_r, vc_r, cq, ck, max(q_lens), max(k_lens), causal=True, block_table=bt)
    o_r = f4()
    err4 = None
    if world > 1:
        gathered = [torch.empty_like(o_r) for _ in range(world)]
        dist.all_gather(gathered, o_r)             # check only: the product path has no collective
    else:
        gathered = [o_r]
    if rank == 0:
        o_full = fa.flash_attn_varlen_func(q_all.to(dev), kc_all.to(dev), vc_all.to(dev), cq, ck, max(q_lens), max(k_lens),
                                           causal=True, block_table=bt)
        err4 = (torch.cat(gathered, dim=1).float() - o_full.float()).abs().max().item()
    ms4 = timed(f4, 20, world, dev)
    if rank == 0:
        flops = 2.0 * Hq * D * n_pre * s_pre * s_pre
        nbytes = 2.0 * float(sum(k_lens[n_pre:])) * Hk * D * 2
        print(json.dumps({"case": "C4_head_sharded_mixed", "n_gpus": world, "ms": ms4,
                          "prefill_TFLOPs_if_alone": flops / (ms4 * 1e-3) / 1e12,
                          "decode_GBps_if_alone": nbytes / (ms4 * 1e-3) / 1e9, "max_abs_err_vs_all_heads": err4}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
