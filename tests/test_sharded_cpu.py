"""CPU (gloo, world_size 2) tests of the host-side multi-GPU plumbing in vllm_flash_attn.sharded:
range arithmetic, packed (out|lse) record layout and the ONE all-gather of the sequence-sharded
decode path.  The CUDA compute steps are replaced through the two hooks by the CPU oracle
(test infrastructure) so the collective + merge algebra can be checked against attention_ref."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import attention_oracle as orc


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def test_head_shard_partitions_heads():
    from vllm_flash_attn import sharded
    seen_q, seen_k = [], []
    for r in range(4):
        qs, ks = sharded.head_shard(r, 4, 64, 8)
        seen_q += list(range(64))[qs]; seen_k += list(range(8))[ks]
        assert (qs.stop - qs.start) == 16 and (ks.stop - ks.start) == 2
        assert qs.start // 8 == ks.start          # query head h uses kv head h // g
    assert seen_q == list(range(64)) and seen_k == list(range(8))
    with pytest.raises(ValueError):
        sharded.head_shard(0, 3, 64, 8)


def test_seq_shard_ranges_cover_and_align():
    from vllm_flash_attn import sharded
    for ctx, world, page in ((131072, 8, 16), (1000, 4, 16), (70, 2, 256), (4096, 1, 16)):
        prev = 0
        for r in range(world):
            s, e = sharded.seq_shard_range(r, world, ctx, page)
            assert s == prev and s % page == 0 and e > s
            prev = e
        assert prev >= ctx
    lens = torch.tensor([0, 5, 100, 517, 1000], dtype=torch.int32)
    tot = torch.zeros_like(lens)
    for r in range(4):
        s, e = sharded.seq_shard_range(r, 4, 1000, 16)
        tot += sharded.seq_shard_lengths(lens, s, e)
    assert torch.equal(tot, lens)


def _worker(rank, world, port, q, k, v, lens, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from vllm_flash_attn import sharded
    b, sq, h, d = q.shape
    s, e = sharded.seq_shard_range(rank, world, k.shape[1], 16)
    kl, vl = k[:, s:e], v[:, s:e]
    local_lens = sharded.seq_shard_lengths(lens, s, e)

    def local_fn(buf):        # stands in for the CUDA decode kernel in partial_out mode
        rows = b * h * sq
        o32 = buf[: rows * d].view(b, h, sq, d); lse = buf[rows * d: rows * (d + 1)].view(b, h, sq)
        for i in range(b):
            n = int(local_lens[i])
            if n == 0:
                o32[i] = 0; lse[i] = float("-inf"); continue
            o, _ = orc.attention_ref(q[i:i + 1], kl[i:i + 1, :n], vl[i:i + 1, :n])
            o32[i] = o[0].float().transpose(0, 1)
            lse[i] = orc.lse_ref(q[i:i + 1], kl[i:i + 1, :n])[0]

    def merge_fn(records, b_, h_, sq_, d_, dtype):
        rows = b_ * h_ * sq_
        o, l = orc.combine_ref(records[:, : rows * d_].view(-1, rows, d_), records[:, rows * d_: rows * (d_ + 1)])
        return o.view(b_, h_, sq_, d_).transpose(1, 2).to(dtype), l.view(b_, h_, sq_)

    out, lse = sharded.flash_attn_with_kvcache_seq_sharded(
        q, kl, vl, lens, None, shard_start=s, shard_end=e, return_softmax_lse=True,
        _local_fn=local_fn, _merge_fn=merge_fn)
    ret[rank] = (out.clone(), lse.clone())
    dist.barrier()
    dist.destroy_process_group()


def test_seq_sharded_decode_gloo_world2():
    torch.manual_seed(0)
    b, h, hk, d, ctx = 3, 4, 2, 32, 160
    q = torch.randn(b, 1, h, d)
    k = torch.randn(b, ctx, hk, d); v = torch.randn(b, ctx, hk, d)
    lens = torch.tensor([160, 70, 9], dtype=torch.int32)      # third sequence lives entirely on rank 0
    mgr = mp.Manager(); ret = mgr.dict()
    port = _free_port()
    mp.spawn(_worker, args=(2, port, q, k, v, lens, ret), nprocs=2, join=True)
    kpm = torch.arange(ctx).expand(b, ctx) < lens.view(-1, 1)
    ref, _ = orc.attention_ref(q, k, v, None, kpm)
    lref = orc.lse_ref(q, k, kpm)
    for r in range(2):
        out, lse = ret[r]
        assert out.shape == (b, 1, h, d)
        assert (out - ref).abs().max() < 1e-5
        assert (lse - lref).abs().max() < 1e-5
    assert torch.equal(ret[0][0], ret[1][0])       # every rank ends with the same result
