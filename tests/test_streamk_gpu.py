"""GPU: the decode kernel's balanced-range form (fa_decode_streamk_kernel: a fixed grid of 2 CTAs per SM, every CTA an equal
share of ALL units' KV chunks, pieces of a unit merged by its last arriver).  It is opt-in (FA_B200_STREAMK=1; measured
wins and losses: profiles/r02_decode_streamk.txt), so the decode / kvcache / append / fuzz tests and a TP-shard-like
mixed batch are re-run in a subprocess with that setting: every mask, page size, head dim, split merge and the in-kernel
append are checked on it as well.  The mixed batch is also checked in-process on the default (classic) grid."""
import os
import subprocess
import sys

import pytest
import torch

import _util  # noqa: F401  (sys.path)
import attention_oracle as orc

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_mixed_batch_decode_half_on_balanced_ranges():
    """C4's TP=8 shard in small: 8 q heads / 1 kv head, 3 prompts + 190 single-token decodes with ragged contexts (one
    empty, one of a single key), paged cache with NaN in every unused row."""
    import vllm_flash_attn as fa
    from vllm_flash_attn import vllm_flash_attn_2_cuda as ext
    g = torch.Generator().manual_seed(3)
    dev, dt, h, hk, d, page = "cuda", torch.bfloat16, 8, 1, 128, 16
    lens_q = [400, 129, 300] + [1] * 190
    lens_k = [400, 129, 300] + torch.randint(1, 3000, (190,), generator=g).tolist()
    lens_k[10], lens_k[11] = 0, 1
    nb = len(lens_q); mb = (max(lens_k) + page - 1) // page
    cu_q = torch.tensor([0] + list(torch.tensor(lens_q).cumsum(0)), dtype=torch.int32)
    cu_k = torch.tensor([0] + list(torch.tensor(lens_k).cumsum(0)), dtype=torch.int32)
    q = torch.randn(sum(lens_q), h, d, generator=g).to(dt).to(dev)
    k = torch.randn(sum(lens_k), hk, d, generator=g).to(dt).to(dev); v = torch.randn(sum(lens_k), hk, d, generator=g).to(dt).to(dev)
    bt = torch.randperm(nb * mb, generator=g).to(torch.int32).view(nb, mb).contiguous().to(dev)
    kc = torch.full((nb * mb, page, hk, d), float("nan"), device=dev, dtype=dt); vc = torch.full_like(kc, float("nan"))
    from vllm_flash_attn.cache_ops import reshape_and_cache_flash
    slots = torch.cat([bt[i, torch.arange(n, device=dev) // page].long() * page + torch.arange(n, device=dev) % page
                       for i, n in enumerate(lens_k) if n > 0])
    reshape_and_cache_flash(k, v, kc, vc, slots)
    out = fa.flash_attn_varlen_func(q, kc, vc, cu_q.to(dev), cu_k.to(dev), max(lens_q), max(lens_k), causal=True, block_table=bt)
    info = dict(ext.last_call_info)
    assert info["kernel_used"] == 1 and info["num_splits"] > 1, info           # the decode half is split (classic grid unless forced)
    ref32, _ = orc.attention_ref_varlen(q, k, v, cu_q, cu_k, causal=True)
    pt, _ = orc.attention_ref_varlen(q, k, v, cu_q, cu_k, causal=True, upcast=False, reorder_ops=True)
    err, err_pt = (out.float() - ref32.float()).abs().max().item(), (pt.float() - ref32.float()).abs().max().item()
    assert not torch.isnan(out.float()).any()
    assert err <= 2 * err_pt + 1e-5, (err, err_pt)
    for _ in range(3):                                                          # arrival counters back at zero, pool reused: same bits
        assert torch.equal(fa.flash_attn_varlen_func(q, kc, vc, cu_q.to(dev), cu_k.to(dev), max(lens_q), max(lens_k), causal=True,
                                                     block_table=bt), out)


def test_decode_tests_again_with_balanced_ranges_forced_on():
    env = dict(os.environ, FA_B200_STREAMK="1")
    cmd = [sys.executable, "-m", "pytest", "-x", "-q", "-m", "gpu", "-p", "no:cacheprovider",
           os.path.join(ROOT, "tests", "test_parity_gpu.py"), os.path.join(ROOT, "tests", "test_fuzz_gpu.py"),
           os.path.join(ROOT, "tests", "test_api_gpu.py"), os.path.join(ROOT, "tests", "test_serving_loop_gpu.py"),
           os.path.join(ROOT, "tests", "test_zz_no_alloc.py"), os.path.join(ROOT, "tests", "test_torch_ext.py"),
           os.path.join(ROOT, "tests", "test_cabi.py"), os.path.join(ROOT, "tests", "test_streamk_gpu.py"),
           "-k", "not four_threads and not forced_on"]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=1200, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
