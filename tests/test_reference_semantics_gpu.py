"""GPU: the BASELINE.json configurations at FULL size and the corners of the reference's host glue, each held to
north_star's criterion — max|out - attention_ref_fp32| <= 2 x the reference kernel's own error on the same inputs
(3 x for kvcache, tests/test_flash_attn.py:2075-2076) — with the reference's FA2 kernel (oracle/_ref, built for sm_100
from /root/reference by oracle/Makefile) as the second checker.  attention_ref materialises (B,H,Sq,Sk) scores, so at
these sizes it runs per sequence / per KV-head group (SURVEY 8d's chunking)."""
import pytest
import torch

import _util
import attention_oracle as orc

pytestmark = pytest.mark.gpu
DEV, BF16 = "cuda", torch.bfloat16


def _need_refk():
    if _util.ref_lib() is None:
        pytest.skip("oracle/_ref/libfa2_ref_sm100.so not built")


def _errs(out, ref32, pt, refk=None):
    e = (out.float() - ref32.float()).abs().max().item()
    e_pt = (pt.float() - ref32.float()).abs().max().item()
    e_rk = (refk.float() - ref32.float()).abs().max().item() if refk is not None else None
    return e, e_pt, e_rk


# ------------------------------------------------------------------------------------------------ C5, one GPU
@pytest.mark.parametrize("batch", [1, 8])
def test_c5_long_context_decode_131072_keys(batch):
    """BASELINE.json configs[4] on one GPU: flash_attn_with_kvcache, ctx 131072, paged block 16, 32q/8kv D=128, automatic
    split count (the reference's suite runs kvcache at seqlen_k = 128*1024, tests/test_flash_attn.py:1855-1856, and
    split-KV at 100 000, :1714)."""
    _need_refk()
    import vllm_flash_attn as fa
    from vllm_flash_attn import vllm_flash_attn_2_cuda as ext
    torch.manual_seed(5)
    ctx, page, h, hk, d = 131072, 16, 32, 8, 128
    nblk = batch * ctx // page
    kc = torch.randn(nblk, page, hk, d, device=DEV, dtype=BF16); vc = torch.randn_like(kc)
    bt = torch.randperm(nblk, device=DEV).to(torch.int32).view(batch, -1).contiguous()
    q = torch.randn(batch, 1, h, d, device=DEV, dtype=BF16)
    lens = torch.full((batch,), ctx, dtype=torch.int32, device=DEV)
    if batch > 1:
        lens[1] = ctx // 3 + 5
        lens[2] = 17
    out, lse = fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=lens, block_table=bt, causal=True, return_softmax_lse=True)
    info = dict(ext.last_call_info)
    assert info["kernel_used"] == 3 and info["launches"] == 1
    assert info["num_splits"] > 1 or batch > 1
    rk, lrk = _util.ref_kernel_kvcache(q, kc, vc, lens, bt, causal=True)
    # fp32 oracle per (sequence, kv head): K/V of one kv head, its 4 query heads
    g = h // hk
    ref32 = torch.empty(batch, 1, h, d, device=DEV, dtype=torch.float32); pt = torch.empty_like(ref32)
    lref = torch.empty(batch, h, 1, device=DEV, dtype=torch.float32)
    for b in range(batch):
        n = int(lens[b])
        blocks = bt[b, : (n + page - 1) // page].long()
        for kh in range(hk):
            ks = kc[blocks, :, kh].reshape(1, -1, 1, d)[:, :n]; vs = vc[blocks, :, kh].reshape(1, -1, 1, d)[:, :n]
            qs = q[b:b + 1, :, kh * g:(kh + 1) * g]
            ref32[b, :, kh * g:(kh + 1) * g] = orc.attention_ref(qs, ks, vs)[0][0]
            pt[b, :, kh * g:(kh + 1) * g] = orc.attention_ref(qs, ks, vs, upcast=False, reorder_ops=True)[0][0]
            lref[b, kh * g:(kh + 1) * g] = orc.lse_ref(qs, ks)[0]
    e, e_pt, e_rk = _errs(out, ref32, pt, rk)
    assert not torch.isnan(out.float()).any()
    assert e <= 3 * e_rk + 1e-5, (e, e_rk, e_pt, info)
    assert e <= 3 * e_pt + 1e-5, (e, e_pt)
    assert (lse - lref).abs().max().item() < 2e-3 and (lse - lrk).abs().max().item() < 2e-3


# ------------------------------------------------------------------------------------------------ C4, full size
def test_c4_mixed_batch_full_size():
    """BASELINE.json configs[3] on one GPU (all 64q/8kv heads): 4 x 2048-token prompts + 192 single-token decodes with
    contexts randint(1024, 4097), paged block 16, ONE flash_attn_varlen_func(block_table=...) call — the call in which
    the tcgen05 prefill kernel and the decode kernel share a batch (fa_api.cu, mixed-batch routing)."""
    _need_refk()
    import vllm_flash_attn as fa
    from vllm_flash_attn import vllm_flash_attn_2_cuda as ext
    Hq, Hk, d, page = 64, 8, 128, 16
    n_pre, s_pre, n_dec = 4, 2048, 192
    gc = torch.Generator(device="cpu").manual_seed(0)
    dec_ctx = torch.randint(1024, 4097, (n_dec,), generator=gc).tolist()
    q_lens, k_lens = [s_pre] * n_pre + [1] * n_dec, [s_pre] * n_pre + dec_ctx
    cu_q = torch.tensor([0] + q_lens).cumsum(0).to(torch.int32); cu_k = torch.tensor([0] + k_lens).cumsum(0).to(torch.int32)
    nb = len(q_lens); mb = (max(k_lens) + page - 1) // page; nblk = nb * mb
    g = torch.Generator(device=DEV).manual_seed(0)
    q = torch.randn(sum(q_lens), Hq, d, device=DEV, dtype=BF16, generator=g)
    kc = torch.randn(nblk, page, Hk, d, device=DEV, dtype=BF16, generator=g)
    vc = torch.randn(nblk, page, Hk, d, device=DEV, dtype=BF16, generator=g)
    bt = torch.randperm(nblk, device=DEV, generator=g).to(torch.int32).view(nb, mb).contiguous()
    # pool rows no sequence owns hold NaN: they must never reach the output
    own = torch.zeros(nblk, page, dtype=torch.bool, device=DEV)
    for i in range(nb):
        nfull, rem = divmod(k_lens[i], page)
        own[bt[i, :nfull].long()] = True
        if rem:
            own[bt[i, nfull].long(), :rem] = True
    kp = torch.where(own.view(nblk, page, 1, 1), kc, torch.full_like(kc, float("nan")))
    vp = torch.where(own.view(nblk, page, 1, 1), vc, torch.full_like(vc, float("nan")))
    cq, ck = cu_q.to(DEV), cu_k.to(DEV)
    out, lse, _ = fa.flash_attn_varlen_func(q, kp, vp, cq, ck, max(q_lens), max(k_lens), causal=True, block_table=bt,
                                             return_attn_probs=True)
    info = dict(ext.last_call_info)
    assert info["kernel_used"] == 1 and not torch.isnan(out.float()).any()
    rk, _ = _util.ref_kernel_varlen(q, kc, vc, cq, ck, max(q_lens), max(k_lens), causal=True, block_table=bt)
    kg, vg = orc.gather_paged(kc, bt), orc.gather_paged(vc, bt)
    kd = torch.cat([kg[i, :k_lens[i]] for i in range(nb)]); vd = torch.cat([vg[i, :k_lens[i]] for i in range(nb)])
    del kg, vg
    ref32, lref = orc.attention_ref_varlen(q, kd, vd, cu_q, cu_k, causal=True)
    pt, _ = orc.attention_ref_varlen(q, kd, vd, cu_q, cu_k, causal=True, upcast=False, reorder_ops=True)
    for name, sl in (("prefill", slice(0, n_pre * s_pre)), ("decode", slice(n_pre * s_pre, None))):
        e, e_pt, e_rk = _errs(out[sl], ref32[sl], pt[sl], rk[sl])
        assert e <= 2 * e_rk + 1e-5, (name, e, e_rk, e_pt)
        assert e <= 2 * e_pt + 1e-5, (name, e, e_pt)
    for i in (0, n_pre - 1, n_pre, nb - 1):
        n = q_lens[i]
        assert (lse[i, :, :n].cpu() - lref[i, :, :n]).abs().max().item() < 2e-3


# ------------------------------------------------------------------------------------------------ C2, the long sequence
def test_c2_8192_token_sequence_2x_rule():
    """BASELINE.json configs[1] at full size: the whole varlen batch in one call, and its 8192-token sequence — where the
    lazy O rescale and the max-free softmax path matter most — held to the 2x rule against the fp32 oracle, per KV-head
    group (4 query heads: 1 GiB of fp32 scores)."""
    _need_refk()
    import vllm_flash_attn as fa
    seqlens, h, hk, d = [8192, 4096, 2048, 1024, 512, 256, 128, 128], 32, 8, 128
    g = torch.Generator(device=DEV).manual_seed(0)
    tot = sum(seqlens)
    q = torch.randn(tot, h, d, device=DEV, dtype=BF16, generator=g)
    k = torch.randn(tot, hk, d, device=DEV, dtype=BF16, generator=g); v = torch.randn(tot, hk, d, device=DEV, dtype=BF16, generator=g)
    cu = torch.tensor([0] + seqlens).cumsum(0).to(torch.int32).to(DEV)
    out = fa.flash_attn_varlen_func(q, k, v, cu, cu, 8192, 8192, causal=True)
    rk, _ = _util.ref_kernel_varlen(q, k, v, cu, cu, 8192, 8192, causal=True)
    grp = h // hk
    worst = (0.0, 0.0, 0.0)
    for kh in range(hk):
        qs, ks, vs = q[:8192, kh * grp:(kh + 1) * grp].unsqueeze(0), k[:8192, kh:kh + 1].unsqueeze(0), v[:8192, kh:kh + 1].unsqueeze(0)
        ref32 = orc.attention_ref(qs, ks, vs, causal=True)[0][0]
        pt = orc.attention_ref(qs, ks, vs, causal=True, upcast=False, reorder_ops=True)[0][0]
        e, e_pt, e_rk = _errs(out[:8192, kh * grp:(kh + 1) * grp], ref32, pt, rk[:8192, kh * grp:(kh + 1) * grp])
        assert e <= 2 * e_rk + 1e-5, (kh, e, e_rk, e_pt)
        assert e <= 2 * e_pt + 1e-5, (kh, e, e_pt)
        worst = max(worst, (e, e_rk, e_pt))
    # the rest of the batch against the reference kernel
    assert (out.float() - rk.float()).abs().max().item() < 2e-2, worst


# ------------------------------------------------------------------------------------------------ window >= seqlen_k
def _normalised_window(window, causal, seqlen_k, seqlen_q, alibi=False):
    """The reference glue's rewrite of (window, causal) before the kernel sees them (flash_api.cpp:289-294 and
    set_params_fprop :119-126), restated for the oracle: returns (causal, window) to hand to attention_ref.  Note that
    a one-sided window gets seqlen_k — not "unbounded" — on its open side (:123-124), which is NOT the same thing when
    seqlen_q > seqlen_k."""
    wl, wr = window
    if wl >= seqlen_k:
        wl = -1
    if wr >= seqlen_k:
        wr = -1
    if seqlen_q == 1 and not alibi:
        causal = False
    if causal:
        wr = 0
    if wl < 0 and wr == 0:
        return True, (-1, -1)
    if wl < 0 and wr >= 0:
        wl = seqlen_k
    if wl >= 0 and wr < 0:
        wr = seqlen_k
    return False, (wl, wr)


@pytest.mark.parametrize("sq,sk,window,causal", [
    (300, 200, (250, 10), False),     # left window >= seqlen_k: dropped -> (-1, 10)
    (300, 200, (20, 200), False),     # right window >= seqlen_k: dropped -> (20, -1)
    (300, 200, (200, 200), False),    # both dropped: plain attention
    (128, 500, (600, 0), False),      # (left dropped, right 0) IS the causal mask
    (700, 300, (300, -1), True),      # causal with a left window >= seqlen_k and sq > sk (rows without keys)
    (257, 257, (257, 5), False),
])
@pytest.mark.parametrize("d", [64, 128])
def test_window_at_or_above_seqlen_k_matches_the_reference_kernel(sq, sk, window, causal, d):
    """A window size >= seqlen_k is dropped by the reference's glue before the kernel runs (flash_api.cpp:289-290), which
    changes the mask when seqlen_q > seqlen_k and which attention_ref (called with the raw window) does not model.
    Our glue (fa_api.cu, make_plan) must make the same rewrite: same output as the reference kernel launched with the
    reference's normalised parameters, and the 2x rule against attention_ref evaluated with the normalised mask."""
    _need_refk()
    import vllm_flash_attn as fa
    torch.manual_seed(7)
    b, h, hk = 2, 4, 2
    q = torch.randn(b, sq, h, d, device=DEV, dtype=BF16)
    k = torch.randn(b, sk, hk, d, device=DEV, dtype=BF16); v = torch.randn_like(k)
    out, lse, _ = fa.flash_attn_func(q, k, v, causal=causal, window_size=window, return_attn_probs=True)
    rk, lrk = _util.ref_kernel_fwd(q, k, v, causal, window)
    c2, w2 = _normalised_window(window, causal, sk, sq)
    ref32, _ = orc.attention_ref(q, k, v, causal=c2, window_size=w2)
    pt, _ = orc.attention_ref(q, k, v, causal=c2, window_size=w2, upcast=False, reorder_ops=True)
    e, e_pt, e_rk = _errs(out, ref32, pt, rk)
    assert not torch.isnan(out.float()).any()
    assert e <= 2 * e_rk + 1e-5 and e <= 2 * e_pt + 1e-5, (e, e_rk, e_pt)
    assert (out.float() - rk.float()).abs().max().item() < 2e-2
    fin = torch.isfinite(lrk)
    assert bool((torch.isfinite(lse) == fin).all()) and (lse[fin] - lrk[fin]).abs().max().item() < 2e-3


def test_varlen_window_is_compared_with_the_longest_sequence():
    """mha_varlen_fwd compares the window with max_seqlen_k (flash_api.cpp:515-516), i.e. with the
    LONGEST sequence of the batch: a window between a short sequence's length and the longest one's stays in force for
    the short sequence.  Checked against the reference kernel on the same packed batch."""
    _need_refk()
    import vllm_flash_attn as fa
    torch.manual_seed(8)
    h, hk, d = 4, 2, 128
    lens_q, lens_k = [300, 150, 400], [200, 150, 600]
    cu_q = torch.tensor([0] + lens_q).cumsum(0).to(torch.int32).to(DEV); cu_k = torch.tensor([0] + lens_k).cumsum(0).to(torch.int32).to(DEV)
    q = torch.randn(sum(lens_q), h, d, device=DEV, dtype=BF16)
    k = torch.randn(sum(lens_k), hk, d, device=DEV, dtype=BF16); v = torch.randn_like(k)
    for window in ((250, 30), (600, 10), (40, 700)):
        out = fa.flash_attn_varlen_func(q, k, v, cu_q, cu_k, max(lens_q), max(lens_k), window_size=window)
        rk, _ = _util.ref_kernel_varlen(q, k, v, cu_q, cu_k, max(lens_q), max(lens_k), False, window)
        assert not torch.isnan(out.float()).any()
        assert (out.float() - rk.float()).abs().max().item() < 2e-2, window


# ------------------------------------------------------------------------------------------------ paged overflow
@pytest.mark.parametrize("causal", [False, True])
def test_paged_kvcache_overflow(causal):
    """Port of test_flash_attn_paged_kvcache_overflow (tests/test_flash_attn.py:2466-2507): seqlen 170 on 16-row pages
    with a block table of exactly ceil(170/16) = 11 columns, new K/V appended into an empty paged cache, 10 rounds.
    The reference's 128-key tiles compute table addresses up to 7 columns past the end; here the index is
    clamped (fa_decode.cu page_of, fa_prefill_tc5.cu paged producer).  The reference test only requires the call not
    to fault; this one also checks the cache contents and the output (block ids drawn WITHOUT replacement, so that the
    result is defined), with every pool row outside the table's blocks poisoned with NaN."""
    import vllm_flash_attn as fa
    torch.manual_seed(9)
    dt, page, h, d, b, sq = torch.float16, 16, 32, 128, 4, 170
    num_blocks = 1000
    ncol = (sq + page - 1) // page
    for rnd in range(10):
        kc = torch.full((num_blocks, page, h, d), float("nan"), dtype=dt, device=DEV); vc = torch.full_like(kc, float("nan"))
        lens = torch.zeros(b, dtype=torch.int32, device=DEV)
        q = torch.rand(b, sq, h, d, dtype=dt, device=DEV)
        kn = torch.rand(b, sq, h, d, dtype=dt, device=DEV); vn = torch.rand(b, sq, h, d, dtype=dt, device=DEV)
        bt = torch.randperm(num_blocks, device=DEV)[: b * ncol].to(torch.int32).view(b, ncol).contiguous()
        out = fa.flash_attn_with_kvcache(q, kc, vc, k=kn, v=vn, cache_seqlens=lens, block_table=bt, causal=causal)
        torch.cuda.synchronize()
        assert not torch.isnan(out.float()).any(), rnd
        kg, vg = orc.gather_paged(kc, bt)[:, :sq], orc.gather_paged(vc, bt)[:, :sq]
        assert torch.equal(kg, kn) and torch.equal(vg, vn)                      # appended rows landed where the table says
        touched = torch.zeros(num_blocks, dtype=torch.bool, device=DEV); touched[bt.view(-1).long()] = True
        assert bool(torch.isnan(kc[~touched].float()).all())                    # and nowhere else
        if rnd in (0, 9):
            ref32, _ = orc.attention_ref(q, kn, vn, causal=causal)
            pt, _ = orc.attention_ref(q, kn, vn, causal=causal, upcast=False, reorder_ops=True)
            e, e_pt, _ = _errs(out, ref32, pt)
            assert e <= 3 * e_pt + 1e-5, (e, e_pt)


# ------------------------------------------------------------------------------------------------ varlen packed wrappers
def test_varlen_packed_wrappers():
    """flash_attn_varlen_qkvpacked_func / flash_attn_varlen_kvpacked_func (flash_attn_interface.py:870-1023) are thin
    slicers over flash_attn_varlen_func: same result bit for bit."""
    import vllm_flash_attn as fa
    torch.manual_seed(10)
    h, d = 4, 64
    lens = [100, 257, 31]
    cu = torch.tensor([0] + lens).cumsum(0).to(torch.int32).to(DEV)
    qkv = torch.randn(sum(lens), 3, h, d, device=DEV, dtype=torch.float16)
    o0 = fa.flash_attn_varlen_func(qkv[:, 0], qkv[:, 1], qkv[:, 2], cu, cu, max(lens), max(lens), causal=True)
    o1 = fa.flash_attn_varlen_qkvpacked_func(qkv, cu, max(lens), causal=True)
    o2 = fa.flash_attn_varlen_kvpacked_func(qkv[:, 0], qkv[:, 1:], cu, cu, max(lens), max(lens), causal=True)
    assert torch.equal(o0, o1) and torch.equal(o0, o2)
    # GQA through the kv-packed wrapper, different q / k lengths, with the lse returned
    lens_k = [300, 257, 64]
    cu_k = torch.tensor([0] + lens_k).cumsum(0).to(torch.int32).to(DEV)
    kv = torch.randn(sum(lens_k), 2, 2, d, device=DEV, dtype=torch.float16)
    o3, l3, _ = fa.flash_attn_varlen_kvpacked_func(qkv[:, 0], kv, cu, cu_k, max(lens), max(lens_k), return_attn_probs=True)
    o4, l4, _ = fa.flash_attn_varlen_func(qkv[:, 0], kv[:, 0], kv[:, 1], cu, cu_k, max(lens), max(lens_k), return_attn_probs=True)
    assert torch.equal(o3, o4)
    for i, n in enumerate(lens):        # lse is (b, h, max_seqlen_q): columns past a sequence's length are never written
        assert torch.equal(l3[i, :, :n], l4[i, :, :n])
    ref32, _ = orc.attention_ref_varlen(qkv[:, 0], kv[:, 0], kv[:, 1], cu.cpu(), cu_k.cpu())
    pt, _ = orc.attention_ref_varlen(qkv[:, 0], kv[:, 0], kv[:, 1], cu.cpu(), cu_k.cpu(), upcast=False, reorder_ops=True)
    e, e_pt, _ = _errs(o3, ref32, pt)
    assert e <= 2 * e_pt + 1e-5
