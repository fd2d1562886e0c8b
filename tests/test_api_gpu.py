"""GPU: behaviour of the drop-in boundary — auto dispatch, out=, errors, determinism, combine,
append-KV + rotary, full-size configs against the reference kernel."""
import ctypes as C
import math

import pytest
import torch

import _util
import attention_oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _auto_kernel():
    from vllm_flash_attn import vllm_flash_attn_2_cuda as ext
    ext._force_kernel = 0
    yield
    ext._force_kernel = 0


def _ext():
    from vllm_flash_attn import vllm_flash_attn_2_cuda as ext
    return ext


def test_auto_dispatch_picks_expected_kernels():
    import vllm_flash_attn as fa
    dev, dt = "cuda", torch.bfloat16
    q = torch.randn(2, 512, 8, 64, device=dev, dtype=dt); k = torch.randn_like(q); v = torch.randn_like(q)
    fa.flash_attn_func(q, k, v)
    assert _ext().last_call_info["kernel_used"] == 1          # tcgen05 prefill
    q1 = torch.randn(4, 1, 32, 128, device=dev, dtype=dt)
    kc = torch.randn(4, 1024, 8, 128, device=dev, dtype=dt); vc = torch.randn_like(kc)
    fa.flash_attn_with_kvcache(q1, kc, vc, cache_seqlens=torch.full((4,), 1000, dtype=torch.int32, device=dev))
    assert _ext().last_call_info["kernel_used"] == 3          # decode
    q2 = torch.randn(1, 200, 4, 96, device=dev, dtype=dt); k2 = torch.randn_like(q2)
    fa.flash_attn_func(q2, k2, k2)
    assert _ext().last_call_info["kernel_used"] == 1          # tcgen05 prefill (d=96 in the D=128 instantiation)
    q3 = torch.randn(1, 200, 4, 160, device=dev, dtype=dt); k3 = torch.randn_like(q3)
    fa.flash_attn_func(q3, k3, k3)
    assert _ext().last_call_info["kernel_used"] == 1          # tcgen05 prefill, one-tile units (d > 128)
    kc3 = torch.randn(3, 400, 4, 160, device=dev, dtype=dt)
    fa.flash_attn_with_kvcache(q3, kc3, kc3, cache_seqlens=torch.tensor([300], dtype=torch.int32, device=dev),
                               cache_batch_idx=torch.tensor([2], dtype=torch.int32, device=dev))
    assert _ext().last_call_info["kernel_used"] == 1          # non-paged cache as one page per sequence (cache_batch_idx = its block table)
    q4 = torch.randn(2, 1, 8, 192, device=dev, dtype=dt); k4 = torch.randn(2, 300, 2, 192, device=dev, dtype=dt)
    fa.flash_attn_with_kvcache(q4, k4, k4)
    assert _ext().last_call_info["kernel_used"] == 3          # decode, every head dim of the reference build


def test_c1_full_and_out_argument():
    import vllm_flash_attn as fa
    torch.manual_seed(0)
    q = torch.randn(2, 512, 8, 64, device="cuda", dtype=torch.bfloat16); k = torch.randn_like(q); v = torch.randn_like(q)
    out_buf = torch.empty_like(q)
    out = fa.flash_attn_func(q, k, v, out=out_buf)
    assert out.data_ptr() == out_buf.data_ptr()
    ref32, _ = orc.attention_ref(q, k, v)
    pt, _ = orc.attention_ref(q, k, v, upcast=False, reorder_ops=True)
    assert (out.float() - ref32.float()).abs().max() <= 2 * (pt.float() - ref32.float()).abs().max() + 1e-5
    if _util.ref_lib() is not None:
        rk, _ = _util.ref_kernel_fwd(q, k, v)
        assert (out.float() - ref32.float()).abs().max() <= 2 * (rk.float() - ref32.float()).abs().max() + 1e-5


def test_errors_raise_runtime_error():
    import vllm_flash_attn as fa
    q = torch.randn(1, 8, 4, 64, device="cuda", dtype=torch.float32)
    with pytest.raises(RuntimeError, match="fp16 and bf16"):
        fa.flash_attn_func(q, q, q)
    q = q.half()
    with pytest.raises(RuntimeError, match="must divide"):
        fa.flash_attn_func(q, q[:, :, :3], q[:, :, :3])
    with pytest.raises(RuntimeError, match="dropout"):
        fa.flash_attn_func(q, q, q, dropout_p=0.1)
    kc = torch.randn(6, 24, 4, 64, device="cuda", dtype=torch.float16)
    bt = torch.zeros(1, 2, dtype=torch.int32, device="cuda")
    with pytest.raises(RuntimeError, match="divisible by 16"):
        fa.flash_attn_with_kvcache(q, kc, kc, cache_seqlens=torch.tensor([20], dtype=torch.int32, device="cuda"), block_table=bt)


def test_head_dim_padding_and_packed_wrappers():
    import vllm_flash_attn as fa
    torch.manual_seed(1)
    qkv = torch.randn(2, 100, 3, 4, 40, device="cuda", dtype=torch.float16)     # d=40 -> padded to 64? (d%32 rule)
    with pytest.raises(RuntimeError):
        fa.flash_attn_qkvpacked_func(qkv)          # reference build: headdim must be a multiple of 32
    qkv = torch.randn(2, 100, 3, 4, 64, device="cuda", dtype=torch.float16)
    o1 = fa.flash_attn_qkvpacked_func(qkv, causal=True)
    o2 = fa.flash_attn_func(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], causal=True)
    assert torch.equal(o1, o2)
    o3 = fa.flash_attn_kvpacked_func(qkv[:, :, 0], qkv[:, :, 1:], causal=True)
    assert torch.equal(o1, o3)


@pytest.mark.parametrize("kind", ["prefill", "decode_split"])
def test_bitwise_deterministic(kind):
    """tests/test_flash_attn.py:2107-2173 (race condition test): repeated calls are bit-identical."""
    import vllm_flash_attn as fa
    torch.manual_seed(0)
    dev, dt = "cuda", torch.bfloat16
    if kind == "prefill":
        q = torch.randn(2, 1000, 8, 128, device=dev, dtype=dt); k = torch.randn(2, 1000, 2, 128, device=dev, dtype=dt); v = torch.randn_like(k)
        f = lambda: fa.flash_attn_func(q, k, v, causal=True, return_attn_probs=True)[:2]
    else:
        q = torch.randn(1, 1, 32, 128, device=dev, dtype=dt)
        kc = torch.randn(1, 20000, 8, 128, device=dev, dtype=dt); vc = torch.randn_like(kc)
        cs = torch.tensor([19999], dtype=torch.int32, device=dev)
        f = lambda: fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=cs, return_softmax_lse=True)
    o0, l0 = f()
    for _ in range(30):
        o, l = f()
        assert torch.equal(o, o0) and torch.equal(l, l0)


def test_combine_entry_point():
    from vllm_flash_attn._cabi import lib, check
    torch.manual_seed(0)
    n, rows, d = 5, 37, 128
    o_p = torch.randn(n, rows, d, device="cuda")
    l_p = torch.randn(n, rows, device="cuda") * 3
    l_p[2] = float("-inf")
    l_p[:, 5] = float("-inf")
    o = torch.empty(rows, d, device="cuda", dtype=torch.bfloat16); lse = torch.empty(rows, device="cuda")
    check(lib.fa_b200_combine(o_p.data_ptr(), l_p.data_ptr(), o.data_ptr(), lse.data_ptr(), n, rows, d, rows * d, rows, 0, 1,
                              torch.cuda.current_stream().cuda_stream))
    oc, lc = orc.combine_ref(o_p.cpu(), l_p.cpu())
    assert (o.float().cpu() - oc).abs().max() < 2e-2
    assert torch.isinf(lse[5]) and lse[5] > 0 and (o[5] == 0).all()
    fin = torch.isfinite(lc)
    assert (lse.cpu()[fin] - lc[fin]).abs().max() < 1e-4


@pytest.mark.parametrize("mode", ["fused", "two_step"])
@pytest.mark.parametrize("paged", [False, True])
@pytest.mark.parametrize("interleaved", [True, False])
@pytest.mark.parametrize("rotary_frac", [0.0, 0.5, 1.0])
def test_append_kv_and_rotary(paged, interleaved, rotary_frac, mode, monkeypatch):
    """tests/test_flash_attn.py:1944-2074: new k/v are appended at cache_seqlens (rotary applied),
    cache contents are checked afterwards, output matches attention over the updated cache.  `fused`: the decode
    kernel appends and rotates itself (the reference's fused path, flash_fwd_kernel.h:672-824); `two_step`: the
    stand-alone append kernel runs first (what non-decode shapes use)."""
    import vllm_flash_attn as fa
    from vllm_flash_attn import vllm_flash_attn_2_cuda as ext
    if mode == "two_step":
        monkeypatch.setattr(ext, "_two_step_append", True)
    torch.manual_seed(0)
    dev, dt = "cuda", torch.float16
    b, sq, sk, skn, h, hk, d = 2, 3, 160, 3, 4, 2, 64
    q = torch.randn(b, sq, h, d, device=dev, dtype=dt)
    kn = torch.randn(b, skn, hk, d, device=dev, dtype=dt); vn = torch.randn_like(kn)
    lens = torch.tensor([100, 37], dtype=torch.int32, device=dev)
    bt = None
    if paged:
        nblk = (sk // 16) * b * 2
        kc = torch.randn(nblk, 16, hk, d, device=dev, dtype=dt); vc = torch.randn_like(kc)
        bt = torch.randperm(nblk, device=dev).to(torch.int32).view(b, -1)[:, :sk // 16].contiguous()
        kfull, vfull = orc.gather_paged(kc, bt).clone(), orc.gather_paged(vc, bt).clone()
    else:
        kc = torch.randn(b, sk, hk, d, device=dev, dtype=dt); vc = torch.randn_like(kc)
        kfull, vfull = kc.clone(), vc.clone()
    rd = int(rotary_frac * d) // 16 * 16
    cos = sin = None
    q_ro, kn_ro = q, kn
    if rd > 0:
        ang = torch.rand(sk, rd // 2, device=dev) * 2 * math.pi
        cos, sin = torch.cos(ang).to(dt), torch.sin(ang).to(dt)

        def rot(x, pos):                     # x (b,s,h,d), pos (b,s)
            c = cos.float()[pos].unsqueeze(2); s_ = sin.float()[pos].unsqueeze(2)
            xr = x.float().clone()
            if interleaved:
                x0, x1 = xr[..., 0:rd:2], xr[..., 1:rd:2]
                xr[..., 0:rd:2], xr[..., 1:rd:2] = x0 * c - x1 * s_, x0 * s_ + x1 * c
            else:
                x0, x1 = xr[..., :rd // 2].clone(), xr[..., rd // 2:rd].clone()
                xr[..., :rd // 2], xr[..., rd // 2:rd] = x0 * c - x1 * s_, x0 * s_ + x1 * c
            return xr.to(dt)
        pos_k = lens.long().view(b, 1) + torch.arange(skn, device=dev).view(1, -1)
        kn_ro = rot(kn, pos_k)
        pos_q = lens.long().view(b, 1) + torch.arange(sq, device=dev).view(1, -1)   # causal: by position
        q_ro = rot(q, pos_q)
    for i in range(b):
        L = int(lens[i])
        kfull[i, L:L + skn] = kn_ro[i]; vfull[i, L:L + skn] = vn[i]
    out = fa.flash_attn_with_kvcache(q, kc, vc, kn, vn, rotary_cos=cos, rotary_sin=sin, cache_seqlens=lens,
                                     block_table=bt, causal=True, rotary_interleaved=interleaved)
    kafter = orc.gather_paged(kc, bt) if paged else kc
    vafter = orc.gather_paged(vc, bt) if paged else vc
    assert torch.allclose(kafter.float(), kfull.float(), atol=2e-3, rtol=1e-3)
    assert torch.equal(vafter, vfull)
    S = kfull.shape[1]
    kpm = torch.arange(S, device=dev).expand(b, S) < (lens + skn).view(-1, 1)
    ref32, _ = orc.attention_ref(q_ro, kfull, vfull, None, kpm, causal=True)
    pt, _ = orc.attention_ref(q_ro, kfull, vfull, None, kpm, causal=True, upcast=False, reorder_ops=True)
    assert (out.float() - ref32.float()).abs().max() <= 3 * (pt.float() - ref32.float()).abs().max() + 1e-5


@pytest.mark.parametrize("sq,skn,causal", [(1, 1, True), (2, 2, True), (1, 3, False)])
def test_fused_append_long_context_with_splits(sq, skn, causal, monkeypatch):
    """In-kernel append on a split decode: the new rows are written by the ONE split CTA whose chunk range holds them
    (here the last split, or — for the length that ends on a chunk boundary — a CTA that starts with them), every other
    CTA never touches them.  Output, LSE and the cache afterwards are bit-identical to the two-step path (same fp32
    rotation, same attention kernel over the same cache contents); q and the new k/v themselves are left untouched."""
    import vllm_flash_attn as fa
    from vllm_flash_attn import vllm_flash_attn_2_cuda as ext
    torch.manual_seed(5)
    dev, dt, page = "cuda", torch.bfloat16, 16
    b, h, hk, d, ctx = 3, 16, 2, 128, 8192
    mb = ctx // page
    lens = torch.tensor([5000, ctx - skn, 4096], dtype=torch.int32, device=dev)      # mid-page, cache exactly filled, chunk boundary
    kc0 = torch.randn(b * mb, page, hk, d, device=dev, dtype=dt); vc0 = torch.randn_like(kc0)
    bt = torch.randperm(b * mb, device=dev).to(torch.int32).view(b, mb).contiguous()
    q = torch.randn(b, sq, h, d, device=dev, dtype=dt); q_keep = q.clone()
    kn = torch.randn(b, skn, hk, d, device=dev, dtype=dt); vn = torch.randn_like(kn); kn_keep = kn.clone()
    ang = torch.rand(ctx, d // 2, device=dev) * 2 * math.pi
    cos, sin = torch.cos(ang).to(dt), torch.sin(ang).to(dt)

    def run(two_step):
        monkeypatch.setattr(ext, "_two_step_append", two_step)
        kc, vc = kc0.clone(), vc0.clone()
        o, l = fa.flash_attn_with_kvcache(q, kc, vc, kn, vn, rotary_cos=cos, rotary_sin=sin, cache_seqlens=lens, block_table=bt,
                                          causal=causal, rotary_interleaved=False, return_softmax_lse=True)
        return o, l, kc, vc, dict(ext.last_call_info)
    of, lf, kf, vf, info_f = run(False)
    ot, lt, kt, vt, info_t = run(True)
    assert info_f["kernel_used"] == 3 and info_f["num_splits"] > 1 and info_f == info_t, (info_f, info_t)
    assert torch.equal(kf, kt) and torch.equal(vf, vt) and not torch.equal(kf, kc0)
    assert torch.equal(of, ot) and torch.equal(lf, lt)
    assert torch.equal(q, q_keep) and torch.equal(kn, kn_keep)


def test_fused_append_with_cache_batch_idx_and_alibi(monkeypatch):
    """In-kernel append into a NON-paged cache addressed through cache_batch_idx (the rows must land in the indexed cache
    entry), with ALiBi and interleaved partial rotary, fp16: bit-identical to the two-step path, other entries untouched."""
    import vllm_flash_attn as fa
    from vllm_flash_attn import vllm_flash_attn_2_cuda as ext
    torch.manual_seed(9)
    dev, dt = "cuda", torch.float16
    b, bc, sq, skn, h, hk, d, sk = 3, 7, 1, 2, 8, 4, 64, 384
    idx = torch.tensor([5, 0, 3], dtype=torch.int32, device=dev)
    lens = torch.tensor([100, 382, 0], dtype=torch.int32, device=dev)       # middle, cache exactly filled, empty cache
    kc0 = torch.randn(bc, sk, hk, d, device=dev, dtype=dt); vc0 = torch.randn_like(kc0)
    q = torch.randn(b, sq, h, d, device=dev, dtype=dt)
    kn = torch.randn(b, skn, hk, d, device=dev, dtype=dt); vn = torch.randn_like(kn)
    ang = torch.rand(sk, 16, device=dev) * 2 * math.pi                         # rotary_dim 32 of 64
    cos, sin = torch.cos(ang).to(dt), torch.sin(ang).to(dt)
    slopes = torch.rand(h, device=dev, dtype=torch.float32) * 0.2

    def run(two_step):
        monkeypatch.setattr(ext, "_two_step_append", two_step)
        kc, vc = kc0.clone(), vc0.clone()
        o, l = fa.flash_attn_with_kvcache(q, kc, vc, kn, vn, rotary_cos=cos, rotary_sin=sin, cache_seqlens=lens, cache_batch_idx=idx,
                                          causal=True, rotary_interleaved=True, alibi_slopes=slopes, return_softmax_lse=True)
        return o, l, kc, vc
    of, lf, kf, vf = run(False)
    ot, lt, kt, vt = run(True)
    assert torch.equal(kf, kt) and torch.equal(vf, vt) and torch.equal(of, ot) and torch.equal(lf, lt)
    untouched = [i for i in range(bc) if i not in (5, 0, 3)]
    assert torch.equal(kf[untouched], kc0[untouched]) and torch.equal(vf[untouched], vc0[untouched])
    for i, e in enumerate((5, 0, 3)):
        L = int(lens[i])
        assert torch.equal(vf[e, L:L + skn], vn[i]) and torch.equal(vf[e, :L], vc0[e, :L])


def test_append_on_a_prefill_shape_takes_the_two_step_path():
    """64 new tokens per sequence against a cache (chunked prefill through flash_attn_with_kvcache): not a decode shape,
    so fa_b200_fused_append_supported says no and the bindings run the append kernel first; a sliding window does the
    same on a decode shape.  Both host bindings, same results."""
    import vllm_flash_attn as fa
    from vllm_flash_attn import _cabi
    from vllm_flash_attn import vllm_flash_attn_2_cuda as ext
    torch.manual_seed(6)
    dev, dt = "cuda", torch.bfloat16
    b, sq, h, hk, d, sk = 2, 64, 8, 2, 128, 512
    q = torch.randn(b, sq, h, d, device=dev, dtype=dt)
    kn = torch.randn(b, sq, hk, d, device=dev, dtype=dt); vn = torch.randn_like(kn)
    kc0 = torch.randn(b, sk, hk, d, device=dev, dtype=dt); vc0 = torch.randn_like(kc0)
    lens = torch.tensor([300, 17], dtype=torch.int32, device=dev)
    outs = []
    for native in (True, False):
        if native and ext._NATIVE is None:
            continue
        saved = ext._N
        ext._N = ext._NATIVE if native else None
        try:
            for window in ((-1, -1), (100, 0)):
                qq = q if window == (-1, -1) else q[:, :1]
                kk, vv = (kn, vn) if window == (-1, -1) else (kn[:, :1], vn[:, :1])
                kc, vc = kc0.clone(), vc0.clone()
                o = fa.flash_attn_with_kvcache(qq, kc, vc, kk, vv, cache_seqlens=lens, causal=True, window_size=window)
                full_k, full_v = kc0.clone(), vc0.clone()
                for i in range(b):
                    L = int(lens[i]); full_k[i, L:L + kk.shape[1]] = kk[i]; full_v[i, L:L + kk.shape[1]] = vv[i]
                assert torch.equal(kc, full_k) and torch.equal(vc, full_v)
                outs.append((native, window, o))
        finally:
            ext._N = saved
    if len(outs) == 4:
        assert torch.equal(outs[0][2], outs[2][2]) and torch.equal(outs[1][2], outs[3][2])


def test_cache_batch_idx_and_seqlen_k_zero():
    import vllm_flash_attn as fa
    torch.manual_seed(0)
    dev, dt = "cuda", torch.bfloat16
    q = torch.randn(2, 1, 8, 128, device=dev, dtype=dt)
    kc = torch.randn(5, 300, 2, 128, device=dev, dtype=dt); vc = torch.randn_like(kc)
    idx = torch.tensor([3, 1], dtype=torch.int32, device=dev)
    lens = torch.tensor([300, 123], dtype=torch.int32, device=dev)
    out = fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=lens, cache_batch_idx=idx)
    kpm = torch.arange(300, device=dev).expand(2, 300) < lens.view(-1, 1)
    ref32, _ = orc.attention_ref(q, kc[idx.long()], vc[idx.long()], None, kpm)
    assert (out.float() - ref32.float()).abs().max() < 2e-2
    o = fa.flash_attn_func(q, kc[:2, :0], vc[:2, :0])
    assert (o == 0).all()


@pytest.mark.parametrize("cfg", ["C2_varlen_prefill", "C3_paged_decode"])
def test_full_size_configs_vs_reference_kernel(cfg):
    """BASELINE.json configs at full size: our output vs the reference FA2 kernel built for sm_100
    (size-independent check: both must agree within bf16 rounding of outputs of magnitude <~ 1)."""
    if _util.ref_lib() is None:
        pytest.skip("oracle/_ref not built")
    import vllm_flash_attn as fa
    torch.manual_seed(0)
    dev, dt = "cuda", torch.bfloat16
    if cfg == "C2_varlen_prefill":
        lens = [8192, 4096, 2048, 1024, 512, 256, 128, 128]
        cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device=dev)
        q = torch.randn(16384, 32, 128, device=dev, dtype=dt)
        k = torch.randn(16384, 8, 128, device=dev, dtype=dt); v = torch.randn_like(k)
        out, lse, _ = fa.flash_attn_varlen_func(q, k, v, cu, cu, 8192, 8192, causal=True, return_attn_probs=True)
        assert _ext().last_call_info["kernel_used"] == 1
        rk, rl = _util.ref_kernel_varlen(q, k, v, cu, cu, 8192, 8192, causal=True)
        # spot-check the shortest sequence against the fp32 oracle too
        s0 = int(cu[-2]); qs, ks, vs = q[s0:].unsqueeze(0), k[s0:].unsqueeze(0), v[s0:].unsqueeze(0)
        ref32, _ = orc.attention_ref(qs, ks, vs, causal=True)
        e_new = (out[s0:].float() - ref32[0].float()).abs().max(); e_rk = (rk[s0:].float() - ref32[0].float()).abs().max()
        assert e_new <= 2 * e_rk + 1e-5
    else:
        b, ctx, page = 256, 4096, 16
        nblk = b * (ctx // page) * 5 // 4
        kc = torch.randn(nblk, page, 8, 128, device=dev, dtype=dt); vc = torch.randn_like(kc)
        bt = torch.randperm(nblk, device=dev)[: b * ctx // page].to(torch.int32).view(b, -1).contiguous()
        q = torch.randn(b, 1, 32, 128, device=dev, dtype=dt)
        lens = torch.randint(2048, ctx + 1, (b,), device=dev).to(torch.int32)
        out, lse = fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=lens, block_table=bt, causal=True, return_softmax_lse=True)
        assert _ext().last_call_info["kernel_used"] == 3
        rk, rl = _util.ref_kernel_kvcache(q, kc, vc, lens, bt, causal=True)
    assert not torch.isnan(out.float()).any()
    assert (out.float() - rk.float()).abs().max() < 2e-2
    n = min(lse.shape[-1], rl.shape[-1])
    fin = torch.isfinite(rl[..., :n])
    if cfg == "C2_varlen_prefill":
        # varlen lse is (b,h,max_seqlen_q), padded: positions past each sequence's length are
        # never written by either kernel (flash_api.cpp:578 allocates with torch::empty)
        valid = torch.arange(n, device=dev).view(1, 1, -1) < torch.tensor(lens, device=dev).view(-1, 1, 1)
        fin = fin & valid
    assert (lse[..., :n][fin] - rl[..., :n][fin]).abs().max() < 2e-3


@pytest.mark.parametrize("d", [96, 192, 256])
def test_full_size_other_head_dims_two_kernels_agree(d):
    """Head dims for which oracle/_ref holds no reference build: at full size (16k-token varlen causal prefill,
    batch-128 paged decode) the tcgen05 / decode kernels must agree with the independent mma.sync tile kernel
    (different tiling, different softmax schedule) within bf16 rounding, LSE within 2e-3; the shortest sequence is
    also checked against the fp32 oracle with the reference's 2x criterion."""
    import vllm_flash_attn as fa
    ext = _ext()
    torch.manual_seed(0)
    dev, dt = "cuda", torch.bfloat16
    h, hk = 16, 4
    lens = [8192, 4096, 2048, 1024, 512, 256, 128, 128]
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device=dev)
    q = torch.randn(16384, h, d, device=dev, dtype=dt)
    k = torch.randn(16384, hk, d, device=dev, dtype=dt); v = torch.randn_like(k)
    out, lse, _ = fa.flash_attn_varlen_func(q, k, v, cu, cu, 8192, 8192, causal=True, return_attn_probs=True)
    assert ext.last_call_info["kernel_used"] == 1
    ext._force_kernel = 2
    out_t, lse_t, _ = fa.flash_attn_varlen_func(q, k, v, cu, cu, 8192, 8192, causal=True, return_attn_probs=True)
    ext._force_kernel = 0
    assert not torch.isnan(out.float()).any()
    assert (out.float() - out_t.float()).abs().max() < 2e-2
    valid = torch.arange(lse.shape[-1], device=dev).view(1, 1, -1) < torch.tensor(lens, device=dev).view(-1, 1, 1)
    assert (lse - lse_t)[valid.expand_as(lse)].abs().max() < 2e-3
    s0 = int(cu[-2])
    ref32, _ = orc.attention_ref(q[s0:].unsqueeze(0), k[s0:].unsqueeze(0), v[s0:].unsqueeze(0), causal=True)
    pt, _ = orc.attention_ref(q[s0:].unsqueeze(0), k[s0:].unsqueeze(0), v[s0:].unsqueeze(0), causal=True, upcast=False, reorder_ops=True)
    assert (out[s0:].float() - ref32[0].float()).abs().max() <= 2 * (pt[0].float() - ref32[0].float()).abs().max() + 1e-5

    b, ctx, page = 128, 4096, 16
    nblk = b * (ctx // page) * 5 // 4
    kc = torch.randn(nblk, page, hk, d, device=dev, dtype=dt); vc = torch.randn_like(kc)
    bt = torch.randperm(nblk, device=dev)[: b * ctx // page].to(torch.int32).view(b, -1).contiguous()
    q1 = torch.randn(b, 1, h, d, device=dev, dtype=dt)
    lens1 = torch.randint(2048, ctx + 1, (b,), device=dev).to(torch.int32)
    o1, l1 = fa.flash_attn_with_kvcache(q1, kc, vc, cache_seqlens=lens1, block_table=bt, causal=True, return_softmax_lse=True)
    assert ext.last_call_info["kernel_used"] == 3
    ext._force_kernel = 2
    o2, l2 = fa.flash_attn_with_kvcache(q1, kc, vc, cache_seqlens=lens1, block_table=bt, causal=True, return_softmax_lse=True)
    ext._force_kernel = 0
    assert (o1.float() - o2.float()).abs().max() < 1e-2 and (l1 - l2).abs().max() < 2e-3


def _varlen_check(lens_q, lens_k, h=8, hk=2, d=128, causal=True, seed=0, tol_mult=2):
    """flash_attn_varlen_func on explicit (possibly empty) sequence lengths vs the fp32 oracle."""
    import vllm_flash_attn as fa
    from vllm_flash_attn import vllm_flash_attn_2_cuda as ext
    torch.manual_seed(seed)
    dev, dt = "cuda", torch.bfloat16
    cu_q = torch.tensor([0] + list(torch.tensor(lens_q).cumsum(0)), dtype=torch.int32)
    cu_k = torch.tensor([0] + list(torch.tensor(lens_k).cumsum(0)), dtype=torch.int32)
    q = torch.randn(int(cu_q[-1]), h, d, device=dev, dtype=dt)
    k = torch.randn(max(int(cu_k[-1]), 1), hk, d, device=dev, dtype=dt)[: int(cu_k[-1])]
    v = torch.randn_like(k)
    out = fa.flash_attn_varlen_func(q, k, v, cu_q.to(dev), cu_k.to(dev), max(lens_q), max(lens_k), causal=causal)
    info = dict(ext.last_call_info)
    ref32, _ = orc.attention_ref_varlen(q, k, v, cu_q, cu_k, causal=causal)
    pt, _ = orc.attention_ref_varlen(q, k, v, cu_q, cu_k, causal=causal, upcast=False, reorder_ops=True)
    err = (out.float() - ref32.float()).abs().max().item()
    err_pt = (pt.float() - ref32.float()).abs().max().item()
    assert not torch.isnan(out.float()).any()
    assert err <= tol_mult * err_pt + 1e-5, (err, err_pt, info)
    return info


def test_varlen_empty_sequences_through_the_work_list():
    """Zero-length q and/or k sequences inside a varlen batch (reference BlockInfo semantics,
    src/block_info.h:11-42): no work unit for empty q, O = 0 for empty k; the batch is large enough
    (> 148 units) that the persistent kernel runs from the device-built work list."""
    info = _varlen_check([600, 0, 1, 257, 0, 128, 300, 513, 40], [600, 50, 700, 257, 0, 0, 300, 513, 900], h=16, hk=4)
    assert info["kernel_used"] == 1 and info["launches"] == 4     # schedule + prefill + decode (16-row form) + decode (32-row form)
    info = _varlen_check([130, 0, 70], [130, 64, 0], h=4, hk=4, causal=False)
    assert info["kernel_used"] == 1
    # one-tile units (d > 128): 128-row work units, Q through the K/V ring, empty rows / sequences in between
    info = _varlen_check([600, 0, 1, 257, 0, 128, 300, 513, 40], [600, 50, 700, 257, 0, 0, 300, 513, 900], h=16, hk=4, d=256)
    assert info["kernel_used"] == 1 and info["launches"] == 3     # d > 128: the decode kernel holds 16 packed rows, one form
    info = _varlen_check([130, 0, 70, 300], [130, 64, 0, 200], h=4, hk=2, d=192)
    assert info["kernel_used"] == 1


def test_mixed_batch_with_few_kv_heads_splits_its_decode_half():
    """A tensor-parallel shard of a mixed batch (C4 at TP = 8: 8 q heads / 1 kv head): one decode CTA per (sequence,
    kv head) would leave most SMs idle, so the decode half is split along the context and merged inside the decode
    kernel — same launches as before, partials behind the work list in the same workspace.  Paged cache with NaN in
    every row past a sequence's length; empty and one-key contexts next to 4096-key ones."""
    import vllm_flash_attn as fa
    from vllm_flash_attn import vllm_flash_attn_2_cuda as ext
    torch.manual_seed(11)
    dev, dt, h, hk, d, page = "cuda", torch.bfloat16, 8, 1, 128, 16
    lens_q = [300, 1, 1, 2, 1, 700, 1, 1, 1, 2]
    lens_k = [300, 3000, 4096, 1500, 17, 700, 1, 0, 2500, 4095]
    nb = len(lens_q); mb = (max(lens_k) + page - 1) // page
    cu_q = torch.tensor([0] + list(torch.tensor(lens_q).cumsum(0)), dtype=torch.int32)
    cu_k = torch.tensor([0] + list(torch.tensor(lens_k).cumsum(0)), dtype=torch.int32)
    q = torch.randn(sum(lens_q), h, d, device=dev, dtype=dt)
    k = torch.randn(sum(lens_k), hk, d, device=dev, dtype=dt); v = torch.randn_like(k)
    bt = torch.randperm(nb * mb, device=dev).to(torch.int32).view(nb, mb).contiguous()
    kc = torch.full((nb * mb, page, hk, d), float("nan"), device=dev, dtype=dt); vc = torch.full_like(kc, float("nan"))
    for i in range(nb):
        for j in range(lens_k[i]):
            blk = int(bt[i, j // page])
            kc[blk, j % page] = k[int(cu_k[i]) + j]; vc[blk, j % page] = v[int(cu_k[i]) + j]
    out = fa.flash_attn_varlen_func(q, kc, vc, cu_q.to(dev), cu_k.to(dev), max(lens_q), max(lens_k), causal=True, block_table=bt)
    info = dict(ext.last_call_info)
    assert info["kernel_used"] == 1 and info["launches"] == 3 and info["num_splits"] > 1, info
    ref32, _ = orc.attention_ref_varlen(q, k, v, cu_q, cu_k, causal=True)
    pt, _ = orc.attention_ref_varlen(q, k, v, cu_q, cu_k, causal=True, upcast=False, reorder_ops=True)
    err = (out.float() - ref32.float()).abs().max().item(); err_pt = (pt.float() - ref32.float()).abs().max().item()
    assert not torch.isnan(out.float()).any()
    assert err <= 2 * err_pt + 1e-5, (err, err_pt, info)
    # twice more through the same cached workspace: the arrival counters must be back at zero, results identical
    for _ in range(2):
        again = fa.flash_attn_varlen_func(q, kc, vc, cu_q.to(dev), cu_k.to(dev), max(lens_q), max(lens_k), causal=True, block_table=bt)
        assert torch.equal(again, out)


def test_varlen_many_sequences_mixed():
    """1500 sequences (more than the schedule kernel's thread count), mostly short: single-token and
    few-token sequences go to the decode kernel, the rest to the tcgen05 kernel, a handful of long ones
    dominate the cost ordering."""
    g = torch.Generator().manual_seed(5)
    lens_q = torch.randint(1, 41, (1500,), generator=g).tolist()
    lens_k = [x + int(y) for x, y in zip(lens_q, torch.randint(0, 200, (1500,), generator=g))]
    for i, s in ((7, 1100), (400, 700), (1499, 300)):
        lens_q[i] = s; lens_k[i] = s + 100
    info = _varlen_check(lens_q, lens_k, h=8, hk=2, seed=1)
    assert info["kernel_used"] == 1 and info["launches"] == 4


def test_fused_exchange_single_rank_protocol():
    """fa_b200_exchange_combine with world = 1 (plain allocation): the push / flag / parity protocol over
    several consecutive calls, and the merge result against the ordinary kvcache call."""
    import vllm_flash_attn as fa
    from vllm_flash_attn import sharded
    torch.manual_seed(0)
    dev, dt = torch.device("cuda", 0), torch.bfloat16
    b, h, hk, d, page, ctx = 3, 8, 2, 128, 16, 2048
    nblk = b * ctx // page
    kc = torch.randn(nblk, page, hk, d, device=dev, dtype=dt); vc = torch.randn_like(kc)
    bt = torch.randperm(nblk, device=dev).to(torch.int32).view(b, -1).contiguous()
    lens = torch.tensor([ctx, 777, 0], dtype=torch.int32, device=dev)     # one empty sequence: O = 0, LSE = +inf
    xch = sharded.PeerExchange(sharded.partial_record_numel(b, h, 1, d), dev)
    for it in range(5):                                                   # parity flips every call
        q = torch.randn(b, 1, h, d, device=dev, dtype=dt)
        ref, lref = fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=lens, block_table=bt, causal=True,
                                                return_softmax_lse=True)
        out, lse = sharded.flash_attn_with_kvcache_seq_sharded(q, kc, vc, lens, bt, shard_start=0, shard_end=ctx,
                                                                causal=True, return_softmax_lse=True, exchange=xch)
        assert (out.float() - ref.float()).abs().max().item() <= 4e-3, it
        fin = torch.isfinite(lref)
        assert (lse[fin] - lref[fin]).abs().max().item() < 1e-4 and bool((torch.isinf(lse) == torch.isinf(lref)).all())
        assert bool((out[2] == 0).all())


@pytest.mark.parametrize("d,h,hk", [(128, 8, 2), (64, 16, 2), (256, 8, 1)])
def test_one_kernel_sharded_decode_single_rank(d, h, hk):
    """fwd_kvcache(fused_exchange=...) with world = 1: the decode kernel merges its own splits (arrival ticket), runs the
    push / flag / wait protocol against itself and writes the final result — ONE launch.  Batch size and split count
    change from call to call through one PeerExchange (slots and call numbers are per (sequence, kv head) unit), an
    empty sequence gives O = 0 / LSE = +inf, and the result equals the ordinary kvcache call's."""
    import vllm_flash_attn as fa
    from vllm_flash_attn import sharded
    ext = _ext()
    torch.manual_seed(0)
    dev, dt, page, ctx = torch.device("cuda", 0), torch.bfloat16, 16, 4096
    xch = sharded.PeerExchange.for_decode(6, h, hk, d, dev)
    for it, (b, splits) in enumerate([(1, 0), (6, 3), (1, 7), (3, 1), (6, 0), (2, 16), (1, 0), (1, 1)]):
        nblk = b * ctx // page
        kc = torch.randn(nblk, page, hk, d, device=dev, dtype=dt); vc = torch.randn_like(kc)
        bt = torch.randperm(nblk, device=dev).to(torch.int32).view(b, -1).contiguous()
        lens = torch.randint(1, ctx + 1, (b,), device=dev).to(torch.int32)
        lens[0] = ctx
        if b >= 3:
            lens[2] = 0
        q = torch.randn(b, 1, h, d, device=dev, dtype=dt)
        ref, lref = fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=lens, block_table=bt, causal=True, num_splits=splits,
                                                return_softmax_lse=True)
        out, lse = sharded.flash_attn_with_kvcache_seq_sharded(q, kc, vc, lens, bt, shard_start=0, shard_end=ctx, causal=True,
                                                                num_splits=splits, return_softmax_lse=True, exchange=xch, one_kernel=True)
        assert ext.last_call_info["launches"] == 1 and ext.last_call_info["kernel_used"] == 3, (it, ext.last_call_info)
        assert not torch.isnan(out.float()).any()
        assert (out.float() - ref.float()).abs().max().item() <= 4e-3, it
        fin = torch.isfinite(lref)
        assert (lse[fin] - lref[fin]).abs().max().item() < 1e-4 and bool((torch.isinf(lse) == torch.isinf(lref)).all()), it
        if b >= 3:
            assert bool((out[2] == 0).all()) and bool(torch.isinf(lse[2]).all() and (lse[2] > 0).all())


def test_split_decode_merges_in_kernel():
    """Split-KV decode through the ordinary API is one launch now (the last CTA of each (sequence, kv head) merges the
    splits; the reference launches flash_fwd_splitkv_combine_kernel, flash_fwd_launch_template.h:131-155) and equals
    the merge of the same partials by the separate combine entry point."""
    import vllm_flash_attn as fa
    from vllm_flash_attn import sharded
    ext = _ext()
    torch.manual_seed(4)
    dev, dt = "cuda", torch.bfloat16
    q = torch.randn(2, 1, 32, 128, device=dev, dtype=dt)
    kc = torch.randn(2, 20000, 8, 128, device=dev, dtype=dt); vc = torch.randn_like(kc)
    cs = torch.tensor([19999, 5], dtype=torch.int32, device=dev)
    out, lse = fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=cs, return_softmax_lse=True)
    assert ext.last_call_info["launches"] == 1 and ext.last_call_info["num_splits"] > 1
    # the partial-result mode with the same split count, merged per split count 1 => the plain kernel result
    one, lone = fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=cs, num_splits=1, return_softmax_lse=True)
    assert (out.float() - one.float()).abs().max().item() <= 4e-3 and (lse - lone).abs().max().item() < 1e-4
    ref32, _ = orc.attention_ref(q, kc, vc, None, torch.arange(20000, device=dev).expand(2, 20000) < cs.view(-1, 1))
    pt, _ = orc.attention_ref(q, kc, vc, None, torch.arange(20000, device=dev).expand(2, 20000) < cs.view(-1, 1), upcast=False, reorder_ops=True)
    assert (out.float() - ref32.float()).abs().max().item() <= 3 * (pt.float() - ref32.float()).abs().max().item() + 1e-5


@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("paged,sq,splits,window,alibi", [
    (True, 1, 1, (-1, -1), False), (True, 1, 3, (-1, -1), True), (False, 3, 3, (-1, -1), False),
    (True, 3, 2, (3000, 0), False), (False, 1, 4, (-1, -1), True), (True, 4, 1, (-1, -1), False)])
def test_decode_one_wave_of_long_pieces_takes_the_12_warp_form(paged, sq, splits, window, alibi, dt):
    """d = 128, at most 16 packed rows, every piece resident at once and >= 4096 keys long: fa_api.cu sets wide_cta and
    the decode kernel runs 6 warps x 2 stages per CTA (profiles/r02_decode_wide_cta.txt).  Same contract as every kvcache
    call: ragged lengths (one of 16 keys: empty pieces), masks, ALiBi, both cache layouts, in-kernel split merge."""
    import vllm_flash_attn as fa
    torch.manual_seed(11)
    dev, b, h, hk, d, page, smax = "cuda", 3, 32, 8, 128, 16, 17008
    lens = torch.tensor([17000, 4097, 16], dtype=torch.int32, device=dev)
    q = torch.randn(b, sq, h, d, device=dev, dtype=dt)
    k = torch.randn(b, smax, hk, d, device=dev, dtype=dt); v = torch.randn_like(k)
    slopes = (torch.rand(b, h, device=dev) * 0.3) if alibi else None
    if paged:
        nblk = b * smax // page
        bt = torch.randperm(nblk, device=dev).to(torch.int32).view(b, -1).contiguous()
        kc = torch.empty(nblk, page, hk, d, device=dev, dtype=dt); vc = torch.empty_like(kc)
        kc[bt.long().flatten()] = k.view(b * smax // page, page, hk, d); vc[bt.long().flatten()] = v.view(b * smax // page, page, hk, d)
        out, lse = fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=lens, block_table=bt, causal=True, window_size=window,
                                              alibi_slopes=slopes, num_splits=splits, return_softmax_lse=True)
    else:
        out, lse = fa.flash_attn_with_kvcache(q, k, v, cache_seqlens=lens, causal=True, window_size=window,
                                              alibi_slopes=slopes, num_splits=splits, return_softmax_lse=True)
    kmask = torch.arange(smax, device=dev).expand(b, smax) < lens.view(-1, 1)
    bias = orc.attn_bias_from_alibi_slopes(slopes, sq, smax, None, kmask, causal=True) if alibi else None
    ref32, _ = orc.attention_ref(q, k, v, None, kmask, bias, causal=True, window_size=window)
    pt, _ = orc.attention_ref(q, k, v, None, kmask, bias, causal=True, window_size=window, upcast=False, reorder_ops=True)
    err, err_pt = (out.float() - ref32.float()).abs().max().item(), (pt.float() - ref32.float()).abs().max().item()
    assert not torch.isnan(out.float()).any() and torch.isfinite(lse).all()
    assert err <= 3 * err_pt + 1e-5, (err, err_pt)


@pytest.mark.parametrize("paged,alibi,sq", [(True, False, 1), (True, True, 2), (False, False, 1)])
def test_decode_many_waves_takes_the_three_cta_per_sm_form(paged, alibi, sq):
    """d = 128, >= 10 units per SM, one split: 4 warps x 2 stages, three CTAs per SM (fa_api.cu wide_cta = 2;
    profiles/r02_decode_wide_cta.txt).  1536 units of ragged short contexts (some empty, some of one key)."""
    import vllm_flash_attn as fa
    torch.manual_seed(12)
    dev, dt, b, h, hk, d, page, smax = "cuda", torch.bfloat16, 192, 16, 8, 128, 16, 320
    lens = torch.randint(0, smax + 1, (b,), device=dev).to(torch.int32)
    lens[0], lens[1], lens[2] = 0, 1, smax
    if sq > 1: lens = lens.clamp(min=sq)          # a causal row needs its own key
    q = torch.randn(b, sq, h, d, device=dev, dtype=dt)
    k = torch.randn(b, smax, hk, d, device=dev, dtype=dt); v = torch.randn_like(k)
    slopes = (torch.rand(b, h, device=dev) * 0.3) if alibi else None
    if paged:
        nblk = b * smax // page
        bt = torch.randperm(nblk, device=dev).to(torch.int32).view(b, -1).contiguous()
        kc = torch.empty(nblk, page, hk, d, device=dev, dtype=dt); vc = torch.empty_like(kc)
        kc[bt.long().flatten()] = k.view(nblk, page, hk, d); vc[bt.long().flatten()] = v.view(nblk, page, hk, d)
        out = fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=lens, block_table=bt, causal=True, alibi_slopes=slopes)
    else:
        out = fa.flash_attn_with_kvcache(q, k, v, cache_seqlens=lens, causal=True, alibi_slopes=slopes)
    info = dict(_ext().last_call_info)
    assert info["kernel_used"] == 3 and info["num_splits"] == 1, info
    kmask = torch.arange(smax, device=dev).expand(b, smax) < lens.view(-1, 1)
    bias = orc.attn_bias_from_alibi_slopes(slopes, sq, smax, None, kmask, causal=True) if alibi else None
    ref32, _ = orc.attention_ref(q, k, v, None, kmask, bias, causal=True)
    pt, _ = orc.attention_ref(q, k, v, None, kmask, bias, causal=True, upcast=False, reorder_ops=True)
    live = (lens > 0).view(-1, 1, 1, 1)            # a sequence without keys: the reference leaves zeros, the oracle NaN
    err = ((out.float() - ref32.float()) * live).nan_to_num(0).abs().max().item()
    err_pt = ((pt.float() - ref32.float()) * live).nan_to_num(0).abs().max().item()
    assert not torch.isnan(out.float()).any() and (sq > 1 or (out[0] == 0).all())
    assert err <= 3 * err_pt + 1e-5, (err, err_pt)


def test_varlen_seqused_k():
    """varlen_fwd(seqused_k=...) (flash_api.cpp:428,:553-556; BlockInfo src/block_info.h:24): keys beyond
    seqused_k[b] are ignored although cu_seqlens_k spans them.  Prefill rows run on the tcgen05 kernel,
    the single-token sequence on the decode kernel."""
    ext = _ext()
    torch.manual_seed(3)
    dev, dt, h, hk, d = "cuda", torch.bfloat16, 8, 2, 128
    lens_q, lens_k, used = [300, 1, 200], [500, 400, 350], [420, 333, 200]
    cu_q = torch.tensor([0] + lens_q).cumsum(0).to(torch.int32); cu_k = torch.tensor([0] + lens_k).cumsum(0).to(torch.int32)
    q = torch.randn(sum(lens_q), h, d, device=dev, dtype=dt)
    k = torch.randn(sum(lens_k), hk, d, device=dev, dtype=dt); v = torch.randn_like(k)
    su = torch.tensor(used, dtype=torch.int32, device=dev)
    kp, vp = k.clone(), v.clone()
    for i in range(3):          # rows past seqused_k hold NaN: they must never reach the output (0 * NaN in P.V)
        kp[int(cu_k[i]) + used[i]:int(cu_k[i + 1])] = float("nan"); vp[int(cu_k[i]) + used[i]:int(cu_k[i + 1])] = float("nan")
    res = ext.varlen_fwd(q, kp, vp, None, cu_q.to(dev), cu_k.to(dev), su, None, None, max(lens_q), max(lens_k), 0.0,
                         d ** -0.5, False, True, -1, -1, False, None)
    out = res[0]
    assert not torch.isnan(out.float()).any()
    assert ext.last_call_info["kernel_used"] == 1 and ext.last_call_info["launches"] == 3     # prefill + decode (16- and 32-row forms)
    # oracle: the same batch with k, v physically cut to the used lengths
    kc = torch.cat([k[int(cu_k[i]):int(cu_k[i]) + used[i]] for i in range(3)])
    vc = torch.cat([v[int(cu_k[i]):int(cu_k[i]) + used[i]] for i in range(3)])
    cu_u = torch.tensor([0] + used).cumsum(0).to(torch.int32)
    ref32, _ = orc.attention_ref_varlen(q, kc, vc, cu_q, cu_u, causal=True)
    pt, _ = orc.attention_ref_varlen(q, kc, vc, cu_q, cu_u, causal=True, upcast=False, reorder_ops=True)
    err, err_pt = (out.float() - ref32.float()).abs().max().item(), (pt.float() - ref32.float()).abs().max().item()
    assert err <= 2 * err_pt + 1e-5, (err, err_pt)


@pytest.mark.parametrize("d", [64, 128, 256])
@pytest.mark.parametrize("causal", [True, False])
def test_varlen_neighbour_rows_do_not_leak_into_pv(d, causal):
    """Packed varlen K/V: the 128-key V tile that holds the end of a sequence also holds the NEXT sequence's rows.
    Their P is 0, but 0 * NaN = NaN, so one request's non-finite V would poison its neighbour's output unless the
    kernel keeps those rows out of P.V (the reference zero-fills them: Clear_OOB_MN=true, flash_fwd_kernel.h:292).
    Sequences 1 and 3 have no query rows and all-NaN K/V; every other sequence ends inside a tile they share."""
    import vllm_flash_attn as fa
    torch.manual_seed(11)
    dev, dt, h, hk = "cuda", torch.bfloat16, 4, 2
    lens_q = [200, 0, 333, 0, 129]
    lens_k = [200, 100, 333, 300, 129]
    cu_q = torch.tensor([0] + lens_q).cumsum(0).to(torch.int32); cu_k = torch.tensor([0] + lens_k).cumsum(0).to(torch.int32)
    q = torch.randn(sum(lens_q), h, d, device=dev, dtype=dt)
    k = torch.randn(sum(lens_k), hk, d, device=dev, dtype=dt); v = torch.randn_like(k)
    kp, vp = k.clone(), v.clone()
    for i in (1, 3):
        kp[int(cu_k[i]):int(cu_k[i + 1])] = float("nan"); vp[int(cu_k[i]):int(cu_k[i + 1])] = float("nan")
    out = fa.flash_attn_varlen_func(q, kp, vp, cu_q.to(dev), cu_k.to(dev), max(lens_q), max(lens_k), causal=causal)
    assert _ext().last_call_info["kernel_used"] == 1
    assert not torch.isnan(out.float()).any()
    clean = fa.flash_attn_varlen_func(q, k, v, cu_q.to(dev), cu_k.to(dev), max(lens_q), max(lens_k), causal=causal)
    assert torch.equal(out, clean)          # the neighbours' contents make no difference at all
    ref32, _ = orc.attention_ref_varlen(q, k, v, cu_q, cu_k, causal=causal)
    pt, _ = orc.attention_ref_varlen(q, k, v, cu_q, cu_k, causal=causal, upcast=False, reorder_ops=True)
    err, err_pt = (out.float() - ref32.float()).abs().max().item(), (pt.float() - ref32.float()).abs().max().item()
    assert err <= 2 * err_pt + 1e-5, (err, err_pt)


@pytest.mark.parametrize("d,use_idx,s_cache", [(128, False, 700), (128, True, 300), (64, True, 100), (256, False, 513), (96, True, 40)])
def test_kvcache_prefill_over_a_non_paged_cache(d, use_idx, s_cache):
    """flash_attn_with_kvcache with many query rows against a NON-paged cache addressed through cache_seqlens and
    cache_batch_idx (tests/test_flash_attn.py:1816-2076): runs on the tcgen05 kernel with the cache seen as one page per
    sequence.  Rows past cache_seqlens hold NaN: they must never reach the output."""
    import vllm_flash_attn as fa
    torch.manual_seed(2)
    dev, dt, b, h, hk = "cuda", torch.bfloat16, 3, 8, 2
    b_c = 7 if use_idx else b
    lens = torch.tensor([s_cache, max(s_cache // 2, 1), max(s_cache - 17, 1)], dtype=torch.int32, device=dev)
    sq = min(60, s_cache // 2) if s_cache < 128 else 200
    lens = torch.maximum(lens, torch.tensor(sq, dtype=torch.int32, device=dev))
    kc = torch.randn(b_c, s_cache, hk, d, device=dev, dtype=dt); vc = torch.randn_like(kc)
    idx = torch.tensor([5, 0, 3], dtype=torch.int32, device=dev) if use_idx else None
    sel = idx.long() if use_idx else torch.arange(b, device=dev)
    dead = torch.arange(s_cache, device=dev).view(1, -1) >= lens.view(-1, 1).long()          # (b, s_cache)
    kc[sel] = torch.where(dead.view(b, s_cache, 1, 1), torch.full_like(kc[sel], float("nan")), kc[sel])
    vc[sel] = torch.where(dead.view(b, s_cache, 1, 1), torch.full_like(vc[sel], float("nan")), vc[sel])
    q = torch.randn(b, sq, h, d, device=dev, dtype=dt)
    out, lse = fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=lens, cache_batch_idx=idx, causal=True, return_softmax_lse=True)
    assert _ext().last_call_info["kernel_used"] == 1
    kpm = ~dead
    kg, vg = torch.nan_to_num(kc[sel]), torch.nan_to_num(vc[sel])
    ref32, _ = orc.attention_ref(q, kg, vg, None, kpm, causal=True)
    pt, _ = orc.attention_ref(q, kg, vg, None, kpm, causal=True, upcast=False, reorder_ops=True)
    assert not torch.isnan(out.float()).any()
    err, err_pt = (out.float() - ref32.float()).abs().max().item(), (pt.float() - ref32.float()).abs().max().item()
    assert err <= 3 * err_pt + 1e-5, (err, err_pt)
    lref = orc.lse_ref(q, kg, kpm, None, None, causal=True)
    fin = torch.isfinite(lref)
    assert (lse[fin] - lref[fin]).abs().max().item() < 2e-3
