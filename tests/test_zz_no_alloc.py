"""GPU: fa_b200_fwd_kvcache_static_scratch (the allocation-free variant of fwd_kvcache whose split-KV
scratch is a static device array inside libfa_b200.so) against the existing fa_b200_fwd_kvcache entry
point, both called straight through the C ABI (ctypes), on two calls in a row with different shapes so
that the second call finds the first call's contents in the shared scratch.  Collected last on purpose
(file name).  No counterpart in the reference's test-suite."""
import ctypes as C

import pytest
import torch

import _util  # noqa: F401  (sys.path)
from vllm_flash_attn import _cabi
from vllm_flash_attn._cabi import FaFwdParams, lib

pytestmark = pytest.mark.gpu
TOL = 2e-2          # the max-abs bound tests/test_api_gpu.py uses between two kernels' bf16 outputs


def _params(q, kc, vc, bt, lens, out, lse, num_splits):
    b, sq, h, d = q.shape
    page, hk = kc.shape[1], kc.shape[2]
    P = FaFwdParams()
    P.q, P.k, P.v, P.o, P.softmax_lse = q.data_ptr(), kc.data_ptr(), vc.data_ptr(), out.data_ptr(), lse.data_ptr()
    P.q_batch_stride, P.q_row_stride, P.q_head_stride = q.stride(0), q.stride(1), q.stride(2)
    P.k_batch_stride, P.k_row_stride, P.k_head_stride = kc.stride(0), kc.stride(1), kc.stride(2)
    P.v_batch_stride, P.v_row_stride, P.v_head_stride = vc.stride(0), vc.stride(1), vc.stride(2)
    P.o_batch_stride, P.o_row_stride, P.o_head_stride = out.stride(0), out.stride(1), out.stride(2)
    P.b, P.seqlen_q, P.seqlen_k, P.h, P.h_k, P.d = b, sq, bt.shape[1] * page, h, hk, d
    P.is_bf16, P.softmax_scale = 1, d ** -0.5
    P.is_causal, P.window_left, P.window_right = 1, -1, -1
    P.cu_seqlens_k, P.is_seqlens_k_cumulative = lens.data_ptr(), 0
    P.block_table, P.block_table_batch_stride = bt.data_ptr(), bt.stride(0)
    P.page_block_size, P.max_blocks_per_seq, P.num_blocks = page, bt.shape[1], kc.shape[0]
    P.num_splits = num_splits
    return P


def _case(seed, b, ctx, h, hk, d, num_splits, page=16):
    g = torch.Generator(device="cuda").manual_seed(seed)
    dt = torch.bfloat16
    blocks = ctx // page
    nblk = b * blocks + 7
    kc = torch.randn(nblk, page, hk, d, device="cuda", dtype=dt, generator=g)
    vc = torch.randn(nblk, page, hk, d, device="cuda", dtype=dt, generator=g)
    bt = torch.randperm(nblk, device="cuda", generator=g)[: b * blocks].to(torch.int32).view(b, blocks).contiguous()
    q = torch.randn(b, 1, h, d, device="cuda", dtype=dt, generator=g)
    lens = torch.randint(ctx // 2, ctx + 1, (b,), device="cuda", generator=g).to(torch.int32)
    stream = torch.cuda.current_stream().cuda_stream

    # existing entry point: caller-allocated workspace of the size the library reports
    out_a, lse_a = torch.empty_like(q), torch.empty(b, h, 1, device="cuda", dtype=torch.float32)
    Pa = _params(q, kc, vc, bt, lens, out_a, lse_a, num_splits)
    need = lib.fa_b200_workspace_bytes(C.byref(Pa), _cabi.OP_KVCACHE)
    assert need != C.c_size_t(-1).value and 0 < need <= _cabi.STATIC_SCRATCH_BYTES
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    Pa.workspace, Pa.workspace_bytes = ws.data_ptr(), need
    _cabi.check(lib.fa_b200_fwd_kvcache(C.byref(Pa), stream))
    assert Pa.num_splits > 1, "the case must exercise the split-KV scratch"

    # new entry point: no workspace at all.  Outputs pre-filled with NaN: every element must be written.
    out_b = torch.full_like(q, float("nan"))
    lse_b = torch.full((b, h, 1), float("nan"), device="cuda", dtype=torch.float32)
    Pb = _params(q, kc, vc, bt, lens, out_b, lse_b, num_splits)
    _cabi.check(lib.fa_b200_fwd_kvcache_static_scratch(C.byref(Pb), stream))
    torch.cuda.synchronize()
    assert Pb.num_splits == Pa.num_splits and Pb.kernel_used == Pa.kernel_used
    assert Pb.workspace is None and Pb.workspace_bytes == 0      # the caller's block comes back as passed
    assert not torch.isnan(out_b.float()).any() and not torch.isnan(lse_b).any()
    assert (out_b.float() - out_a.float()).abs().max().item() <= TOL
    assert (lse_b - lse_a).abs().max().item() <= 1e-3
    # same kernels, same split count, fixed summation order: in fact the results are bit-identical
    assert torch.equal(out_b, out_a) and torch.equal(lse_b, lse_a)


def test_static_scratch_matches_workspace_entry_point_two_calls_in_a_row():
    torch.cuda.set_device(0)
    # call 1: long context, heuristic split count (fills a large part of the scratch)
    _case(seed=1, b=2, ctx=16384, h=32, hk=8, d=128, num_splits=0)
    # call 2: other shape and a forced split count — it finds call 1's partials in the shared scratch
    _case(seed=2, b=3, ctx=2048, h=16, hk=2, d=64, num_splits=5)


def test_static_scratch_rejects_oversize_and_caller_workspace():
    torch.cuda.set_device(0)
    dt = torch.bfloat16
    b, h, hk, d, page, blocks = 64, 32, 8, 128, 16, 8
    kc = torch.zeros(b * blocks, page, hk, d, device="cuda", dtype=dt)
    bt = torch.arange(b * blocks, device="cuda", dtype=torch.int32).view(b, blocks)
    q = torch.zeros(b, 1, h, d, device="cuda", dtype=dt)
    lens = torch.full((b,), page * blocks, device="cuda", dtype=torch.int32)
    out = torch.full_like(q, 7.0)
    lse = torch.zeros(b, h, 1, device="cuda", dtype=torch.float32)
    stream = torch.cuda.current_stream().cuda_stream
    P = _params(q, kc, kc, bt, lens, out, lse, num_splits=8)     # 8 * 64 * 32 * 129 * 4 B = 8.5 MB > 4 MiB
    rc = lib.fa_b200_fwd_kvcache_static_scratch(C.byref(P), stream)
    assert rc == _cabi.FA_ERR_WORKSPACE
    msg = _cabi.last_error()
    assert "static maximum" in msg and str(_cabi.STATIC_SCRATCH_BYTES) in msg
    torch.cuda.synchronize()
    assert torch.all(out == 7.0)                                  # nothing was launched
    P = _params(q, kc, kc, bt, lens, out, lse, num_splits=1)
    P.workspace, P.workspace_bytes = kc.data_ptr(), 1024
    assert lib.fa_b200_fwd_kvcache_static_scratch(C.byref(P), stream) != 0
    assert "takes no workspace" in _cabi.last_error()
