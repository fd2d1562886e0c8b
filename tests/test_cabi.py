"""CPU: the C-ABI library loads, exports every symbol include/fa_b200.h declares, its parameter
block matches the ctypes mirror, and the host-only logic (split heuristic, validation) behaves
like the reference's glue.  No kernel is launched here."""
import ctypes as C
import os
import re

import pytest

from vllm_flash_attn import _cabi
from vllm_flash_attn._cabi import FaFwdParams, lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "fa_b200.h")).read()
    declared = set(re.findall(r"\b(fa_b200_\w+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in fa_b200.h but not exported"
    assert declared == set(_cabi.EXPORTS)


def test_params_layout_and_version():
    assert lib.fa_b200_abi_version() == 5
    assert lib.fa_b200_params_size() == C.sizeof(FaFwdParams)


def ref_heuristic(work, sms, nblk, max_splits):
    """Independent restatement of num_splits_heuristic (flash_api.cpp:158-192)."""
    import math
    if work >= 0.8 * sms:
        return 1
    max_splits = min(max_splits, sms, nblk)
    cdiv = lambda a, b: (a + b - 1) // b
    elig = lambda s: s == 1 or cdiv(nblk, s) != cdiv(nblk, s - 1)
    eff = {}
    for s in range(1, max_splits + 1):
        if elig(s):
            w = work * s / sms
            eff[s] = w / math.ceil(w)
    best = max(eff.values()) if eff else 0
    for s in range(1, max_splits + 1):
        if elig(s) and eff[s] >= 0.85 * best:
            return s
    return 1


@pytest.mark.parametrize("work,sms,nblk", [(2048, 296, 32), (8, 296, 1024), (128, 296, 2), (48, 216, 64),
                                            (1, 296, 1), (1, 296, 500), (100, 264, 7), (3, 296, 129)])
def test_split_heuristic_matches_reference_rule(work, sms, nblk):
    assert lib.fa_b200_num_splits_heuristic(work, sms, nblk, 128) == ref_heuristic(work, sms, nblk, 128)


def test_survey_split_examples():
    # C3 (b=256,h_k=8) -> 1 split; C5 (b=1,h_k=8, 1024 n-blocks, 148 SMs) -> 32 splits
    assert lib.fa_b200_num_splits_heuristic(256 * 8, 296, 32, 128) == 1
    assert lib.fa_b200_num_splits_heuristic(8, 296, 1024, 128) == 32


def _base():
    P = FaFwdParams()
    P.q = P.k = P.v = P.o = 0x1000
    P.b, P.seqlen_q, P.seqlen_k, P.h, P.h_k, P.d = 2, 1, 512, 8, 2, 128
    P.q_row_stride = P.k_row_stride = P.v_row_stride = P.o_row_stride = 1024
    P.q_head_stride = P.k_head_stride = P.v_head_stride = P.o_head_stride = 128
    P.is_bf16, P.softmax_scale, P.num_sms = 1, 0.1, 148
    return P


@pytest.mark.parametrize("mut,msg", [
    (lambda P: setattr(P, "q", None), "non-NULL"),
    (lambda P: setattr(P, "d", 264), "at most 256"),
    (lambda P: setattr(P, "h_k", 3), "must divide"),
    (lambda P: setattr(P, "d", 20), "multiple of 8"),
    (lambda P: setattr(P, "b", 0), "batch size"),
])
def test_validation_errors(mut, msg):
    P = _base()
    mut(P)
    assert lib.fa_b200_workspace_bytes(C.byref(P), 2) == C.c_size_t(-1).value
    assert msg in _cabi.last_error()


def test_paged_block_size_rule():
    P = _base()
    P.block_table, P.page_block_size, P.max_blocks_per_seq = 0x2000, 24, 4
    assert lib.fa_b200_workspace_bytes(C.byref(P), 2) == C.c_size_t(-1).value
    assert "divisible by 16" in _cabi.last_error()          # flash_api.cpp:491,735
    P.page_block_size = 16
    assert lib.fa_b200_workspace_bytes(C.byref(P), 2) != C.c_size_t(-1).value


@pytest.mark.parametrize("batch,ctx,want", [
    (24, 1024, 2),      # 192 units: 2 pieces each fill one wave of 3 CTAs per SM (the wave rule alone said 4 splits on 2 per SM: 38.5 us; now 24.6)
    (24, 4096, 2),
    (20, 4096, 2),
    (48, 1024, 1),      # 384 units: one wave of the three-per-SM form, no split (60 -> 45 us)
    (48, 4096, 1),
    (40, 4096, 1),
    (64, 4096, 1), (256, 4096, 1),
])
def test_decode_split_choice_follows_the_time_model(batch, ctx, want):
    """fa_api.cu decode_choose(): grid form and split count of the d = 128 decode kernel from the time model fitted to
    profiles/r02_decode_form_sweep.jsonl, read back through the workspace size (partials of `s` pieces per unit)."""
    if os.environ.get("FA_B200_STREAMK", "") not in ("", "0"):
        pytest.skip("balanced ranges forced on: no split count")
    P = _base()
    P.b, P.h, P.h_k, P.seqlen_k, P.num_splits = batch, 32, 8, ctx, 0
    P.q_row_stride, P.k_row_stride = 32 * 128, 8 * 128
    P.v_row_stride, P.o_row_stride = 8 * 128, 32 * 128
    P.block_table, P.page_block_size, P.max_blocks_per_seq = 0x2000, 16, ctx // 16
    need = lib.fa_b200_workspace_bytes(C.byref(P), 2)
    assert need != C.c_size_t(-1).value, _cabi.last_error()
    got = 1 if need == 0 else (need - 16384 - 16) // (batch * 32 * (128 + 1) * 4)
    assert got == want, (got, want, need)


@pytest.mark.parametrize("batch,ctx,want", [(24, 1024, 1), (24, 4096, 3), (48, 4096, 2), (48, 1024, 1), (64, 4096, 1)])
def test_decode_split_choice_other_head_dims_keep_the_corrected_wave_rule(batch, ctx, want):
    """d != 128 has one grid form: decode_splits() — the reference's wave rule with the two corrections of
    profiles/r02_decode_split_sweep.jsonl (split at 1.0-1.5 waves when pieces stay long; short pieces stay in one wave)."""
    if os.environ.get("FA_B200_STREAMK", "") not in ("", "0"):
        pytest.skip("balanced ranges forced on: no split count")
    P = _base()
    P.b, P.h, P.h_k, P.seqlen_k, P.num_splits, P.d = batch, 32, 8, ctx, 0, 64
    P.q_row_stride, P.k_row_stride = 32 * 64, 8 * 64
    P.v_row_stride, P.o_row_stride = 8 * 64, 32 * 64
    P.q_head_stride = P.k_head_stride = P.v_head_stride = P.o_head_stride = 64
    P.block_table, P.page_block_size, P.max_blocks_per_seq = 0x2000, 16, ctx // 16
    need = lib.fa_b200_workspace_bytes(C.byref(P), 2)
    assert need != C.c_size_t(-1).value, _cabi.last_error()
    got = 1 if need == 0 else (need - 16384 - 16) // (batch * 32 * (64 + 1) * 4)
    assert got == want, (got, want, need)


def test_workspace_matches_split_choice():
    P = _base()                         # b*h_k = 4 CTAs on 148 SMs, 512 keys -> splits
    P.num_splits = 0
    need = lib.fa_b200_workspace_bytes(C.byref(P), 2)
    s = lib.fa_b200_num_splits_heuristic(4, 296, 2, 128)
    if os.environ.get("FA_B200_STREAMK", "") not in ("", "0"):     # balanced ranges forced on: a pool of piece slots instead
        assert need == 16384 + (2 * 148 + 4) * 16 * (128 + 1) * 4 + 16
    else:
        assert s == 2 and need == 16384 + s * 2 * 8 * 1 * (128 + 1) * 4 + 16     # counter head + partials + slack
    P.num_splits = 1
    assert lib.fa_b200_workspace_bytes(C.byref(P), 2) == 0
    P.num_splits = 129
    assert lib.fa_b200_workspace_bytes(C.byref(P), 2) == C.c_size_t(-1).value
    assert "128" in _cabi.last_error()


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "vllm-flash-attention_b200")
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".inc")):
                src = open(os.path.join(dp, fn)).read()
                assert "attention_oracle" not in src and "oracle/" not in src, fn
