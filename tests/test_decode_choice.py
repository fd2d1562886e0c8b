"""CPU: the split count the library plans for a d = 128 decode equals an independent restatement of its time model
(fa_api.cu decode_choose(): waves of resident CTAs, per-CTA stream rate by CTAs per SM, fixed cost per wave, merge cost)
over a grid of batch sizes and cache capacities — read back through fa_b200_workspace_bytes, no GPU needed."""
import ctypes as C
import math
import os

import numpy as np
import pytest

import _util  # noqa: F401  (sys.path)
from vllm_flash_attn import _cabi
from vllm_flash_attn._cabi import FaFwdParams, lib

SMS = 148


def model_choice(units, seqlen_k, sms=SMS):
    rate = ((34.0, 27.0, 27.0), (34.0, 27.0, 20.0))
    hbm, fix, base, merge, per_split = 6900.0, (3.0, 3.5), 6.0, 2.0, 0.15
    f32 = np.float32                                   # the library computes in float: same rounding, same ties
    best, best_t = (0, 1), f32(1e30)
    for s in range(1, min(128, max(1, (seqlen_k + 255) // 256)) + 1):
        for f in (0, 1):
            slots, pieces = (3 if f else 2) * sms, units * s
            piece_bytes = f32(f32(seqlen_k) * f32(512.0)) / f32(s)

            def wave(n):
                c = min((n + sms - 1) // sms, 3 if f else 2)
                r = min(f32(rate[f][c - 1]), f32(hbm) / f32(n))
                return piece_bytes / r * f32(1e-3) + f32(fix[f])
            t = f32(base) + (f32(merge) + f32(per_split) * f32(s) if s > 1 else f32(0))
            if pieces // slots:
                t = t + f32(pieces // slots) * wave(slots)
            if pieces % slots:
                t = t + wave(pieces % slots)
            if t < best_t - f32(1e-4):
                best_t, best = t, (f, s)
    return best


@pytest.mark.parametrize("h_k", [1, 8])
def test_planned_split_count_equals_the_model(h_k):
    if os.environ.get("FA_B200_STREAMK", "") not in ("", "0"):
        pytest.skip("balanced ranges forced on: no split count")
    g, d = 4, 128
    for batch in (1, 2, 4, 8, 12, 16, 20, 24, 32, 40, 48, 56, 64, 96, 128, 192, 256, 384):
        for cap in (256, 1024, 2560, 4096, 16384, 65536, 131072):
            P = FaFwdParams()
            P.q = P.k = P.v = P.o = 0x1000
            P.b, P.seqlen_q, P.seqlen_k, P.h, P.h_k, P.d = batch, 1, cap, h_k * g, h_k, d
            P.q_row_stride = P.o_row_stride = h_k * g * d
            P.k_row_stride = P.v_row_stride = h_k * d
            P.q_head_stride = P.k_head_stride = P.v_head_stride = P.o_head_stride = d
            P.is_bf16, P.softmax_scale, P.num_sms, P.num_splits = 1, 0.1, SMS, 0
            P.block_table, P.page_block_size, P.max_blocks_per_seq = 0x2000, 16, cap // 16
            need = lib.fa_b200_workspace_bytes(C.byref(P), 2)
            assert need != C.c_size_t(-1).value, _cabi.last_error()
            units = batch * h_k
            counters_behind = 0 if units * 4 <= 16384 else units * 4
            got = 1 if need == 0 else (need - 16384 - 16 - counters_behind) / (batch * h_k * g * (d + 1) * 4)
            want = model_choice(units, cap)[1]
            assert abs(got - want) < 0.01, (batch, h_k, cap, got, want, need)
