"""GPU: the library is re-entrant across host threads and CUDA streams, as the reference's extension is (plain functions
on the caller's current stream, no global mutable state: flash_api.cpp:344,399).  Four host threads, each on its own
stream, hammer a different call type — a mixed paged varlen batch (work list + side-stream decode half), a split decode
(arrival counters in the per-thread workspace), a dense causal prefill, a varlen prefill — and every result must be
bit-identical to the one the same call gave single-threaded; each thread's last_call_info must describe ITS call.
Run through both host bindings (compiled pybind11 module and ctypes shim)."""
import threading

import pytest
import torch

import _util  # noqa: F401  (sys.path)

pytestmark = pytest.mark.gpu


def _make_calls():
    import vllm_flash_attn as fa
    dev, dt = "cuda", torch.bfloat16
    g = torch.Generator(device="cuda").manual_seed(7)
    R = lambda *s: torch.randn(*s, device=dev, dtype=dt, generator=g)
    page = 16
    # 1. mixed paged varlen batch: 2 prompts + 6 decodes, 8q/1kv (split decode half), > 148 units (work list)
    lens_q, lens_k = [1200, 1, 1, 900, 1, 2, 1, 1], [1200, 3000, 777, 900, 4096, 1500, 33, 2048]
    cu_q = torch.tensor([0] + list(torch.tensor(lens_q).cumsum(0)), dtype=torch.int32, device=dev)
    cu_k = torch.tensor([0] + list(torch.tensor(lens_k).cumsum(0)), dtype=torch.int32, device=dev)
    mb = (max(lens_k) + page - 1) // page
    q1, kc1, vc1 = R(sum(lens_q), 32, 128), R(len(lens_q) * mb, page, 4, 128), R(len(lens_q) * mb, page, 4, 128)
    bt1 = torch.randperm(len(lens_q) * mb, device=dev).to(torch.int32).view(len(lens_q), mb).contiguous()
    mixed = lambda: fa.flash_attn_varlen_func(q1, kc1, vc1, cu_q, cu_k, max(lens_q), max(lens_k), causal=True, block_table=bt1)
    # 2. split decode: 2 sequences x 8192 keys, auto splits
    q2, kc2, vc2 = R(2, 1, 32, 128), R(2 * 512, page, 8, 128), R(2 * 512, page, 8, 128)
    bt2 = torch.randperm(2 * 512, device=dev).to(torch.int32).view(2, 512).contiguous()
    cl2 = torch.tensor([8192, 5000], dtype=torch.int32, device=dev)
    decode = lambda: fa.flash_attn_with_kvcache(q2, kc2, vc2, cache_seqlens=cl2, block_table=bt2, causal=True)
    # 3. dense causal prefill, GQA (head-pair units)
    q3, k3, v3 = R(2, 1024, 16, 128), R(2, 1024, 4, 128), R(2, 1024, 4, 128)
    dense = lambda: fa.flash_attn_func(q3, k3, v3, causal=True)
    # 4. packed varlen prefill, d = 64, MHA (row-pair units), non-causal
    lens4 = [700, 130, 512, 64]
    cu4 = torch.tensor([0] + list(torch.tensor(lens4).cumsum(0)), dtype=torch.int32, device=dev)
    q4, k4, v4 = R(sum(lens4), 8, 64), R(sum(lens4), 8, 64), R(sum(lens4), 8, 64)
    varlen = lambda: fa.flash_attn_varlen_func(q4, k4, v4, cu4, cu4, max(lens4), max(lens4), causal=False)
    return [("mixed", mixed, 1), ("decode", decode, 3), ("dense", dense, 1), ("varlen", varlen, 1)]


@pytest.mark.parametrize("binding", ["pybind11", "ctypes"])
def test_four_threads_four_streams_bit_identical(binding):
    from vllm_flash_attn import vllm_flash_attn_2_cuda as ext
    if binding == "pybind11" and ext._NATIVE is None:
        pytest.skip("compiled host binding not built")
    saved = ext._N
    ext._N = ext._NATIVE if binding == "pybind11" else None
    try:
        calls = _make_calls()
        expect = []
        for name, f, kern in calls:
            o = f().clone()
            info = dict(ext.last_call_info)
            assert info["kernel_used"] == kern, (name, info)
            expect.append((o, info))
        torch.cuda.synchronize()
        errors = []
        barrier = threading.Barrier(len(calls))

        def worker(i):
            name, f, kern = calls[i]
            try:
                s = torch.cuda.Stream()
                barrier.wait()
                with torch.cuda.stream(s):
                    for it in range(40):
                        o = f()
                        info = dict(ext.last_call_info)
                        if info != expect[i][1]:
                            errors.append((name, it, "info", info, expect[i][1]))
                            return
                        if it % 8 == 7:
                            s.synchronize()
                            if not torch.equal(o, expect[i][0]):
                                errors.append((name, it, "output differs", (o.float() - expect[i][0].float()).abs().max().item()))
                                return
                s.synchronize()
            except Exception as e:  # noqa: BLE001
                errors.append((name, "exception", repr(e)))

        threads = [threading.Thread(target=worker, args=(i,)) for i in range(len(calls))]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        torch.cuda.synchronize()
        assert not errors, errors
    finally:
        ext._N = saved
