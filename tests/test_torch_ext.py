"""The compiled pybind11 / ATen host binding (csrc/fa_torch_ext.cpp -> vllm_flash_attn/_fa_torch*.so): the form the
reference's extension has (PYBIND11_MODULE, flash_api.cpp:939-946), over the same C ABI as the ctypes shim.
CPU: it loads, exports the three ops with the reference's arities, agrees with the library's ABI and raises
RuntimeError like TORCH_CHECK does.  GPU: both host bindings give bit-identical results on every op (same library
call underneath), including the append / rotary path, the out= argument, head-dim padding and CUDA-graph capture."""
import inspect
import math
import os
import re

import pytest
import torch

import _util  # noqa: F401  (sys.path)

ext = pytest.importorskip("vllm_flash_attn.vllm_flash_attn_2_cuda")
N = pytest.importorskip("vllm_flash_attn._fa_torch", reason="build with vllm-flash-attention_b200/build_torch_ext.sh")


def test_module_surface_and_abi():
    from vllm_flash_attn import _cabi
    assert N.abi_version() == _cabi.ABI_VERSION
    assert ext.host_binding == ("ctypes" if os.environ.get("FA_B200_HOST") == "ctypes" else "pybind11")
    for name, n_args in (("fwd", 12), ("varlen_fwd", 19), ("fwd_kvcache", 18)):     # flash_api.cpp:240-252,422-441,671-690
        doc = getattr(N, name).__doc__
        assert len(re.findall(r"arg\d+:", doc)) == n_args, doc
        # the ctypes implementation of the same op takes the same positional list
        params = [p for p in inspect.signature(getattr(ext, name)).parameters.values() if p.kind != p.KEYWORD_ONLY]
        assert len(params) == n_args
    assert N.last_call_info() == (0, 0, 0) or len(N.last_call_info()) == 3


def test_errors_are_runtime_errors_without_a_gpu():
    x = torch.zeros(1, 4, 2, 64)
    with pytest.raises(RuntimeError, match="only support fp16 and bf16"):
        N.fwd(x, x, x, None, None, 0.0, 0.1, False, -1, -1, False, None)
    h = x.half()
    with pytest.raises(RuntimeError, match="must be on CUDA"):
        N.fwd(h, h, h, None, None, 0.0, 0.1, False, -1, -1, False, None)


def _both(f):
    """Run f() through the compiled binding and through ctypes; returns the two results."""
    assert ext._N is not None
    a = f()
    info_a = dict(ext.last_call_info)
    saved = ext._N
    ext._N = None
    try:
        b = f()
        info_b = dict(ext.last_call_info)
    finally:
        ext._N = saved
    assert info_a == info_b, (info_a, info_b)
    return a, b


@pytest.mark.gpu
def test_both_host_bindings_agree_bitwise():
    torch.manual_seed(3)
    dev, dt = "cuda", torch.bfloat16
    # fwd: GQA causal, ALiBi, out=, and a head dim that needs padding (d = 60 -> 64)
    q = torch.randn(2, 300, 8, 128, device=dev, dtype=dt); k = torch.randn(2, 300, 2, 128, device=dev, dtype=dt); v = torch.randn_like(k)
    slopes = torch.rand(8, device=dev, dtype=torch.float32) * 0.3
    a, b = _both(lambda: ext.fwd(q, k, v, None, slopes, 0.0, 128 ** -0.5, True, -1, -1, False, None))
    assert len(a) == len(b) == 8 and torch.equal(a[0], b[0]) and torch.equal(a[5], b[5])
    with pytest.raises(RuntimeError, match="does not support dropout"):
        N.fwd(q, k, v, None, None, 0.1, 0.1, False, -1, -1, False, None)
    with pytest.raises(RuntimeError, match="Number of heads in key/value must divide"):
        N.fwd(q, k[:, :, :1].expand(-1, -1, 3, -1).contiguous(), v[:, :, :1].expand(-1, -1, 3, -1).contiguous(), None, None, 0.0, 0.1,
              False, -1, -1, False, None)
    assert a[5].shape == (2, 8, 300) and a[6].numel() == 0 and a[7].shape == (2,)
    o_arg = torch.empty_like(q)
    a, b = _both(lambda: ext.fwd(q, k, v, o_arg, None, 0.0, 128 ** -0.5, False, 100, 20, False, None))
    assert a[0].data_ptr() == o_arg.data_ptr() and torch.equal(a[0], b[0])
    q72 = torch.randn(2, 130, 4, 60, device=dev, dtype=torch.float16); k72 = torch.randn_like(q72); v72 = torch.randn_like(q72)
    a, b = _both(lambda: ext.fwd(q72, k72, v72, None, None, 0.0, 60 ** -0.5, True, -1, -1, False, None))
    assert a[0].shape == q72.shape and a[4].shape[-1] == 64 and torch.equal(a[0], b[0])
    # varlen: packed K/V with seqused_k, and paged K/V (mixed prefill + decode batch)
    lens = [257, 1, 64, 513]
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device=dev)
    qv = torch.randn(sum(lens), 8, 128, device=dev, dtype=dt); kv = torch.randn(sum(lens), 2, 128, device=dev, dtype=dt); vv = torch.randn_like(kv)
    used = torch.tensor([200, 1, 64, 500], dtype=torch.int32, device=dev)
    a, b = _both(lambda: ext.varlen_fwd(qv, kv, vv, None, cu, cu, used, None, None, max(lens), max(lens), 0.0, 128 ** -0.5, False,
                                        True, -1, -1, False, None))
    assert torch.equal(a[0], b[0]) and a[5].shape == (4, 8, 513)
    page, mb = 16, 40
    kc = torch.randn(4 * mb, page, 2, 128, device=dev, dtype=dt); vc = torch.randn_like(kc)
    bt = torch.randperm(4 * mb, device=dev).to(torch.int32).view(4, mb).contiguous()
    ck = torch.tensor([0, 257, 257 + 600, 257 + 600 + 64, 257 + 600 + 64 + 513], dtype=torch.int32, device=dev)
    a, b = _both(lambda: ext.varlen_fwd(qv, kc, vc, None, cu, ck, None, bt, None, max(lens), 600, 0.0, 128 ** -0.5, False,
                                        True, -1, -1, False, None))
    assert torch.equal(a[0], b[0])
    # fwd_kvcache: paged split decode; append + rotary into a dense cache
    qd = torch.randn(4, 1, 8, 128, device=dev, dtype=dt)
    cl = torch.tensor([600, 17, 640, 333], dtype=torch.int32, device=dev)
    a, b = _both(lambda: ext.fwd_kvcache(qd, kc, vc, None, None, cl, None, None, None, bt, None, None, 128 ** -0.5, True, -1, -1, False, 0))
    assert len(a) == 2 and torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    cache_k = torch.randn(4, 256, 2, 128, device=dev, dtype=dt); cache_v = torch.randn_like(cache_k)
    kn = torch.randn(4, 1, 2, 128, device=dev, dtype=dt); vn = torch.randn_like(kn)
    ang = torch.rand(256, 32, device=dev) * 2 * math.pi
    cos, sin = ang.cos().to(dt), ang.sin().to(dt)
    cl2 = torch.tensor([10, 100, 255, 0], dtype=torch.int32, device=dev)

    def appended():
        ck_, cv_ = cache_k.clone(), cache_v.clone()
        r = ext.fwd_kvcache(qd, ck_, cv_, kn, vn, cl2, cos, sin, None, None, None, None, 128 ** -0.5, True, -1, -1, True, 1)
        return r[0], r[1], ck_, cv_
    a, b = _both(appended)
    assert all(torch.equal(x, y) for x, y in zip(a, b))


@pytest.mark.gpu
def test_compiled_binding_under_cuda_graph_capture():
    torch.manual_seed(4)
    dev, dt, page, mb = "cuda", torch.bfloat16, 16, 64
    kc = torch.randn(2 * mb, page, 2, 128, device=dev, dtype=dt); vc = torch.randn_like(kc)
    bt = torch.randperm(2 * mb, device=dev).to(torch.int32).view(2, mb).contiguous()
    q = torch.randn(2, 1, 8, 128, device=dev, dtype=dt)
    cl = torch.tensor([1000, 1024], dtype=torch.int32, device=dev)
    f = lambda: ext.fwd_kvcache(q, kc, vc, None, None, cl, None, None, None, bt, None, None, 128 ** -0.5, True, -1, -1, False, 4)
    eager = f()[0].clone()
    s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        f()
    torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        out = f()[0]
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, eager)
