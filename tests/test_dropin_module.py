"""The drop-in boundary one level below the Python package: a module importable as
``vllm_flash_attn_2_cuda`` with the reference's three ops, so that the reference's own unmodified
flash_attn_interface.py can sit on top of libfa_b200.so.
* CPU: the module imports under that name and its ops take exactly the positional lists the reference's call sites
  pass (parsed from the reference file with ``ast`` when /root/reference is present).
* GPU: the three ops called with the reference's positional lists (C1 + varlen + paged kvcache), and — wherever the
  reference's interface file is reachable ($FA_REF_INTERFACE or /root/reference; it does not travel to the GPU box by
  itself) — the reference's unmodified Python layer running C1 and a kvcache call on top of this module."""
import ast
import importlib.util
import inspect
import os
import sys

import pytest
import torch

import _util  # noqa: F401  (sys.path: the package directory)

REF_IFACE = os.environ.get("FA_REF_INTERFACE", "/root/reference/vllm_flash_attn/flash_attn_interface.py")


def _load_reference_interface():
    spec = importlib.util.spec_from_file_location("_reference_flash_attn_interface", REF_IFACE)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)          # executes `import vllm_flash_attn_2_cuda as flash_attn_cuda`: resolves to ours
    return mod


def test_top_level_module_exports_the_three_ops():
    import vllm_flash_attn_2_cuda as m
    from vllm_flash_attn import vllm_flash_attn_2_cuda as shim
    assert m.fwd is shim.fwd and m.varlen_fwd is shim.varlen_fwd and m.fwd_kvcache is shim.fwd_kvcache
    # positional arity of the pybind signatures (flash_api.cpp:240-252, :422-441, :671-690)
    npos = lambda f: sum(p.kind == p.POSITIONAL_OR_KEYWORD for p in inspect.signature(f).parameters.values())
    assert (npos(m.fwd), npos(m.varlen_fwd), npos(m.fwd_kvcache)) == (12, 19, 18)


@pytest.mark.skipif(not os.path.exists(REF_IFACE), reason="reference interface file not reachable")
def test_reference_call_sites_match_our_positional_lists():
    import vllm_flash_attn_2_cuda as m
    tree = ast.parse(open(REF_IFACE).read())
    seen = {}
    for node in ast.walk(tree):
        if isinstance(node, ast.Call) and isinstance(node.func, ast.Attribute) and isinstance(node.func.value, ast.Name) \
                and node.func.value.id == "flash_attn_cuda":
            assert not node.keywords, "the reference passes everything positionally"
            seen.setdefault(node.func.attr, set()).add(len(node.args))
    assert {"fwd", "varlen_fwd", "fwd_kvcache"} <= set(seen), seen          # the forward call sites (bwd ones are out of scope)
    npos = lambda f: sum(p.kind == p.POSITIONAL_OR_KEYWORD for p in inspect.signature(f).parameters.values())
    for name in ("fwd", "varlen_fwd", "fwd_kvcache"):
        assert seen[name] == {npos(getattr(m, name))}, (name, seen[name])
    ref = _load_reference_interface()
    for fn in ("flash_attn_func", "flash_attn_varlen_func", "flash_attn_with_kvcache", "flash_attn_qkvpacked_func",
               "flash_attn_kvpacked_func", "flash_attn_varlen_qkvpacked_func", "flash_attn_varlen_kvpacked_func"):
        assert callable(getattr(ref, fn))
    assert ref.flash_attn_cuda is sys.modules["vllm_flash_attn_2_cuda"]


@pytest.mark.gpu
def test_ops_with_the_reference_positional_lists():
    """fwd / varlen_fwd / fwd_kvcache exactly as reference flash_attn_interface.py:51-64, :88-108, :1233-1252 call them."""
    import attention_oracle as orc
    import vllm_flash_attn_2_cuda as m
    torch.manual_seed(0)
    dev, dt = "cuda", torch.bfloat16
    q = torch.randn(2, 512, 8, 64, device=dev, dtype=dt); k = torch.randn_like(q); v = torch.randn_like(q)       # C1
    out, q_, k_, v_, out_padded, lse, S_dmask, rng_state = m.fwd(q, k, v, None, None, 0.0, 64 ** -0.5, False, -1, -1, False, None)
    ref32, _ = orc.attention_ref(q, k, v)
    pt, _ = orc.attention_ref(q, k, v, upcast=False, reorder_ops=True)
    assert (out.float() - ref32.float()).abs().max() <= 2 * (pt.float() - ref32.float()).abs().max() + 1e-5
    assert lse.shape == (2, 8, 512) and out_padded.shape == out.shape and rng_state.numel() == 2
    cu = torch.tensor([0, 300, 512], dtype=torch.int32, device=dev)
    qv, kv_, vv = q[0], k[0], v[0]
    res = m.varlen_fwd(qv, kv_, vv, None, cu, cu, None, None, None, 300, 300, 0.0, 64 ** -0.5, False, True, -1, -1, False, None)
    r32, _ = orc.attention_ref_varlen(qv, kv_, vv, cu.cpu(), cu.cpu(), causal=True)
    rpt, _ = orc.attention_ref_varlen(qv, kv_, vv, cu.cpu(), cu.cpu(), causal=True, upcast=False, reorder_ops=True)
    assert len(res) == 8 and (res[0].float() - r32.float()).abs().max() <= 2 * (rpt.float() - r32.float()).abs().max() + 1e-5
    kc = torch.randn(64, 16, 2, 64, device=dev, dtype=dt); vc = torch.randn_like(kc)
    bt = torch.randperm(64, device=dev).to(torch.int32).view(2, 32)
    lens = torch.tensor([500, 37], dtype=torch.int32, device=dev)
    q1 = torch.randn(2, 1, 8, 64, device=dev, dtype=dt)
    o1, l1 = m.fwd_kvcache(q1, kc, vc, None, None, lens, None, None, None, bt, None, None, 64 ** -0.5, True, -1, -1, True, 0)
    kf, vf = orc.gather_paged(kc, bt), orc.gather_paged(vc, bt)
    kpm = torch.arange(512, device=dev).expand(2, 512) < lens.view(-1, 1)
    r32, _ = orc.attention_ref(q1, kf, vf, None, kpm)
    rpt, _ = orc.attention_ref(q1, kf, vf, None, kpm, upcast=False, reorder_ops=True)
    assert (o1.float() - r32.float()).abs().max() <= 3 * (rpt.float() - r32.float()).abs().max() + 1e-5 and l1.shape == (2, 8, 1)


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(REF_IFACE), reason="reference interface file not reachable on this box")
def test_reference_python_layer_runs_on_this_library():
    """The reference's UNMODIFIED flash_attn_interface.py (autograd wrappers and all) on top of our extension module:
    C1 through flash_attn_func and one paged decode through flash_attn_with_kvcache."""
    import attention_oracle as orc
    ref = _load_reference_interface()
    torch.manual_seed(0)
    dev, dt = "cuda", torch.bfloat16
    q = torch.randn(2, 512, 8, 64, device=dev, dtype=dt); k = torch.randn_like(q); v = torch.randn_like(q)
    out = ref.flash_attn_func(q, k, v)
    ref32, _ = orc.attention_ref(q, k, v)
    pt, _ = orc.attention_ref(q, k, v, upcast=False, reorder_ops=True)
    assert (out.float() - ref32.float()).abs().max() <= 2 * (pt.float() - ref32.float()).abs().max() + 1e-5
    out_c, lse, _ = ref.flash_attn_func(q, k, v, causal=True, return_attn_probs=True)
    assert lse.shape == (2, 8, 512)
    kc = torch.randn(64, 16, 2, 64, device=dev, dtype=dt); vc = torch.randn_like(kc)
    bt = torch.randperm(64, device=dev).to(torch.int32).view(2, 32)
    lens = torch.tensor([500, 37], dtype=torch.int32, device=dev)
    q1 = torch.randn(2, 1, 8, 64, device=dev, dtype=dt)
    o1 = ref.flash_attn_with_kvcache(q1, kc, vc, cache_seqlens=lens, block_table=bt, causal=True)
    kf, vf = orc.gather_paged(kc, bt), orc.gather_paged(vc, bt)
    kpm = torch.arange(512, device=dev).expand(2, 512) < lens.view(-1, 1)
    r32, _ = orc.attention_ref(q1, kf, vf, None, kpm)
    rpt, _ = orc.attention_ref(q1, kf, vf, None, kpm, upcast=False, reorder_ops=True)
    assert (o1.float() - r32.float()).abs().max() <= 3 * (rpt.float() - r32.float()).abs().max() + 1e-5
