import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "vllm-flash-attention_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


def _ensure_library_built():
    """A fresh checkout has no libfa_b200.so (build outputs are git-ignored).  The tests are about the library, so
    build it in-tree once (nvcc cross-compiles sm_100a without a GPU, ~2 min) instead of failing at import; the
    package itself still refuses to import without it."""
    import subprocess
    pkg = os.path.join(ROOT, "vllm-flash-attention_b200")
    so = os.path.join(pkg, "vllm_flash_attn", "libfa_b200.so")
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    if not os.path.exists(so) and os.path.exists(nvcc):
        subprocess.run(["bash", os.path.join(pkg, "build.sh")], check=False)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    _ensure_library_built()


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(autouse=True)
def _unpinned_kernel_choice():
    """Tests that pin a kernel family (ext._force_kernel) must not leak the pin into the next test."""
    yield
    mod = sys.modules.get("vllm_flash_attn.vllm_flash_attn_2_cuda")
    if mod is not None:
        mod._force_kernel = 0
