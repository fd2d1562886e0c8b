"""Top-level ``vllm_flash_attn_2_cuda`` — the name under which the reference's own Python layer imports its
pybind11 extension (reference vllm_flash_attn/flash_attn_interface.py:10, ``import vllm_flash_attn_2_cuda as
flash_attn_cuda``; module definition csrc/flash_attn/flash_api.cpp:939-946).  With this directory on sys.path the
reference's UNMODIFIED flash_attn_interface.py runs on libfa_b200.so: the same three ops, the same positional argument
lists and return lists (flash_api.cpp:240-252, :422-441, :671-690).  The implementation is the ctypes shim over the C
ABI in vllm_flash_attn/vllm_flash_attn_2_cuda.py; nothing is duplicated here."""
from vllm_flash_attn.vllm_flash_attn_2_cuda import fwd, fwd_kvcache, varlen_fwd  # noqa: F401

__all__ = ["fwd", "varlen_fwd", "fwd_kvcache"]
