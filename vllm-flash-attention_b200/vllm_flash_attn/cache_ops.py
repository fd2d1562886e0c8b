"""Caller-side companion of the attention ops: the writer of the paged KV cache the ops read.  Not
part of the reference repository — vLLM brings its own (``reshape_and_cache_flash``); this one has the same contract, so
that a serving loop can be built and tested on this package alone (tests/test_serving_loop_gpu.py)."""
from __future__ import annotations

import torch

from ._cabi import check, lib


def reshape_and_cache_flash(key: torch.Tensor, value: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                            slot_mapping: torch.Tensor) -> None:
    """key, value: (num_tokens, nheads_k, headdim); caches: (num_blocks, page_block_size, nheads_k, headdim);
    slot_mapping: (num_tokens,) int64 on the same device, slot = block * page_block_size + offset, negative = skip.
    Runs on the current stream of key's device."""
    if key.dim() != 3 or value.shape != key.shape:
        raise RuntimeError("key and value must be (num_tokens, nheads_k, headdim)")
    if key_cache.dim() != 4 or value_cache.shape != key_cache.shape or key_cache.shape[2:] != key.shape[1:]:
        raise RuntimeError("caches must be (num_blocks, page_block_size, nheads_k, headdim) matching key")
    if key.dtype not in (torch.float16, torch.bfloat16) or not (value.dtype == key_cache.dtype == value_cache.dtype == key.dtype):
        raise RuntimeError("key, value and the caches must share one of fp16 / bf16")
    if slot_mapping.dtype != torch.int64 or slot_mapping.numel() != key.shape[0] or not slot_mapping.is_contiguous():
        raise RuntimeError("slot_mapping must be a contiguous int64 tensor of num_tokens elements")
    for t in (key, value, key_cache, value_cache, slot_mapping):
        if not t.is_cuda or t.device != key.device:
            raise RuntimeError("all tensors must be on the same CUDA device")
    if key.stride(-1) != 1 or value.stride(-1) != 1 or key_cache.stride(-1) != 1 or value_cache.stride(-1) != 1:
        raise RuntimeError("last dimension must be contiguous")
    n, hk, d = key.shape
    with torch.cuda.device(key.device):
        check(lib.fa_b200_reshape_and_cache(
            key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(), slot_mapping.data_ptr(),
            n, hk, d, key_cache.shape[1], key_cache.shape[0],
            key.stride(0), key.stride(1), value.stride(0), value.stride(1),
            key_cache.stride(0), key_cache.stride(1), key_cache.stride(2),
            value_cache.stride(0), value_cache.stride(1), value_cache.stride(2),
            torch.cuda.current_stream(key.device).cuda_stream))
