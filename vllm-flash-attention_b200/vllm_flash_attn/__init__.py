"""vllm_flash_attn — B200-native (sm_100a) drop-in for Yard1/vllm-flash-attention's Python
package (reference vllm_flash_attn/__init__.py:1-11): same seven functions, same version tag."""
__version__ = "2.5.8.post3"

from vllm_flash_attn.flash_attn_interface import (
    flash_attn_func,
    flash_attn_kvpacked_func,
    flash_attn_qkvpacked_func,
    flash_attn_varlen_func,
    flash_attn_varlen_kvpacked_func,
    flash_attn_varlen_qkvpacked_func,
    flash_attn_with_kvcache,
)

__all__ = [
    "flash_attn_func", "flash_attn_kvpacked_func", "flash_attn_qkvpacked_func",
    "flash_attn_varlen_func", "flash_attn_varlen_kvpacked_func",
    "flash_attn_varlen_qkvpacked_func", "flash_attn_with_kvcache",
]
