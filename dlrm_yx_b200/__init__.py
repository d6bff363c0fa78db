"""dlrm_yx_b200 -- B200-native sparse hot path for DLRM (embedding-bag forward / sorted fused backward,
dot interaction, model-parallel exchange) behind the API surface of YuxinxinChen/dlrm-yx.

Host code is Python/PyTorch; all arithmetic on the path is hand-written sm_100a CUDA in libdlrmx.so,
reached through the C ABI of include/dlrmx.h.  There is no CPU fallback.
"""
from . import _lib  # noqa: F401

__all__ = ["_lib"]
__version__ = "0.1.0"
