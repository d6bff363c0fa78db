"""ctypes binding of libdlrmx.so (C ABI in include/dlrmx.h).

There is no CPU fallback and no alternative backend: if the shared library is missing or a call
fails, this raises.  The library is built in-tree by dlrm_yx_b200/build.py (nvcc, sm_100a).
"""
import ctypes
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "libdlrmx.so")

c_void_p, c_int, c_int64, c_float, c_size_t = (ctypes.c_void_p, ctypes.c_int, ctypes.c_int64,
                                                ctypes.c_float, ctypes.c_size_t)

# name -> (restype, argtypes).  Must list every symbol include/dlrmx.h declares
# (tests/test_cabi.py checks the header against this table and against the .so).
SIGNATURES = {
    "dlrmx_version": (c_int, []),
    "dlrmx_last_error": (ctypes.c_char_p, []),
    "dlrmx_launch_count": (c_int64, []),
    "dlrmx_set_tunable": (c_int, [c_int, c_int]),
    "dlrmx_embbag_fwd_i32": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_int64,
                                     c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p]),
    "dlrmx_embbag_fwd_i64": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_int64,
                                     c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p]),
    "dlrmx_embbag_fwd_scatter_i32": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int,
                                             c_int64, c_void_p, c_void_p, c_int, c_int64, c_int64,
                                             c_void_p, c_void_p]),
    "dlrmx_embbag_fwd_scatter_i64": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int,
                                             c_int64, c_void_p, c_void_p, c_int, c_int64, c_int64,
                                             c_void_p, c_void_p]),
    "dlrmx_embbag_fwd_bf16_i32": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_int64,
                                          c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p]),
    "dlrmx_embbag_fwd_bf16_i64": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_int64,
                                          c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p]),
    "dlrmx_bwd_sort_workspace_bytes": (c_size_t, [c_int64, c_int64]),
    "dlrmx_bwd_sort_i32": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p, c_int64, c_void_p,
                                   c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "dlrmx_bwd_sort_i64": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p, c_int64, c_void_p,
                                   c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "dlrmx_bwd_segments_workspace_bytes": (c_size_t, [c_int64]),
    "dlrmx_bwd_segments": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
                                   c_void_p]),
    "dlrmx_bwd_update_workspace_bytes": (c_size_t, [c_int64, c_int]),
    "dlrmx_bwd_update": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int64, c_void_p, c_int64,
                                 c_int64, c_void_p, c_void_p, c_int64, c_int64, c_float, c_float, c_void_p, c_size_t,
                                 c_void_p]),
    "dlrmx_bwd_update_bf16": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int64, c_void_p, c_int64,
                                      c_int64, c_void_p, c_void_p, c_int64, c_int64, c_float, c_float, c_void_p, c_size_t,
                                      c_void_p]),
    "dlrmx_bwd_update_ex": (c_int, [c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int64, c_void_p,
                                    c_int64, c_int64, c_void_p, c_void_p, c_int64, c_int64, c_float, c_float, c_void_p,
                                    c_size_t, c_void_p]),
    "dlrmx_bwd_coalesce": (c_int, [c_int, c_int, c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p,
                                   c_int64, c_void_p, c_void_p]),
    "dlrmx_rows_update": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_float,
                                  c_float, c_void_p]),
    "dlrmx_dense_rows_update": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int64, c_float, c_float,
                                        c_void_p]),
    "dlrmx_interact_dot_fwd": (c_int, [c_void_p, c_int64, c_int, c_int, c_int, c_int, c_void_p, c_int64,
                                       c_void_p]),
    "dlrmx_interact_dot_bwd": (c_int, [c_void_p, c_int64, c_int, c_int, c_int, c_int, c_void_p, c_int64,
                                       c_void_p, c_int64, c_void_p]),
    "dlrmx_interact_dot_fwd_mixed": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_int, c_int, c_int, c_int, c_void_p,
                                             c_int64, c_void_p]),
    "dlrmx_interact_dot_bwd_mixed": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_int, c_int, c_int, c_int, c_void_p,
                                             c_int64, c_void_p, c_int64, c_void_p, c_int, c_void_p]),
    "dlrmx_relu_bwd_bias_grad": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                         c_int, c_int, c_void_p]),
    "dlrmx_relu_bwd_bias_grad_noalloc": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "dlrmx_gemm_bf16": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int, c_void_p, c_int64, c_int, c_int, c_int, c_int,
                                c_void_p, c_void_p, c_int64, c_void_p, c_void_p]),
    "dlrmx_gemm_wgrad_workspace_bytes": (c_size_t, [c_int, ctypes.POINTER(c_int), ctypes.POINTER(c_int), c_int]),
    "dlrmx_gemm_wgrad_bf16": (c_int, [c_int, ctypes.POINTER(c_void_p), ctypes.POINTER(c_int64), ctypes.POINTER(c_void_p),
                                      ctypes.POINTER(c_int64), ctypes.POINTER(c_void_p), ctypes.POINTER(c_int64),
                                      ctypes.POINTER(c_void_p), ctypes.POINTER(c_int), ctypes.POINTER(c_int), c_int,
                                      c_void_p, c_size_t, c_void_p]),
    "dlrmx_head_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "dlrmx_head_bwd_workspace_bytes": (c_size_t, [c_int, c_int]),
    "dlrmx_head_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                               c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "dlrmx_dense_opt": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_float, c_float, c_float,
                                c_void_p]),
    "dlrmx_a2a_unpack": (c_int, [c_void_p, c_int, ctypes.POINTER(c_int), c_int, c_int, c_void_p, c_int64, c_int64,
                                 c_int, c_void_p]),
    "dlrmx_a2a_pack_grad": (c_int, [c_void_p, c_int64, c_int64, c_int, c_int, ctypes.POINTER(c_int), c_int, c_int,
                                    c_void_p, c_void_p]),
    "dlrmx_a2a_push_grad": (c_int, [c_void_p, c_int64, c_int64, c_int, c_int, ctypes.POINTER(c_int), c_int, c_int,
                                    ctypes.POINTER(c_void_p), c_void_p]),
    "dlrmx_a2a_push_grad_bf16": (c_int, [c_void_p, c_int64, c_int64, c_int, c_int, ctypes.POINTER(c_int), c_int, c_int,
                                         ctypes.POINTER(c_void_p), c_void_p]),
    "dlrmx_peer_allreduce": (c_int, [ctypes.POINTER(c_void_p), c_int, c_int, c_int64, c_int, c_void_p]),
    "dlrmx_bce_mean_workspace_bytes": (c_size_t, [c_int64]),
    "dlrmx_bce_mean_fwd": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p]),
    "dlrmx_bce_mean_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p]),
    "dlrmx_embbag_psw_grad_i32": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p,
                                          c_int64, c_void_p, c_void_p]),
    "dlrmx_embbag_psw_grad_i64": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p,
                                          c_int64, c_void_p, c_void_p]),
    "dlrmx_bwd_coalesce_weighted": (c_int, [c_int, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                            c_int64, c_void_p, c_void_p]),
    "dlrmx_criteo_bin_batch_i32": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                           c_void_p]),
    "dlrmx_criteo_bin_batch_i64": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                           c_void_p]),
}

OPT_SGD = 0
OPT_RWSADAGRAD = 1
# test hooks (include/dlrmx.h: dlrmx_set_tunable)
TUNE_IX_CUDACORE = 2   # 1: force the CUDA-core interaction kernels (default: 3xTF32 tensor-core kernels)
TUNE_BWD_SEG = 3       # sorted positions per group of the fused update (32/64/128/256; 0: chosen per optimizer)
TUNE_IX_STAGE = 10     # 1: per-thread cp.async staging instead of TMA bulk copies (interaction kernels)
TUNE_IX_GRID = 11      # CTAs per SM of the interaction kernels' persistent grid

_lib = None


class DlrmxError(RuntimeError):
    pass


def load():
    """Loads libdlrmx.so (once).  Raises if it has not been built -- never falls back."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise DlrmxError(
            "libdlrmx.so not found at %s: build it with `python -m dlrm_yx_b200.build` "
            "(or __graft_entry__.build()); there is no CPU or PyTorch fallback." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc, what):
    if rc != 0:
        msg = load().dlrmx_last_error()
        raise DlrmxError("%s failed (%d): %s" % (what, rc, msg.decode() if msg else "?"))


def launch_count():
    return int(load().dlrmx_launch_count())


_TUNABLES = {}


def set_tunable(which, value):
    check(load().dlrmx_set_tunable(which, value), "dlrmx_set_tunable")
    _TUNABLES[which] = value


def get_tunable_py(which):
    """Value last set through set_tunable (0 = default)."""
    return _TUNABLES.get(which, 0)
