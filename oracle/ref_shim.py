"""TEST INFRASTRUCTURE ONLY -- imports the UNMODIFIED reference from /root/reference.

This module exists so that the CPU oracle (oracle/oracle_np.py, oracle/oracle_c.c) can be
pinned against the reference's own Python path and so that golden vectors under
tests/golden/ can be (re)generated.  It only works in the build container, where
/root/reference is mounted; the GPU box has no /root/reference and nothing on the
`-m gpu` tests / smoke() / bench.py path imports this file.

Three import shims are needed: the reference imports `h5py`
(dlrm_data_pytorch.py:27) and `torchviz` (dlrm_s_pytorch.py:113), which are not installed,
and `torch.profiler.ExecutionGraphObserver` (dlrm_s_pytorch.py:90), which torch 2.11 removed.
Nothing in the reference is edited or copied.
"""
import importlib
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("DLRM_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "dlrm_s_pytorch.py"))


class _H5File(dict):
    """Dict-backed stand-in for h5py.File, enough for import (never used for I/O here)."""

    def __init__(self, *a, **k):
        super().__init__()

    def create_dataset(self, name, data=None, **k):
        self[name] = data
        return data

    def close(self):
        pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


def _install_shims():
    import torch
    import torch.profiler

    if "h5py" not in sys.modules:
        try:
            importlib.import_module("h5py")
        except ImportError:
            m = types.ModuleType("h5py")
            m.File = _H5File
            sys.modules["h5py"] = m
    if "torchviz" not in sys.modules:
        try:
            importlib.import_module("torchviz")
        except ImportError:
            m = types.ModuleType("torchviz")
            m.make_dot = lambda *a, **k: None
            sys.modules["torchviz"] = m
    if not hasattr(torch.profiler, "ExecutionGraphObserver"):
        torch.profiler.ExecutionGraphObserver = getattr(
            torch.profiler, "ExecutionTraceObserver", object
        )


_cache = {}


def load_reference():
    """Returns a namespace with the reference's modules:
    .dlrm (dlrm_s_pytorch), .dp (dlrm_data_pytorch), .ext_dist (extend_distributed),
    .sharders, .rwsadagrad (optim.rwsadagrad)."""
    if "ns" in _cache:
        return _cache["ns"]
    if not reference_available():
        raise RuntimeError("reference not mounted at %s" % REFERENCE_ROOT)
    _install_shims()
    import builtins

    saved_print = builtins.print
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    try:
        ns = types.SimpleNamespace()
        ns.ext_dist = importlib.import_module("extend_distributed")
        ns.sharders = importlib.import_module("sharders")
        ns.dp = importlib.import_module("dlrm_data_pytorch")
        ns.rwsadagrad = importlib.import_module("optim.rwsadagrad")
        ns.dlrm = importlib.import_module("dlrm_s_pytorch")
    finally:
        # extend_distributed.py:654-662 replaces builtins.print with a rank-0 filter at import.
        builtins.print = saved_print
    # single-process defaults as init_distributed would leave them (extend_distributed.py:198-202)
    if ns.ext_dist.my_size < 0:
        ns.ext_dist.my_size = 1
        ns.ext_dist.my_rank = 0
        ns.ext_dist.my_local_rank = 0
        ns.ext_dist.my_local_size = 1
    _cache["ns"] = ns
    return ns
