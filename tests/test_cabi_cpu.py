"""The C-ABI library builds, loads, and exports exactly what include/dlrmx.h declares (no GPU needed: only
argument-free / host-only entry points are CALLED here)."""
import ctypes
import os
import re

import pytest

from dlrm_yx_b200 import _lib, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "dlrmx.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(dlrmx_[a-z0-9_]+)\s*\(", text)))


def test_library_builds_and_loads():
    path = build.build()
    assert os.path.exists(path)
    lib = _lib.load()
    assert lib.dlrmx_version() == 100
    assert lib.dlrmx_launch_count() >= 0


def test_every_declared_symbol_is_exported_and_bound():
    build.build()
    syms = header_symbols()
    assert len(syms) >= 19
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for s in syms:
        assert hasattr(lib, s), "libdlrmx.so does not export " + s
        assert s in _lib.SIGNATURES, "dlrm_yx_b200/_lib.py does not bind " + s
    assert sorted(_lib.SIGNATURES) == syms, "binding table and header disagree"


def test_argument_errors_are_reported_not_crashed():
    lib = _lib.load()
    # unsupported D is rejected before any launch, with a message
    rc = lib.dlrmx_embbag_fwd_i32(0, 0, 0, 0, 0, 1, 1, 5000, 0, 0, 0, 0, 0, 0, 0)
    assert rc == -1
    assert b"unsupported D" in lib.dlrmx_last_error()
    assert lib.dlrmx_set_tunable(99, 1) == -1
    assert lib.dlrmx_bwd_sort_workspace_bytes(1 << 20, 1 << 10) > (1 << 23)
    # empty problems succeed without touching the device
    assert lib.dlrmx_embbag_fwd_i64(0, 0, 0, 0, 0, 0, 0, 128, 0, 0, 0, 0, 0, 0, 0) == 0


def test_no_cpu_fallback():
    """CPU tensors are refused loudly by the product path."""
    import torch
    from dlrm_yx_b200 import embedding, interaction
    E = embedding.EmbeddingBagSum(4, 8)
    with pytest.raises(_lib.DlrmxError):
        E(torch.tensor([0, 1]), torch.tensor([0, 1]))
    with pytest.raises(_lib.DlrmxError):
        interaction.interact_features(torch.zeros(2, 8), [torch.zeros(2, 8)])


def test_product_code_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "dlrm_yx_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("SURVEY", ""), "%s mentions the oracle" % f
