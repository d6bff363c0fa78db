"""Builds dlrm_yx_b200/libdlrmx.so (the C-ABI library declared in include/dlrmx.h) in-tree with nvcc.

sm_100a only: `-gencode arch=compute_100a,code=sm_100a -lineinfo`.  nvcc cross-compiles without a GPU, so
this also is the "does it build" check of __graft_entry__.build().  The .so is git-ignored but travels
with the repo snapshot to the GPU box.  Replaces yx_modfs/setup.py (which passes only -O2 and needs a
visible GPU and NCCL_HOME, yx_modfs/setup.py:16-38).
"""
import concurrent.futures
import hashlib
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
OBJ = os.path.join(ROOT, "build", "obj")
LIB = os.path.join(PKG, "libdlrmx.so")
SOURCES = ["api.cu", "embbag_fwd_i32.cu", "embbag_fwd_i64.cu", "embbag_fwd_bf16_i32.cu", "embbag_fwd_bf16_i64.cu", "radix_sort.cu", "segments.cu", "embbag_bwd.cu", "interact.cu", "interact_mma.cu", "mlp_fused.cu", "gemm_tc.cu",
           "a2a_layout.cu", "peer_allreduce.cu", "criteo_bin.cu", "embbag_psw.cu", "loss.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: cannot build libdlrmx.so")


def _digest():
    h = hashlib.sha256()
    names = sorted(os.listdir(CSRC)) + ["../../include/dlrmx.h"]
    for n in names:
        with open(os.path.join(CSRC, n), "rb") as f:
            h.update(n.encode())
            h.update(f.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def needs_build():
    stamp = LIB + ".stamp"
    if not (os.path.exists(LIB) and os.path.exists(stamp)):
        return True
    with open(stamp) as f:
        return f.read().strip() != _digest()


def build(force=False, verbose=False):
    """Compile every .cu and link the shared library.  Returns the library path."""
    if not force and not needs_build():
        return LIB
    nvcc = _nvcc()
    os.makedirs(OBJ, exist_ok=True)

    def compile_one(src):
        obj = os.path.join(OBJ, src.replace(".cu", ".o"))
        cmd = [nvcc] + NVCC_FLAGS + ["-c", os.path.join(CSRC, src), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
        if verbose and r.stderr.strip():
            sys.stderr.write(r.stderr)
        return obj

    with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    tmp = LIB + ".tmp"
    cmd = [nvcc, "-shared", "-o", tmp] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    os.replace(tmp, LIB)
    with open(LIB + ".stamp", "w") as f:
        f.write(_digest())
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
