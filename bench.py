"""bench.py -- DLRM train samples/s on the Terabyte/MLPerf shape (BASELINE.json configs[3]/[4]) through the
B200 sparse hot path, plus the roofline of the embedding-bag kernels and the CPU baseline.

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference            # the reference's CPU path (oracle port) on the host cores

A "step" = one full training iteration on one synthetic mini-batch: bottom MLP, table-batched embedding-bag
lookup, (all-to-all,) dot interaction, top MLP, BCE loss, backward, fused row-wise-Adagrad update of the touched
table rows, Adagrad step of the MLPs.  One JSON line on stdout (rank 0).
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="terabyte", choices=["terabyte", "bench", "kaggle"])
    ap.add_argument("--batch", type=int, default=0, help="per-GPU mini-batch (default: the config's)")
    ap.add_argument("--global-batch", type=int, default=0, help="fix the GLOBAL batch (strong scaling)")
    ap.add_argument("--rows-cap", type=int, default=0, help="cap rows per table (debug only; marks the run)")
    ap.add_argument("--mlp", default="tf32", choices=["tf32", "fp32", "bf16"])
    ap.add_argument("--sharder", default="greedy")
    ap.add_argument("--optimizer", default="rwsadagrad", choices=["rwsadagrad", "sgd"])
    ap.add_argument("--nbatches", type=int, default=8, help="distinct synthetic batches cycled through")
    ap.add_argument("--cpu-batch", type=int, default=4096)
    ap.add_argument("--cpu-rows-cap", type=int, default=1000000)
    ap.add_argument("--cpu-steps", type=int, default=4)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy read+write)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""
    BITS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
            0x80: "hw_power_brake_slowdown"}

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self.ok = index, [], set(), None, False
        self._stop_ev = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception as e:  # NVML missing: clocks are reported as unknown, never invented
            self.err = repr(e)

    def sample(self):
        if not self.ok:
            return
        try:
            self.samples.append(int(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
            r = int(self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
            for bit, name in self.BITS.items():
                if r & bit:
                    self.reasons.add(name)
        except Exception:
            pass

    def run(self):
        while not self._stop_ev.is_set():
            self.sample()
            self._stop_ev.wait(0.02)

    def stop(self):
        self._stop_ev.set()
        self.join(timeout=2)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "note": "no NVML samples"}
        return {"sm_mhz": int(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------------
# synthetic workload
# ------------------------------------------------------------------------------------------------------
def config_of(args):
    from dlrm_yx_b200 import data
    cfg = dict(data.CONFIGS[args.config])
    cfg["rows"] = [min(r, args.rows_cap) if args.rows_cap else r for r in cfg["rows"]]
    if args.batch:
        cfg["B"] = args.batch
    return cfg


def make_batches(cfg, B_global, local_tables, lb_slice, n, seed):
    """n pinned host batches for one rank: dense X / targets for the LOCAL batch slice, packed int32 lookups of
    the LOCAL tables for the FULL global batch (dlrm_wrap's split, dlrm_s_pytorch.py:145-148)."""
    rng = np.random.default_rng(seed)          # same seed on every rank -> every rank draws the same global batch
    T_all, L = len(cfg["rows"]), cfg["L"]
    out = []
    for _ in range(n):
        X = rng.random((B_global, cfg["bot"][0]), dtype=np.float32)
        tgt = np.round(rng.random((B_global, 1))).astype(np.float32)
        idx_parts = []
        for t in range(T_all):
            E = cfg["rows"][t]
            ids = np.round(rng.random((B_global, L)) * (E - 1)).astype(np.int32)
            if L > 1:
                ids.sort(axis=1)
            if t in local_tables:
                idx_parts.append(ids.reshape(-1))
        idx = np.concatenate(idx_parts)
        off = (np.arange(len(local_tables) * B_global + 1, dtype=np.int64) * L).astype(np.int32)
        host = [torch.from_numpy(np.ascontiguousarray(X[lb_slice])), torch.from_numpy(idx), torch.from_numpy(off),
                torch.from_numpy(np.ascontiguousarray(tgt[lb_slice]))]
        out.append([t.pin_memory() for t in host])
    return out


def algorithmic_bytes(N, TB, D, U, ib=4):
    """forward = rows + indices + offsets + pooled rows; backward+row-wise Adagrad = pooled-row
    gradients + (key, bag) + read/write of the U distinct rows and their accumulators."""
    fwd = N * D * 4 + N * ib + TB * ib + TB * D * 4
    bwd = TB * D * 4 + N * 8 + 2 * U * D * 4 + 2 * U * 4
    return fwd, bwd
