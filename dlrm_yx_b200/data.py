"""Synthetic inputs of the reference's shapes and distributions, generated on the host with numpy/torch
(host-side plumbing either side of the hot path; nothing here is timed as kernel work).

Follows dlrm_data_pytorch.py: dense X ~ U[0,1) fp32 [B, m_den] (1178); per lookup a set of row ids
`round(U * (E_t - 1))`, sorted and de-duplicated within the bag (1195-1208); with a fixed number of indices
per lookup the draw is repeated until the bag has exactly L distinct ids (1131-1138 semantics); targets
round(U) fp32 [B, 1] (1095-1106).  `pack` gives the table-major packed int32 layout of the batched op
(dlrm_data_pytorch.py:748-753): indices [sum N_t], offsets [T*B+1].
"""
import numpy as np
import torch

TERABYTE_ROWS = [39884406, 39043, 17289, 7420, 20263, 3, 7120, 1543, 63, 38532951, 2953546, 403346, 10,
                 2208, 11938, 155, 4, 976, 14, 39979771, 25641295, 39664984, 585935, 12972, 108, 36]
KAGGLE_ROWS = [1460, 583, 10131227, 2202608, 305, 24, 12517, 633, 3, 93145, 5683, 8351593, 3194, 27,
               14992, 5461306, 10, 5652, 2173, 4, 7046547, 18, 15, 286181, 105, 142572]

# BASELINE.json configs as shapes.  rows / D / B / L / MLPs.
CONFIGS = {
    "tiny": dict(rows=[4, 3, 2], D=2, B=2, L=2, fixed=False, bot=[4, 3, 2], top=[8, 4, 2, 1]),
    "kaggle": dict(rows=KAGGLE_ROWS, D=16, B=128, L=1, fixed=True, bot=[13, 512, 256, 64, 16],
                   top=[367, 512, 256, 1]),
    "bench": dict(rows=[1000000] * 8, D=64, B=2048, L=100, fixed=True, bot=[512, 512, 64],
                  top=[100, 1024, 1024, 1024, 1]),
    "terabyte": dict(rows=TERABYTE_ROWS, D=128, B=32768, L=1, fixed=True, bot=[13, 512, 256, 128],
                     top=[479, 1024, 1024, 512, 256, 1]),
}


def num_interactions(T, D, itself=False):
    """dlrm_s_pytorch.py:1441-1450: width of the top-MLP input for the dot interaction."""
    F = T + 1
    return (F * (F + 1) // 2 if itself else F * (F - 1) // 2) + D


def random_batch(rows, B, L, m_den, fixed=True, rng=None):
    """One mini-batch on the host: (X [B,m_den] f32, lS_o list of T int64 [B], lS_i list of T int64 [N_t],
    target [B,1] f32), drawn like generate_uniform_input_batch."""
    rng = np.random.default_rng() if rng is None else rng
    X = rng.random((B, m_den), dtype=np.float32)
    lS_o, lS_i = [], []
    for E in rows:
        if L == 1 and fixed:
            idx = np.round(rng.random(B) * (E - 1)).astype(np.int64)
            off = np.arange(B, dtype=np.int64)
        elif fixed:
            idx = np.sort(np.round(rng.random((B, L)) * (E - 1)).astype(np.int64), axis=1)
            need = min(L, E)
            for _ in range(64):        # re-draw the few bags that lost entries to de-duplication
                bad = np.nonzero((np.diff(idx, axis=1) == 0).sum(axis=1) > L - need)[0]
                if bad.size == 0:
                    break
                idx[bad] = np.sort(np.round(rng.random((bad.size, L)) * (E - 1)).astype(np.int64), axis=1)
            if need < L:
                bags = [np.unique(r) for r in idx]
            else:
                bags = None
            if bags is None:
                off = np.arange(B, dtype=np.int64) * L
                idx = idx.reshape(-1)
            else:
                off = np.concatenate([[0], np.cumsum([b.size for b in bags])[:-1]]).astype(np.int64)
                idx = np.concatenate(bags)
        else:
            sizes = np.maximum(1, np.round(rng.random(B) * min(L, E)).astype(np.int64))
            bags = [np.unique(np.round(rng.random(s) * (E - 1)).astype(np.int64)) for s in sizes]
            off = np.concatenate([[0], np.cumsum([b.size for b in bags])[:-1]]).astype(np.int64)
            idx = np.concatenate(bags)
        lS_o.append(torch.from_numpy(off))
        lS_i.append(torch.from_numpy(idx))
    target = torch.from_numpy(np.round(rng.random((B, 1))).astype(np.float32))
    return torch.from_numpy(X), lS_o, lS_i, target


def pack(lS_o, lS_i, dtype=torch.int32):
    """Table-major packed layout of the batched op: (indices [sum N_t], offsets [T*B+1])."""
    sizes = [int(i.numel()) for i in lS_i]
    starts = np.concatenate([[0], np.cumsum(sizes)])
    offsets = torch.cat([o.to(torch.int64) + int(s) for o, s in zip(lS_o, starts[:-1])]
                        + [torch.tensor([int(starts[-1])])])
    return torch.cat([i.reshape(-1) for i in lS_i]).to(dtype), offsets.to(dtype)


def local_packed(lS_o, lS_i, local_tables, dtype=torch.int32):
    """Packed batch restricted to this rank's tables (what dlrm_wrap + distribute_batched_emb_data hand to a
    rank, dlrm_s_pytorch.py:147-148, 772-802)."""
    return pack([lS_o[k] for k in local_tables], [lS_i[k] for k in local_tables], dtype)


class CriteoBinDataset:
    """Device-side twin of the reference's CriteoBinDataset (data_loader_terabyte.py:205-252; same constructor
    arguments and item layout): item `idx` is batch `idx` of the binary file (records of 40 int32: label, 13 dense,
    26 categorical).  The raw bytes of the batch are read into a pinned staging buffer, copied to the device as they
    are (160 B per sample) and split / log-transformed / taken modulo max_ind_range / transposed by ONE kernel
    (csrc/criteo_bin.cu), so the tensors the model consumes are born on the GPU:
        (X [B,13] fp32, lS_o, lS_i, T [B,1] fp32)   with lS_o / lS_i int64 [26, B], or, with
        batched_or_fbgemm_emb, offsets int32 [26*B+1] / indices int32 [26*B].
    counts_file is optional here (the reference loads it and never uses it on this path)."""

    tar_fea, den_fea, spa_fea = 1, 13, 26

    def __init__(self, data_file, counts_file=None, batch_size=1, max_ind_range=-1, bytes_per_feature=4,
                 batched_or_fbgemm_emb=False, device="cuda"):
        import math
        import os
        if bytes_per_feature != 4:
            raise ValueError("CriteoBinDataset: the binary format has 4 bytes per feature")
        self.tad_fea = self.tar_fea + self.den_fea
        self.tot_fea = self.tad_fea + self.spa_fea
        self.batch_size = batch_size
        self.max_ind_range = max_ind_range
        self.bytes_per_entry = bytes_per_feature * self.tot_fea * batch_size
        self.num_entries = math.ceil(os.path.getsize(data_file) / self.bytes_per_entry)
        self.file = open(data_file, "rb")
        self.counts = None
        if counts_file is not None:
            with np.load(counts_file) as data:
                self.counts = data["counts"]
        self.m_den = 13
        self.batched_or_fbgemm_emb = batched_or_fbgemm_emb
        self.device = torch.device(device)
        self._stage = torch.empty((batch_size, self.tot_fea), dtype=torch.int32).pin_memory()

    def __len__(self):
        return self.num_entries

    def __getitem__(self, idx):
        from . import ops
        if idx < 0 or idx >= self.num_entries:
            raise IndexError(idx)
        self.file.seek(idx * self.bytes_per_entry, 0)
        buf = self._stage.numpy().reshape(-1).view(np.uint8)
        n = self.file.readinto(memoryview(buf))
        nrec = n // (4 * self.tot_fea)                       # the last batch of a file may be short
        rec = self._stage[:nrec].to(self.device, non_blocking=True)
        out = ops.criteo_bin_batch(rec, self.max_ind_range, batched=self.batched_or_fbgemm_emb)
        torch.cuda.current_stream().synchronize()            # the staging buffer is reused by the next item
        return out

    def __del__(self):
        f = getattr(self, "file", None)
        if f is not None:
            f.close()
