"""Summarises an ncu launch list (`--metrics gpu__time_duration.sum --csv`) of bench.py: kernels of ONE training step
(between two consecutive embbag_fwd launches, last full step), grouped by name, with each group's share of the step.
    python tools/launch_summary.py gpurun_out/launches.csv > profiles/rNN_launches_summary.txt"""
import collections
import csv
import re
import sys

lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
rows = []
for x in csv.DictReader(lines):
    if x.get("Metric Name") == "gpu__time_duration.sum":
        v = float(x["Metric Value"].replace(",", ""))
        v = v / 1000 if x["Metric Unit"] == "ns" else (v * 1000 if x["Metric Unit"] == "ms" else v)
        rows.append((x["Kernel Name"], v))
idx = [i for i, (n, _) in enumerate(rows) if "embbag_fwd" in n]
a, b = idx[-2], idx[-1]
step = rows[a:b]
tot = sum(v for _, v in step)
print("launches in one step: %d   sum of kernel durations: %.1f us (cold-cache, serialised under ncu)" % (len(step), tot))
agg = collections.OrderedDict()
for n, v in step:
    n = re.sub(r"\(.*", "", re.sub(r"<.*", "", n))
    n = n.replace("void ", "").replace("dlrmx::", "dlrmx::")[:64]
    c = agg.setdefault(n, [0, 0.0])
    c[0] += 1
    c[1] += v
print("%-66s %5s %9s %7s" % ("kernel", "calls", "us", "share"))
for n, (c, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-66s %5d %9.1f %6.1f%%" % (n, c, v, 100 * v / tot))
lib = lambda n: "nvjet" in n or "cublas" in n.lower() or "cutlass" in n.lower()          # noqa: E731  library GEMMs
other = lambda n: "at::" in n or "nccl" in n.lower() or "c10d" in n   # noqa: E731  torch / NCCL
ours = sum(v for n, v in step if not lib(n) and not other(n))
print("this repo's kernels (libdlrmx: dlrmx:: and dlrmx::tc::): %.1f us (%.1f%% of the step); library GEMMs (nvjet / cuBLASLt): "
      "%.1f us; other torch kernels (fills, copies): %.1f us"
      % (ours, 100 * ours / tot, sum(v for n, v in step if lib(n)), sum(v for n, v in step if other(n) and not lib(n))))
