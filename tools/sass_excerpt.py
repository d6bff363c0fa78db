"""SASS evidence for profiles/: per kernel of the built objects, the counts of the instructions that show HOW it moves data and
computes (tcgen05 MMA = UTC*MMA, TMEM loads = LDTM, TMA = UTMALDG / UTMASTG / UBLKCP, mbarrier = SYNCS, legacy MMA = HMMA,
vector global accesses), plus the first occurrence of each as a sample line.

    python tools/sass_excerpt.py > profiles/rNN_sass_excerpt.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "build", "obj")
KEYS = ["UTCHMMA", "UTCQMMA", "UTCBAR", "UTCATOMSWS", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMACCTL", "UBLKCP", "SYNCS", "HMMA",
        "LDSM", "LDG.E.128", "LDG.E.64", "STG.E.128", "STG.E.64", "LDGSTS", "UCGABAR", "MEMBAR", "ATOMG", "RED.E", "REDG"]
WANT = re.compile(r"gemm_pair_kernel|wgrad_splitk|wgrad_reduce|embbag_fwd_kernel|bwd_update_kernel|interact_fwd_mma|interact_bwd_mma|"
                  r"radix_scatter|build_keys|a2a_push|head_bwd|relu_bwd_bias|interact_fwd_bulk|peer_allreduce_kernel")


def main():
    for obj in sorted(os.listdir(OBJ)):
        if not obj.endswith(".o"):
            continue
        out = subprocess.run(["cuobjdump", "-sass", os.path.join(OBJ, obj)], capture_output=True, text=True).stdout
        name, counts, first = None, None, None
        blocks = []
        for line in out.splitlines():
            m = re.search(r"Function : (\S+)", line)
            if m:
                if name:
                    blocks.append((name, counts, first))
                name, counts, first = m.group(1), collections.Counter(), {}
                continue
            if name is None:
                continue
            m = re.search(r"/\*[0-9a-f]{4}\*/\s+(.*?);", line)
            if not m:
                continue
            ins = m.group(1).strip()
            op = ins.split()[1] if ins.startswith("@") and len(ins.split()) > 1 else ins.split()[0]
            for k in KEYS:
                if op.startswith(k):
                    full = op if k in ("UTCHMMA", "UTMALDG", "UTCBAR", "SYNCS", "HMMA", "UBLKCP", "LDTM", "LDSM") else k
                    counts[full] += 1
                    first.setdefault(full, ins[:110])
        if name:
            blocks.append((name, counts, first))
        seen = collections.Counter()
        for name, counts, first in blocks:
            dem = subprocess.run(["cu++filt", name], capture_output=True, text=True).stdout.strip() or name
            if not WANT.search(dem) or not counts:
                continue
            base = dem.split("<")[0].split("(")[0]
            if ("embbag_fwd_kernel" in base or "bwd_update_kernel" in base) and "<(int)4, (int)32, (int)1," not in dem:
                continue                      # the instantiations of the 512-byte rows (D = 128) stand for their templates
            seen[base] += 1
            if seen[base] > 6:
                continue
            print("== %s :: %s" % (obj, dem[:150]))
            for k, v in sorted(counts.items(), key=lambda kv: -kv[1]):
                print("   %5d  %-40s %s" % (v, k, first[k]))
    return 0


if __name__ == "__main__":
    sys.exit(main())
