#!/usr/bin/env python
"""Per-kernel SASS evidence of the Blackwell-native path: counts of the tcgen05 / TMEM / TMA mnemonics
(UTCHMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTMALDG / UTMASTG = cp.async.bulk.tensor, UTCBAR =
tcgen05.commit, HMMA = legacy mma.sync) in every kernel of libuva_head.so, plus a short excerpt around the first
UTCHMMA of each tensor-core kernel.  Runs without a GPU:  python scripts/sass_summary.py > profiles/r02_sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "unified_video_action_b200", "libuva_head.so")
KEYS = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTCCP", "HMMA", "SYNCS", "FFMA", "MUFU"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    filt = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True, text=True).stdout.split("\n")
    names = iter(filt)
    cur, counts, lines = None, collections.OrderedDict(), {}
    for ln in sass.split("\n"):
        m = re.search(r"Function : (\S+)", ln)
        if m:
            cur = next(names)
            counts[cur] = collections.Counter()
            lines[cur] = []
            continue
        if cur is None:
            continue
        m = re.search(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
        if m:
            op = m.group(1)
            counts[cur]["_n"] += 1
            for k in KEYS:
                if op.startswith(k):
                    counts[cur][k] += 1
            lines[cur].append(ln.rstrip())
    print("# SASS mnemonic counts per kernel of", os.path.relpath(LIB, ROOT), "(cuobjdump -sass, sm_100a)")
    print("# " + " ".join(f"{k:>8s}" for k in ["instrs"] + KEYS) + "  kernel")
    tot = collections.Counter()
    for name, c in counts.items():
        tot.update(c)
        short = re.sub(r"\(.*", "", name)[:150]
        print("  " + " ".join(f"{c[k]:8d}" for k in ["_n"] + KEYS) + "  " + short)
    print("  " + " ".join(f"{tot[k]:8d}" for k in ["_n"] + KEYS) + "  TOTAL")
    print()
    shown = 0
    for name, ls in lines.items():
        idx = [i for i, l in enumerate(ls) if "UTCHMMA" in l]
        if not idx or shown >= 6:
            continue
        shown += 1
        print("## excerpt:", re.sub(r"\(.*", "", name)[:150])
        for l in ls[max(0, idx[0] - 6): idx[0] + 8]:
            print("   ", re.sub(r"\s+/\* 0x[0-9a-f]+ \*/", "", l))
        print()


if __name__ == "__main__":
    sys.exit(main())
