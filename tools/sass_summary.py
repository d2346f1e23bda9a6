#!/usr/bin/env python
"""Per-kernel counts of the SASS mnemonics that prove (or disprove) a Blackwell-native kernel
(/opt/skills/guides/B200_PROFILING.md "What proves a Blackwell-native kernel"), from `cuobjdump -sass` of the built
library.  Runs without a GPU.

    python tools/sass_summary.py > profiles/sass_summary.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "nano-vllm-v1_b200", "b200attn", "libb200attn.so")
COLS = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UBLKCP", "SYNCS", "HMMA", "LDGSTS", "LDSM", "MUFU.EX2", "LDGMC", "ST/RED.SYS",
        "LDG", "STG", "ATOM/RED", "BAR"]


def main():
    lib = sys.argv[1] if len(sys.argv) > 1 else LIB
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    demangle = {}
    counts = collections.OrderedDict()
    cur = None
    for ln in sass.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
        if not m:
            continue
        op = m.group(1)
        c = counts[cur]
        c["total"] += 1
        if op.startswith("UTCHMMA") or op.startswith("UTCQMMA") or op.startswith("UTCOMMA"):
            c["UTCHMMA"] += 1
        elif op.startswith("UTCBAR"):
            c["UTCBAR"] += 1
        elif op.startswith("LDTM"):
            c["LDTM"] += 1
        elif op.startswith("STTM"):
            c["STTM"] += 1
        elif op.startswith("UTMALDG"):
            c["UTMALDG"] += 1
        elif op.startswith("UBLKCP"):
            c["UBLKCP"] += 1
        elif op.startswith("SYNCS"):
            c["SYNCS"] += 1
        elif op.startswith("HMMA"):
            c["HMMA"] += 1
        elif op.startswith("LDGSTS"):
            c["LDGSTS"] += 1
        elif op.startswith("LDSM"):
            c["LDSM"] += 1
        elif op.startswith("MUFU.EX2"):
            c["MUFU.EX2"] += 1
        elif op.startswith("LDGMC"):
            c["LDGMC"] += 1
        elif (op.startswith("STG") or op.startswith("RED")) and op.endswith(".STRONG.SYS"):
            c["ST/RED.SYS"] += 1
        elif op.startswith("LDG"):
            c["LDG"] += 1
        elif op.startswith("STG"):
            c["STG"] += 1
        elif op.startswith("ATOM") or op.startswith("RED"):
            c["ATOM/RED"] += 1
        elif op.startswith("BAR"):
            c["BAR"] += 1
    names = list(counts)
    try:
        dem = subprocess.run(["c++filt"] + names, capture_output=True, text=True, check=True).stdout.splitlines()
        demangle = dict(zip(names, dem))
    except Exception:
        pass
    print(f"# cuobjdump -sass {os.path.relpath(lib, ROOT)}: instruction counts per kernel (static SASS, sm_100a)")
    print("# tcgen05.mma -> UTCHMMA, tcgen05.commit -> UTCBAR, tcgen05.ld/st -> LDTM/STTM, TMA -> UTMALDG, mbarrier -> SYNCS,")
    print("# mma.sync -> HMMA, cp.async -> LDGSTS, ldmatrix -> LDSM, multimem.ld_reduce -> LDGMC (the only distinct multimem mnemonic);")
    print("# multimem.st / multimem.red and peer stores are ordinary system-scope STG / RED (ST/RED.SYS) on a multicast or peer address")
    print(f"{'kernel':78s} " + " ".join(f"{c:>9s}" for c in COLS) + f" {'total':>7s}")
    for n in names:
        c = counts[n]
        short = re.sub(r"\(.*$", "", demangle.get(n, n)).replace("void b200::", "").replace("b200::", "")
        print(f"{short[:78]:78s} " + " ".join(f"{c[k]:9d}" for k in COLS) + f" {c['total']:7d}")


if __name__ == "__main__":
    main()
