#!/usr/bin/env python
"""Per-instruction warp-stall samples of one kernel from an .ncu-rep captured with --import-source on:
overall stall mix and the N hottest SASS instructions.   python tools/ncu_source_hot.py rep.ncu-rep [N]"""
import csv
import io
import subprocess
import sys

rep, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 30
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
print("# source:", rep, "(ncu --set full --import-source on; warp-state samples per SASS instruction)")
print("# kernel:", rows[0][1] if rows and len(rows[0]) > 1 else "?")
hdr = rows[1]
idx = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
recs = []
for r in rows[2:]:
    try:
        n = int(r[idx["# Samples"]])
    except (ValueError, IndexError):
        continue
    recs.append((n, r))
total = sum(n for n, _ in recs)
agg = {s: sum(int(r[idx[s]] or 0) for _, r in recs) for s in stalls}
print(f"total samples {total}")
print("stall mix:", ", ".join(f"{k[6:]} {100 * v / total:.1f}%" for k, v in sorted(agg.items(), key=lambda x: -x[1])[:8]))
print(f"{'samples':>8} {'share':>6}  {'executed':>9}  instruction   [top stall reasons]")
for n, r in sorted(recs, key=lambda x: -x[0])[:top]:
    st = sorted(((s[6:], int(r[idx[s]] or 0)) for s in stalls), key=lambda x: -x[1])[:2]
    print(f"{n:8d} {100 * n / total:5.1f}%  {r[idx['Instructions Executed']]:>9}  {r[idx['Source']][:64]:64s} {st}")
