#!/usr/bin/env python3
"""Per-source-line summary of an ncu report's source page (cuda,sass view): instruction share, SIMT efficiency, stall samples.
usage: ncu -i X.ncu-rep --page source --print-source cuda,sass --csv > src.csv ; python tools/ncu_lines.py src.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
N = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur = None; hdr = None; out = []
for r in rows:
    if len(r) == 2 and r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if len(r) > 5 and r[0] == "Line No": hdr = r; continue
    if hdr and len(r) == len(hdr) and r[0] not in ("", "Line No"):
        d = dict(zip(hdr[4:], r[4:]))
        try:
            out.append((cur, int(r[0]), r[1].strip()[:70], int(d["# Samples"]), int(d["Instructions Executed"]), int(d["Thread Instructions Executed"])))
        except ValueError:
            pass
ti = sum(o[4] for o in out); ts = sum(o[3] for o in out); tt = sum(o[5] for o in out)
print(f"total inst {ti:.3e} thread-inst {tt:.3e} avg lanes {tt/ti:.2f} samples {ts}")
byfile = {}
for o in out:
    f = byfile.setdefault(o[0], [0, 0, 0]); f[0] += o[3]; f[1] += o[4]; f[2] += o[5]
for k, v in sorted(byfile.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:22s} inst {100*v[1]/ti:5.1f}%  samples {100*v[0]/ts:5.1f}%  lanes {v[2]/max(v[1],1):5.2f}")
print()
for o in sorted(out, key=lambda o: -o[4])[:N]:
    print(f"{o[0]:18s}:{o[1]:4d} inst {100*o[4]/ti:5.2f}% smp {100*o[3]/ts:5.2f}% lanes {o[5]/max(o[4],1):5.2f} | {o[2]}")
