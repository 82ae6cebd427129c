#!/usr/bin/env python
"""Per-instruction stall samples of an ncu report's source page, grouped by the warp-specialised
regions of a kernel (split at USETMAXREG).  usage: ncu_hot.py report.ncu-rep [top]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                     stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
ci, cs, ce = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
stall_cols = [(i, x) for i, x in enumerate(h) if x.startswith("stall_")]
data = []
for r in rows[hi + 1:]:
    try:
        data.append((int(r[cs]), int(r[ce]), r[ci].strip(), r))
    except (ValueError, IndexError):
        pass
tot = sum(d[0] for d in data)
print("total samples", tot, "warp instructions", sum(d[1] for d in data))
idx = [i for i, d in enumerate(data) if "USETMAXREG" in d[2]]
pts = [0] + idx + [len(data)]
for a, b in zip(pts[:-1], pts[1:]):
    print("region %5d..%5d samples %7d (%.1f%%) warp-instr %d" % (a, b, sum(d[0] for d in data[a:b]),
          100.0 * sum(d[0] for d in data[a:b]) / max(tot, 1), sum(d[1] for d in data[a:b])))
for i, d in sorted(sorted(enumerate(data), key=lambda t: -t[1][0])[:top_n]):
    st = sorted(((int(d[3][c] or 0), n) for c, n in stall_cols), reverse=True)[:2]
    print("%5d %6d %9d  %-70s %s" % (i, d[0], d[1], d[2][:70], " ".join("%s=%d" % (n[6:], v) for v, n in st if v)))

# per-region totals by stall reason (regions split further at big gaps is left to the reader)
if len(sys.argv) > 3:
    bounds = [int(x) for x in sys.argv[3].split(",")]
    for a, b in zip(bounds[:-1], bounds[1:]):
        agg = {}
        for d in data[a:b]:
            for c, n in stall_cols:
                if "Not Issued" in n:
                    continue
                v = int(d[3][c] or 0)
                if v:
                    agg[n[6:]] = agg.get(n[6:], 0) + v
        s = sum(d[0] for d in data[a:b])
        print("rows %d..%d samples %d instr %d:" % (a, b, s, sum(d[1] for d in data[a:b])),
              " ".join("%s=%d" % kv for kv in sorted(agg.items(), key=lambda t: -t[1])[:8]))
