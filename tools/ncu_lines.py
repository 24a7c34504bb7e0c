#!/usr/bin/env python
"""Per-source-line instruction counts / stall samples of one kernel launch from an .ncu-rep (needs -lineinfo)."""
import csv, subprocess, sys
rep = sys.argv[1]; launch = sys.argv[2] if len(sys.argv) > 2 else "0"; top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--launch-skip", launch,
                      "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
fname, hdr, rec = None, None, []
for r in rows:
    if not r: continue
    if r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; continue
    if hdr and r[0].isdigit():
        d = dict(zip(hdr[4:], r[4:]))
        I = lambda v: int(v) if str(v).lstrip("-").isdigit() else 0
        rec.append((fname, int(r[0]), I(d.get("Instructions Executed", 0)), I(d.get("# Samples", 0)), r[1].strip()[:100]))
ti = sum(x[2] for x in rec); ts = sum(x[3] for x in rec)
print("total warp-instructions %d, samples %d" % (ti, ts))
for x in sorted(rec, key=lambda x: -x[2])[:top]:
    print("%-18s %4d inst %9d (%.3f) samples %6d (%.3f) | %s" % (x[0], x[1], x[2], x[2] / max(ti, 1), x[3], x[3] / max(ts, 1), x[4]))
