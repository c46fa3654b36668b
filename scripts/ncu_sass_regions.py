#!/usr/bin/env python
"""ncu_sass_regions.py <report.ncu-rep> [bucket_bytes] — executed warp instructions and stall samples per SASS address bucket,
with the dominant source line of each bucket (so the state machine's regions can be costed)."""
import csv, collections, subprocess, sys, io
rep = sys.argv[1]; B = int(sys.argv[2], 0) if len(sys.argv) > 2 else 0x400
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--launch-count", "1"], capture_output=True, text=True).stdout
hdr = None; rows = []
for r in csv.reader(io.StringIO(raw)):
    if not r: continue
    if r[0] == "Address":
        hdr = r; continue
    if hdr is None: continue
    rows.append(r)
ixA = hdr.index('Address'); ixS = hdr.index('# Samples'); ixE = hdr.index('Instructions Executed'); ixT = hdr.index('Thread Instructions Executed'); ixSrc = hdr.index('Source')
base = None
agg = collections.defaultdict(lambda: [0, 0, 0, 0])
for r in rows:
    try:
        a = int(r[ixA], 0)
    except ValueError:
        continue
    if base is None: base = a
    off = a - base
    b = off // B
    g = agg[b]
    g[0] += int(r[ixE] or 0); g[1] += int(r[ixS] or 0); g[2] += int(r[ixT] or 0); g[3] += 1
ti = sum(g[0] for g in agg.values()); ts = sum(g[1] for g in agg.values())
print("total warp inst %d samples %d" % (ti, ts))
for b in sorted(agg):
    g = agg[b]
    print("0x%05x-0x%05x inst %5.2f%% samples %5.2f%% thr/inst %4.1f  exec/static %.0f" % (b * B, (b + 1) * B, 100 * g[0] / ti, 100 * g[1] / max(1, ts), g[2] / max(1, g[0]), g[0] / max(1, g[3])))
