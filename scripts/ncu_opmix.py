#!/usr/bin/env python
"""ncu_opmix.py <report.ncu-rep> — dynamic SASS opcode mix (warp instructions executed, stall samples)
of the first kernel in the report, and the hottest SASS addresses that were never or rarely executed
(static footprint vs dynamic footprint)."""
import collections, csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--launch-count", "1"],
                     capture_output=True, text=True).stdout
hdr = None
ops = collections.Counter(); samp = collections.Counter(); rows = []
for r in csv.reader(io.StringIO(raw)):
    if not r or r[0] == "Kernel Name": continue
    if r[0] == "Address":
        hdr = r; iS = hdr.index("# Samples"); iE = hdr.index("Instructions Executed"); continue
    if hdr is None: continue
    try: e = int(r[iE] or 0); s = int(r[iS] or 0)
    except ValueError: continue
    txt = r[1].strip()
    if txt.startswith("@"): txt = txt.split(None, 1)[1]
    op = txt.split()[0].rstrip(";")
    base = op.split(".")[0]
    ops[base] += e; samp[base] += s; rows.append((e, s, r[1].strip()))
tot = sum(ops.values()); ts = sum(samp.values())
print("warp instructions", tot, "samples", ts, "static instrs", len(rows))
for k, v in ops.most_common(30):
    print("%-10s %6.2f%% inst %6.2f%% samples" % (k, 100 * v / tot, 100 * samp[k] / max(1, ts)))
mx = max(e for e, _, _ in rows)
hot = [r for r in rows if r[0] > 0.02 * mx]; warm = [r for r in rows if r[0] > 0]
print("static: %d instrs; executed at all: %d; executed >2%% of max count: %d (covering %.1f%% of dynamic instrs)" %
      (len(rows), len(warm), len(hot), 100 * sum(e for e, _, _ in hot) / tot))
