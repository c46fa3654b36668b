#!/usr/bin/env python
"""ncu_lines.py <report.ncu-rep> [N] — stall samples and executed instructions per CUDA source line."""
import csv, collections, subprocess, sys, io
rep = sys.argv[1]; N = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--launch-count", "1"], capture_output=True, text=True).stdout
cur_file = cur_line = None; hdr = None
agg = collections.Counter(); inst = collections.Counter(); thr = collections.Counter(); src = {}
stall_cols = {}
st = collections.defaultdict(collections.Counter)
for r in csv.reader(io.StringIO(raw)):
    if not r: continue
    if r[0] == "File Path": cur_file = r[1].split('/')[-1]; continue
    if r[0] == "Line No":
        hdr = r; ixS = hdr.index('# Samples'); ixE = hdr.index('Instructions Executed'); ixT = hdr.index('Thread Instructions Executed')
        stall_cols = {i: h for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h}
        continue
    if hdr is None or r[0] in ("Kernel Name",): continue
    if r[0] != "": cur_line = r[0]; src[(cur_file, cur_line)] = r[1]
    if len(r) > ixT and r[2] != "":
        try:
            k = (cur_file, cur_line)
            agg[k] += int(r[ixS] or 0); inst[k] += int(r[ixE] or 0); thr[k] += int(r[ixT] or 0)
            for i, h in stall_cols.items():
                if r[i]: st[k][h] += int(r[i])
        except ValueError: pass
tot = sum(agg.values()); ti = sum(inst.values())
byfile = collections.Counter()
for (f, l), s in agg.items(): byfile[f] += s
print("total samples", tot, "warp inst", ti, "avg threads/inst %.1f" % (sum(thr.values()) / max(1, ti)))
print({f: "%.1f%%" % (100 * s / tot) for f, s in byfile.most_common(8)})
for k, s in agg.most_common(N):
    top = ",".join("%s:%d%%" % (h.replace('stall_', ''), 100 * v / max(1, s)) for h, v in st[k].most_common(3))
    print("%5.2f%% inst %5.2f%% thr/inst %4.1f %s:%s [%s] %s" % (100 * s / tot, 100 * inst[k] / ti, thr[k] / max(1, inst[k]), k[0], k[1], top, src[k].strip()[:90]))
