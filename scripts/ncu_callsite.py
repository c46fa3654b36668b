#!/usr/bin/env python
"""ncu_callsite.py <sass csv from ncu --page source --print-source sass> <nvdisasm -gi listing of the kernel>
Dynamic instructions / stall samples per inlined call site: every SASS instruction is attributed to the
chain of source frames nvdisasm prints for it; the table groups by the outermost DEPTH frames."""
import collections, csv, re, sys
csv_path, sass_path = sys.argv[1], sys.argv[2]
depth = int(sys.argv[3]) if len(sys.argv) > 3 else 3
rows = []
hdr = None
for r in csv.reader(open(csv_path)):
    if not r or r[0] == "Kernel Name": continue
    if r[0] == "Address":
        if hdr is not None: break          # second launch
        hdr = r; iS = hdr.index("# Samples"); iE = hdr.index("Instructions Executed"); continue
    try: rows.append((int(r[iE] or 0), int(r[iS] or 0)))
    except (ValueError, IndexError): pass
frames, chain, per_instr = [], [], []
for l in open(sass_path):
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)( inlined at "([^"]+)", line (\d+))?', l)
    if m:
        f = (m.group(1).split('/')[-1].replace('.cuh', ''), int(m.group(2)))
        if not chain or not chain[-1][1]: chain = [[f, bool(m.group(3))]]     # new chain starts after a chain that ended
        else: chain.append([f, bool(m.group(3))])
        continue
    if re.match(r'\s+/\*[0-9a-f]{4,6}\*/\s+\S', l):
        per_instr.append(tuple(c[0] for c in chain)); 
        # keep chain for following instrs until a new File line arrives
        if chain: chain[-1][1] = False
n = min(len(rows), len(per_instr))
agg = collections.Counter(); sagg = collections.Counter()
for (e, s), ch in zip(rows[:n], per_instr[:n]):
    key = tuple(reversed(ch))[:depth]      # outermost first
    agg[key] += e; sagg[key] += s
tot = sum(agg.values()); ts = sum(sagg.values())
print("instrs", n, "dynamic", tot, "samples", ts)
for k, v in agg.most_common(60):
    print("%6.2f%% inst %6.2f%% samp  %s" % (100 * v / tot, 100 * sagg[k] / max(1, ts), " > ".join("%s:%d" % f for f in k)))
