#!/usr/bin/env python
"""ncu_source_top.py rep [n] -- instruction mix and the hottest SASS instructions (stall samples) of an .ncu-rep"""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]; n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
h = rows[1]; ix = {k: i for i, k in enumerate(h)}
body = [r for r in rows[2:] if len(r) == len(h)]
mix = collections.Counter(); tot = 0; samp = collections.Counter(); tags = collections.Counter()
for r in body:
    op = r[ix['Source']].split()[0] if not r[ix['Source']].strip().startswith('@') else r[ix['Source']].split()[1]
    e = int(r[ix['Instructions Executed']]); mix[op.split('.')[0]] += e; tot += e
    samp[op.split('.')[0]] += int(r[ix['# Samples']])
    tags[op.split('.')[0]] += int(r[ix['L1 Tag Requests Global']])
print("total warp instructions", tot, " rows", len(body))
for op, e in mix.most_common(24): print(f"  {op:12s} {e:12d} {100*e/tot:5.1f}%  samples {samp[op]:7d}  l1tags {tags[op]}")
print("hottest instructions by samples:")
for r in sorted(body, key=lambda r: -int(r[ix['# Samples']]))[:n]:
    st = {k: int(r[ix[k]]) for k in h if k.startswith('stall_') and r[ix[k]].isdigit() and int(r[ix[k]]) > 0}
    top = sorted(st.items(), key=lambda kv: -kv[1])[:3]
    print(f"  {r[ix['# Samples']]:>6s} {r[ix['Instructions Executed']]:>9s}  {r[ix['Source']].strip()[:70]:70s} {top}")
