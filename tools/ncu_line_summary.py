#!/usr/bin/env python
"""Per-source-line totals from `ncu -i X.ncu-rep --page source --csv --print-source cuda,sass`."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
inst = collections.Counter(); samp = collections.Counter(); text = {}
fname = "?"
hdr = None
for r in rows:
    if r and r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if r and r[0] == "Line No": hdr = {h: i for i, h in enumerate(r)}; continue
    if hdr is None or len(r) < 10: continue
    try:
        key = (fname, int(r[0]))
        inst[key] += float(r[hdr["Instructions Executed"]] or 0); samp[key] += float(r[hdr["# Samples"]] or 0)
        text[key] = r[1].strip()
    except Exception:
        pass
ti, ts = sum(inst.values()), sum(samp.values())
print("total warp inst %.4g, samples %d" % (ti, ts))
for key, n in inst.most_common(top):
    print("%5.2f%% inst %5.2f%% samp  %s:%d  %s" % (100 * n / ti, 100 * samp[key] / max(ts, 1), key[0], key[1], text[key][:90]))
