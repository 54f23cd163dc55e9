#!/usr/bin/env python
"""Summarise an ncu SASS source page (ncu -i X.ncu-rep --page source --csv --print-source sass):
instruction mix by opcode, stall reasons, and the hottest SASS instructions."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hdr_i]
col = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[hdr_i + 1:] if len(r) == len(hdr)]
def num(r, k):
    try: return float(r[col[k]])
    except Exception: return 0.0
tot_inst = sum(num(r, "Instructions Executed") for r in data)
tot_samp = sum(num(r, "# Samples") for r in data)
print("SASS lines %d, warp instructions executed %.4g, samples %d" % (len(data), tot_inst, tot_samp))
ops = collections.Counter(); opsamp = collections.Counter()
for r in data:
    src = r[col["Source"]].strip()
    toks = src.split()
    op = toks[1] if toks and toks[0].startswith("@") and len(toks) > 1 else (toks[0] if toks else "?")
    op = op.split(".")[0]
    ops[op] += num(r, "Instructions Executed"); opsamp[op] += num(r, "# Samples")
print("\nopcode           inst%%   samples%%")
for op, n in ops.most_common(28):
    print("%-14s %6.2f   %6.2f" % (op, 100 * n / tot_inst, 100 * opsamp[op] / max(tot_samp, 1)))
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
print("\nstall reason (all samples)")
for s, v in sorted(((s, sum(num(r, s) for r in data)) for s in stalls), key=lambda x: -x[1])[:10]:
    print("%-26s %6.2f%%" % (s, 100 * v / max(tot_samp, 1)))
thr = sum(num(r, "Thread Instructions Executed") for r in data)
print("\navg active threads per warp instruction: %.2f" % (thr / tot_inst))
if len(sys.argv) > 2:
    print("\nhottest instructions by samples")
    for r in sorted(data, key=lambda r: -num(r, "# Samples"))[:int(sys.argv[2])]:
        print("%6d %10.3g  %s" % (num(r, "# Samples"), num(r, "Instructions Executed"), r[col["Source"]].strip()[:100]))
