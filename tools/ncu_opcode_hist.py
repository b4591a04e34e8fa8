"""Opcode histogram + stall summary from `ncu --page source --csv` output.
usage: python tools/ncu_opcode_hist.py file.csv [warp_slots]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
slots = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
hdr = rows[1]
si, ei, ss = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
stall_cols = [(j, c) for j, c in enumerate(hdr) if c.startswith("stall_") and "Not Issued" not in c]
agg, samp, tot, stot = collections.Counter(), collections.Counter(), 0, 0
stalls = collections.Counter()
for r in rows[2:]:
    try:
        n = int(r[ei]); s = int(r[ss])
    except (ValueError, IndexError):
        continue
    op = r[si].strip().split()
    if not op:
        continue
    o = op[1] if op[0].startswith("@") else op[0]
    o = ".".join(o.split(".")[:2]) if o.startswith(("LD", "ST", "IMAD", "SHFL")) else o.split(".")[0]
    agg[o] += n; samp[o] += s; tot += n; stot += s
    for j, c in stall_cols:
        try:
            stalls[c] += int(r[j])
        except ValueError:
            pass
print(f"total warp instructions {tot}, per warp-slot {tot / slots:.1f}; samples {stot}")
for k, v in agg.most_common(28):
    print(f"{k:14s} {v:12d} {v / tot * 100:5.1f}%  per-slot {v / slots:7.1f}  samples {samp[k] / max(stot, 1) * 100:5.1f}%")
print("stall samples:", ", ".join(f"{k[6:]}={v / max(stot, 1) * 100:.1f}%" for k, v in stalls.most_common(9)))
