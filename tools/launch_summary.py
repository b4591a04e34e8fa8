"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list: which kernel holds which SHARE
of the bench step (per-launch times under ncu are cold-cache and serialised; the shares are what must agree with
bench.py's own CUDA-event stage times).  usage: python tools/launch_summary.py launches.csv [skip_first_n_launches]"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if r and not r[0].startswith("==")]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
tot, cnt = collections.Counter(), collections.Counter()
for r in rows[1 + skip:]:
    v = float(r[vi].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r[ui], 1.0)
    name = r[ki].split("(")[0]
    tot[name] += v; cnt[name] += 1
all_us = sum(tot.values())
print(f"{sum(cnt.values())} launches, {all_us / 1e3:.3f} ms in kernels")
for k, v in tot.most_common():
    print(f"{v / all_us * 100:6.2f} %  {v / 1e3:9.3f} ms  {cnt[k]:4d} launches  {v / cnt[k]:9.1f} us each  {k}")
