#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name.
Usage: python scripts/ncu_launch_table.py <launches.csv> <out.md> "<title / command>" """
import collections
import csv
import re
import sys

src, out, title = sys.argv[1], sys.argv[2], sys.argv[3]
lines = open(src).read().splitlines()
start = [i for i, l in enumerate(lines) if l.startswith('"ID"')][0]
rows = list(csv.DictReader(lines[start:]))
acc = collections.OrderedDict()
for r in rows:
    n = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "")[-70:]
    a = acc.setdefault(n, [0, 0.0])
    a[0] += 1
    a[1] += float(r["Metric Value"])
tot = sum(v[1] for v in acc.values())
with open(out, "w") as f:
    f.write(f"# {title}\n\n(per-launch times under ncu are cold-cache and serialised: compare shares, not absolutes)\n\n")
    f.write("| kernel | launches | total ms | ms / launch | share |\n|---|---|---|---|---|\n")
    for n, (c, t) in sorted(acc.items(), key=lambda kv: -kv[1][1])[:40]:
        f.write(f"| `{n}` | {c} | {t / 1e6:.3f} | {t / 1e6 / c:.4f} | {100 * t / tot:.1f}% |\n")
print(open(out).read())
