"""Aggregates a bench.py --timeline file by kernel name: calls, total and mean duration, share of the step span."""
import collections
import re
import sys

rows = collections.OrderedDict()
span = None
for line in open(sys.argv[1]):
    if line.startswith("#"):
        m = re.search(r"span ([0-9.]+) us", line)
        if m:
            span = float(m.group(1))
        continue
    parts = line.split(None, 3)
    if len(parts) < 4:
        continue
    dur, name = float(parts[1]), parts[3].strip()
    name = re.sub(r"\(.*", "", name).replace("void ", "").replace("hstd::", "")
    c = rows.setdefault(name, [0, 0.0])
    c[0] += 1
    c[1] += dur
tot = sum(v[1] for v in rows.values())
print(f"step span {span} us; sum of kernel durations {tot:.1f} us (streams overlap)")
for name, (n, t) in sorted(rows.items(), key=lambda kv: -kv[1][1]):
    print(f"{t:9.1f} us {100*t/tot:5.1f}%  n={n:3d} avg={t/n:7.1f}  {name[:90]}")
