"""Markdown table from an `ncu --metrics gpu__time_duration.sum --csv` launch list.  python tools/launch_table.py file.csv"""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1], errors="replace")))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
hdr = rows[hi]
ci = {h: i for i, h in enumerate(hdr)}
agg = collections.OrderedDict()
for r in rows[hi + 1:]:
    if len(r) < len(hdr) or r[ci["Metric Name"]] != "gpu__time_duration.sum":
        continue
    name = re.sub(r"^(void )?(b200::)?((\(anonymous namespace\)|unnamed>)::)?", "", r[ci["Kernel Name"]])
    name = re.sub(r"\(.*$", "", name)
    key = (name, r[ci["Grid Size"]], r[ci["Block Size"]])
    v = float(r[ci["Metric Value"]].replace(",", ""))
    unit = r[ci["Metric Unit"]]
    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1.0)      # -> us
    a = agg.setdefault(key, [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
print("| kernel | grid | block | launches | total ms | share | avg us |\n|---|---|---|---|---|---|---|")
for (name, g, b), (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"| `{name}` | {g} | {b} | {n} | {us / 1e3:.2f} | {100 * us / tot:.1f}% | {us / n:.1f} |")
