"""Summarise an `ncu --page source --csv` dump: total stall samples by reason and the hottest
SASS lines.  python tools/ncu_stalls.py file.csv [N]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
data = [r for r in rows[hi + 1:] if len(r) >= len(hdr) and r[0].startswith("0x")]
ci = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = {s: 0 for s in stalls}
lines = []
for k, d in enumerate(data):
    n = int(d[ci["# Samples"]] or 0)
    lines.append((n, k, d))
    for s in stalls:
        tot[s] += int(d[ci[s]] or 0)
print("total samples", sum(n for n, _, _ in lines))
print({k: v for k, v in sorted(tot.items(), key=lambda kv: -kv[1]) if v})
lines.sort(key=lambda x: -x[0])
for n, k, d in lines[:topn]:
    st = {s: int(d[ci[s]]) for s in stalls if int(d[ci[s]] or 0) > 0}
    top = sorted(st.items(), key=lambda kv: -kv[1])[:3]
    print(n, k, d[ci["Source"]].strip()[:80], top)
