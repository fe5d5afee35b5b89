"""Summarise `ncu -i X.ncu-rep --page source --csv` (stdin or file): stall totals and the hottest SASS lines."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1]) if len(sys.argv) > 1 else sys.stdin))
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
idx = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[hi + 1:] if len(r) == len(hdr) and r[0].startswith("0x")]
tot = sum(int(r[idx["# Samples"]]) for r in data)
print("kernel:", rows[0][1][:100] if rows[0] else "?")
print("total samples", tot, "SASS instrs", len(data))
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
agg = {s: sum(int(r[idx[s]]) for r in data) for s in stalls}
print("stalls:", [(k, v) for k, v in sorted(agg.items(), key=lambda x: -x[1])[:8]])
for r in sorted(data, key=lambda r: -int(r[idx["# Samples"]]))[:topn]:
    st = {s: int(r[idx[s]]) for s in stalls if int(r[idx[s]]) > 0}
    print(r[idx["# Samples"]].rjust(6), r[idx["Source"]].strip()[:72].ljust(72), r[idx["Instructions Executed"]].rjust(8),
          sorted(st.items(), key=lambda x: -x[1])[:2])
