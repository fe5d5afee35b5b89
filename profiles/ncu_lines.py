"""Per-CUDA-source-line stall samples: ncu -i X.ncu-rep --page source --csv --print-source cuda,sass | python profiles/ncu_lines.py"""
import csv
import sys

rows = list(csv.reader(sys.stdin))
cur_file = ""
out = []
hdr = None
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
    elif r[0] == "Line No":
        hdr = r
    elif hdr and r[0].strip().isdigit():
        d = dict(zip(hdr, r))
        # the header has two "Source" columns; csv->dict keeps the last, use positional for the first
        out.append((int(r[6]) if r[6].isdigit() else 0, cur_file, int(r[0]), r[1].strip()[:110], r))
tot = sum(o[0] for o in out)
print("total samples", tot)
stall_idx = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
for n, f, ln, src, r in sorted(out, key=lambda o: -o[0])[: int(sys.argv[1]) if len(sys.argv) > 1 else 30]:
    st = sorted(((int(r[i]), h[6:]) for i, h in stall_idx if r[i].isdigit() and int(r[i]) > 0), reverse=True)[:2]
    print(f"{n:6d} {100.0 * n / max(1, tot):5.1f}%  {f}:{ln:<4d} {src:110s} {st}")
