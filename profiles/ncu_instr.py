"""Per-CUDA-source-line executed instructions: python profiles/ncu_instr.py src.csv [N]   (csv from
ncu -i X.ncu-rep --page source --csv --print-source cuda,sass)"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
cur, hdr, out = "", None, []
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif r[0] == "Line No":
        hdr = r
        ii, ti = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
    elif hdr and r[0].strip().isdigit():
        ie = int(r[ii]) if r[ii].isdigit() else 0
        te = int(r[ti]) if r[ti].isdigit() else 0
        out.append((ie, te, cur, int(r[0]), r[1].strip()[:100]))
tot = sum(o[0] for o in out)
print("total warp-instr", tot)
for ie, te, f, ln, src in sorted(out, reverse=True)[: int(sys.argv[2]) if len(sys.argv) > 2 else 25]:
    print(f"{ie:9d} {100 * ie / max(1, tot):5.1f}% thr/inst {te / max(1, ie):5.1f}  {f}:{ln:<4d} {src}")
