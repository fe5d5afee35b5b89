"""Per-CUDA-source-line instruction counts and stall samples of a capture made with --import-source on:
    python profiles/ncu_lines.py X.ncu-rep [top N] [inst|samples]"""
import csv
import subprocess
import sys

out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
key = sys.argv[3] if len(sys.argv) > 3 else "inst"
cur_file, hdr, rows = "", None, []
for r in csv.reader(out.splitlines()):
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
    elif r[0] == "Line No":
        hdr = r
    elif hdr and r[0].strip().isdigit():
        i_s, i_i = hdr.index("# Samples"), hdr.index("Instructions Executed")
        samples = int(r[i_s]) if r[i_s].isdigit() else 0
        inst = int(r[i_i]) if r[i_i].isdigit() else 0
        rows.append((inst, samples, cur_file, int(r[0]), r[1].strip()[:100], r))
ti, ts = sum(x[0] for x in rows), sum(x[1] for x in rows)
print(f"total warp instructions {ti}, stall samples {ts}")
stall_idx = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
for inst, samples, f, ln, src, r in sorted(rows, key=lambda o: -(o[0] if key == "inst" else o[1]))[:top]:
    st = sorted(((int(r[i]), h[6:]) for i, h in stall_idx if r[i].isdigit() and int(r[i]) > 0), reverse=True)[:2]
    print(f"{inst:9d} {100.0 * inst / max(1, ti):5.1f}% inst  {samples:6d} {100.0 * samples / max(1, ts):5.1f}% smp  {f}:{ln:<4d} {src:100s} {st}")
