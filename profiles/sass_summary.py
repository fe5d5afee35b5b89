"""Register / spill table from the build's ptxas log and an instruction-class summary of the SASS of every kernel in
libgsdf_b200.so (no GPU needed):  python profiles/sass_summary.py > profiles/r2_ptxas_sass_summary.txt"""
import collections
import re
import subprocess
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
log = (ROOT / "gradient-sdf_b200" / "csrc" / "ptxas.log").read_text().splitlines()
print("## ptxas -v (gradient-sdf_b200/csrc/ptxas.log): registers, spills, static shared memory\n")
name = None
for i, l in enumerate(log):
    m = re.search(r"Compiling entry function '(\S+)'", l)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
    m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", l)
    if m and name:
        stack, st, ld = m.groups()
    m = re.search(r"Used (\d+) registers.*?(?:, (\d+) bytes smem)?$", l)
    if m and name:
        smem = re.search(r"(\d+) bytes smem", l)
        print(f"{name[:70]:70s} regs {m.group(1):>3s}  stack {stack:>4s} B  spill st/ld {st:>3s}/{ld:>3s} B  smem {smem.group(1) if smem else '0':>5s} B")
        name = None
sass = subprocess.run(["cuobjdump", "-sass", str(ROOT / "gradient-sdf_b200" / "libgsdf_b200.so")], capture_output=True, text=True).stdout
classes = [("REDG (fire-and-forget reductions)", r"\bREDG"), ("ATOMG (atomics with return)", r"\bATOMG"), ("ATOMS", r"\bATOMS"),
           ("LDG", r"\bLDG"), ("STG", r"\bSTG"), ("LDL+STL (local memory)", r"\b(LDL|STL)"), ("LDS+STS", r"\b(LDS|STS)"),
           ("SHFL", r"\bSHFL"), ("MATCH", r"\bMATCH"), ("VOTE", r"\bVOTE"), ("BAR.SYNC", r"\bBAR"), ("MEMBAR/CCTL", r"\b(MEMBAR|CCTL)"),
           ("DADD/DMUL/DFMA (fp64)", r"\bD(ADD|MUL|FMA)"), ("FFMA", r"\bFFMA"), ("FADD+FMUL", r"\bF(ADD|MUL)\b"), ("MUFU", r"\bMUFU"),
           ("UTMALDG/UBLKCP (TMA)", r"\b(UTMALDG|UTMASTG|UBLKCP)"), ("UTC*MMA / HMMA (tensor)", r"\b(UTC\w*MMA|HMMA|DMMA)")]
print("\n## SASS instruction classes per kernel (cuobjdump -sass libgsdf_b200.so; static counts)\n")
cur, counts, total = None, collections.OrderedDict(), {}
for l in sass.splitlines():
    m = re.search(r"Function : (\S+)", l)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        counts[cur] = collections.Counter(); total[cur] = 0
        continue
    if cur and re.match(r"\s+/\*[0-9a-f]{4,5}\*/", l):
        total[cur] += 1
        for nm, pat in classes:
            if re.search(pat, l):
                counts[cur][nm] += 1
for k, c in counts.items():
    print(f"{k[:90]}: {total[k]} instructions")
    print("    " + ", ".join(f"{nm} {c[nm]}" for nm, _ in classes if c[nm]))
