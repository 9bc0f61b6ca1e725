"""Per-kernel registers / stack / spills from csrc/ptxas.log (nvcc -Xptxas -v)."""
import re, subprocess, sys
log = open(sys.argv[1] if len(sys.argv) > 1 else "mitsuba_customization_b200/csrc/ptxas.log").read()
rows = []
for m in re.finditer(r"Compiling entry function '(\S+)' for 'sm_100a'\n.*?Function properties for \S+\n\s+(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\n.*?Used (\d+) registers", log, re.S):
    name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
    name = re.sub(r"\(.*", "", name).replace("void merl::", "").replace("merl::", "")
    rows.append((name, int(m.group(5)), int(m.group(2)), int(m.group(3)), int(m.group(4))))
for r in sorted(rows):
    print(f"{r[0]:55s} regs {r[1]:3d}  stack {r[2]:4d}  spill st/ld {r[3]:3d}/{r[4]:3d}")
