"""Key metrics + hottest SASS lines of every kernel in an ncu report: python tools/ncu_multi.py rep [nlines]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; nl = int(sys.argv[2]) if len(sys.argv) > 2 else 8
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw))); hdr = rows[0]
keep = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__registers_per_thread",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio"]
for k in keep:
    if k in hdr:
        i = hdr.index(k); print(k[:58].ljust(60), [r[i][:26] for r in rows[2:]], rows[1][i])
for r in rows[2:]:
    st = [(float(r[i]), h.split("issue_stalled_")[1].replace("_per_issue_active.ratio", "")) for i, h in enumerate(hdr)
          if "issue_stalled" in h and h.endswith("per_issue_active.ratio")]
    print([f"{n} {v:.1f}" for v, n in sorted(st, reverse=True)[:6]])
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks = src.split('"Kernel Name"')
rows = list(csv.reader(io.StringIO(src)))
# the source page concatenates kernels; split on header rows
cur = None; h2 = None; data = []
def flush():
    if not data: return
    iS, iE, iN = h2.index("Source"), h2.index("Instructions Executed"), h2.index("# Samples")
    tot = sum(int(r[iN]) for r in data if r[iN].isdigit()) or 1
    print(f"--- {cur}: {tot} samples")
    for r in sorted(data, key=lambda r: -int(r[iN]) if r[iN].isdigit() else 0)[:nl]:
        print(f"  {100 * int(r[iN]) / tot:5.1f} %  exec {r[iE]:>9s}  {r[iS].strip()[:110]}")
for r in rows:
    if "Source" in r and "# Samples" in r:
        flush(); h2 = r; data = []; continue
    if h2 is None or len(r) != len(h2):
        if len(r) >= 1 and h2 is None: cur = ",".join(r)[:80]
        elif len(r) >= 1 and len(r) != len(h2 or []): flush(); data = []; cur = ",".join(r)[:80]; 
        continue
    data.append(r)
flush()
