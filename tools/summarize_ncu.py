"""Summarise an ncu report (--set full) into profiles/: key metrics, stall mix, hottest SASS lines,
and traffic.json (dram bytes per launch, read by bench.py's roofline.traffic)."""
import csv, io, json, subprocess, sys, collections

rep, tag = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
m = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
keep = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__cycles_elapsed.avg.per_second"]
out = [f"# ncu summary: {tag}", "", f"source report: `{rep}` (ncu --set full --clock-control none; per-launch, cold-cache, serialised)", "",
       "| metric | value | unit |", "|---|---|---|"]
for k in keep:
    if k in m:
        out.append(f"| {k} | {m[k][0]} | {m[k][1]} |")
out += ["", "## warp stall mix (cycles per issued instruction)", "", "| reason | ratio |", "|---|---|"]
for h, (v, u) in sorted(m.items(), key=lambda kv: -float(kv[1][0]) if "issue_stalled" in kv[0] and kv[0].endswith("per_issue_active.ratio") else 0):
    if "issue_stalled" in h and h.endswith("per_issue_active.ratio") and float(v) > 0.02:
        out.append(f"| {h.split('issue_stalled_')[1].replace('_per_issue_active.ratio','')} | {float(v):.3f} |")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h2, data = rows[1], rows[2:]
iS, iE, iN = h2.index("Source"), h2.index("Instructions Executed"), h2.index("# Samples")
tot = sum(int(r[iN]) for r in data if r[iN].isdigit())
by = collections.Counter()
for r in data:
    if r[iE].isdigit():
        op = [o for o in r[iS].split() if not o.startswith("@")][0].split(".")[0]
        by[op] += int(r[iE])
ti = sum(by.values())
out += ["", f"## dynamic instruction mix ({ti} warp instructions)", "", "| opcode | share |", "|---|---|"]
out += [f"| {op} | {100 * c / ti:.1f} % |" for op, c in by.most_common(14)]
out += ["", f"## hottest SASS lines by stall samples ({tot} samples)", "", "| samples | executed | SASS |", "|---|---|---|"]
for r in sorted(data, key=lambda r: -int(r[iN]) if r[iN].isdigit() else 0)[:12]:
    out.append(f"| {r[iN]} ({100 * int(r[iN]) / tot:.1f} %) | {r[iE]} | `{r[iS].strip()}` |")
open(f"profiles/{tag}_ncu_summary.md", "w").write("\n".join(out) + "\n")
if "eval_pdf" in tag:
    rd, wr = float(m["dram__bytes_read.sum"][0]), float(m["dram__bytes_write.sum"][0])
    scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
    b = rd * scale[m["dram__bytes_read.sum"][1]] + wr * scale[m["dram__bytes_write.sum"][1]]
    json.dump({"eval_pdf_dram_bytes_per_launch": b, "source": f"profiles/{tag}_ncu_summary.md"}, open("profiles/traffic.json", "w"))
print("\n".join(out[:40]))
