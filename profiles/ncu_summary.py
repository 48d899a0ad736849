"""Text summary of an ncu --set full report (first profiled launch): the metrics DESIGN.md quotes plus the top
stall instructions.  usage: python profiles/ncu_summary.py X.ncu-rep > profiles/<name>.txt   (needs ncu on PATH)"""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
print("report:", rep.split("/")[-1], " launches in report:", len(data))
keys = ["Kernel Name", "gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "sm__inst_executed.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "sm__cycles_active.avg", "sm__cycles_active.min", "sm__cycles_active.max", "sm__warps_active.avg.pct_of_peak_sustained_active"]
keys += [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
for d in data[:1]:
    for k in keys:
        if k in hdr:
            i = hdr.index(k)
            print(f"{k:95s} {d[i]} {units[i]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h = rows[1]
ws, so, ex = h.index("Warp Stall Sampling (All Samples)"), h.index("Source"), h.index("Instructions Executed")
stall = [i for i, x in enumerate(h) if x.startswith("stall_") and "Not Issued" not in x]
body = []
for r in rows[2:]:
    if len(r) != len(h) or r[ws] == h[ws]:
        if body:
            break
        continue
    body.append(r)
tot = sum(float(r[ws]) for r in body)
print("\ntop stall instructions (first launch), total samples", tot, " SASS instructions", len(body))
for idx, r in sorted(enumerate(body), key=lambda t: -float(t[1][ws]))[:25]:
    top = sorted(((float(r[i]), h[i]) for i in stall), reverse=True)[:2]
    print(f"{float(r[ws]):7.0f} {100*float(r[ws])/tot:5.1f}% line {idx:5d} exec {r[ex]:>9s} {r[so].strip()[:64]:64s} {top[0][1]}={top[0][0]:.0f} {top[1][1]}={top[1][0]:.0f}")
