"""Top warp-stall instructions from an `ncu --page source --csv` dump.
usage: ncu -i X.ncu-rep --page source --csv > src.csv; python profiles/top_stalls.py src.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = rows[1]
ws, src, ex = hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Source"), hdr.index("Instructions Executed")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
body = []
for r in rows[2:]:          # a report with several launches repeats the header block: keep the first launch
    if len(r) != len(hdr):
        if body:
            break
        continue
    if r[ws] == hdr[ws]:
        if body:
            break
        continue
    body.append(r)
tot = sum(float(r[ws]) for r in body)
print("total samples", tot)
for idx, r in sorted(enumerate(body), key=lambda t: -float(t[1][ws]))[:n]:
    top = sorted(((float(r[i]), hdr[i]) for i in stall_cols), reverse=True)[:2]
    print(f"{float(r[ws]):7.0f} {100*float(r[ws])/tot:5.1f}% line {idx:4d} exec {r[ex]:>9s} {r[src].strip()[:70]:70s} {top[0][1]}={top[0][0]:.0f} {top[1][1]}={top[1][0]:.0f}")
