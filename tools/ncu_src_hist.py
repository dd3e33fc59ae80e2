"""Histogram of an `ncu --page source --csv` export: samples, stall_no_inst and executed instructions per block of SASS."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
step = int(sys.argv[2]) if len(sys.argv) > 2 else 128
sect = int(sys.argv[3]) if len(sys.argv) > 3 else 0
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"] + [len(rows)]
print(rows[starts[sect]][1], f"(section {sect} of {len(starts) - 1})")
h = rows[starts[sect] + 1]
ci = {n: i for i, n in enumerate(h)}
body = [r for r in rows[starts[sect] + 2:starts[sect + 1]] if len(r) >= len(h) - 1]
tot = sum(int(r[ci["# Samples"]]) for r in body); totx = sum(int(r[ci["Instructions Executed"]]) for r in body)
print("instr", len(body), "samples", tot, "warp-instr executed", totx)
keys = ["stall_no_inst", "stall_wait", "stall_long_sb", "stall_short_sb", "stall_branch_resolving", "stall_math", "stall_not_selected", "stall_selected"]
print("block  samples%  exec%  avg_thr " + " ".join(k[6:14] for k in keys) + "   first-instr")
for b in range(0, len(body), step):
    blk = body[b:b + step]
    s = sum(int(r[ci["# Samples"]]) for r in blk); x = sum(int(r[ci["Instructions Executed"]]) for r in blk)
    tx = sum(int(r[ci["Thread Instructions Executed"]]) for r in blk)
    if s == 0 and x == 0: continue
    ks = [sum(int(r[ci[k]]) for r in blk) for k in keys]
    print(f"{b:5d} {100*s/tot:7.2f} {100*x/max(totx,1):7.2f} {tx/max(x,1):6.1f}  " + " ".join(f"{100*k/max(s,1):7.1f}" for k in ks) + "   " + blk[0][1].strip()[:40])
