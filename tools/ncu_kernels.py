"""Per-launch table of an .ncu-rep: duration, DRAM bytes, instructions, issue utilisation, occupancy, top stalls."""
import csv, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
i = [k for k, r in enumerate(rows) if r and r[0] == "ID"][0]
h = {n: k for k, n in enumerate(rows[i])}
def g(r, n):
    try: return float(r[h[n]].replace(",", ""))
    except Exception: return float("nan")
stalls = [n for n in h if n.startswith("smsp__average_warps_issue_stalled_") and n.endswith("_per_issue_active.ratio") and "not_issued" not in n]
print(f"{'kernel':34s} {'grid':>8s} {'ms':>8s} {'rdMB':>7s} {'wrMB':>7s} {'Minst':>7s} {'thr/i':>5s} {'issue%':>6s} {'warps%':>6s} {'regs':>4s}  top stalls")
for r in rows[i + 2:]:
    name = r[h["Kernel Name"]].replace("void <unnamed>::", "").split("(RenderCtx")[0]
    rd, wr = g(r, "dram__bytes_read.sum"), g(r, "dram__bytes_write.sum")
    ru, wu = rows[i + 1][h["dram__bytes_read.sum"]], rows[i + 1][h["dram__bytes_write.sum"]]
    sc = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}
    top = sorted(((g(r, n), n.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")) for n in stalls), reverse=True)[:4]
    print(f"{name[:34]:34s} {r[h['Grid Size']].split(',')[0].strip('('):>8s} {g(r,'gpu__time_duration.sum'):8.1f} {rd*sc.get(ru,1):7.1f} {wr*sc.get(wu,1):7.1f} "
          f"{g(r,'smsp__inst_executed.sum')/1e6:7.1f} {g(r,'smsp__thread_inst_executed_per_inst_executed.ratio'):5.1f} {g(r,'smsp__issue_active.avg.pct_of_peak_sustained_active'):6.1f} "
          f"{g(r,'sm__warps_active.avg.pct_of_peak_sustained_active'):6.1f} {int(g(r,'launch__registers_per_thread')):4d}  " + " ".join(f"{n}={v:.2f}" for v, n in top))
