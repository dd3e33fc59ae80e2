"""Summarise an `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --csv` launch list:
per kernel: launches, total time (cold-cache, serialised: use the SHARE), DRAM bytes.  With --last-half only the second
half of the launches is used (tools/render_once.py renders twice: warm-up, then the render that counts).
With --from-last NAME only the launches from the last launch whose kernel name contains NAME onward are used (the scene's BVH
build precedes the renders in the list, so halving the list no longer separates them: --from-last k_light_paths = the
second render).
With --version V --workload W --scale S the summary is stamped so that bench.py's roofline.traffic only uses it for the very
build (b2_version()) and workload it was captured on; S = bench step / captured render (e.g. 16 for a 2-spp capture of the 32-spp step).
usage: ncu_traffic.py launches.csv out.json [--last-half | --from-last NAME] [--version V --workload W --scale S]"""
import csv, json, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
i = [k for k, r in enumerate(rows) if r and r[0] == "ID"][0]
h = {n: k for k, n in enumerate(rows[i])}
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}
recs = collections.OrderedDict()
for r in rows[i + 1:]:
    if len(r) <= h["Metric Value"] or not r[0].isdigit(): continue
    k = int(r[0]); d = recs.setdefault(k, {"name": r[h["Kernel Name"]]})
    d[r[h["Metric Name"]]] = float(r[h["Metric Value"]].replace(",", "")) * scale.get(r[h["Metric Unit"]], 1.0)
launches = list(recs.values())
if "--last-half" in sys.argv: launches = launches[len(launches) // 2:]
if "--from-last" in sys.argv:
    name = sys.argv[sys.argv.index("--from-last") + 1]
    launches = launches[max(k for k, d in enumerate(launches) if name in d["name"]):]
out = collections.OrderedDict(); tot = 0.0
for d in launches:
    n = d["name"].replace("void <unnamed>::", "").replace("void ", "").split("(")[0]
    o = out.setdefault(n, {"launches": 0, "us": 0.0, "dram_read_bytes": 0.0, "dram_write_bytes": 0.0, "warp_inst": 0.0})
    o["launches"] += 1; o["us"] += d.get("gpu__time_duration.sum", 0.0); tot += d.get("gpu__time_duration.sum", 0.0)
    o["dram_read_bytes"] += d.get("dram__bytes_read.sum", 0.0); o["dram_write_bytes"] += d.get("dram__bytes_write.sum", 0.0)
    o["warp_inst"] += d.get("smsp__inst_executed.sum", 0.0)
for o in out.values(): o["share"] = o["us"] / tot
def opt(name, default=None):
    return sys.argv[sys.argv.index(name) + 1] if name in sys.argv else default
res = {"source": sys.argv[1], "version": opt("--version"), "workload": opt("--workload"), "per_step_scale": float(opt("--scale", 1.0)),
       "n_launches": len(launches), "total_us_under_ncu": tot, "kernels": out}
json.dump(res, open(sys.argv[2], "w"), indent=1)
for n, o in out.items():
    print(f"{n:28s} x{o['launches']:3d} {o['us']:9.1f} us {100*o['share']:5.1f}%  rd {o['dram_read_bytes']/1e6:8.1f} MB  wr {o['dram_write_bytes']/1e6:8.1f} MB  {o['warp_inst']/1e6:8.1f} Minst")
