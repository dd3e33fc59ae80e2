"""Top source lines of an `ncu --page source --csv --print-source cuda,sass` export (samples and executed instructions)."""
import csv, sys, collections, os
rows = list(csv.reader(open(sys.argv[1])))
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
fn = None; hdr = None; lines = []
for r in rows:
    if not r: continue
    if r[0] == "File Path": fn = os.path.basename(r[1]); continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = {n: i for i, n in enumerate(r)}; continue
    if hdr and r[0].isdigit():
        try:
            lines.append((fn, int(r[0]), r[1].strip(), int(r[hdr["# Samples"]]), int(r[hdr["Instructions Executed"]]), int(r[hdr["stall_no_inst"]])))
        except (ValueError, IndexError):
            pass
ts = sum(l[3] for l in lines); tx = sum(l[4] for l in lines)
byf = collections.Counter(); byfx = collections.Counter()
for l in lines: byf[l[0]] += l[3]; byfx[l[0]] += l[4]
print("total samples", ts, "warp-instr", tx)
for f, s in byf.most_common(): print(f"  {f:22s} samples {100*s/ts:5.1f}%  instr {100*byfx[f]/tx:5.1f}%")
print("top lines by executed instructions:")
for l in sorted(lines, key=lambda l: -l[4])[:topn]:
    print(f"  {100*l[4]/tx:5.2f}% x  {100*l[3]/ts:5.2f}% s  noinst {l[5]:5d}  {l[0]}:{l[1]}  {l[2][:110]}")
