"""Attribute the per-instruction counters of one `ncu --set full --import-source on` capture to source lines.

    python tools/ncu_by_line.py <report.ncu-rep> <object.o> <mangled-kernel-substring> [--match <demangled substring>] [--top N] [--mem]

(<mangled-kernel-substring> selects the function in the object, e.g. k_connectILb0ELb1; --match selects the kernel in the
report, e.g. "k_connect<(bool)0, (bool)1>"; default: the first kernel of the report)

ncu's CSV source page carries counters only in its SASS view; this maps each SASS address back to source through
`nvdisasm --print-line-info-inline` of the SAME object (build/obj/*.o, -lineinfo) and prints, for executed warp instructions and
for stall samples, the share of (a) every line of the kernel's own file an instruction is (transitively) inlined into -
i.e. the phase of the kernel - and (b) the innermost file:line."""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict


def line_map(obj, kernel):
    with tempfile.TemporaryDirectory() as d:
        subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=d, stdout=subprocess.DEVNULL)
        cubin = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
        txt = subprocess.run(["nvdisasm", "--print-line-info-inline", os.path.join(d, cubin)], capture_output=True, text=True).stdout
    amap, cur, pending, on = {}, [("?", 0)], [], False
    for ln in txt.splitlines():
        if ln.startswith("\t.section") or ln.startswith(".text."):
            on = (".text." in ln) and (kernel in ln)
            continue
        if not on:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)      # a chain: innermost first ("... inlined at ..."), the kernel's own line last
        if m:
            pending.append((m.group(1), int(m.group(2))))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(\S.*?);", ln)
        if m:
            if pending:
                cur, pending = pending, []
            amap[int(m.group(1), 16)] = (cur, m.group(2).strip())
    return amap


def main():
    rep, obj, kernel = sys.argv[1:4]
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 25
    amap = line_map(obj, kernel)
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    blocks = out.split('"Kernel Name",')
    match = sys.argv[sys.argv.index("--match") + 1] if "--match" in sys.argv else None
    blk = [b for b in blocks[1:] if match is None or match in b.split("\n", 1)[0]]
    print("report kernel:", blk[0].split("\n", 1)[0].strip().strip(",").strip('"'))
    rows = list(csv.reader(io.StringIO(blk[0].split("\n", 1)[1])))
    hdr = rows[0]
    ia, ii, isamp = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples")
    ith = hdr.index("Thread Instructions Executed")
    base = None
    outer_i, outer_s, inner_i, inner_s, file_i, thr = defaultdict(float), defaultdict(float), defaultdict(float), defaultdict(float), defaultdict(float), defaultdict(float)
    tot_i = tot_s = 0.0
    main_file = None
    for r in rows[1:]:
        if len(r) <= max(ii, isamp) or not r[ia]:
            continue
        a = int(r[ia], 16) if not r[ia].startswith("0x") else int(r[ia], 16)
        if base is None:
            base = a
        n, s = float(r[ii] or 0), float(r[isamp] or 0)
        chain, _ = amap.get(a - base, ([("?", 0)], ""))
        if main_file is None and chain[-1][0] != "?":
            main_file = chain[-1][0]
        inner = "%s:%d" % (os.path.basename(chain[0][0]), chain[0][1])
        outer = "%s:%d" % (os.path.basename(chain[-1][0]), chain[-1][1])
        outer_i[outer] += n; outer_s[outer] += s; inner_i[inner] += n; inner_s[inner] += s
        file_i[os.path.basename(chain[0][0])] += n; thr[outer] += float(r[ith] or 0)
        tot_i += n; tot_s += s
    print("kernel %s: %.4g warp instructions, %d stall samples" % (kernel, tot_i, tot_s))
    print("-- by innermost file (share of executed warp instructions)")
    for k, v in sorted(file_i.items(), key=lambda kv: -kv[1]):
        print("  %-28s %6.2f %%" % (k, 100 * v / tot_i))
    print("-- by line of the kernel body (outermost frame): inst %%, stall-sample %%, active threads / instruction")
    for k, v in sorted(outer_i.items(), key=lambda kv: -kv[1])[:top]:
        print("  %-28s %6.2f %%  %6.2f %%  %5.1f" % (k, 100 * v / tot_i, 100 * outer_s[k] / max(tot_s, 1), thr[k] / max(v, 1)))
    if "--mem" in sys.argv:
        # memory instructions by innermost source line: L1 tag requests and the sectors they ask L2 for ("theoretical": before
        # the L1 hit rate), global and local; x 32 B = bytes.  Names the buffer behind every load of the kernel.
        col = lambda name: hdr.index(name) if name in hdr else -1          # a kernel without local memory has no local column
        il1, il2g, il2l = col("L1 Tag Requests Global"), col("L2 Theoretical Sectors Global"), col("L2 Theoretical Sectors Local")
        mem = defaultdict(lambda: [0.0, 0.0, 0.0, ""])
        for r in rows[1:]:
            if len(r) <= max(il1, il2g, il2l) or not r[ia]:
                continue
            a = int(r[ia], 16)
            chain, ins = amap.get(a - base, ([("?", 0)], ""))
            v = [float(r[k] or 0) if k >= 0 else 0.0 for k in (il1, il2g, il2l)]
            if sum(v) == 0:
                continue
            key = "%s:%d" % (os.path.basename(chain[0][0]), chain[0][1])
            m = mem[key]; m[0] += v[0]; m[1] += v[1]; m[2] += v[2]; m[3] = ins.split()[0] if ins else m[3]
        tg = sum(m[1] for m in mem.values()); tl = sum(m[2] for m in mem.values())
        print("-- memory instructions by innermost line: L1 tag requests, L2 theoretical sectors global (share) / local (share), bytes = sectors x 32")
        print("   total: global %.3f GB, local %.3f GB" % (tg * 32 / 1e9, tl * 32 / 1e9))
        for k, m in sorted(mem.items(), key=lambda kv: -(kv[1][1] + kv[1][2]))[:top]:
            print("  %-26s %-8s tags %12.0f  global %8.3f GB (%5.1f %%)  local %8.3f GB (%5.1f %%)" % (k, m[3], m[0], m[1] * 32 / 1e9, 100 * m[1] / max(tg, 1), m[2] * 32 / 1e9, 100 * m[2] / max(tl, 1)))
    print("-- by innermost line: inst %%, stall-sample %%")
    for k, v in sorted(inner_i.items(), key=lambda kv: -kv[1])[:top]:
        print("  %-28s %6.2f %%  %6.2f %%" % (k, 100 * v / tot_i, 100 * inner_s[k] / max(tot_s, 1)))


if __name__ == "__main__":
    main()
