#!/usr/bin/env python
"""Digest of one ncu report: headline metrics, stall mix and the hottest SASS lines of the first kernel in it.
Usage: python tools/ncu_digest.py <report.ncu-rep> [tiles_for_normalisation]
       python tools/ncu_digest.py --lines <report.ncu-rep> <object.o> <mangled kernel name substring> [top N]
         per-SOURCE-LINE instruction and stall-sample counts: joins the SASS page of the report (per-instruction counters)
         with nvdisasm's line table of the object file (assumes the report was taken from the same build)"""
import collections, csv, io, os, re, subprocess, sys, tempfile


def per_line(argv):
    sys.argv = ["sass_lines"] + argv

    rep, obj, key = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
    sass = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout.splitlines()
    line_of = {}  # offset -> (file, line)
    inside, cur = False, ("?", 0)
    for ln in sass:
        if ln.startswith(".text."):
            inside = key in ln
            continue
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/", ln)
        if m:
            line_of[int(m.group(1), 16)] = cur
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr = rows[1]
    ia, ii, isamp = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples")
    data = [r for r in rows[2:] if len(r) == len(hdr)]
    base = int(data[0][ia], 16)
    agg = collections.defaultdict(lambda: [0, 0])
    tot_i = tot_s = 0
    for r in data:
        off = int(r[ia], 16) - base
        if off in line_of and int(r[ia], 16) >= base:
            k = line_of[off]
        else:
            k = ("?", 0)
        try:
            n, s = int(r[ii]), int(r[isamp])
        except ValueError:
            continue
        agg[k][0] += n; agg[k][1] += s
        tot_i += n; tot_s += s
        if off > max(line_of) and off != 0 and r is not data[0] and int(r[ia], 16) == base:
            break
    print(f"total warp instructions {tot_i}, samples {tot_s}")
    for k, (n, s) in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
        print(f"{100*n/tot_i:5.1f}% inst {100*s/max(tot_s,1):5.1f}% smp  {k[0]}:{k[1]}")


if len(sys.argv) > 1 and sys.argv[1] == "--lines":
    per_line(sys.argv[2:])
    sys.exit(0)

rep = sys.argv[1]
tiles = float(sys.argv[2]) if len(sys.argv) > 2 else None
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, r = rows[0], rows[2]
def g(k): return r[hdr.index(k)] if k in hdr else None
print(g("Kernel Name"))
for k in ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size", "launch__registers_per_thread",
          "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
          "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "lts__t_sector_hit_rate.pct",
          "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]:
    print(f"  {k:70s} {g(k)}")
for i, h in enumerate(hdr):
    if "issue_stalled" in h and h.endswith(".ratio") and "not_issued" not in h:
        try:
            v = float(r[i])
        except ValueError:
            continue
        if v > 0.15:
            print("  stall/issue", h.replace("smsp__average_warps_issue_stalled_", "").replace("smsp__average_warp_latency_issue_stalled_", ""), round(v, 2))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]
data = [x for x in rows[2:] if len(x) == len(hdr)]
ix = {h: i for i, h in enumerate(hdr)}
first = data[0][ix["Address"]]  # a report with several launches repeats the listing: keep the first copy
for i in range(1, len(data)):
    if data[i][ix["Address"]] == first:
        data = data[:i]
        break
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
def n(x, k):
    try:
        return int(x[ix[k]])
    except ValueError:
        return 0
T = sum(n(x, "# Samples") for x in data)
E = sum(n(x, "Instructions Executed") for x in data)
print(f"  SASS lines {len(data)}, warp instructions {E}" + (f", per tile-warp {E / tiles:.0f}" if tiles else ""))
ops = collections.Counter()
for x in data:
    s = x[ix["Source"]].strip().split()
    op = s[1] if s[0].startswith("@") else s[0]
    ops[op.split(".")[0]] += n(x, "Instructions Executed")
print("  opcode mix: " + ", ".join(f"{k} {100 * v / E:.1f}%" for k, v in ops.most_common(14)))
tot = {s: sum(n(x, s) for x in data) for s in stalls}
print("  samples: " + ", ".join(f"{s[6:]} {100 * v / T:.1f}%" for s, v in sorted(tot.items(), key=lambda kv: -kv[1])[:9]))
for i in sorted(range(len(data)), key=lambda i: -n(data[i], "# Samples"))[:16]:
    x = data[i]
    ss = sorted([(s[6:], n(x, s)) for s in stalls if n(x, s) > 0], key=lambda kv: -kv[1])
    print(f"  {i:5d} {100 * n(x, '# Samples') / T:4.1f}%  {x[ix['Source']].strip()[:52]:52s} {ss[:2]}")
