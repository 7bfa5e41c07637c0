#!/usr/bin/env python
"""Per-source-line instruction and stall-sample counts of one kernel: joins the SASS page of an ncu report (per-instruction
counters) with nvdisasm's line table of the object file.  Assumes the report was taken from the same build.
Usage: python tools/sass_lines.py <report.ncu-rep> <object.o> <mangled kernel name substring> [top N]"""
import collections, csv, io, os, re, subprocess, sys, tempfile

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
