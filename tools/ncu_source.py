"""Aggregate the source page of a --set full --import-source on report by CUDA source line:
   python tools/ncu_source.py REPORT.ncu-rep [top]
prints, per (file, line): warp instructions executed, average active threads, stall samples."""
import csv, io, subprocess, sys
from collections import defaultdict
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
agg = defaultdict(lambda: [0, 0, 0, ""])
cur_file = ""; hdr = None; line_key = None
tot_i = tot_t = tot_s = 0
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = r; iI = r.index("Instructions Executed"); iT = r.index("Thread Instructions Executed"); iS = r.index("# Samples"); continue
    if hdr is None: continue
    if r[0] != "":   # a CUDA source line; SASS rows follow
        line_key = (cur_file, int(r[0])); agg[line_key][3] = r[1].strip()[:90]; continue
    try: i, t, s = int(r[iI]), int(r[iT]), int(r[iS])
    except ValueError: continue
    a = agg[line_key]; a[0] += i; a[1] += t; a[2] += s
    tot_i += i; tot_t += t; tot_s += s
print("# %s: %d warp instructions, %.2f threads/instruction, %d stall samples" % (rep, tot_i, tot_t / max(tot_i, 1), tot_s))
print("%-22s %5s %7s %7s %7s  %s" % ("file", "line", "inst%", "thr/in", "smpl%", "source"))
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    if a[0] == 0: continue
    print("%-22s %5d %6.2f%% %7.2f %6.2f%%  %s" % (k[0], k[1], 100 * a[0] / tot_i, a[1] / a[0], 100 * a[2] / max(tot_s, 1), a[3]))
