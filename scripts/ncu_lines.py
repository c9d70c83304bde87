#!/usr/bin/env python
"""Top CUDA source lines by stall samples for one kernel of an .ncu-rep.  Usage: ncu_lines.py rep kernel_regex [n]"""
import csv, io, subprocess, sys
rep, kre = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + kre],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[2]
def I(x):
    try: return int(x)
    except Exception: return 0
lines = [r for r in rows[3:] if r and r[0].strip() and r[0] != 'Line No' and len(r) > 7]
ie = hdr.index('Instructions Executed')
tot = sum(I(r[4]) for r in lines); toti = sum(I(r[ie]) for r in lines)
print("kernel", rows[1][1][:80], "samples", tot, "warp-instr", toti)
for r in sorted(lines, key=lambda r: -I(r[4]))[:n]:
    print("%5s %7d %5.1f%% %11d  %s" % (r[0], I(r[4]), 100 * I(r[4]) / max(tot, 1), I(r[ie]), r[1][:120]))
