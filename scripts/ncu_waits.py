#!/usr/bin/env python
"""Attribute stall samples of inlined mbarrier waits to their call sites: walks the SASS of one kernel in
address order and reports, for every hot spot, the enclosing non-helper source lines.
Usage: ncu_waits.py rep kernel_regex [min_samples]"""
import csv, io, subprocess, sys
rep, kre = sys.argv[1], sys.argv[2]
thr = int(sys.argv[3]) if len(sys.argv) > 3 else 1500
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + kre],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
def I(x):
    try: return int(x)
    except Exception: return 0
# rows: source-line rows (Line No set) followed by their sass rows (Address set)
sass = []   # (addr, line_no, src_text, sass_text, samples)
cur = (None, None)
for r in rows[3:]:
    if len(r) < 6: continue
    if r[0].strip() and r[0] != 'Line No':
        cur = (r[0], r[1])
    elif r[2].startswith('0x'):
        sass.append((int(r[2], 16), cur[0], cur[1], r[3], I(r[4])))
sass.sort()
tot = sum(s[4] for s in sass)
print("total samples", tot)
helper = lambda txt: any(k in txt for k in ("mbar_try_wait", "spins", "asm volatile", "__device__ __forceinline__", "NV_IF", "smem_u32", "cvta"))
for i, s in enumerate(sass):
    if s[4] >= thr:
        # nearest following non-helper line
        nxt = next((t for t in sass[i + 1:i + 400] if not helper(t[2])), None)
        prv = next((t for t in reversed(sass[max(0, i - 400):i]) if not helper(t[2])), None)
        print("%6d %4.1f%%  %-28s | after L%s: %-60s | before L%s: %s" % (s[4], 100 * s[4] / tot, s[3][:28], prv[1] if prv else '-', (prv[2] if prv else '')[:60],
                                                                nxt[1] if nxt else '-', (nxt[2] if nxt else '')[:60]))
