#!/usr/bin/env python
"""SASS / ptxas summary of libpntf_b200.so for profiles/: per kernel, registers / stack / spills (ptxas -v) and the counts
of the Blackwell-specific instructions (cuobjdump -sass).  Usage: python scripts/sass_summary.py profiles/<tag>_sass_summary.md"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "p-ntfields_b200", "pntf", "libpntf_b200.so")
KEYS = ("UTCHMMA", "UTCBAR", "LDTM", "STTM", "UBLKCP", "SYNCS", "UTCATOM", "MUFU", "FFMA", "HFMA2", "F2FP", "STS", "LDS", "LDG", "STG", "ATOM", "RED", "BAR")


def demangle(n):
    return subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip().split("(")[0]


def main():
    out = sys.argv[1]
    build = subprocess.run([sys.executable, os.path.join(ROOT, "p-ntfields_b200", "build.py"), "-v", "--force"],
                           capture_output=True, text=True).stdout
    res, cur = {}, None
    for line in build.splitlines():
        m = re.search(r"Function properties for (\S+)", line)
        if m:
            cur = m.group(1)
            continue
        m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", line)
        if m and cur:
            res.setdefault(cur, {})["spill"] = (int(m.group(1)), int(m.group(2)), int(m.group(3)))
        m = re.search(r"Used (\d+) registers", line)
        if m and cur:
            res.setdefault(cur, {})["regs"] = int(m.group(1))
    sass = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
    hist, cur = {}, None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            hist[cur] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m and cur:
            hist[cur][m.group(1)] += 1
    with open(out, "w") as f:
        f.write("# SASS / ptxas summary of libpntf_b200.so (sm_100a)\n\n`python scripts/sass_summary.py` : `nvcc -Xptxas -v` "
                "and `cuobjdump -sass` of the in-tree library.  Counts are static instruction counts per kernel.\n\n")
        f.write("| kernel | regs | stack | spill st/ld | total | " + " | ".join(KEYS) + " |\n")
        f.write("|---|---|---|---|---|" + "---|" * len(KEYS) + "\n")
        for k in sorted(hist, key=demangle):
            h = hist[k]
            r = res.get(k, {})
            sp = r.get("spill", (0, 0, 0))
            f.write("| `%s` | %s | %d | %d / %d | %d | " % (demangle(k).replace("pntf::", ""), r.get("regs", "?"), sp[0], sp[1], sp[2], sum(h.values())))
            f.write(" | ".join(str(sum(v for op, v in h.items() if op.startswith(key))) for key in KEYS) + " |\n")
        tot = collections.Counter()
        for h in hist.values():
            tot.update(h)
        f.write("\nLibrary totals: " + ", ".join("%s %d" % (key, sum(v for op, v in tot.items() if op.startswith(key))) for key in KEYS[:7]) + "\n")
    print(open(out).read()[:1500])


if __name__ == "__main__":
    main()
