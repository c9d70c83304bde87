#!/usr/bin/env python
"""DRAM bytes per 128-row tile and kernel class from an `ncu --set full` report -> profiles/<tag>_dram_bytes_per_tile.json,
the table bench.py's roofline.traffic is computed from (so that it cannot go stale against hard-coded constants).

Usage: python scripts/ncu_traffic.py gpurun_out/tc_full_<tag>.ncu-rep profiles/<tag>_dram_bytes_per_tile.json <tiles of the captured launch> "<command>"
The report must hold one launch of each training kernel over the same chunk of tiles (scripts/gpu_round.sh: -k regex:tc_ -c 3)."""
import csv
import io
import json
import subprocess
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
CLASS = {"tc_forward_kernel": "tc_forward", "tc_reverse_kernel": "tc_reverse", "tc_dw_kernel": "tc_dw", "tc_backward_kernel": "tc_backward"}


def main():
    rep, out, tiles, cmd = sys.argv[1], sys.argv[2], int(sys.argv[3]), (sys.argv[4] if len(sys.argv) > 4 else "")
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    head, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(head)}
    res, detail = {}, {}
    for r in data:
        name = r[col["Kernel Name"]]
        cls = next((v for k, v in CLASS.items() if k in name), None)
        if cls is None or cls in res:
            continue
        tot = 0.0
        d = {}
        for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            val = float(r[col[m]].replace(",", "")) * UNIT[units[col[m]]]
            d[m] = val
            tot += val
        d["gpu__time_duration.sum_ms"] = float(r[col["gpu__time_duration.sum"]].replace(",", "")) * \
            {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "nsecond": 1e-6, "s": 1e3, "second": 1e3}[units[col["gpu__time_duration.sum"]]]
        res[cls] = tot / tiles
        detail[cls] = d
    json.dump({"source": rep, "command": cmd, "what": "dram__bytes_read.sum + dram__bytes_write.sum of one launch / tiles of that "
               "launch (128-row tiles; 12 pairs per tile at d = 3, 8 at d = 6 in the loss kernels)",
               "tiles_per_launch": tiles, "bytes_per_tile": res, "per_launch": detail}, open(out, "w"), indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
