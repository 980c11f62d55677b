#!/usr/bin/env python3
"""DRAM traffic per launch of every kernel in an `ncu --set full` report, as the JSON bench.py's roofline reads
(profiles/*ncu_traffic*.json): usage  ncu_traffic.py report.ncu-rep out.json "source note" """
import collections
import csv
import json
import re
import subprocess
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
TIME = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}


def main(rep, out, note):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    h, units = rows[0], rows[1]
    kn = h.index("Kernel Name")
    rd, wr, tm = h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum"), h.index("gpu__time_duration.sum")
    kernels = collections.OrderedDict()
    for r in rows[2:]:
        if len(r) <= max(rd, wr, tm):
            continue
        sig = r[kn]
        name = re.sub(r"[<(].*", "", sig).split()[-1].split("::")[-1]
        b = float(r[rd].replace(",", "")) * UNIT.get(units[rd], 1.0) + float(r[wr].replace(",", "")) * UNIT.get(units[wr], 1.0)
        us = float(r[tm].replace(",", "")) * TIME.get(units[tm], 1.0)
        kernels.setdefault(name, []).append({"dram_bytes": int(b), "ncu_us": round(us, 3), "signature": sig[:120]})
    # the library's profile names the two users of the generic sort kernels separately (LODGS_LAUNCHED labels): per step the launch order
    # is depth histogram, 4 depth passes, tile histogram, 2 tile passes
    if "sort_onesweep_kernel" in kernels and len(kernels["sort_onesweep_kernel"]) % 6 == 0:
        v = kernels["sort_onesweep_kernel"]
        kernels["sort_depth_pass_kernel"] = [e for i, e in enumerate(v) if i % 6 < 4]
        kernels["sort_tile_pass_kernel"] = [e for i, e in enumerate(v) if i % 6 >= 4]
    if "sort_histogram_kernel" in kernels and len(kernels["sort_histogram_kernel"]) % 2 == 0:
        v = kernels["sort_histogram_kernel"]
        kernels["sort_depth_hist_kernel"] = v[0::2]
        kernels["sort_tile_hist_kernel"] = v[1::2]
    json.dump({"source": note, "kernels": kernels}, open(out, "w"), indent=1)
    for k, v in kernels.items():
        print("%-28s launches %3d  max dram bytes %12d  last ncu_us %9.3f" % (k, len(v), max(e["dram_bytes"] for e in v), v[-1]["ncu_us"]))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "")
