#!/usr/bin/env python
"""DRAM bytes per launch from the CSVs of scripts/profile_round2.sh (ncu --cache-control none over
consecutive launches on rotating buffers).  Prints one summary per workload and, with --write,
updates profiles/traffic.json (what bench.py reports as roofline.traffic).

    python scripts/traffic_summary.py gpurun_out/r2_traffic_*.csv [--write]
"""
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def to_bytes(value, unit):
    v = float(value.replace(",", ""))
    u = unit.lower()
    return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}[u]


def main():
    files = [a for a in sys.argv[1:] if not a.startswith("--")]
    out = {}
    for path in files:
        wl = re.search(r"r2_traffic_(.+)\.csv", path).group(1)
        rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
        hdr = rows[0]
        iid, iname, iunit, ival = hdr.index("ID"), hdr.index("Metric Name"), hdr.index("Metric Unit"), hdr.index("Metric Value")
        per = {}
        for r in rows[1:]:
            per.setdefault(r[iid], {})[r[iname]] = (r[ival], r[iunit])
        rd = [to_bytes(*m["dram__bytes_read.sum"]) for m in per.values()]
        wr = [to_bytes(*m["dram__bytes_write.sum"]) for m in per.values()]
        pct = [float(m["dram__throughput.avg.pct_of_peak_sustained_elapsed"][0]) for m in per.values()]
        us = []
        for m in per.values():
            v, u = m["gpu__time_duration.sum"]
            us.append(float(v.replace(",", "")) * {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "nsecond": 1e-3, "ms": 1e3, "msecond": 1e3}[u])
        n = len(rd)
        s = {"launches": n, "dram_read_bytes_per_launch": sum(rd) / n, "dram_write_bytes_per_launch": sum(wr) / n,
             "dram_bytes_per_launch": (sum(rd) + sum(wr)) / n, "dram_throughput_pct_of_peak_mean": sum(pct) / n,
             "gpu_time_us_mean_serialised": sum(us) / n}
        out[wl] = s
        print(wl, json.dumps(s))
    if "--write" in sys.argv:
        tj = {wl: int(round(s["dram_bytes_per_launch"])) for wl, s in out.items()}
        tj["_detail"] = out
        tj["_note"] = ("dram__bytes_read.sum + dram__bytes_write.sum per launch of pva_tiled_kernel, mean over consecutive launches "
                       "(ncu --cache-control none --clock-control none, rotating buffer sets, scripts/profile_round2.sh): dirty lines a "
                       "launch leaves in L2 are written back during the next one, so the mean over consecutive launches is the "
                       "steady-state figure.  Raw CSVs: profiles/r2_traffic_*.csv")
        json.dump(tj, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
