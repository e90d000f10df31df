#!/usr/bin/env python
"""profiles/dram_traffic.json from `ncu --set full` reports: dram__bytes_read.sum + dram__bytes_write.sum of
one launch, keyed by (kernel, K, N, M).  bench.py emits roofline.traffic only from entries of this file whose
kernel / shape / M equal what it timed.

    python scripts/make_dram_traffic.py REPORT.ncu-rep:K:N:M:profiles/SUMMARY.txt [...]"""
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    dst = os.path.join(ROOT, "profiles", "dram_traffic.json")
    entries = json.load(open(dst)) if os.path.exists(dst) else []
    for spec in sys.argv[1:]:
        rep, K, N, M, src = spec.split(":")
        out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
        rows = list(csv.reader(io.StringIO(out)))
        hdr, units, r = rows[0], rows[1], rows[2]
        val = lambda k: float(r[hdr.index(k)].replace(",", "")) * UNIT[units[hdr.index(k)]]  # noqa: E731
        name = re.search(r"(\w+_kernel)", r[hdr.index("Kernel Name")]).group(1)
        e = {"kernel": "awq::" + name, "K": int(K), "N": int(N), "M": int(M),
             "dram_bytes": val("dram__bytes_read.sum") + val("dram__bytes_write.sum"),
             "dram_bytes_read": val("dram__bytes_read.sum"), "dram_bytes_write": val("dram__bytes_write.sum"),
             "duration_us_under_ncu": val("gpu__time_duration.sum") if units[hdr.index("gpu__time_duration.sum")] in UNIT
             else float(r[hdr.index("gpu__time_duration.sum")]), "source": src}
        entries = [x for x in entries if (x["kernel"], x["K"], x["N"], x["M"]) != (e["kernel"], e["K"], e["N"], e["M"])]
        entries.append(e)
    json.dump(entries, open(dst, "w"), indent=1)
    print("wrote", dst, len(entries), "entries")


if __name__ == "__main__":
    main()
