#!/usr/bin/env python3
"""profiles/dram_traffic.json from the CSV of `tools/profile_bench.sh dram <tag>` and the library it was
captured from (bench.py quotes roofline.traffic only while the device sources are the ones it was captured with).

    python tools/dram_traffic.py gpurun_out/<tag>_dram_bench_kernel.csv gpurun_out/<tag>_sources.sha256 profiles/<tag>_dram_bench_kernel.csv
"""
import csv
import json
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    src, sha_file, keep = sys.argv[1:4]
    rows = [r for r in csv.reader(open(src)) if len(r) > 10]
    h = rows[0]
    name, value = h.index("Metric Name"), h.index("Metric Value")
    m = {r[name]: float(r[value].replace(",", "")) for r in rows[1:]}
    unit = {r[name]: r[h.index("Metric Unit")] for r in rows[1:]}
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    shutil.copy(src, os.path.join(ROOT, keep))
    out = {
        "kernel": "render_kernel<false, false>",
        "workload": "scenes/scene5.json 1920x1080 -g path -n 16 (defaults: -b 10 -a 0.01 -s phong -l sqr)",
        "n_gpus": 1,
        "dram_bytes_read": int(m["dram__bytes_read.sum"] * scale[unit["dram__bytes_read.sum"]]),
        "dram_bytes_write": int(m["dram__bytes_write.sum"] * scale[unit["dram__bytes_write.sum"]]),
        "kernel_ns_under_ncu": int(m["gpu__time_duration.sum"] * {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9, "nsecond": 1, "usecond": 1e3, "msecond": 1e6, "second": 1e9}[unit["gpu__time_duration.sum"]]),
        "device_sources_sha256": open(sha_file).read().split()[0],
        "source": f"{keep}: ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none, "
                  "the fifth render_kernel launch of `python bench.py --no-cpu-baseline` (tools/profile_bench.sh dram)",
    }
    json.dump(out, open(os.path.join(ROOT, "profiles", "dram_traffic.json"), "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
