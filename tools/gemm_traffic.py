"""DRAM traffic of the dominant kernel from an ncu launch list:

    ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
        -k regex:gemm_tc --csv --log-file gpurun_out/r2_gemm_dram.csv python tools/profile_step.py --steps 1
    python tools/gemm_traffic.py gpurun_out/r2_gemm_dram.csv profiles/r2_gemm_dram.json

Writes {"launches", "dram_bytes_per_launch", "time_us_per_launch", ...}: what ``bench.py`` reports as
``roofline.traffic`` (per launch, like ``achieved``)."""
import csv
import json
import sys


def main(src, dst):
    rows = []
    with open(src) as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    rd = csv.DictReader(lines)
    per = {}
    for r in rd:
        if "Metric Name" not in r or not r.get("Metric Value"):
            continue
        key = r["ID"]
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "")
        scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "us": 1.0, "ms": 1e3, "ns": 1e-3, "s": 1e6}.get(unit, 1.0)
        per.setdefault(key, {"kernel": r["Kernel Name"]})[r["Metric Name"]] = v * scale
    rows = [p for p in per.values() if "gemm_tc" in p["kernel"]]
    n = len(rows)
    rdb = sum(p.get("dram__bytes_read.sum", 0.0) for p in rows)
    wrb = sum(p.get("dram__bytes_write.sum", 0.0) for p in rows)
    us = sum(p.get("gpu__time_duration.sum", 0.0) for p in rows)
    out = {"kernel": "gemm_tc_kernel", "launches": n, "dram_read_bytes_per_launch": rdb / n, "dram_write_bytes_per_launch": wrb / n,
           "dram_bytes_per_launch": (rdb + wrb) / n, "time_us_per_launch": us / n, "total_ms": us / 1e3,
           "source": f"ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum over every gemm_tc launch of one step ({src})"}
    json.dump(out, open(dst, "w"), indent=1)
    print(json.dumps(out))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
