"""Digest of an `ncu --page raw --csv` export: one line per launch with the metrics the roofline discussion uses.

    python tools/ncu_digest.py gpurun_out/r2_gemm_raw.csv > profiles/r2_gemm_ncu.md
"""
import csv
import re
import sys

COLS = [
    ("us", "gpu__time_duration.sum", 1.0),
    ("tensor pipe %", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", 1.0),
    ("sm %", "sm__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
    ("issue %", "smsp__issue_active.avg.pct_of_peak_sustained_active", 1.0),
    ("xu %", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", 1.0),
    ("fma %", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", 1.0),
    ("alu %", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", 1.0),
    ("dram rd MB", "dram__bytes_read.sum", 1.0),
    ("dram wr MB", "dram__bytes_write.sum", 1.0),
    ("dram %", "dram__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
    ("l2 %", "lts__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
    ("lts rd MB", "lts__t_bytes_equiv_l1sectormiss_pipe_lsu_mem_global_op_ld.sum", 1.0),
    ("sm MHz", "sm__cycles_elapsed.avg.per_second", 1.0),
    ("regs", "launch__registers_per_thread", 1.0),
    ("grid", "launch__grid_size", 1.0),
    ("warps %", "sm__warps_active.avg.pct_of_peak_sustained_active", 1.0),
]


def main(path):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    use = [(n, ix[k]) for n, k, _ in COLS if k in ix]
    print("| # | kernel | " + " | ".join(f"{n} [{units[i]}]" if units[i] else n for n, i in use) + " |")
    print("|---:|---|" + "---:|" * len(use))
    for r in rows[2:]:
        name = re.sub(r"\(.*", "", r[ix["Kernel Name"]].replace("void ", "").replace("<unnamed>::", ""))
        vals = []
        for _, i in use:
            try:
                vals.append(f"{float(r[i].replace(',', '')):.1f}")
            except ValueError:
                vals.append(r[i])
        print(f"| {r[ix['ID']]} | `{name}` | " + " | ".join(vals) + " |")


if __name__ == "__main__":
    main(sys.argv[1])
