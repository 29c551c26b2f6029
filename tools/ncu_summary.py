#!/usr/bin/env python
"""Condense an ``ncu --page raw --csv`` export into one line per kernel launch with the metrics the roofline discussion uses.

    python tools/ncu_summary.py gpurun_out/prof_r2_tc_raw.csv > profiles/r2/ncu_tc_summary.csv
"""
import csv
import sys

COLUMNS = [
    ("kernel", "Kernel Name"),
    ("grid", "launch__grid_size"),
    ("block", "launch__block_size"),
    ("regs", "launch__registers_per_thread"),
    ("smem_dyn_kb", "launch__shared_mem_per_block_dynamic"),
    ("time_ms", "gpu__time_duration.sum"),
    ("sm_clock_ghz", "sm__cycles_elapsed.avg.per_second"),
    ("tensor_pipe_pct", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
    ("sm_throughput_pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("dram_throughput_pct", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
    ("dram_read_gb", "dram__bytes_read.sum"),
    ("dram_write_gb", "dram__bytes_write.sum"),
    ("l2_hit_pct", "lts__t_sector_hit_rate.pct"),
    ("warps_active_pct", "sm__warps_active.avg.pct_of_peak_sustained_active"),
    ("nvlink_rx_bytes", "nvlrx__bytes.sum"),
    ("nvlink_tx_bytes", "nvltx__bytes.sum"),
]


def main(path):
    rows = list(csv.reader(open(path)))
    header, units, body = rows[0], rows[1], rows[2:]
    at = {name: i for i, name in enumerate(header)}
    out = csv.writer(sys.stdout)
    present = [(short, full) for short, full in COLUMNS if full in at]
    out.writerow([short + (f" [{units[at[full]]}]" if units[at[full]] else "") for short, full in present])
    for row in body:
        name = row[at["Kernel Name"]]
        if name.startswith("void at::") or "at::native" in name:
            continue  # torch's own fill / copy kernels that happened to run inside the range
        out.writerow([row[at[full]][:90] if short == "kernel" else row[at[full]] for short, full in present])


if __name__ == "__main__":
    main(sys.argv[1])
