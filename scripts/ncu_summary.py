#!/usr/bin/env python
"""Condense `ncu -i X.ncu-rep --page raw --csv` into one row per kernel launch with the metrics the roofline discussion uses
(duration, DRAM bytes and % of peak, tensor-pipe activity, SM throughput, occupancy, registers, L2 hit rate).
    python scripts/ncu_summary.py raw.csv > profiles/ncu_<what>.csv"""
import csv
import sys

WANT = [("Kernel Name", "kernel"), ("Grid Size", "grid"), ("Block Size", "block"), ("gpu__time_duration.sum", "dur"),
        ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_thr_pct"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor_pipe_active_pct"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor_pipe_elapsed_pct"),
        ("sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tmem_pct"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy_pct"), ("launch__registers_per_thread", "regs"),
        ("lts__t_sector_hit_rate.pct", "l2_hit_pct"), ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem_wavefronts"),
        ("smsp__inst_executed.sum", "warp_insts")]


def main() -> None:
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units = rows[0], rows[1]
    cols = []
    for name, short in WANT:
        ix = [i for i, h in enumerate(hdr) if h == name]
        if ix:
            cols.append((ix[0], short, units[ix[0]]))
    w = csv.writer(sys.stdout)
    w.writerow([f"{s} [{u}]" if u else s for _, s, u in cols])
    for r in rows[2:]:
        if len(r) < len(hdr):
            continue
        out = []
        for i, s, _ in cols:
            v = r[i]
            if s == "kernel":
                v = v.split("(")[0][-70:]
            out.append(v)
        w.writerow(out)


if __name__ == "__main__":
    main()
