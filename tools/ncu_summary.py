#!/usr/bin/env python
"""Summarise `ncu --page raw --csv` exports: one line per captured launch with the counters the
roofline tables quote (duration, DRAM bytes and %, L2 %, tensor pipe %, SM %, registers, smem)."""
import csv
import sys

KEYS = [
    ("gpu__time_duration.sum", "dur"),
    ("dram__bytes_read.sum", "dram_rd"),
    ("dram__bytes_write.sum", "dram_wr"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1%"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
    ("sm__inst_executed_pipe_tensor.sum", "tensor_inst"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor%"),
    ("sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", "hmma%"),
    ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "fma%"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
    ("launch__registers_per_thread", "regs"),
    ("launch__shared_mem_per_block_dynamic", "dsmem"),
    ("launch__grid_size", "grid"),
    ("lts__t_sector_hit_rate.pct", "l2hit%"),
]


def main(paths):
    for p in paths:
        rows = list(csv.reader(open(p)))
        hdr = None
        for i, r in enumerate(rows):
            if "Kernel Name" in r:
                hdr = i
                break
        if hdr is None:
            print(p, ": no kernel table")
            continue
        names, units = rows[hdr], rows[hdr + 1]
        col = {n: j for j, n in enumerate(names)}
        tens = [n for n in names if "tensor" in n and "pct" in n]
        for r in rows[hdr + 2:]:
            if len(r) < len(names):
                continue
            out = ["%s" % r[col["Kernel Name"]][:40]]
            for k, lab in KEYS:
                if k in col:
                    out.append("%s=%s%s" % (lab, r[col[k]], units[col[k]] if lab in ("dur", "dram_rd", "dram_wr", "dsmem") else ""))
            for k in tens:
                if k not in dict(KEYS) and "--all" in sys.argv:
                    out.append("%s=%s" % (k, r[col[k]]))
            print(p.split("/")[-1], "|", " ".join(out))


if __name__ == "__main__":
    main([a for a in sys.argv[1:] if not a.startswith("--")])
