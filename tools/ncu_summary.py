#!/usr/bin/env python
"""Markdown summary of an ncu report: `python tools/ncu_summary.py rep.ncu-rep > profiles/x.md`.

Reads `ncu -i rep --page raw --csv`; one table per distinct kernel (first captured launch of each,
launch-to-launch spread noted), with duration, DRAM traffic, pipe utilisation, occupancy and the
top warp-stall reasons.
"""
import csv
import io
import re
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM write"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput % of peak"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput % of peak"),
    ("sm__instruction_throughput.avg.pct_of_peak_sustained_active", "instruction throughput % (active)"),
    ("sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed", "FMA-heavy pipe (IMAD) cycles active % (elapsed)"),
    ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "FMA pipe cycles active %"),
    ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "FMA pipe inst % of peak"),
    ("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "ALU pipe cycles active %"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "LSU inst % of peak"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("launch__registers_per_thread", "registers/thread"),
    ("launch__shared_mem_per_block_dynamic", "dynamic smem/block"),
    ("launch__shared_mem_per_block_static", "static smem/block"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__waves_per_multiprocessor", "waves/SM"),
    ("sm__sass_inst_executed_op_local_ld.sum", "local loads (spills)"),
    ("sm__sass_inst_executed_op_local_st.sum", "local stores (spills)"),
    ("smsp__cycles_active.avg", "SMSP active cycles"),
    ("sm__cycles_elapsed.max", "SM cycles elapsed"),
]


def short(name):
    m = re.search(r"(k_\w+)", name)
    base = m.group(1) if m else name[:60]
    for c in ("alt_bn128", "bls12_377", "bw6_761"):
        if c in name:
            base += f" <{c}{' Fp2' if 'Fp2' in name else ''}>"
            break
    return base


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    kn = col["Kernel Name"]
    seen = {}
    for d in data:
        seen.setdefault(short(d[kn]), []).append(d)
    print(f"source: `{rep}` ({len(data)} captured launches; ncu --set full --clock-control none)")
    for name, launches in seen.items():
        d = launches[0]
        durs = [float(x[col["gpu__time_duration.sum"]]) for x in launches]
        print(f"\n## `{name}` ({len(launches)} launches captured; duration {min(durs):.4g}–{max(durs):.4g} {units[col['gpu__time_duration.sum']]})\n")
        print("| metric | value |")
        print("|---|---|")
        for k, label in KEYS:
            if k in col and d[col[k]] != "":
                v = d[col[k]]
                try:
                    v = f"{float(v):,.4g}" if "." in v else v
                except ValueError:
                    pass
                print(f"| {label} (`{k}`) | {v} {units[col[k]]} |")
        stalls = []
        for h, i in col.items():
            m = re.match(r"smsp__average_warps_issue_stalled_(\w+)_per_issue_active\.ratio", h)
            if m and d[i] != "":
                stalls.append((float(d[i]), m.group(1)))
        stalls.sort(reverse=True)
        print("\nwarp stalls per issued instruction (top 6): " +
              ", ".join(f"{n} {v:.2f}" for v, n in stalls[:6]))


if __name__ == "__main__":
    main()
