#!/usr/bin/env python
"""Condense an .ncu-rep into the text summary kept under profiles/:
   python tools/ncu_summary.py gpurun_out/x.ncu-rep profiles/x_raw_metrics.txt [launch-index]
Keeps launch geometry, occupancy limits, duration, DRAM/L2/shared traffic, pipe utilisation, issue/warp-state figures."""
import csv
import io
import re
import subprocess
import sys

KEEP = re.compile(r"^(gpu__time_duration|dram__bytes_(read|write)\.sum($|\.per_second)|launch__(block_size|grid_size|"
                  r"registers_per_thread$|shared_mem_per_block$|shared_mem_per_block_dynamic|occupancy_limit|waves_per_multiprocessor)|"
                  r"sm__inst_executed_pipe_(alu|fma|fp64|lsu|xu|adu|cbu|uniform)\.avg\.pct_of_peak_sustained_active|"
                  r"sm__inst_executed\.(sum|avg\.per_cycle_active)|sm__inst_issued\.avg\.(pct_of_peak_sustained_active|per_cycle_active)|"
                  r"sm__warps_active\.avg\.pct_of_peak_sustained_active|sm__throughput|sm__cycles_(active|elapsed)\.(avg|max)$|"
                  r"smsp__issue_active\.avg\.(pct|per_cycle_active)|smsp__thread_inst_executed_per_inst_executed\.ratio|"
                  r"smsp__average_warps?_issue_stalled_[a-z_]+_per_issue_active\.ratio|smsp__average_warp_latency_issue_stalled|"
                  r"l1tex__data_bank_conflicts_pipe_lsu_mem_shared(_op_(ld|st|atom))?\.sum|l1tex__data_pipe_lsu_wavefronts_mem_shared(_op_(ld|st|atom))?\.sum|"
                  r"smsp__inst_executed_op_shared_(ld|st|atom)\.sum|lts__t_bytes\.sum$|lts__t_sector_hit_rate|l1tex__t_sector_hit_rate|"
                  r"sm__sass_thread_inst_executed_op_(d|f)(add|mul|fma)_pred_on\.sum|achieved_occupancy|sm__maximum_warps_per_active_cycle_pct)")


def main():
    rep, out = sys.argv[1], sys.argv[2]
    idx = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units, data = rows[0], rows[1], rows[2 + idx]
    with open(out, "w") as f:
        for h, u, v in zip(hdr, units, data):
            if h in ("Kernel Name", "Block Size", "Grid Size") or KEEP.search(h):
                f.write(f"{h} {u} {v}\n".replace("  ", " "))
    print(open(out).read())


if __name__ == "__main__":
    main()
