"""Summarise an .ncu-rep (read with `ncu -i <rep> --page raw --csv`) into the per-kernel text form kept under
profiles/: python tools/ncu_summary.py <rep> [metric-regex]  (default: the metrics DESIGN.md quotes)."""
import csv, io, re, subprocess, sys
rep = sys.argv[1]
pat = re.compile(sys.argv[2] if len(sys.argv) > 2 else
                 r"dram__bytes_(read|write)\.sum$|gpu__time_duration\.sum|dram_throughput.*elapsed|l1tex__data_pipe_lsu_wavefronts|"
                 r"l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum$|launch__(grid|block)_size|launch__registers|launch__shared_mem_per_block_dynamic|"
                 r"launch__occupancy_limit|lts__t_sector_hit_rate|lts__throughput.*elapsed|sm__inst_executed\.avg\.per_cycle_active|"
                 r"sm__inst_executed_pipe_lsu.*active|sm__pipe_fp64_cycles_active.*active|sm__warps_active.*active|sm__cycles_active\.avg$|"
                 r"smsp__average_warps_issue_stalled.*per_issue_active|sm__throughput.*elapsed|smsp__inst_executed\.sum$|sm__sass_inst_executed_op_shared")
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
for r in rows[2:]:
    name = r[hdr.index("Kernel Name")]
    print("==", name, "grid", r[hdr.index("Grid Size")], "block", r[hdr.index("Block Size")])
    for h, u, v in zip(hdr, units, r):
        if pat.search(h):
            print(f"{h} [{u}] = {v}")
