"""Reads an .ncu-rep (ncu -i ... --page raw --csv) and prints `metric [unit] = value` lines for the metrics the roofline
discussion uses.  python scripts/ncu_summary.py gpurun_out/prof.ncu-rep ["header line" ...]"""
import csv
import io
import re
import subprocess
import sys

KEEP = re.compile(r"^(dram__bytes_(read|write)\.sum|gpu__dram_throughput|gpu__time_duration\.sum|launch__(block_size|grid_size|"
                  r"registers_per_thread|occupancy_limit|shared_mem_per_block_static)|lts__t_sector_hit_rate\.pct|"
                  r"l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum|sm__inst_executed_pipe_(alu|fma|xu)\.avg\.pct|"
                  r"sm__throughput\.avg\.pct|sm__warps_active\.avg\.pct|smsp__average_warps_issue_stalled_.*_per_issue_active|"
                  r"smsp__thread_inst_executed_per_inst_executed|smsp__issue_active\.avg\.pct|sm__cycles_active\.avg|"
                  r"smsp__inst_executed\.sum$|sm__inst_executed\.sum$)")
rep = sys.argv[1]
for h in sys.argv[2:]:
    print("# " + h)
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
for r in rows[2:]:
    name = r[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
    print("# kernel: " + name)
    for k, u, v in sorted(zip(hdr, units, r)):
        if KEEP.match(k):
            print("%s [%s] = %s" % (k, u, v))
