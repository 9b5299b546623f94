"""usage: ncu -i X.ncu-rep --page raw --csv > raw.csv; python tools/ncu_summary.py raw.csv "<header note>" > profiles/NAME.txt"""
import csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
want = re.compile(r'^(Kernel Name|Grid Size|Block Size)$|dram__bytes_(read|write)\.sum$|gpu__time_duration\.sum|lts__t_sector_hit_rate\.pct|l1tex__t_sector_hit_rate\.pct|sm__warps_active\.avg\.pct|launch__registers_per_thread$|launch__shared_mem_per_block_static|launch__occupancy_limit|smsp__thread_inst_executed_per_inst|smsp__inst_executed\.sum$|smsp__issue_active\.avg\.pct|smsp__average_warps_issue_stalled.*_per_issue_active|dram__throughput\.avg\.pct|lts__throughput\.avg\.pct|sm__throughput\.avg\.pct|lts__t_bytes\.sum$|sm__icc_request_hit_rate|sm__icc_requests\.sum$|gcc__cache_requests_type_instruction\.sum|sm__inst_executed_pipe_(alu|fma|lsu|xu)\.avg\.pct_of_peak_sustained_active|l1tex__data_pipe_lsu_wavefronts(_mem_shared|_mem_lgds)?\.avg\.pct|l1tex__data_pipe_lsu_wavefronts(_mem_shared|_mem_lgds)?\.sum$')
print("#", sys.argv[2] if len(sys.argv) > 2 else "")
for vals in rows[2:]:
    print("----")
    for h, u, v in zip(hdr, units, vals):
        if want.search(h):
            print(f"{h} [{u}] = {v}")
