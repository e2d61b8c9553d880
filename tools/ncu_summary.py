"""Selected metrics of one kernel from `ncu -i X.ncu-rep --page raw --csv` -> metric,unit,value rows (profiles/*.csv)."""
import csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
kernel_re = re.compile(sys.argv[2])
hdr, units = rows[0], rows[1]
ki = hdr.index("Kernel Name")
keep = re.compile(r"^(gpu__time_duration|dram__bytes_(read|write)\.sum|dram__throughput|lts__t_(bytes|sectors)(_op_(read|write|red|atom))?\.sum|lts__t_sector_hit_rate|"
                  r"l1tex__data_bank_conflicts|l1tex__t_bytes|sm__throughput|sm__inst_executed_pipe_|sm__pipe_|sm__inst_executed\.sum|sm__inst_issued|"
                  r"sm__issue_active|smsp__issue_active|smsp__inst_executed\.sum|sm__warps_active|sm__cycles_active|sm__cycles_elapsed|launch__|"
                  r"smsp__average_warp.*issue_stalled.*ratio$|smsp__warp_issue_stalled.*_per_warp_active|sm__sass_inst_executed_op_(shared|global)|TPC)")
for r in rows[2:]:
    if len(r) <= ki or not kernel_re.search(r[ki]):
        continue
    print("metric,unit,value")
    print(f"kernel,,\"{r[ki]}\"")
    for name, unit, val in zip(hdr, units, r):
        if keep.search(name):
            print(f"{name},{unit},{val}")
    break
