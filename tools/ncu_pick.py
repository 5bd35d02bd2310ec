"""Print the handful of ncu raw-page metrics we steer by.  python tools/ncu_pick.py raw.csv [row]"""
import csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
pat = re.compile(r'Kernel Name|gpu__time_duration.sum|pipe_tensor_cycles_active.avg.pct|pipe_xu.avg.pct_of_peak_sustained_active|pipe_fma.avg.pct.*active'
                 r'|pipe_alu.avg.pct.*active|mem_shared.*pct|sm__throughput.avg.pct|lts__throughput.avg.pct|dram__bytes_(read|write).sum$'
                 r'|smsp__inst_executed.sum$|sm__cycles_elapsed.avg$|registers_per_thread|smsp__cycles_active.avg$|l1tex__throughput.avg.pct'
                 r'|smsp__issue_active.avg.pct|l1tex__data_pipe.*pct|lts__t_sectors.*(red|atom).*sum$|dram__throughput.avg.pct|bank_conflicts.*sum$')
r = rows[2 + which]
for k in hdr:
    if pat.search(k):
        print(k, r[hdr.index(k)], units[hdr.index(k)])
