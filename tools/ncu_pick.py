"""Print selected metrics of an `ncu --page raw --csv` export: python tools/ncu_pick.py raw.csv [substr ...]"""
import csv
import sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
keys = sys.argv[2:] or ['gpu__time_duration.sum', 'dram__bytes_read.sum ', 'dram__bytes_write.sum ', 'smsp__issue_active.avg.pct',
                        'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum ',
                        'smsp__inst_executed.sum ', 'launch__registers_per_thread ', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum ',
                        'smsp__average_warps_issue_stalled', 'sm__inst_executed_pipe', 'sm__pipe_fma', 'sm__pipe_alu', 'sm__warps_active.avg.pct']
for vals in rows[2:]:
    print("==", vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "")
    for h, u, v in zip(hdr, units, vals):
        if any(k in h + ' ' for k in keys) and '.min' not in h and '.max' not in h and 'elapsed' not in h.replace('l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed', ''):
            print(f"  {h} [{u}] {v}")
