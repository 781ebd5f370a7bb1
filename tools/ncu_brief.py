"""Brief of one ncu --set full capture: python tools/ncu_brief.py <report.ncu-rep> — time, dram bytes, pipes, stalls."""
import csv, re, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, u = rows[0], rows[1]
pat = re.compile(r"gpu__time_duration.sum|^dram__bytes_(read|write).sum$|^smsp__inst_executed.sum$|smsp__issue_active.avg.pct|"
                 r"sm__warps_active.avg.pct|registers_per_thread$|occupancy_limit_(registers|shared_mem)|launch__grid_size|"
                 r"thread_inst_executed_per_inst|lts__t_sector_hit_rate.pct|issue_stalled.*per_issue_active.ratio|"
                 r"sm__inst_executed_pipe_(adu|alu|fma|lsu|xu|cbu|uniform).avg.pct_of_peak_sustained_active|"
                 r"l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum|l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct|"
                 r"dram__throughput.avg.pct_of_peak_sustained_elapsed|Kernel Name")
for v in rows[2:]:
    for i, k in enumerate(h):
        if pat.search(k):
            try:
                f = float(v[i].replace(",", ""))
                if "stalled" in k and f < 0.3: continue
            except ValueError:
                pass
            print("%-92s %-10s %s" % (k, u[i], v[i][:110]))
    print()
