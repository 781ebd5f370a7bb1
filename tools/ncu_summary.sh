#!/bin/bash
# Summary of one `ncu --set full --import-source on` capture of findNeighborsKernel: headline metrics + per-line shares.
# usage: tools/ncu_summary.sh <report.ncu-rep> <lib with the same build> "<header line>"
set -e
rep=$1; lib=$(realpath $2); hdr=$3
tmp=$(mktemp -d)
ncu -i $rep --page raw --csv 2>/dev/null > $tmp/raw.csv
ncu -i $rep --page source --csv 2>/dev/null > $tmp/sass.csv
(cd $tmp && cuobjdump -xelf neighbors $lib >/dev/null)
echo "# $hdr"
python - $tmp/raw.csv <<'PY'
import csv,sys
rows=list(csv.reader(open(sys.argv[1])))
h,u,v=rows[0],rows[1],rows[2]
keys=['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','smsp__inst_executed.sum','smsp__issue_active.avg.pct_of_peak_sustained_active','sm__warps_active.avg.pct_of_peak_sustained_active','launch__registers_per_thread','launch__grid_size','launch__block_size','launch__occupancy_limit_registers','launch__occupancy_limit_shared_mem','smsp__thread_inst_executed_per_inst_executed.ratio','lts__t_sector_hit_rate.pct','l1tex__t_sector_hit_rate.pct','sm__throughput.avg.pct_of_peak_sustained_elapsed','dram__throughput.avg.pct_of_peak_sustained_elapsed','sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active','smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio','smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio','smsp__average_warps_issue_stalled_wait_per_issue_active.ratio','smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio','smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio','smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio','smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio','smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio']
for k in keys:
    if k in h:
        i=h.index(k); print("%-90s %-16s %s"%(k,u[i],v[i]))
PY
echo
echo "# per source line (tools/ncu_lines.py on --page source --csv; cubin of the same build)"
python tools/ncu_lines.py $tmp/sass.csv $tmp/neighbors.sm_100a.cubin ${KERN:-_ZN3csb19findNeighborsKernelIdLb0ELb1EEEvNS_8NbParamsIT_EE} 45 | cut -c1-170
rm -rf $tmp
