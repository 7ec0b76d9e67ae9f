#!/bin/bash
# usage: tools/ncu_summary.sh <file.ncu-rep> > profiles/<name>.csv   (per-launch metrics the rooflines cite)
ncu -i "$1" --page raw --csv --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,lts__throughput.avg.pct_of_peak_sustained_elapsed,lts__t_sector_hit_rate.pct,sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed,sm__inst_executed_pipe_tc.sum,sm__inst_executed_pipe_tmem.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed,sm__warps_active.avg.pct_of_peak_sustained_active,launch__grid_size,launch__block_size,launch__registers_per_thread,sm__cycles_elapsed.max,smsp__inst_executed.sum 2>/dev/null \
 | python3 -c "
import csv,sys
r=list(csv.reader(sys.stdin))
drop={'Process ID','Process Name','Host Name','Context','Stream','Device','CC','Section Name','Metric Name'}
keep=[i for i,h in enumerate(r[0]) if h not in drop]
w=csv.writer(sys.stdout)
for row in r: w.writerow([row[i] for i in keep])
"
