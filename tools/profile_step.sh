#!/bin/bash
# ncu --set full of one step_kernel launch (block column SKIP of the first evaluation) for the path selected by
# GPSAT_B200_STEP, reduced on the box to text summaries under gpurun_out/ (copy to profiles/).
#   GPSAT_B200_STEP=i8 bash tools/profile_step.sh r01g_i8 5
TAG=${1:-r01g}
SKIP=${2:-5}
OUT=gpurun_out
python tools/ncu_target.py > $OUT/${TAG}_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/${TAG}_launches.csv python tools/ncu_target.py > $OUT/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:step_kernel -s $SKIP -c 1 -o /tmp/${TAG}_step python tools/ncu_target.py > $OUT/ncu_s.log 2>&1
k=step
ncu -i /tmp/${TAG}_$k.ncu-rep --page details --csv > $OUT/${TAG}_${k}_details.csv 2>/dev/null
ncu -i /tmp/${TAG}_$k.ncu-rep --page raw --csv > /tmp/${k}_raw.csv 2>/dev/null
python - <<PY > $OUT/${TAG}_${k}_raw_selected.txt
import csv, re
rows = list(csv.reader(open("/tmp/${k}_raw.csv")))
h, u, v = rows[0], rows[1], rows[2]
pat = r"gpu__time_duration.sum|dram__bytes_read.sum$|dram__bytes_write.sum$|dram__throughput.avg.pct|lts__throughput.avg.pct|sm__pipe_tensor_cycles_active_realtime|sm__pipe_fp64_cycles_active_realtime|smsp__pipe_tensor_subpipe_dmma_cycles_active|sm__throughput.avg.pct|smsp__issue_active.avg.pct|launch__registers_per_thread$|launch__grid_size|launch__shared_mem_per_block_dynamic|sm__warps_active.avg.pct|smsp__inst_executed.sum$|lts__t_sector_hit_rate.pct|sm__cycles_elapsed.max|sm__inst_executed_pipe_tensor|tensor_op"
for a, b, c in zip(h, u, v):
    if re.search(pat, a): print(f"{a:95s} {b:14s} {c}")
PY
ncu -i /tmp/${TAG}_$k.ncu-rep --page source --csv > /tmp/${k}_source.csv 2>/dev/null
python tools/ncu_source_summary.py /tmp/${k}_source.csv 30 > $OUT/${TAG}_${k}_source_summary.txt
