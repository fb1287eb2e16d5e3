#!/bin/bash
# ncu evidence for the int8-path gradient kernels on one B200 (run under gpurun), reduced to small text.
#   bash tools/profile_i8.sh r01f
TAG=${1:-r01f}
OUT=gpurun_out
set -x
python tools/ncu_target.py 296 800 1200 2 > $OUT/${TAG}_i8_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"lauum|u_slice|u_rowmax" --csv --log-file $OUT/${TAG}_i8_launches.csv python tools/ncu_target.py 296 800 1200 2 > $OUT/ncu_il.log 2>&1
for k in lauum_grad_i8p_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 1 -o /tmp/${TAG}_$k python tools/ncu_target.py 296 800 1200 2 > $OUT/ncu_$k.log 2>&1
  ncu -i /tmp/${TAG}_$k.ncu-rep --page raw --csv > /tmp/${k}_raw.csv 2>/dev/null
  python - <<PY > $OUT/${TAG}_${k}_raw_selected.txt
import csv, re
rows = list(csv.reader(open("/tmp/${k}_raw.csv")))
h, u, v = rows[0], rows[1], rows[2]
pat = r"gpu__time_duration.sum|dram__bytes_read.sum$|dram__bytes_write.sum$|dram__throughput.avg.pct|sm__pipe_tensor|sm__inst_executed_pipe_tensor|sm__pipe_fp64_cycles_active_realtime|sm__throughput.avg.pct|smsp__issue_active.avg.pct|launch__registers_per_thread$|launch__grid_size|launch__shared_mem_per_block_dynamic|sm__warps_active.avg.pct|smsp__inst_executed.sum$|lts__t_sector_hit_rate.pct|lts__throughput.avg.pct|sm__cycles_elapsed.max|l1tex__throughput.avg.pct|smsp__average_warp.*stall|smsp__warp_issue_stalled.*per_warp_active"
for a, b, c in zip(h, u, v):
    if re.search(pat, a): print(f"{a:95s} {b:14s} {c}")
PY
  ncu -i /tmp/${TAG}_$k.ncu-rep --page source --csv > /tmp/${k}_source.csv 2>/dev/null
  python tools/ncu_source_summary.py /tmp/${k}_source.csv 30 > $OUT/${TAG}_${k}_source_summary.txt
done
ls -la $OUT
