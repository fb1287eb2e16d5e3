#!/bin/bash
# ncu evidence for the sparse-GPR GEMM kernel on one B200 (run under gpurun), reduced to small text.
#   bash tools/profile_sgpr.sh r01e
TAG=${1:-r01e}
OUT=gpurun_out
set -x
python tools/ncu_target_sgpr.py > $OUT/${TAG}_sgpr_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/${TAG}_sgpr_launches.csv python tools/ncu_target_sgpr.py > $OUT/ncu_sl.log 2>&1
# second evaluation: launch 8 + 1 = SYRK, 8 + 4 = GC (gradient contraction epilogue)
ncu --set full --clock-control none --import-source on -k regex:sgpr_gemm_kernel -s 9 -c 1 -o /tmp/${TAG}_syrk python tools/ncu_target_sgpr.py > $OUT/ncu_ss.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:sgpr_gemm_kernel -s 12 -c 1 -o /tmp/${TAG}_gc python tools/ncu_target_sgpr.py > $OUT/ncu_sg.log 2>&1
for k in syrk gc; do
  ncu -i /tmp/${TAG}_$k.ncu-rep --page raw --csv > /tmp/${k}_raw.csv 2>/dev/null
  python - <<PY > $OUT/${TAG}_sgpr_${k}_raw_selected.txt
import csv, re
rows = list(csv.reader(open("/tmp/${k}_raw.csv")))
h, u, v = rows[0], rows[1], rows[2]
pat = r"gpu__time_duration.sum|dram__bytes_read.sum$|dram__bytes_write.sum$|sm__pipe_tensor_cycles_active_realtime|sm__pipe_fp64_cycles_active_realtime|smsp__pipe_tensor_subpipe_dmma_cycles_active|sm__throughput.avg.pct|smsp__issue_active.avg.pct|launch__registers_per_thread$|launch__grid_size|launch__shared_mem_per_block_dynamic|sm__warps_active.avg.pct|smsp__inst_executed.sum$|lts__t_sector_hit_rate.pct|sm__cycles_elapsed.max"
for a, b, c in zip(h, u, v):
    if re.search(pat, a): print(f"{a:95s} {b:14s} {c}")
PY
  ncu -i /tmp/${TAG}_$k.ncu-rep --page source --csv > /tmp/${k}_source.csv 2>/dev/null
  python tools/ncu_source_summary.py /tmp/${k}_source.csv 25 > $OUT/${TAG}_sgpr_${k}_source_summary.txt
done
ls -la $OUT
