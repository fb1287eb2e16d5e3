#!/bin/bash
# Round profile on one B200 (run under gpurun): bench lines, ncu launch list, ncu --set full of the
# two DMMA kernels, reduced on the box to small text summaries under gpurun_out/ (copy to profiles/).
#   bash tools/profile_round.sh r01c
TAG=${1:-r01}
OUT=gpurun_out
set -x
python bench.py --steps 3 --warmup 3 > $OUT/${TAG}_bench_c2.json 2> $OUT/${TAG}_bench_c2.err; echo "bench c2 rc=$?"
python bench.py --workload ns --experts-per-gpu 256 --steps 2 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_bench_ns256.json 2> $OUT/${TAG}_bench_ns.err; echo "bench ns rc=$?"
python tools/ncu_target.py > $OUT/${TAG}_ncu_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/${TAG}_launches.csv python tools/ncu_target.py > $OUT/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 11 -c 1 -o /tmp/${TAG}_step python tools/ncu_target.py > $OUT/ncu_s.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:lauum_grad_kernel -s 1 -c 1 -o /tmp/${TAG}_lauum python tools/ncu_target.py > $OUT/ncu_u.log 2>&1
for k in step lauum; do
  ncu -i /tmp/${TAG}_$k.ncu-rep --page details --csv > $OUT/${TAG}_${k}_details.csv 2>/dev/null
  ncu -i /tmp/${TAG}_$k.ncu-rep --page raw --csv > /tmp/${k}_raw.csv 2>/dev/null
  python - <<PY > $OUT/${TAG}_${k}_raw_selected.txt
import csv, re
rows = list(csv.reader(open("/tmp/${k}_raw.csv")))
h, u, v = rows[0], rows[1], rows[2]
pat = r"gpu__time_duration.sum|dram__bytes_read.sum$|dram__bytes_write.sum$|sm__pipe_tensor_cycles_active_realtime|sm__pipe_fp64_cycles_active_realtime|smsp__pipe_tensor_subpipe_dmma_cycles_active|sm__throughput.avg.pct|smsp__issue_active.avg.pct|launch__registers_per_thread$|launch__grid_size|launch__shared_mem_per_block_dynamic|sm__warps_active.avg.pct|smsp__inst_executed.sum$|lts__t_sector_hit_rate.pct|sm__cycles_elapsed.max"
for a, b, c in zip(h, u, v):
    if re.search(pat, a): print(f"{a:95s} {b:14s} {c}")
PY
  ncu -i /tmp/${TAG}_$k.ncu-rep --page source --csv > /tmp/${k}_source.csv 2>/dev/null
  python tools/ncu_source_summary.py /tmp/${k}_source.csv 30 > $OUT/${TAG}_${k}_source_summary.txt
done
ls -la $OUT
