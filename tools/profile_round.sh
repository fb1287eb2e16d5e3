#!/bin/bash
# Round profile on one B200 (run under gpurun): bench lines, ncu launch list, ncu --set full of the step and
# LAUUM kernels, reduced on the box to small text summaries under gpurun_out/ (copy to profiles/).
#   bash tools/profile_round.sh r01g
TAG=${1:-r01}
OUT=gpurun_out
set -x
python bench.py --steps 3 --warmup 3 > $OUT/${TAG}_bench_c2.json 2> $OUT/${TAG}_bench_c2.err; echo "bench c2 rc=$?"
python bench.py --workload ns --experts-per-gpu 256 --steps 2 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_bench_ns256.json 2> $OUT/${TAG}_bench_ns.err; echo "bench ns rc=$?"
python bench.py --workload c3 --experts-per-gpu 96 --steps 2 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_bench_c3_96.json 2> $OUT/${TAG}_bench_c3.err; echo "bench c3 rc=$?"
python bench.py --impl reference --steps 1 --warmup 1 > $OUT/${TAG}_bench_reference_arm.json 2> $OUT/${TAG}_bench_ref.err; echo "reference arm rc=$?"
python tools/ncu_target.py > $OUT/${TAG}_ncu_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/${TAG}_launches.csv python tools/ncu_target.py > $OUT/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 5 -c 1 -o /tmp/${TAG}_step python tools/ncu_target.py > $OUT/ncu_s.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:lauum_grad -s 1 -c 1 -o /tmp/${TAG}_lauum python tools/ncu_target.py > $OUT/ncu_u.log 2>&1
for k in step lauum; do
  ncu -i /tmp/${TAG}_$k.ncu-rep --page details --csv > $OUT/${TAG}_${k}_details.csv 2>/dev/null
  ncu -i /tmp/${TAG}_$k.ncu-rep --page raw --csv > /tmp/${k}_raw.csv 2>/dev/null
  python - <<PY > $OUT/${TAG}_${k}_raw_selected.txt
import csv, re
rows = list(csv.reader(open("/tmp/${k}_raw.csv")))
h, u, v = rows[0], rows[1], rows[2]
pat = r"gpu__time_duration.sum|dram__bytes_read.sum$|dram__bytes_write.sum$|dram__throughput.avg.pct_of_peak_sustained_elapsed$|sm__pipe_tensor_cycles_active_realtime|sm__pipe_fp64_cycles_active_realtime|smsp__pipe_tensor_subpipe_dmma_cycles_active|sm__throughput.avg.pct|smsp__issue_active.avg.pct|launch__registers_per_thread$|launch__grid_size|launch__shared_mem_per_block_dynamic|sm__warps_active.avg.pct|smsp__inst_executed.sum$|lts__t_sector_hit_rate.pct|sm__cycles_elapsed.max|sm__inst_executed_pipe_tensor_subpipe|utcimma_src_int8_sparsity_off.sum.pct_of_peak_sustained_elapsed|sm__inst_executed_pipe_uniform|sm__inst_executed_pipe_tc"
for a, b, c in zip(h, u, v):
    if re.search(pat, a): print(f"{a:95s} {b:14s} {c}")
PY
  ncu -i /tmp/${TAG}_$k.ncu-rep --page source --csv > /tmp/${k}_source.csv 2>/dev/null
  python tools/ncu_source_summary.py /tmp/${k}_source.csv 30 > $OUT/${TAG}_${k}_source_summary.txt
done
cuobjdump -sass gpsat_b200/libgpsat_b200.so 2>/dev/null | grep -oE "UTCIMMA[A-Z0-9_.]*|UTCBAR[A-Z0-9_.]*|UBLKCP[A-Z0-9_.]*|DMMA[A-Z0-9_.]*|LDTM[A-Z0-9_.]*" | sort | uniq -c > $OUT/${TAG}_sass_mnemonics.txt
ls -la $OUT
