#!/bin/bash
# Round-2 profile on one B200 (run under gpurun): ncu launch list of the bench command, ncu --set full of step_kernel,
# lauum_grad_i8_kernel and of the HBM-bound kernels (predict_reduce_kernel, kuf_build_kernel, vff_kuf_build_kernel),
# reduced on the box to small text summaries under gpurun_out/ (copy to profiles/).
#   bash tools/profile_r02.sh r02e
TAG=${1:-r02}
OUT=gpurun_out
mkdir -p $OUT
set -x
# the bench command itself (one step after one warm-up, no CPU arm / extras) under the launch-list pass
python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-extra --no-e2e > $OUT/${TAG}_bench_short.json 2> $OUT/${TAG}_bench_short.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $OUT/${TAG}_bench_launches.csv \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-extra --no-e2e > $OUT/ncu_b.log 2>&1
python tools/ncu_target.py > $OUT/${TAG}_ncu_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 5 -c 1 -o /tmp/${TAG}_step python tools/ncu_target.py > $OUT/ncu_s.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:lauum_grad_i8 -s 1 -c 1 -o /tmp/${TAG}_lauum python tools/ncu_target.py > $OUT/ncu_u.log 2>&1
ncu --set full --clock-control none -k regex:predict_reduce_kernel -s 1 -c 1 -o /tmp/${TAG}_predred python tools/ncu_target.py 296 2000 2000 1 > $OUT/ncu_p.log 2>&1
python tools/ncu_target_sgpr.py > $OUT/${TAG}_sgpr_plain.log 2>&1 && \
ncu --set full --clock-control none -k regex:kuf_build_kernel -s 1 -c 1 -o /tmp/${TAG}_kuf python tools/ncu_target_sgpr.py > $OUT/ncu_k.log 2>&1
python tools/ncu_target_vff.py > $OUT/${TAG}_vff_plain.log 2>&1 && \
ncu --set full --clock-control none -k regex:vff_kuf_build_kernel -s 0 -c 1 -o /tmp/${TAG}_vffkuf python tools/ncu_target_vff.py > $OUT/ncu_v.log 2>&1
for k in step lauum predred kuf vffkuf; do
  [ -f /tmp/${TAG}_$k.ncu-rep ] || continue
  ncu -i /tmp/${TAG}_$k.ncu-rep --page raw --csv > /tmp/${k}_raw.csv 2>/dev/null
  python - <<PY > $OUT/${TAG}_${k}_raw_selected.txt
import csv, re
rows = list(csv.reader(open("/tmp/${k}_raw.csv")))
h, u, v = rows[0], rows[1], rows[2]
pat = r"Kernel Name|gpu__time_duration.sum|dram__bytes_read.sum$|dram__bytes_write.sum$|dram__throughput.avg.pct_of_peak_sustained_elapsed$|sm__pipe_tensor_cycles_active_realtime|sm__pipe_fp64_cycles_active_realtime|smsp__pipe_tensor_subpipe_dmma_cycles_active|sm__throughput.avg.pct|smsp__issue_active.avg.pct|launch__registers_per_thread$|launch__grid_size|launch__block_size|launch__shared_mem_per_block_dynamic|sm__warps_active.avg.pct|smsp__inst_executed.sum$|lts__t_sector_hit_rate.pct|sm__cycles_elapsed.max|sm__inst_executed_pipe_tensor_subpipe|utcimma_src_int8_sparsity_off.sum.pct_of_peak_sustained_elapsed|sm__inst_executed_pipe_uniform|sm__inst_executed_pipe_tc|lts__t_bytes.sum$|l1tex__t_bytes.sum$"
for a, b, c in zip(h, u, v):
    if re.search(pat, a): print(f"{a:95s} {b:14s} {c}")
PY
done
for k in step lauum; do
  ncu -i /tmp/${TAG}_$k.ncu-rep --page details --csv > $OUT/${TAG}_${k}_details.csv 2>/dev/null
  ncu -i /tmp/${TAG}_$k.ncu-rep --page source --csv > /tmp/${k}_source.csv 2>/dev/null
  python tools/ncu_source_summary.py /tmp/${k}_source.csv 30 > $OUT/${TAG}_${k}_source_summary.txt
done
# achieved HBM GB/s of the memory-bound kernels from the captures: (dram read + write) / duration
python - <<PY > $OUT/${TAG}_hbm_kernels.json
import json, re
out = {}
for k, name in (("predred", "predict_reduce_kernel"), ("kuf", "kuf_build_kernel"), ("vffkuf", "vff_kuf_build_kernel")):
    try:
        txt = open("$OUT/${TAG}_%s_raw_selected.txt" % k).read()
    except OSError:
        continue
    def val(metric):
        m = re.search(r"^%s\s+(\S+)\s+(\S+)" % re.escape(metric), txt, re.M)
        if not m: return None
        x = float(m.group(2).replace(",", "")); u = m.group(1)
        scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1e-3, "us": 1e-6, "ns": 1e-9, "s": 1.0, "usecond": 1e-6, "msecond": 1e-3, "nsecond": 1e-9, "second": 1.0}.get(u)
        return None if scale is None else x * scale
    rd, wr, t = val("dram__bytes_read.sum"), val("dram__bytes_write.sum"), val("gpu__time_duration.sum")
    if None in (rd, wr, t): continue
    out[name] = {"dram_bytes_read": rd, "dram_bytes_write": wr, "duration_s": t, "achieved_GBps": (rd + wr) / t / 1e9,
                 "peak_GBps": 6541.8, "frac": (rd + wr) / t / 1e9 / 6541.8,
                 "note": "one launch under ncu --set full --clock-control none (cold cache, serialised)"}
print(json.dumps(out, indent=1))
PY
cuobjdump -sass gpsat_b200/libgpsat_b200.so 2>/dev/null | grep -oE "UTCIMMA[A-Z0-9_.]*|UTCBAR[A-Z0-9_.]*|UBLKCP[A-Z0-9_.]*|DMMA[A-Z0-9_.]*|LDTM[A-Z0-9_.]*" | sort | uniq -c > $OUT/${TAG}_sass_mnemonics.txt
ls -la $OUT
