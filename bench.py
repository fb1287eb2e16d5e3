#!/usr/bin/env python
"""bench.py - local experts/sec (fit + predict, FP64) on N B200s of one node.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2]

One "step" = one full pass of the hot path over one batch of synthetic experts: batched
L-BFGS hyper-parameter fit (every objective evaluation = kernel build + Cholesky + inverse +
analytic gradient) followed by prediction at the optimised hyper-parameters.
Workload at N=1: BASELINE.json configs[1] (`configs/example_local_expert_oi.json` shape,
~500 experts x ~1k observations).  For N>1 every rank gets its own 500-expert shard (weak
scaling, expert seeds continue across ranks), ranks run independently and rank 0 gathers the
per-expert results once per step.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "local experts/sec (fit+predict, FP64)"
UNIT = "experts/s"
FP64_PEAK_FILE = os.path.join(ROOT, "profiles", "r01_fp64_peak.json")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c1", "c2", "c3", "ns"])
    ap.add_argument("--experts-per-gpu", type=int, default=0, help="override the workload's expert count")
    ap.add_argument("--cpu-sample", type=int, default=0, help="experts in the CPU baseline sample (0 = one per core)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def workload_spec(args):
    from gpsat_b200 import synthetic as syn
    from dataclasses import replace
    spec = syn.WORKLOADS[args.workload]
    if args.experts_per_gpu > 0:
        spec = replace(spec, E=args.experts_per_gpu)
    return spec


def fp64_peak_tflops():
    """FP64 pipe peak measured on this pool by tools/fp64_peak.cu (MEASURED_PEAKS.json has no FP64
    entry and the profiling guide states no FP64 fallback)."""
    try:
        d = json.load(open(FP64_PEAK_FILE))
        return float(d["dmma_m8n8k4_sustained3s_tflops"]), "measured: tools/fp64_peak.cu DMMA m8n8k4 sustained 3 s (profiles/r01_fp64_peak.json)"
    except Exception:
        return 37.1, "fallback: 148 SM x 64 FMA/clk x 2 x 1.965 GHz"


def arithmetic_note():
    step = os.environ.get("GPSAT_B200_STEP", "i8") != "dmma"
    lauum = os.environ.get("GPSAT_B200_LAUUM", "i8") != "dmma"
    if not step and not lauum:
        return "FP64 mma.sync throughout"
    parts = []
    if step:
        parts.append("the k-sums of the potrf / trtri tiles are FP64 values split into 8 exact int8 slices with a-priori "
                     "row scales and multiplied on the int8 tensor path (tcgen05, TMEM accumulators; truncation at "
                     "2^-56 of the row scales, the level of the FP64 accumulation error)")
    if lauum:
        parts.append("the K^-1 tile products of the gradient use 7 slices (error 5e-14 of sum|a||b|)")
    return ("FP64 data and results; " + "; ".join(parts) + "; covariance evaluation, diagonal-tile factorisation, "
            "C W^T / SYRK epilogues, L-BFGS and prediction reductions in FP64; GPSAT_B200_STEP=dmma "
            "GPSAT_B200_LAUUM=dmma run everything as FP64 mma.sync")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.lines, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def cpu_jobs(spec, experts, seed=0):
    from gpsat_b200 import synthetic as syn
    th0, tk, ta, tb, tr = syn.default_fit_setup(spec, 1)
    jobs = []
    for e in experts:
        X, y, Xs = syn.make_expert(spec, e, seed)
        jobs.append((spec.kernel, X, y, Xs, th0[0], tk[0], ta[0], tb[0], tr[0], 10_000))
    return jobs


def run_reference(args):
    """--impl reference: the oracle port of the reference path on all host cores (the reference
    itself - GPflow/TensorFlow - is not installable here)."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    from oracle.cpu_baseline import CpuPool
    spec = workload_spec(args)
    cores = os.cpu_count() or 1
    sample = args.cpu_sample or cores
    pool = CpuPool(min(cores, sample))
    times, done = [], 0
    for s in range(args.warmup + args.steps):
        experts = [(s * sample + k) % spec.E for k in range(sample)]
        res, wall = pool.run(cpu_jobs(spec, experts))
        if s >= args.warmup:
            times.append(wall); done += len(res)
    pool.close()
    total = float(np.sum(times))
    value = done / total
    blas = CpuPool.blas_report(res)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / max(args.steps, 1),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": spec.name, "experts_per_step": sample, "kernel": spec.kernel,
                       "n_obs": [spec.n_lo, spec.n_hi], "n_pred": [spec.m_lo, spec.m_hi], "optimise": True},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": min(cores, sample), "kind": "port",
                             "sample": f"{sample} experts of {spec.name} per step, NumPy/SciPy oracle, "
                                       f"one process per core, 1 BLAS thread each", "blas": blas},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def run_ours(args):
    import torch
    import torch.distributed as dist
    from gpsat_b200 import parallel, synthetic as syn
    from gpsat_b200.engine import Engine

    rank, world, local = parallel.init_process_group()
    assert world == args.gpus or world == 1, f"WORLD_SIZE={world} but --gpus {args.gpus}"
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py --impl ours needs a B200: the CUDA path has no CPU fallback")
    torch.cuda.set_device(local)
    spec = workload_spec(args)
    E = spec.E
    experts = range(rank * E, (rank + 1) * E)          # weak scaling: a fresh shard per rank
    host = syn.make_batch(spec, seed=0, experts=experts, with_pred=True)
    th0, tk, ta, tb, tr = syn.default_fit_setup(spec, E)

    # optional CPU baseline (rank 0, N=1 only) - before CUDA is busy, worker processes are spawned
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle.cpu_baseline import CpuPool
        cores = os.cpu_count() or 1
        sample = args.cpu_sample or cores
        pool = CpuPool(min(cores, sample))
        res, wall = pool.run(cpu_jobs(spec, range(sample)))
        pool.close()
        cpu_baseline = {"value": len(res) / wall, "unit": UNIT, "cores": min(cores, sample), "kind": "port",
                        "sample": f"first {sample} experts of {spec.name}, NumPy/SciPy oracle (SciPy L-BFGS-B + "
                                  f"LAPACK), one process per core, 1 BLAS thread each, {wall:.1f} s wall",
                        "blas": CpuPool.blas_report(res),      # raises unless every BLAS in every worker had 1 thread
                        "mean_evals_per_expert": float(np.mean([r["nev"] for r in res]))}
        cpu_check = res

    eng = Engine(local)
    stream = torch.cuda.ExternalStream(eng.stream_ptr, device=f"cuda:{local}")
    dev = host.to_device(f"cuda:{local}")
    d_args = [torch.as_tensor(a, device=f"cuda:{local}") for a in (th0, tk, ta, tb, tr)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=f"cuda:{local}")   # > 126 MB L2

    def pin(a):
        return None if a is None else torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()

    host.X, host.y, host.Xs = pin(host.X), pin(host.y), pin(host.Xs)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_device(collect=None):
        with torch.cuda.stream(stream):
            flush.zero_()
        res = eng.fit(dev, *d_args)
        if collect is not None:
            s = eng.stats(); collect["fit"].append(s)
        pred = eng.predict(dev, res.theta)
        if collect is not None:
            s = eng.stats(); collect["predict"].append(s)
        return res, pred

    def step_e2e():
        res = eng.fit(host, th0, tk, ta, tb, tr)
        s1 = eng.stats()
        pred = eng.predict(host, res.theta)
        s2 = eng.stats()
        if world > 1:   # the single final gather
            parallel.gather_results({"theta": res.theta, "lml": res.lml, "status": res.status},
                                    np.arange(rank * E, (rank + 1) * E), world * E)
        return res, pred, s1["h2d_bytes"] + s2["h2d_bytes"], s1["d2h_bytes"] + s2["d2h_bytes"]

    # ---- device-resident timing (value) ----
    for _ in range(args.warmup):
        step_device()
    eng.set_profile(True)
    collect = {"fit": [], "predict": []}
    barrier()
    clocks = ClockSampler(local)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        res, pred = step_device(collect)
    e1.record(stream)
    barrier()
    clk = clocks.stop()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    eng.set_profile(False)
    value = world * E * args.steps / (ms_total * 1e-3)

    # ---- per-kernel roofline numbers from the timed region ----
    n = host.n.astype(np.float64)
    m = host.m.astype(np.float64)
    nev = res.n_evals.cpu().numpy().astype(np.float64)
    # algorithmic FP64 flops of the step kernels: 2 n^3/3 per objective evaluation
    # (potrf + trtri tiles), n^3/3 + n^2 m per prediction pass
    flops_step = args.steps * float(np.sum(nev * 2.0 * n ** 3 / 3.0) + np.sum(n ** 3 / 3.0 + n ** 2 * m))
    ms_step = sum(s["ms_step"] for k in collect for s in collect[k])
    n_step = sum(s["n_step"] for k in collect for s in collect[k])
    ms_lauum = sum(s["ms_lauum"] for s in collect["fit"])
    ms_diag = sum(s["ms_diag"] for k in collect for s in collect[k])
    ms_opt = sum(s["ms_opt"] for s in collect["fit"])
    launches = sum(s["kernel_launches"] for k in collect for s in collect[k])
    peak, peak_src = fp64_peak_tflops()
    achieved = flops_step / (ms_step * 1e-3) / 1e12
    total_flops = args.steps * float(np.sum(nev * n ** 3) + np.sum(n ** 3 / 3.0 + n ** 2 * m))
    traffic, traffic_src = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        traffic, traffic_src = tj["dram_bytes_read"] + tj["dram_bytes_write"], tj["source"]
    except Exception:
        pass
    # executed int8 work of the sliced k-sums: 36 slice-pair MMAs of 2 * 128^3 ops per k-block of a
    # potrf / trtri / diagonal-sum tile
    int8 = None
    if os.environ.get("GPSAT_B200_STEP", "i8") != "dmma":
        T = np.ceil((n + 1) / 128.0)
        TR = np.ceil((n + 1 + m) / 128.0)
        kb_pot = T * (T - 1) * (T - 2) / 6.0
        kb_tr = (T - 1) * T * (T + 1) / 6.0
        kb_dg = np.maximum((T - 1) * (T - 2) / 2.0, 0.0)
        kb_pred = kb_pot + (TR - T) * T * (T - 1) / 2.0 + kb_dg
        ops = args.steps * 2.0 * 128.0 ** 3 * 36.0 * float(np.sum(nev * (kb_pot + kb_tr + kb_dg)) + np.sum(kb_pred))
        try:
            pk = json.load(open(os.path.join(ROOT, "profiles", "r01f_i8_rate.json")))
            i8peak = float(pk["rate"]["n128"]["int8_tops"]) / 1e3
        except Exception:
            i8peak = 3.48
        int8 = {"achieved": ops / (ms_step * 1e-3) / 1e15, "peak": i8peak, "unit": "POP/s",
                "frac": ops / (ms_step * 1e-3) / 1e15 / i8peak,
                "note": "executed int8 tensor ops of step_kernel's sliced k-sums / its CUDA-event time (which also holds "
                        "the FP64 epilogues); peak = measured tcgen05 kind::i8 rate at M = N = 128 (profiles/r01f_i8_rate.json)"}
    roofline = {"bound": "tensor", "kernel": "step_kernel (potrf/trtri tiles: int8-sliced tcgen05 main loops, FP64 DMMA epilogues, fused diagonal factorisation)"
                if os.environ.get("GPSAT_B200_STEP", "i8") != "dmma" else "step_kernel (FP64 DMMA potrf/trtri tiles + fused diagonal factorisation)",
                "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak, "traffic": traffic,
                "traffic_source": traffic_src, "peak_source": peak_src,
                "launches": int(n_step), "avg_launch_ms": ms_step / max(n_step, 1),
                "int8_tensor": int8,
                "note": "achieved = algorithmic FP64 flops / kernel time, peak = measured FP64 (DMMA) pipe peak; with the "
                        "k-sums on the int8 tensor path the fraction is an FP64-equivalent rate and may exceed 1"
                        if int8 else None,
                "whole_job_tflops": total_flops / (ms_total * 1e-3) / 1e12,
                "whole_job_frac": total_flops / (ms_total * 1e-3) / 1e12 / peak,
                "kernel_share_of_step": {"step_kernel": ms_step / ms_total, "lauum_grad_kernel": ms_lauum / ms_total,
                                         "diag_kernel": ms_diag / ms_total, "lbfgs_kernel": ms_opt / ms_total}}

    # ---- end to end through the host-facing call ----
    e2e = None
    if not args.no_e2e:
        step_e2e()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            res_h, pred_h, h2d, d2h = step_e2e()
        barrier()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=f"cuda:{local}")
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": world * E * args.steps / float(dt.item()), "unit": UNIT,
               "h2d_bytes_per_step": float(h2d), "d2h_bytes_per_step": float(d2h)}
        # device-resident and host paths are the same arithmetic: results must agree bit for bit
        assert np.array_equal(res_h.theta, res.theta.cpu().numpy())

    if rank == 0:
        if cpu_baseline is not None:   # the GPU results agree with the CPU sample (checker role of the oracle)
            th = res.theta.cpu().numpy()
            lm = res.lml.cpu().numpy()
            dth, dth_all, dl, pinned = 0.0, 0.0, 0.0, 0
            for k, r in enumerate(cpu_check):
                if r["ok"] and int(res.status[k]) == 0:
                    rel = np.abs(th[k] - r["theta"]) / np.abs(r["theta"])
                    dth_all = max(dth_all, float(np.max(rel)))
                    # components pinned at a bound +- tol (tol = 1e-2, local_experts.py:1115) are skipped
                    sig = tk[k] == 1
                    at_bound = sig & ((r["theta"] >= tb[k] - 1e-2) | (r["theta"] <= ta[k] + 1e-2))
                    pinned += int(at_bound.sum())
                    if (~at_bound).any():
                        dth = max(dth, float(np.max(rel[~at_bound])))
                    dl = max(dl, abs(lm[k] - r["lml"]) / abs(r["lml"]))
            cpu_baseline["parity_vs_gpu"] = {"experts": len(cpu_check), "max_rel_theta": dth, "max_rel_lml": dl,
                                             "max_rel_theta_incl_pinned": dth_all, "components_pinned_at_bound": pinned}
            assert dl < 1e-4, f"optimised LML differs from the CPU oracle: {dl}"
            assert dth < 1e-4, f"optimised hyper-parameters differ from the CPU oracle: {dth}"
        st = np.bincount(res.status.cpu().numpy(), minlength=4)
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": spec.name, "experts_per_gpu": E, "kernel": spec.kernel,
                           "n_obs": [spec.n_lo, spec.n_hi], "n_pred": [spec.m_lo, spec.m_hi], "optimise": True,
                           "parallelism": f"experts sharded x{world}, no collective on the hot path",
                           "l2": "256 MB flush written between steps; per-step working set "
                                 f"{int(collect['fit'][-1]['slots'])} slot matrices >> 126 MB L2",
                           "arithmetic": arithmetic_note(),
                           "mean_evals_per_expert": float(nev.mean()), "max_evals_per_expert": int(nev.max()),
                           "ticks_per_fit": int(collect["fit"][-1]["ticks"]), "slots": int(collect["fit"][-1]["slots"]),
                           "status_counts": {"ok": int(st[0]), "not_pd": int(st[1]), "linesearch": int(st[2]), "maxiter": int(st[3])}},
                "lml_grad_evals_per_s": world * float(nev.sum()) * args.steps / (ms_total * 1e-3),
                "roofline": roofline, "cpu_baseline": cpu_baseline, "e2e": e2e, "gpu_launches": int(launches),
                "clocks": clk}
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def emit(line):
    """The one JSON line goes to the real stdout; everything else printed to fd 1 during the run (NCCL's version
    banner, library chatter) was redirected to stderr in main()."""
    _REAL_STDOUT.write(json.dumps(line) + "\n")
    _REAL_STDOUT.flush()


_REAL_STDOUT = sys.stdout


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
