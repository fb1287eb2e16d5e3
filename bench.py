#!/usr/bin/env python
"""bench.py - local experts/sec (fit + predict, FP64) on N B200s of one node.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c1|c2|c3|ns|c4|c5]

One "step" = one full pass of the hot path over one batch of synthetic experts: batched
L-BFGS hyper-parameter fit (every objective evaluation = kernel build + Cholesky + inverse +
analytic gradient) followed by prediction at the optimised hyper-parameters.

Default workload (no --workload): BASELINE.json configs[1] (`configs/example_local_expert_oi.json` shape,
~500 experts x ~1k observations per GPU).  At N GPUs ONE global set of 500 N experts is partitioned over the
ranks by `gpsat_b200.parallel.lpt_partition` (longest-processing-time-first on n^3 evals + n^2 m), every rank
generates and uploads only its shard, runs independently, and rank 0 gathers the per-expert results once per
step: weak scaling (per-GPU work fixed), no collective on the hot path.
Every default line also carries `extra.ns2048_strong`: the north-star pass - 2048 experts x 2000 observations,
fit + predict - strong-scaled over the same N ranks through the same partitioner; at N = 1 `extra` adds short
runs of the other BASELINE.json configs (c1, c3, c4 predict-only, c5 sparse GPR), at N >= 4 the full-size c3 set
(3 000 experts x 2 000-4 000 observations) strong-scaled.

Workloads: c1 c2 c3 ns = exact GPR fit + predict; c4 = predict only at fixed hyper-parameters
(GPSat/postprocessing.py:365-395); c5 = GPflowSGPRModel shape, n = 20 000, M = 500 (gpflow_models.py:704).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from dataclasses import replace

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "local experts/sec (fit+predict, FP64)"
UNIT = "experts/s"
FP64_PEAK_FILE = os.path.join(ROOT, "profiles", "r01_fp64_peak.json")
FTOL = 2.220446049250313e-09


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=["c1", "c2", "c3", "ns", "c4", "c5"],
                    help="default: c2 with the extra block (see the module docstring)")
    ap.add_argument("--experts-per-gpu", type=int, default=0, help="override the workload's expert count per GPU")
    ap.add_argument("--strong", action="store_true", help="fixed global expert set (workload's E) split over the ranks")
    ap.add_argument("--cpu-sample", type=int, default=0, help="experts in the CPU baseline sample (0 = one per core)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extra", action="store_true")
    return ap.parse_args()


def workload_spec(name, experts_per_gpu=0):
    from gpsat_b200 import synthetic as syn
    spec = syn.WORKLOADS[name]
    if experts_per_gpu > 0:
        spec = replace(spec, E=experts_per_gpu)
    return spec


def workload_mode(name):
    return {"c4": "predict", "c5": "sgpr"}.get(name, "fit")


def fp64_peak_tflops():
    """FP64 pipe peak measured on this pool by tools/fp64_peak.cu (MEASURED_PEAKS.json has no FP64
    entry and the profiling guide states no FP64 fallback)."""
    try:
        d = json.load(open(FP64_PEAK_FILE))
        return float(d["dmma_m8n8k4_sustained3s_tflops"]), "measured: tools/fp64_peak.cu DMMA m8n8k4 sustained 3 s (profiles/r01_fp64_peak.json)"
    except Exception:
        return 37.1, "fallback: 148 SM x 64 FMA/clk x 2 x 1.965 GHz"


def hbm_peak_gbs():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "MEASURED_PEAKS.json (burst copy bandwidth)"
    except Exception:
        return 6542.0, "fallback: 6 542 GB/s measured on this pool in round 1 (B200_PROFILING.md recipe)"


def arithmetic_note():
    step = os.environ.get("GPSAT_B200_STEP", "i8") != "dmma"
    lauum = os.environ.get("GPSAT_B200_LAUUM", "i8") != "dmma"
    if not step and not lauum:
        return "FP64 mma.sync throughout"
    parts = []
    if step:
        parts.append("the k-sums of the potrf / trtri tiles are FP64 values split into 8 exact int8 slices with a-priori "
                     "row scales and multiplied on the int8 tensor path (tcgen05, TMEM accumulators; truncation at "
                     "2^-56 of the row scales)")
    if lauum:
        parts.append("the K^-1 tile products of the gradient use 7 slices")
    return ("FP64 data and results; " + "; ".join(parts) + "; per expert and evaluation only while (var + noise) / noise "
            "<= 4096, FP64 mma.sync above (precision rule, profiles/r02_precision_envelope.md); covariance evaluation, "
            "diagonal-tile factorisation, C W^T / SYRK epilogues, L-BFGS and prediction reductions in FP64")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe: `-lms 200`).

    One nvidia-smi process per bench process, started before the warm-up steps and read by a thread that stamps every
    line with its arrival time; a timed region is the slice [start(), stop()] of that stream.  (Starting nvidia-smi at
    the top of every timed region put its start-up - NVML initialisation under the driver lock, 0.1 s - inside the first
    timed step: visible as a doubled time per pass on the 40 ms c1 passes.)"""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    _shared = {}

    @classmethod
    def get(cls, index):
        if index not in cls._shared:
            cls._shared[index] = cls(index)
        return cls._shared[index]

    def __init__(self, index):
        self.lines, self.proc, self.t0 = [], None, None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
            import atexit
            atexit.register(self.close)
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def close(self):
        if self.proc is not None and self.proc.poll() is None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def start(self):
        self.t0 = time.time()
        return self

    def stop(self):
        """Summary of the samples that arrived since start(); a region shorter than the sampling interval takes the
        first sample after it (waits up to 0.5 s for one)."""
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        t1 = time.time()
        sel = [ln for (ts, ln) in list(self.lines) if self.t0 <= ts <= t1]
        if not sel:
            deadline = time.time() + 0.5
            while time.time() < deadline and not any(ts > t1 for ts, _ in list(self.lines)):
                time.sleep(0.02)
            after = [ln for (ts, ln) in list(self.lines) if ts > t1]
            before = [ln for (ts, ln) in list(self.lines) if ts < self.t0]
            sel = after[:1] or before[-1:]
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in sel:
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


# ---- CPU arm -----------------------------------------------------------------------------------------------
def cpu_jobs(spec, experts, mode="fit", seed=0):
    from gpsat_b200 import synthetic as syn
    th0, tk, ta, tb, tr = syn.default_fit_setup(spec, 1)
    jobs = []
    for e in experts:
        X, y, Xs = syn.make_expert(spec, e, seed)
        if mode == "sgpr":
            jobs.append((spec.kernel, X, y, Xs, syn.inducing_points(X, e, seed), th0[0], tk[0], ta[0], tb[0], tr[0], 10_000))
        elif mode == "predict":
            jobs.append((spec.kernel, X, y, Xs, np.asarray(syn.FIXED_THETA), tk[0], ta[0], tb[0], tr[0], 0))
        else:
            jobs.append((spec.kernel, X, y, Xs, th0[0], tk[0], ta[0], tb[0], tr[0], 10_000))
    return jobs


def cpu_fn(mode):
    from oracle import cpu_baseline as cb
    return cb.sgpr_fit_predict_one if mode == "sgpr" else cb.fit_predict_one


def cpu_sample_size(args, spec, mode, cores):
    if args.cpu_sample:
        return args.cpu_sample
    # about 10-30 s of CPU work: one expert per core for c1 / c2, fewer of the big ones
    if mode == "sgpr" or spec.n_hi > 1500:
        return max(1, min(cores, 8))
    return cores


def run_reference(args):
    """--impl reference: the oracle port of the reference path on all host cores (the reference
    itself - GPflow/TensorFlow - is not installable here)."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    from oracle.cpu_baseline import CpuPool
    name = args.workload or "c2"
    spec, mode = workload_spec(name, args.experts_per_gpu), workload_mode(name)
    cores = os.cpu_count() or 1
    sample = cpu_sample_size(args, spec, mode, cores)
    pool = CpuPool(min(cores, sample))
    times, done, res = [], 0, []
    for s in range(args.warmup + args.steps):
        experts = [(s * sample + k) % spec.E for k in range(sample)]
        res, wall = pool.run(cpu_jobs(spec, experts, mode), cpu_fn(mode))
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
                       "n_obs": [spec.n_lo, spec.n_hi], "n_pred": [spec.m_lo, spec.m_hi], "optimise": mode != "predict"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": min(cores, sample), "kind": "port",
                             "sample": f"{sample} experts of {spec.name} per step, NumPy/SciPy oracle, "
                                       f"one process per core, 1 BLAS thread each", "blas": blas},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


# ---- one workload on one rank ------------------------------------------------------------------------------
class Job:
    """The shard of one workload that this rank owns: host (pinned) and device copies of the packed batch, the fit
    set-up, and the two ways of running one step (inputs resident in HBM / through the host-facing call)."""

    def __init__(self, name, spec, idx, E_total, rank, world, local, seed=0):
        import torch
        from gpsat_b200 import synthetic as syn
        from gpsat_b200.engine import Engine
        self.name, self.spec, self.mode = name, spec, workload_mode(name)
        self.idx, self.E_total, self.rank, self.world, self.local = np.asarray(idx), E_total, rank, world, local
        self.E = len(self.idx)
        self.dev = f"cuda:{local}"
        self.host = syn.make_batch(spec, seed=seed, experts=[int(e) for e in self.idx], with_pred=True)
        self.setup = syn.default_fit_setup(spec, self.E)
        self.Z = None
        if self.mode == "sgpr":
            self.Z = [syn.inducing_points(self.host.X[self.host.off_n[k]:self.host.off_n[k + 1]], int(e), seed)
                      for k, e in enumerate(self.idx)]
        self.theta_fixed = np.tile(np.asarray(syn.FIXED_THETA), (self.E, 1))
        self.eng = Engine(local)
        self.stream = torch.cuda.ExternalStream(self.eng.stream_ptr, device=self.dev)
        self.devb = self.host.to_device(self.dev)
        self.d_setup = [torch.as_tensor(a, device=self.dev) for a in self.setup]
        pin = lambda a: None if a is None else torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
        self.host.X, self.host.y, self.host.Xs = pin(self.host.X), pin(self.host.y), pin(self.host.Xs)

    def close(self):
        self.eng.close()
        self.devb = None

    def _run(self, batch, setup, theta_fixed, collect):
        eng = self.eng
        stats = []
        if self.mode == "predict":
            res, pred = None, eng.predict(batch, theta_fixed)
            stats.append(("predict", eng.stats()))
        elif self.mode == "sgpr":
            res = eng.fit(batch, *setup, inducing=self.Z)
            stats.append(("fit", eng.stats()))
            pred = eng.sgpr_predict(batch, self.Z, res.theta)
            stats.append(("predict", eng.stats()))
        else:
            res = eng.fit(batch, *setup)
            stats.append(("fit", eng.stats()))
            pred = eng.predict(batch, res.theta)
            stats.append(("predict", eng.stats()))
        if collect is not None:
            for k, s in stats:
                collect[k].append(s)
        return res, pred, stats

    def step_device(self, flush, collect=None):
        import torch
        with torch.cuda.stream(self.stream):
            flush.zero_()
        res, pred, _ = self._run(self.devb, self.d_setup, self.theta_fixed, collect)
        return res, pred

    def step_e2e(self):
        """Host buffers in, host results out, then the single final gather to rank 0."""
        from gpsat_b200 import parallel
        res, pred, stats = self._run(self.host, self.setup, self.theta_fixed, None)
        if self.world > 1:
            local = {"lml": pred["lml"] if "lml" in pred else pred["elbo"], "status": pred["status"]}
            if res is not None:
                local.update({"theta": res.theta, "n_evals": res.n_evals})
            parallel.gather_results(local, self.idx, self.E_total)
        return res, pred, sum(s["h2d_bytes"] for _, s in stats), sum(s["d2h_bytes"] for _, s in stats)

    # ---- algorithmic work ----
    def flops(self, nev):
        """(flops of the sweep's step kernels, whole-job flops) of one step."""
        from gpsat_b200 import synthetic as syn
        n, m = self.host.n.astype(np.float64), self.host.m.astype(np.float64)
        if self.mode == "sgpr":
            M = float(syn.SGPR_M)
            per_eval = 5.0 * M * M * n + 14.0 * M ** 3 / 3.0          # ELBO + gradient (tools/sgpr_bench.py)
            tot = float(np.sum(nev * per_eval) + np.sum(2.0 * M * M * n + 2.0 * M ** 3 / 3.0 + 2.0 * m * M * M))
            return None, tot
        pred = n ** 3 / 3.0 + n ** 2 * m
        if self.mode == "predict":
            return float(np.sum(pred)), float(np.sum(pred))
        return float(np.sum(nev * 2.0 * n ** 3 / 3.0) + np.sum(pred)), float(np.sum(nev * n ** 3) + np.sum(pred))

    def int8_ops(self, nev):
        """Executed int8 tensor ops of the sliced k-sums: 36 slice-pair MMAs of 2 * 128^3 ops per k-block of a
        potrf / trtri / diagonal-sum tile."""
        n, m = self.host.n.astype(np.float64), self.host.m.astype(np.float64)
        T = np.ceil((n + 1) / 128.0)
        TR = np.ceil((n + 1 + m) / 128.0)
        kb_pot = T * (T - 1) * (T - 2) / 6.0
        kb_tr = (T - 1) * T * (T + 1) / 6.0
        kb_dg = np.maximum((T - 1) * (T - 2) / 2.0, 0.0)
        kb_pred = kb_pot + (TR - T) * T * (T - 1) / 2.0 + kb_dg
        fit = float(np.sum(nev * (kb_pot + kb_tr + kb_dg))) if self.mode == "fit" else 0.0
        return 2.0 * 128.0 ** 3 * 36.0 * (fit + float(np.sum(kb_pred)))


def partition(spec, E_total, world, mode):
    """Index arrays per rank: LPT on the cost model n^3 evals + n^2 m (gpsat_b200.parallel.lpt_partition)."""
    from gpsat_b200 import parallel, synthetic as syn
    n, m = syn.expert_sizes(spec, range(E_total))
    cost = parallel.expert_cost(n, m, evals=0.34 if mode == "predict" else 30.0)
    parts = parallel.lpt_partition(cost, world)
    load = np.array([cost[p].sum() for p in parts])
    return parts, {"rule": "lpt_partition on n^3 * evals + n^2 * m", "experts_per_rank": [int(len(p)) for p in parts],
                   "cost_imbalance_max_over_mean": float(load.max() / load.mean())}


def timed_region(job, steps, warmup, flush, dist, world):
    """W warm-up steps, then K steps between barrier + synchronize, timed with CUDA events on the engine's stream;
    max over ranks.  Returns (ms_total, last results, per-call stats, clocks)."""
    import torch

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler.get(job.local)          # running before the warm-up, so that its start-up is outside the timed region
    for _ in range(warmup):
        job.step_device(flush)
    job.eng.set_profile(True)
    collect = {"fit": [], "predict": []}
    barrier()
    clocks = sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(job.stream)
    res = pred = None
    for _ in range(steps):
        res, pred = job.step_device(flush, collect)
    e1.record(job.stream)
    barrier()
    clk = clocks.stop()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=job.dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    job.eng.set_profile(False)
    return float(ms.item()), res, pred, collect, clk


def timed_e2e(job, steps, dist, world):
    import torch

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    job.step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        res_h, pred_h, h2d, d2h = job.step_e2e()
    barrier()
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=job.dev)
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    return float(dt.item()), res_h, pred_h, float(h2d), float(d2h)


def kernel_numbers(job, steps, ms_total, res, collect):
    """Roofline numbers of the timed region from the per-kernel CUDA-event times the engine collected."""
    nev = res.n_evals.cpu().numpy().astype(np.float64) if res is not None else np.zeros(job.E)
    f_step, f_total = job.flops(nev)
    peak, peak_src = fp64_peak_tflops()
    calls = [s for k in collect for s in collect[k]]
    ms_step = sum(s["ms_step"] for s in calls)
    n_step = sum(s["n_step"] for s in calls)
    ms_lauum = sum(s["ms_lauum"] for s in calls)
    ms_diag = sum(s["ms_diag"] for s in calls)
    ms_opt = sum(s["ms_opt"] for s in calls)
    ms_reduce = sum(s["ms_reduce"] for s in calls)
    out = {"nev": nev, "launches": int(sum(s["kernel_launches"] for s in calls)), "peak": peak, "peak_src": peak_src,
           "whole_job_tflops": steps * f_total / (ms_total * 1e-3) / 1e12, "ms_step": ms_step, "n_step": n_step,
           "share": {"step_kernel": ms_step / ms_total, "lauum_grad_kernel": ms_lauum / ms_total,
                     "diag_kernel": ms_diag / ms_total, "lbfgs_kernel": ms_opt / ms_total,
                     "predict_reduce_kernel": ms_reduce / ms_total}}
    out["whole_job_frac"] = out["whole_job_tflops"] / peak
    if f_step is not None and ms_step > 0:
        out["step_tflops"] = steps * f_step / (ms_step * 1e-3) / 1e12
        out["step_frac"] = out["step_tflops"] / peak
        if os.environ.get("GPSAT_B200_STEP", "i8") != "dmma":
            try:
                pk = json.load(open(os.path.join(ROOT, "profiles", "r01f_i8_rate.json")))
                i8peak = float(pk["rate"]["n128"]["int8_tops"]) / 1e3
            except Exception:
                i8peak = 3.48
            ach = steps * job.int8_ops(nev) / (ms_step * 1e-3) / 1e15
            out["int8"] = {"achieved": ach, "peak": i8peak, "unit": "POP/s", "frac": ach / i8peak,
                           "note": "executed int8 tensor ops of step_kernel's sliced k-sums / its CUDA-event time (which "
                                   "also holds the FP64 epilogues); peak = measured tcgen05 kind::i8 rate at M = N = 128 "
                                   "(profiles/r01f_i8_rate.json)"}
    if ms_reduce > 0 and job.mode != "sgpr":
        # predict_reduce_kernel reads the factored prediction rows once: 8 n m bytes per expert
        hbm, hbm_src = hbm_peak_gbs()
        n, m = job.host.n.astype(np.float64), job.host.m.astype(np.float64)
        gbs = steps * float(np.sum(8.0 * n * m)) / (ms_reduce * 1e-3) / 1e9
        out["hbm"] = {"kernel": "predict_reduce_kernel", "bound": "hbm", "achieved": gbs, "peak": hbm, "unit": "GB/s",
                      "frac": gbs / hbm, "peak_source": hbm_src, "algorithmic_bytes": "8 n m per expert"}
    return out


def run_extra(name, spec, world, rank, local, dist, flush, passes=1):
    """One short measurement of another workload for the `extra` block: the global set `spec.E` partitioned over the
    ranks (strong scaling), `passes` timed device-resident steps after one warm-up step, and one end-to-end pass."""
    import torch
    mode = workload_mode(name)
    parts, pinfo = partition(spec, spec.E, world, mode)
    job = Job(name, spec, parts[rank], spec.E, rank, world, local)
    try:
        ms, res, pred, collect, clk = timed_region(job, passes, 1, flush, dist, world)
        kn = kernel_numbers(job, passes, ms, res, collect)
        dt, _, _, h2d, d2h = timed_e2e(job, 1, dist, world)
        # whole-job numbers need the evaluation counts of every rank
        tot = torch.tensor([kn["whole_job_tflops"], float(kn["nev"].sum())], dtype=torch.float64, device=job.dev)
        if world > 1:
            dist.all_reduce(tot)
        st = np.bincount((res.status if res is not None else pred["status"]).cpu().numpy(), minlength=4)
        out = {"workload": spec.name, "mode": {"fit": "fit + predict", "predict": "predict only at fixed theta",
                                               "sgpr": "sparse GPR (M = 500) fit + predict"}[mode],
               "experts": spec.E, "n_obs": [spec.n_lo, spec.n_hi], "scaling": "strong",
               "value": spec.E * passes / (ms * 1e-3), "unit": UNIT, "seconds_per_pass": ms * 1e-3 / passes,
               "e2e_value": spec.E / dt, "e2e_seconds_per_pass": dt, "h2d_bytes_rank0": h2d, "d2h_bytes_rank0": d2h,
               "whole_job_tflops_all_ranks": float(tot[0].item()),
               "whole_job_frac": float(tot[0].item()) / (world * kn["peak"]),
               "evals_all_ranks": int(tot[1].item()), "partition": pinfo, "clocks": clk,
               "rank0": {"step_kernel_frac": kn.get("step_frac"), "int8_tensor_frac": kn.get("int8", {}).get("frac"),
                         "kernel_share_of_step": kn["share"], "hbm": kn.get("hbm"), "gpu_launches": kn["launches"],
                         "status_counts": {"ok": int(st[0]), "not_pd": int(st[1]), "linesearch": int(st[2]), "maxiter": int(st[3])}}}
    finally:
        job.close()
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist
    from gpsat_b200 import parallel, synthetic as syn

    rank, world, local = parallel.init_process_group()
    assert world == args.gpus or world == 1, f"WORLD_SIZE={world} but --gpus {args.gpus}"
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py --impl ours needs a B200: the CUDA path has no CPU fallback")
    torch.cuda.set_device(local)
    default_run = args.workload is None
    name = args.workload or "c2"
    spec, mode = workload_spec(name, args.experts_per_gpu), workload_mode(name)
    # the global expert set and this rank's shard of it
    E_total = spec.E if args.strong else spec.E * world
    parts, pinfo = partition(spec, E_total, world, mode)

    # optional CPU baseline (rank 0, N=1 only) - before CUDA is busy, worker processes are spawned
    cpu_baseline, cpu_check, cpu_idx = None, None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle.cpu_baseline import CpuPool
        cores = os.cpu_count() or 1
        sample = cpu_sample_size(args, spec, mode, cores)
        pool = CpuPool(min(cores, sample))
        cpu_idx = list(range(sample))
        cpu_check, wall = pool.run(cpu_jobs(spec, cpu_idx, mode), cpu_fn(mode))
        pool.close()
        cpu_baseline = {"value": len(cpu_check) / wall, "unit": UNIT, "cores": min(cores, sample), "kind": "port",
                        "sample": f"first {sample} experts of {spec.name}, NumPy/SciPy oracle (SciPy L-BFGS-B + "
                                  f"LAPACK), one process per core, 1 BLAS thread each, {wall:.1f} s wall",
                        "blas": CpuPool.blas_report(cpu_check),   # raises unless every BLAS in every worker had 1 thread
                        "mean_evals_per_expert": float(np.mean([r["nev"] for r in cpu_check]))}

    job = Job(name, spec, parts[rank], E_total, rank, world, local)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=job.dev)   # > 126 MB L2

    # ---- device-resident timing (value) ----
    ms_total, res, pred, collect, clk = timed_region(job, args.steps, args.warmup, flush, dist, world)
    value = E_total * args.steps / (ms_total * 1e-3)
    kn = kernel_numbers(job, args.steps, ms_total, res, collect)
    nev = kn["nev"]
    tot = torch.tensor([kn["whole_job_tflops"], float(nev.sum()), float(kn["launches"])], dtype=torch.float64, device=job.dev)
    if world > 1:
        dist.all_reduce(tot)
    traffic, traffic_src = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        traffic = tj["dram_bytes_read"] + tj["dram_bytes_write"]
        traffic_src = tj["source"] + " - a builder constant from one ncu --set full capture, not taken from this run"
    except Exception:
        pass
    i8 = os.environ.get("GPSAT_B200_STEP", "i8") != "dmma"
    if mode == "sgpr":
        roofline = {"bound": "tensor", "kernel": "sgpr_gemm_kernel (FP64 DMMA: A = W Kuf, B = A A^T, gradient products)",
                    "achieved": kn["whole_job_tflops"], "peak": kn["peak"], "unit": "TFLOP/s",
                    "frac": kn["whole_job_frac"], "traffic": None, "peak_source": kn["peak_src"],
                    "note": "achieved = algorithmic flops of the whole step (5 M^2 n + 14 M^3 / 3 per ELBO + gradient "
                            "evaluation) / step time on rank 0"}
    else:
        roofline = {"bound": "tensor", "kernel": "step_kernel (potrf/trtri tiles: int8-sliced tcgen05 main loops, FP64 DMMA "
                    "epilogues, fused diagonal factorisation)" if i8 else "step_kernel (FP64 DMMA potrf/trtri tiles + fused diagonal factorisation)",
                    "achieved": kn.get("step_tflops"), "peak": kn["peak"], "unit": "TFLOP/s", "frac": kn.get("step_frac"),
                    "traffic": traffic, "traffic_source": traffic_src, "peak_source": kn["peak_src"],
                    "launches": int(kn["n_step"]), "avg_launch_ms": kn["ms_step"] / max(kn["n_step"], 1),
                    "int8_tensor": kn.get("int8"),
                    "note": "achieved = algorithmic FP64 flops / kernel time on rank 0, peak = measured FP64 (DMMA) pipe "
                            "peak; with the k-sums on the int8 tensor path the fraction is an FP64-equivalent rate and "
                            "may exceed 1" if i8 else None,
                    "hbm_kernels": [kn["hbm"]] if "hbm" in kn else []}
    roofline.update({"whole_job_tflops": float(tot[0].item()), "whole_job_frac": float(tot[0].item()) / (world * kn["peak"]),
                     "kernel_share_of_step": kn["share"]})

    # ---- end to end through the host-facing call ----
    e2e = None
    if not args.no_e2e:
        dt, res_h, pred_h, h2d, d2h = timed_e2e(job, args.steps, dist, world)
        e2e = {"value": E_total * args.steps / dt, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "includes": "pinned host buffers -> device, fit, predict, results -> host, gather to rank 0"}
        # device-resident and host paths are the same arithmetic: results must agree bit for bit
        if res is not None:
            assert np.array_equal(res_h.theta, res.theta.cpu().numpy())
        assert np.array_equal(pred_h["f*"], pred["f*"].cpu().numpy())

    # ---- parity of the timed results with the CPU sample (checker role of the oracle) ----
    if rank == 0 and cpu_baseline is not None:
        cpu_baseline["parity_vs_gpu"] = check_parity(job, spec, mode, res, pred, cpu_check, cpu_idx)

    st = np.bincount((res.status if res is not None else pred["status"]).cpu().numpy(), minlength=4)
    slots = int(collect["fit"][-1]["slots"]) if collect["fit"] else int(collect["predict"][-1]["slots"])
    ticks = int(collect["fit"][-1]["ticks"]) if collect["fit"] else 0

    # ---- the other configs ----
    extra = None
    if default_run and not args.no_extra:
        job.close()
        extra = {"ns2048_strong": run_extra("ns", syn.WORKLOADS["ns"], world, rank, local, dist, flush)}
        if world >= 4:      # BASELINE.json configs[2] at full size: ~3k experts x 2-4k observations (a few seconds per pass)
            extra["c3_3000_strong"] = run_extra("c3", syn.WORKLOADS["c3"], world, rank, local, dist, flush)
        if world == 1:
            extra["c3_96"] = run_extra("c3", replace(syn.WORKLOADS["c3"], E=96), world, rank, local, dist, flush)
            extra["c4_96_predict_only"] = run_extra("c4", replace(syn.WORKLOADS["c4"], E=96), world, rank, local, dist, flush, passes=3)
            extra["c5_8_sgpr"] = run_extra("c5", replace(syn.WORKLOADS["c5"], E=8), world, rank, local, dist, flush)
            extra["c1"] = run_extra("c1", syn.WORKLOADS["c1"], world, rank, local, dist, flush, passes=3)
            # Matern-3/2 is the headline; the other kernels of the plugin as a sweep at c2
            sweep = {}
            for kind in ("Matern12", "Matern52", "RBF"):
                r = run_extra("c2", replace(syn.WORKLOADS["c2"], kernel=kind), world, rank, local, dist, flush)
                sweep[kind] = {k: r[k] for k in ("value", "unit", "seconds_per_pass", "e2e_value", "whole_job_frac", "evals_all_ranks")}
                sweep[kind]["status_counts"] = r["rank0"]["status_counts"]
                sweep[kind]["step_kernel_frac"] = r["rank0"]["step_kernel_frac"]
            extra["c2_kernel_sweep"] = sweep
            # the reference-facing top-level call: LocalExpertOI.run fed from DataFrames (selection on the GPU, one fit + one
            # predict call for all experts, result tables), wall clock of the second run - unpinned pandas inputs included
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import driver_bench
            extra["driver_local_expert_oi_500"] = driver_bench.run_driver(500, None)

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
                "scaling": "strong" if args.strong else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": spec.name, "experts_total": E_total, "experts_per_gpu": pinfo["experts_per_rank"],
                           "kernel": spec.kernel, "n_obs": [spec.n_lo, spec.n_hi], "n_pred": [spec.m_lo, spec.m_hi],
                           "optimise": mode != "predict", "partition": pinfo,
                           "parallelism": f"one global expert set partitioned x{world} by lpt_partition, no collective on the "
                                          f"hot path, one gather of the results per step (inside e2e)",
                           "l2": f"256 MB flush written between steps; per-step working set {slots} slot matrices >> 126 MB L2",
                           "arithmetic": arithmetic_note(),
                           "mean_evals_per_expert": float(tot[1].item()) / E_total,
                           "max_evals_per_expert_rank0": int(nev.max()) if len(nev) else 0,
                           "ticks_per_fit_rank0": ticks, "slots": slots,
                           "status_counts_rank0": {"ok": int(st[0]), "not_pd": int(st[1]), "linesearch": int(st[2]), "maxiter": int(st[3])}},
                "lml_grad_evals_per_s": float(tot[1].item()) * args.steps / (ms_total * 1e-3),
                "roofline": roofline, "cpu_baseline": cpu_baseline, "e2e": e2e, "gpu_launches": int(tot[2].item()),
                "clocks": clk, "extra": extra}
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def check_parity(job, spec, mode, res, pred, cpu_check, cpu_idx):
    """GPU results of the timed region against the CPU sample (the first experts of the global set: with one rank
    the shard is the whole set in index order)."""
    from gpsat_b200 import synthetic as syn
    th0, tk, ta, tb, tr = job.setup
    off = job.host.off_m
    fm = pred["f*"].cpu().numpy()
    pos = {int(e): k for k, e in enumerate(job.idx)}
    if mode == "predict":
        lm = pred["lml"].cpu().numpy()
        dl = dp = 0.0
        for e, r in zip(cpu_idx, cpu_check):
            k = pos[e]
            dl = max(dl, abs(lm[k] - r["lml"]) / abs(r["lml"]))
            dp = max(dp, float(np.max(np.abs(fm[off[k]:off[k + 1]] - r["fmean"])) / np.max(np.abs(r["fmean"]))))
        assert dl < 1e-8 and dp < 1e-8, f"fixed-theta parity (1e-8): lml {dl}, f* {dp}"
        return {"experts": len(cpu_check), "max_rel_lml": dl, "max_rel_fmean": dp, "tolerance": 1e-8}
    th = res.theta.cpu().numpy()
    lm = res.lml.cpu().numpy()
    nev = res.n_evals.cpu().numpy()
    status = res.status.cpu().numpy()
    if mode == "sgpr":
        from oracle import sgpr as _model
    else:
        from oracle import gpr as _model      # checker role only
    dth, dth_all, dl, pinned, flat, nev_diff = 0.0, 0.0, 0.0, 0, [], 0
    for e, r in zip(cpu_idx, cpu_check):
        k = pos[e]
        if not (r["ok"] and int(status[k]) == 0):
            continue
        rel = np.abs(th[k] - r["theta"]) / np.abs(r["theta"])
        dth_all = max(dth_all, float(np.max(rel)))
        nev_diff += int(int(nev[k]) != int(r["nev"]))
        # components pinned at a bound +- tol (tol = 1e-2, local_experts.py:1115) are skipped
        sig = tk[k] == 1
        at_bound = sig & ((r["theta"] >= tb[k] - 1e-2) | (r["theta"] <= ta[k] + 1e-2))
        pinned += int(at_bound.sum())
        dk = float(np.max(rel[~at_bound])) if (~at_bound).any() else 0.0
        if dk >= 1e-4:
            # L-BFGS-B stops when the relative decrease of the objective falls below ftol = 2.2e-9; on a flat
            # likelihood that leaves theta undetermined beyond 1e-4, and an LML that differs in the 9th digit (inside
            # the 1e-8 parity bar) can stop the search one iteration apart.  Such an end point counts as equal if
            # the ORACLE's objective at the GPU's theta is within 10 ftol of the oracle's own optimum - the same
            # statement as "SciPy could have stopped there".
            X, y, _ = syn.make_expert(spec, e, 0)
            if mode == "sgpr":
                l_at_gpu = _model.elbo(spec.kernel, X, y, syn.inducing_points(X, e, 0), th[k])
            else:
                l_at_gpu = _model.lml(spec.kernel, X, y, th[k])
            gap = (r["lml"] - l_at_gpu) / abs(r["lml"])
            flat.append({"expert": int(e), "max_rel_theta": dk, "oracle_objective_gap_at_gpu_theta": float(gap),
                         "nev_gpu": int(nev[k]), "nev_cpu": int(r["nev"])})
            assert gap <= 10 * FTOL, f"expert {e}: the GPU end point is a worse optimum: {flat[-1]}"
        else:
            dth = max(dth, dk)
        dl = max(dl, abs(lm[k] - r["lml"]) / abs(r["lml"]))
    assert dl < 1e-4, f"optimised objective differs from the CPU oracle: {dl}"
    assert dth < 1e-4, f"optimised hyper-parameters differ from the CPU oracle: {dth}"
    return {"experts": len(cpu_check), "max_rel_theta": dth, "max_rel_lml": dl, "tolerance": 1e-4,
            "max_rel_theta_incl_pinned_and_flat": dth_all, "components_pinned_at_bound": pinned,
            "flat_end_points": flat, "experts_with_different_nev": nev_diff}


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
