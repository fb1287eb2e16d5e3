"""Expert sharding across GPUs: one process per GPU, no collective on the hot path.

Experts are independent (the reference builds one model per expert from its own data,
``GPSat/local_experts.py:1043``), so the batch is partitioned by expert index with a
longest-processing-time-first rule on the cost model ``n^3 * evals + n^2 * m`` and every rank
runs its shard through its own engine.  The only communication is one gather of the per-expert
results (theta, objective, status, packed predictions) to rank 0 at the end.
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Sequence

import numpy as np


def expert_cost(n: np.ndarray, m: Optional[np.ndarray] = None, evals: float = 30.0) -> np.ndarray:
    n = np.asarray(n, dtype=np.float64)
    c = evals * n ** 3
    if m is not None:
        c = c + n ** 2 * np.asarray(m, dtype=np.float64)
    return c


def lpt_partition(cost: Sequence[float], world: int) -> List[np.ndarray]:
    """Deterministic LPT: sort by cost (descending, ties by index), give each expert to the
    currently least-loaded rank (ties by rank).  Returns ascending index arrays, one per rank."""
    cost = np.asarray(cost, dtype=np.float64)
    order = np.lexsort((np.arange(len(cost)), -cost))
    load = np.zeros(world)
    parts: List[List[int]] = [[] for _ in range(world)]
    for e in order:
        r = int(np.argmin(load))
        parts[r].append(int(e))
        load[r] += cost[e]
    return [np.array(sorted(p), dtype=np.int64) for p in parts]


def dist_env():
    """(rank, world, local_rank) from the torchrun environment (1 process if absent)."""
    return (int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)),
            int(os.environ.get("LOCAL_RANK", 0)))


def init_process_group(backend: Optional[str] = None):
    """Initialise torch.distributed from the environment (NCCL on GPUs, gloo otherwise)."""
    import torch
    import torch.distributed as dist
    rank, world, local = dist_env()
    if world == 1 or dist.is_initialized():
        return rank, world, local
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if backend == "nccl":
        torch.cuda.set_device(local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group(backend=backend, rank=rank, world_size=world)
    return rank, world, local


def gather_results(local: Dict[str, np.ndarray], idx: np.ndarray, E_total: int,
                   pred: Optional[Dict[str, np.ndarray]] = None, off_m_total: Optional[np.ndarray] = None,
                   dst: int = 0):
    """The single final gather.  ``local[k]`` holds per-expert rows for the experts ``idx`` of
    this rank; ``pred[k]`` holds this rank's packed predictions in ``idx`` order.  Rank ``dst``
    gets arrays for all ``E_total`` experts in global order (others get ``None``)."""
    import torch.distributed as dist
    rank, world, _ = dist_env()
    payload = {"idx": np.asarray(idx), "local": local, "pred": pred}
    if world == 1 or not dist.is_initialized():
        gathered = [payload]
    else:
        gathered = [None] * world if rank == dst else None
        dist.gather_object(payload, gathered, dst=dst)
        if rank != dst:
            return None, None
    out: Dict[str, np.ndarray] = {}
    for part in gathered:
        for k, v in part["local"].items():
            v = np.asarray(v)
            if k not in out:
                out[k] = np.zeros((E_total,) + v.shape[1:], dtype=v.dtype)
            if len(part["idx"]):
                out[k][part["idx"]] = v
    pout = None
    if off_m_total is not None:
        pout = {}
        for part in gathered:
            if part["pred"] is None:
                continue
            pos = 0
            for e in part["idx"]:
                k_m = int(off_m_total[e + 1] - off_m_total[e])
                for k, v in part["pred"].items():
                    if k not in pout:
                        pout[k] = np.zeros(int(off_m_total[-1]), dtype=np.asarray(v).dtype)
                    pout[k][off_m_total[e]:off_m_total[e + 1]] = np.asarray(v)[pos:pos + k_m]
                pos += k_m
    return out, pout
