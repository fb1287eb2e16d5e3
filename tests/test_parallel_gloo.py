"""Host logic of the multi-GPU path on CPU: world-size-2 gloo process group, LPT sharding and the
single final gather (gpsat_b200/parallel.py).  No engine call: each rank fabricates per-expert rows
from the expert index, so the gathered arrays must come back in global order."""
import os
import socket
import sys

import numpy as np
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_path):
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch.distributed as dist
    from gpsat_b200 import parallel
    r, w, _ = parallel.init_process_group("gloo")
    assert (r, w) == (rank, world)
    rng = np.random.default_rng(0)
    E = 23
    n = rng.integers(50, 400, E)
    m = rng.integers(0, 9, E)
    off_m = np.concatenate([[0], np.cumsum(m)])
    parts = parallel.lpt_partition(parallel.expert_cost(n, m), world)
    idx = parts[rank]
    local = {"theta": np.stack([np.full(5, float(e)) for e in idx]) if len(idx) else np.zeros((0, 5)),
             "status": np.asarray(idx, np.int32) % 3}
    pred = {"f*": np.concatenate([np.full(m[e], e + 0.5) for e in idx]) if len(idx) else np.zeros(0)}
    out, pout = parallel.gather_results(local, idx, E, pred, off_m)
    if rank == 0:
        np.savez(out_path, theta=out["theta"], status=out["status"], f=pout["f*"], off_m=off_m,
                 sizes=np.array([len(p) for p in parts]))
    else:
        assert out is None and pout is None
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_shard_and_gather(tmp_path):
    out_path = str(tmp_path / "gathered.npz")
    mp.spawn(_worker, args=(2, _free_port(), out_path), nprocs=2, join=True)
    g = np.load(out_path)
    E = 23
    np.testing.assert_array_equal(g["theta"], np.repeat(np.arange(E, dtype=float)[:, None], 5, axis=1))
    np.testing.assert_array_equal(g["status"], np.arange(E) % 3)
    off_m = g["off_m"]
    for e in range(E):
        np.testing.assert_array_equal(g["f"][off_m[e]:off_m[e + 1]], e + 0.5)
    assert g["sizes"].sum() == E and g["sizes"].min() >= 1


def test_lpt_partition_is_deterministic_and_balanced():
    sys.path.insert(0, ROOT)
    from gpsat_b200 import parallel
    rng = np.random.default_rng(1)
    cost = parallel.expert_cost(rng.integers(800, 1200, 500), rng.integers(200, 1300, 500))
    a = parallel.lpt_partition(cost, 8)
    b = parallel.lpt_partition(cost, 8)
    assert all(np.array_equal(x, y) for x, y in zip(a, b))
    assert sorted(np.concatenate(a).tolist()) == list(range(500))
    loads = np.array([cost[p].sum() for p in a])
    assert loads.max() / loads.mean() < 1.01
    assert [len(p) for p in parallel.lpt_partition(cost[:3], 8)].count(0) == 5     # more ranks than experts
