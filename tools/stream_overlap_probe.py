"""Developer probe: does running the c2 fit as G independent slot groups on G streams (here: G engine
handles driven by G host threads) recover the per-launch tails of the lock-step tick loop?

  python tools/stream_overlap_probe.py [E] [groups ...]

Prints the time of one fit of E c2-shaped experts for every group count, and the number of experts still
active at every optimiser tick (from n_evals), which bounds what any scheduling change can recover."""
import json
import sys
import threading
import time

import numpy as np

sys.path.insert(0, ".")
import torch
from gpsat_b200 import synthetic as syn
from gpsat_b200.engine import Engine
from dataclasses import replace

E = int(sys.argv[1]) if len(sys.argv) > 1 else 500
groups = [int(a) for a in sys.argv[2:]] or [1, 2, 4]
spec = replace(syn.WORKLOADS["c2"], E=E)
out = {"E": E, "runs": []}


def run_group(eng, dev, args, res):
    res.append(eng.fit(dev, *args))


for G in groups:
    bounds = np.linspace(0, E, G + 1).astype(int)
    # interleave sizes: expert e goes to group e % G so every group sees the same size mix
    members = [list(range(g, E, G)) for g in range(G)]
    engs, devs, dargs = [], [], []
    for g in range(G):
        host = syn.make_batch(spec, seed=0, experts=members[g], with_pred=True)
        th0, tk, ta, tb, tr = syn.default_fit_setup(spec, len(members[g]))
        engs.append(Engine(0))
        devs.append(host.to_device("cuda:0"))
        dargs.append([torch.as_tensor(a, device="cuda:0") for a in (th0, tk, ta, tb, tr)])
    best = None
    for rep in range(4):
        results = [[] for _ in range(G)]
        ths = [threading.Thread(target=run_group, args=(engs[g], devs[g], dargs[g], results[g])) for g in range(G)]
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if rep > 0:
            best = dt if best is None else min(best, dt)
    nev = np.concatenate([r[0].n_evals.cpu().numpy() for r in results])
    run = {"groups": G, "fit_ms": best * 1e3, "mean_evals": float(nev.mean()), "max_evals": int(nev.max())}
    if G == 1:
        run["active_per_tick"] = [int((nev >= t).sum()) for t in range(1, int(nev.max()) + 1)]
    out["runs"].append(run)
    print(json.dumps(run), flush=True)
    del engs, devs, dargs
    torch.cuda.empty_cache()
json.dump(out, open("gpurun_out/stream_overlap_probe.json", "w"))
