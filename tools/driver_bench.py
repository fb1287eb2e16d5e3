"""Wall-clock split of the batched driver (LocalExpertOI.run) on a c2-shaped problem: host selection,
engine (fit + predict), result tables.  A measurement aid for the rows either side of the engine.

  python tools/driver_bench.py [--experts 400] [--device-select]
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import pandas as pd

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run_driver(experts=400, device_select=None):
    """One warm-up run and one timed run of LocalExpertOI.run on `experts` c2-like experts fed from DataFrames."""
    from gpsat_b200.local_experts import LocalExpertOI

    rng = np.random.default_rng(0)
    side = int(np.ceil(np.sqrt(experts)))
    spacing = 200e3
    half = side * spacing / 2 + 300e3
    # ~1000 observations inside every 300 km disc over a 9-day window
    dens = 1000.0 / (np.pi * 300e3 ** 2)
    N = int(dens * (2 * half) ** 2)
    xy = rng.uniform(-half, half, (N, 2))
    t = rng.integers(-4, 5, N).astype(float)
    obs = np.sin(xy[:, 0] / 3e5) + 0.5 * np.cos(xy[:, 1] / 2e5) + 0.02 * t + 0.1 * rng.standard_normal(N)
    data = pd.DataFrame({"x": xy[:, 0], "y": xy[:, 1], "t": t, "obs": obs})
    g = (np.arange(side) - (side - 1) / 2) * spacing
    locs = pd.DataFrame([(a, b, 0.0) for a in g for b in g][:experts], columns=["x", "y", "t"])
    pg = np.arange(-half + 300e3, half - 300e3 + 1, 25e3)
    pred = pd.DataFrame([(a, b) for a in pg for b in pg], columns=["x", "y"])
    oi = LocalExpertOI(locs, data,
                       {"obs_col": "obs", "coords_col": ["x", "y", "t"],
                        "local_select": [{"col": "t", "comp": "<=", "val": 4}, {"col": "t", "comp": ">=", "val": -4},
                                         {"col": ["x", "y"], "comp": "<", "val": 300e3}]},
                       {"oi_model": "B200GPRModel", "init_params": {"coords_scale": [5e4, 5e4, 1], "obs_mean": "local"},
                        "constraints": {"lengthscales": {"low": [1e-8, 1e-8, 1e-8], "high": [6e5, 6e5, 9]}}},
                       {"df": pred, "max_dist": 200e3})
    kw = {"device_select": device_select}
    oi.run(**kw)                      # warm-up (allocations, first CUDA call)
    t0 = time.time()
    tabs = oi.run(**kw)
    wall = time.time() - t0
    rd = tabs["run_details"]
    return {"experts": int(len(rd)), "observations": N, "mean_obs_per_expert": float(rd["num_obs"].mean()),
            "pred_rows": int(len(tabs["preds"])), "wall_s": wall, **oi.timings,
            "experts_per_s_end_to_end": len(rd) / wall, "device_select": device_select}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--experts", type=int, default=400)
    ap.add_argument("--device-select", action="store_true")
    args = ap.parse_args()
    print(json.dumps(run_driver(args.experts, bool(args.device_select))))


if __name__ == "__main__":
    main()
