"""Conditioning sweep on the GPU: LML, gradient and predictions of every case of
tests/envelope.py through the C ABI on (a) the default engine configuration, (b) the int8-sliced arithmetic forced on
every slot, (c) FP64 mma.sync everywhere, each against the extended-precision golden values
(tests/golden/precision_envelope.npz), with the float64 LAPACK oracle's own error beside them.

  python tools/precision_envelope.py [--out gpurun_out/precision_envelope.json] [--sizes 1000,2000,4000]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import envelope as env                                   # noqa: E402
from gpsat_b200.engine import Engine, RaggedBatch        # noqa: E402


def run_config(eng, kind, items):
    b = RaggedBatch.from_lists(kind, [it[0] for it in items], [it[1] for it in items], [it[2] for it in items])
    theta = np.stack([it[3] for it in items])
    lml, grad, st = eng.lml_grad(b, theta)
    p = eng.predict(b, theta)
    return lml, grad, st, p, b.off_m


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "precision_envelope.json"))
    ap.add_argument("--sizes", default="1000,2000,4000")
    args = ap.parse_args()
    sizes = tuple(int(s) for s in args.sizes.split(","))
    gold = np.load(os.path.join(ROOT, "tests", "golden", "precision_envelope.npz"))
    eng = Engine(0)
    configs = {"default": None, "i8_forced": ("i8", "i8", 0), "dmma": ("dmma", "dmma", 0)}
    rows = []
    all_cases = env.cases(sizes)
    groups = {}
    for c in all_cases:
        groups.setdefault((c["kind"], c["D"]), []).append(c)
    groups[("RBF", 2)] = [dict(name="smoke", kind="RBF", D=2)]
    for (kind, D), cs in groups.items():
        items = [env.smoke_case() if c["name"] == "smoke" else env.make_case(c) for c in cs]
        res = {}
        for cname, cfg in configs.items():
            if cfg is None:
                eng.set_step_path("i8"); eng.set_lauum_path("i8"); eng.set_precision_rule(True)
            else:
                eng.set_step_path(cfg[0]); eng.set_lauum_path(cfg[1]); eng.set_precision_rule(False)
            res[cname] = run_config(eng, kind, items)
        for k, c in enumerate(cs):
            ref = {q: gold[f"{c['name']}/{q}"] for q in ("lml", "grad", "fmean", "fvar")}
            ref["lml"] = float(ref["lml"])
            row = dict(case=c["name"], kind=kind, n=int(len(items[k][0])), noise=float(items[k][3][-1]),
                       var=float(items[k][3][-2]), f64_oracle=dict(zip(("lml", "grad", "fmean", "fvar"),
                                                                     map(float, gold[f"{c['name']}/f64_err"]))))
            for cname in configs:
                lml, grad, st, p, off = res[cname]
                sl = slice(off[k], off[k + 1])
                e = env.rel_errors(lml[k], grad[k], p["f*"][sl], p["f*_var"][sl], ref, items[k][3])
                e["status"] = int(st[k])
                row[cname] = {q: (float(v) if q != "status" else v) for q, v in e.items()}
            rows.append(row)
            print(json.dumps(row), flush=True)
    eng.set_step_path("i8"); eng.set_lauum_path("i8"); eng.set_precision_rule(True)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(rows, open(args.out, "w"), indent=1)
    # compact table
    print(f"\n{'case':30s} {'quantity':6s} {'f64 LAPACK':>11s} {'dmma':>9s} {'i8 forced':>9s} {'default':>9s}")
    for r in rows:
        for q in ("lml", "grad", "fmean", "fvar"):
            print(f"{r['case']:30s} {q:6s} {r['f64_oracle'][q]:11.1e} {r['dmma'][q]:9.1e} {r['i8_forced'][q]:9.1e} "
                  f"{r['default'][q]:9.1e}")


if __name__ == "__main__":
    main()
