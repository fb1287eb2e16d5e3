"""``python -m gpsat_b200.local_expert_oi <config.json>`` - the reference's command-line entry
(``examples/local_expert_oi.py:62-145``) on the B200 engine.

The JSON file has the reference's layout (``configs/example_local_expert_oi.json``): a dict, or a list of dicts, with
the keys ``locations`` / ``data`` / ``model`` / ``pred_loc`` (the four configs of ``LocalExpertOI``,
``GPSat/config_dataclasses.py:591-595``), ``run_kwargs`` (legacy name ``misc``: arguments of ``LocalExpertOI.run``),
``results`` (legacy: ``{"dir", "file"}`` instead of ``run_kwargs.store_path``) and ``comment``.
``model.oi_model`` may name the reference's classes (``GPflowGPRModel`` -> ``B200GPRModel``, ...).

Under ``torchrun --nproc-per-node N`` the experts are partitioned over N GPUs (``LocalExpertOI.run(world=...)``)
and rank 0 writes the store.  A ``store_path`` ending in ``.h5`` is an HDF5 store as in the reference (needs
PyTables); any other path is a directory with one parquet file per table.
"""
from __future__ import annotations

import json
import os
import sys
import time

VALID_KEYS = ["comment", "run_kwargs", "misc", "locations", "data", "model", "pred_loc", "results", "skip_valid_checks_on"]


def get_local_expert_oi_config(argv=None) -> list:
    """``get_config_from_sysargv`` (``GPSat/utils.py:139-176``): the first command-line argument is a JSON file."""
    argv = sys.argv if argv is None else argv
    if len(argv) < 2 or not argv[1].endswith(".json"):
        raise SystemExit("usage: python -m gpsat_b200.local_expert_oi <config.json>")
    with open(argv[1], "r") as f:
        configs = json.load(f)
    if not isinstance(configs, list):
        configs = [configs]
    return configs


def run_config(config: dict, verbose: bool = True) -> dict:
    """One config: build ``LocalExpertOI`` and run it (``examples/local_expert_oi.py:89-142``).  Returns the tables."""
    from .local_experts import LocalExpertOI
    from .parallel import dist_env
    config = dict(config)
    invalid = [k for k in config if k not in VALID_KEYS]
    assert len(invalid) == 0, f"the following invalid keys were provided in the config:\n{invalid}\nremove and try again"
    comment = config.pop("comment", None)
    comment = "\n".join(comment) if isinstance(comment, list) else comment
    rank = dist_env()[0]
    if verbose and rank == 0:
        print(f"\nconfig 'comment':\n\n{comment}\n\n")
    run_kwargs = dict(config.get("run_kwargs", config.get("misc", {})) or {})
    if "skip_valid_checks_on" not in run_kwargs:      # legacy handling (examples/local_expert_oi.py:112-115)
        run_kwargs["skip_valid_checks_on"] = ["skip_valid_checks_on"] + config.get("skip_valid_checks_on", [])
    locexp = LocalExpertOI(expert_loc_config=config["locations"], data_config=config["data"],
                           model_config=config["model"], pred_loc_config=config.get("pred_loc", None))
    if "results" in config:                             # legacy handling of how to get store_path (:131-136)
        store_path = os.path.join(config["results"]["dir"], config["results"]["file"])
        run_kwargs.pop("store_path", None)
    else:
        store_path = run_kwargs.pop("store_path")
    run_kwargs.setdefault("device_select", None)        # the array-built path whenever the configuration allows it
    tables = locexp.run(store_path=store_path, **run_kwargs)
    if verbose and rank == 0:
        print(f"results were written to:\n{store_path}")
    return tables


def main(argv=None):
    t1 = time.time()
    configs = get_local_expert_oi_config(argv)
    for count, config in enumerate(configs):
        print(f"config count: {count + 1}/{len(configs)}")
        run_config(config)
    print(f"Total run time: {time.time() - t1:.2f} seconds")


if __name__ == "__main__":
    main()
