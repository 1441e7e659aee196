import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import pkgload  # noqa: E402

pkgload.load()


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


def pytest_collection_modifyitems(config, items):
    import torch

    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def pytest_sessionfinish(session, exitstatus):
    """Parity ledger (tests/helpers.py): one JSON file per GPU session + the worst entry on stdout."""
    import json

    try:
        from helpers import LEDGER
    except Exception:
        return
    if not LEDGER:
        return
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    path = os.path.join(out, "parity_ledger.json")
    rows = []
    if os.environ.get("MDE_LEDGER_APPEND") == "1" and os.path.exists(path):  # tests/run_isolated.py: one process per test
        try:
            rows = json.load(open(path))["entries"]
        except Exception:
            rows = []
    rows += LEDGER
    worst = max(rows, key=lambda e: e["frac_of_bar"])
    json.dump({"worst": worst, "n": len(rows), "entries": rows}, open(path, "w"), indent=1)
    print(f"\nparity ledger: {len(rows)} entries -> {path}; worst {worst['case']} {worst['key']} {worst['metric']} "
          f"{worst['value']:.3e} ({worst['frac_of_bar']:.2f} of the {worst['bar']:g} bar)")
