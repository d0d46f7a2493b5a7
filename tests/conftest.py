import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "micro-vllm_b200")
for p in (PKG, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:  # noqa: BLE001
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


def pytest_sessionfinish(session, exitstatus):
    """Write the tolerance-tail records of every assert_close call of this session (GPU runs only)."""
    try:
        import json
        from _cases import PARITY_LOG
    except Exception:  # noqa: BLE001
        return
    if not PARITY_LOG:
        return
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    worst = max(PARITY_LOG, key=lambda r: r["worst_err_over_bound"])
    most = max(PARITY_LOG, key=lambda r: r["out_of_bound"] / max(1, r["numel"]))
    summary = {"calls": len(PARITY_LOG), "calls_with_any_out_of_bound": sum(1 for r in PARITY_LOG if r["out_of_bound"]),
               "max_out_of_bound": max(r["out_of_bound"] for r in PARITY_LOG),
               "max_out_of_bound_fraction": most["out_of_bound"] / max(1, most["numel"]),
               "max_worst_err_over_bound": worst["worst_err_over_bound"], "worst_case": worst, "largest_tail_case": most}
    with open(os.path.join(out, "parity_tail.json"), "w") as f:
        json.dump({"summary": summary, "records": [r for r in PARITY_LOG if r["out_of_bound"]]}, f, indent=1)
