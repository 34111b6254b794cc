import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session", autouse=True)
def _built():
    """The CUDA library and the oracle are built in-tree before any test runs."""
    import __graft_entry__
    __graft_entry__.build()


def pytest_sessionfinish(session, exitstatus):
    """The size-parity tests record what they measured (errors next to the reference's run-to-run
    spread); dump it where gpurun brings it back from."""
    try:
        import json
        import util
        if util._PARITY_RECORD:
            out_dir = os.path.join(ROOT, "gpurun_out")
            os.makedirs(out_dir, exist_ok=True)
            path = os.path.join(out_dir, "r02_parity.json")
            # start from the committed record (a 1-GPU run must not drop the multi-GPU sections and
            # vice versa), then from what an earlier session of this run left
            old = {}
            for base in (os.path.join(ROOT, "profiles", "r02_parity.json"), path):
                if os.path.exists(base):
                    try:
                        old.update(json.load(open(base)))
                    except ValueError:
                        pass
            old.update(util._PARITY_RECORD)
            json.dump(old, open(path, "w"), indent=1, sort_keys=True)
    except Exception as ex:  # noqa: BLE001 - never turn a green run red from the reporting hook
        sys.stderr.write(f"parity record not written: {ex}\n")
