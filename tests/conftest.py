import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "vings-mono_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


# Collection order: the parity tests against the oracle come first, trajectory / churn tests last, so that under the
# driver's `pytest -x` a failure in a caller-side test can never hide the hot-path parity suite again (round 1).
_ORDER = ("test_oracle_golden", "test_parity_gpu", "test_parity_full_gpu", "test_ops_gpu", "test_mapping",
          "test_sh_gpu", "test_host_cpu", "test_block_cull_cpu", "test_formats", "test_mapper_gloo", "test_multigpu",
          "test_churn", "test_zz_no_alloc")   # test_zz_no_alloc stays last: nothing can hide behind a failure in it


def _rank(item):
    name = os.path.splitext(os.path.basename(str(item.fspath)))[0]
    return _ORDER.index(name) if name in _ORDER else len(_ORDER) - 2


def pytest_collection_modifyitems(config, items):
    items.sort(key=_rank)   # stable: order inside a file is kept
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)
