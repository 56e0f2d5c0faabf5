import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

REFERENCE = Path("/root/reference")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 CUDA devices")


def pytest_collection_modifyitems(config, items):
    # keep tests/test_zz_* last whatever order files were passed in
    items.sort(key=lambda it: (Path(str(it.fspath)).name.startswith("test_zz_"),))
    try:
        import torch

        have = torch.cuda.is_available()
        ngpu = torch.cuda.device_count() if have else 0
    except Exception:  # pragma: no cover
        have, ngpu = False, 0
    skip_gpu = pytest.mark.skip(reason="no CUDA device")
    skip_multi = pytest.mark.skip(reason="needs >= 2 CUDA devices")
    for it in items:
        if "gpu" in it.keywords and not have:
            it.add_marker(skip_gpu)
        if "multigpu" in it.keywords and ngpu < 2:
            it.add_marker(skip_multi)


@pytest.fixture(scope="session")
def reference_dir():
    return REFERENCE if REFERENCE.exists() else None


@pytest.fixture(scope="session")
def english_text():
    """A few hundred KB of English-like text built from a fixed paragraph (no dataset on the box)."""
    from cme213_sp18_b200.utils.text import synthetic_english

    return synthetic_english(400_000, seed=7)
