import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def _build_if_missing():
    """Build the libraries if they are missing (CPU container); the GPU box gets them prebuilt.  Runs at configure
    time, i.e. BEFORE collection: the skipif marks that look for oracle/_ref are evaluated while collecting."""
    from libsplinter_b200 import LIB_PATH
    ref_missing = Path("/root/reference/splinter.c").exists() and not (ROOT / "oracle" / "_ref" / "libsplinter_p.so").exists()
    if not LIB_PATH.exists() or not (ROOT / "oracle" / "libspl_oracle.so").exists() or ref_missing:
        import __graft_entry__ as g
        g.build()


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    _build_if_missing()
