"""CPU checks of the drop-in boundary: the C-ABI library loads without a GPU, exports every symbol
include/splinter_gpu.h declares, and refuses to compute without a device (no CPU fallback)."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

import libsplinter_b200 as sg

ROOT = Path(__file__).resolve().parent.parent


def _declared_symbols():
    text = (ROOT / "include" / "splinter_gpu.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(spl_gpu_[a-z0-9_]+)\s*\(", text)))


def test_library_loads_and_exports_every_declared_symbol():
    lib = sg.load()
    declared = _declared_symbols()
    assert len(declared) >= 16
    for sym in declared:
        assert hasattr(lib, sym), f"{sym} declared in include/splinter_gpu.h but not exported"
    assert sorted(sg.ABI_SYMBOLS) == declared


def test_struct_layouts_match_the_header():
    assert C.sizeof(sg.SearchParams) == 48
    assert C.sizeof(sg.binding.Opts) == 24
    assert sg.RESULT_DTYPE.itemsize == 32
    assert C.sizeof(sg.SyncStats) == 96
    assert C.sizeof(sg.Counters) == 48
    assert C.sizeof(sg.Info) == 64


def test_only_sm100a_code_is_embedded():
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", str(sg.LIB_PATH)], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_product_does_not_link_or_name_the_oracle():
    import subprocess
    deps = subprocess.run(["ldd", str(sg.LIB_PATH)], capture_output=True, text=True).stdout
    assert "oracle" not in deps and "splinter.so" not in deps
    for src in (ROOT / "libsplinter_b200").rglob("*"):
        if src.suffix in {".cu", ".cuh", ".py", ".cpp", ".h"}:
            text = src.read_text()
            assert "libspl_oracle" not in text and "oracle/_ref" not in text, src


@pytest.mark.skipif(__import__("torch").cuda.is_available(), reason="needs a box WITHOUT a GPU")
def test_no_cpu_fallback():
    with pytest.raises(sg.SplinterGpuError, match="no CUDA device|no CPU fallback"):
        sg.Mirror.open_dense(16, 768)
    img = np.zeros(5440 + 16 * 3200 + 16 * 64, np.uint8)
    with pytest.raises(sg.SplinterGpuError):
        sg.Mirror.attach_image(img)
