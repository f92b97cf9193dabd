"""the `search` command shim.  The reference's real cmd_search (oracle/_ref/ref_search_cli)
and splinter_gpu_search run against the same live store with the same flags; their stdout must be
byte-identical (the shim scores in SPL_MODE_EXACT, so even the %.4f rounding agrees)."""
import os
import subprocess

import numpy as np
import pytest

import support

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not support.have_ref(), reason="needs oracle/_ref")]

DIM = 768
SHIM = support.ROOT / "libsplinter_b200" / "cli" / "splinter_gpu_search"


def test_cli_output_is_byte_identical_to_the_reference(tmp_path):
    ref = support.ref_store()
    name = f"/splb200_cli_{os.getpid()}"
    path = "/dev/shm" + name
    if os.path.exists(path):
        os.unlink(path)
    rng = np.random.default_rng(12)
    assert ref.splinter_create(name.encode(), 1024, 64) == 0
    try:
        for i in range(700):
            key = (f"doc{i:06d}" if i % 3 else f"note{i:06d}").encode()
            assert ref.splinter_set(key, f"value {i}".encode(), len(f"value {i}")) == 0
            if i % 9:
                v = rng.standard_normal(DIM).astype(np.float32)
                assert ref.splinter_set_embedding(key, support._p(v)) == 0
            m = int(rng.integers(0, 8))
            if m:
                ref.splinter_set_label(key, m)
        ref.splinter_close()
        q = rng.standard_normal(DIM).astype(np.float32)
        qfile = tmp_path / "q.f32"
        q.tofile(qfile)
        env = {**os.environ, "SPLINTER_LIB": str(support.REF_DIR / "libsplinter.so")}
        cases = [["--limit", "10"], ["--json", "--limit", "7"], ["--bloom", "3", "--json"], ["--bloom", "1", "--limit", "5"],
                 ["--regex", "^note", "--limit", "12", "--json"], ["--similarity", "0.05", "--json"], ["--distance", "38.5"],
                 ["--json"]]
        for flags in cases:
            want = subprocess.run([str(support.REF_DIR / "ref_search_cli"), name, str(qfile), "what is splinter?"] + flags,
                                  capture_output=True, text=True, timeout=60)
            assert want.returncode == 0, want.stderr
            stub = subprocess.Popen([str(support.REF_DIR / "ref_embed_stub"), name, str(qfile), "20000"])
            got = subprocess.run([str(SHIM), "--use", name, "what is splinter?"] + flags, capture_output=True, text=True,
                                 timeout=120, env=env)
            stub.wait(timeout=30)
            assert got.returncode == 0, got.stderr
            assert got.stdout == want.stdout, f"flags {flags}:\\n--- reference\\n{want.stdout[:600]}\\n--- shim\\n{got.stdout[:600]}"
            assert len(want.stdout) > 80
    finally:
        if os.path.exists(path):
            os.unlink(path)
