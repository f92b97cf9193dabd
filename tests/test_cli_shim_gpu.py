"""the `search` module behind the reference's own dispatcher.

oracle/_ref/splinterctl is the reference CLI built from its own sources; oracle/_ref/gpu/splinterctl is the SAME
dispatcher and command table with one file swapped -- libsplinter_b200/cli/splinter_cli_cmd_search_gpu.c in place of
splinter_cli_cmd_search.c (command_modules[] entry 25, splinter_cli_main.c:268-277, untouched).  Both run against
one live store with the same flags; stdout must be byte-identical (the module scores in SPL_MODE_EXACT, so even the
%.4f rounding agrees).  Also: the REPL keeps the GPU context between queries, and the stand-alone
splinter_gpu_search wraps the same module."""
import os
import subprocess

import numpy as np
import pytest

import support

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not (support.REF_DIR / "gpu" / "splinterctl").exists(), reason="needs oracle/_ref")]

DIM = 768
SHIM = support.ROOT / "libsplinter_b200" / "cli" / "splinter_gpu_search"
STOCK, GPU = support.REF_DIR / "splinterctl", support.REF_DIR / "gpu" / "splinterctl"
CASES = [["--limit", "10"], ["--json", "--limit", "7"], ["--bloom", "3", "--json"], ["--bloom", "1", "--limit", "5"],
         ["--regex", "^note", "--limit", "12", "--json"], ["--similarity", "0.05", "--json"], ["--distance", "38.5"],
         ["--json"]]


def _make_store(name):
    """700 keys (two key families, most with an embedding, random labels) through the reference's own library"""
    ref = support.ref_store()
    path = "/dev/shm" + name
    if os.path.exists(path):
        os.unlink(path)
    rng = np.random.default_rng(12)
    assert ref.splinter_create(name.encode(), 1024, 64) == 0
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
    return rng


@pytest.fixture()
def store(tmp_path):
    name = f"/splb200_cli_{os.getpid()}"
    rng = _make_store(name)
    qfile = tmp_path / "q.f32"
    rng.standard_normal(DIM).astype(np.float32).tofile(qfile)
    yield name, str(qfile)
    if os.path.exists("/dev/shm" + name):
        os.unlink("/dev/shm" + name)


def _one_shot(binary, name, qfile, flags, pre=("--use",)):
    stub = subprocess.Popen([str(support.REF_DIR / "ref_embed_stub"), name, qfile, "20000"])
    env = {**os.environ, "LD_LIBRARY_PATH": str(support.REF_DIR) + ":" + os.environ.get("LD_LIBRARY_PATH", "")}
    cmd = [str(binary), pre[0], name] + (["search"] if binary != SHIM else []) + ["what is splinter?"] + flags
    got = subprocess.run(cmd, capture_output=True, text=True, timeout=120, env=env)
    stub.wait(timeout=30)
    return got


def test_reference_dispatcher_with_the_gpu_module_prints_the_reference_s_bytes(store):
    name, qfile = store
    for flags in CASES:
        want = _one_shot(STOCK, name, qfile, flags)
        assert want.returncode == 0, want.stderr
        got = _one_shot(GPU, name, qfile, flags)
        assert got.returncode == 0, got.stderr
        assert got.stdout == want.stdout, f"flags {flags}:\n--- reference\n{want.stdout[:600]}\n--- gpu module\n{got.stdout[:600]}"
        assert len(want.stdout) > 80
    # the module list of the swapped build is the reference's: `search` is still entry 25, nothing else moved
    a = subprocess.run([str(STOCK), "--list-modules"], capture_output=True, text=True).stdout
    b = subprocess.run([str(GPU), "--list-modules"], capture_output=True, text=True).stdout
    assert a == b and "search" in a
    ha = subprocess.run([str(STOCK), "help", "search"], capture_output=True, text=True).stdout
    hb = subprocess.run([str(GPU), "help", "search"], capture_output=True, text=True).stdout
    assert ha == hb and "--bloom" in ha


def test_repl_session_keeps_the_mirror_between_queries(store):
    """Three searches typed into one REPL process (the binary under its REPL name): the GPU context is attached by the
    first and reused by the others; a write between two queries is picked up by the incremental sync."""
    name, qfile = store
    script = ('search "what is splinter?" --limit 5 --json\n'
              'set zzz_new_key hello\n'
              'search "what is splinter?" --bloom 3 --json\n'
              'search "what is splinter?" --regex ^zzz --json\n')
    outs = []
    for binary in (support.REF_DIR / "splinter_cli", support.REF_DIR / "gpu" / "splinter_cli"):
        if not binary.exists():   # the REPL name is a symlink; make it if the snapshot dropped it
            os.symlink("splinterctl", binary)
        stub = subprocess.Popen([str(support.REF_DIR / "ref_embed_stub"), name, qfile, "30000", "3"])
        r = subprocess.run([str(binary), "--use", name], input=script, capture_output=True, text=True, timeout=180)
        stub.wait(timeout=60)
        assert r.returncode == 0, r.stderr
        outs.append(r.stdout)
        _make_store(name)   # the session wrote to the store: give the other binary the same starting point
    assert outs[0] == outs[1]
    assert outs[0].count('"results"') == 3 and "zzz_new_key" in outs[0]


def test_standalone_command_wraps_the_same_module(store):
    name, qfile = store
    for flags in (CASES[0], CASES[2], CASES[4]):
        want = _one_shot(STOCK, name, qfile, flags)
        got = _one_shot(SHIM, name, qfile, flags)
        assert got.returncode == 0, got.stderr
        assert got.stdout == want.stdout
