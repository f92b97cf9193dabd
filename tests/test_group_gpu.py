"""Several GPUs behind the C ABI (spl_gpu_group_*): a C program -- no Python on its path -- searches one live store
through shards and must get the single-context bytes (tests/c/group_search_test.c).  On a one-GPU box the shards share
the device ("0,0,0"); with more GPUs each shard gets its own.  A ctypes leg checks the same against the CPU oracle."""
import os
import subprocess

import numpy as np
import pytest

import libsplinter_b200 as sg
import support

torch = pytest.importorskip("torch")
pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not support.have_ref(), reason="needs oracle/_ref")]
DIM = 768
PROG = support.ROOT / "tests" / "c" / "group_search_test"


@pytest.fixture()
def store(tmp_path):
    ref = support.ref_store()
    name = f"/splb200_grp_{os.getpid()}"
    path = "/dev/shm" + name
    if os.path.exists(path):
        os.unlink(path)
    rng = np.random.default_rng(21)
    assert ref.splinter_create(name.encode(), 8192, 64) == 0
    for i in range(6000):
        key = f"doc{i:06d}".encode()
        assert ref.splinter_set(key, b"v", 1) == 0
        if i % 11:
            v = rng.standard_normal(DIM).astype(np.float32)
            assert ref.splinter_set_embedding(key, support._p(v)) == 0
        lab = int(rng.integers(0, 4))
        if lab:
            ref.splinter_set_label(key, lab)
    ref.splinter_close()
    q = rng.standard_normal((16, DIM)).astype(np.float32)
    qfile = tmp_path / "q.f32"
    q.tofile(qfile)
    yield name, str(qfile), q
    if os.path.exists(path):
        os.unlink(path)


def _devices(n):
    have = torch.cuda.device_count()
    return [i % have for i in range(n)]


@pytest.mark.parametrize("shards", [2, 3])
def test_c_host_searches_through_a_group(store, shards):
    name, qfile, _ = store
    devs = ",".join(str(d) for d in _devices(shards))
    r = subprocess.run([str(PROG), name, qfile, "16", devs], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert f"GROUP_OK shards={shards}" in r.stdout, r.stdout + r.stderr


def test_group_answer_is_the_oracle_s(store):
    name, _, q = store
    img = np.frombuffer(open("/dev/shm" + name, "rb").read(), np.uint8).copy()
    with sg.Group.attach(name, _devices(4), flags=sg.OPT_BF16_MIRROR) as g:
        assert g.size() == 4
        for mask in (0, 0x2):
            out, cnt = g.search(q[:3], 25, bloom_mask=mask, mode=sg.MODE_FAST)
            for i in range(3):
                ref_hits, _ = support.oracle_search_image(img, q[i], 25, bloom_mask=mask)
                support.assert_topk_matches(out[i][:cnt[i]], ref_hits)
        out, cnt = g.search(q, 10, mode=sg.MODE_BATCH_BF16)
        for i in range(16):
            ref_hits, _ = support.oracle_search_image(img, q[i], 10)
            support.assert_topk_matches(out[i][:cnt[i]], ref_hits)
