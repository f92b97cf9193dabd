"""spl_gpu_wait: the reference's event bus as the wake-up, H->epoch as the gate.
The store and its bus are created through the reference's own library (oracle/_ref/libsplinter.so:
splinter_create + splinter_event_bus_init, splinter.c:991-999), so the header offsets the wait reads and
the eventfd hand-over are checked against the real thing, not against a restatement."""
import os
import subprocess
import threading
import time

import numpy as np
import pytest

import libsplinter_b200 as sg
import support

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not support.have_ref(), reason="needs oracle/_ref")]

DIM = 768


def test_wait_wakes_on_bus_writes_label_writes_and_foreign_writers():
    ref = support.ref_store()
    name = f"/splb200_wait_{os.getpid()}"
    path = "/dev/shm" + name
    if os.path.exists(path):
        os.unlink(path)
    assert ref.splinter_create(name.encode(), 256, 64) == 0      # this process owns the store ...
    assert ref.splinter_event_bus_init() == 0                     # ... and its event bus
    vec = np.zeros(DIM, np.float32)
    for k in range(16):
        key = f"doc{k:08d}".encode()
        assert ref.splinter_set(key, b"v", 1) == 0
        vec[0] = k; vec[1] = 1.0
        vec[2:] = np.float32(1.0) + np.arange(2, DIM, dtype=np.float32) / np.float32(1024.0)
        assert ref.splinter_set_embedding(key, support._p(vec)) == 0
    try:
        m = sg.Mirror.attach(name, sync_threads=2)
        # nothing happened since the attach's sync: the wait runs out
        t0 = time.perf_counter()
        assert m.wait(60) is False
        assert 0.05 <= time.perf_counter() - t0 < 2.0
        # a write by the bus owner wakes it long before the timeout
        threading.Timer(0.03, lambda: ref.splinter_set(b"doc00000001", b"vv", 2)).start()
        t0 = time.perf_counter()
        assert m.wait(5000) is True
        assert time.perf_counter() - t0 < 3.0
        assert m.wait(5000) is True                  # still dirty until the mirror is synced
        st = m.sync()
        assert st.rows_uploaded >= 1
        assert m.wait(20) is False
        # a label write moves no slot epoch but does move H->epoch (splinter.c:843-845)
        assert ref.splinter_set_label(b"doc00000002", 4) == 0
        assert m.wait(1000) is True
        st = m.sync()
        assert st.meta_updated == 1 and st.rows_uploaded == 0
        assert m.wait(20) is False
        # a writer PROCESS never signals the owner's eventfd (g_event_fd is per process, splinter.c:929-939):
        # the periodic H->epoch check catches it
        writer = str(support.REF_DIR / "ref_writer")
        p = subprocess.Popen([writer, name, "0", "8", "5", "100"], stdout=subprocess.PIPE, text=True)
        t0 = time.perf_counter()
        assert m.wait(5000) is True
        assert time.perf_counter() - t0 < 4.0
        p.communicate()
        m.sync()
        assert m.wait(20) is False
        m.close()
    finally:
        ref.splinter_close()
        if os.path.exists(path):
            os.unlink(path)


def test_wait_without_a_bus_degrades_to_the_timed_epoch_check():
    img = support.new_image(64, 64, DIM)                  # oracle-built image: owner_fd = -1, no bus
    support.image_set(img, "a")
    m = sg.Mirror.attach_image(img)
    assert m.wait(30) is False
    support.image_set(img, "b")                          # bumps H->epoch like splinter_set
    assert m.wait(1000) is True
    m.sync()
    assert m.wait(10) is False
    m.close()
    d = sg.Mirror.open_dense(256, DIM)
    with pytest.raises(sg.SplinterGpuError):
        d.wait(1)
    d.close()
