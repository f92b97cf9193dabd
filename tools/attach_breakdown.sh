#!/bin/bash
# developer helper (GPU box): where the one-shot CLI's time goes at 4096 slots
python - <<'PY'
import os, sys, subprocess, time
sys.path.insert(0, "tools"); sys.path.insert(0, "tests")
import cli_wallclock as w, support
q = support.synth_queries(4321, 1, 768)[0]; q.tofile("/tmp/q.f32")
for slots in (4096, 1 << 20):
    name = f"/splb200_brk_{slots}"
    path = w.make_store(name, slots, slots - 6)
    try:
        for rep in range(2):
            stub = subprocess.Popen([str(w.REF / "ref_embed_stub"), name, "/tmp/q.f32", "60000"])
            time.sleep(0.05)
            t0 = time.perf_counter()
            r = subprocess.run([str(w.ARMS["gpu_module"]), "--use", name, "search", "what", "--limit", "10", "--json"],
                               capture_output=True, text=True, env={**os.environ, "SPL_GPU_DEBUG_ATTACH": "1"})
            dt = (time.perf_counter() - t0) * 1e3
            stub.wait()
            print(slots, f"{dt:.0f} ms", r.stderr.strip().splitlines()[-1:])
    finally:
        os.unlink(path)
PY
