"""Wall-clock of the command a user runs: the reference's stock `splinterctl search` against the same
dispatcher with the GPU search module swapped in (oracle/_ref/gpu/splinterctl), on a 4096-slot store (config C1) and
on a 2^20-slot store (config C2's size), one-shot and inside one REPL session (first query pays CUDA context + full
upload, later ones an incremental sync).

Both arms talk to the same stand-in for the splinference daemon (oracle/_ref/ref_embed_stub), which answers a
handshake 30 ms after it sees it (so that the client is inside splinter_poll) and polls the store every 2 ms: that
constant sits in every number below and is reported as `handshake_floor_ms` (a search over an empty result set).

    python tools/cli_wallclock.py [--slots 4096,1048576] [--reps 5] > profiles/r2_cli_wallclock.json
"""
import argparse
import json
import os
import subprocess
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tests"))
import support  # noqa: E402

REF = support.REF_DIR
ARMS = {"reference_cpu": REF / "splinterctl", "gpu_module": REF / "gpu" / "splinterctl"}
REPL = {"reference_cpu": REF / "splinter_cli", "gpu_module": REF / "gpu" / "splinter_cli"}


def make_store(name, slots, nkeys):
    img = support.new_image(slots, 64, 768)
    support.image_fill_synthetic(img, nkeys)
    path = "/dev/shm" + name
    with open(path, "wb") as f:
        f.write(img.tobytes())
    return path


def one_shot(binary, name, qfile, flags, reps):
    times = []
    out = None
    for _ in range(reps):
        stub = subprocess.Popen([str(REF / "ref_embed_stub"), name, qfile, "60000"])
        time.sleep(0.05)
        t0 = time.perf_counter()
        r = subprocess.run([str(binary), "--use", name, "search", "what is splinter?"] + flags, capture_output=True, text=True)
        times.append((time.perf_counter() - t0) * 1e3)
        stub.wait()
        assert r.returncode == 0, r.stderr
        out = r.stdout
    return times, out


def repl_session(binary, name, qfile, flags, nqueries):
    """Type nqueries searches into ONE REPL process and time each from the keystroke to the closing brace of its JSON.
    The child's stdout is a pty so that its libc flushes line by line (into a pipe it would hold everything back
    until exit)."""
    import pty
    import select
    stub = subprocess.Popen([str(REF / "ref_embed_stub"), name, qfile, "120000", str(nqueries)])
    master, slave = pty.openpty()
    p = subprocess.Popen([str(binary), "--use", name], stdin=subprocess.PIPE, stdout=slave, stderr=subprocess.DEVNULL)
    os.close(slave)
    times, outs = [], []
    line = ('search "what is splinter?" ' + " ".join(flags) + " --json\n").encode()
    try:
        for _ in range(nqueries):
            t0 = time.perf_counter()
            p.stdin.write(line)
            p.stdin.flush()
            buf = b""
            while not buf.replace(b"\r\n", b"\n").endswith(b"\n}\n"):
                r, _, _ = select.select([master], [], [], 120)
                if not r:
                    raise RuntimeError("REPL did not answer within 120 s: " + buf[-200:].decode(errors="replace"))
                buf += os.read(master, 65536)
            times.append((time.perf_counter() - t0) * 1e3)
            outs.append(buf.replace(b"\r\n", b"\n").decode())
        p.stdin.close()
        p.wait(timeout=60)
    finally:
        if p.poll() is None:
            p.kill()
        os.close(master)
        stub.wait(timeout=130)
    return times, outs


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--slots", default="4096,1048576")
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    res = {"what": "wall-clock ms of `search \"...\" --limit 10 --json` through the reference's dispatcher; median of reps",
           "host_cores": os.cpu_count(), "cases": []}
    q = support.synth_queries(4321, 1, 768)[0]
    qfile = "/tmp/spl_cli_q.f32"
    q.tofile(qfile)
    for slots in [int(s) for s in a.slots.split(",")]:
        name = f"/splb200_wall_{slots}"
        path = make_store(name, slots, slots - 6)
        try:
            case = {"slots": slots, "store_bytes": os.path.getsize(path)}
            for flags, tag in ((["--limit", "10"], "limit10"), (["--limit", "10", "--bloom", "3"], "limit10_bloom3")):
                entry = {}
                outs = {}
                for arm, binary in ARMS.items():
                    t, outs[arm] = one_shot(binary, name, qfile, flags + ["--json"], a.reps)
                    entry[arm + "_one_shot_ms"] = round(float(np.median(t)), 2)
                entry["identical_stdout"] = outs["reference_cpu"] == outs["gpu_module"]
                for arm, binary in REPL.items():
                    t, o = repl_session(binary, name, qfile, flags, 4)
                    entry[arm + "_repl_first_ms"] = round(t[0], 2)
                    entry[arm + "_repl_later_ms"] = round(float(np.median(t[1:])), 2)
                case[tag] = entry
            # what the handshake alone costs: a bloom mask (almost) no slot carries -> nothing to score in either arm
            t, _ = one_shot(ARMS["reference_cpu"], name, qfile, ["--bloom", "7fffffffffffffff", "--limit", "1", "--json"], a.reps)
            case["handshake_floor_ms_reference"] = round(float(np.median(t)), 2)
            res["cases"].append(case)
        finally:
            os.unlink(path)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
