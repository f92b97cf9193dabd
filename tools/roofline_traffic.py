#!/usr/bin/env python
"""Write profiles/roofline_traffic.json from an `ncu --set full` capture of the C2 headline kernel.

    python tools/roofline_traffic.py gpurun_out/r2_scan_tma.ncu-rep

The file records which kernel SOURCES the capture belongs to (sha16 over the scan kernel's files); bench.py reports the
traffic only while that still matches the tree it runs from, so a stale capture reads as `traffic: null`, never as a
measurement of a kernel it was not taken from."""
import csv, hashlib, json, subprocess, sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
KERNEL_FILES = ["libsplinter_b200/csrc/spl_scan_tma.cu", "libsplinter_b200/csrc/spl_scan_epilogue.cuh",
                "libsplinter_b200/csrc/spl_device.cuh", "libsplinter_b200/csrc/spl_scan.cuh"]


def kernel_sha16():
    h = hashlib.sha256()
    for f in KERNEL_FILES:
        h.update((ROOT / f).read_bytes())
    return h.hexdigest()[:16]


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr = rows[0]
    per = []
    for r in rows[2:]:
        rd, wr = float(r[hdr.index("dram__bytes_read.sum")]), float(r[hdr.index("dram__bytes_write.sum")])
        unit_r, unit_w = rows[1][hdr.index("dram__bytes_read.sum")], rows[1][hdr.index("dram__bytes_write.sum")]
        scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        per.append((r[hdr.index("Kernel Name")], rd * scale[unit_r] + wr * scale[unit_w],
                    float(r[hdr.index("gpu__time_duration.sum")])))
    tp = ROOT / "profiles" / "roofline_traffic.json"
    d = json.loads(tp.read_text()) if tp.exists() else {}
    d["c2_mask3_dram_bytes_per_launch"] = int(sum(p[1] for p in per) / len(per))
    d["c2_mask3_capture"] = {"report": Path(rep).name, "launches": len(per), "kernel": per[0][0][:60],
                             "kernel_source_sha16": kernel_sha16(), "files": KERNEL_FILES}
    tp.write_text(json.dumps(d, indent=2))
    print(json.dumps(d["c2_mask3_capture"]), d["c2_mask3_dram_bytes_per_launch"])


if __name__ == "__main__":
    main()
