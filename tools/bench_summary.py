import json,sys
# usage: bench_summary.py FILE   (or pipe the bench output in)
src = open(sys.argv[1]) if len(sys.argv) > 1 else sys.stdin
for ln in src:
    if ln.startswith("{"):
        d=json.loads(ln)
        print("head ms", d["ms_per_step"], "GB/s", d["value"], "frac", d["roofline"]["frac"], "| e2e ms", d["e2e"]["ms_per_step"], "| parity", d["parity"].get("ok"), d["parity"].get("indices_identical"))
        for k,v in d.get("variants",{}).items(): print("   ", k, v["ms_per_step"], "ms", v["scanned_GB_per_s"], "GB/s frac", v["roofline_frac"])
        if "cpu_baseline" in d: print("    cpu", d["cpu_baseline"]["value"], "GB/s", d["cpu_baseline"]["cores"], "cores")
    else: print(ln.rstrip())
