#!/usr/bin/env python
"""Records the ncu counters bench.py quotes in its roofline (warp instructions, lanes per instruction, DRAM bytes per
launch of the trace kernel and of one round's scatter kernels) in profiles/traffic.json, keyed by workload and stamped
with the hash of the CUDA sources they were captured from (path_tracer_b200._build.source_hash).  bench.py refuses to
quote an entry whose hash differs from the build it runs.

usage: python tools/ncu_traffic.py C2 gpurun_out/r02_c2.ncu-rep [more: C3 gpurun_out/r02_c3.ncu-rep ...]
(the .ncu-rep is `ncu --set full --clock-control none` over a few launches of `bench.py --workload <key>`)"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def rows_of(path):
    """An .ncu-rep, or the CSV that `ncu -i REP --page raw --csv --print-units base` printed from it (the reports of five
    workloads together are too big to bring back from the GPU box; the CSVs are not)."""
    if path.endswith(".csv"):
        with open(path) as f:
            raw = f.read()
    else:
        raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv", "--print-units", "base"], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr = rows[0]
    return hdr, rows[2:]


def entry(hdr, r):
    g = lambda k: float(r[hdr.index(k)].replace(",", ""))
    return {"kernel": r[hdr.index("Kernel Name")], "warp_instructions_per_launch": g("smsp__inst_executed.sum"),
            "lanes_per_instruction": g("smsp__thread_inst_executed_per_inst_executed.ratio"),
            "dram_bytes_per_launch": g("dram__bytes_read.sum") + g("dram__bytes_write.sum"),
            "issue_active_pct": g("smsp__issue_active.avg.pct_of_peak_sustained_active"),
            "duration_ns_under_ncu": g("gpu__time_duration.sum")}


def main(argv):
    from path_tracer_b200 import _build
    path = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(path) as f:
            data = json.load(f)
    except Exception:
        data = {}
    for key, rep in zip(argv[0::2], argv[1::2]):
        hdr, rows = rows_of(rep)
        names = [r[hdr.index("Kernel Name")] for r in rows]
        traces = [i for i, n in enumerate(names) if "k_trace_coop" in n]
        if not traces:
            print("no k_trace_coop launch in", rep)
            continue
        first = traces[0]
        nxt = traces[1] if len(traces) > 1 else len(rows)
        scat = [entry(hdr, rows[i]) for i in range(first + 1, nxt) if "k_scatter" in names[i] or "k_classify" in names[i]]
        e = {"build_hash": _build.source_hash(), "source": "%s (ncu --set full --clock-control none, one round of bench.py --workload %s)" % (os.path.relpath(rep, ROOT), key),
             "trace": entry(hdr, rows[first])}
        if scat:
            e["scatter"] = {"kernels": [s["kernel"] for s in scat], "warp_instructions_per_launch": sum(s["warp_instructions_per_launch"] for s in scat),
                            "dram_bytes_per_launch": sum(s["dram_bytes_per_launch"] for s in scat),
                            "duration_ns_under_ncu": sum(s["duration_ns_under_ncu"] for s in scat)}
        data[key] = e
        print(key, json.dumps(e)[:300])
    with open(path, "w") as f:
        json.dump(data, f, indent=1)


if __name__ == "__main__":
    main(sys.argv[1:])
