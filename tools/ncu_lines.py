#!/usr/bin/env python
"""Per-source-line summary of one kernel from an .ncu-rep captured with --import-source on.
usage: python tools/ncu_lines.py REPORT KERNEL_REGEX [TOP]
Prints, per source line, warp instructions executed, average active lanes and stall samples."""
import csv
import io
import os
import subprocess
import sys


def load(path, kernel):
    raw = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "cuda,sass", "-k", "regex:" + kernel],
                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    rows, cur, hdr = [], None, None
    for r in csv.reader(io.StringIO(raw)):
        if not r:
            continue
        if r[0] == "File Path":
            cur = os.path.basename(r[1]); continue
        if r[0] == "Line No":
            hdr = r; continue
        if hdr is None or cur is None or r[0] == "":
            continue            # SASS rows repeat the line's numbers
        try:
            rows.append({"file": cur, "line": int(r[0]), "src": r[1].strip(), "samples": int(r[hdr.index("# Samples")]),
                         "inst": int(r[hdr.index("Instructions Executed")]), "tinst": int(r[hdr.index("Thread Instructions Executed")])})
        except ValueError:
            pass
    return rows


def main():
    rows = load(sys.argv[1], sys.argv[2])
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    ti = sum(r["inst"] for r in rows) or 1
    ts = sum(r["samples"] for r in rows) or 1
    print("# kernel %s: %d warp instructions, %.1f lanes/instr, %d samples" % (sys.argv[2], ti, sum(r["tinst"] for r in rows) / ti, ts))
    byfile = {}
    for r in rows:
        f = byfile.setdefault(r["file"], [0, 0, 0]); f[0] += r["inst"]; f[1] += r["tinst"]; f[2] += r["samples"]
    for f, v in sorted(byfile.items(), key=lambda kv: -kv[1][2]):
        print("file %-28s inst %5.1f%%  lanes %5.1f  samples %5.1f%%" % (f, 100.0 * v[0] / ti, v[1] / max(1, v[0]), 100.0 * v[2] / ts))
    for r in sorted(rows, key=lambda r: -r["samples"])[:top]:
        print("%-18s %4d inst %5.2f%% lanes %5.1f samples %5.2f%% | %s" % (r["file"], r["line"], 100.0 * r["inst"] / ti, r["tinst"] / max(1, r["inst"]),
                                                                          100.0 * r["samples"] / ts, r["src"][:90]))


if __name__ == "__main__":
    main()
