#!/usr/bin/env python3
"""Per CUDA source line: warp instructions, active threads and stall samples, from
`ncu -i X.ncu-rep --page source --csv --print-source cuda,sass` (first kernel launch of the report).
usage: ncu_cuda_lines.py report.ncu-rep [top_n]"""
import csv
import io
import subprocess
import sys


def main(rep, n=40):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    fname = None
    hdr = None
    lines = []
    seen_files = set()
    for r in rows:
        if not r:
            continue
        if r[0] == "File Name":
            if r[1] in seen_files:
                break  # second launch
            fname = r[1].split("/")[-1]
            seen_files.add(r[1])
            hdr = None
            continue
        if r[0] == "Line No":
            hdr = {h: i for i, h in enumerate(r)}
            continue
        if hdr and r[0].isdigit() and len(r) > 8:
            try:
                ex = int(r[hdr["Instructions Executed"]])
                th = int(r[hdr["Thread Instructions Executed"]])
                sm = int(r[hdr["# Samples"]])
            except (ValueError, KeyError):
                continue
            lines.append((fname, int(r[0]), r[1].strip()[:90], ex, th, sm))
    tot = sum(l[3] for l in lines)
    tots = sum(l[5] for l in lines)
    print("warp instructions %d, samples %d" % (tot, tots))
    for f, ln, src, ex, th, sm in sorted(lines, key=lambda l: -l[3])[:n]:
        print("%-16s %4d inst %5.2f%% thr %5.1f smp %5.2f%%  %s" % (f, ln, 100.0 * ex / tot, th / max(ex, 1), 100.0 * sm / max(tots, 1), src))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
