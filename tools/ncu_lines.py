#!/usr/bin/env python3
"""Per-source-line summary of an .ncu-rep captured with --import-source on (read here,
no GPU needed): warp instructions executed, active lanes per instruction and stall
samples for every CUDA source line of the top kernel, hottest first, plus totals per
file.  This is where "instructions per node visit" and "active lanes in the NODE /
PRIM step" in DESIGN.md come from.

    python tools/ncu_lines.py gpurun_out/prof.ncu-rep [--top 40] [--ranges file:lo-hi:name ...]
"""
import argparse
import csv
import io
import subprocess
import sys


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep")
    ap.add_argument("--top", type=int, default=40)
    ap.add_argument("--ranges", nargs="*", default=[], help="file-suffix:first-last:name groups of source lines to total")
    a = ap.parse_args()
    raw = subprocess.run(["ncu", "-i", a.rep, "--page", "source", "--print-source", "cuda,sass", "--csv"],
                         check=True, capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    cur_file, hdr = None, None
    lines = []  # (file, line, source, samples, inst, thread_inst)
    for r in rows:
        if len(r) == 2 and r[0] == "File Path":
            cur_file = r[1]
            continue
        if r and r[0] == "Line No":
            hdr = r
            i_s, i_i, i_t = hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
            continue
        if hdr is None or not r or not r[0].isdigit():
            continue
        try:
            lines.append((cur_file, int(r[0]), r[1].strip(), int(r[i_s]), int(r[i_i]), int(r[i_t])))
        except ValueError:
            continue
    tot_s = sum(x[3] for x in lines) or 1
    tot_i = sum(x[4] for x in lines) or 1
    tot_t = sum(x[5] for x in lines)
    print(f"# {a.rep}: {tot_i} warp instructions, {tot_t / tot_i:.2f} active lanes per instruction, {tot_s} stall samples")
    per_file = {}
    for f, ln, src, s, i, t in lines:
        d = per_file.setdefault(f, [0, 0, 0])
        d[0] += s
        d[1] += i
        d[2] += t
    for f, (s, i, t) in sorted(per_file.items(), key=lambda kv: -kv[1][0]):
        print(f"file {f}: {100 * s / tot_s:5.1f}% samples, {100 * i / tot_i:5.1f}% instructions, {t / max(i, 1):5.2f} lanes")
    for spec in a.ranges:
        suffix, rng, name = spec.split(":")
        lo, hi = (int(x) for x in rng.split("-"))
        s = i = t = 0
        for f, ln, src, ss, ii, tt in lines:
            if f and f.endswith(suffix) and lo <= ln <= hi:
                s += ss
                i += ii
                t += tt
        print(f"range {name:28s} {suffix}:{lo}-{hi}: {100 * s / tot_s:5.1f}% samples, {100 * i / tot_i:5.1f}% instructions ({i}), {t / max(i, 1):5.2f} lanes")
    print(f"\n{'samples%':>8s} {'inst%':>6s} {'lanes':>6s}  line")
    for f, ln, src, s, i, t in sorted(lines, key=lambda x: -x[3])[:a.top]:
        short = (f or "?").split("/")[-1]
        print(f"{100 * s / tot_s:8.2f} {100 * i / tot_i:6.2f} {t / max(i, 1):6.2f}  {short}:{ln}: {src[:110]}")


if __name__ == "__main__":
    main()
