#!/usr/bin/env python
"""Summarise an .ncu-rep per CUDA source line: executed warp instructions, stall samples.

  python scripts/ncu_lines.py gpurun_out/prof.ncu-rep [--top 40] [--kernel-index 0] [--kernel-name REGEX]

Reads `ncu --page source --csv --print-source cuda,sass` (needs -lineinfo at compile time and
--import-source on at capture time) and prints the hottest source lines of one profiled launch.
"""
import csv
import subprocess
import sys


def main():
    rep = sys.argv[1]
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 40
    kidx = int(sys.argv[sys.argv.index("--kernel-index") + 1]) if "--kernel-index" in sys.argv else 0
    flt = ["-k", "regex:" + sys.argv[sys.argv.index("--kernel-name") + 1]] if "--kernel-name" in sys.argv else []
    out = subprocess.run(["ncu", "-i", rep, *flt, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    # split into kernels: a kernel restarts when a (file, function) header pair repeats for the first file
    lines = {}   # (kernel_no, file, line) -> [src, inst, samples, thread_inst]
    cur_file, kernel_no, seen = None, -1, set()
    hdr = None
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            cur_file = r[1].split("/")[-1]
            continue
        if r[0] == "Function Name":
            key = (cur_file, r[1])
            if key in seen or kernel_no < 0:
                kernel_no += 1
                seen = set()
            seen.add(key)
            continue
        if r[0] == "Line No":
            hdr = r
            i_inst = hdr.index("Instructions Executed")
            i_samp = hdr.index("# Samples")
            i_thr = hdr.index("Thread Instructions Executed")
            continue
        if hdr is None or r[0] in ("", "..."):
            continue
        try:
            ln = int(r[0])
            inst = int(r[i_inst])
            samp = int(r[i_samp])
            thr = int(r[i_thr])
        except ValueError:
            continue
        k = (kernel_no, cur_file, ln)
        if k in lines:
            lines[k][1] += inst
            lines[k][2] += samp
            lines[k][3] += thr
        else:
            lines[k] = [r[1], inst, samp, thr]
    sel = {k: v for k, v in lines.items() if k[0] == kidx}
    tot_i = sum(v[1] for v in sel.values()) or 1
    tot_s = sum(v[2] for v in sel.values()) or 1
    print(f"kernel #{kidx}: {tot_i} warp instructions, {tot_s} stall samples, "
          f"avg active threads {sum(v[3] for v in sel.values()) / tot_i:.1f}")
    print(f"{'file:line':28s} {'inst%':>6s} {'samp%':>6s} {'thr/inst':>8s}  source")
    for (_, f, ln), v in sorted(sel.items(), key=lambda kv: -kv[1][2])[:top]:
        print(f"{f + ':' + str(ln):28s} {100 * v[1] / tot_i:6.2f} {100 * v[2] / tot_s:6.2f} {v[3] / max(v[1], 1):8.1f}  {v[0].strip()[:90]}")


if __name__ == "__main__":
    main()
