#!/usr/bin/env python
"""Per-CUDA-source-line totals from `ncu -i rep --page source --print-source cuda,sass --csv`: warp instructions
executed, share of the kernel's instruction stream, average active lanes, stall samples.  Inlined device functions
appear under the file/line they were written at.
    ncu -i prof.ncu-rep --page source --print-source cuda,sass --csv > src.csv
    python tools/ncu_cuda_lines.py src.csv [top] [sort: inst|samples]"""
import csv
import sys


def main():
    path = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 50
    key = sys.argv[3] if len(sys.argv) > 3 else "inst"
    rows, cur_file, hdr = [], None, None
    for r in csv.reader(open(path)):
        if not r:
            continue
        if r[0] == "File Path":
            cur_file = r[1].split("/")[-1]
        elif r[0] == "Line No":
            hdr = {h: i for i, h in enumerate(r)}
        elif hdr and r[0] != "" and r[0].isdigit():
            try:
                inst = int(r[hdr["Instructions Executed"]]); thr = int(r[hdr["Thread Instructions Executed"]]); smp = int(r[hdr["# Samples"]])
            except (ValueError, IndexError):
                continue
            l2 = r[hdr["L2 Theoretical Sectors Global"]] if "L2 Theoretical Sectors Global" in hdr else "0"
            rows.append((cur_file, int(r[0]), r[1].strip(), inst, thr, smp, int(l2) if l2.isdigit() else 0))
    tot_i = sum(x[3] for x in rows); tot_s = sum(x[5] for x in rows); tot_t = sum(x[4] for x in rows)
    print(f"-- {len(rows)} source lines, {tot_i/1e9:.3f} G warp-instructions, {tot_t/max(tot_i,1):.2f} lanes/instruction, {tot_s} stall samples")
    by_file = {}
    for f, ln, src, inst, thr, smp, l2 in rows:
        a = by_file.setdefault(f, [0, 0, 0]); a[0] += inst; a[1] += smp; a[2] += l2
    for f, (i, s_, l2) in sorted(by_file.items(), key=lambda kv: -kv[1][0]):
        print(f"   {f:28s} inst {100*i/tot_i:5.1f}%  samples {100*s_/max(tot_s,1):5.1f}%  L2 sectors {l2/1e6:9.1f} M")
    k = 3 if key == "inst" else 5
    for f, ln, src, inst, thr, smp, l2 in sorted(rows, key=lambda x: -x[k])[:top]:
        print(f"{100*inst/tot_i:5.2f}% inst {inst/1e6:8.1f}M lanes {thr/max(inst,1):5.1f} | {100*smp/max(tot_s,1):5.2f}% smp | L2 {l2/1e6:7.1f}M | {f}:{ln}: {src[:110]}")


if __name__ == "__main__":
    main()
