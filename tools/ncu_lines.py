#!/usr/bin/env python
"""Join an `ncu --page source --csv` export (per-SASS-instruction counters) with the line table of the cubin
(`nvdisasm --print-line-info`) and print per-source-line totals: warp instructions, active lanes, stall samples.

    cuobjdump -xelf all libicon_b200.so; nvdisasm --print-line-info kernels.sm_100a.cubin > k.sass
    ncu -i prof.ncu-rep --page source --csv > src.csv
    python tools/ncu_lines.py k.sass src.csv renderKernelILb0 [top]
"""
import collections
import csv
import re
import sys


def line_table(sass_path, func_substr):
    table, cur, inside = {}, None, False
    for ln in open(sass_path, errors="replace"):
        if ln.startswith(".text."):
            inside = func_substr in ln
            continue
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            table[int(m.group(1), 16)] = (cur, m.group(2).strip())
    return table


def main():
    sass_path, csv_path, func = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    table = line_table(sass_path, func)
    rows = list(csv.reader(open(csv_path)))
    hdr, data = rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    base = int(data[0][ix["Address"]], 16) if data[0][ix["Address"]].startswith("0x") else int(data[0][ix["Address"]])
    agg = collections.defaultdict(lambda: [0, 0, 0, 0])   # warp inst, thread inst, samples, long_sb samples
    tot = [0, 0, 0]
    for r in data:
        a = r[ix["Address"]]
        off = (int(a, 16) if a.startswith("0x") else int(a)) - base
        key = table.get(off, (None, ""))[0] or ("?", 0)
        wi, ti, sm = int(r[ix["Instructions Executed"]]), int(r[ix["Thread Instructions Executed"]]), int(r[ix["# Samples"]])
        lsb = int(r[ix["stall_long_sb"]] or 0) if "stall_long_sb" in ix else 0
        v = agg[key]
        v[0] += wi; v[1] += ti; v[2] += sm; v[3] += lsb
        tot[0] += wi; tot[1] += ti; tot[2] += sm
    print(f"total: {tot[0]/1e9:.2f} G warp-inst, {tot[1]/1e9:.1f} G thread-inst ({tot[1]/max(tot[0],1):.1f} lanes/inst), {tot[2]} samples")
    print("-- by warp instructions")
    for key, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{key[0]}:{key[1]:<5d} inst {v[0]/1e6:9.1f}M ({v[0]/tot[0]*100:5.2f}%) lanes {v[1]/max(v[0],1):5.1f}  samples {v[2]/tot[2]*100:5.2f}% (long_sb {v[3]/max(tot[2],1)*100:5.2f}%)")
    print("-- by stall samples")
    for key, v in sorted(agg.items(), key=lambda kv: -kv[1][2])[:top]:
        print(f"{key[0]}:{key[1]:<5d} samples {v[2]/tot[2]*100:5.2f}% (long_sb {v[3]/max(tot[2],1)*100:5.2f}%) inst {v[0]/1e6:9.1f}M lanes {v[1]/max(v[0],1):5.1f}")


if __name__ == "__main__":
    main()
