#!/usr/bin/env python
"""L2 sectors requested per GATHER of the hot path (north_star: cell / vertex(plane) / height / scalar gathers), from the
source-level page of an ncu capture: every global-memory instruction is attributed to the source line it was written at
(`ncu -i rep --page source --print-source cuda,sass --csv`), the line to its enclosing function (from the sources of the
captured build), and the functions are grouped by what they fetch.  ncu exposes hit/miss per kernel only
(lts__t_sector_hit_rate.pct), not per instruction; the per-gather figure it does expose is the sectors each load site asks
L2 for, and how far that is from the ideal (coalesced) count.
    python tools/ncu_gathers.py src.csv locator.cuh kernels.cu [kernel-level L2 hit rate %] [prism|logtable] > json"""
import csv
import json
import re
import sys

GATHERS = [
    ("node fetch (RecBlock: side planes, radius range, neighbour links)", ["loadNode"]),
    ("line fetch (heights | values of a record, cp.async into the lane's slot)", ["linesFetch"]),
    ("prism mask word (travels with the lines)", ["recLinesFetch"]),
    ("slot copies: SLOTCOPY", ["slotCopy", "slotCopy4"]),
    ("prism tail / vertex scalar rows (prism sampler)", ["prismFetchTail", "prismFetchScalars", "vertexScalarAt", "prismSample"]),
    ("log table (free-flight distances)", ["woodcockPixel", "wgStageLog", "trackStep"]),
    ("majorants / occupancy bits", ["wgStep", "wsStep"]),
    ("walk seeds / candidate bins / column scans (exact query, seeds, gaps)", ["seedState", "exactQueryId", "stageFetch", "columnGap", "globalsQuery", "exactQuery"]),
    ("tile order / work cursor / tile list / transfer function staging", ["renderKernel", "woodcockKernel", "woodcockGridKernel", "woodcockSphereKernel"]),
    ("frame buffer, accumulation buffer", ["writePixel"]),
]


def spans(path):
    out, cur = [], None
    pat = re.compile(r"^(?:template.*>\s*)?(?:__device__|__global__|static|inline)[^;(]*?\b(\w+)\(")
    lines = open(path, errors="replace").read().split("\n")
    for i, ln in enumerate(lines, 1):
        m = pat.match(ln)
        if m and not ln.rstrip().endswith(";"):
            cur = m.group(1)
        elif re.match(r"^(\w+)\(const DeviceGrid", ln):            # kernel name on its own line after __launch_bounds__
            cur = re.match(r"^(\w+)\(", ln).group(1)
        out.append(cur)
    return out


def main():
    path, loc, ker = sys.argv[1:4]
    hit = float(sys.argv[4]) if len(sys.argv) > 4 else None
    slotcopy_is = sys.argv[5] if len(sys.argv) > 5 else "prism"
    fn = {"locator.cuh": spans(loc), "kernels.cu": spans(ker)}
    cur_file, hdr, rows = None, None, []
    for r in csv.reader(open(path)):
        if not r:
            continue
        if r[0] == "File Path":
            cur_file = r[1].split("/")[-1]
        elif r[0] == "Line No":
            hdr = {h: i for i, h in enumerate(r)}
        elif hdr and r[0].isdigit():
            def num(k):
                v = r[hdr[k]] if k in hdr and hdr[k] < len(r) else "0"
                return int(v) if v.isdigit() else 0
            sec, ideal, req = num("L2 Theoretical Sectors Global"), num("L2 Theoretical Sectors Global Ideal"), num("L1 Tag Requests Global")
            if sec or req:
                ln = int(r[0])
                f = fn.get(cur_file)
                func = f[ln - 1] if f and ln - 1 < len(f) else None
                rows.append((cur_file, ln, r[1].strip(), sec, ideal, req, num("# Samples"), func or cur_file))
    tot = sum(x[3] for x in rows)
    out = {"total_l2_sectors_requested": tot, "kernel_lts_t_sector_hit_rate_pct": hit, "gathers": [], "other": []}
    used = set()
    for name, funcs in GATHERS:
        if "SLOTCOPY" in name:
            name = "vertex scalar rows (prism sampler, cp.async into the slot)" if slotcopy_is == "prism" else "log table entries staged into the slot (cp.async)"
        sec = ideal = req = smp = 0
        for i, row in enumerate(rows):
            if i not in used and row[7] in funcs:
                used.add(i); sec += row[3]; ideal += row[4]; req += row[5]; smp += row[6]
        if sec or req:
            out["gathers"].append({"gather": name, "functions": funcs, "l2_sectors": sec, "share_of_sectors": sec / max(tot, 1), "ideal_sectors": ideal,
                                   "l1_tag_requests": req, "stall_samples_at_these_lines": smp})
    for i, row in enumerate(rows):
        if i not in used and row[3] > tot * 0.002:
            out["other"].append({"line": f"{row[0]}:{row[1]}", "function": row[7], "source": row[2][:100], "l2_sectors": row[3]})
    json.dump(out, sys.stdout, indent=1)


if __name__ == "__main__":
    main()
