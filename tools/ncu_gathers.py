#!/usr/bin/env python
"""L2 sectors requested per GATHER of the hot path (north_star: cell / vertex(plane) / height / scalar gathers), from the
source-level page of an ncu capture: every global-memory instruction is attributed to the source line it was written at
(`ncu -i rep --page source --print-source cuda,sass --csv`), and the lines are grouped by what they fetch.  ncu exposes
hit/miss per kernel only (lts__t_sector_hit_rate.pct), not per instruction; the per-gather figure it does expose is the
sectors each load site asks L2 for, and how far that is from the ideal (coalesced) count.
    python tools/ncu_gathers.py src.csv [kernel-level hit rate %] > json"""
import csv
import json
import re
import sys

GATHERS = [
    ("node fetch (RecBlock: side planes, radius range, neighbour links)", [r"__ldg\(q \+ \d\)", r"reinterpret_cast<const int4 \*>\(q\) \+ 4", r"loadNode"]),
    ("line fetch (heights | values of a record, cp.async into the lane's slot)", [r"cp\.async\.ca\.shared\.global \[%0\], \[%1\], 16", r"linesFetch"]),
    ("prism mask word (travels with the lines)", [r"cp\.async\.ca\.shared\.global \[%0\], \[%1\], 4"]),
    ("vertex scalars (prism sampler)", [r"prismFetchScalars", r"slotCopy", r"vtxScalars"]),
    ("log table (free-flight distances)", [r"logTable"]),
    ("majorants / occupancy bits", [r"maxOpacities", r"occupied"]),
    ("walk seeds / candidate bins (exact query)", [r"g\.seed", r"binOff", r"binList", r"stageFetch", r"recs \+ r0", r"qv \+ \d"]),
    ("tile order / work cursor / tile list", [r"order \+", r"tileList \+", r"tileCounter", r"atomicAdd"]),
    ("frame buffer, accumulation buffer, counters", [r"writePixel", r"out\.fb", r"accum", r"perPixel", r"counters"]),
]


def main():
    path = sys.argv[1]
    hit = float(sys.argv[2]) if len(sys.argv) > 2 else None
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
                rows.append((cur_file, int(r[0]), r[1].strip(), sec, ideal, req, num("# Samples")))
    tot = sum(x[3] for x in rows)
    out = {"total_l2_sectors_requested": tot, "kernel_lts_t_sector_hit_rate_pct": hit, "gathers": [], "unattributed": []}
    used = set()
    for name, pats in GATHERS:
        sec = ideal = req = smp = 0
        lines = []
        for i, (f, ln, src, s_, id_, rq, sm) in enumerate(rows):
            if i in used:
                continue
            if any(re.search(p, src) for p in pats):
                used.add(i); sec += s_; ideal += id_; req += rq; smp += sm; lines.append(f"{f}:{ln}")
        if sec or req:
            out["gathers"].append({"gather": name, "l2_sectors": sec, "share_of_sectors": sec / max(tot, 1), "ideal_sectors": ideal,
                                   "l1_tag_requests": req, "stall_samples_at_these_lines": smp, "lines": lines[:12]})
    for i, (f, ln, src, s_, id_, rq, sm) in enumerate(rows):
        if i not in used and s_ > tot * 0.002:
            out["unattributed"].append({"line": f"{f}:{ln}", "source": src[:100], "l2_sectors": s_})
    json.dump(out, sys.stdout, indent=1)


if __name__ == "__main__":
    main()
