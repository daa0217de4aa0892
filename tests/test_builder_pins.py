"""The host grid builder's arrays, pinned: FNV-1a checksums of every array buildGrid produces (nodes, columns, bins, seeds,
permutation, vertex lists, column table, bounds) for four synthetic grids -- plain, shuffled record order, and shuffled with
degenerate records (zero layers, NaN corners, clockwise and repeated corners, unsorted/overlapping heights, duplicates).
tests/golden/builder_checksums.txt was written by round 1's builder (commit 96f64d0) through tools/time_build.cpp; the
round-2 builder (planes computed in place, sample sort, parallel linking) must reproduce it bit for bit."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "icon-ray-tracing_b200")


@pytest.fixture(scope="module")
def tool(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("tb") / "time_build")
    objs = [os.path.join(PKG, "build", o) for o in ("grid_build.o", "gen_grid.o")]
    for o in objs:
        assert os.path.exists(o), f"{o} missing: run __graft_entry__.build()"
    subprocess.run(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(PKG, "csrc"),
                    os.path.join(ROOT, "tools", "time_build.cpp"), *objs, "-o", out, "-lpthread"], check=True)
    return out


def _golden():
    cases, cur = {}, None
    with open(os.path.join(ROOT, "tests", "golden", "builder_checksums.txt")) as f:
        for line in f:
            if line.startswith("# "):
                cur = line[2:].strip()
                cases[cur] = []
            elif line.strip():
                cases[cur].append(line.rstrip("\n"))
    return cases


@pytest.mark.parametrize("case", list(_golden()))
def test_builder_arrays_are_bit_identical_to_the_pinned_builder(tool, case):
    got = subprocess.run([tool, *case.split()], check=True, capture_output=True, text=True).stdout.splitlines()
    assert got == _golden()[case]
