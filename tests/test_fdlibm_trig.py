"""CPU: csrc/fdlibm_trig.cuh (the asinf / atanf / atan2f the SPHERE_ACCEL_MODE ray-gen evaluates on the device) compiled
for the host without FMA contraction and swept against this machine's libm -- the functions the reference's own CPU build
calls in toSpherical (ICONGrid.h:36-42).  Same operation sequence, IEEE round-to-nearest single operations on both sides,
so host equality is device equality (the device build uses __fmul_rn/__fadd_rn/__fdiv_rn/__fsqrt_rn for every operation).
Default: every 251st float bit pattern + 8M atan2f pairs (seconds); ICB_TRIG_FULL=1 sweeps all 2^32 (about a minute on
8 cores; done once per change of the header, result recorded in DESIGN.md)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def sweep(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("trig") / "trig_sweep")
    subprocess.run(["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-pthread", "-I", os.path.join(ROOT, "icon-ray-tracing_b200", "csrc"),
                    os.path.join(ROOT, "tests", "host", "trig_sweep.cpp"), "-o", exe, "-lm"], check=True)
    return exe


def test_device_trig_sequence_is_glibc_bit_for_bit(sweep):
    full = os.environ.get("ICB_TRIG_FULL") == "1"
    out = subprocess.run([sweep, "1" if full else "251", "400000000" if full else "8000000"], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    rows = dict((l.split()[0], (int(l.split()[1]), int(l.split()[2]))) for l in out.stdout.strip().splitlines())
    for fn in ("asinf", "atanf", "atan2f"):
        tested, bad = rows[fn]
        assert bad == 0 and tested > 8_000_000


@pytest.fixture(scope="module")
def logf_sweep(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("logf") / "logf_sweep")
    subprocess.run(["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-pthread", "-I", os.path.join(ROOT, "icon-ray-tracing_b200", "csrc"),
                    os.path.join(ROOT, "tests", "host", "logf_sweep.cpp"), "-o", exe, "-lm"], check=True)
    return exe


def test_device_logf_sequence_is_glibc_bit_for_bit(logf_sweep):
    """csrc/glibc_logf.cuh (the free-flight logf(1-u) of the Woodcock ray-gens) against this machine's libm: all 2^24 arguments
    1 - k/2^24 the LCG can produce, and every 61st positive normal float (ICB_TRIG_FULL=1: all 2.13e9 of them); both of
    glibc's builds (FMA / SSE2) are restated, the kernels use the FMA form"""
    full = os.environ.get("ICB_TRIG_FULL") == "1"
    out = subprocess.run([logf_sweep, "1" if full else "61"], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    rows = dict((l.split()[0], [int(v) for v in l.split()[1:]]) for l in out.stdout.strip().splitlines())
    assert rows["draws"][0] == 1 << 24 and rows["draws"][1] == 0          # the form the kernels use
    assert rows["normals"][0] > 30_000_000 and rows["normals"][1] == 0
