"""bench.py's roofline object (no GPU needed): the fraction quoted is a measured quantity over the limit of the unit that
bounds the kernel -- never the algorithmic 60 B x samples over the HBM peak, which exceeds 1 on R2B8 -- and nothing is
quoted from an N = 1 capture when there is none for the line's workload / GPU count."""
import importlib.util
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
K = "renderKernel<false,false,false,false>"


def _bench():
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_fraction_never_exceeds_one_and_names_the_bound():
    b = _bench()
    # C3 as BENCH_r02 measured it: 1.40 G samples per launch, 10.6 ms per launch with four frames in flight
    r = b.roofline_block(b.profile_facts("C3", K), K, 1.4012e9, 10.62, 14.78, 6549.8, "measured", 1965.0)
    assert r["algorithmic_gbs"] > r["peak_hbm_gbs"]                    # the figure that used to be printed as frac 1.2
    assert r["bound"] == "issue" and 0.0 < r["frac"] <= 1.0 and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-12
    assert r["traffic"] < 60.0 * 1.4012e9 and 0.0 < r["dram_frac"] < 0.2
    # C5: measured DRAM bytes exceed the algorithmic bytes -> HBM bound, achieved = measured bytes / measured time
    r = b.roofline_block(b.profile_facts("C5", K), K, 1.766e9, 42.3, 50.0, 6549.8, "measured", 1965.0)
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and r["traffic"] >= 60.0 * 1.766e9
    assert 0.0 < r["frac"] <= 1.0 and abs(r["frac"] - r["traffic"] / 42.3e-3 / 1e9 / 6549.8) < 1e-9


def test_no_capture_no_fraction():
    b = _bench()
    assert b.profile_facts("C2", K) == {}
    r = b.roofline_block({}, K, 1e9, 5.0, 5.0, 6549.8, "measured", None)
    assert r["frac"] is None and r["achieved"] is None and r["traffic"] is None and r["algorithmic_gbs"] > 0
