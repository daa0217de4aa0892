#!/usr/bin/env python
"""bench.py -- samples/s (and frames/s) of the icon_rt hot path on B200.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (C ABI, libicon_b200.so)
    python bench.py --impl reference --gpus N --steps K ...   # the reference's own CPU code (oracle/_ref) on host cores

A step = one frame: every pixel's ray is marched through the ICON prism shells (locate -> sample -> transfer
function -> front-to-back compositing).  Workload (BASELINE.json configs[2], the one `metric` is quoted on):
synthetic R2B8 x 90 levels (5 242 880 columns = 15 728 640 `.ic` records), 3840x2160, sharp/high-opacity
transfer function, orbit camera (a new pose every step).  `value` = samples/s with everything resident in HBM;
`e2e` = the same through icb_render_host with HOST buffers (camera/params up, RGBA8 frame down into pinned memory).
N > 1: sort-first tiles, grid built once on rank 0 and broadcast (replicated per GPU), frame loop + NCCL tile gather behind
the C ABI (icb_group_*), strong scaling: one frame.  The K poses are looped `repeats` times so that the timed region is
at least --min-region seconds at every N; ms_per_step = region / (K * repeats).  Besides the throughput with several
frames in flight, `latency_ms` is one frame at a time.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "icon-ray-tracing_b200"))

WORKLOADS = {
    # name: (n, k, levels, width, height, tf, description)
    "C3": (2, 8, 90, 3840, 2160, "sharp", "R2B8x90 (5.24M columns, 15.7M records) 3840x2160 DVR, sharp TF"),
    "R2B9": (2, 9, 90, 3840, 2160, "sharp", "R2B9x90 (21.0M columns, 62.9M records; a quarter of C5's R2B10) 3840x2160 DVR, sharp TF"),
    "C5": (2, 10, 90, 3840, 2160, "sharp", "R2B10x90 (83.9M columns, 251.7M records, 71.5 GB of records -> 96.6 GB in HBM) 3840x2160 DVR, sharp TF"),
    "C2": (2, 6, 90, 1920, 1080, "smooth", "R2B6x90 (327 680 columns, 983 040 records) 1920x1080 DVR, smooth TF"),
    "C1": (2, 4, 20, 512, 512, "smooth", "R2B4x20 (20 480 records) 512x512 orbit view, smooth TF"),
}
BYTES_PER_SAMPLE = 60.0      # algorithmic bytes per located sample


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def traffic_per_launch(workload, kernel="renderKernel<false,false>"):
    """dram__bytes_read.sum + dram__bytes_write.sum of the render kernel from the committed ncu capture, if any"""
    p = os.path.join(ROOT, "profiles", "render_traffic.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        if d.get("workload") == workload:
            if kernel in d.get("kernels", {}):
                return d["kernels"][kernel]
            if kernel == "renderKernel<false,false>":
                return d.get("dram_bytes_per_launch")
    return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md clocks line)"""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


ZOOM_FOVY_DEG = 35.0     # from the viewAll eye the globe (angular radius 16.8 deg) then fills the frame height


def orbit_cameras(icb, b, W, H, n, zoom=True):
    """SURVEY §8d camera: Camera::viewAll eye (default 90 deg pose) orbited about +y from 30 deg, 2 deg per step;
    zoom=True narrows fovy to 35 deg from the same eye so that the volume fills the frame (viewAll's own 90 deg
    leaves the globe on ~6 % of the pixels)."""
    fovy = np.float32(np.deg2rad(ZOOM_FOVY_DEG)) if zoom else icb.FOVY_DEFAULT
    return [icb.camera_orbit(b["volLo"], b["volHi"], W, H, 30.0 + 2.0 * i, fovy=fovy, pose_fovy=icb.FOVY_DEFAULT) for i in range(n)]


def sample_positions(icb, b, cams, W, H, dt, n, seed=12345):
    """positions the marcher would sample for this workload: random pixels of the bench cameras, random step
    indices inside the pixel's shell segments (host arithmetic; only used to feed the CPU baseline)."""
    rng = np.random.default_rng(seed)
    r_in, r_out = float(b["sphLo"][0]), float(b["sphHi"][0])
    out = []
    while len(out) < n:
        cam = cams[rng.integers(len(cams))]
        x, y = rng.integers(W), rng.integers(H)
        d = cam["d00"].astype(np.float64) + (x + 0.5) * cam["du"] + (y + 0.5) * cam["dv"]
        d /= np.linalg.norm(d)
        o = cam["org"].astype(np.float64)
        bq, cq_out, cq_in = np.dot(o, d), np.dot(o, o) - r_out ** 2, np.dot(o, o) - r_in ** 2
        disc = bq * bq - cq_out
        if disc <= 0:
            continue
        t0, t1 = -bq - np.sqrt(disc), -bq + np.sqrt(disc)
        disc_in = bq * bq - cq_in
        if disc_in > 0:
            t1 = -bq - np.sqrt(disc_in)            # front segment only (what the sharp TF lets a ray see)
        if t1 <= max(t0, 0):
            continue
        t = max(t0, 0) + (rng.integers(max(1, int((t1 - max(t0, 0)) / dt))) + 0.5) * dt
        out.append(o + d * t)
    return np.asarray(out, np.float32)


def cpu_reference_rate(cells, pos, lut, budget_s, threads=0):
    """the reference's own CPU code (oracle/_ref: icon_rt/deviceCode.cu sampleVolume loop + postClassify compiled
    in place), or the C port when _ref is absent, over `pos`, all host threads; returns samples/s and a description"""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    if ol.Ref.available():
        oracle, kind = ol.Ref(), "reference"
    else:
        oracle, kind = ol.Port(), "port"
    cores = oracle.num_threads()
    done, t_used, chunk = 0, 0.0, max(cores * 2, 16)
    all_ids, all_vals = [], []
    while done < len(pos) and t_used < budget_s:
        p = pos[done:done + chunk]
        t0 = time.perf_counter()
        ids, vals = oracle.locate(cells, p)
        oracle.post_classify(lut, 0.0, 1.0, 1.0, vals)
        dt = time.perf_counter() - t0
        t_used += dt
        done += len(p)
        all_ids.append(ids); all_vals.append(vals)
        if dt < 0.5:
            chunk *= 2
    return done / max(t_used, 1e-9), kind, cores, done, t_used, np.concatenate(all_ids), np.concatenate(all_vals)


def run_reference(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import icb
    n, k, levels, W, H, tf, desc = WORKLOADS[wl]
    cells = icb.gen_grid(n, k, levels)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    oracle = ol.Ref() if ol.Ref.available() else ol.Port()
    b = oracle.grid_bounds(cells[:: max(1, len(cells) // 200000)])      # radius range only feeds the sample positions
    b = {"sphLo": b["sphLo"], "sphHi": b["sphHi"], "volLo": -np.full(3, b["sphHi"][0], np.float32), "volHi": np.full(3, b["sphHi"][0], np.float32)}
    cams = orbit_cameras(icb, b, W, H, args.steps + args.warmup)
    lut = icb.sharp_lut() if tf == "sharp" else icb.smooth_lut()
    cores = oracle.num_threads()
    # size a step from a probe so that the whole run stays near 90 s whatever the grid size and core count
    probe_pos = sample_positions(icb, b, cams, W, H, 500.0, 2 * cores, seed=1)
    t0 = time.perf_counter()
    oracle.locate(cells, probe_pos)
    rate = len(probe_pos) / max(time.perf_counter() - t0, 1e-6)
    step_s = min(5.0, 90.0 / (args.steps + args.warmup))
    per_step = int(min(1 << 20, max(cores, rate * step_s)))
    pos = sample_positions(icb, b, cams, W, H, 500.0, per_step * (args.steps + args.warmup))
    times = []
    for s in range(args.steps + args.warmup):
        p = pos[s * per_step:(s + 1) * per_step]
        t0 = time.perf_counter()
        ids, vals = oracle.locate(cells, p)
        oracle.post_classify(lut, 0.0, 1.0, 1.0, vals)
        dt = time.perf_counter() - t0
        if s >= args.warmup:
            times.append(dt)
    total = sum(times)
    value = per_step * len(times) / total
    kind = "reference" if ol.Ref.available() else "port"
    line = {"impl": "reference", "metric": "samples/s", "value": value, "unit": "samples/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{wl}: {desc}", "dt_m": 500.0},
            "cpu_baseline": {"value": value, "unit": "samples/s", "cores": cores, "kind": kind,
                             "sample": f"{per_step} sample positions per step drawn from the bench frames' rays, located by the "
                                       f"reference's brute-force sampleVolume loop over all {len(cells)} records + postClassify"},
            "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def flythrough_cameras(icb, b, W, H, n):
    """BASELINE configs[3] / SURVEY §8d C4: eye inside the atmosphere shell at mid height, moving along the equator and
    looking along the tangent (up = the local vertical), one pose per timestep"""
    r = 0.5 * (float(b["sphLo"][0]) + float(b["sphHi"][0]))
    cams = []
    for i in range(n):
        phi = 2.0 * np.pi * i / 96.0
        eye = np.array([r * np.cos(phi), r * np.sin(phi), 0.0], np.float32)
        tang = np.array([-np.sin(phi), np.cos(phi), 0.0], np.float32)
        cams.append(icb.camera_orient(eye, eye + tang * np.float32(1e6), eye / np.float32(r), W, H))
    return cams


def _dist_init(torch, world, local):
    if world <= 1:
        return None
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return dist


def _reduce(torch, dist, vals, op="max"):
    t = torch.tensor([float(v) for v in vals], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.SUM)
    return [float(x) for x in t]


def _gather(torch, dist, vals, world):
    t = torch.tensor([float(v) for v in vals], dtype=torch.float64, device="cuda")
    if dist is None:
        return [[float(x) for x in t]]
    out = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(out, t)
    return [[float(x) for x in o] for o in out]


def _repeats(K, est_step_s, min_region_s=0.5):
    """the K poses are looped R times so that the timed region is at least half a second at every N (pipeline fill and
    drain, clock sampling); ms_per_step stays region / (K * R)"""
    return max(1, int(np.ceil(min_region_s / max(K * est_step_s, 1e-6))))


def run_c4(args):
    """Time series (BASELINE configs[3]): a step = one timestep = scalar swap (streamed from pinned host memory, 1/N per rank
    over its own PCIe link, equal slabs exchanged by one ncclAllGather, double-buffered against rendering) + one sort-first
    frame of the fly-through landing in pinned host memory.  All of it behind the C ABI (icb_group_*)."""
    import torch
    import icb
    import sortfirst
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = _dist_init(torch, world, local)
    n, k, levels, W, H, tf, desc = WORKLOADS["C3"]
    t_setup = time.perf_counter()
    ctx = icb.Context(local)
    grp = sortfirst.make_group(ctx, torch, dist)
    cells = icb.gen_grid(n, k, levels) if rank == 0 else None
    nrec = int(icb.lib().icb_gen_num_records(n, k, levels))
    grp.set_grid(cells)
    ctx.set_sampler(args.sampler)
    T = args.timesteps
    # every rank keeps ONLY ITS SLAB of the T distinct timesteps in pinned host memory (the run cycles through them):
    # icb_group_stream_scalars reads nothing else of the array it is handed, so the pointer passed is slab - first_float
    a_f, n_f = grp.scalar_slab(nrec)
    r0, r1 = a_f // 32, (a_f + n_f + 31) // 32
    series = torch.empty(T, (r1 - r0) * 32, dtype=torch.float32, pin_memory=True)
    icb.gen_values(n, k, levels, 0, T, r0, r1 - r0, series.data_ptr())
    if cells is not None and r0 == 0:      # timestep 0 of the series is the grid's own scalar field, bit for bit
        assert np.array_equal(series[0].numpy().reshape(-1, 32)[:4096], cells["value"][:4096])
    steps_ptr = [series[t].data_ptr() - r0 * 32 * 4 for t in range(T)]
    del cells
    b = ctx.bounds()
    grid_build_s = ctx.stats()["setupSeconds"]
    grp.set_transfunc(icb.sharp_lut())
    grp.set_params(jitter=True)
    K, Wm = args.steps, max(args.warmup, 3)
    inflight = 2
    ctx.set_frames_in_flight(inflight)
    grp.begin(W, H, inflight)
    npx = W * H
    hosts = [ctx.host_alloc(npx * 4).view(np.uint32) for _ in range(inflight)] if rank == 0 else [None] * inflight
    t_setup = time.perf_counter() - t_setup
    R = 1
    cams = flythrough_cameras(icb, b, W, H, K + Wm)

    def step(i, pose):
        grp.stream_scalars(steps_ptr[(i + 1) % T], nrec)          # the next timestep travels while this one renders
        grp.swap_scalars()
        grp.set_camera(cams[pose])
        slot = i % inflight
        if i >= inflight:
            grp.wait(slot)                                                      # the frame this slot produced two steps ago has arrived
        grp.render(hosts[slot])

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    grp.stream_scalars(steps_ptr[0], nrec)
    tw = time.perf_counter()
    for i in range(Wm):
        step(i, i)
    grp.wait(-1)
    barrier()
    est = _reduce(torch, dist, [(time.perf_counter() - tw) / Wm])[0]
    R = _repeats(K, est)
    ctx.reset_totals()
    grp.reset_stats()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    barrier()
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    start.record()
    for j in range(K * R):
        step(Wm + j, Wm + j % K)
    grp.wait(-1)
    end.record()
    barrier()
    wall = time.perf_counter() - t0
    clk = clocks.stop() if rank == 0 else None
    dev_ms = start.elapsed_time(end)
    st = ctx.stats()
    gs = grp.stats()
    dev_ms = _reduce(torch, dist, [dev_ms])[0]
    samples = _reduce(torch, dist, [float(st["totalSteps"])], "sum")[0]
    if rank == 0:
        peak, peak_src = peaks()
        value = samples / (dev_ms * 1e-3)
        nsteps = K * R
        h2d = (nrec * 128 + world - 1) // world + 112
        ach = BYTES_PER_SAMPLE * samples / world / (dev_ms * 1e-3) / 1e9
        line = {"metric": "samples/s", "value": value, "unit": "samples/s", "n_gpus": world, "steps": K, "warmup": Wm, "repeats": R,
                "ms_per_step": dev_ms / nsteps, "frames_per_s": 1e3 * nsteps / dev_ms, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": "C4: R2B8x90 time series, one scalar swap per frame streamed from pinned host memory, camera flying "
                                       "through the atmosphere shell along the equator, 3840x2160, sharp TF",
                           "records": nrec, "width": W, "height": H, "distinct_timesteps_in_host_memory": T, "pinned_bytes_per_rank": int(series.numel()) * 4, "sampler": args.sampler,
                           "scalar_bytes_per_timestep": nrec * 128, "h2d_bytes_per_step_per_rank": h2d - 112,
                           "h2d_gb_per_s_per_rank": (h2d - 112) / (dev_ms / nsteps * 1e-3) / 1e9, "frames_in_flight": inflight,
                           "samples_per_step": samples / nsteps, "wall_s_timed_region": wall, "setup_s": t_setup, "grid_build_s": grid_build_s,
                           "parallelism": "single GPU" if world == 1 else f"sort-first, {world} GPUs (icb_group_*); scalars: 1/{world} per rank over PCIe + one ncclAllGather of equal slabs",
                           "host_ms": {"per_frame": gs["hostMsPerFrame"], "of_which_waiting_for_device": gs["waitMsPerFrame"]},
                           "l2_policy": "every step swaps 2 GB of scalars and moves the camera"},
                "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": npx * 4, "ms_per_step": dev_ms / nsteps,
                        "path": "the timed region IS end to end: scalars come from pinned host memory and every frame lands in pinned host memory"},
                "gpu_launches": (5 if world == 1 else 5) * nsteps, "clocks": clk,
                "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None, "kernel": "renderKernel",
                             "bytes_per_sample": BYTES_PER_SAMPLE, "peak_source": peak_src,
                             "note": "the step is bound by the PCIe upload of the scalars (see h2d_gb_per_s_per_rank), not by the kernel"}}
        print(json.dumps(line), flush=True)
    grp.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def profile_facts(workload, kernel_name):
    """static facts about the dominant kernel from the committed ncu summaries (profiles/render_traffic.json): DRAM bytes
    per launch, warp instructions per launch, issue-active %, active lanes per instruction -- measured under ncu at N = 1,
    so never timings, and never quoted on a multi-GPU line (a rank's launch there renders a share of the frame)"""
    p = os.path.join(ROOT, "profiles", "render_traffic.json")
    if not os.path.exists(p):
        return {}
    with open(p) as f:
        d = json.load(f)
    if d.get("workload") != workload:
        d = d.get("workloads", {}).get(workload)
        if d is None:
            return {}
    out = dict(d.get("facts", {}).get(kernel_name, {}))
    tr = d.get("kernels", {}).get(kernel_name)
    if tr is not None:
        out["traffic"] = tr
    out["capture"] = d.get("source")
    return out


def roofline_block(facts, kernel_name, samples_per_launch, kern_ms, kernel_ms_alone, peak, peak_src, sm_mhz, num_sms=148):
    """The roofline object of the bench line, frac <= 1 by construction: a MEASURED quantity of the kernel over the limit of
    the unit that bounds it.  Which unit that is comes from the ncu capture of the same workload (profile_facts):
      - DRAM bytes per launch >= the algorithmic bytes (60 B x samples): the gathers miss L2, the kernel is HBM
        bound -> bound "hbm", achieved = measured DRAM bytes / launch duration, peak = the measured HBM bandwidth (R2B10);
      - DRAM bytes far below the algorithmic bytes (coherent rays: same-prism samples and skip-ahead touch no memory, the
        rest hits L2): the kernel is bound by warp-instruction issue under divergence -> bound "issue", achieved = warp
        instructions per launch / launch duration, peak = 4 schedulers x SMs x the SM clock sampled during the run (R2B8).
    The algorithmic figure stays beside it as algorithmic_gbs -- it may exceed the HBM peak and is not a roofline fraction.
    Without a capture for this workload / GPU count nothing is quoted: achieved and frac are null."""
    alg_bytes = BYTES_PER_SAMPLE * samples_per_launch
    dur = kern_ms * 1e-3
    traffic, winst = facts.get("traffic"), facts.get("warp_instructions")
    r = {"bound": "hbm", "achieved": None, "peak": peak, "unit": "GB/s", "frac": None, "traffic": traffic, "kernel": kernel_name,
         "algorithmic_gbs": alg_bytes / dur / 1e9, "bytes_per_sample": BYTES_PER_SAMPLE, "samples_per_launch": samples_per_launch,
         "kernel_ms": kern_ms, "kernel_ms_alone": kernel_ms_alone,
         "kernel_ms_is": "timed region / launches (launches of the frames in flight overlap); kernel_ms_alone = one launch with the whole device",
         "dram_gbs": None, "dram_frac": None, "peak_source": peak_src, "capture": facts.get("capture")}
    if traffic:
        r["dram_gbs"] = traffic / dur / 1e9
        r["dram_frac"] = min(1.0, r["dram_gbs"] / peak)
        r["dram_bytes_per_sample"] = traffic / max(samples_per_launch, 1.0)
    issue_peak = 4 * num_sms * (sm_mhz or 1965.0) * 1e6 / 1e9
    if winst:
        r["issue"] = {"achieved": winst / dur / 1e9, "peak": issue_peak, "unit": "G warp-instructions/s",
                      "frac": min(1.0, winst / dur / 1e9 / issue_peak), "warp_instructions_per_launch": winst,
                      "issue_active_pct_under_ncu": facts.get("issue_active_pct"), "lanes_per_instruction": facts.get("lanes_per_instruction"),
                      "l2_hit_rate_pct": facts.get("l2_hit_rate_pct")}
    if traffic and traffic >= alg_bytes:
        r.update(achieved=r["dram_gbs"], frac=r["dram_frac"],
                 reading="HBM bound: measured DRAM bytes per launch (ncu, same workload) exceed the algorithmic 60 B x samples; "
                         "achieved = those bytes / the launch duration measured in this run")
    elif traffic and winst:
        r.update(bound="issue", achieved=r["issue"]["achieved"], peak=issue_peak, unit="G warp-instructions/s", frac=r["issue"]["frac"],
                 reading="not HBM bound: measured DRAM bytes per launch are %.1f %% of the algorithmic 60 B x samples (dram_frac = share of the "
                         "measured HBM peak), so the fraction quoted is warp instructions issued per second (ncu count of one launch / the "
                         "launch duration measured in this run) over 4 schedulers x %d SMs x the sampled SM clock; peak_hbm_gbs is the HBM "
                         "figure the dram_* fields use" % (100.0 * traffic / alg_bytes, num_sms), peak_hbm_gbs=peak)
    else:
        r["reading"] = ("no ncu capture of this kernel for this workload and GPU count under profiles/: only the algorithmic rate "
                        "(60 B x samples / time, which is not bounded by the HBM peak) is known, so no fraction is quoted")
    return r


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C3", choices=sorted(WORKLOADS) + ["C4"])
    ap.add_argument("--timesteps", type=int, default=4, help="C4: distinct timesteps held in pinned host memory (the run cycles through them)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="CPU-baseline budget (0 disables)")
    ap.add_argument("--no-jitter", action="store_true")
    ap.add_argument("--min-region", type=float, default=0.5, help="the K poses are repeated until the timed region is at least this long (s)")
    ap.add_argument("--in-flight", type=int, default=0, choices=list(range(17)),
                    help="frames in flight (marcher): k alternates k streams, frame-state slots and sets of frame buffers, so that the "
                         "tail of a frame (a few long limb rays) is filled by the head of the next; 1 renders frame after frame; 0 (default): 4")
    ap.add_argument("--integrator", default="march", choices=["march", "woodcock", "woodcock_grid", "woodcock_sphere"],
                    help="march: the ray marcher (headline); woodcock / woodcock_grid / woodcock_sphere: the reference's own ray-gens "
                         "woodcockTrackingAE / woodcockTrackingWithAccel in GRID / SPHERE accel mode (one stochastic collision per pixel per "
                         "frame; a sample = one tracking step)")
    ap.add_argument("--sampler", default="cell", choices=["cell", "prism"],
                    help="cell: value[layer] (the reference's shipped sampler, the headline); prism: barycentric x linear-in-height")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    wl = args.workload
    if args.impl == "reference":
        return run_reference(args, "C3" if wl == "C4" else wl)
    if wl == "C4":
        return run_c4(args)

    import torch
    import icb
    import sortfirst

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dist = _dist_init(torch, world, local)
    n, k, levels, W, H, tf, desc = WORKLOADS[wl]
    if args.integrator != "march" and world > 1:
        raise SystemExit("bench.py: the Woodcock ray-gens run on one GPU (sort-first tile lists apply to the marcher)")
    t_setup = time.perf_counter()
    ctx = icb.Context(local)
    grp = sortfirst.make_group(ctx, torch, dist) if world > 1 else None
    # the grid is generated and built ONCE (rank 0); the other ranks receive the finished device arrays over NVLink
    cells = icb.gen_grid(n, k, levels) if rank == 0 else None
    nrec = int(icb.lib().icb_gen_num_records(n, k, levels))
    if grp is not None:
        grp.set_grid(cells)
    else:
        ctx.set_grid(cells)
    ctx.set_sampler(args.sampler)
    ctx.set_integrator(args.integrator)
    b = ctx.bounds()
    st0 = ctx.stats()
    lut = icb.sharp_lut() if tf == "sharp" else icb.smooth_lut()
    ctx.set_transfunc(lut)
    jitter = not args.no_jitter
    ctx.set_params(jitter=jitter)                  # unitDistance 1000, dt 500
    K, Wm = max(1, args.steps), max(args.warmup, 3)      # (timing rule: at least 3 warm-up steps, whatever was asked for)
    cams = orbit_cameras(icb, b, W, H, K + Wm)
    t_setup = time.perf_counter() - t_setup
    npx = W * H
    stream = torch.cuda.current_stream().cuda_stream

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    if args.in_flight == 0:
        # measured best (profiles/r02_scale_sweep_n*.jsonl): four frames on the device up to four GPUs; at eight a rank's share of
        # a frame is so short that eight frames are needed to fill the tails of each other's longest rays
        args.in_flight = 8 if world >= 8 else 4
    inflight = args.in_flight if args.integrator == "march" else 1
    # each frame's kernel takes a share of the device's CTA slots (measured best: all four frames co-resident on one GPU,
    # two by two with sort-first, where a rank's kernels are short)
    share = int(os.environ.get("ICB_BENCH_SHARE", inflight if world == 1 else (4 if world >= 8 else min(inflight, 2))))
    ctx.set_frames_in_flight(share)

    # ---------------- device-resident arm: value ----------------
    # (always created streams, never the legacy default stream: the library's L2 access-policy window for the Woodcock
    # ray-gens' logf table is a stream attribute)
    streams = [torch.cuda.Stream() for _ in range(inflight)]
    if world == 1:
        fbs = [torch.zeros(npx, dtype=torch.int32, device="cuda") for _ in range(inflight)]
        accums = [torch.zeros(npx, 4, dtype=torch.float32, device="cuda") for _ in range(inflight)]
        depths = [torch.zeros(npx, dtype=torch.float32, device="cuda") for _ in range(inflight)]

        def frame(i, pose):
            s_ = i % inflight
            ctx.set_frame_slot(s_)
            ctx.set_camera(cams[pose])
            if args.integrator != "march":
                ctx.set_params(accum_id=i, jitter=True)      # progressive accumulation, a new RNG stream per frame
            ctx.render(fbs[s_].data_ptr(), accums[s_].data_ptr(), W, H, depth_ptr=depths[s_].data_ptr(), stream=streams[s_].cuda_stream)

        def drain():
            for st_ in streams:
                torch.cuda.current_stream().wait_stream(st_)
        launches_per_step = 3 if args.integrator == "march" else 2   # orderTiles + prepareFrame + render | prepareFrame + ray-gen
    else:
        grp.begin(W, H, inflight)

        def frame(i, pose):
            grp.set_camera(cams[pose])
            grp.render(None)                # no host synchronisation inside except for the tile costs of `inflight` frames ago

        def drain():
            grp.wait(-1)
        launches_per_step = 4               # + scatterTilesKernel on rank 0 (NCCL's own kernels not counted)

    for i in range(Wm):
        frame(i, i)
    drain()
    barrier()
    tw = time.perf_counter()                 # (a second, warm pass sizes the timed region: the first one pays first-use set-up)
    for i in range(Wm):
        frame(i, i)
    drain()
    barrier()
    est = _reduce(torch, dist, [(time.perf_counter() - tw) / Wm])[0]
    R = _repeats(K, est, args.min_region)
    if grp is not None:
        grp.reset_stats()
    ctx.reset_totals()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    barrier()
    t0 = time.perf_counter()
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for j in range(K * R):
        frame(Wm + j, Wm + j % K)
    drain()
    end.record()
    barrier()
    wall = time.perf_counter() - t0
    clk = clocks.stop() if rank == 0 else None
    dev_ms = start.elapsed_time(end)
    nsteps = K * R
    st = ctx.stats()
    my_samples, my_hits = float(st["totalSteps"]), float(st["totalHits"])
    dev_ms = _reduce(torch, dist, [dev_ms])[0]
    samples, hits = _reduce(torch, dist, [my_samples, my_hits], "sum")
    # last frame's render-kernel time and samples on every rank (load balance evidence)
    allr = _gather(torch, dist, [float(st["lastKernelMs"]), float(st["steps"])], world)
    per_rank = {"kernel_ms": [round(a[0], 3) for a in allr], "samples": [int(a[1]) for a in allr]}
    host_ms = None
    if grp is not None:       # host time per timed frame inside icb_group_render, and how much of it was waiting for the device
        gs = grp.stats()
        host_ms = {"per_frame": gs["hostMsPerFrame"], "of_which_waiting_for_device": gs["waitMsPerFrame"]}
    ms_per_step = dev_ms / nsteps
    value = samples / (dev_ms * 1e-3)

    # ---------------- latency: one frame at a time (the full device per frame) ----------------
    ctx.set_frames_in_flight(1)
    KL = min(K, 8)
    lat, kern_alone = [], []
    if world == 1:
        ctx.set_frame_slot(0)
        for i in range(KL + 1):
            ctx.set_camera(cams[Wm + i % K])
            if args.integrator != "march":
                ctx.set_params(accum_id=i, jitter=True)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(streams[0])
            ctx.render(fbs[0].data_ptr(), accums[0].data_ptr(), W, H, depth_ptr=depths[0].data_ptr(), stream=streams[0].cuda_stream)
            e1.record(streams[0])
            torch.cuda.synchronize()
            if i > 0:
                lat.append(e0.elapsed_time(e1))
                kern_alone.append(ctx.stats()["lastKernelMs"])
    else:
        grp.begin(W, H, 1)
        for i in range(KL + 1):
            grp.set_camera(cams[Wm + i % K])
            barrier()
            tl = time.perf_counter()
            grp.render(None)
            grp.wait(0)
            if i > 0:
                lat.append(1e3 * (time.perf_counter() - tl))
                kern_alone.append(ctx.stats()["lastKernelMs"])
    latency_ms = _reduce(torch, dist, [float(np.mean(lat))])[0]
    kernel_ms_alone = _reduce(torch, dist, [float(np.mean(kern_alone))])[0]
    ctx.set_frames_in_flight(share)

    # ---------------- end-to-end arm: host buffers through the C ABI ----------------
    fb_hosts = [ctx.host_alloc(npx * 4).view(np.uint32) for _ in range(inflight)] if rank == 0 else [None] * inflight
    barrier()
    ctx.set_frame_slot(0)
    if grp is not None:
        grp.begin(W, H, inflight)

    host_t = {"wait": 0.0, "enqueue": 0.0}         # where the host thread's time goes (perf_counter around the C ABI calls)

    def frame_e2e(i, pose):
        s_ = i % inflight
        ta = time.perf_counter()
        if world == 1:
            ctx.set_camera(cams[pose])
            if args.integrator != "march":
                ctx.set_params(accum_id=i, jitter=True)
            if inflight > 1:
                tb = time.perf_counter()
                ctx.frame_wait(s_)                 # the frame this slot produced `inflight` steps ago is in host memory: hand it on
                host_t["wait"] += time.perf_counter() - tb
                ctx.set_frame_slot(s_)
                ctx.render_host_async(W, H, fb_hosts[s_])
            else:
                tb = time.perf_counter()
                ctx.render_host(W, H, fb=fb_hosts[0], want_accum=False)      # synchronous: render + copy + wait in one call
                host_t["wait"] += time.perf_counter() - tb
        else:
            grp.set_camera(cams[pose])
            if i >= inflight:
                tb = time.perf_counter()
                grp.wait(s_)                       # (rank 0: that frame is in pinned host memory)
                host_t["wait"] += time.perf_counter() - tb
            grp.render(fb_hosts[s_])
        host_t["enqueue"] += time.perf_counter() - ta

    def drain_e2e():
        tb = time.perf_counter()
        if world == 1:
            for s_ in range(inflight):
                ctx.frame_wait(s_)
        else:
            grp.wait(-1)
        host_t["wait"] += time.perf_counter() - tb

    for i in range(min(Wm, 3)):
        frame_e2e(i, i)
    drain_e2e()
    barrier()
    ctx.reset_totals()
    barrier()
    host_t["wait"] = host_t["enqueue"] = 0.0
    t0 = time.perf_counter()
    for j in range(nsteps):
        frame_e2e(min(Wm, 3) + j, Wm + j % K)
    t_loop = time.perf_counter() - t0
    drain_e2e()                                    # every one of the frames has arrived in host memory
    barrier()
    e2e_s = time.perf_counter() - t0
    # the frame's trip home alone: one RGBA8 frame, device -> page-locked host memory, nothing else on the device
    d2h_ms_alone = None
    if world == 1:
        src = torch.zeros(npx, dtype=torch.int32, device="cuda")
        dst = torch.empty(npx, dtype=torch.int32, pin_memory=True)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        ev[0].record()
        for _ in range(5):
            dst.copy_(src, non_blocking=True)
        ev[1].record()
        torch.cuda.synchronize()
        d2h_ms_alone = ev[0].elapsed_time(ev[1]) / 5
        del src, dst
    e2e_breakdown = {"host_ms_per_step_in_calls": 1e3 * (host_t["enqueue"] - (host_t["wait"] - (e2e_s - t_loop))) / nsteps,
                     "host_ms_per_step_waiting_for_frames": 1e3 * host_t["wait"] / nsteps,
                     "drain_ms_after_last_enqueue": 1e3 * (e2e_s - t_loop), "d2h_ms_alone_per_frame": d2h_ms_alone,
                     "device_ms_per_step_resident": ms_per_step,
                     "reading": "ms_per_step (end to end) - device_ms_per_step_resident = what the host path adds; waiting = the host blocked on a frame "
                                "(device or copy still busy: the device is the limit), in_calls = host time inside the enqueueing C ABI calls "
                                "(host is the limit when this approaches ms_per_step); d2h alone = one frame's PCIe copy with the device idle"}
    st_e = ctx.stats()
    e2e_s = _reduce(torch, dist, [e2e_s])[0]
    e2e_samples = _reduce(torch, dist, [float(st_e["totalSteps"])], "sum")[0]
    e2e_value = e2e_samples / e2e_s

    # sort-first must reproduce the single-GPU frame bit for bit (no cross-tile blending): check it on the last pose
    frame_identical = None
    if world > 1:
        grp.set_camera(cams[-1])
        slot = grp.render(None)
        grp.wait(-1)
        barrier()
        if rank == 0:
            got = np.empty(npx, np.uint32)
            ctx.dev_download(got, grp.frame_ptr(slot), npx * 4)
            ctx.set_frame_slot(0)
            ctx.set_tiles(np.zeros(0, np.int32))
            full = torch.zeros(npx, dtype=torch.int32, device="cuda")
            acc_full = torch.zeros(npx, 4, dtype=torch.float32, device="cuda")
            ctx.render(full.data_ptr(), acc_full.data_ptr(), W, H, stream=stream)
            torch.cuda.synchronize()
            frame_identical = bool(np.array_equal(full.cpu().numpy().view(np.uint32), got))
        dist.barrier()

    if rank != 0:
        if grp is not None:
            grp.close()
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return

    peak, peak_src = peaks()
    kernel_name = {"march": "renderKernel<false," + ("true" if args.sampler == "prism" else "false") + ",false,false>", "woodcock": "woodcockKernel",
                   "woodcock_grid": "woodcockGridKernel", "woodcock_sphere": "woodcockSphereKernel"}[args.integrator]
    # a launch's duration: with frames in flight launches overlap, so a launch's share of the device is the timed region /
    # launches (kernel_ms); kernel_ms_alone is one launch with the whole device to itself (the library's own events)
    # (every rank launches one render kernel per step, so this holds at every N; the per-rank kernel times of the last frame
    # are in per_rank and are a different window: one frame's launch, overlapped by its neighbours in flight)
    kern_ms = dev_ms / nsteps
    facts = profile_facts(wl, kernel_name) if world == 1 else {}
    line = {
        "metric": "samples/s", "value": value, "unit": "samples/s", "n_gpus": world, "steps": K, "warmup": Wm, "repeats": R,
        "ms_per_step": ms_per_step, "frames_per_s": 1e3 / ms_per_step, "latency_ms": latency_ms,
        "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{wl}: {desc}", "records": nrec, "width": W, "height": H, "dt_m": 500.0,
                   "unit_distance_m": float(b["unitDistance"]), "transfer_function": tf, "jitter": jitter, "sampler": args.sampler, "integrator": args.integrator,
                   "frames_in_flight": inflight, "timed_frames": nsteps,
                   "latency": "latency_ms = one frame at a time with the whole device (frames_in_flight 1); ms_per_step = throughput with frames_in_flight frames on the device",
                   "camera": "Camera::viewAll eye orbited about +y, 30deg + 2deg per step (every step a new view), fovy %g deg (globe fills the frame height)" % ZOOM_FOVY_DEG,
                   "l2_policy": "inputs larger than L2: the grid is %.2f GB in HBM and every step is a different view" % (st0["deviceBytes"] / 1e9),
                   "parallelism": "single GPU" if world == 1 else f"sort-first, {world} GPUs, grid built once and broadcast, frame loop + NCCL tile gather behind the C ABI (icb_group_*)",
                   "samples_per_step": samples / nsteps, "hit_fraction": hits / max(samples, 1.0),
                   "slow_path_fraction": st["slowSteps"] / max(st["steps"], 1), "exact_query_fraction": st["exactQueries"] / max(st["steps"], 1),
                   "frame_identical_to_single_gpu": frame_identical, "setup_s": t_setup, "grid_build_s": st0["setupSeconds"], "wall_s_timed_region": wall},
        "per_rank": {"last_frame_kernel_ms": per_rank["kernel_ms"], "last_frame_samples": per_rank["samples"],
                     "kernel_ms_max_over_min": max(per_rank["kernel_ms"]) / max(min(per_rank["kernel_ms"]), 1e-9),
                     "samples_max_over_min": max(per_rank["samples"]) / max(min(per_rank["samples"]), 1)},
        "host_ms": host_ms,
        "e2e": {"value": e2e_value, "unit": "samples/s", "h2d_bytes_per_step": 12 * 4 + 64, "d2h_bytes_per_step": npx * 4,
                "ms_per_step": 1e3 * e2e_s / nsteps, "frames_per_s": nsteps / e2e_s, "breakdown": e2e_breakdown,
                "path": ("icb_set_camera + icb_render_host -> pinned host RGBA8 frame" if inflight == 1 else
                         f"icb_set_camera + icb_render_host_async ({inflight} frames in flight, icb_frame_wait before a slot is reused) -> pinned host RGBA8 frames") if world == 1 else
                        f"icb_group_set_camera + icb_group_render ({inflight} frames in flight): render + NCCL gather + scatter + asynchronous D2H of the full frame into pinned memory on rank 0"},
        "gpu_launches": launches_per_step * nsteps,
        "clocks": clk,
        "roofline": roofline_block(facts, kernel_name, my_samples / nsteps, kern_ms, kernel_ms_alone, peak, peak_src, (clk or {}).get("sm_mhz")),
    }
    if world == 1 and args.cpu_seconds > 0:
        pos = sample_positions(icb, b, cams[Wm:], W, H, 500.0, 4096)
        v, kind, cores, done, used, ref_ids, ref_vals = cpu_reference_rate(cells, pos, lut, args.cpu_seconds)
        # parity at the size the headline is quoted on: the ids / values the reference just located, against the CUDA path
        ids, vals = ctx.locate_host(pos[:done])
        ids_w, vals_w = ctx.locate_host(pos[:done], icb.LOCATE_WALK)
        mism = int((ids != ref_ids).sum() + (ids_w != ref_ids).sum() + (vals != ref_vals).sum() + (vals_w != ref_vals).sum())
        line["cpu_baseline"] = {"value": v, "unit": "samples/s", "cores": cores, "kind": kind,
                                "sample": f"{done} sample positions drawn from the timed frames' rays, located with the reference's brute-force "
                                          f"sampleVolume loop (deviceCode.cu:116-123) over all {nrec} records + postClassify, "
                                          f"{used:.1f} s on {cores} threads",
                                "parity_checked": {"n": int(done), "located": int((ref_ids >= 0).sum()), "mismatch": mism,
                                                   "what": "record ids and values of these positions: icb_locate (plain and walk mode) vs the reference's loop"}}
    print(json.dumps(line), flush=True)
    if grp is not None:
        grp.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
