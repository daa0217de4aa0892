#!/usr/bin/env python
"""bench.py -- samples/s (and frames/s) of the icon_rt hot path on B200.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (C ABI, libicon_b200.so)
    python bench.py --impl reference --gpus N --steps K ...   # the reference's own CPU code (oracle/_ref) on host cores

A step = one frame: every pixel's ray is marched through the ICON prism shells (locate -> sample -> transfer
function -> front-to-back compositing).  Workload (BASELINE.json configs[2], the one `metric` is quoted on):
synthetic R2B8 x 90 levels (5 242 880 columns = 15 728 640 `.ic` records), 3840x2160, sharp/high-opacity
transfer function, orbit camera (a new pose every step).  `value` = samples/s with everything resident in HBM;
`e2e` = the same through icb_render_host with HOST buffers (camera/params up, RGBA8 frame down into pinned memory).
N > 1: sort-first tiles, grid replicated per GPU, tiles gathered on rank 0 with NCCL (strong scaling: one frame).
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


def make_grid(icb, n, k, levels, rank, world, dist):
    """rank 0 generates the `.ic` records once into /dev/shm; the other ranks of the box map the same file"""
    if world == 1:
        return icb.gen_grid(n, k, levels)
    path = f"/dev/shm/icb_bench_r{n}b{k}l{levels}_{os.environ.get('MASTER_PORT', '0')}.ic"
    if rank == 0:
        icb.gen_grid(n, k, levels).tofile(path)
    dist.barrier()
    cells = np.fromfile(path, icb.CELL_DTYPE)
    dist.barrier()
    if rank == 0:
        os.unlink(path)
    return cells


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
    while done < len(pos) and t_used < budget_s:
        p = pos[done:done + chunk]
        t0 = time.perf_counter()
        ids, vals = oracle.locate(cells, p)
        oracle.post_classify(lut, 0.0, 1.0, 1.0, vals)
        dt = time.perf_counter() - t0
        t_used += dt
        done += len(p)
        if dt < 0.5:
            chunk *= 2
    return done / max(t_used, 1e-9), kind, cores, done, t_used


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


def run_c4(args):
    """Time series: a step = one timestep = scalar swap (streamed from pinned host memory, 1/N per rank over PCIe, slabs
    exchanged over NVLink, double-buffered against rendering) + one frame of the fly-through."""
    import torch
    import icb
    import sortfirst
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n, k, levels, W, H, tf, desc = WORKLOADS["C3"]
    cells = make_grid(icb, n, k, levels, rank, world, dist)
    ctx = icb.Context(local)
    ctx.set_grid(cells)
    ctx.set_sampler(args.sampler)
    nrec = len(cells)
    T = args.timesteps
    steps_host = []
    for s_ in range(T):                      # distinct timesteps held in pinned host memory (the run cycles through them)
        v = torch.empty(nrec * 32, dtype=torch.float32, pin_memory=True)
        v.copy_(torch.from_numpy(np.ascontiguousarray((cells if s_ == 0 else icb.gen_grid(n, k, levels, timestep=s_))["value"]).reshape(-1)))
        steps_host.append(v)
    del cells
    b = ctx.bounds()
    ctx.set_transfunc(icb.sharp_lut())
    ctx.set_params(jitter=True)
    K, Wm = args.steps, max(args.warmup, 3)
    cams = flythrough_cameras(icb, b, W, H, K + Wm)
    npx = W * H
    inflight = 2
    ctx.set_frames_in_flight(inflight)
    streams = [torch.cuda.Stream() for _ in range(inflight)]
    stream_ts = sortfirst.ScalarStream(ctx, nrec, rank, world, torch, dist)
    if world == 1:
        sf = None
    else:
        sf = sortfirst.SortFirstFrame(ctx, W, H, rank, world, torch, dist, "cuda")
        sf.in_flight = inflight
        host_fbs = [torch.zeros(npx, dtype=torch.int32, pin_memory=True) for _ in range(inflight)] if rank == 0 else None
    fb_ev = [None] * inflight

    def step(i):
        slot = i % inflight
        if i + 1 < K + Wm:
            stream_ts.upload(steps_host[(i + 1) % T], (i + 1) & 1)          # next timestep travels while this one renders
        with torch.cuda.stream(streams[slot]):
            stream_ts.apply(i & 1, streams)
            ctx.set_camera(cams[i])
            if world == 1:
                if fb_ev[slot] is not None:
                    fb_ev[slot].synchronize()            # the frame this slot produced two steps ago has arrived: hand it on
                ctx.set_frame_slot(slot)
                # (render on the torch stream so that the scalar swap's stream ordering applies, then bring the frame down)
                ctx.render(fbs[slot].data_ptr(), accums[slot].data_ptr(), W, H, stream=streams[slot].cuda_stream)
                hosts_t[slot].copy_(fbs[slot], non_blocking=True)
                fb_ev[slot] = torch.cuda.Event()
                fb_ev[slot].record()
            else:
                if rank == 0 and fb_ev[slot] is not None:
                    fb_ev[slot].synchronize()
                sf.render_async()
                if rank == 0:
                    host_fbs[slot].copy_(sf.fb, non_blocking=True)
                    fb_ev[slot] = torch.cuda.Event()
                    fb_ev[slot].record()

    if world == 1:
        fbs = [torch.zeros(npx, dtype=torch.int32, device="cuda") for _ in range(inflight)]
        accums = [torch.zeros(npx, 4, dtype=torch.float32, device="cuda") for _ in range(inflight)]
        hosts_t = [torch.zeros(npx, dtype=torch.int32, pin_memory=True) for _ in range(inflight)]

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    stream_ts.upload(steps_host[0], 0)
    for i in range(Wm):
        step(i)
    barrier()
    ctx.reset_totals()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    barrier()
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    start.record()
    for i in range(K):
        step(Wm + i)
    for st_ in streams + [stream_ts.copy_stream]:
        torch.cuda.current_stream().wait_stream(st_)
    end.record()
    barrier()
    wall = time.perf_counter() - t0
    clk = clocks.stop() if rank == 0 else None
    dev_ms = start.elapsed_time(end)
    st = ctx.stats()
    red = torch.tensor([dev_ms, float(st["totalSteps"])], dtype=torch.float64, device="cuda")
    if dist is not None:
        mx = red.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = red.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        dev_ms, samples = float(mx[0]), float(sm[1])
    else:
        samples = float(st["totalSteps"])
    if rank == 0:
        peak, peak_src = peaks()
        value = samples / (dev_ms * 1e-3)
        h2d = stream_ts.bytes_per_rank + 112
        line = {"metric": "samples/s", "value": value, "unit": "samples/s", "n_gpus": world, "steps": K, "warmup": Wm,
                "ms_per_step": dev_ms / K, "frames_per_s": 1e3 * K / dev_ms, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": "C4: R2B8x90 time series, one scalar swap per frame streamed from pinned host memory, camera flying "
                                       "through the atmosphere shell along the equator, 3840x2160, sharp TF",
                           "records": nrec, "width": W, "height": H, "distinct_timesteps_in_host_memory": T, "sampler": args.sampler,
                           "scalar_bytes_per_timestep": nrec * 128, "h2d_bytes_per_step_per_rank": stream_ts.bytes_per_rank,
                           "h2d_gb_per_s_per_rank": stream_ts.bytes_per_rank / (dev_ms / K * 1e-3) / 1e9, "frames_in_flight": inflight,
                           "samples_per_step": samples / K, "wall_s_timed_region": wall,
                           "parallelism": "single GPU" if world == 1 else f"sort-first, {world} GPUs; scalars: 1/{world} per rank over PCIe + NCCL slab exchange",
                           "l2_policy": "every step swaps 2 GB of scalars and moves the camera"},
                "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": npx * 4, "ms_per_step": dev_ms / K,
                        "path": "the timed region IS end to end: scalars come from pinned host memory and every frame lands in pinned host memory"},
                "gpu_launches": (4 if world == 1 else 5) * K, "clocks": clk,
                "roofline": {"bound": "hbm", "achieved": BYTES_PER_SAMPLE * samples / world / (dev_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                             "frac": BYTES_PER_SAMPLE * samples / world / (dev_ms * 1e-3) / 1e9 / peak, "traffic": None, "kernel": "renderKernel",
                             "bytes_per_sample": BYTES_PER_SAMPLE, "peak_source": peak_src,
                             "note": "the step is bound by the PCIe upload of the scalars (see h2d_gb_per_s_per_rank), not by the kernel"}}
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


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
    ap.add_argument("--in-flight", type=int, default=0, choices=list(range(9)),
                    help="frames in flight (marcher): 2 alternates two streams, two frame-state slots and two sets of frame buffers, so "
                         "that the tail of a frame (a few long limb rays) is filled by the head of the next; 1 renders frame after frame; "
                         "0 (default): 4")
    ap.add_argument("--integrator", default="march", choices=["march", "woodcock", "woodcock_grid"],
                    help="march: the ray marcher (headline); woodcock / woodcock_grid: the reference's own ray-gens woodcockTrackingAE / "
                         "woodcockTrackingWithAccel (one stochastic collision per pixel per frame; a sample = one tracking step)")
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
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n, k, levels, W, H, tf, desc = WORKLOADS[wl]
    t_setup = time.perf_counter()
    cells = make_grid(icb, n, k, levels, rank, world, dist)
    ctx = icb.Context(local)
    ctx.set_grid(cells)
    ctx.set_sampler(args.sampler)
    ctx.set_integrator(args.integrator)
    if args.integrator != "march" and world > 1:
        raise SystemExit("bench.py: the Woodcock ray-gens run on one GPU (sort-first tile lists apply to the marcher)")
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

    # ---------------- device-resident arm: value ----------------
    if args.in_flight == 0:
        args.in_flight = 4
    inflight = args.in_flight if args.integrator == "march" else 1
    # (always a created stream, never the legacy default stream: the library's L2 access-policy window for the Woodcock
    # ray-gens' logf table is a stream attribute)
    streams = [torch.cuda.Stream() for _ in range(inflight)]
    # each frame's kernel takes a share of the device's CTA slots (measured best: all four frames co-resident on one GPU,
    # two by two with sort-first, where a rank's kernels are short)
    ctx.set_frames_in_flight(int(os.environ.get("ICB_BENCH_SHARE", inflight if world == 1 else min(inflight, 2))))
    if world == 1:
        fbs = [torch.zeros(npx, dtype=torch.int32, device="cuda") for _ in range(inflight)]
        accums = [torch.zeros(npx, 4, dtype=torch.float32, device="cuda") for _ in range(inflight)]
        depths = [torch.zeros(npx, dtype=torch.float32, device="cuda") for _ in range(inflight)]
        fb, accum = fbs[0], accums[0]

        def frame(i):
            b = i % inflight
            ctx.set_frame_slot(b)
            ctx.set_camera(cams[i])
            if args.integrator != "march":
                ctx.set_params(accum_id=i, jitter=True)      # progressive accumulation, a new RNG stream per frame
            ctx.render(fbs[b].data_ptr(), accums[b].data_ptr(), W, H, depth_ptr=depths[b].data_ptr(), stream=streams[b].cuda_stream)
        launches_per_step = 3 if args.integrator == "march" else 2   # orderTiles + prepareFrame + render | prepareFrame + woodcock
        sf = None
    else:
        sf = sortfirst.SortFirstFrame(ctx, W, H, rank, world, torch, dist, "cuda")
        sf.in_flight = inflight
        sf.rebalance_every = int(os.environ.get("ICB_REBALANCE_EVERY", "1"))

        def frame(i):
            ctx.set_camera(cams[i])
            with torch.cuda.stream(streams[getattr(sf, "frame_no", 0) % inflight]):
                sf.render_async()          # no host synchronisation inside (sortfirst.py)
        launches_per_step = 4          # + scatterTilesKernel on rank 0

    for i in range(Wm):
        frame(i)
    barrier()
    if sf is not None:
        sf.t_host = sf.t_wait = 0.0
    ctx.reset_totals()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    kern_ev = []
    barrier()
    t0 = time.perf_counter()
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for i in range(K):
        if inflight == 1:
            ev[i][0].record(streams[0])
        frame(Wm + i)
        if inflight == 1:
            ev[i][1].record(streams[0])
    for st_ in streams:
        torch.cuda.current_stream().wait_stream(st_)
    end.record()
    barrier()
    wall = time.perf_counter() - t0
    clk = clocks.stop() if rank == 0 else None
    dev_ms = start.elapsed_time(end)
    st = ctx.stats()
    my_samples, my_hits = float(st["totalSteps"]), float(st["totalHits"])
    red = torch.tensor([dev_ms, my_samples, my_hits], dtype=torch.float64, device="cuda")
    if dist is not None:
        mx = red.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = red.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        dev_ms, samples, hits = float(mx[0]), float(sm[1]), float(sm[2])
    else:
        samples, hits = my_samples, my_hits
    per_rank = None
    if dist is not None:                    # last frame's render-kernel time and samples on every rank (load balance evidence)
        mine_t = torch.tensor([float(st["lastKernelMs"]), float(st["steps"])], dtype=torch.float64, device="cuda")
        allr = [torch.zeros_like(mine_t) for _ in range(world)]
        dist.all_gather(allr, mine_t)
        per_rank = {"kernel_ms": [round(float(a[0]), 3) for a in allr], "samples": [int(a[1]) for a in allr]}
    host_ms = None
    if sf is not None:        # rank 0's host time per timed frame, and how much of it was waiting for the device (the rest is host work)
        host_ms = {"per_frame": 1e3 * sf.t_host / K, "of_which_waiting_for_device": 1e3 * sf.t_wait / K}
    ms_per_step = dev_ms / K
    value = samples / (dev_ms * 1e-3)
    # dominant kernel: the render kernel; at N=1 a step is prepare (a few us) + render, so the per-step events
    # bracket the render launch itself
    # (two frames in flight: launches overlap, so a launch's share of the device is the timed region / launches)
    kern_ms = float(np.mean([a.elapsed_time(bb) for a, bb in ev])) if inflight == 1 else dev_ms / K
    lib_kernel_ms = st["lastKernelMs"] if inflight == 1 else dev_ms / K   # the library's own events around the last launch

    # ---------------- end-to-end arm: host buffers through the C ABI ----------------
    fb_host = ctx.host_alloc(npx * 4).view(np.uint32)
    fb_hosts = [fb_host] + [ctx.host_alloc(npx * 4).view(np.uint32) for _ in range(inflight - 1)]
    barrier()
    ctx.reset_totals()

    ctx.set_frame_slot(0)
    if world > 1:
        host_fbs = [torch.zeros(npx, dtype=torch.int32, pin_memory=True) for _ in range(inflight)] if rank == 0 else None
        e2e_ev = [None] * inflight

    def frame_e2e(i):
        ctx.set_camera(cams[i])
        if args.integrator != "march":
            ctx.set_params(accum_id=i, jitter=True)
        if world == 1 and inflight > 1:
            b = i % inflight
            ctx.frame_wait(b)                  # the frame this slot produced two steps ago is in host memory: hand it on
            ctx.set_frame_slot(b)
            ctx.render_host_async(W, H, fb_hosts[b])
        elif world == 1:
            ctx.render_host(W, H, fb=fb_host, want_accum=False)
        else:
            slot = sf.frame_no % inflight
            if rank == 0 and e2e_ev[slot] is not None:
                e2e_ev[slot].synchronize()     # the frame this slot produced `inflight` steps ago is in host memory: hand it on
            with torch.cuda.stream(streams[slot]):
                sf.render_async()
                if rank == 0:
                    host_fbs[slot].copy_(sf.fb, non_blocking=True)      # full RGBA8 frame -> pinned host memory, behind the scatter
                    e2e_ev[slot] = torch.cuda.Event()
                    e2e_ev[slot].record()
    for i in range(min(Wm, 3)):
        frame_e2e(i)
    barrier()
    ctx.reset_totals()
    barrier()
    t0 = time.perf_counter()
    for i in range(K):
        frame_e2e(Wm + i)
    if world == 1 and inflight > 1:
        for slot in range(inflight):
            ctx.frame_wait(slot)               # every one of the K frames has arrived in host memory
    barrier()
    e2e_s = time.perf_counter() - t0
    st_e = ctx.stats()
    red = torch.tensor([e2e_s, float(st_e["totalSteps"])], dtype=torch.float64, device="cuda")
    if dist is not None:
        mx = red.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = red.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        e2e_s, e2e_samples = float(mx[0]), float(sm[1])
    else:
        e2e_samples = float(st_e["totalSteps"])
    e2e_value = e2e_samples / e2e_s

    # sort-first must reproduce the single-GPU frame bit for bit (no cross-tile blending): check it on the last pose
    frame_identical = None
    if world > 1:
        ctx.set_camera(cams[-1])
        sf.render_async()
        torch.cuda.synchronize()
        if rank == 0:
            got = sf.fb.clone()
            ctx.set_frame_slot(0)
            ctx.set_tiles(np.zeros(0, np.int32))
            full = torch.zeros(npx, dtype=torch.int32, device="cuda")
            acc_full = torch.zeros(npx, 4, dtype=torch.float32, device="cuda")
            ctx.render(full.data_ptr(), acc_full.data_ptr(), W, H, stream=stream)
            torch.cuda.synchronize()
            frame_identical = bool(torch.equal(full, got))
        dist.barrier()

    if rank != 0:
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return

    peak, peak_src = peaks()
    kernel_name = {"march": "renderKernel<false," + ("true>" if args.sampler == "prism" else "false>"), "woodcock": "woodcockKernel",
                   "woodcock_grid": "woodcockGridKernel"}[args.integrator]
    samples_per_launch = samples / K / world if world > 1 else samples / K
    achieved = BYTES_PER_SAMPLE * (my_samples / K) / (kern_ms * 1e-3) / 1e9 if world == 1 else \
        BYTES_PER_SAMPLE * (my_samples / K) / (lib_kernel_ms * 1e-3) / 1e9
    line = {
        "metric": "samples/s", "value": value, "unit": "samples/s", "n_gpus": world, "steps": K, "warmup": Wm,
        "ms_per_step": ms_per_step, "frames_per_s": 1e3 / ms_per_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{wl}: {desc}", "records": int(len(cells)), "width": W, "height": H, "dt_m": 500.0,
                   "unit_distance_m": float(b["unitDistance"]), "transfer_function": tf, "jitter": jitter, "sampler": args.sampler, "integrator": args.integrator, "frames_in_flight": inflight,
                   "camera": "Camera::viewAll eye orbited about +y, 30deg + 2deg per step (every step a new view), fovy %g deg (globe fills the frame height)" % ZOOM_FOVY_DEG,
                   "l2_policy": "inputs larger than L2: the grid is %.2f GB in HBM and every step is a different view" % (st0["deviceBytes"] / 1e9),
                   "parallelism": "single GPU" if world == 1 else f"sort-first, {world} GPUs, grid replicated, NCCL tile gather",
                   "samples_per_step": samples / K, "hit_fraction": hits / max(samples, 1.0),
                   "slow_path_fraction": st["slowSteps"] / max(st["steps"], 1), "exact_query_fraction": st["exactQueries"] / max(st["steps"], 1),
                   "frame_identical_to_single_gpu": frame_identical, "last_frame_per_rank": per_rank, "host_ms": host_ms, "setup_s": t_setup, "grid_build_s": st0["setupSeconds"], "wall_s_timed_region": wall},
        "e2e": {"value": e2e_value, "unit": "samples/s", "h2d_bytes_per_step": 12 * 4 + 64, "d2h_bytes_per_step": npx * 4,
                "ms_per_step": 1e3 * e2e_s / K, "frames_per_s": K / e2e_s,
                "path": ("icb_set_camera + icb_render_host -> pinned host RGBA8 frame" if inflight == 1 else
                         "icb_set_camera + icb_render_host_async (two frames in flight, icb_frame_wait before a slot is reused) -> pinned host RGBA8 frames") if world == 1 else
                        "sort-first frames in flight: render + NCCL gather + scatter + asynchronous D2H of the full frame into pinned memory on rank 0"},
        "gpu_launches": launches_per_step * K,
        "clocks": clk,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic_per_launch(wl, kernel_name), "kernel": kernel_name, "bytes_per_sample": BYTES_PER_SAMPLE,
                     "samples_per_launch": my_samples / K, "kernel_ms": kern_ms if world == 1 else lib_kernel_ms, "peak_source": peak_src},
    }
    if world == 1 and args.cpu_seconds > 0:
        pos = sample_positions(icb, b, cams[Wm:], W, H, 500.0, 4096)
        v, kind, cores, done, used = cpu_reference_rate(cells, pos, lut, args.cpu_seconds)
        line["cpu_baseline"] = {"value": v, "unit": "samples/s", "cores": cores, "kind": kind,
                                "sample": f"{done} sample positions drawn from the timed frames' rays, located with the reference's brute-force "
                                          f"sampleVolume loop (deviceCode.cu:116-123) over all {len(cells)} records + postClassify, "
                                          f"{used:.1f} s on {cores} threads"}
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
