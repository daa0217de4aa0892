"""Sort-first parameter sweep on N GPUs (torchrun): the C3 grid is built once and broadcast, then every configuration
(frames in flight, CTA share per frame, split-ray mode / thresholds) renders the bench's orbit poses through icb_group_*.
One JSON line per configuration on rank 0.  Usage:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29517 tools/sweep_scale.py [--workload C3]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "icon-ray-tracing_b200"))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C3")
    ap.add_argument("--frames", type=int, default=20)
    ap.add_argument("--configs", default="")
    args = ap.parse_args()
    import torch
    import icb
    import sortfirst
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = bench._dist_init(torch, world, local)
    n, k, levels, W, H, tf, desc = bench.WORKLOADS[args.workload]
    ctx = icb.Context(local)
    grp = sortfirst.make_group(ctx, torch, dist)
    t0 = time.perf_counter()
    cells = icb.gen_grid(n, k, levels) if rank == 0 else None
    t_gen = time.perf_counter() - t0
    t0 = time.perf_counter()
    grp.set_grid(cells)
    t_grid = time.perf_counter() - t0
    del cells
    b = ctx.bounds()
    ctx.set_transfunc(icb.sharp_lut() if tf == "sharp" else icb.smooth_lut())
    ctx.set_params(jitter=True)
    K = args.frames
    cams = bench.orbit_cameras(icb, b, W, H, K + 4)
    if rank == 0:
        print(json.dumps({"world": world, "gen_s": t_gen, "group_set_grid_s": t_grid, "grid_build_s": ctx.stats()["setupSeconds"]}), flush=True)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # (in flight, CTA share, split mode, max heavy tiles, min tile cost log2, chunk log2)
    configs = [(4, 2, 0, 384, 16, 10), (4, 2, 1, 384, 16, 10), (4, 4, 1, 384, 16, 10), (8, 4, 1, 384, 16, 10), (8, 4, 0, 384, 16, 10),
               (8, 2, 1, 384, 16, 10), (4, 2, 1, 384, 15, 10), (4, 2, 1, 1024, 14, 10), (4, 2, 1, 384, 16, 9), (2, 2, 1, 384, 16, 10),
               (1, 1, 1, 384, 16, 10), (1, 1, 0, 384, 16, 10)]
    if args.configs:
        configs = [tuple(int(v) for v in c.split(",")) for c in args.configs.split(";")]
    for (inflight, share, mode, heavy, costlog2, chunklog2) in configs:
        ctx.set_ray_split(chunklog2, heavy, costlog2, mode)
        ctx.set_frames_in_flight(share)
        grp.begin(W, H, inflight)
        for i in range(4):
            grp.set_camera(cams[i]); grp.render(None)
        grp.wait(-1)
        barrier()
        tw = time.perf_counter()
        for i in range(4):
            grp.set_camera(cams[i]); grp.render(None)
        grp.wait(-1)
        barrier()
        est = bench._reduce(torch, dist, [(time.perf_counter() - tw) / 4])[0]
        R = bench._repeats(K, est)
        grp.reset_stats()
        ctx.reset_totals()
        barrier()
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        start.record()
        for j in range(K * R):
            grp.set_camera(cams[4 + j % K]); grp.render(None)
        grp.wait(-1)
        end.record()
        barrier()
        ms = bench._reduce(torch, dist, [start.elapsed_time(end)])[0]
        st = ctx.stats()
        samples = bench._reduce(torch, dist, [float(st["totalSteps"])], "sum")[0]
        allr = bench._gather(torch, dist, [float(st["lastKernelMs"]), float(st["steps"])], world)
        gs = grp.stats()
        if rank == 0:
            print(json.dumps({"in_flight": inflight, "share": share, "split_mode": mode, "max_heavy": heavy, "cost_log2": costlog2, "chunk_log2": chunklog2,
                              "ms_per_frame": ms / (K * R), "G_samples_per_s": samples / ms / 1e6, "frames": K * R,
                              "host_ms": gs["hostMsPerFrame"], "wait_ms": gs["waitMsPerFrame"],
                              "kernel_ms": [round(a[0], 2) for a in allr], "samples_M": [round(a[1] / 1e6, 1) for a in allr]}), flush=True)
    grp.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
