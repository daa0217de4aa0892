"""Sort-first multi-GPU frames -- the launcher-side glue only.

The frame loop itself (partition from the all-reduced per-tile costs, tile lists, render, ncclSend/ncclRecv gather on
rank 0, ncclAllReduce of the costs, scatter, time-series slab exchange by ncclAllGather) lives behind the C ABI:
`icb_group_*` in include/icon_b200.h, implemented in csrc/icb_group.cu, bound as `icb.Group`.  What is left here is how
one-rank-per-process launchers hand the NCCL unique id around: under torchrun, through torch.distributed (plumbing).

The reference has no multi-GPU path; there is nothing to mirror except the frame itself, which must be
bit-identical to the single-GPU frame (no cross-tile blending).
"""
import icb


def make_group(ctx, torch=None, dist=None, device="cuda"):
    """this process's icb.Group: a group of one without a process group, else one rank per process with the unique id
    made on rank 0 (icb_group_unique_id = ncclGetUniqueId) and broadcast through the launcher's own process group"""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return icb.Group(ctx, 0, 1)
    rank, world = dist.get_rank(), dist.get_world_size()
    buf = torch.zeros(icb.GROUP_ID_BYTES, dtype=torch.uint8, device=device)
    if rank == 0:
        buf.copy_(torch.frombuffer(bytearray(icb.group_unique_id()), dtype=torch.uint8))
    dist.broadcast(buf, 0)
    return icb.Group(ctx, rank, world, bytes(buf.cpu().numpy().tobytes()))
