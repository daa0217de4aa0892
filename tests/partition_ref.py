"""TEST INFRASTRUCTURE: numpy statement of the sort-first partition (Z-ordered tiles, cost-sorted blocks dealt to the ranks
in snake order) that the library computes in C++ (csrc/icb_group.cu dealTiles, exported as icb_partition_tiles).  The
C++ one is the product; this one only checks it."""
import numpy as np

TILE = 16


def morton_order(tiles_x, tiles_y):
    """tile ids (row-major id = ty*tiles_x + tx) sorted along a Z-curve: neighbouring tiles in the list are
    neighbours on screen, so a contiguous slice of the list is a compact screen region (rays that share cells)."""
    ty, tx = np.divmod(np.arange(tiles_x * tiles_y, dtype=np.int64), tiles_x)

    def spread(v):
        v = v & 0xFFFF
        v = (v | (v << 8)) & 0x00FF00FF
        v = (v | (v << 4)) & 0x0F0F0F0F
        v = (v | (v << 2)) & 0x33333333
        v = (v | (v << 1)) & 0x55555555
        return v

    key = spread(tx) | (spread(ty) << 1)
    return np.argsort(key, kind="stable").astype(np.int32)


def assign_tiles(order, costs, n_ranks):
    """Cut the Z-ordered tile list into n_ranks contiguous slices of (nearly) equal cost.

    costs: per-tile sample counts of the previous frame indexed by tile id, or None for the first frame
    (equal-count slices).  Deterministic: every rank computes the same partition from the same all-reduced
    costs.  Returns a list of int32 arrays (tile ids per rank); every tile appears exactly once.
    """
    n = len(order)
    if costs is None:
        w = np.ones(n, np.float64)
    else:
        w = np.asarray(costs, np.float64)[order] + 1.0      # +1: a tile is never free (ray set-up, empty pixels)
    cum = np.cumsum(w)
    total = cum[-1]
    cuts = np.searchsorted(cum, total * np.arange(1, n_ranks) / n_ranks, side="left")
    cuts = np.concatenate([[0], np.minimum(cuts + 1, n), [n]])
    cuts = np.maximum.accumulate(cuts)
    return [order[cuts[r]:cuts[r + 1]].copy() for r in range(n_ranks)]


def assign_tiles_dealt(order, costs, n_ranks, block=16):
    """Deal blocks of `block` Z-consecutive tiles (16 = a 64x64-pixel square) to the ranks: blocks sorted by cost, heaviest
    first, dealt in snake order (0..n-1, n-1..0, ...), so that every rank receives the same mix of heavy (limb), medium and
    empty screen regions.  Unlike contiguous slices this also spreads the frame's longest rays -- serial chains of
    dependent fetches that bound a rank's frame time from below -- over all ranks.  Deterministic (stable sorts); each
    rank's tiles stay in Z order.  Returns a list of int32 arrays; every tile appears exactly once."""
    n = len(order)
    nb = (n + block - 1) // block
    if costs is None:
        w = np.ones(n, np.float64)
    else:
        w = np.asarray(costs, np.float64)[order] + 1.0
    pad = nb * block - n
    bw = np.concatenate([w, np.zeros(pad)]).reshape(nb, block).sum(1)
    rank_of_sorted = np.arange(nb) % (2 * n_ranks)
    rank_of_sorted = np.where(rank_of_sorted < n_ranks, rank_of_sorted, 2 * n_ranks - 1 - rank_of_sorted)
    by_cost = np.argsort(-bw, kind="stable")
    rank_of_block = np.empty(nb, np.int64)
    rank_of_block[by_cost] = rank_of_sorted
    # rank-major tile list, Z order kept inside a rank: sort the BLOCKS by rank (stable), then expand to tiles
    blocks = np.argsort(rank_of_block.astype(np.int16), kind="stable")
    idx = (blocks[:, None] * block + np.arange(block)[None, :]).reshape(-1)
    ranks = np.repeat(rank_of_block[blocks], block)
    keep = idx < n                                               # (the last block may be ragged)
    tiles, ranks = order[idx[keep]], ranks[keep]
    cuts = np.concatenate([[0], np.cumsum(np.bincount(ranks, minlength=n_ranks))])
    return [tiles[cuts[r]:cuts[r + 1]] for r in range(n_ranks)]


