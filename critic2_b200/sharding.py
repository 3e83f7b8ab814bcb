"""Slab decomposition of a grid over the GPUs of one box.

The evaluation is embarrassingly parallel over points: no halo and no data-path collective.  The slowest output
index (z for CUBE's lf(ix,iy,iz), x for NCIPLOT's crho(k,j,i)) is cut into layers of `lh` planes (the kernel's
super-tile height) that are dealt round-robin to the ranks -- thin interleaved slabs instead of N thick ones, because a
molecule in its vacuum border makes thick equal slabs badly imbalanced.  The only communication is the gather of the
finished layers to the rank that writes the cube and the sum of the NCIPLOT void integrals."""
from __future__ import annotations


def rank_layers(nslow: int, lh: int, rank: int, world: int):
    """[(c0, c1)] plane ranges owned by `rank` (layer index mod world == rank)."""
    nl = (nslow + lh - 1) // lh
    return [(L * lh, min(nslow, (L + 1) * lh)) for L in range(rank, nl, world)]


def rank_rows(nslow: int, nmid: int, lh: int, rank: int, world: int):
    """The finer decomposition of c2gpu_eval_grid_dev_part_rows: [(c0, c1, b0, b1)] blocks owned by `rank`, one per row
    of super-tiles: unit = layer * ceil(nmid / lh) + row, unit mod world == rank (common.cuh part_item)."""
    nl, nr = (nslow + lh - 1) // lh, (nmid + lh - 1) // lh
    out = []
    for u in range(rank, nl * nr, world):
        L, r = divmod(u, nr)
        out.append((L * lh, min(nslow, (L + 1) * lh), r * lh, min(nmid, (r + 1) * lh)))
    return out


def rank_points(nslow: int, lh: int, rank: int, world: int, plane_points: int) -> int:
    return sum(c1 - c0 for c0, c1 in rank_layers(nslow, lh, rank, world)) * plane_points


def gather_layers(full, nslow: int, lh: int, rank: int, world: int, dist, dst: int = 0):
    """Gather the interleaved layers held in each rank's full-size array `full` (shape (nslow, plane) or flat) on rank
    `dst`.  Layers are disjoint and every rank's array is zero outside its own layers, so a sum-reduce to `dst` is an
    exact gather (x + 0 = x in IEEE arithmetic); over NVLink/NVSwitch NCCL reduces it in one pass."""
    dist.reduce(full, dst=dst, op=dist.ReduceOp.SUM)
    return full


def zero_foreign_layers(full, nslow: int, lh: int, rank: int, world: int):
    """Zero every plane of `full` (viewed as (nslow, -1)) that this rank does not own."""
    v = full.view(nslow, -1)
    own = [False] * nslow
    for c0, c1 in rank_layers(nslow, lh, rank, world):
        for c in range(c0, c1):
            own[c] = True
    for c in range(nslow):
        if not own[c]:
            v[c].zero_()
    return full


def reduce_void_sums(raw_sums, dist):
    """NCIPLOT void integrals sumq, sumqp, sumv (nci@proc.f90:573-582, 600-603) summed over the ranks."""
    dist.all_reduce(raw_sums, op=dist.ReduceOp.SUM)
    return raw_sums
