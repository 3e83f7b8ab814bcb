"""Slab decomposition of a grid over the GPUs of one box.

The evaluation is embarrassingly parallel over points: no halo and no data-path collective.  The slowest output
index (z for CUBE's lf(ix,iy,iz), x for NCIPLOT's crho(k,j,i)) is cut into layers of `lh` planes (the kernel's
super-tile height) that are dealt round-robin to the ranks -- thin interleaved slabs instead of N thick ones, because a
molecule in its vacuum border makes thick equal slabs badly imbalanced.  The only communication is the gather of the
finished layers to the rank that writes the cube (every rank copies its blocks into one shared page-locked host buffer:
SharedPinned + c2gpu_gather_part_host, no inter-GPU traffic) and the sum of the NCIPLOT void integrals."""
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


class SharedPinned:
    """Page-locked host buffers shared by the ranks of one box: the destination of c2gpu_gather_part_host in the
    one-process-per-GPU model ("each GPU D2H-copies straight into its slice of the pinned host array").

    Rank 0 creates an anonymous memory file (memfd: not limited by the size of /dev/shm), the other ranks map the same
    pages through /proc/<pid>/fd/<fd>, and every rank registers its mapping with cudaHostRegister so its own DMA engine
    can write there.  Afterwards the writer rank (0) sees the complete arrays without any inter-GPU traffic."""

    def __init__(self, rank: int, world: int, dist=None, register: bool = True):
        self.rank, self.world, self.dist, self.register = rank, world, dist, register
        self.bufs = []

    def get(self, nbytes: int, tag: str = ""):
        import mmap
        import os

        import numpy as np
        import torch
        if self.rank == 0:
            fd = os.memfd_create("c2gpu_" + tag)
            os.ftruncate(fd, nbytes)
            info = [os.getpid(), fd]
        else:
            fd, info = -1, [0, 0]
        if self.world > 1:
            self.dist.broadcast_object_list(info, src=0)
            if self.rank != 0:
                fd = os.open(f"/proc/{info[0]}/fd/{info[1]}", os.O_RDWR)
        mm = mmap.mmap(fd, nbytes)
        arr = np.frombuffer(mm, dtype=np.float64)
        if self.register:
            rc = torch.cuda.cudart().cudaHostRegister(arr.ctypes.data, nbytes, 0)
            if int(rc) != 0:
                raise RuntimeError(f"cudaHostRegister of the shared host buffer failed: {rc}")
        if self.world > 1:
            self.dist.barrier()          # every rank holds its mapping before rank 0 may drop the file descriptor
        self.bufs.append((arr, mm, fd))
        return arr

    def close(self):
        import os

        import torch
        for arr, mm, fd in self.bufs:
            if self.register:
                torch.cuda.cudart().cudaHostUnregister(arr.ctypes.data)
            try:
                os.close(fd)
            except OSError:
                pass
        self.bufs = []          # the mappings go away with the last numpy view


def copy_owned_blocks(dst, src, nslow: int, nmid: int, lh: int, rank: int, world: int, rows: bool = False):
    """Host mirror of c2gpu_gather_part_host's block list (csrc part_jobs): copy the blocks `rank` owns from its full-size
    array `src` into the shared full-size array `dst` (both viewed as (nslow, nmid, nfast))."""
    d, s_ = dst.reshape(nslow, nmid, -1), src.reshape(nslow, nmid, -1)
    if rows:
        for c0, c1, b0, b1 in rank_rows(nslow, nmid, lh, rank, world):
            d[c0:c1, b0:b1] = s_[c0:c1, b0:b1]
    else:
        for c0, c1 in rank_layers(nslow, lh, rank, world):
            d[c0:c1] = s_[c0:c1]
    return dst


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
