"""Multi-GPU partitioning of the hot path (one process per GPU, torch.distributed for plumbing).

* config 5 (batch of pairs): pair p -> rank p % world, no collective (`pairs_of_rank`).
* config 4 (one very large pair): rank r owns the output rows `band_rows(h, r, world)`; every rank
  holds the whole pair in HBM (the 2h-1 halo rows are simply read from it), runs the search on
  its band and the bands are gathered to rank 0 with point-to-point sends over NCCL/NVLink
  (`gather_bands`).  The same code runs over gloo on CPU tensors (tests/test_sharding_cpu.py).
"""
import torch
import torch.distributed as dist

from . import capi


def band_rows(h, rank, world):
    """Band of output rows of `rank` (sr_band_rows: equal rows; the search is data independent)."""
    return capi.band_rows(h, rank, world)


def pairs_of_rank(n_pairs, rank, world):
    return list(range(rank, n_pairs, world))


class _DevArray:
    """Exposes a raw device pointer through __cuda_array_interface__ (no copy, no ownership)."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2, "strides": None}


def result_tensor(pair, device):
    """The CV_64F result buffer of a sr_pair as a torch tensor view (rows x width, float64)."""
    return torch.as_tensor(_DevArray(pair.result_ptr(), (pair.h2, pair.w2), "<f8"), device=device)


def gather_bands(full, h, rank, world, group=None):
    """`full` is an (h, w) tensor on every rank whose rows band_rows(h, rank, world) hold this rank's
    band.  After the call rank 0's `full` holds every band.  Point-to-point, no padding."""
    if world == 1:
        return full
    ops = []
    if rank == 0:
        for r in range(1, world):
            r0, r1 = band_rows(h, r, world)
            if r1 > r0:
                ops.append(dist.P2POp(dist.irecv, full[r0:r1], r, group))
    else:
        r0, r1 = band_rows(h, rank, world)
        if r1 > r0:
            ops.append(dist.P2POp(dist.isend, full[r0:r1], 0, group))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    return full
