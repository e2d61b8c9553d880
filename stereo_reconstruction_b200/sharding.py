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


def result16_tensor(pair, device):
    """The int16 wire-format buffer of a sr_pair (sr_search_result16_dev) as a torch tensor view."""
    return torch.as_tensor(_DevArray(pair.result16_ptr(), (pair.h2, pair.w2), "<i2"), device=device)


def gather_bands16(pair, full16, h, rank, world, group=None):
    """Band gather in the int16 wire format (a quarter of the CV_64F bytes): every rank narrows its band,
    the bands travel as int16, rank 0 widens the received rows back into its CV_64F result.  Call with the
    pair's stream current (torch.cuda.stream(ExternalStream(ctx.stream)))."""
    if world == 1:
        return
    r0, r1 = band_rows(h, rank, world)
    if rank != 0:
        pair.pack16(r0, r1)
    gather_bands(full16.view(torch.uint8), h, rank, world, group)   # NCCL has no int16: the rows travel as bytes
    if rank == 0:
        pair.unpack16(r1, h)       # rank 0 owns rows [0, r1); the rest arrived as int16


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
