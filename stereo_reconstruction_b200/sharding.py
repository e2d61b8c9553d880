"""Multi-GPU partitioning of the hot path (one process per GPU, torch.distributed for plumbing).

* config 5 (batch of pairs): pair p -> rank p % world, no collective (`pairs_of_rank`).
* config 4 (one very large pair): rank r owns the output rows `band_rows(h, r, world)`; every rank
  holds the whole pair in HBM (the 2h-1 halo rows are simply read from it), runs the search on
  its band and the bands are gathered to rank 0 -- by stores into rank 0's HBM over NVLink (`PeerGather`,
  capi_peer.cu) or, where peer mapping is unavailable, by point-to-point sends over NCCL (`gather_bands16`).  The same code runs over gloo on CPU tensors (tests/test_sharding_cpu.py).
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


class PeerGather:
    """Band gather over peer memory (capi_peer.cu): rank 0 exports a gather window of its pair, the other ranks map
    it through CUDA IPC and store their int16 bands straight into rank 0's HBM over NVLink; rank 0's collect
    kernel widens each band as its flag arrives.  torch.distributed only carries the 64-byte handle, once.
    `gather()` is asynchronous on the pair's context stream; every rank must call it once per step."""

    def __init__(self, pair, h, rank, world, group=None):
        self.pair, self.h, self.rank, self.world, self.seq = pair, h, rank, world, 0
        if world == 1:
            return
        box = [pair.peer_export() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0, group=group)
        err = 0
        if rank != 0:
            try:
                pair.peer_attach(box[0])
            except capi.SrError:
                err = 1
        # every rank takes the same route: one failed mapping sends all of them back to the NCCL gather
        flag = torch.tensor([err], dtype=torch.int32, device=torch.device("cuda", torch.cuda.current_device()))
        dist.all_reduce(flag, op=dist.ReduceOp.MAX, group=group)
        if int(flag.item()):
            if rank != 0 and not err:
                pair.peer_detach()
            raise RuntimeError("peer memory mapping is not available between these ranks")

    def gather(self):
        if self.world == 1:
            return
        self.seq += 1
        if self.rank == 0:
            self.pair.collect_bands(self.world, self.seq)
        else:
            r0, r1 = band_rows(self.h, self.rank, self.world)
            self.pair.push_band(r0, r1, self.rank, self.seq)

    def run_and_gather(self, params):
        """One step: this rank's band search + the gather.  Ranks > 0 run the search with the gather fused into its
        resolve pass (sr_search_run_push: int16 stores into rank 0's window, no CV_64F band, no push kernel); rank 0
        searches its own band and widens the others' as their flags arrive."""
        r0, r1 = band_rows(self.h, self.rank, self.world)
        if self.world == 1:
            self.pair.run(params, r0, r1)
            return
        self.seq += 1
        if self.rank == 0:
            self.pair.run(params, r0, r1)
            self.pair.collect_bands(self.world, self.seq)
        else:
            self.pair.run_push(params, r0, r1, self.rank, self.seq)


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
