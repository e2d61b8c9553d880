"""World-size-2 gloo test of the row-band sharding path (CPU): each rank computes its band with the
oracle (the checker stands in for the GPU search here), bands are gathered with the same
gather_bands() the NCCL path uses, and rank 0 must hold exactly the unsharded result."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import orc
    from stereo_reconstruction_b200 import sharding
    w, h, D, B = 320, 75, 40, 9
    L, R = orc.synth_pair(0xB200, 0, w, h, 0, D)
    r0, r1 = sharding.band_rows(h, rank, world)
    band = orc.block_search_right_fast(L, R, B, 0, D, 1.0, rows=(r0, r1), nthreads=1)
    full = torch.zeros((h, w), dtype=torch.float64)
    full[r0:r1] = torch.from_numpy(band[r0:r1])
    sharding.gather_bands(full, h, rank, world)
    # the int16 wire-format gather (gather_bands16) with a CPU stand-in for capi.Pair's pack16 / unpack16
    class FakePair:
        def __init__(self, t):
            self.full = t
            self.full16 = torch.zeros(t.shape, dtype=torch.int16)

        def pack16(self, a, b):
            self.full16[a:b] = self.full[a:b].to(torch.int16)

        def unpack16(self, a, b):
            self.full[a:b] = self.full16[a:b].to(torch.float64)

    full2 = torch.zeros((h, w), dtype=torch.float64)
    full2[r0:r1] = torch.from_numpy(band[r0:r1])
    fp = FakePair(full2)
    sharding.gather_bands16(fp, fp.full16, h, rank, world)
    if rank == 0:
        want = orc.block_search_right_fast(L, R, B, 0, D, 1.0, nthreads=1)
        q.put(bool(np.array_equal(full.numpy(), want)) and bool(np.array_equal(full2.numpy(), want)))
    assert sharding.pairs_of_rank(5, rank, world) == list(range(rank, 5, world))
    dist.barrier()
    dist.destroy_process_group()


def test_band_rows_cover():
    sys.path.insert(0, ROOT)
    from stereo_reconstruction_b200 import sharding
    for h in (1, 7, 75, 2160):
        for world in (1, 2, 3, 8):
            bands = [sharding.band_rows(h, r, world) for r in range(world)]
            assert bands[0][0] == 0 and bands[-1][1] == h
            assert all(bands[i][1] == bands[i + 1][0] for i in range(world - 1))


def test_gather_bands_gloo_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True
