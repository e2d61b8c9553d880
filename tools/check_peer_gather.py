"""Multi-process check of the band gather over peer memory (run under torchrun, one rank per GPU):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/check_peer_gather.py
Every step searches a new synthetic pair by row bands, gathers the bands into rank 0's HBM (sharding.PeerGather)
and rank 0 compares the gathered map with its own search of the whole pair.  No barrier between steps: the
acknowledgement word is what keeps a fast rank from overwriting rows rank 0 has not consumed yet."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
import stereo_reconstruction_b200 as s
from stereo_reconstruction_b200 import capi, sharding, synth

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ctx = capi.Context(local)
w, h, D, block = (1283 if os.environ.get("SR_CHECK_ODD") else 1284), 721, 64, 9   # even width: gather fused into the resolve pass; odd: push kernel
pair = capi.Pair(ctx, w, h, w, h)
p = capi.search_params(block=block, min_disp=0, max_disp=D, smooth=1.0, cost=capi.COST_SAD)
peer = sharding.PeerGather(pair, h, rank, world)
r0, r1 = sharding.band_rows(h, rank, world)
steps, bad = 12, 0
gathered = []
for step in range(steps):
    L, R = synth.synth_pair(0xB200, step, w, h, 0, D)
    pair.upload(L, R)
    if step % 4 == 3:
        pair.run(p, r0, r1)           # the two-kernel form of the same protocol step
        peer.gather()
    else:
        peer.run_and_gather(p)        # ranks > 0: sr_search_run_push
    if rank == 0:
        gathered.append((L, R, pair.download(p)))
    elif step % 3 == rank % 3:
        time.sleep(0.02)              # ranks drift apart
if rank == 0:
    for step, (L, R, got) in enumerate(gathered):
        pair.upload(L, R)
        pair.run(p)
        if not np.array_equal(got, pair.download(p)):
            bad += 1
            print(f"step {step}: gathered map differs", flush=True)
    print(f"peer gather over {world} ranks: {steps - bad}/{steps} steps identical to the single-GPU map", flush=True)
pair.peer_status()                    # a failed protocol step is an error on every rank
ctx.sync()
dist.barrier()
dist.destroy_process_group()
sys.exit(1 if bad else 0)
