"""2-rank NCCL probe: does an asymmetric batched P2P pattern (rank 0 groups send+recv, rank 1 issues them as
separate batches) complete?  Run: torchrun --nproc-per-node 2 tools/p2p_probe.py"""
import datetime
import os
import sys

import torch
import torch.distributed as dist

rank = int(os.environ["RANK"])
torch.cuda.set_device(rank)
dev = torch.device("cuda", rank)
dist.init_process_group("nccl", timeout=datetime.timedelta(seconds=25), device_id=dev)
peer = 1 - rank


def ex(sends, recvs):
    ops = [dist.P2POp(dist.isend, t, peer) for t in sends] + [dist.P2POp(dist.irecv, t, peer) for t in recvs]
    for w in dist.batch_isend_irecv(ops):
        w.wait()


a, b, c, d = (torch.full((65536,), float(i + 1), device=dev) for i in range(4))
ra, rb, rc, rd = (torch.zeros(65536, device=dev) for _ in range(4))
mode = sys.argv[1] if len(sys.argv) > 1 else "asym"
if rank == 0:
    ex([a], [])
    if mode == "asym":
        ex([b], [rc])
    else:
        ex([b], [])
        ex([], [rc])
    ex([], [rd])
else:
    ex([], [ra])
    ex([], [rb])
    torch.cuda._sleep(int(2e8))
    ex([c], [])
    ex([d], [])
torch.cuda.synchronize()
print(f"rank {rank} mode {mode}: ok", flush=True)
dist.destroy_process_group()
