"""Micro-benchmark of the fused compute+collective kernels against their NCCL counterparts and the NVLink roofline.
    torchrun --nproc-per-node N tools/comm_bench.py
DP: reduce-scatter of the 89.6 M-parameter fp32 gradient (multimem.ld_reduce) and sharded AdamW + bf16 all-gather
(multimem.st) vs one NCCL all-reduce of the same buffer.  TP: row-parallel GEMM whose epilogue all-reduces through
multimem.red (+ flag barrier) vs the same GEMM followed by an NCCL all-reduce.  CUDA events, max over ranks."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_training_compare_jax_b200.ops import functional as Fn  # noqa: E402
from distributed_training_compare_jax_b200.ops import native  # noqa: E402
from distributed_training_compare_jax_b200.parallel.symm import FlagBarrier, SymmBuffer  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dev = torch.device("cuda", rank)
dist.init_process_group("nccl", device_id=dev)
grp = dist.group.WORLD
LINK = 770e9  # measured peer copy bandwidth per direction per GPU (B200_PROFILING.md)


def timed(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item() * 1e3  # us


out = {"n_gpus": world}
# ---------------------------------------------------------------- DP
n = 89_633_792  # flat buffer of the reference model (64-element aligned)
grad = SymmBuffer(n, torch.float32, dev, grp)
comp = SymmBuffer(n, torch.bfloat16, dev, grp)
bar = FlagBarrier(dev, grp)
per = -(-(n // world) // 64) * 64
off = min(rank * per, n)
cnt = max(0, min(n, off + per) - off)
npart = int(native.load().dtc_dp_num_partials(cnt))
partials = torch.zeros(npart, device=dev)
p = torch.randn(n, device=dev)
m = torch.zeros(n, device=dev)
v = torch.zeros(n, device=dev)
scal = torch.tensor([1.0, 1.0], device=dev)
step = torch.ones(1, device=dev, dtype=torch.int32)
grad.tensor.normal_()


def rs():
    bar(0)
    native.call("dtc_dp_reduce_scatter", grad.tensor, grad.mc_ptr, grad.ptr_array, rank, world, off, cnt, partials)


def adam_ag():
    native.call("dtc_dp_adamw_allgather", p, m, v, grad.tensor, comp.mc_ptr, comp.ptr_array, world, off, cnt, n, scal, step,
                3e-4, 0.9, 0.999, 1e-8, 0.1)
    bar(2)


g2 = torch.randn(n, device=dev)
t_rs, t_ag = timed(rs), timed(adam_ag)
t_nccl = timed(lambda: dist.all_reduce(g2))
bytes_rs = (world - 1) / world * n * 4        # every rank ships the (n-1)/n of its gradient it does not own
bytes_ag = (world - 1) / world * n * 2 / 1    # every rank receives the other ranks' bf16 shards
out["dp"] = {"multicast": bool(grad.mc_ptr), "reduce_scatter_us": t_rs, "reduce_scatter_GBps_per_gpu": bytes_rs / t_rs / 1e3,
             "reduce_scatter_frac_of_770GBps": bytes_rs / (t_rs * 1e-6) / LINK,
             "adamw_allgather_us": t_ag, "allgather_GBps_per_gpu": bytes_ag / t_ag / 1e3,
             "nccl_allreduce_fp32_us": t_nccl, "nccl_busbw_GBps": 2 * (world - 1) / world * n * 4 / t_nccl / 1e3}
del p, m, v, g2
# ---------------------------------------------------------------- TP (fc2-like row-parallel GEMM of the reference model)
M, D, F = 4096, 512, 2048 // world
x = (torch.randn(M, F, device=dev)).to(torch.bfloat16)
w = (torch.randn(F, D, device=dev) * 0.05).to(torch.bfloat16)
outb = torch.empty(M, D, device=dev, dtype=torch.bfloat16)


def unfused():
    Fn.gemm(x, w, b_mn=True, out=outb)
    dist.all_reduce(outb)


def graphed(fn, reps=20):
    """The GPU box's host needs ~15-40 us per launch: compare device time by replaying a captured graph."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    return lambda: g.replay(), reps


ru, reps = graphed(unfused)
t_u = timed(ru, iters=5) / reps
out["tp"] = {"gemm_plus_nccl_allreduce_us": t_u, "wire_time_at_770GBps_us": M * D * 2 / LINK * 1e6, "shape": [M, D, F]}
# ---------------------------------------------------------------- TP, sequence-parallel form (parallel/tp_sp.py)
# one sublayer boundary: row-parallel GEMM -> [sum over ranks + bias + residual] -> LayerNorm -> operand of the next GEMM
#   plain : GEMM(+bias+residual epilogue) -> NCCL all-reduce of [M, D] -> LayerNorm over all M rows on every rank
#   seqpar: GEMM with the scatter epilogue (P2P stores into the owner's slots) -> flag barrier -> reduce + bias + residual +
#           LayerNorm on the M/n owned rows, output pushed to every rank (multimem.st) -> flag barrier
if M % (world * 128) == 0:
    import ctypes as C
    Ms, row0 = M // world, rank * (M // world)
    arena = SymmBuffer(2 * M * D, torch.bfloat16, dev, grp)          # [partial-sum slots | gathered LayerNorm output]
    part, xg = arena.tensor[:M * D].view(world, Ms, D), arena.tensor[M * D:].view(M, D)
    part_ptrs = (C.c_ulonglong * world)(*arena.ptrs)
    xg_ptrs = (C.c_ulonglong * world)(*[q + M * D * 2 for q in arena.ptrs])
    xg_mc = arena.mc_ptr + M * D * 2 if arena.mc_ptr else 0
    res, bias = torch.randn(M, D, device=dev).to(torch.bfloat16), torch.randn(D, device=dev).to(torch.bfloat16)
    gam, bet = torch.ones(D, device=dev, dtype=torch.bfloat16), torch.zeros(D, device=dev, dtype=torch.bfloat16)
    hsum = torch.empty(Ms, D, device=dev, dtype=torch.bfloat16)
    mean, rstd = torch.empty(M, device=dev), torch.empty(M, device=dev)
    lnout = torch.empty(M, D, device=dev, dtype=torch.bfloat16)
    sslot = [0]

    def sp_bar():
        bar(3 + sslot[0] % 12)
        sslot[0] += 1

    def seqpar():
        Fn.gemm_scatter(x, w, part_ptrs, world, rank, D, b_mn=True)
        sp_bar()
        Fn.sp_ln_fwd(res[row0:row0 + Ms], bias, part, world, gam, bet, hsum, mean[:Ms], rstd[:Ms], xg_mc, xg_ptrs, world, row0)
        sp_bar()

    def plain():
        Fn.gemm(x, w, b_mn=True, bias=bias, epi="bias_residual", aux=res, out=outb)
        dist.all_reduce(outb)
        Fn.layernorm_fwd(outb, gam, bet, out=lnout, mean=mean, rstd=rstd)

    rs_, reps = graphed(seqpar)
    t_sp = timed(rs_, iters=5) / reps
    rp_, reps = graphed(plain)
    t_pl = timed(rp_, iters=5) / reps
    out["tp_seqpar"] = {"gemm_scatter_barrier_reduceLN_push_barrier_us": t_sp, "gemm_nccl_allreduce_layernorm_us": t_pl,
                        "multicast": bool(xg_mc), "shape": [M, D, F],
                        "wire_bytes_per_gpu": (world - 1) / world * M * D * 2 + Ms * D * 2,
                        "wire_time_at_770GBps_us": ((world - 1) / world * M * D * 2 + Ms * D * 2) / LINK * 1e6}
    del rs_, rp_
if rank == 0:
    print(json.dumps(out), flush=True)
# CUDA graphs that captured NCCL work must be gone before the process group is destroyed (else destroy can hang)
for name in ("rf", "ru"):
    globals().pop(name, None)
import gc  # noqa: E402

gc.collect()
torch.cuda.synchronize()
dist.barrier()
os._exit(0)
