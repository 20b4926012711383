"""Headline benchmark: training throughput (tokens/s, whole job) of the reference GPT on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--parallel dp|tp|pp] [--impl native|reference]

Metric / config (BASELINE.json): GPT 12L / d512 / 16 heads / ff2048 / seq 512 / vocab 50258
(configs/model_config.yaml), train_config_{dp,tp,pp}.yaml shapes (batch 8 per data-parallel rank — weak scaling:
global batch = 8·N for DP; TP / PP keep batch 8), clip 1.0 + AdamW lr 3e-4 wd 0.1, bf16 compute with fp32 master
weights / optimizer state, synthetic FineWeb-edu-shaped tokens and random-init weights (no network on the box).

Two numbers per run, both over EXACTLY K full training steps (forward, backward, gradient reduction, clip,
AdamW — nothing skipped), bracketed by barrier + synchronize, CUDA events, max over ranks:
  value : the captured step graph replayed back to back with device-resident batches (kernel time)
  e2e   : the public API ``train_step(state, Batch(x, y))`` fed from pinned host memory every step (H2D of x and
          y) with a blocking device→host read of the loss every step — what a user of main.py gets.
L2 is not flushed explicitly: one step touches ≈1.8 GB of parameters / optimizer state and ≈1.2 GB of
activations per rank, far more than the 126 MB L2, so nothing survives between iterations (stated in config).
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

BASELINE_TOKENS_PER_S = {"dp": 27887.0, "tp": 27919.0, "pp": 19978.0}  # BASELINE.md §1 (published logs, n=2, fp32)


def reference_arm(args) -> int:
    """The unmodified reference needs jax/flax/optax/datasets + the HF hub; none is installable offline."""
    why = ("reference cannot be installed offline: pip install of /root/reference fails at metadata generation "
           "(flat layout, no build-system/packages) and jax/jaxlib/flax/optax/datasets are neither on the image nor "
           "in /opt/wheelhouse; its data path also needs the HF hub (gpt2 tokenizer, FineWeb-edu)")
    try:
        sys.path.insert(0, os.path.join(ROOT, "baseline", "_ref"))
        import jax  # noqa: F401
        import flax  # noqa: F401
        import optax  # noqa: F401
    except Exception:
        if int(os.environ.get("RANK", "0")) == 0:
            print(json.dumps({"impl": "reference", "unavailable": why}))
        return 0
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps({"impl": "reference", "unavailable": "jax importable but the reference's HF data path is offline"}))
    return 0


def run(args) -> dict | None:
    import numpy as np
    import torch

    from distributed_training_compare_jax_b200.config.schema import (TrainConfig, load_model_config, load_optim_config,
                                                                  mesh_shape)
    from distributed_training_compare_jax_b200.data.fineweb_edu import VOCAB_SIZE, get_batch_iterator
    from distributed_training_compare_jax_b200.models.param_tree import LocalShape
    from distributed_training_compare_jax_b200.ops import native
    from distributed_training_compare_jax_b200.parallel.mesh import build_mesh, shutdown
    from distributed_training_compare_jax_b200.parallel.pp import stage_layers
    from distributed_training_compare_jax_b200.train.create_train_step import (Batch, create_pp_train_step,
                                                                            create_train_step)
    from distributed_training_compare_jax_b200.train.train import build_state
    from distributed_training_compare_jax_b200.utils.timing import ClockSampler, DeviceTimer

    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run "
                         f"--nproc-per-node {args.gpus}")
    strategy = args.parallel
    per_rank_batch = args.batch
    tc = TrainConfig(seed=0, parallel=strategy, batch=per_rank_batch, steps=args.steps, log_every=10 ** 9,
                     output_dir="outputs/bench", pp_microbatches=args.pp_microbatches, warmup_steps=args.warmup,
                     global_batch_is_per_device=True, sync_every_step=True, backend=args.backend,
                     pp_schedule=args.pp_schedule, fused_comm=not args.no_fused_comm, cuda_graph=not args.no_graph,
                     dp_overlap=not args.no_dp_overlap, tp_fused=args.tp_fused,
                     tp_seq_parallel=not args.no_tp_seq_parallel, pp_cuda_graph=args.pp_graph)
    mc = load_model_config(os.path.join(ROOT, args.model_config), vocab_size=VOCAB_SIZE, parallel=strategy)
    oc = load_optim_config(os.path.join(ROOT, "configs/optim_config.yaml"))
    dp, tp, pp = mesh_shape(tc, world)
    mesh = build_mesh(dp, tp, pp)
    T = mc.max_seq_len
    if strategy == "pp":
        layers = stage_layers(mc.n_layers, pp, False)[mesh.pp.rank]
        local = LocalShape(mc, layer_start=layers.start, layer_end=layers.stop, has_embed=mesh.pp.rank == 0,
                           has_head=mesh.pp.rank == pp - 1)
        global_batch = per_rank_batch
        M = tc.pp_microbatches
        state = build_state(tc, mc, oc, mesh, global_batch // M, M, global_batch * T, local, clip_scope=tc.pp_clip,
                            clip_scale=1.0 / pp, pp_stage_keys=True)
        train_step = create_pp_train_step(M, pp, tc.pp_schedule, cuda_graph=tc.pp_cuda_graph and tc.cuda_graph)
        it = get_batch_iterator(global_batch, T + 1, seed=0)
    else:
        global_batch = per_rank_batch * dp
        local = LocalShape(mc, tp_rank=mesh.tp.rank, tp_size=tp)
        chains = args.chains if tp == 1 else 1
        state = build_state(tc, mc, oc, mesh, per_rank_batch // chains, chains, global_batch * T, local)
        train_step = create_train_step(strategy, cuda_graph=tc.cuda_graph)
        it = get_batch_iterator(global_batch, T + 1, seed=0, rank=mesh.dp.rank, world=dp)
    tokens_per_step = global_batch * T
    dev = mesh.device

    def host_batch() -> Batch:
        a = next(it)
        return Batch(x=np.ascontiguousarray(a[:, :-1]), y=np.ascontiguousarray(a[:, 1:]))

    # ---- warm-up: W real training steps through the public API (also triggers graph capture) ----
    loss = None
    for _ in range(max(args.warmup, 3)):
        state, loss = train_step(state, host_batch(), None)
    first_loss = float(loss)
    if dev.type == "cuda":
        torch.cuda.synchronize()
    launches_before = native.launches()  # graph replays are accounted with the launches seen at capture time

    gpu_index = dev.index if dev.type == "cuda" else 0
    timer = DeviceTimer(dev)

    # ---- (1) device-timed: K steps, batches already resident on the device ----
    dev_batches = []
    for _ in range(min(args.steps, 8)):
        b = host_batch()
        dev_batches.append(Batch(x=torch.as_tensor(b.x).to(dev), y=torch.as_tensor(b.y).to(dev)))
    with ClockSampler(gpu_index) as clocks:
        mesh.barrier()
        timer.start()
        for i in range(args.steps):
            state, loss = train_step(state, dev_batches[i % len(dev_batches)], None)
        ms_dev = timer.stop()
        mesh.barrier()
        ms_dev = mesh.max_over_ranks(ms_dev)

        # ---- (2) end to end: pinned host batch -> H2D every step, blocking loss read every step ----
        host_batches = [host_batch() for _ in range(min(args.steps, 8))]
        pinned = [Batch(x=torch.as_tensor(b.x).pin_memory() if dev.type == "cuda" else torch.as_tensor(b.x),
                        y=torch.as_tensor(b.y).pin_memory() if dev.type == "cuda" else torch.as_tensor(b.y))
                  for b in host_batches]
        mesh.barrier()
        timer.start()
        last = 0.0
        for i in range(args.steps):
            state, loss = train_step(state, pinned[i % len(pinned)], None)
            last = float(loss)  # device -> host read of the step's result
        ms_e2e = timer.stop()
        mesh.barrier()
        ms_e2e = mesh.max_over_ranks(ms_e2e)
    gpu_launches = native.launches() - launches_before

    opt = state.optimizer
    comm_path = {"dp": ("symmetric-memory reduce-scatter + sharded AdamW + all-gather ("
                        + ("multimem/NVLS" if getattr(opt, "uses_multicast", False) else "peer loads/stores") + ")")
                 if type(opt).__name__ == "FusedShardedStep" else ("NCCL bucketed all-reduce" if dp > 1 else "none"),
                 "tp": ("sequence-parallel: GEMM-epilogue reduce-scatter (P2P stores) + LayerNorm all-gather ("
                        + ("multimem.st" if state.engine.sp.use_mc else "peer stores") + ")"
                        if hasattr(state.engine, "sp") else
                        "GEMM epilogue multimem.red all-reduce" if getattr(state.engine, "tp_hooks", None) is not None
                        else ("GEMM + NCCL all-reduce" if tp > 1 else "none")),
                 "pp": "NCCL send/recv, one communicator per direction" if pp > 1 else "none"}
    tps = tokens_per_step * args.steps / (ms_dev / 1e3)
    tps_e2e = tokens_per_step * args.steps / (ms_e2e / 1e3)
    # bytes actually copied host->device per step over the whole job: every DP/TP rank stages its own x and y;
    # in PP only the first stage copies x and only the last stage copies y
    h2d = 2 * per_rank_batch * T * 4 * (1 if strategy == "pp" else world)
    result = {
        "metric": "training throughput, tokens/s (whole job), " + strategy.upper(),
        "value": tps, "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak" if strategy == "dp" else "strong",
        "vs_baseline": tps / BASELINE_TOKENS_PER_S[strategy] if strategy in BASELINE_TOKENS_PER_S else None,
        "dtype": "bf16", "data": "synthetic FineWeb-edu-shaped tokens (Zipf + bigram), random-init weights",
        "impl": "native",
        "config": {"model": f"GPT {mc.n_layers}L/d{mc.d_model}/h{mc.n_heads}/ff{mc.d_ff}/seq{T}/vocab{mc.vocab_size} "
                            f"({args.model_config})", "global_batch": global_batch, "seq_len": T,
                   "parallelism": f"{strategy}{world}", "optimizer": "clip_by_global_norm(1.0)+AdamW(3e-4, wd 0.1)",
                   "l2": "no explicit flush: per-step working set (~3 GB/rank) >> 126 MB L2",
                   "backend": state.engine.backend, "cuda_graph": bool(tc.cuda_graph and state.engine.backend == "native"),
                   "fused_comm": tc.fused_comm, "comm_path": comm_path},
        "e2e": {"value": tps_e2e, "unit": "tokens/s", "ms_per_step": ms_e2e / args.steps,
                "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 4 * world},
        "gpu_launches": int(gpu_launches),  # rank 0, both timed regions (2K steps)
        "clocks": clocks.summary(), "loss_first": first_loss, "loss_last": last,
    }
    return result if mesh.is_main else None


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--parallel", default="dp", choices=["dp", "tp", "pp", "dp_tp"])
    ap.add_argument("--batch", type=int, default=8, help="batch per data-parallel rank (train_config_*.yaml: 8)")
    ap.add_argument("--pp-microbatches", type=int, default=2)
    ap.add_argument("--pp-schedule", default="gpipe")
    ap.add_argument("--model-config", default="configs/model_config.yaml")
    ap.add_argument("--backend", default="auto")
    ap.add_argument("--no-fused-comm", action="store_true")
    ap.add_argument("--no-dp-overlap", action="store_true")
    ap.add_argument("--tp-fused", action="store_true", help="TP: fused GEMM + multimem.red all-reduce epilogue")
    ap.add_argument("--tp-seq-parallel", action="store_true", help="(default) TP: row-sharded residual stream, "
                    "GEMM-epilogue reduce-scatter + LayerNorm all-gather")
    ap.add_argument("--no-tp-seq-parallel", action="store_true", help="TP: replicated activations, GEMM + NCCL all-reduce")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--pp-graph", action="store_true",
                    help="PP: one CUDA graph per (forward | backward, microbatch) of a stage (experimental)")
    ap.add_argument("--chains", type=int, default=1, help="concurrent chains the local batch is split into (DP)")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    if args.gpus > 1 and "RANK" not in os.environ:
        # convenience: re-launch ourselves under torchrun on one node
        import subprocess
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29541", os.path.abspath(__file__), *sys.argv[1:]]
        return subprocess.call(cmd)
    if os.environ.get("DTC_HANG_DUMP"):  # debugging aid: dump every thread's stack if the run stalls
        import faulthandler
        faulthandler.dump_traceback_later(int(os.environ["DTC_HANG_DUMP"]), exit=True)
    os.environ.setdefault("RANK", "0")
    os.environ.setdefault("WORLD_SIZE", "1")
    os.environ.setdefault("LOCAL_RANK", "0")
    res = run(args)  # engine, CUDA graph and communicator users are released when run() returns
    if res is not None:
        print(json.dumps(res), flush=True)
    import gc
    gc.collect()
    from distributed_training_compare_jax_b200.parallel.mesh import shutdown
    shutdown(hard_exit_after_s=20)
    return 0


if __name__ == "__main__":
    sys.exit(main())
