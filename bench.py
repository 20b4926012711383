"""Headline benchmark: training throughput (tokens/s, whole job) of the reference GPT on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--arms dp,dp_weak,tp,pp] [--impl native|reference]

Metric / config (BASELINE.json): GPT 12L / d512 / 16 heads / ff2048 / seq 512 / vocab 50258
(configs/model_config.yaml), train_config_{dp,tp,pp}.yaml shapes, clip 1.0 + AdamW lr 3e-4 wd 0.1, bf16 compute
with fp32 master weights / optimizer state, synthetic FineWeb-edu-shaped tokens and random-init weights (no network).

One invocation at N GPUs times, in this order, the reference's three strategies on the reference's shapes:
  dp       HEADLINE.  Data parallel at the reference's batch semantics: ``batch: 8`` is the GLOBAL batch, sharded over
           the data axis (reference train/create_train_step.py:36-40) -> strong scaling, 8/N sequences per rank.
  dp_weak  data parallel with batch 8 PER RANK (global batch 8·N)                      -> ``extra.dp_weak``
  tp       tensor parallel over all N GPUs, global batch 8 (train_config_tp.yaml)       -> ``extra.tp``
  pp       pipeline parallel, N stages, 2 microbatches (train_config_pp.yaml)           -> ``extra.pp``
At N = 1 the four coincide (one GPU, batch 8) and only the headline arm runs.  ``--parallel X`` runs a single arm
(tools / A-B scripts).  An arm that fails or exceeds its time limit is reported as ``{"error": ...}`` under ``extra``
and does not take the headline number down with it.

Two numbers per arm, both over EXACTLY K full training steps (forward, backward, gradient reduction, clip, AdamW —
nothing skipped), bracketed by barrier + synchronize, CUDA events, max over ranks:
  value : the captured step graph replayed back to back with device-resident batches (kernel time)
  e2e   : the public API ``train_step(state, Batch(x, y))`` fed from pinned host memory every step (H2D of x and
          y) with a blocking device→host read of the loss every step — what a user of main.py gets.
L2 is not flushed explicitly: one step touches ≈1.8 GB of parameters / optimizer state and ≈1.2 GB of
activations per rank, far more than the 126 MB L2, so nothing survives between iterations (stated in config).
SM clock / throttle reasons are sampled through NVML every 10 ms during both timed regions.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# the reference's published logs (n = 2 GPUs inferred, fp32, hardware unstated, host/data-bound)
BASELINE_TOKENS_PER_S = {"dp": 27887.0, "dp_weak": 27887.0, "tp": 27919.0, "pp": 19978.0}
ARM_STRATEGY = {"dp": "dp", "dp_weak": "dp", "tp": "tp", "pp": "pp", "dp_tp": "dp_tp"}


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


def run_arm(args, arm: str) -> dict:
    """Build one strategy on the current world, warm it up, time K steps twice (device-resident and end to end).
    Every rank returns the same dict (times are max over ranks; losses come from rank 0's view of the global loss)."""
    import gc

    import numpy as np
    import torch

    from distributed_training_compare_jax_b200.config.schema import (TrainConfig, load_model_config, load_optim_config,
                                                                  mesh_shape)
    from distributed_training_compare_jax_b200.data.fineweb_edu import VOCAB_SIZE, get_batch_iterator
    from distributed_training_compare_jax_b200.models.param_tree import LocalShape
    from distributed_training_compare_jax_b200.ops import native
    from distributed_training_compare_jax_b200.parallel.mesh import build_mesh
    from distributed_training_compare_jax_b200.parallel.pp import bubble_fraction, stage_layers
    from distributed_training_compare_jax_b200.train.create_train_step import (Batch, create_pp_train_step,
                                                                            create_train_step)
    from distributed_training_compare_jax_b200.train.train import build_state
    from distributed_training_compare_jax_b200.utils.timing import ClockSampler, DeviceTimer

    world = int(os.environ.get("WORLD_SIZE", "1"))
    strategy = ARM_STRATEGY[arm]
    fault = os.environ.get("DTC_BENCH_FAULT", "")   # test hook "<arm>:<rank>": that rank raises at the start of that arm
    if fault and fault.split(":")[0] == arm and int(fault.split(":")[1]) == int(os.environ.get("RANK", "0")):
        raise RuntimeError(f"injected fault in arm `{arm}` (DTC_BENCH_FAULT)")
    weak = arm == "dp_weak" or (arm == "dp_tp")          # batch is per data-parallel rank
    tc = TrainConfig(seed=0, parallel=strategy, batch=args.batch, steps=args.steps, log_every=10 ** 9,
                     output_dir="outputs/bench", pp_microbatches=args.pp_microbatches, warmup_steps=args.warmup,
                     global_batch_is_per_device=weak, sync_every_step=True, backend=args.backend,
                     pp_schedule=args.pp_schedule, fused_comm=not args.no_fused_comm, cuda_graph=not args.no_graph,
                     dp_overlap=not args.no_dp_overlap,
                     tp_seq_parallel=not args.no_tp_seq_parallel, pp_cuda_graph=not args.no_pp_graph)
    mc = load_model_config(os.path.join(ROOT, args.model_config), vocab_size=VOCAB_SIZE, parallel=strategy)
    oc = load_optim_config(os.path.join(ROOT, "configs/optim_config.yaml"))
    dp, tp, pp = mesh_shape(tc, world)
    T = mc.max_seq_len
    global_batch = args.batch * (dp if weak else 1)
    if strategy != "pp" and global_batch % dp:
        raise ValueError(f"global batch {global_batch} is not divisible by {dp} data-parallel ranks")
    mesh = build_mesh(dp, tp, pp)
    if strategy == "pp":
        layers = stage_layers(mc.n_layers, pp, False, 3.5 if args.pp_balanced else 0.0)[mesh.pp.rank]
        local = LocalShape(mc, layer_start=layers.start, layer_end=layers.stop, has_embed=mesh.pp.rank == 0,
                           has_head=mesh.pp.rank == pp - 1)
        M = tc.pp_microbatches
        state = build_state(tc, mc, oc, mesh, global_batch // M, M, global_batch * T, local, clip_scope=tc.pp_clip,
                            clip_scale=1.0 / pp, pp_stage_keys=True)
        train_step = create_pp_train_step(M, pp, tc.pp_schedule, cuda_graph=tc.pp_cuda_graph and tc.cuda_graph)
        it = get_batch_iterator(global_batch, T + 1, seed=0)
        local_batch = global_batch
    else:
        local_batch = global_batch // dp
        local = LocalShape(mc, tp_rank=mesh.tp.rank, tp_size=tp)
        chains = args.chains if (tp == 1 and local_batch % max(args.chains, 1) == 0) else 1
        state = build_state(tc, mc, oc, mesh, local_batch // chains, chains, global_batch * T, local)
        train_step = create_train_step(strategy, cuda_graph=tc.cuda_graph)
        it = get_batch_iterator(global_batch, T + 1, seed=0, rank=mesh.dp.rank, world=dp)
    tokens_per_step = global_batch * T
    dev = mesh.device

    def host_batch() -> Batch:
        a = next(it)
        return Batch(x=np.ascontiguousarray(a[:, :-1]), y=np.ascontiguousarray(a[:, 1:]))

    # ---- warm-up: W real training steps through the public API (also triggers graph capture) ----
    warm = max(args.warmup, 4)   # eager, eager, capture + replay, replay
    loss = None
    for _ in range(warm):
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
    eng = state.engine
    comm_path = {"dp": ("symmetric-memory reduce-scatter + sharded AdamW + all-gather ("
                        + ("multimem/NVLS" if getattr(opt, "uses_multicast", False) else "peer loads/stores") + ")")
                 if type(opt).__name__ == "FusedShardedStep" else ("NCCL bucketed all-reduce" if dp > 1 else "none"),
                 "tp": ("sequence-parallel: GEMM-epilogue reduce-scatter (P2P stores) + LayerNorm all-gather ("
                        + ("multimem.st" if eng.sp.use_mc else "peer stores") + ")"
                        if hasattr(eng, "sp") else
                        ("GEMM + NCCL all-reduce" if tp > 1 else "none")),
                 "pp": ("NCCL send/recv, one communicator per direction"
                        + (", one CUDA graph per (F|B, microbatch)" if tc.pp_cuda_graph and tc.cuda_graph else ", eager launches")
                        ) if pp > 1 else "none"}
    tps = tokens_per_step * args.steps / (ms_dev / 1e3)
    tps_e2e = tokens_per_step * args.steps / (ms_e2e / 1e3)
    # bytes actually copied host->device per step over the whole job: every DP/TP rank stages its own x and y
    # (DP: its slice of the global batch; TP: the replicated batch); in PP every stage stages x and y of the whole batch
    # into pinned memory but only the first stage copies x and only the last stage copies y
    h2d = 2 * local_batch * T * 4 * (1 if strategy == "pp" else world)
    res = {
        "metric": "training throughput, tokens/s (whole job), " + arm.upper(),
        "value": tps, "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": warm,
        "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak" if weak else "strong",
        "vs_baseline": tps / BASELINE_TOKENS_PER_S[arm] if arm in BASELINE_TOKENS_PER_S else None,
        "vs_baseline_note": "published reference log: n = 2 GPUs (inferred), fp32, global batch 8, host/data-bound, "
                            "hardware unstated (BASELINE.md §1) — context, not a like-for-like ratio",
        "dtype": "bf16", "data": "synthetic FineWeb-edu-shaped tokens (Zipf + bigram), random-init weights",
        "impl": "native",
        "config": {"model": f"GPT {mc.n_layers}L/d{mc.d_model}/h{mc.n_heads}/ff{mc.d_ff}/seq{T}/vocab{mc.vocab_size} "
                            f"({args.model_config})", "global_batch": global_batch, "seq_len": T,
                   "parallelism": f"{strategy}{world}" if strategy != "dp_tp" else f"dp{dp}xtp{tp}",
                   "batch_semantics": "batch 8 per data-parallel rank" if weak else "global batch 8 (reference semantics)",
                   "optimizer": "clip_by_global_norm(1.0)+AdamW(3e-4, wd 0.1)",
                   "l2": "no explicit flush: per-step working set (~3 GB/rank) >> 126 MB L2",
                   "backend": eng.backend, "cuda_graph": bool(tc.cuda_graph and eng.backend == "native"),
                   "fused_comm": tc.fused_comm, "comm_path": comm_path,
                   "wgrad_grouped": bool(getattr(eng, "wgrad_grouped", False))},
        "e2e": {"value": tps_e2e, "unit": "tokens/s", "ms_per_step": ms_e2e / args.steps,
                "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 4 * world},
        "gpu_launches": int(gpu_launches),  # rank 0, both timed regions (2K steps)
        "clocks": clocks.summary(), "loss_first": first_loss, "loss_last": last,
    }
    if strategy == "pp":
        res["config"]["pp_microbatches"] = tc.pp_microbatches
        res["config"]["pp_schedule"] = tc.pp_schedule
        res["config"]["pp_layers_per_stage"] = [len(r) for r in stage_layers(mc.n_layers, pp, False,
                                                                             3.5 if args.pp_balanced else 0.0)]
        res["config"]["bubble_fraction"] = bubble_fraction(pp, tc.pp_microbatches)
    # release the arm: graphs first (they reference NCCL work), then the engine / symmetric buffers
    del train_step, state, opt, eng, dev_batches, pinned
    gc.collect()
    if dev.type == "cuda":
        torch.cuda.synchronize()
        torch.cuda.empty_cache()
    mesh.barrier()
    return res


class _Watchdog:
    """Per-arm time limit.  A hung collective (a rank that never arrives) would otherwise hang the whole benchmark and
    lose the arms that already finished: on expiry rank 0 prints the line assembled so far and every rank exits."""

    def __init__(self, emit, fail_marker: str):
        self.emit, self.deadline, self.arm, self.fail_marker = emit, None, "", fail_marker
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()

    def arm_start(self, arm: str, limit_s: float) -> None:
        self.arm, self.deadline = arm, time.time() + limit_s

    def arm_done(self) -> None:
        self.deadline = None

    def _run(self) -> None:
        while True:
            time.sleep(0.5)
            d = self.deadline
            peer_failed = d is not None and os.path.exists(self.fail_marker)
            if peer_failed or (d is not None and time.time() > d):
                why = "a peer rank failed (see stderr)" if peer_failed else "exceeded its time limit (hung collective?)"
                self.emit(f"arm `{self.arm}`: {why}", self.arm)
                sys.stdout.flush()
                sys.stderr.flush()
                os._exit(0)


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--arms", default="", help="comma list of dp,dp_weak,tp,pp,dp_tp (default: all four reference "
                    "arms when N > 1, dp when N = 1); the first one is the headline")
    ap.add_argument("--parallel", default="", choices=["", "dp", "dp_weak", "tp", "pp", "dp_tp"],
                    help="run ONE arm only (tools); `dp` = global batch 8, `dp_weak` = batch 8 per rank")
    ap.add_argument("--batch", type=int, default=8, help="train_config_*.yaml: batch 8")
    ap.add_argument("--arm-timeout", type=float, default=240.0, help="seconds per arm before the watchdog gives up")
    ap.add_argument("--pp-microbatches", type=int, default=2)
    ap.add_argument("--pp-schedule", default="gpipe")
    ap.add_argument("--pp-balanced", action="store_true", help="PP: fewer layers on the last stage (it also runs lm_head + "
                    "CE); default: the reference's equal layer counts")
    ap.add_argument("--model-config", default="configs/model_config.yaml")
    ap.add_argument("--backend", default="auto")
    ap.add_argument("--no-fused-comm", action="store_true")
    ap.add_argument("--no-dp-overlap", action="store_true")
    ap.add_argument("--tp-seq-parallel", action="store_true", help="(default) TP: row-sharded residual stream, "
                    "GEMM-epilogue reduce-scatter + LayerNorm all-gather")
    ap.add_argument("--no-tp-seq-parallel", action="store_true", help="TP: replicated activations, GEMM + NCCL all-reduce")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--pp-graph", action="store_true", help="(default) PP: one CUDA graph per (F | B, microbatch) of a stage")
    ap.add_argument("--no-pp-graph", action="store_true", help="PP: eager kernel launches from Python")
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
    world = int(os.environ["WORLD_SIZE"])
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run "
                         f"--nproc-per-node {args.gpus}")
    rank0 = int(os.environ["RANK"]) == 0
    if args.parallel:
        arms = [args.parallel]
    elif args.arms:
        arms = [a.strip() for a in args.arms.split(",") if a.strip()]
    else:
        arms = ["dp"] if world == 1 else ["dp", "dp_weak", "tp", "pp"]
    for a in arms:
        if a not in ARM_STRATEGY:
            raise SystemExit(f"unknown arm `{a}`")

    headline: dict = {}
    extra: dict = {}
    printed = threading.Event()

    def emit(error: str | None = None, failed_arm: str = "") -> None:
        if printed.is_set():
            return
        printed.set()
        if error and failed_arm:
            extra.setdefault(failed_arm, {"error": error})
        if rank0:
            line = dict(headline) if headline else {"metric": "training throughput, tokens/s (whole job), " + arms[0].upper(),
                                                    "value": None, "error": error or "headline arm did not finish"}
            if extra:
                line["extra"] = extra
            line["arms"] = arms
            print(json.dumps(line), flush=True)

    # a rank that fails drops this marker; the watchdogs of its peers (stuck in a collective waiting for it) see it,
    # rank 0 prints what has been measured so far, and everybody leaves
    fail_marker = f"/tmp/dtc_bench_failed_{os.getppid()}_{os.environ.get('MASTER_PORT', '0')}"
    dog = _Watchdog(emit, fail_marker)
    poisoned = ""
    for i, arm in enumerate(arms):
        if poisoned:
            extra[arm] = {"error": "skipped: " + poisoned}
            continue
        dog.arm_start(arm, args.arm_timeout)
        try:
            res = run_arm(args, arm)
            if i == 0:
                headline.update(res)
            else:
                extra[arm] = {k: res[k] for k in ("value", "unit", "ms_per_step", "scaling", "vs_baseline", "e2e",
                                                  "gpu_launches", "loss_first", "loss_last", "clocks")}
                extra[arm]["config"] = {k: res["config"][k] for k in res["config"]
                                        if k in ("global_batch", "parallelism", "batch_semantics", "comm_path",
                                                 "pp_microbatches", "pp_schedule", "bubble_fraction", "cuda_graph")}
        except BaseException as e:  # noqa: BLE001 — a failed arm must not lose the arms that finished
            if isinstance(e, (KeyboardInterrupt, SystemExit)):
                raise
            import traceback
            traceback.print_exc()
            msg = f"{type(e).__name__}: {str(e).splitlines()[0] if str(e) else ''}"[:300]
            if i == 0:
                headline.update({"metric": "training throughput, tokens/s (whole job), " + arm.upper(), "value": None,
                                 "error": msg})
            else:
                extra[arm] = {"error": msg}
            # a failed rank leaves its peers inside a collective and (for CUDA faults) a dead context: stop here
            poisoned = f"arm `{arm}` failed ({msg})"
            try:
                open(fail_marker, "w").write(msg)
            except OSError:
                pass
            emit()
            sys.stdout.flush()
            sys.stderr.flush()
            time.sleep(3.0)   # let the peers' watchdogs notice the marker and rank 0 print before the launcher reacts
            os._exit(0 if (headline.get("value") or not rank0) else 1)
        finally:
            dog.arm_done()
    emit()
    import gc
    gc.collect()
    from distributed_training_compare_jax_b200.parallel.mesh import shutdown
    shutdown(hard_exit_after_s=20)
    return 0


if __name__ == "__main__":
    sys.exit(main())
