"""Training drivers (component C9; reference: train/train.py:22-104 ``train_dp_tp`` and :107-233 ``train_pp``).

Same signatures and the same observable behaviour as the reference: banner lines ``Warmup`` / ``Start measuring``,
``warmup_steps`` (5) un-timed TRAINING steps that consume data, a timed loop with the ``Step: … | Avg loss: … |
Average step time: …`` line every ``log_every`` steps, ``Total time`` / ``End`` and ``output_dir/log.csv``.
Each process drives ONE device (the reference drives all devices from one process); ``main.py`` spawns them.
Beyond the reference: device-side timing (CUDA events, max over ranks), tokens/s, Flax-layout checkpoints and
resume, optional non-blocking loss fetch, an uneven layer split for PP instead of silently dropping layers.
"""
from __future__ import annotations

import os
from dataclasses import replace

import numpy as np
import torch

from ..config.schema import ModelConfig, OptimConfig, TrainConfig, mesh_shape, validate
from ..data.fineweb_edu import get_batch_iterator
from ..models.param_tree import LocalShape, ParamStore, init_params
from ..parallel.mesh import DeviceMesh, build_mesh
from ..parallel.pp import stage_layers
from ..utils.checkpoint import load_checkpoint, save_checkpoint
from ..utils.logging import RunLogger
from ..utils.timing import DeviceTimer
from .create_optimizer import create_optimizer
from .create_train_step import Batch, TrainState, create_pp_train_step, create_train_step
from .engine import make_engine


def resolve_backend(train_config: TrainConfig, device: torch.device) -> str:
    b = train_config.backend
    if b == "auto":
        b = "native" if device.type == "cuda" else "torch"
    if b == "native":
        from ..ops import native
        native.require_gpu_library()  # fail loudly: no silent eager fallback on a GPU box
    return b


def build_state(train_config: TrainConfig, model_config: ModelConfig, opt_config: OptimConfig, mesh: DeviceMesh,
                micro_batch: int, n_micro: int, global_tokens: int, local: LocalShape, clip_scope: str = "global",
                clip_scale: float = 1.0, pp_stage_keys: bool = False) -> TrainState:
    backend = resolve_backend(train_config, mesh.device)
    native_dp = backend == "native" and mesh.dp.size > 1
    # fused DP path (symmetric-memory reduce-scatter + sharded AdamW + all-gather): pure data parallelism only
    fused_dp = native_dp and train_config.fused_comm and mesh.tp.size == 1 and mesh.pp.size == 1
    store = ParamStore(local, mesh.device, with_compute_copy=(backend == "native"),
                       symm_group=mesh.dp.group if fused_dp else None)
    init_params(store, train_config.seed, pp_stage_keys=pp_stage_keys, stage_index=mesh.pp.rank)
    engine = None
    # (hybrid DP×TP uses it too — validated on 4 and 8 GPUs, profiles/r2; DTC_SP_HYBRID=0 selects GEMM + NCCL all-reduce)
    if (backend == "native" and train_config.tp_seq_parallel and mesh.tp.size > 1 and mesh.pp.size == 1 and n_micro == 1
            and train_config.fused_comm and (mesh.dp.size == 1 or os.environ.get("DTC_SP_HYBRID", "1") != "0")):
        from ..parallel.tp_sp import SeqParallelEngine
        try:   # the conditions checked inside are the same on every rank, so all ranks fall back together
            engine = SeqParallelEngine(model_config, local, mesh, store, micro_batch, model_config.max_seq_len, n_micro,
                                       global_tokens, train_config.seed)
        except (ValueError, RuntimeError) as e:
            if mesh.is_main:
                print(f"[dtc] sequence-parallel TP unavailable ({e}); using the replicated-activation TP path", flush=True)
    if engine is None:
        engine = make_engine(backend, model_config, local, mesh, store, micro_batch, model_config.max_seq_len, n_micro,
                             global_tokens, train_config.seed, seq_parallel=train_config.tp_seq_parallel)
    reducer = None
    if native_dp and not fused_dp:
        from ..parallel.dp import make_grad_reducer
        reducer = make_grad_reducer(engine, mesh)   # NCCL: per-layer buckets overlapped with the backward pass
    if fused_dp:
        from ..parallel.dp import FusedShardedStep
        opt = FusedShardedStep(opt_config, engine, mesh, overlap=train_config.dp_overlap)
    else:
        opt = create_optimizer(opt_config).bind(engine, mesh, clip_scope=clip_scope, clip_scale=clip_scale,
                                                grad_reducer=reducer)
    state = TrainState(engine, opt, mesh)
    if train_config.resume:
        meta = load_checkpoint(train_config.resume, store)
        state.step = int(meta.get("step", 0))
        # a resumed run continues the token stream where the interrupted one stopped and does not repeat the warm-up
        state.batches_consumed = int(meta.get("batches_consumed", state.step))
        state.resumed = True
    return state


def _checkpoint_meta(state: TrainState, mesh: DeviceMesh, strategy: str) -> dict:
    return {"strategy": strategy, "batches_consumed": state.batches_consumed,
            "mesh_shape": [mesh.dp.size, mesh.tp.size, mesh.pp.size]}


def _run_loop(train_config: TrainConfig, model_config: ModelConfig, state: TrainState, train_step, data_iterator,
              mesh: DeviceMesh, tokens_per_step: int, strategy: str) -> dict:
    log = RunLogger(train_config.output_dir, train_config.log_every, mesh.is_main)

    def next_batch() -> Batch:
        x_np = next(data_iterator)
        return Batch(x=np.ascontiguousarray(x_np[:, :-1]), y=np.ascontiguousarray(x_np[:, 1:]))

    log.say("Warmup")
    # the reference's warm-up steps TRAIN and consume data (train/train.py:63-70); a resumed run already did them
    for _ in range(0 if state.resumed else train_config.warmup_steps):
        state, loss = train_step(state, next_batch(), None)
        state.batches_consumed += 1
    if mesh.device.type == "cuda":
        torch.cuda.synchronize()
    mesh.barrier()

    log.say("Start measuring")
    timer = DeviceTimer(mesh.device)
    timer.start()
    log.start_timing()
    pending: list[tuple[int, torch.Tensor]] = []
    fault = os.environ.get("DTC_FAULT", "")     # fault injection for the failure-handling tests: "<rank>:<step>"
    fault_rank, fault_step = (int(v) for v in fault.split(":")) if fault else (-1, -1)
    for step in range(1, train_config.steps + 1):
        if step == fault_step and mesh.rank == fault_rank:
            raise RuntimeError(f"injected fault on rank {mesh.rank} at step {step} (DTC_FAULT)")
        state, loss = train_step(state, next_batch(), None)
        state.batches_consumed += 1
        if train_config.sync_every_step:
            log.record(step, float(loss))          # blocking D2H like the reference (train/train.py:82-83)
        else:
            pending.append((step, loss.detach().clone()))
            if step % train_config.log_every == 0 or step == train_config.steps:
                for s, l in pending:
                    log.record(s, float(l))
                pending = []
        if train_config.checkpoint_dir and train_config.checkpoint_every and step % train_config.checkpoint_every == 0:
            if hasattr(state.optimizer, "sync_full_state"):
                state.optimizer.sync_full_state()
            save_checkpoint(os.path.join(train_config.checkpoint_dir, f"step_{state.step}.npz"), state.engine.store,
                            mesh, state.step, _checkpoint_meta(state, mesh, strategy))
    ms = mesh.max_over_ranks(timer.stop())
    steps = max(train_config.steps, 1)
    metrics = {"strategy": strategy, "world_size": mesh.world_size, "backend": state.engine.backend,
               "steps": train_config.steps, "device_ms_per_step_max_over_ranks": ms / steps,
               "tokens_per_step": tokens_per_step, "tokens_per_s": tokens_per_step * steps / (ms / 1e3) if ms else 0.0,
               "final_loss": log.loss_history[-1] if log.loss_history else None,
               "tp_seq_parallel": bool(getattr(state.engine, "seq_parallel", False) or hasattr(state.engine, "sp"))}
    if mesh.tp.size > 1:   # how far the tensor-parallel copies of the replicated leaves have drifted apart (0.0 = bit-equal)
        st = state.engine.store
        mine = st.flat_master[:st.rep_numel]
        ref = mine.clone()
        mesh.tp.broadcast(ref, 0)
        diff = (mine - ref).abs().max().reshape(1).float()
        mesh.tp.all_reduce(diff, op=torch.distributed.ReduceOp.MAX)
        metrics["tp_replica_max_abs_diff"] = float(diff.item())
    log.finish(metrics)
    if train_config.checkpoint_dir:
        if hasattr(state.optimizer, "sync_full_state"):
            state.optimizer.sync_full_state()
        save_checkpoint(os.path.join(train_config.checkpoint_dir, "final.npz"), state.engine.store, mesh, state.step,
                        _checkpoint_meta(state, mesh, strategy))
    return metrics


def train_dp_tp(train_config: TrainConfig, model_config: ModelConfig, opt_config: OptimConfig, parallel_strategy: str,
                num_devices: int, mesh: DeviceMesh | None = None) -> dict:
    """Data-parallel, tensor-parallel or hybrid DP×TP training on this rank's device."""
    dp, tp, _ = mesh_shape(train_config, num_devices)
    validate(model_config, train_config, num_devices)
    mesh = mesh or build_mesh(dp, tp, 1)
    global_batch = train_config.batch * (dp if train_config.global_batch_is_per_device else 1)
    local_batch = global_batch // dp
    T = model_config.max_seq_len
    local = LocalShape(model_config, tp_rank=mesh.tp.rank, tp_size=tp)
    chains = train_config.micro_chains if (tp == 1 and local_batch % max(train_config.micro_chains, 1) == 0) else 1
    state = build_state(train_config, model_config, opt_config, mesh, local_batch // chains, chains, global_batch * T,
                        local)
    train_step = create_train_step(parallel=parallel_strategy, cuda_graph=train_config.cuda_graph,
                                   tp_resync_every=train_config.tp_resync_every)
    data_iterator = get_batch_iterator(batch_size=global_batch, seq_len=T + 1, seed=train_config.seed, rank=mesh.dp.rank,
                                       world=dp, synthetic=train_config.synthetic_data,
                                       skip_batches=state.batches_consumed)
    return _run_loop(train_config, model_config, state, train_step, data_iterator, mesh, global_batch * T,
                     parallel_strategy)


def train_pp(train_config: TrainConfig, model_config: ModelConfig, opt_config: OptimConfig, parallel_strategy: str,
             num_devices: int, mesh: DeviceMesh | None = None) -> dict:
    assert parallel_strategy == "pp", "Should be called if `PP`"
    validate(model_config, train_config, num_devices)
    mesh = mesh or build_mesh(1, 1, num_devices)
    S, stage = num_devices, mesh.pp.rank
    head_cost = train_config.pp_head_cost if train_config.pp_layer_split == "balanced" else 0.0
    layers = stage_layers(model_config.n_layers, S, train_config.pp_reference_truncate, head_cost)[stage]
    local = LocalShape(model_config, layer_start=layers.start, layer_end=layers.stop, has_embed=(stage == 0),
                       has_head=(stage == S - 1))
    M = train_config.pp_microbatches
    T = model_config.max_seq_len
    clip_scale = 1.0 / S if (train_config.pp_grad_scale_by_stages and train_config.pp_clip == "stage") else 1.0
    state = build_state(train_config, model_config, opt_config, mesh, train_config.batch // M, M,
                        train_config.batch * T, local, clip_scope=train_config.pp_clip, clip_scale=clip_scale,
                        pp_stage_keys=(train_config.pp_init == "reference"))
    train_step = create_pp_train_step(num_microbatches=M, num_stages=S, schedule=train_config.pp_schedule,
                                      cuda_graph=train_config.pp_cuda_graph and train_config.cuda_graph)
    data_iterator = get_batch_iterator(batch_size=train_config.batch, seq_len=T + 1, seed=train_config.seed,
                                       synthetic=train_config.synthetic_data, skip_batches=state.batches_consumed)
    return _run_loop(train_config, model_config, state, train_step, data_iterator, mesh, train_config.batch * T,
                     parallel_strategy)
