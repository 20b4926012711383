"""Configuration schema (component C2; reference: config/schema.py:7-38).

The three frozen dataclasses carry exactly the reference's fields, so the five reference YAML files load
unchanged and an unknown YAML key raises ``TypeError`` (plain ``**dict`` construction, as in the reference
``main.py:22,26,30``).  Framework extensions are additional *defaulted* fields that a reference YAML never
has to mention.
"""
from __future__ import annotations

from dataclasses import dataclass, fields
from typing import Any

import yaml

PyTree = Any

PARALLEL_STRATEGIES = ("dp", "tp", "pp", "dp_tp", "none")


@dataclass(frozen=True)
class ModelConfig:
    vocab_size: int
    d_model: int
    n_layers: int
    n_heads: int
    d_ff: int
    max_seq_len: int
    dropout: float
    parallel: str
    # --- extensions (defaulted) ---
    dtype: str = "bf16"          # compute dtype of the native path (master weights / optimizer state are fp32)
    layernorm_eps: float = 1e-6  # flax.linen.LayerNorm default

    @property
    def head_dim(self) -> int:
        return self.d_model // self.n_heads


@dataclass(frozen=True)
class OptimConfig:
    lr: float
    weight_decay: float
    grad_clip: float
    # --- extensions: optax.adamw defaults ---
    b1: float = 0.9
    b2: float = 0.999
    eps: float = 1e-8


@dataclass(frozen=True)
class TrainConfig:
    seed: int
    parallel: str
    # batching
    batch: int
    # steps
    steps: int
    log_every: int
    # output
    output_dir: str
    # pipeline parallelism
    pp_microbatches: int = 1
    # --- extensions (defaulted; never required) ---
    warmup_steps: int = 5             # un-timed, un-logged TRAINING steps before "step 0" (reference: train/train.py:63-70)
    dp_size: int = 0                  # hybrid mesh (parallel: dp_tp): 0 = infer from world size
    tp_size: int = 0
    pp_schedule: str = "gpipe"        # "gpipe" (reference schedule) | "1f1b"
    pp_clip: str = "stage"            # "stage" = per-stage local norm (reference quirk Q3) | "global"
    pp_grad_scale_by_stages: bool = True   # reference quirk Q4: psum inside the differentiated fn scales grads by S
    pp_layer_split: str = "even"      # "even" (reference: same number of layers per stage) | "balanced" (fewer layers on the last stage, which also runs lm_head + cross-entropy)
    pp_head_cost: float = 3.5         # cost of final LayerNorm + lm_head + CE in transformer-layer units (pp_layer_split: balanced)
    pp_reference_truncate: bool = False    # reference quirk Q2: n_layers // n_stages layers per stage, rest dropped
    pp_init: str = "reference"        # "reference" = per-stage keys (Q6) | "shared" = same weights as DP/TP
    synthetic_data: bool = True       # offline token stream (no HF hub access on the box)
    sync_every_step: bool = True      # blocking loss fetch each step like the reference (train/train.py:82-83)
    backend: str = "auto"             # "auto" (native kernels on CUDA, torch oracle on CPU) | "native" | "torch"
    cuda_graph: bool = True           # replay the captured step graph on the native path
    dp_overlap: bool = True           # fused DP: reduce-scatter each layer's gradient bucket during the backward pass
    tp_seq_parallel: bool = True      # TP: row-sharded residual stream, GEMM-epilogue reduce-scatter + LayerNorm all-gather (parallel/tp_sp.py; +22 % at TP2 vs GEMM + NCCL all-reduce, falls back when batch % tp != 0)
    pp_cuda_graph: bool = True        # PP: one CUDA graph per (forward | backward, microbatch) of a stage (2 GPUs: 7.49 -> 5.59 ms/step)
    tp_resync_every: int = 128        # TP: every N steps re-broadcast the REPLICATED leaves (embeddings, LayerNorm, row biases) and their Adam moments from tp-rank 0: their gradients are accumulated with fp32 atomics whose order differs per rank, so the copies would otherwise drift apart bit by bit (the reference's GSPMD keeps them identical); 0 = never
    micro_chains: int = 1             # DP/TP: split the local batch into this many concurrently executed chains
    fused_comm: bool = True           # fused compute+collective kernels over symmetric memory (vs plain NCCL)
    global_batch_is_per_device: bool = False  # weak scaling: `batch` is per data-parallel rank
    checkpoint_dir: str = ""          # save Flax-layout state here ("" = off)
    checkpoint_every: int = 0
    resume: str = ""                  # path of a checkpoint to resume from
    num_devices: int = 0              # 0 = all visible devices (CUDA) / 1 (CPU)


def _load_yaml(path: str) -> dict:
    with open(path) as f:
        d = yaml.safe_load(f)
    if not isinstance(d, dict):
        raise ValueError(f"{path}: expected a YAML mapping")
    return d


def load_train_config(path: str, **overrides: Any) -> TrainConfig:
    d = _load_yaml(path)
    d.update(overrides)
    return TrainConfig(**d)


def load_model_config(path: str, vocab_size: int, parallel: str, **overrides: Any) -> ModelConfig:
    d = _load_yaml(path)
    d["vocab_size"] = vocab_size
    d["parallel"] = parallel
    d.update(overrides)
    return ModelConfig(**d)


def load_optim_config(path: str, **overrides: Any) -> OptimConfig:
    d = _load_yaml(path)
    d.update(overrides)
    return OptimConfig(**d)


def validate(model: ModelConfig, train: TrainConfig, world_size: int) -> None:
    """Divisibility checks the reference leaves to JAX (or silently ignores)."""
    if train.parallel not in ("dp", "tp", "pp", "dp_tp"):
        raise ValueError(f"Unsupported strategy `{train.parallel}`")
    if model.d_model % model.n_heads:
        raise ValueError("d_model must be divisible by n_heads")
    dp, tp, pp = mesh_shape(train, world_size)
    if tp > 1:
        if model.n_heads % tp or model.d_ff % tp:
            raise ValueError(f"tensor parallel size {tp} must divide n_heads={model.n_heads} and d_ff={model.d_ff}")
    if dp > 1 and not train.global_batch_is_per_device and train.batch % dp:
        raise ValueError(f"batch {train.batch} is not divisible by the data-parallel size {dp}")
    if pp > 1:
        if train.batch % train.pp_microbatches:
            raise ValueError(f"batch {train.batch} is not divisible by pp_microbatches {train.pp_microbatches}")
        if not train.pp_reference_truncate and model.n_layers < pp:
            raise ValueError(f"{pp} pipeline stages need at least {pp} layers")


def mesh_shape(train: TrainConfig, world_size: int) -> tuple[int, int, int]:
    """(dp, tp, pp) sizes for a strategy on `world_size` devices."""
    s = train.parallel
    if s == "dp":
        return world_size, 1, 1
    if s == "tp":
        return 1, world_size, 1
    if s == "pp":
        return 1, 1, world_size
    if s == "dp_tp":
        tp = train.tp_size or (4 if world_size % 4 == 0 else world_size)
        dp = train.dp_size or world_size // tp
        if dp * tp != world_size:
            raise ValueError(f"dp_size*tp_size = {dp}*{tp} != world size {world_size}")
        return dp, tp, 1
    raise ValueError(f"Unsupported strategy `{s}`")


def field_names(cls) -> list[str]:
    return [f.name for f in fields(cls)]
