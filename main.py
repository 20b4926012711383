"""CLI entry point (component C1; reference: main.py:10-61).

    python main.py --train_config_path configs/train_config_dp.yaml

Loads the same three YAML files as the reference (``configs/model_config.yaml`` and ``configs/optim_config.yaml``
are cwd-relative like in the reference, main.py:14,28), injects ``vocab_size = len(tokenizer)`` (50258) and the
strategy into ``ModelConfig``, prints the banner and dispatches ``pp`` → ``train_pp``, ``dp`` / ``tp`` (and the
hybrid ``dp_tp``) → ``train_dp_tp``; any other strategy raises ``ValueError``.

Process model: the reference drives every GPU from one process; here it is one process per GPU.  When started
plainly, ``main.py`` spawns one worker per visible GPU (``--num_devices`` to override; CPU runs use gloo); when
started by ``torchrun`` it uses the RANK / WORLD_SIZE environment as is.
"""
from __future__ import annotations

import os
import sys

import click
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from distributed_training_compare_jax_b200.config.schema import (  # noqa: E402
    load_model_config, load_optim_config, load_train_config)
from distributed_training_compare_jax_b200.data.fineweb_edu import get_tokenizer  # noqa: E402
from distributed_training_compare_jax_b200.parallel.mesh import shutdown  # noqa: E402
from distributed_training_compare_jax_b200.train.train import train_dp_tp, train_pp  # noqa: E402


def _count_devices(requested: int) -> int:
    if requested > 0:
        return requested
    if "WORLD_SIZE" in os.environ:
        return int(os.environ["WORLD_SIZE"])
    return torch.cuda.device_count() if torch.cuda.is_available() else 1


def run(train_config_path: str, model_config_path: str, optim_config_path: str, overrides: dict | None = None) -> dict:
    overrides = overrides or {}
    tokenizer = get_tokenizer()
    train_config = load_train_config(train_config_path, **overrides)
    # Model config with parallel strategy (reference main.py:25-26)
    model_config = load_model_config(model_config_path, vocab_size=len(tokenizer), parallel=train_config.parallel)
    opt_config = load_optim_config(optim_config_path)
    parallel_strategy = train_config.parallel
    num_devices = int(os.environ.get("WORLD_SIZE", "1"))
    if int(os.environ.get("RANK", "0")) == 0:
        print(f"Running `{parallel_strategy}` on {num_devices} devices.", flush=True)
    failed = False
    try:
        if parallel_strategy == "pp":
            return train_pp(train_config=train_config, model_config=model_config, opt_config=opt_config,
                            parallel_strategy=parallel_strategy, num_devices=num_devices)
        elif parallel_strategy in ("dp", "tp", "dp_tp"):
            return train_dp_tp(train_config=train_config, model_config=model_config, opt_config=opt_config,
                               parallel_strategy=parallel_strategy, num_devices=num_devices)
        else:
            raise ValueError(f"Unsupported strategy `{parallel_strategy}`")
    except BaseException:
        failed = True      # leave without the collective teardown: the other ranks are not going to take part in it
        raise
    finally:
        shutdown(abort=failed)


def _worker(rank: int, world: int, port: int, args: tuple) -> None:
    os.environ.update({"RANK": str(rank), "LOCAL_RANK": str(rank), "WORLD_SIZE": str(world),
                       "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port)})
    run(*args)


def _parse_overrides(items: tuple[str, ...]) -> dict:
    import yaml
    out = {}
    for it in items:
        k, _, v = it.partition("=")
        out[k] = yaml.safe_load(v)
    return out


@click.command()
@click.option("--train_config_path", default="configs/train_config_dp.yaml")
@click.option("--model_config_path", default="configs/model_config.yaml")
@click.option("--optim_config_path", default="configs/optim_config.yaml")
@click.option("--num_devices", default=0, help="processes to spawn (0 = one per visible GPU, or 1 on CPU)")
@click.option("--set", "overrides", multiple=True, help="train-config override, e.g. --set steps=100")
def main(train_config_path: str, model_config_path: str, optim_config_path: str, num_devices: int,
         overrides: tuple[str, ...]):
    args = (train_config_path, model_config_path, optim_config_path, _parse_overrides(overrides))
    if "RANK" in os.environ and "WORLD_SIZE" in os.environ:  # launched by torchrun
        run(*args)
        return
    world = _count_devices(num_devices)
    if world == 1:
        os.environ.update({"RANK": "0", "WORLD_SIZE": "1", "LOCAL_RANK": "0"})
        run(*args)
        return
    import socket
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(world, port, args), nprocs=world, join=True)


if __name__ == "__main__":
    main()
