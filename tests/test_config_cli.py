"""T3: configuration / CLI contract (reference: main.py:10-61, config/schema.py:7-38, configs/*.yaml)."""
import dataclasses
import os

import pytest
import yaml

from distributed_training_compare_jax_b200.config import schema
from distributed_training_compare_jax_b200.data.fineweb_edu import VOCAB_SIZE, get_tokenizer

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CFG = os.path.join(ROOT, "configs")


def test_reference_yaml_files_load_unchanged():
    mc = schema.load_model_config(os.path.join(CFG, "model_config.yaml"), vocab_size=VOCAB_SIZE, parallel="dp")
    assert (mc.d_model, mc.n_layers, mc.n_heads, mc.d_ff, mc.max_seq_len, mc.dropout) == (512, 12, 16, 2048, 512, 0.1)
    assert mc.vocab_size == 50258 and mc.head_dim == 32
    oc = schema.load_optim_config(os.path.join(CFG, "optim_config.yaml"))
    assert (oc.lr, oc.weight_decay, oc.grad_clip) == (3e-4, 0.1, 1.0)
    assert (oc.b1, oc.b2, oc.eps) == (0.9, 0.999, 1e-8)           # optax.adamw defaults
    for s in ("dp", "tp", "pp"):
        tc = schema.load_train_config(os.path.join(CFG, f"train_config_{s}.yaml"))
        assert (tc.parallel, tc.batch, tc.steps, tc.log_every, tc.seed) == (s, 8, 5000, 50, 0)
        assert tc.output_dir == f"outputs/{s}"
        assert tc.pp_microbatches == (2 if s == "pp" else 1)
        assert tc.warmup_steps == 5


def test_reference_fields_are_a_prefix_of_ours_and_frozen():
    assert schema.field_names(schema.ModelConfig)[:8] == ["vocab_size", "d_model", "n_layers", "n_heads", "d_ff",
                                                          "max_seq_len", "dropout", "parallel"]
    assert schema.field_names(schema.OptimConfig)[:3] == ["lr", "weight_decay", "grad_clip"]
    assert schema.field_names(schema.TrainConfig)[:7] == ["seed", "parallel", "batch", "steps", "log_every",
                                                          "output_dir", "pp_microbatches"]
    tc = schema.load_train_config(os.path.join(CFG, "train_config_dp.yaml"))
    with pytest.raises(dataclasses.FrozenInstanceError):
        tc.batch = 3


def test_unknown_yaml_key_raises_typeerror(tmp_path):
    d = yaml.safe_load(open(os.path.join(CFG, "train_config_dp.yaml")))
    d["not_a_field"] = 1
    p = tmp_path / "bad.yaml"
    p.write_text(yaml.safe_dump(d))
    with pytest.raises(TypeError):
        schema.load_train_config(str(p))


def test_unsupported_strategy_and_divisibility_checks():
    mc = schema.load_model_config(os.path.join(CFG, "model_config.yaml"), vocab_size=VOCAB_SIZE, parallel="dp")
    tc = schema.load_train_config(os.path.join(CFG, "train_config_dp.yaml"))
    with pytest.raises(ValueError, match="Unsupported strategy"):
        schema.validate(mc, dataclasses.replace(tc, parallel="zz"), 2)
    with pytest.raises(ValueError):
        schema.validate(mc, dataclasses.replace(tc, batch=6), 4)            # batch % dp
    with pytest.raises(ValueError):
        schema.validate(mc, dataclasses.replace(tc, parallel="tp"), 3)      # heads % tp
    schema.validate(mc, dataclasses.replace(tc, parallel="tp"), 8)          # vocab 50258 % 8 != 0 is fine (padded)
    schema.validate(mc, dataclasses.replace(tc, parallel="pp", pp_microbatches=2), 8)
    assert schema.mesh_shape(dataclasses.replace(tc, parallel="dp_tp", dp_size=2, tp_size=4), 8) == (2, 4, 1)


def test_tokenizer_len_is_gpt2_plus_pad():
    assert len(get_tokenizer()) == 50258


def test_main_cli_default_and_dispatch(monkeypatch):
    import main as cli
    from click.testing import CliRunner
    seen = {}

    def fake_run(train_config_path, model_config_path, optim_config_path, overrides):
        seen.update(path=train_config_path, model=model_config_path, optim=optim_config_path, ov=overrides)

    monkeypatch.setattr(cli, "run", fake_run)
    monkeypatch.setenv("RANK", "0")
    monkeypatch.setenv("WORLD_SIZE", "1")
    r = CliRunner().invoke(cli.main, [])
    assert r.exit_code == 0, r.output
    assert seen["path"] == "configs/train_config_dp.yaml"           # reference default (main.py:11)
    assert seen["model"] == "configs/model_config.yaml" and seen["optim"] == "configs/optim_config.yaml"
    r = CliRunner().invoke(cli.main, ["--train_config_path", "configs/train_config_pp.yaml", "--set", "steps=3"])
    assert r.exit_code == 0 and seen["path"].endswith("pp.yaml") and seen["ov"] == {"steps": 3}
