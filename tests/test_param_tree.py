"""T2: parameter-tree compatibility with the reference's Flax state and the sharding rules of
parallel/sharding.py:17-62 — names, shapes, init statistics, shard → gather round trips, checkpoint I/O."""
import math
import os

import numpy as np
import pytest
import torch

from distributed_training_compare_jax_b200.config.schema import ModelConfig
from distributed_training_compare_jax_b200.models import param_tree as pt
from distributed_training_compare_jax_b200.parallel import sharding
from distributed_training_compare_jax_b200.parallel.mesh import single_device_mesh
from distributed_training_compare_jax_b200.utils.checkpoint import load_checkpoint, save_checkpoint

REF = ModelConfig(vocab_size=50258, d_model=512, n_layers=12, n_heads=16, d_ff=2048, max_seq_len=512, dropout=0.1,
                  parallel="dp")
SMALL = ModelConfig(vocab_size=1003, d_model=64, n_layers=4, n_heads=4, d_ff=128, max_seq_len=32, dropout=0.0,
                    parallel="dp")


def test_flax_tree_names_shapes_and_param_count():
    s = pt.flax_param_shapes(REF)
    blk = "ScanScanBlock_0/TransformerBlock_0"
    assert s["wte/embedding"] == (50258, 512) and s["wpe/embedding"] == (512, 512)
    assert s[f"{blk}/LayerNorm_0/scale"] == (12, 512) and s[f"{blk}/LayerNorm_1/bias"] == (12, 512)
    for p in ("q_proj", "k_proj", "v_proj", "out_proj"):
        assert s[f"{blk}/CausalSelfAttention_0/{p}/kernel"] == (12, 512, 512)
        assert s[f"{blk}/CausalSelfAttention_0/{p}/bias"] == (12, 512)
    assert s[f"{blk}/MLP_0/fc1/kernel"] == (12, 512, 2048) and s[f"{blk}/MLP_0/fc2/kernel"] == (12, 2048, 512)
    assert s["LayerNorm_0/scale"] == (512,) and s["lm_head/kernel"] == (512, 50258) and s["lm_head/bias"] == (50258,)
    assert len(s) == 22
    assert pt.count_params(REF) == 89_606_226


def test_local_shapes_vocab_padding_and_tp_split():
    lo = pt.LocalShape(REF)
    assert lo.vocab_local == 50304 and lo.vocab_valid == 50258
    v = [pt.LocalShape(REF, tp_rank=r, tp_size=8) for r in range(8)]
    assert all(x.vocab_local == 6400 for x in v) and sum(x.vocab_valid for x in v) == 50258
    assert v[7].vocab_valid == 50258 - 7 * 6400
    assert v[0].d_local == 64 and v[0].heads_local == 2 and v[0].ff_local == 256


def test_init_statistics_follow_flax_defaults():
    st = pt.ParamStore(pt.LocalShape(SMALL), "cpu", with_compute_copy=False)
    pt.init_params(st, seed=0)
    D, F = SMALL.d_model, SMALL.d_ff
    assert abs(st.p["wte"].std().item() - 1 / math.sqrt(D)) < 0.01
    assert torch.all(st.p["ln1_g"] == 1) and torch.all(st.p["bqkv"] == 0) and torch.all(st.p["lnf_b"] == 0)
    w1 = st.p["w1"]
    assert abs(w1.std().item() - math.sqrt(1 / D)) < 0.01                     # lecun_normal: std = sqrt(1/fan_in)
    assert w1.abs().max().item() <= 2 * math.sqrt(1 / D) / 0.87962566 + 1e-6   # truncated at 2 sigma
    assert abs(st.p["w2"].std().item() - math.sqrt(1 / F)) < 0.01
    assert torch.all(st.p["wlm"][:, SMALL.vocab_size:] == 0)                   # vocab pad columns


@pytest.mark.parametrize("tp", [1, 2, 4])
def test_tp_shards_reassemble_to_the_unsharded_tree(tp):
    full = pt.ParamStore(pt.LocalShape(SMALL), "cpu", with_compute_copy=False)
    pt.init_params(full, seed=3)
    want = pt.export_flax_tree([(full.local, full.p)])
    parts = []
    for r in range(tp):
        st = pt.ParamStore(pt.LocalShape(SMALL, tp_rank=r, tp_size=tp), "cpu", with_compute_copy=False)
        pt.init_params(st, seed=3)
        parts.append((st.local, st.p))
    got = pt.export_flax_tree(parts)
    assert set(got) == set(pt.flax_param_shapes(SMALL))
    for k, shape in pt.flax_param_shapes(SMALL).items():
        assert got[k].shape == shape, k
        np.testing.assert_array_equal(got[k], want[k], err_msg=k)


def test_pp_stages_reassemble_and_import_roundtrip():
    full = pt.ParamStore(pt.LocalShape(SMALL), "cpu", with_compute_copy=False)
    pt.init_params(full, seed=5)
    tree = pt.export_flax_tree([(full.local, full.p)])
    parts = []
    for s, (a, b) in enumerate([(0, 1), (1, 3), (3, 4)]):
        lo = pt.LocalShape(SMALL, layer_start=a, layer_end=b, has_embed=(s == 0), has_head=(s == 2))
        st = pt.ParamStore(lo, "cpu", with_compute_copy=False)
        pt.import_flax_tree(tree, lo, st.p)
        parts.append((lo, st.p))
    got = pt.export_flax_tree(parts)
    for k in tree:
        np.testing.assert_array_equal(got[k], tree[k], err_msg=k)


def test_sharding_rule_table_matches_reference():
    shapes = pt.flax_param_shapes(REF)
    tree = {k: np.empty(v, dtype=np.float32) if np.prod(v) < 10_000 else np.broadcast_to(np.float32(0), v)
            for k, v in shapes.items()}
    spec = sharding.spec_tree(tree, "tp")
    blk = "ScanScanBlock_0/TransformerBlock_0"
    assert spec[f"{blk}/CausalSelfAttention_0/q_proj/kernel"] == (None, None, "data")
    assert spec[f"{blk}/CausalSelfAttention_0/out_proj/kernel"] == (None, "data", None)
    assert spec[f"{blk}/MLP_0/fc1/kernel"] == (None, None, "data") and spec[f"{blk}/MLP_0/fc2/kernel"] == (None, "data", None)
    assert spec["lm_head/kernel"] == (None, "data") and spec["lm_head/bias"] == ("data",)
    for k in ("wte/embedding", "wpe/embedding", f"{blk}/LayerNorm_0/scale", f"{blk}/MLP_0/fc1/bias",
              f"{blk}/CausalSelfAttention_0/q_proj/bias", "LayerNorm_0/bias"):
        assert spec[k] == (None,), k
    assert all(v == (None,) for v in sharding.spec_tree(tree, "dp").values())
    x = np.arange(24, dtype=np.float32).reshape(2, 3, 4)
    np.testing.assert_array_equal(sharding.shard_leaf(x, (None, None, "data"), 1, 2), x[:, :, 2:])
    with pytest.raises(ValueError):
        sharding.shard_leaf(np.zeros((512, 50258), np.float32), (None, "data"), 0, 4)   # reference limitation (Q1)


def test_get_sharded_params_returns_store_and_spec():
    full = pt.ParamStore(pt.LocalShape(SMALL), "cpu", with_compute_copy=False)
    pt.init_params(full, seed=1)
    tree = pt.export_flax_tree([(full.local, full.p)])
    st = pt.ParamStore(pt.LocalShape(SMALL, tp_rank=1, tp_size=2), "cpu", with_compute_copy=False)
    sharded, spec = sharding.get_sharded_params(tree, None, "tp", store=st)
    assert sharded is st and spec["lm_head/bias"] == ("data",)
    Dl = SMALL.d_model // 2
    blk = "ScanScanBlock_0/TransformerBlock_0/CausalSelfAttention_0"
    np.testing.assert_array_equal(st.p["wqkv"][:, :, Dl:2 * Dl].numpy(), tree[f"{blk}/k_proj/kernel"][:, :, Dl:])
    np.testing.assert_array_equal(st.p["wo"].numpy(), tree[f"{blk}/out_proj/kernel"][:, Dl:, :])


def test_checkpoint_is_flax_layout_and_resumes(tmp_path):
    st = pt.ParamStore(pt.LocalShape(SMALL), "cpu", with_compute_copy=False)
    pt.init_params(st, seed=2)
    st.flat_m.normal_()
    st.flat_v.uniform_()
    st.step.fill_(17)
    path = os.path.join(tmp_path, "ck.npz")
    save_checkpoint(path, st, single_device_mesh("cpu"), step=17, meta={"strategy": "dp"})
    z = np.load(path)
    assert z["params/lm_head/kernel"].shape == (SMALL.d_model, SMALL.vocab_size)          # unpadded
    assert z["opt_state/mu/ScanScanBlock_0/TransformerBlock_0/MLP_0/fc1/kernel"].shape == (4, 64, 128)
    assert int(z["opt_state/count"]) == 17
    # resume under a different layout: tensor-parallel rank 1 of 2
    st2 = pt.ParamStore(pt.LocalShape(SMALL, tp_rank=1, tp_size=2), "cpu", with_compute_copy=False)
    meta = load_checkpoint(path, st2)
    assert meta["step"] == 17 and int(st2.step.item()) == 17
    Fl = SMALL.d_ff // 2
    assert torch.equal(st2.p["w1"], st.p["w1"][:, :, Fl:])
    assert torch.equal(st2.views_of(st2.flat_v)["w2"], st.views_of(st.flat_v)["w2"][:, Fl:, :])


def _small_flax_tree():
    st = pt.ParamStore(pt.LocalShape(SMALL), "cpu", with_compute_copy=False)
    pt.init_params(st, seed=3)
    return pt.export_flax_tree([(st.local, st.p)])


def test_pp_state_layout_matches_the_reference_and_round_trips():
    """reference train/train.py:155-167: {"embed","stage","head"} per stage, stacked on a leading S axis."""
    tree = _small_flax_tree()
    blk = "ScanScanBlock_0/TransformerBlock_0"
    pp = pt.export_pp_tree(tree, n_stages=2)
    assert isinstance(pp, dict)
    assert pp["embed/wte/embedding"].shape == (2, 1003, 64) and pp["head/lm_head/kernel"].shape == (2, 64, 1003)
    assert pp[f"stage/{blk}/MLP_0/fc1/kernel"].shape == (2, 2, 64, 128)          # (S, layers_per_stage, D, F)
    assert {k.split("/")[0] for k in pp} == {"embed", "stage", "head"}
    assert np.array_equal(pp[f"stage/{blk}/MLP_0/fc1/kernel"][1, 0], tree[f"{blk}/MLP_0/fc1/kernel"][2])
    assert np.array_equal(pp["embed/wte/embedding"][0], pp["embed/wte/embedding"][1])   # replicated copies
    back = pt.import_pp_tree(pp)
    assert set(back) == set(tree) and all(np.array_equal(back[k], tree[k]) for k in tree)
    # the reference's floor split drops the remainder (12 layers on 8 stages -> 8 layers): 4 layers on 3 stages -> 3
    pp3 = pt.export_pp_tree(tree, n_stages=3)
    assert pp3[f"stage/{blk}/LayerNorm_0/scale"].shape == (3, 1, 64)
    # the framework's uneven split keeps every layer and cannot be stacked: a list of per-stage trees
    from distributed_training_compare_jax_b200.parallel.pp import stage_layers
    uneven = pt.export_pp_tree(tree, 3, layer_ranges=stage_layers(4, 3))
    assert isinstance(uneven, list) and [d[f"stage/{blk}/LayerNorm_0/scale"].shape[0] for d in uneven] == [2, 1, 1]
    back = pt.import_pp_tree(uneven)
    assert all(np.array_equal(back[k], tree[k]) for k in tree)
    # dead copies: unused embed / head replicas shrink by weight decay only
    dead = pt.export_pp_tree(tree, 2, dead_copy_scale=0.5)
    assert np.allclose(dead["embed/wte/embedding"][1], 0.5 * tree["wte/embedding"])
    assert np.array_equal(dead["embed/wte/embedding"][0], tree["wte/embedding"])
    assert np.allclose(dead["head/lm_head/kernel"][0], 0.5 * tree["lm_head/kernel"])
    assert np.array_equal(dead["head/lm_head/kernel"][1], tree["lm_head/kernel"])
