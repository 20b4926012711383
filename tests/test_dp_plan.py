"""Partition plan of the fused data-parallel step (parallel/dp.py): buckets follow the backward order, every gradient element is
owned by exactly one rank, owned sub-ranges respect the 8-element alignment the kernels require."""
import numpy as np
import pytest

from distributed_training_compare_jax_b200.config.schema import ModelConfig
from distributed_training_compare_jax_b200.models.param_tree import LocalShape, ParamStore
from distributed_training_compare_jax_b200.parallel.dp import plan_buckets

REF = ModelConfig(vocab_size=50258, d_model=512, n_layers=12, n_heads=16, d_ff=2048, max_seq_len=512, dropout=0.1,
                  parallel="dp")


@pytest.mark.parametrize("dp", [2, 3, 8])
def test_every_element_owned_exactly_once(dp):
    st = ParamStore(LocalShape(REF), "meta", with_compute_copy=False, with_optimizer_state=False)
    cover = np.zeros(st.numel, dtype=np.int8)
    plans = [plan_buckets(st.offsets, 12, r, dp) for r in range(dp)]
    for plan, index in plans:
        assert len(plan) == 14 and index["head"] == 0 and index[11] == 1 and index[0] == 12 and index["embed"] == 13
        assert all(len(b["owned"]) <= 16 for b in plan)            # one multi-range launch per bucket
        for b in plan:
            for (off, k), (foff, fn) in zip(b["owned"], b["full"]):
                assert off % 8 == 0 and k % 8 == 0 and foff <= off and off + k <= foff + fn
                cover[off:off + k] += 1
    real = np.zeros(st.numel, dtype=bool)
    for name, (off, shape) in st.offsets.items():
        real[off:off + int(np.prod(shape))] = True
    assert np.all(cover[real] == 1)        # each parameter element reduced / updated by exactly one rank
    assert np.all(cover[~real] == 0)       # alignment padding between tensors belongs to nobody


def test_layer_bucket_is_the_layer_slice_of_every_stacked_tensor():
    st = ParamStore(LocalShape(REF), "meta", with_compute_copy=False, with_optimizer_state=False)
    plan, index = plan_buckets(st.offsets, 12, 0, 2)
    b = plan[index[5]]
    names = [n for n in st.offsets if n not in ("wte", "wpe", "wlm", "blm", "lnf_g", "lnf_b")]
    assert len(b["full"]) == len(names) == 12
    for (off, k), name in zip(b["full"], names):
        base, shape = st.offsets[name]
        per = int(np.prod(shape[1:]))
        assert (off, k) == (base + 5 * per, per)
