"""Host-side argument marshalling of the grouped launches (ctypes arrays, shape validation) — no GPU needed."""
import pytest
import torch

from distributed_training_compare_jax_b200.ops import functional as Fn


def test_wgrad_group_args_marshals_pointers_strides_and_shapes():
    xs = [torch.zeros(64, 16, dtype=torch.bfloat16), torch.zeros(64, 8, dtype=torch.bfloat16)]
    dys = [torch.zeros(64, 24, dtype=torch.bfloat16), torch.zeros(64, 32, dtype=torch.bfloat16)[:, :16]]   # strided view
    outs = [torch.zeros(16, 24), torch.zeros(8, 16)]
    a = Fn.wgrad_group_args(xs, dys, outs)
    assert a.count == 2 and a.K == 64
    assert list(a.m) == [16, 8] and list(a.n) == [24, 16]
    assert list(a.lda) == [16, 8] and list(a.ldb) == [24, 32] and list(a.ldo) == [24, 16]
    assert list(a.a) == [t.data_ptr() for t in xs] and list(a.out) == [t.data_ptr() for t in outs]


def test_wgrad_group_args_rejects_inconsistent_problems():
    x, dy = torch.zeros(64, 16, dtype=torch.bfloat16), torch.zeros(64, 24, dtype=torch.bfloat16)
    with pytest.raises(ValueError):   # output shape is not [M, N]
        Fn.wgrad_group_args([x], [dy], [torch.zeros(24, 16)])
    with pytest.raises(ValueError):   # output must be the fp32 accumulator
        Fn.wgrad_group_args([x], [dy], [torch.zeros(16, 24, dtype=torch.bfloat16)])
    with pytest.raises(ValueError):   # different reduction lengths cannot share a launch
        Fn.wgrad_group_args([x, torch.zeros(32, 16, dtype=torch.bfloat16)], [dy, torch.zeros(32, 24, dtype=torch.bfloat16)],
                            [torch.zeros(16, 24), torch.zeros(16, 24)])
    with pytest.raises(ValueError):   # transposed (non unit inner stride) operand
        Fn.wgrad_group_args([x.t()], [dy], [torch.zeros(64, 24)])


def test_colsum_group_args():
    xs = [torch.zeros(10, 16, dtype=torch.bfloat16), torch.zeros(20, 8, dtype=torch.bfloat16)]
    outs = [torch.zeros(16), torch.zeros(8)]
    a = Fn.colsum_group_args(xs, outs)
    assert a.count == 2 and list(a.m) == [10, 20] and list(a.n) == [16, 8] and list(a.ld) == [16, 8]
