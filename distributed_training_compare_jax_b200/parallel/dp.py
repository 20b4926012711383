"""Data-parallel gradient reduction (components R10/R11, collectives X1/X2; reference: the all-reduce GSPMD
inserts for ``P(None)`` parameters with a ``P("data", None)`` batch — parallel/sharding.py:25-27,
train/create_train_step.py:34-40).

``make_grad_reducer`` returns the callable the optimizer invokes to make ``flat_grad`` hold the DP-summed
gradient.  Baseline: one NCCL all-reduce over the flat gradient buffer.  The bucketed, backward-overlapped
variant hooks the engine's ``on_*_grads_ready`` callbacks and reduces each bucket (head, layer L-1 … 0, embed)
on a side stream as soon as its gradients are final.
"""
from __future__ import annotations

import torch


class BucketedAllReduce:
    """Per-layer gradient buckets reduced on a communication stream while the backward pass continues."""

    def __init__(self, engine, mesh):
        self.engine, self.mesh = engine, mesh
        self.store = engine.store
        self.stream = torch.cuda.Stream(device=engine.device)
        self.done = torch.cuda.Event()
        engine.on_layer_grads_ready = self._layer_ready
        engine.on_head_grads_ready = self._head_ready
        st = self.store
        L = engine.local.L
        # contiguous element ranges per bucket inside the flat buffer: the stacked (L, ...) tensors are not
        # contiguous per layer, so a bucket is the list of per-layer slices of every stacked tensor
        self.layer_slices = []
        for l in range(L):
            sl = []
            for name, (off, shape) in st.offsets.items():
                if len(shape) >= 2 and shape[0] == L and name not in ("wte", "wpe", "wlm"):
                    per = 1
                    for s in shape[1:]:
                        per *= s
                    sl.append((off + l * per, per))
            self.layer_slices.append(sl)
        self.head_slices = [(st.offsets[n][0], int(torch.tensor(st.offsets[n][1]).prod())) for n in
                            ("lnf_g", "lnf_b", "wlm", "blm") if n in st.offsets]
        self.embed_slices = [(st.offsets[n][0], int(torch.tensor(st.offsets[n][1]).prod())) for n in ("wte", "wpe")
                             if n in st.offsets]

    def _reduce(self, slices) -> None:
        cur = torch.cuda.current_stream()
        self.stream.wait_stream(cur)
        with torch.cuda.stream(self.stream):
            for off, n in slices:
                self.mesh.dp.all_reduce(self.store.flat_grad[off:off + n])

    def _layer_ready(self, l: int) -> None:
        self._reduce(self.layer_slices[l])

    def _head_ready(self) -> None:
        self._reduce(self.head_slices)

    def __call__(self) -> None:
        """Called by the optimizer after backward: reduce the embedding bucket, then join the side stream."""
        self._reduce(self.embed_slices)
        torch.cuda.current_stream().wait_stream(self.stream)


def make_grad_reducer(engine, mesh):
    return BucketedAllReduce(engine, mesh)
