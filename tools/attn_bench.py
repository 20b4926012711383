"""Micro-benchmark of the flash-attention kernels on the model's shapes (device time from a captured graph of 12 launches
over distinct buffers, CUDA events; causal FLOPs = 4·B·H·T²·hd / 2 forward, 2.5× that backward).
Usage: python tools/attn_bench.py   (DTC_ATTN_FWD_SERIAL=1 / DTC_ATTN_BWD_V1=1 select the older kernels for A/B)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_training_compare_jax_b200.ops import functional as Fn  # noqa: E402

bf = torch.bfloat16


def timeit(fn, n):
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(n):
            fn(i)
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


def main():
    for (B, H, T, hd) in [(8, 16, 512, 32), (4, 16, 512, 32), (1, 16, 512, 32), (8, 16, 1024, 64)]:
        n = 12
        qkvs = [torch.randn(B * T, 3 * H * hd, device="cuda").to(bf) for _ in range(n)]
        dos = [torch.randn(B * T, H * hd, device="cuda").to(bf) for _ in range(n)]
        outs = [Fn.flash_attn_fwd(q, B, H, T, hd) for q in qkvs]
        o_buf = [torch.empty_like(o[0]) for o in outs]
        l_buf = [torch.empty_like(o[1]) for o in outs]
        dq = [torch.empty_like(q) for q in qkvs]
        dqa = [torch.zeros(B * T, H * hd, device="cuda") for _ in range(n)]
        dl = [torch.empty(B, H, T, device="cuda") for _ in range(n)]
        for i in range(3):
            Fn.flash_attn_fwd(qkvs[i], B, H, T, hd, out=o_buf[i], lse2=l_buf[i])
            Fn.flash_attn_bwd(qkvs[i], outs[i][0], dos[i], outs[i][1], B, H, T, hd, dqkv=dq[i], dq_accum=dqa[i], delta=dl[i])
        torch.cuda.synchronize()
        tf = timeit(lambda i: Fn.flash_attn_fwd(qkvs[i], B, H, T, hd, out=o_buf[i], lse2=l_buf[i]), n)
        tb = timeit(lambda i: Fn.flash_attn_bwd(qkvs[i], outs[i][0], dos[i], outs[i][1], B, H, T, hd, dqkv=dq[i],
                                                dq_accum=dqa[i], delta=dl[i]), n)
        fl = 4.0 * B * H * T * T * hd / 2
        print(f"attn B{B} H{H} T{T} hd{hd}: fwd {tf:7.1f} us ({fl / tf / 1e6:6.1f} TFLOP/s causal)   "
              f"bwd (zero + prep + main + cast) {tb:7.1f} us ({2.5 * fl / tb / 1e6:6.1f} TFLOP/s)", flush=True)


if __name__ == "__main__":
    main()
