"""Stand-in "library kernels only" baseline: the same GPT, data shapes and optimizer written in plain
PyTorch — cuBLAS GEMMs (nn.Linear), SDPA flash attention, F.layer_norm / F.gelu / F.cross_entropy, fused
torch.optim.AdamW, clip_grad_norm_, NCCL all-reduce through DistributedDataParallel — bf16 autocast with fp32
master weights.  It shares NO code with the product path (none of this repo's kernels, engine or models) and
exists only to be measured on the same box, because the reference's own JAX build cannot be installed offline.

    python baseline/torch_baseline.py [--steps 50] [--warmup 5] [--graph]
``--graph`` (single GPU): the whole step — forward, backward, clip, fused AdamW — is captured into one CUDA graph and
replayed, which removes the host launch overhead that bounds the eager variant on this box; it is the fairer bar for the
kernels (same library kernels, no launch gaps), the eager number is the fairer bar for an out-of-the-box framework.
    torchrun --nproc-per-node N baseline/torch_baseline.py     (data parallel, batch 8 per rank)
Prints one JSON line with device-timed (CUDA events, max over ranks) tokens/s.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F


class Block(nn.Module):
    def __init__(self, d, h, f):
        super().__init__()
        self.h = h
        self.ln1, self.ln2 = nn.LayerNorm(d, eps=1e-6), nn.LayerNorm(d, eps=1e-6)
        self.qkv, self.out = nn.Linear(d, 3 * d), nn.Linear(d, d)
        self.fc1, self.fc2 = nn.Linear(d, f), nn.Linear(f, d)

    def forward(self, x):
        B, T, D = x.shape
        q, k, v = self.qkv(self.ln1(x)).view(B, T, 3, self.h, D // self.h).permute(2, 0, 3, 1, 4)
        a = F.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(B, T, D)
        x = x + self.out(a)
        return x + self.fc2(F.gelu(self.fc1(self.ln2(x)), approximate="tanh"))


class GPT(nn.Module):
    def __init__(self, V=50258, d=512, L=12, h=16, f=2048, T=512, p=0.1):
        super().__init__()
        self.wte, self.wpe, self.drop = nn.Embedding(V, d), nn.Embedding(T, d), nn.Dropout(p)
        self.blocks = nn.ModuleList(Block(d, h, f) for _ in range(L))
        self.lnf, self.head = nn.LayerNorm(d, eps=1e-6), nn.Linear(d, V)

    def forward(self, x, y):
        h = self.drop(self.wte(x) + self.wpe(torch.arange(x.shape[1], device=x.device))[None])
        for b in self.blocks:
            h = b(h)
        logits = self.head(self.lnf(h))
        return F.cross_entropy(logits.float().view(-1, logits.shape[-1]), y.reshape(-1))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--graph", action="store_true", help="capture the whole training step into one CUDA graph (with DDP: the "
                    "PyTorch recipe — DDP built and warmed up for 11 iterations on a side stream, NCCL async error "
                    "handling off; falls back to eager launches if the capture fails)")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    use_graph = args.graph
    if world > 1:
        if use_graph:
            os.environ["TORCH_NCCL_ASYNC_ERROR_HANDLING"] = "0"
        dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(0)
    model = GPT().to(dev)
    if world > 1:
        ddp_stream = torch.cuda.Stream()
        ddp_stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(ddp_stream):   # (graph recipe: DDP must be constructed on a side stream)
            net = nn.parallel.DistributedDataParallel(model, device_ids=[dev.index])
        torch.cuda.current_stream().wait_stream(ddp_stream)
    else:
        net = model
    opt = torch.optim.AdamW(model.parameters(), lr=3e-4, weight_decay=0.1, eps=1e-8, fused=True, capturable=use_graph)
    T = 512
    g = torch.Generator(device=dev).manual_seed(rank)
    static_tok = torch.zeros(args.batch, T + 1, device=dev, dtype=torch.int64)
    graph, static_loss = None, None

    def body():
        with torch.autocast("cuda", dtype=torch.bfloat16):
            loss = net(static_tok[:, :-1], static_tok[:, 1:])
        loss.backward()
        torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
        opt.step()
        return loss

    def step():
        if graph is not None:
            static_tok.copy_(torch.randint(0, 50257, (args.batch, T + 1), device=dev, generator=g))
            graph.replay()
            return static_loss
        tok = torch.randint(0, 50257, (args.batch, T + 1), device=dev, generator=g)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            loss = net(tok[:, :-1], tok[:, 1:])
        loss.backward()
        torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
        opt.step()
        opt.zero_grad(set_to_none=True)
        return loss

    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    if use_graph:   # whole-step capture (PyTorch "whole network capture" recipe): warm up on a side stream, then capture
        static_tok.copy_(torch.randint(0, 50257, (args.batch, T + 1), device=dev, generator=g))
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(11 if world > 1 else 3):
                opt.zero_grad(set_to_none=True)
                body()
        torch.cuda.current_stream().wait_stream(side)
        opt.zero_grad(set_to_none=True)
        try:
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                static_loss = body()
            for _ in range(3):
                step()
            torch.cuda.synchronize()
        except Exception as e:   # report the eager number rather than nothing
            graph, static_loss = None, None
            if rank == 0:
                print(f"[torch_baseline] CUDA-graph capture failed ({type(e).__name__}: {str(e)[:200]}); timing eager launches",
                      file=sys.stderr, flush=True)
            torch.cuda.synchronize()
            opt.zero_grad(set_to_none=True)
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        loss = step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = ms.item()
    if rank == 0:
        print(json.dumps({"impl": ("torch CUDA-graph" if graph is not None else "torch-eager")
                          + " stand-in baseline (cuBLAS + SDPA + fused AdamW + NCCL DDP)",
                          "metric": "training throughput, tokens/s (whole job), DP", "n_gpus": world,
                          "value": args.batch * world * T * args.steps / (ms / 1e3), "unit": "tokens/s",
                          "ms_per_step": ms / args.steps, "steps": args.steps, "dtype": "bf16 autocast",
                          "loss_last": float(loss)}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
