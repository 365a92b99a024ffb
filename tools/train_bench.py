#!/usr/bin/env python
"""Data-parallel training throughput of the transformer part (BASELINE.json configs[3], SURVEY.md C4 shapes):
6-layer MSDeformAttn encoder + 6-layer decoder (d = 256, 800x1333 pyramid, S = 22 223, 300 queries),
batch 2 per GPU (cfgs/train.yaml:9), bf16 autocast, forward + backward + AdamW step per iteration.
One process per GPU; the only exchange step is the gradient all-reduce (NCCL over NVLink), which is outside
the op -- the reference wraps its model in DistributedDataParallel (src/train.py:141-144; --eager does the
same here).  Default: forward+backward and the optimizer step are CUDA graphs, with one all-reduce of a flat
gradient buffer between them.
samples/s is the sum over GPUs ("weak" scaling: per-GPU batch fixed).
The backbone, matcher and losses are the reference's and are not part of this number; the loss here is a
plain squared error on the decoder output so that every parameter of the stack receives a gradient.

    python tools/train_bench.py [--steps 20]
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/train_bench.py
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from _harness import FusedStack, Stack  # noqa: E402  (encoder/decoder stack around kinet_b200.MSDeformAttn)
from kinet_b200.workloads import PYRAMIDS  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--batch", type=int, default=2, help="samples per GPU (cfgs/train.yaml:9)")
    ap.add_argument("--fp32", action="store_true", help="no autocast (the reference's only mode)")
    ap.add_argument("--eager", action="store_true", help="DistributedDataParallel + eager launches instead of CUDA graphs")
    ap.add_argument("--layers", choices=("fused", "torch"), default="fused",
                    help="fused: kinet_b200.modules encoder/decoder classes; torch: the same layers as plain PyTorch ops")
    args = ap.parse_args()
    make_stack = FusedStack if args.layers == "fused" else Stack
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        torch.distributed.init_process_group("nccl", device_id=dev)
    levels = PYRAMIDS["detr_800x1333"]
    S = sum(h * w for h, w in levels)
    torch.manual_seed(0)                      # same initial weights on every rank, as DDP requires
    model = make_stack(levels, n_queries=300, d=256).to(dev).train()
    n_params = sum(p.numel() for p in model.parameters())
    g = torch.Generator(device=dev).manual_seed(100 + rank)          # a different batch on every rank
    batches = [(torch.randn(args.batch, S, 256, device=dev, generator=g),
                torch.randn(args.batch, S, 256, device=dev, generator=g) * 0.1,
                torch.randn(args.batch, 300, 256, device=dev, generator=g)) for _ in range(4)]

    def fwd_bwd(net, src, pos, target):
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=not args.fp32):
            out, _ = net(src, pos)
        loss = (out.float() - target).square().mean()
        loss.backward()
        return loss

    if args.eager:
        # stock recipe: DistributedDataParallel (bucketed all-reduce overlapped with the backward), eager launches
        net = model
        if world > 1:
            net = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local], gradient_as_bucket_view=True)
        opt = torch.optim.AdamW(model.parameters(), lr=1e-4, weight_decay=1e-4, fused=True)

        def step(i):
            opt.zero_grad(set_to_none=True)
            loss = fwd_bwd(net, *batches[i & 3])
            opt.step()
            return loss
    else:
        # B200 recipe: the step is two CUDA graphs (forward+backward into ONE flat fp32 gradient buffer; AdamW on ONE
        # flat parameter + one cast kernel refreshing the bf16 weight mirror) with a single NCCL all-reduce of the
        # gradient buffer between them -- no per-kernel host cost, no bucket bookkeeping, no per-layer weight casts
        from kinet_b200.training import GraphedTrainStep

        def loss_fn(src, pos, target):
            with torch.autocast("cuda", dtype=torch.bfloat16, enabled=not args.fp32):
                out, _ = model(src, pos)
            return (out.float() - target).square().mean()
        graphed = GraphedTrainStep(model, loss_fn,
                                   lambda p: torch.optim.AdamW(p, lr=1e-4, weight_decay=1e-4, fused=True, capturable=True),
                                   batches[0])

        def step(i):
            return graphed(*batches[i & 3])

    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize()
    if world > 1:
        torch.distributed.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(args.steps):
        loss = step(i)
    b.record()
    torch.cuda.synchronize()
    ms = torch.tensor([a.elapsed_time(b)], device=dev)
    losses = loss.detach().reshape(1).clone()
    if world > 1:
        torch.distributed.all_reduce(ms, op=torch.distributed.ReduceOp.MAX)
        torch.distributed.all_reduce(losses, op=torch.distributed.ReduceOp.SUM)
    if rank == 0:
        ms = float(ms)
        print(json.dumps({"metric": "training_transformer_samples_per_s",
                          "value": world * args.batch * args.steps / ms * 1e3,
                          "n_gpus": world, "ms_per_step": ms / args.steps, "steps": args.steps,
                          "batch_per_gpu": args.batch, "dtype": "f32" if args.fp32 else "bf16 autocast",
                          "layers": args.layers, "mode": "eager_ddp" if args.eager else "cuda_graphs", "scaling": "weak", "mean_final_loss": float(losses) / world,
                          "grad_allreduce_bytes_per_step": 4 * n_params if world > 1 else 0,
                          "config": "800x1333 pyramid, S=22223, d=256 (D=32), 6 enc + 6 dec layers, 300 queries, "
                                    "fwd+bwd+AdamW; gradient all-reduce over NCCL when n_gpus > 1"}))
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
