"""Comparison baseline: stock PyTorch (DDP + NCCL + cuDNN/cuBLAS) on the identical model.

The reference's own TensorFlow MirroredStrategy run is unavailable on this box (TensorFlow is not
installable offline), so this is the measured stand-in: torchvision ResNet-50 /
the README MNIST MLP, bf16 autocast, channels_last, ``torch.optim.Adam(fused=True)``, DDP over NCCL.
Nothing from cloud_b200 is used on this path.

    python bench/baseline_ddp.py --model resnet50 --batch 64 [--graph]
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 bench/baseline_ddp.py ...
"""
import argparse
import json
import os
import time

import torch
import torch.distributed as dist
import torch.nn as nn


def build_model(name, classes):
    if name == "resnet50":
        import torchvision
        return torchvision.models.resnet50(num_classes=classes)
    if name == "resnet18":
        import torchvision
        return torchvision.models.resnet18(num_classes=classes)
    if name == "mlp":
        return nn.Sequential(nn.Linear(784, 512), nn.ReLU(), nn.Dropout(0.2), nn.Linear(512, 10))
    raise ValueError(name)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resnet50")
    ap.add_argument("--batch", type=int, default=64, help="per-GPU batch")
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--classes", type=int, default=1000)
    ap.add_argument("--graph", action="store_true", help="capture the step in a CUDA graph")
    ap.add_argument("--fp32", action="store_true")
    args = ap.parse_args()

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    torch.manual_seed(0)
    model = build_model(args.model, args.classes).to(dev)
    if args.model != "mlp":
        model = model.to(memory_format=torch.channels_last)
        x = torch.randn(args.batch, 3, 224, 224, device=dev).to(memory_format=torch.channels_last)
        y = torch.randint(0, args.classes, (args.batch,), device=dev)
    else:
        x = torch.randn(args.batch, 784, device=dev)
        y = torch.randint(0, 10, (args.batch,), device=dev)
    opt = torch.optim.Adam(model.parameters(), lr=1e-3, fused=True, capturable=args.graph)
    net = model
    if world > 1:
        net = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local],
                                                        gradient_as_bucket_view=True)
    loss_fn = nn.CrossEntropyLoss()

    def step():
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=not args.fp32):
            out = net(x)
            loss = loss_fn(out.float(), y)
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()
        return loss

    graph = None
    if args.graph:
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(11 if world > 1 else 3):
                step()
        torch.cuda.current_stream().wait_stream(s)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            static_loss = step()

    def run():
        if graph is not None:
            graph.replay()
            return static_loss
        return step()

    for _ in range(args.warmup):
        run()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    e0.record()
    for _ in range(args.steps):
        run()
    e1.record()
    torch.cuda.synchronize()
    wall = time.time() - t0
    ms = e0.elapsed_time(e1) / args.steps
    t = torch.tensor([ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = t.item()
    if rank == 0:
        unit = "img/s" if args.model != "mlp" else "samples/s"
        print(json.dumps({"impl": "torch_ddp_nccl", "model": args.model, "n_gpus": world,
                          "batch_per_gpu": args.batch, "graph": bool(args.graph),
                          "ms_per_step": ms, "value": world * args.batch / (ms * 1e-3),
                          "unit": unit, "wall_ms_per_step": wall / args.steps * 1e3}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
