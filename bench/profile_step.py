"""Eager (no CUDA graph) ResNet-50 training steps for kernel-level profiling under ncu:
every launch is visible with its own name.  ``--impl torch`` profiles the stock torchvision model
for comparison."""
import argparse
import os
import sys

os.environ.setdefault("TFC_CUDA_GRAPH", "0")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--impl", default="cloud_b200")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    if args.impl == "torch":
        import torchvision
        m = torchvision.models.resnet50().to(dev).to(memory_format=torch.channels_last)
        opt = torch.optim.Adam(m.parameters(), fused=True)
        x = torch.randn(args.batch, 3, 224, 224, device=dev).to(memory_format=torch.channels_last)
        y = torch.randint(0, 1000, (args.batch,), device=dev)
        for _ in range(args.steps):
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss = torch.nn.functional.cross_entropy(m(x).float(), y)
            opt.zero_grad(set_to_none=True)
            loss.backward()
            opt.step()
        torch.cuda.synchronize()
        return
    import bench as B
    import cloud_b200 as tfc
    tfc.distribute.experimental_set_strategy(tfc.distribute.OneDeviceStrategy("cuda:0"))
    model = B.build_model(tfc.keras, 1000)
    x = torch.randint(0, 256, (args.batch, 224, 224, 3), dtype=torch.uint8, device=dev)
    y = torch.randint(0, 1000, (args.batch,), device=dev)
    model._ensure_engine(x)
    model.train(True)
    xi = model._to_internal(x)
    for _ in range(args.steps):
        model._run_step(xi, y)
    torch.cuda.synchronize()


if __name__ == "__main__":
    main()
