#!/usr/bin/env python
"""tools/config5.py -- BASELINE.json configs[4]: one end-to-end EM-Adapt training step.

DeepLab-LargeFOV on VGG16 exactly as the reference wires it (deeplab.py:38-107,130-147: 13 3x3 convolutions,
conv5_x at rate 2, pools 1-3 of stride 2 and 4-5 of stride 1, fc6 4x4 at rate 4 -> 4096, fc7 1x1, fc8 1x1 -> 21,
dropout 0.5), random-initialised, on a synthetic 321x321 VOC-shaped batch (fc8 is 41x41), momentum SGD with the
reference's learning-rate multipliers 1/2/10/20 and weight decay (deeplab.py:180-208).  The backbone stays on
the framework's convolutions (cuDNN through stock PyTorch): it is out of the scope of this repository.  What is
measured is the step with the loss done two ways:

  device : EMAdaptLoss -- ground-truth shrink + label sets, E-step, labels, softmax cross-entropy and its
           gradient all queued on the GPU by libweaklabel_cuda.so (one autograd Function);
  pyfunc : what deeplab.py:109-124 does -- fc8 leaves the device, the E-step runs behind the host-pointer
           `e_step` ABI (here: this library's drop-in, so the comparison isolates the ROUND TRIP, not the CPU
           algorithm), argmax on the host, labels return to the device, F.cross_entropy there.

    python tools/config5.py [--batch 6] [--steps 10] [--warmup 3] [--loss device|pyfunc|both]
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/config5.py   # DDP over NCCL

Prints one JSON line on rank 0 (milliseconds per step, max over ranks, CUDA events)."""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


class DeepLabLargeFOV(nn.Module):
    def __init__(self, classes=21):
        super().__init__()
        cfg = [(3, 64, 1), (64, 64, 1), "p2", (64, 128, 1), (128, 128, 1), "p2", (128, 256, 1), (256, 256, 1), (256, 256, 1), "p2",
               (256, 512, 1), (512, 512, 1), (512, 512, 1), "p1", (512, 512, 2), (512, 512, 2), (512, 512, 2), "p1"]
        layers = []
        for c in cfg:
            if isinstance(c, str):
                layers.append(nn.MaxPool2d(3, stride=int(c[1]), padding=1))
            else:
                layers += [nn.Conv2d(c[0], c[1], 3, padding=c[2], dilation=c[2]), nn.ReLU(inplace=True)]
        self.features = nn.Sequential(*layers)
        # fc6: 4x4 kernel at rate 4 with SAME padding (deeplab.py:91, 144): effective extent 13 -> pad 6 before, 6 after
        self.fc6 = nn.Conv2d(512, 4096, 4, dilation=4)
        self.fc7 = nn.Conv2d(4096, 4096, 1)
        self.fc8 = nn.Conv2d(4096, classes, 1)
        self.drop = nn.Dropout(0.5)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.normal_(m.weight, std=0.01)
                nn.init.zeros_(m.bias)

    def forward(self, x):
        x = self.features(x)
        x = F.pad(x, (6, 6, 6, 6))
        x = self.drop(F.relu(self.fc6(x)))
        x = self.drop(F.relu(self.fc7(x)))
        return self.fc8(x)                       # [B, 21, 41, 41] for 321x321 inputs


def param_groups(model, lr, wd):
    """deeplab.py:149-176, 191-198: weights x1, biases x2, fc8 weights x10, fc8 biases x20."""
    g = {1: [], 2: [], 10: [], 20: []}
    for name, p in model.named_parameters():
        fc8 = ".fc8." in "." + name or name.startswith("fc8.") or name.startswith("module.fc8.")
        g[(10 if fc8 else 1) * (2 if name.endswith("bias") else 1)].append(p)
    return [dict(params=v, lr=lr * k, weight_decay=wd if k in (1, 10) else 0.0) for k, v in g.items() if v]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=6, help="images per GPU per step (deeplab.py:288)")
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--loss", default="both", choices=["device", "pyfunc", "both"])
    ap.add_argument("--tf32", action="store_true")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "config5 needs CUDA devices (libweaklabel_cuda has no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = bool(args.tf32)
    torch.backends.cudnn.benchmark = True
    import torch.distributed as dist
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")      # stdout carries one JSON line: library chatter (NCCL) goes to stderr
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import cases
    import em_adapt_tensorflow_b200 as wl
    from em_adapt_tensorflow_b200.loss import EMAdaptLoss

    torch.manual_seed(1 + rank)
    model = DeepLabLargeFOV().to(dev)       # NCHW: cuDNN's fp32 channels-last path is ~14x slower for the rate-4 fc6
    net = nn.parallel.DistributedDataParallel(model, device_ids=[local]) if world > 1 else model
    opt = torch.optim.SGD(param_groups(net, 1e-3, 1e-5), lr=1e-3, momentum=0.9)
    B = args.batch
    rng = np.random.default_rng(7 + rank)
    sets = cases.voc_label_sets(B, rng)
    gt_h = np.stack([cases.paint_label_map(321, 321, s, rng) for s in sets]).astype(np.uint8)
    gt = torch.from_numpy(gt_h).to(dev)
    x = torch.randn(B, 3, 321, 321, device=dev)
    crit = EMAdaptLoss()                          # bg 0.4 fg 0.2 num_iter 5 suppress margin 1e-5 (deeplab.py:181)

    def loss_device(fc8):
        loss, _ = crit(fc8.permute(0, 2, 3, 1).contiguous(), gt)
        return loss

    def loss_pyfunc(fc8):
        # deeplab.py:110-123: shrink, py_func(estep) on host arrays, argmax; labels carry no gradient
        with torch.no_grad():
            f = fc8.detach().permute(0, 2, 3, 1).contiguous().cpu().numpy()
            l = F.interpolate(gt[:, None].float(), size=f.shape[1:3], mode="nearest")[:, 0].cpu().numpy()
            shifted = wl.estep(f, l, True, 5, 1e-5, 0.4, 0.2)
            labels = torch.from_numpy(np.argmax(shifted, axis=3)).to(dev)
        return F.cross_entropy(fc8.permute(0, 2, 3, 1).reshape(-1, fc8.shape[1]), labels.reshape(-1))

    def step(loss_fn):
        opt.zero_grad(set_to_none=True)
        loss = loss_fn(net(x))
        loss.backward()
        opt.step()
        return loss

    out = {}
    for mode in (["device", "pyfunc"] if args.loss == "both" else [args.loss]):
        fn = loss_device if mode == "device" else loss_pyfunc
        for _ in range(args.warmup):
            step(fn)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(args.steps):
            last = step(fn)
        b.record()
        torch.cuda.synchronize()
        t = torch.tensor([a.elapsed_time(b) / args.steps], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out[mode] = dict(ms_per_step=float(t.item()), loss=float(last.item()))
        # the loss part alone (forward of the criterion on a fixed fc8), same two ways
        with torch.no_grad():
            fc8 = net(x)
        fc8 = fc8.detach().requires_grad_(True)
        for _ in range(3):
            fn(fc8).backward()
        torch.cuda.synchronize()
        a.record()
        for _ in range(5 * args.steps):
            fn(fc8).backward()
        b.record()
        torch.cuda.synchronize()
        out[mode]["loss_fwd_bwd_ms"] = a.elapsed_time(b) / (5 * args.steps)
    if rank == 0:
        line = dict(config="EM-Adapt step: DeepLab-LargeFOV VGG16 random-init, synthetic 321x321, batch %d per GPU, fp32%s, "
                           "momentum SGD, DDP/NCCL x%d" % (B, " (tf32)" if args.tf32 else "", world),
                    n_gpus=world, steps=args.steps, warmup=args.warmup, **out)
        if "device" in out and "pyfunc" in out:
            line["step_speedup_device_over_pyfunc"] = out["pyfunc"]["ms_per_step"] / out["device"]["ms_per_step"]
            line["loss_speedup_device_over_pyfunc"] = out["pyfunc"]["loss_fwd_bwd_ms"] / out["device"]["loss_fwd_bwd_ms"]
        print(json.dumps(line), file=real_stdout, flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
