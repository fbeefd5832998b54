"""Micro-benchmark of ditsep_conv_tc on the score network's convolution shapes (GPU).
usage: python scripts/bench_conv.py [B] [reps] [shape-filter]   (DITSEP_CONV_MODE=tap|halo)
Each line: shape, us per launch (CUDA events over `reps` back-to-back launches), TFLOP/s, GB/s of
algorithmic traffic (read input once + write output once)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from ditsep_b200 import ops

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
filt = sys.argv[3] if len(sys.argv) > 3 else ""
SHAPES = [  # H, W, Cin, Cout, k, flags (r = residual+stats, b = bias_b+stats)
    (256, 256, 64, 64, 3, "b"), (256, 256, 64, 64, 3, "r"), (256, 256, 128, 64, 3, "b"), (256, 256, 128, 64, 1, ""),
    (256, 256, 64, 6, 3, ""), (128, 128, 64, 64, 3, "b"), (128, 128, 128, 128, 3, "b"), (128, 128, 192, 64, 3, "b"),
    (64, 64, 128, 128, 3, "r"), (64, 64, 256, 128, 3, "b"), (32, 32, 128, 128, 3, "r"), (16, 16, 128, 128, 3, "r"),
    (8, 8, 128, 128, 3, "r"), (4, 4, 128, 128, 3, "r"), (16, 16, 128, 384, 1, ""),
]
dev = "cuda"
g = torch.Generator().manual_seed(0)
for (H, W, Cin, Cout, k, fl) in SHAPES:
    name = f"{H}x{W}:{Cin}->{Cout}:k{k}:{fl}"
    if filt and filt not in name:
        continue
    x = torch.randn(B, H, W, Cin, generator=g).to(dev).to(torch.bfloat16)
    cp = (Cout + 15) // 16 * 16
    w = (torch.randn(cp, k * k, Cin, generator=g) / (k * k * Cin) ** 0.5).to(dev).to(torch.bfloat16)
    bias = torch.randn(Cout, generator=g).to(dev)
    out = torch.empty(B, H, W, Cout, device=dev, dtype=torch.bfloat16)
    res = torch.randn(B, H, W, Cout, generator=g).to(dev).to(torch.bfloat16) if "r" in fl else None
    bb = torch.randn(B, Cout, generator=g).to(dev) if "b" in fl else None
    stats = None
    plan = ops.ConvTcPlan(x, w, Cout, k)
    stats = torch.empty(plan.stats_floats(), device=dev) if (fl and plan.stats_floats()) else None
    kw = dict(bias_b=bb, residual=res, out_scale=0.7071, stats=stats)
    for _ in range(3):
        plan.run(bias, out, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        plan.run(bias, out, **kw)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / reps
    flops = 2.0 * B * H * W * Cout * k * k * Cin
    byts = B * H * W * (Cin + Cout * (2 if res is not None else 1)) * 2
    print(f"{name:28s} {us:9.1f} us  {flops / us / 1e6:8.1f} TFLOP/s  {byts / us / 1e3:8.1f} GB/s", flush=True)
