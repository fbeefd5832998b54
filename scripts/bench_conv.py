"""conv_tc (this repo's tcgen05 implicit GEMM) beside torch + cuDNN on every convolution shape of one NCSN++ score
step (SURVEY.md 8a census), same GPU, same run.
usage: python scripts/bench_conv.py [B] [reps] [shape-filter]
Columns: shape, count per score step, algorithmic GFLOP per launch, then us / TFLOP/s for
  ours   ditsep conv_tc, bf16 NHWC, epilogue as in the network (bias_b + GroupNorm statistics, or residual)
  cudnn  F.conv2d(bf16, channels_last) + bias, torch.backends.cudnn.benchmark = True
  tf32   F.conv2d(fp32, channels_last) + bias with TF32 allowed (the reference's default arithmetic on a GPU)
cuDNN is a LIBRARY BASELINE ("the kernel to beat"); nothing of it is on the product path.
L2 policy: the 256x256 / 128x128 tensors exceed the 126 MB L2 at batch 32; smaller levels are L2-resident for both
arms alike."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
import torch.nn.functional as F
from ditsep_b200 import ops

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
filt = sys.argv[3] if len(sys.argv) > 3 else ""
# (H, Cin, Cout, k, count per score step, epilogue) -- C1/C2 census, W = H
CENSUS = [
    (256, 64, 64, 3, 9, "b"), (256, 128, 64, 3, 3, "b"), (256, 64, 6, 3, 1, ""), (256, 128, 64, 1, 3, ""),
    (128, 64, 64, 3, 9, "b"), (128, 128, 128, 3, 2, "b"), (128, 192, 64, 3, 1, "b"), (128, 128, 64, 3, 2, "b"),
    (128, 64, 6, 3, 1, ""),
    (64, 64, 64, 3, 2, "b"), (64, 64, 128, 3, 1, "b"), (64, 128, 128, 3, 8, "r"), (64, 256, 128, 3, 2, "b"),
    (64, 192, 128, 3, 1, "b"), (64, 128, 6, 3, 1, ""), (64, 256, 128, 1, 2, ""),
    (32, 128, 128, 3, 11, "r"), (32, 256, 128, 3, 3, "b"), (16, 128, 128, 3, 11, "r"), (16, 256, 128, 3, 3, "b"),
    (16, 128, 384, 1, 3, ""), (8, 128, 128, 3, 11, "r"), (8, 256, 128, 3, 3, "b"), (4, 128, 128, 3, 13, "r"),
    (4, 256, 128, 3, 3, "b"),
]
dev = "cuda"
torch.backends.cudnn.benchmark = True
g = torch.Generator().manual_seed(0)


def timeit(fn):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps


print(f"# batch {B}, {reps} back-to-back launches per number, CUDA events; {torch.cuda.get_device_name(0)}; "
      f"cudnn {torch.backends.cudnn.version()}")
print(f"{'shape':24s} {'n':>3s} {'GFLOP':>8s} | {'ours us':>9s} {'TF/s':>7s} | {'cudnn us':>9s} {'TF/s':>7s} | "
      f"{'tf32 us':>9s} {'TF/s':>7s} | ours/cudnn")
tot = {"ours": 0.0, "cudnn": 0.0, "tf32": 0.0}
for (H, Cin, Cout, k, cnt, fl) in CENSUS:
    W = H
    name = f"{H}x{W}:{Cin}->{Cout}:k{k}:{fl}"
    if filt and filt not in name:
        continue
    x = torch.randn(B, H, W, Cin, generator=g).to(dev).to(torch.bfloat16)
    cp = (Cout + 15) // 16 * 16
    wf = torch.randn(Cout, Cin, k, k, generator=g) / (k * k * Cin) ** 0.5
    w = torch.zeros(cp, k * k, Cin)
    w[:Cout] = wf.permute(0, 2, 3, 1).reshape(Cout, k * k, Cin)
    w = w.to(dev).to(torch.bfloat16)
    bias = torch.randn(Cout, generator=g).to(dev)
    odt = torch.bfloat16 if Cout % 16 == 0 else torch.float32
    out = torch.empty(B, H, W, Cout, device=dev, dtype=odt)
    res = torch.randn(B, H, W, Cout, generator=g).to(dev).to(odt) if "r" in fl else None
    bb = torch.randn(B, Cout, generator=g).to(dev) if "b" in fl else None
    plan = ops.ConvTcPlan(x, w, Cout, k)
    stats = ops.new_sums(B, Cout, dev) if (fl and plan.tile) else None
    kw = dict(bias_b=bb, residual=res, out_scale=0.7071, stats=stats)
    us = timeit(lambda: plan.run(bias, out, **kw))
    # library arms: NCHW-shaped tensors in channels_last memory format (= NHWC), bias in the call
    xl = x.permute(0, 3, 1, 2)
    wl = wf.to(dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    bl = bias.to(torch.bfloat16)
    us_c = timeit(lambda: F.conv2d(xl, wl, bl, padding=k // 2))
    torch.backends.cudnn.allow_tf32 = True
    torch.backends.cuda.matmul.allow_tf32 = True
    x32 = x.float().permute(0, 3, 1, 2)
    w32 = wf.to(dev).contiguous(memory_format=torch.channels_last)
    us_t = timeit(lambda: F.conv2d(x32, w32, bias, padding=k // 2))
    flops = 2.0 * B * H * W * Cout * k * k * Cin
    tot["ours"] += cnt * us; tot["cudnn"] += cnt * us_c; tot["tf32"] += cnt * us_t
    print(f"{name:24s} {cnt:3d} {flops / 1e9:8.1f} | {us:9.1f} {flops / us / 1e6:7.1f} | {us_c:9.1f} "
          f"{flops / us_c / 1e6:7.1f} | {us_t:9.1f} {flops / us_t / 1e6:7.1f} | {us / us_c:5.2f}", flush=True)
    del x, x32, xl, out, res, plan
print(f"# sum over one score step (count-weighted): ours {tot['ours'] / 1e3:.2f} ms, cudnn bf16 {tot['cudnn'] / 1e3:.2f} ms, "
      f"cudnn tf32 {tot['tf32'] / 1e3:.2f} ms")
