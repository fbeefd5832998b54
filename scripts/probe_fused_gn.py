"""Which conv shapes survive the fused input GroupNorm at full batch?  usage: python scripts/probe_fused_gn.py [B]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from ditsep_b200 import ops
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
dev = "cuda"
SHAPES = [(256, 64, 64, 3, 0), (256, 64, 64, 3, 64), (256, 128, 64, 3, 0), (256, 64, 64, 3, 128), (128, 192, 64, 3, 0),
          (128, 128, 128, 3, 0), (64, 256, 128, 3, 0), (64, 128, 128, 3, 128), (32, 128, 128, 3, 0), (16, 128, 128, 3, 0),
          (16, 128, 384, 1, 0), (256, 64, 6, 3, 0), (128, 128, 64, 3, 128)]
for (H, Cin, Cout, k, Cx) in SHAPES:
    x = torch.randn(B, H, H, Cin, device=dev).to(torch.bfloat16)
    cp = (Cout + 15) // 16 * 16
    w = (torch.randn(cp, k * k * Cin + Cx, device=dev) / (k * k * Cin) ** 0.5).to(torch.bfloat16)
    xs = torch.randn(B, H, H, Cx, device=dev).to(torch.bfloat16) if Cx else None
    ss = torch.stack([1 + 0.1 * torch.randn(B, Cin, device=dev), 0.1 * torch.randn(B, Cin, device=dev)], 1).contiguous()
    bias = torch.zeros(Cout, device=dev)
    out = torch.empty(B, H, H, Cout, device=dev, dtype=torch.bfloat16)
    try:
        plan = ops.ConvTcPlan(x, w, Cout, k, shortcut=xs)
    except ops.DitsepError as e:
        print(f"{H}x{H} {Cin}->{Cout} k{k} Cx={Cx}: no plan ({str(e)[:60]})"); continue
    for mode in (None, ss):
        torch.cuda.synchronize(); t0 = time.time()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        plan.run(bias, out, in_ss=mode); torch.cuda.synchronize()
        e0.record()
        for _ in range(5): plan.run(bias, out, in_ss=mode)
        e1.record(); torch.cuda.synchronize()
        print(f"{H}x{H} {Cin}->{Cout} k{k} Cx={Cx} {'fusedGN' if mode is not None else 'plain  '}: {e0.elapsed_time(e1) * 200:.1f} us", flush=True)
