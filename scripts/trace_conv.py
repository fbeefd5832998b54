"""Per-tile timeline of conv_tc_tile_kernel (clock64 stamps, ditsep_debug_conv_trace).
usage: python scripts/trace_conv.py [Cin] [Cout] [H] [B] [residual 0/1] [Cx] [f32: fp32 output/residual, no statistics]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from ditsep_b200 import _lib, ops
Cin = int(sys.argv[1]) if len(sys.argv) > 1 else 64
Cout = int(sys.argv[2]) if len(sys.argv) > 2 else 64
H = int(sys.argv[3]) if len(sys.argv) > 3 else 256
B = int(sys.argv[4]) if len(sys.argv) > 4 else 32
use_res = len(sys.argv) > 5 and sys.argv[5] == "1"
Cx = int(sys.argv[6]) if len(sys.argv) > 6 else 0   # fused shortcut channels (0 = none)
dev = "cuda"
x = torch.randn(B, H, H, Cin, device=dev).to(torch.bfloat16)
w = (torch.randn(Cout, 9 * Cin + Cx, device=dev) / (9 * Cin) ** 0.5).to(torch.bfloat16)
xs = torch.randn(B, H, H, Cx, device=dev).to(torch.bfloat16) if Cx else None
bias = torch.zeros(Cout, device=dev)
f32 = len(sys.argv) > 7 and sys.argv[7] == "f32"   # the pyramid output convolutions: fp32 output and residual, no statistics
odt = torch.float32 if f32 else torch.bfloat16
out = torch.empty(B, H, H, Cout, device=dev, dtype=odt)
res = torch.randn(B, H, H, Cout, device=dev).to(odt) if use_res else None
plan = ops.ConvTcPlan(x, w, Cout, 3, shortcut=xs)
stats = None if (f32 or not plan.tile) else ops.new_sums(B, Cout, dev)
for _ in range(3):
    plan.run(bias, out, residual=res, stats=stats)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    plan.run(bias, out, residual=res, stats=stats)
e1.record()
torch.cuda.synchronize()
print(f"untraced: {e0.elapsed_time(e1) * 100:.1f} us per launch")
buf = torch.zeros(4, 96, 16, dtype=torch.int64, device=dev)
lib = _lib.load()
lib.ditsep_debug_conv_trace(buf.data_ptr())
plan.run(bias, out, residual=res, stats=stats)
torch.cuda.synchronize()
lib.ditsep_debug_conv_trace(None)
t = buf.cpu()
names = ["acc_free", "A_landed", "mma_issued", "acc_full", "tmem_read", "stored", "tma_issue"]
for cta in range(2):
    tc = t[cta]
    t0 = int(tc[0, 6]) if int(tc[0, 6]) else int(tc[0, 0])
    print(f"--- CTA {cta}: cycles since its first TMA issue; {Cin}->{Cout} {H}x{H} B={B} res={use_res}")
    print("tile " + " ".join(n.rjust(10) for n in names) + "   mma(acc_full-acc_free)  epi(stored-acc_full)  period(acc_full)")
    prev = None
    for i in range(0, 40):
        row = [int(v) for v in tc[i, :7]]
        if row[3] == 0:
            break
        rel = [v - t0 if v else 0 for v in row]
        per = row[3] - prev if prev else 0
        prev = row[3]
        print(f"{i:4d} " + " ".join(str(v).rjust(10) for v in rel) + f"   {row[3]-row[0]:10d} {row[5]-row[3]:10d} {per:10d}")
    n = int((tc[:, 3] > 0).sum())
    if n > 10:
        d = (tc[5:n, 3] - tc[4:n - 1, 3]).double()
        print(f"steady-state period between accumulator-full events: mean {d.mean():.0f} cycles over {n-5} tiles; "
              f"mma window mean {(tc[4:n,3]-tc[4:n,0]).double().mean():.0f}; issue time mean {(tc[4:n,2]-tc[4:n,0]).double().mean():.0f}; "
              f"epilogue mean {(tc[4:n,5]-tc[4:n,3]).double().mean():.0f}; tmem-read mean {(tc[4:n,4]-tc[4:n,3]).double().mean():.0f}")
