"""Per-tile timeline of conv_ts_kernel's epilogue warp 0 and MMA warps (needs a build with DITSEP_NVCC_EXTRA=-DDITSEP_TS_TRACE).
usage: python scripts/trace_ts.py [H] [B] [Cx]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from ditsep_b200 import _lib, ops
H = int(sys.argv[1]) if len(sys.argv) > 1 else 256
B = int(sys.argv[2]) if len(sys.argv) > 2 else 32
Cx = int(sys.argv[3]) if len(sys.argv) > 3 else 0
dev = "cuda"
x = torch.randn(B, H, H, 64, device=dev).to(torch.bfloat16)
w = (torch.randn(64, 576 + Cx, device=dev) / 24).to(torch.bfloat16)
xs = torch.randn(B, H, H, Cx, device=dev).to(torch.bfloat16) if Cx else None
bias = torch.zeros(64, device=dev)
out = torch.empty(B, H, H, 64, device=dev, dtype=torch.bfloat16)
plan = ops.ConvTcPlan(x, w, 64, 3, shortcut=xs)
stats = ops.new_sums(B, 64, dev)
for _ in range(3):
    plan.run(bias, out, stats=stats)
torch.cuda.synchronize()
buf = torch.zeros(4, 96, 16, dtype=torch.int64, device=dev)
lib = _lib.load()
lib.ditsep_debug_conv_trace(buf.data_ptr())
plan.run(bias, out, stats=stats)
torch.cuda.synchronize()
lib.ditsep_debug_conv_trace(None)
t = buf.cpu()[0]
names = ["loop_top", "acc_full", "tmem_rd", "exch_done", "row0_done", "stored", "|", "acc_free", "B_landed", "issued", "tma"]
slots = [9, 3, 4, 7, 8, 5, None, 0, 1, 2, 6]
t0 = int(t[0, 6])
print("tile " + " ".join(n.rjust(9) for n in names) + "   epi_period  full->full")
prev = prevf = None
for i in list(range(4, 24)) + list(range(70, 96)):
    row = [(int(t[i, s]) - t0) if s is not None else 0 for s in slots]
    per = row[0] - prev if prev else 0
    pf = row[1] - prevf if prevf else 0
    prev, prevf = row[0], row[1]
    print(f"{i:4d} " + " ".join((str(v) if s is not None else "|").rjust(9) for v, s in zip(row, slots)) + f"   {per:9d} {pf:9d}")

import numpy as np
for cta in range(4):
    tt = buf.cpu()[cta]
    a0 = tt[:, 0].numpy().astype(np.int64)
    ok = a0 > 0
    idx = np.nonzero(ok)[0]
    d = np.diff(a0[idx])
    print(f"CTA {cta}: acc_free period per tile: tiles 5-30 mean {d[5:30].mean():.0f}, 30-60 mean {d[30:60].mean():.0f}, 60-95 mean {d[60:94].mean():.0f} cycles")

for cta in range(4):
    tt = buf.cpu()[cta]
    cyc = int(tt[0, 11] - tt[0, 10]); ns = int(tt[0, 13] - tt[0, 12])
    first_tma = int(tt[0, 6] - tt[0, 10]); first_free = int(tt[0, 0] - tt[0, 10])
    print(f"CTA {cta}: kernel body {cyc} cycles = {ns / 1e3:.1f} us -> SM clock {cyc / max(ns, 1):.3f} GHz; first TMA issue at {first_tma}, first MMA start at {first_free} cycles")
