"""One score-network evaluation at batch B (default 8) in bf16: the command profiled with ncu.
usage: python scripts/profile_step.py [B] [n_calls] [precision]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

from common import build_model  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
n = int(sys.argv[2]) if len(sys.argv) > 2 else 2
prec = sys.argv[3] if len(sys.argv) > 3 else "bf16"
model, _, _ = build_model(prec, "cuda")
g = torch.Generator().manual_seed(0)
xt = torch.randn(B, 2, 32000, generator=g).cuda()
mix = torch.randn(B, 1, 32000, generator=g).cuda()
t = torch.full((B,), 0.5).cuda()
with torch.no_grad():
    for _ in range(n):
        out = model(xt, t, mix)
torch.cuda.synchronize()
print("ok", float(out.norm()))
