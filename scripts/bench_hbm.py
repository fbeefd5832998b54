"""Micro-benchmark of the HBM-bound kernels (GPU): achieved GB/s of ALGORITHMIC bytes against the measured
copy bandwidth in MEASURED_PEAKS.json.  usage: python scripts/bench_hbm.py [B] [reps]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from ditsep_b200 import ops
from ditsep_b200._lib import SdeParams

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    peak = 6650.0
dev = "cuda"
bf = torch.bfloat16


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


def report(name, us, byts):
    gbs = byts / us / 1e3
    print(f"{name:44s} {us:9.1f} us  {gbs:8.1f} GB/s  {gbs / peak:6.1%} of measured HBM peak ({peak:.0f} GB/s)", flush=True)


# GroupNorm apply + SiLU, statistics pass, FIR resampling (bf16 activations)
for (H, W, C) in [(256, 256, 64), (256, 256, 128), (128, 128, 128), (64, 64, 256)]:
    x = torch.randn(B, H, W, C, device=dev).to(bf)
    y = torch.empty_like(x)
    ss = torch.randn(B, 2, C, device=dev)
    n = x.numel() * 2
    report(f"gn_apply+silu {H}x{W}x{C}", timeit(lambda: ops.gn_apply(x, ss, True, y)), 2 * n)
    part = ops.new_sums(B, C, dev)
    report(f"gn_sums {H}x{W}x{C}", timeit(lambda: ops.gn_sums(x, part)), n)
for (H, W, C) in [(256, 256, 64), (128, 128, 128)]:
    x = torch.randn(B, H, W, C, device=dev).to(bf)
    ss = torch.randn(B, 2, C, device=dev)
    ox, oh = torch.empty(B, H // 2, W // 2, C, device=dev, dtype=bf), torch.empty(B, H // 2, W // 2, C, device=dev, dtype=bf)
    n = x.numel() * 2
    report(f"fir down dual {H}x{W}x{C}", timeit(lambda: ops.fir_resample(x, False, out_x=ox, scale_shift=ss, out_h=oh)), n + 2 * n // 4)
    xs = torch.randn(B, H // 2, W // 2, C, device=dev).to(bf)
    ux, uh = torch.empty(B, H, W, C, device=dev, dtype=bf), torch.empty(B, H, W, C, device=dev, dtype=bf)
    report(f"fir up dual {H // 2}x{W // 2}x{C}", timeit(lambda: ops.fir_resample(xs, True, out_x=ux, scale_shift=ss, out_h=uh)), n // 4 + 2 * n)
# spectral front / back end and the fused SDE updates (fp32 state)
T = 32000
xt, mix = torch.randn(B, 2, T, device=dev), torch.randn(B, 1, T, device=dev)
win = torch.hann_window(510, periodic=True, device=dev)
feat = torch.empty(B, 256, 256, 6, device=dev)
report("stft+compress (B,3,32000)->(B,256,256,6) f32", timeit(lambda: ops.stft_compress(xt, mix, win, feat, 0.5, 0.15, True)),
       3 * B * (T + 382) * 4 + feat.numel() * 4)
spec = torch.randn(B, 256, 256, 4, device=dev)
out = torch.empty(B, 2, T, device=dev)
report("expand+istft (B,256,256,4)->(B,2,32000)", timeit(lambda: ops.expand_istft(spec, win, out, 0.5, 0.15)),
       spec.numel() * 4 + out.numel() * 4)
p = SdeParams(0, 2, 2.0, 0.05, 0.5, 30, 0.5)
x = torch.randn(B, 2, T, device=dev)
sc = torch.randn(B, 2, T, device=dev)
t = torch.full((B,), 0.5, device=dev)
report("pc_corrector (Philox in registers)", timeit(lambda: ops.pc_corrector(p, x, sc, t, seed=1, draw=3)), 3 * x.numel() * 4)
report("pc_predictor (Philox in registers)", timeit(lambda: ops.pc_predictor(p, x, sc, t, mix, seed=1, draw=4)), 3 * x.numel() * 4)
