"""attention: tcgen05 kernel vs the CUDA-core kernel vs torch SDPA on the shapes of C2 / C3 / C4.
usage: python scripts/bench_attention.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from ditsep_b200 import ops


def timeit(fn, reps=20):
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


print(f"{'shape':28s} {'GFLOP':>7s} | {'tc us':>8s} {'TF/s':>6s} | {'cuda-core us':>12s} | {'torch sdpa us':>13s}")
for (name, B, S, C) in [("C2 16x16 B=32", 32, 256, 128), ("C3 16x80 B=8", 8, 1280, 128), ("C4 16x20 B=32 C=256", 32, 320, 256),
                        ("C1 16x16 B=1", 1, 256, 128)]:
    qkv = torch.randn(B, 1, S, 3 * C, device="cuda").to(torch.bfloat16)
    q, k, v = qkv[..., :C], qkv[..., C:2 * C], qkv[..., 2 * C:]
    out = torch.empty(B, 1, S, C, device="cuda", dtype=torch.bfloat16)
    os.environ["DITSEP_ATTN_TC"] = "1"
    t_tc = timeit(lambda: ops.attention(q, k, v, out))
    os.environ["DITSEP_ATTN_TC"] = "0"
    t_cc = timeit(lambda: ops.attention(q, k, v, out))
    qq, kk, vv = (a.reshape(B, 1, S, C).contiguous() for a in (q, k, v))
    t_sd = timeit(lambda: torch.nn.functional.scaled_dot_product_attention(qq, kk, vv))
    fl = 4.0 * B * S * S * C
    print(f"{name:28s} {fl / 1e9:7.2f} | {t_tc:8.1f} {fl / t_tc / 1e6:6.1f} | {t_cc:12.1f} | {t_sd:13.1f}")
