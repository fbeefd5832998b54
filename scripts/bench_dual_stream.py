"""Experiment: does running the score network as two half batches on two CUDA streams (the small, latency-bound levels of
one half overlapping the large levels of the other) beat one full batch?  Two model instances (own buffers / graphs).
    python scripts/bench_dual_stream.py [B]"""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from common import build_model, score_inputs  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
dev = torch.device("cuda")
full, _, _ = build_model("bf16", dev, N=30)
halves = [build_model("bf16", dev, N=30)[0] for _ in range(2)]
xt, t, mix = (a.to(dev) for a in score_inputs(B, 32000))
hx = [(xt[i * B // 2:(i + 1) * B // 2].contiguous(), t[i * B // 2:(i + 1) * B // 2].contiguous(),
       mix[i * B // 2:(i + 1) * B // 2].contiguous()) for i in range(2)]
streams = [torch.cuda.Stream() for _ in range(2)]


def run_full():
    return full(xt, t, mix)


def run_dual():
    cur = torch.cuda.current_stream()
    outs = []
    for s, m, a in zip(streams, halves, hx):
        s.wait_stream(cur)
        with torch.cuda.stream(s):
            outs.append(m(*a))
    for s in streams:
        cur.wait_stream(s)
    return outs


def run_seq_halves():
    return [m(*a) for m, a in zip(halves, hx)]


def timeit(fn, n=20):
    for _ in range(4):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


with torch.no_grad():
    a = run_full()
    b = torch.cat(run_dual(), dim=0)
    print(f"B={B}: full vs dual outputs rel-L2 {float((a - b).norm() / a.norm()):.3e}")
    for name, fn in (("one batch of B", run_full), ("two halves, one stream", run_seq_halves), ("two halves, two streams", run_dual),
                     ("one batch of B", run_full), ("two halves, two streams", run_dual)):
        print(f"{name:28s} {timeit(fn):8.3f} ms per score evaluation")
