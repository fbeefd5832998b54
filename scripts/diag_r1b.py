"""Diagnostics (GPU): bf16 parity per conv mode / init scale, batch independence, per-launch times."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, ROOT + "/tests")
import torch
from common import rel_l2, score_inputs
from ditsep_b200.diffsep import DiffSepModel, default_config
from oracle import ncsnpp_ref as R
import bench

dev = torch.device("cuda", 0)
xt, t, mixn = score_inputs(1, 7810)
xt8, t8, mixn8 = score_inputs(8, 32000)
for scale in (1.0, 0.01):
    cfg = R.ScoreConfig()
    sd = R.make_state_dict(cfg.backbone, seed=0, init_scale=scale)
    with torch.no_grad():
        ref = R.score_forward_ref(sd, cfg, xt, t, mixn)
    for mode in ("tap", "halo"):
        os.environ["DITSEP_CONV_MODE"] = mode
        for prec in ("fp32", "bf16"):
            m = DiffSepModel(default_config(precision=prec)); m.score_model.load_state_dict(sd); m = m.to(dev).eval()
            with torch.no_grad():
                out = m(xt.to(dev), t.to(dev), mixn.to(dev))
                full = m(xt8.to(dev), t8.to(dev), mixn8.to(dev)).clone()
                one = m(xt8[5:6].to(dev), t8[5:6].to(dev), mixn8[5:6].to(dev))
            print(f"init_scale={scale} mode={mode} {prec}: rel-L2 vs oracle {rel_l2(out, ref):.3e}  batch-indep {rel_l2(full[5:6], one):.3e}", flush=True)
            if prec == "bf16" and scale == 1.0:
                rows = bench.profile_score_step(m, 32, dev)
                agg = {}
                for name, ms, fl in rows:
                    a = agg.setdefault(name, [0, 0.0, 0.0]); a[0] += 1; a[1] += ms; a[2] += fl
                with open(f"{ROOT}/gpurun_out/launch_times_{mode}.txt", "w") as f:
                    tot = sum(v[1] for v in agg.values())
                    for name, (n, ms, fl) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
                        f.write(f"{name:40s} n={n:3d} ms={ms:8.3f} share={ms / tot:6.3f} TFLOP/s={fl / ms / 1e9 if ms else 0:8.1f}\n")
                    f.write(f"total {tot:.3f} ms\n")
            del m
