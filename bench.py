"""Benchmark of the DiffSep PC separation sampler on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--batch 32]

One "step" = one full predictor-corrector sampler run (N=30, snr=0.5, 1 corrector step,
denoise => 60 score-network evaluations) over one batch of synthetic 4 s @ 8 kHz 2-speaker
mixtures (config C2 of BASELINE.json: batch 32 per GPU, bf16 score network).  Prints ONE JSON
line (rank 0): whole-job mixtures/s with inputs resident in HBM (`value`), the same through
the public API from pinned host buffers (`e2e`), the roofline of the dominant kernel class
(the tcgen05 convolution) measured live with CUDA events, and the CPU baseline (the oracle =
port of the reference's PyTorch CPU sampler, timed on this host's cores on a bounded sample).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

T_SAMPLES = 32000          # 4 s @ 8 kHz
N_STEPS_SDE = 30
NFE = 60
GFLOP_PER_SCORE_STEP = 134.7   # per mixture, C1/C2 shape, FlopCounter 2*M*N*K
METRIC = "mixtures/sec, PC sampler N=30 4s@8kHz 2-spk"


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            pk = json.load(f)
        return pk.get("hbm_gbs", 6650.0), pk.get("bf16_tflops_sustained", 1400.0), pk.get("bf16_tflops", 1590.0), "measured"
    except Exception:
        return 6650.0, 1400.0, 1590.0, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons while the timed region runs (profiling recipe)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        super().__init__(daemon=True)
        self.index, self.rows, self._halt = index, [], threading.Event()

    def run(self):
        while not self._halt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            self._halt.wait(0.2)

    def stop(self):
        self._halt.set()
        self.join(timeout=5)
        sm = sorted(float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit())
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) >= 7:
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
        mx = max((float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()), default=None)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(self.rows)}


# ----------------------------------------------------------------------------------- CPU baseline
def cpu_baseline(n_score_steps: int = 60):
    """The reference's PyTorch CPU sampler restated by the oracle (kind "port"), all host cores,
    bounded sample: 1 warm-up + n score-network evaluations of ONE C1 mixture inside the real
    sampler loop (prior draw, corrector and predictor updates included); the default n = 60 is one
    COMPLETE N=30 sampler run of one mixture (~15 s on 16 cores), smaller n is extrapolated."""
    import torch

    from common import sampler_inputs
    from oracle import ncsnpp_ref as R
    from oracle import sampler_ref as S

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = R.ScoreConfig()
    sd = R.make_state_dict(cfg.backbone, seed=0)
    mixn, _ = sampler_inputs(1, T_SAMPLES)
    with torch.no_grad():
        R.score_forward_ref(sd, cfg, torch.cat([mixn, mixn], 1) * 0.5, torch.full((1,), 0.5), mixn)  # warm-up
        n_sde = max(1, n_score_steps // 2)
        t0 = time.perf_counter()
        S.pc_sampler_ref(lambda x, t, y: R.score_forward_ref(sd, cfg, x, t, y), S.MixSDERef(2, 2.0, 0.05, 0.5, N=n_sde),
                         mixn, S.NoiseTape(4321), corrector="ald2", eps=0.03, snr=0.5)
        dt = time.perf_counter() - t0
    per_step = dt / (2 * n_sde)
    return {"value": 1.0 / (per_step * NFE), "unit": "mixtures/s", "cores": cores, "kind": "port",
            "ms_per_score_step": per_step * 1e3,
            "sample": f"{2 * n_sde} of the {NFE} score evaluations of one 4 s mixture (B=1, fp32, sampler loop with "
                      f"N={n_sde}), extrapolated x{NFE / (2 * n_sde):.1f}"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    per = []
    base = None
    # One step = one complete mixture (60 score evaluations) when the whole run then stays within ~3 minutes of CPU
    # time; otherwise a bounded, extrapolated sample of it (the sample is named in cpu_baseline.sample).
    probe = cpu_baseline(n_score_steps=4)
    budget_s = 180.0 / (args.warmup + args.steps)
    n_eval = int(budget_s / (probe["ms_per_score_step"] * 1e-3)) // 2 * 2
    n_eval = max(2, min(NFE, n_eval))
    for i in range(args.warmup + args.steps):
        base = cpu_baseline(n_score_steps=n_eval)
        if i >= args.warmup:
            per.append(base["value"])
    v = sum(per) / len(per)
    base["value"] = v
    line = {"metric": METRIC, "value": v, "unit": "mixtures/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 / v, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": {"workload": "C1-shaped mixture on host cores: reference PyTorch CPU sampler (oracle port), "
                                   "bounded sample per step", "n_samples": T_SAMPLES, "N": N_STEPS_SDE, "nfe": NFE},
            "cpu_baseline": base,
            "e2e": {"value": v, "unit": "mixtures/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------- ours
def run_ours(args):
    import torch
    import torch.distributed as dist

    from common import build_model
    from ditsep_b200 import ops
    from ditsep_b200.utils import normalize_batch
    from ditsep_b200.utils.synth import synth_mixtures

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B = args.batch
    precision = args.precision
    model, _, _ = build_model(precision, dev, N=N_STEPS_SDE)

    # synthetic mixtures: a handful of distinct ones tiled over the batch (generation is host-side and slow)
    mix8, _ = synth_mixtures(min(B, 8), 2, T_SAMPLES, seed=1234 + rank)
    mix = mix8.repeat((B + mix8.shape[0] - 1) // mix8.shape[0], 1, 1)[:B].contiguous()
    mix_host = mix.pin_memory()
    (mix_dev, _), _, _ = normalize_batch((mix.to(dev), None))
    out_host = torch.empty(B, 2, T_SAMPLES).pin_memory()

    def step_resident(seed):
        sampler = model.get_pc_sampler("reverse_diffusion", "ald2", mix_dev, N=N_STEPS_SDE, snr=0.5, corrector_steps=1,
                                       denoise=True, seed=seed)
        x, nfe = sampler()
        return x

    def step_e2e(seed):
        m = mix_host.to(dev, non_blocking=True)
        (mn, _), mean, std = normalize_batch((m, None))
        sampler = model.get_pc_sampler("reverse_diffusion", "ald2", mn, N=N_STEPS_SDE, snr=0.5, corrector_steps=1,
                                       denoise=True, seed=seed)
        x, nfe = sampler()
        out_host.copy_(x, non_blocking=True)
        return x

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for i in range(warmup):
            fn(i)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n0 = ops.launch_count()
        e0.record()
        for i in range(steps):
            fn(1000 + i)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), ops.launch_count() - n0

    clocks = ClockSampler(local) if rank == 0 else None
    if clocks:
        clocks.start()
    ms_res, launches = timed(step_resident, args.steps, args.warmup)
    ms_e2e, _ = timed(step_e2e, args.steps, max(1, args.warmup // 2))
    clk = clocks.stop() if clocks else None

    value = world * B * args.steps / (ms_res * 1e-3)
    e2e = world * B * args.steps / (ms_e2e * 1e-3)

    # ---- roofline of the dominant kernel class (the convolutions), measured live: one extra score step with
    # CUDA events around every launch of the plan on the launching stream
    hbm, tf_sust, tf_burst, peak_src = load_peaks()
    prof = profile_score_step(model, B, dev)
    conv_ms = sum(ms for name, ms, fl in prof if name.startswith("conv"))
    conv_fl = sum(fl for name, ms, fl in prof if name.startswith("conv"))
    total_ms = sum(ms for name, ms, fl in prof)
    achieved = conv_fl / (conv_ms * 1e-3) / 1e12 if conv_ms > 0 else 0.0
    roofline = {"bound": "tensor", "kernel": "conv (tcgen05 implicit GEMM + CUDA-core narrow layers)",
                "achieved": achieved, "peak": tf_sust, "unit": "TFLOP/s", "frac": achieved / tf_sust,
                "peak_source": f"{peak_src} bf16_tflops_sustained", "traffic": None,
                "conv_share_of_score_step": conv_ms / total_ms if total_ms else None,
                "score_step_ms_event_sum": total_ms,
                "whole_path_frac": value / world * NFE * GFLOP_PER_SCORE_STEP * 1e9 / (tf_sust * 1e12)}
    # the single dominant kernel of the step: the 3x3 64->64 convolution at 256x256 (with or without the fused shortcut
    # tap).  traffic: dram__bytes_read.sum + dram__bytes_write.sum of one such launch from the committed ncu --set full
    # capture (profiles/r1o_ncu_full_conv_tc_256x256_B32.txt: 303 MB read + 230 MB written at batch 32, against
    # 2 * B * H * W * 64 * 2 B = 537 MB algorithmic), scaled to this run's batch.
    dom = [(ms, fl) for name, ms, fl in prof if name.startswith("conv_tc:256x256:") and name.endswith("->64")
           and name.split(":")[2].split("->")[0] in ("576", "640")]
    if dom:
        roofline["dominant_kernel"] = {
            "name": "conv_tc_tile_kernel<3,resident,bf16> 256x256 64->64 3x3 (+identity shortcut tap)",
            "launches_per_score_step": len(dom), "avg_us": 1e3 * sum(m for m, _ in dom) / len(dom),
            "achieved_tflops": sum(f for _, f in dom) / sum(m for m, _ in dom) / 1e9,
            "frac_of_sustained_peak": sum(f for _, f in dom) / sum(m for m, _ in dom) / 1e9 / tf_sust,
            "ss_mode_floor_note": "tcgen05.mma SS mode costs 43 + N/2 cycles per 128xNx16 MMA (profiles/r1j_umma_probe.txt): "
                                  "N = Cout = 64 caps this layer at 42.6 % of the 2.29 PFLOP/s nominal rate = 976 TFLOP/s"}
        roofline["traffic"] = 533.1e6 * B / 32.0
    by_kind = {}
    for name, ms, fl in prof:
        k = name.split(":")[0]
        by_kind[k] = by_kind.get(k, 0.0) + ms
    roofline["ms_by_kernel_class"] = {k: round(v, 3) for k, v in sorted(by_kind.items(), key=lambda kv: -kv[1])}
    dump = os.environ.get("DITSEP_BENCH_DUMP")
    if dump and rank == 0:  # per-launch-shape table of the profiled score step (count, total ms, TFLOP/s)
        agg = {}
        for name, ms, fl in prof:
            c = agg.setdefault(name, [0, 0.0, 0.0])
            c[0] += 1; c[1] += ms; c[2] += fl
        with open(dump, "w") as f:
            for name, (cnt, ms, fl) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
                f.write(f"{name:44s} n={cnt:3d} total {ms:8.3f} ms  avg {1e3 * ms / cnt:8.1f} us"
                        + (f"  {fl / ms / 1e9:7.1f} TFLOP/s" if fl else "") + "\n")

    line = None
    if rank == 0:
        # the CPU baseline is timed at N=1 only: under torchrun the other ranks' processes share the host cores
        base = cpu_baseline() if (world == 1 and not args.no_cpu_baseline) else None
        line = {"metric": METRIC, "value": value, "unit": "mixtures/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_res / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "bf16" if precision == "bf16" else "f32", "data": "synthetic",
                "config": {"workload": f"C2: batch {B} per GPU of 4 s @ 8 kHz 2-speaker mixtures, NCSN++ nf=64, PC "
                                       f"sampler N=30 snr=0.5 corrector_steps=1 denoise (60 NFE), {precision} score net, "
                                       "in-kernel Philox noise, random-init weights",
                           "batch_per_gpu": B, "n_samples": T_SAMPLES, "N": N_STEPS_SDE, "nfe": NFE,
                           "ms_per_score_step": ms_res / args.steps / NFE,
                           "l2_policy": "inputs larger than L2: activations of one score step exceed 126 MB at batch 32"},
                "e2e": {"value": e2e, "unit": "mixtures/s", "h2d_bytes_per_step": B * T_SAMPLES * 4,
                        "d2h_bytes_per_step": B * 2 * T_SAMPLES * 4},
                "gpu_launches": launches, "clocks": clk, "roofline": roofline, "cpu_baseline": base}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def profile_score_step(model, B, dev):
    """(name, ms, flops) for every launch of one score step, CUDA events on the launching stream."""
    import torch

    from ditsep_b200 import ops

    xt = torch.randn(B, 2, T_SAMPLES, device=dev)
    mix = torch.randn(B, 1, T_SAMPLES, device=dev)
    t = torch.full((B,), 0.5, device=dev)
    with torch.no_grad():
        model(xt, t, mix)
        plan = model.score_model.backbone.engine().plan(B, 256, 256, dev)
        rows = []
        evs = []
        # park the GPU behind a ~30 ms spin so that the host (ctypes launch path, ~15 us per launch) runs ahead
        # and every event interval below is pure device time, not launch latency
        torch.cuda._sleep(int(30e-3 * 1.9e9))
        for fn, a, kw in plan.steps:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn(*a, **kw)
            e1.record()
            evs.append((fn, a, kw, e0, e1))
        torch.cuda.synchronize()
        for fn, a, kw, e0, e1 in evs:
            name = getattr(fn, "__name__", "op")
            fl = 0.0
            if name == "run":  # ConvTcPlan.run
                pl = fn.__self__
                Bq, H, W = pl.shape
                fl = 2.0 * Bq * H * W * pl.Cout * pl.K
                name = f"conv_tc:{H}x{W}:{pl.K}->{pl.Cout}"
            elif name == "conv_direct":
                x, w = a[0], a[1]
                fl = 2.0 * x.shape[0] * x.shape[1] * x.shape[2] * w.shape[0] * w.shape[1] * w.shape[2]
                name = f"conv_direct:{x.shape[1]}x{x.shape[2]}:{w.shape[1] * w.shape[2]}->{w.shape[0]}"
            elif a and hasattr(a[0], "shape") and len(a[0].shape) == 4:
                name = f"{name}:{a[0].shape[1]}x{a[0].shape[2]}x{a[0].shape[3]}"
            rows.append((name, e0.elapsed_time(e1), fl))
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
