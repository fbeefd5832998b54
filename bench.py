"""Benchmark of the DiffSep PC separation sampler on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config C1|C2|C3|C4|C5]

One "step" = one full predictor-corrector sampler run (N=30, snr=0.5, 1 corrector step, denoise => 60 score-network
evaluations) over one batch of synthetic mixtures.  Default workload = config C2 of BASELINE.json (the one the
metric is quoted on): batch 32 per GPU of 4 s @ 8 kHz 2-speaker mixtures, bf16 score network.  The other configs
(`--config`): C1 = batch 1 (the reference's CPU-runnable case), C3 = 10 s @ 16 kHz (W = 1280 frames), C4 = latent
sampler (nf=128, OUVE SDE + ald corrector, L = 80), C5 = variable-length sweep through `separate_sharded` (strong
scaling, NCCL gather of rows and waveforms inside the timed region).

Prints ONE JSON line (rank 0):
  value            whole-job mixtures/s with inputs resident in HBM
  e2e              the same through the public API from pinned host buffers (H2D + D2H inside the timed region)
  roofline         tensor roofline of the dominant kernel class (the convolutions), ALGORITHMIC FLOPs (nn.Conv2d's
                   2*M*N*K with the real Cin / Cout; fused identity shortcuts and operand padding count zero) over
                   CUDA-event time, measured live on one eager score step
  roofline_hbm     per HBM-bound kernel: algorithmic bytes, measured us, fraction of measured HBM peak
  library_baseline the oracle's torch code moved to the same GPU (torch eager + cuDNN / cuFFT, TF32 and bf16
                   autocast): the existing-library bar of SURVEY.md 2.2 / 8d, timed in the same run
  cpu_baseline     the oracle (port of the reference's PyTorch CPU sampler) on this host's cores, one whole mixture
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

N_STEPS_SDE = 30
NFE = 60
METRIC = "mixtures/sec, PC sampler N=30 4s@8kHz 2-spk"
# per-config workload: batch per GPU, samples, precision, algorithmic GFLOP per score step per mixture
CONFIGS = {
    # (C1 runs the bf16 tensor-core network like C2; --precision fp32 selects the exact CUDA-core path the parity
    #  tests use as their fp32 reference: 20.6 ms per score step at batch 1 against 2.5 ms in bf16, r2n)
    "C1": dict(batch=1, T=32000, fs=8000, precision="bf16", gflop=134.7, kind="stft",
               desc="C1: ONE 4 s @ 8 kHz 2-speaker mixture (batch 1), NCSN++ nf=64"),
    "C2": dict(batch=32, T=32000, fs=8000, precision="bf16", gflop=134.7, kind="stft",
               desc="C2: batch 32 per GPU of 4 s @ 8 kHz 2-speaker mixtures, NCSN++ nf=64"),
    "C3": dict(batch=8, T=160000, fs=16000, precision="bf16", gflop=675.0, kind="stft",
               desc="C3: batch 8 per GPU of 10 s @ 16 kHz 2-speaker mixtures (W = 1280 frames, the longest bucket), "
                    "NCSN++ nf=64"),
    "C4": dict(batch=32, L=80, precision="bf16", gflop=67.0, kind="latent",
               desc="C4: batch 32 per GPU of (1, 64, 80) latents, LatentScoreModelNCSNpp nf=128 ch_mult [1,2,2] "
                    "centered, OUVESDE theta=1.5 sigma [0.96, 10] + ald"),
    "C5": dict(batch=16, precision="bf16", gflop=None, kind="sweep",
               desc="C5: variable-length 16 kHz 2-speaker sweep (Libri2Mix length histogram, 3-10 s), length-bucketed "
                    "batches sharded over the ranks, NCCL gather of rows + waveforms in the timed region"),
}


def workload_config(name, B, precision):
    c = CONFIGS[name]
    d = {"workload": f"{c['desc']}, PC sampler N=30 snr=0.5 corrector_steps=1 denoise (60 NFE), {precision} score net, "
                     "in-kernel Philox noise, random-init weights", "name": name, "batch_per_gpu": B, "N": N_STEPS_SDE,
         "nfe": NFE}
    if "T" in c:
        d["n_samples"] = c["T"]
    return d


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
def cpu_mixture(n_score_steps: int = NFE, T: int = 32000):
    """The reference's PyTorch CPU sampler restated by the oracle (kind "port"), all host cores: n score-network
    evaluations of ONE mixture inside the real sampler loop (prior draw, corrector and predictor updates included);
    n = 60 is one COMPLETE N=30 sampler run.  Returns seconds."""
    import torch

    from common import sampler_inputs
    from oracle import ncsnpp_ref as R
    from oracle import sampler_ref as S

    cfg = R.ScoreConfig()
    sd = cpu_mixture.sd if hasattr(cpu_mixture, "sd") else R.make_state_dict(cfg.backbone, seed=0)
    cpu_mixture.sd = sd
    mixn, _ = sampler_inputs(1, T)
    n_sde = max(1, n_score_steps // 2)
    with torch.no_grad():
        t0 = time.perf_counter()
        S.pc_sampler_ref(lambda x, t, y: R.score_forward_ref(sd, cfg, x, t, y), S.MixSDERef(2, 2.0, 0.05, 0.5, N=n_sde),
                         mixn, S.NoiseTape(4321), corrector="ald2", eps=0.03, snr=0.5)
        return time.perf_counter() - t0


def cpu_baseline(T: int = 32000):
    import torch

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cpu_mixture(2, T)  # warm-up: two score evaluations
    dt = cpu_mixture(NFE, T)
    return {"value": 1.0 / dt, "unit": "mixtures/s", "cores": cores, "kind": "port",
            "ms_per_score_step": dt / NFE * 1e3,
            "sample": f"one whole mixture of the workload (all {NFE} score evaluations, B=1, fp32, T={T}) in {dt:.1f} s"}


def run_reference(args):
    """Reference arm: the reference's own CPU implementation of the path (the oracle port: the reference is Python
    and cannot travel to the GPU box) on the host cores.  Every TIMED step is one WHOLE mixture of the workload (all
    60 score evaluations: nothing is extrapolated); the untimed warm-up steps are two score evaluations each."""
    import torch

    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    c = CONFIGS[args.config if args.config in ("C1", "C2", "C3") else "C2"]
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    for _ in range(args.warmup):
        cpu_mixture(2, c["T"])
    # whole mixtures as long as the run stays inside the budget (default 9 minutes of timed work); if this host is
    # too slow for that, the REMAINING steps shrink to a bounded number of score evaluations -- every reported time
    # is a measured one (ms_per_step = mean measured step), the value counts mixture-equivalents actually computed
    budget = float(os.environ.get("DITSEP_REF_BUDGET_S", "540"))
    spent, done_eval, n_eval, plan = 0.0, 0, NFE, []
    for i in range(args.steps):
        dt_i = cpu_mixture(n_eval, c["T"])
        spent += dt_i
        done_eval += n_eval
        plan.append(n_eval)
        left = args.steps - 1 - i
        if left and spent + left * dt_i > budget:
            per_eval = spent / done_eval
            n_eval = max(2, min(NFE, int((budget - spent) / left / per_eval) // 2 * 2))
    dt = spent / args.steps
    v = (done_eval / NFE) / spent
    whole = sum(1 for n in plan if n == NFE)
    base = {"value": v, "unit": "mixtures/s", "cores": cores, "kind": "port", "ms_per_score_step": spent / done_eval * 1e3,
            "sample": f"{whole} of the {args.steps} timed steps = one WHOLE mixture of the workload (all {NFE} score "
                      f"evaluations, fp32, T={c['T']}, 1 of the {args.batch or c['batch']} mixtures of a batch)"
                      + ("" if whole == args.steps else f"; the other steps ran {sorted(set(plan) - {NFE})} score "
                         f"evaluations to stay inside {budget:.0f} s; value = mixture-equivalents computed / time "
                         "measured") + "; warm-up steps = 2 score evaluations"}
    B = args.batch or c["batch"]
    line = {"metric": METRIC, "value": v, "unit": "mixtures/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": workload_config(args.config, B, args.precision or c["precision"]),
            "cpu_baseline": base,
            "e2e": {"value": v, "unit": "mixtures/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------- builders
def build_workload(name, precision, dev, B, rank):
    """Returns (model, mix (B, 1, ...) on the host, sampler factory (mix_dev, seed) -> sampler, out shape)."""
    import torch

    from common import build_model
    from ditsep_b200.utils.synth import synth_mixtures

    c = CONFIGS[name]
    if c["kind"] == "stft":
        model, _, _ = build_model(precision, dev, N=N_STEPS_SDE)
        mix8, _ = synth_mixtures(min(B, 4), 2, c["T"], fs=c["fs"], seed=1234 + rank)
        mix = mix8.repeat((B + mix8.shape[0] - 1) // mix8.shape[0], 1, 1)[:B].contiguous()

        def make(y, seed):
            return model.get_pc_sampler("reverse_diffusion", "ald2", y, N=N_STEPS_SDE, snr=0.5, corrector_steps=1,
                                        denoise=True, seed=seed)
        return model, mix, make, (B, 2, c["T"])
    # C4: latent sampler (config/latent_diffsep_ouve/model/default.yaml:16-46); VAE encode / decode excluded
    from ditsep_b200 import sdes
    from ditsep_b200.models.diffsep.score_models import LatentScoreModelNCSNpp
    from oracle import ncsnpp_ref as R

    bcfg = R.NCSNppConfig(nf=128, ch_mult=(1, 2, 2), num_res_blocks=2, attn_resolutions=(16,), image_size=64,
                          num_channels_in=3, num_channels_out=2, centered=True)
    sd = R.make_state_dict(bcfg, seed=5, with_window=None)
    lat = LatentScoreModelNCSNpp(2, dict(_target_="models.diffsep.ncsnpp.NCSNpp", nf=128, ch_mult=[1, 2, 2],
                                         num_res_blocks=2, attn_resolutions=[16], resamp_with_conv=True, image_size=64,
                                         centered=True, precision=precision), max_latent_length=4).eval()
    lat.load_state_dict(sd, strict=True)
    lat = lat.to(dev)
    mix = torch.randn(B, 1, 64, c["L"], generator=torch.Generator().manual_seed(77 + rank))

    def make(y, seed):
        return sdes.get_pc_sampler("reverse_diffusion", "ald", sde=sdes.OUVESDE(theta=1.5, sigma_min=0.96, sigma_max=10.0,
                                                                                  N=N_STEPS_SDE),
                                   score_fn=lat, y=y, eps=0.03, snr=0.5, corrector_steps=1, denoise=True, n_spkrs=2,
                                   seed=seed)
    return lat, mix, make, (B, 2, 64, c["L"])


# ----------------------------------------------------------------------------------- ours
def run_ours(args):
    import torch
    import torch.distributed as dist

    from ditsep_b200 import ops
    from ditsep_b200.utils import normalize_batch

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    c = CONFIGS[args.config]
    if c["kind"] == "sweep":
        return run_sweep(args, dev, world, rank)
    B = args.batch or c["batch"]
    precision = args.precision or c["precision"]
    model, mix, make_sampler, out_shape = build_workload(args.config, precision, dev, B, rank)
    latent = c["kind"] == "latent"

    mix_host = mix.pin_memory()
    if latent:
        mix_dev = mix.to(dev)
    else:
        (mix_dev, _), _, _ = normalize_batch((mix.to(dev), None))
    out_host = torch.empty(out_shape).pin_memory()

    def step_resident(seed):
        x, nfe = make_sampler(mix_dev, seed)()
        return x

    def step_e2e(seed):
        m = mix_host.to(dev, non_blocking=True)
        if not latent:
            (m, _), mean, std = normalize_batch((m, None))
        x, nfe = make_sampler(m, seed)()
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
    ms_e2e, _ = timed(step_e2e, args.steps, max(3, args.warmup // 2))
    clk = clocks.stop() if clocks else None

    value = world * B * args.steps / (ms_res * 1e-3)
    e2e = world * B * args.steps / (ms_e2e * 1e-3)

    line = None
    if rank == 0:
        hbm, tf_sust, tf_burst, peak_src = load_peaks()
        roofline, roofline_hbm = rooflines(model, args.config, B, dev, latent, value / world, hbm, tf_sust, peak_src)
        lib = None if args.no_library_baseline else library_baseline(args.config, B, dev)
        # the CPU baseline is timed at N=1 only: under torchrun the other ranks' processes share the host cores
        base = cpu_baseline(c.get("T", 32000)) if (world == 1 and not args.no_cpu_baseline and not latent) else None
        cfg = workload_config(args.config, B, precision)
        cfg["ms_per_score_step"] = ms_res / args.steps / NFE
        act_mb = B * 256 * ((c.get("T", 0) + 382) // 128 + 1) * 64 * 2 / 1e6 if not latent else B * 64 * c["L"] * 128 * 2 / 1e6
        cfg["l2_policy"] = (f"inputs larger than L2: ONE top-level activation tensor is {act_mb:.0f} MB (L2: 126 MB)"
                            if act_mb > 126 else
                            f"working set of a score step partly L2-resident (top-level tensor {act_mb:.0f} MB): the "
                            "sampler's own access pattern, no flush between score steps")
        n_in = mix_host.numel() * 4
        n_out = out_host.numel() * 4
        line = {"metric": METRIC, "value": value, "unit": "mixtures/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_res / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "bf16" if precision == "bf16" else "f32", "data": "synthetic",
                "config": cfg,
                "e2e": {"value": e2e, "unit": "mixtures/s", "h2d_bytes_per_step": n_in, "d2h_bytes_per_step": n_out},
                "gpu_launches": launches, "clocks": clk, "roofline": roofline, "roofline_hbm": roofline_hbm,
                "library_baseline": lib, "cpu_baseline": base}
        if latent and world == 1:
            line["vae"] = time_vae(B, c["L"], dev)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def time_vae(B, L, dev):
    """C4 companion number: the Oobleck VAE either side of the latent sampler (LatentDiffSep.encode / decode,
    oobleck_finetune.json widths, random weights) on B mixtures of L * 2048 samples; not part of `value` (the metric is
    the sampler), reported so that the latent pipeline's whole cost is visible."""
    import torch

    from ditsep_b200.models.oobleck import OobleckVAE
    from oracle import oobleck_ref as O

    ch, cm, st, lat = 128, [1, 2, 4, 8, 16], [2, 4, 4, 8, 8], 128
    vae = OobleckVAE(ch, cm, st, lat, device=dev).load_state_dict(O.make_state_dict(ch, cm, st, lat, seed=7))
    x = 0.3 * torch.randn(B, 1, L * 2048, device=dev)
    z = torch.randn(2 * B, lat // 2, L, device=dev)
    out = {}
    for name, fn in (("encode", lambda: vae.encode(x, seed=1)), ("decode_2src", lambda: vae.decode(z))):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            fn()
        e1.record()
        torch.cuda.synchronize()
        out[name + "_ms_per_mixture"] = e0.elapsed_time(e1) / 3 / B
    out["note"] = "weights: oracle.oobleck_ref.make_state_dict (the weight generator is test infrastructure; the VAE runs in the library)"
    return out


# ----------------------------------------------------------------------------------- C5: the sweep, as designed
def sweep_lengths(n: int, seed: int = 7):
    """n utterance lengths (samples @ 16 kHz) following the Libri2Mix test-set histogram the reference's results/
    directory records (mean 8.4 s, 3.1-21.3 s), clipped to the 3-10 s range config C3 names."""
    import torch

    g = torch.Generator().manual_seed(seed)
    u = torch.rand(8 * n, generator=g)
    secs = 3.1 + (-torch.log(1 - u * 0.98)) * 4.6   # exponential tail above the 3.1 s minimum (mean ~8 s unclipped)
    secs = secs[secs <= 10.0][:n]
    return [int(s * 16000) for s in secs.tolist()]


def run_sweep(args, dev, world, rank):
    """Strong scaling: a FIXED set of variable-length mixtures through `separate_sharded` -- length-bucketed ragged
    batches per rank, no data-path collective, then ONE all-gather of the metric rows and one of the padded
    separated waveforms, both inside the timed region.  Reports per-rank busy time so imbalance is visible."""
    import torch
    import torch.distributed as dist

    from common import build_model
    from ditsep_b200 import evaluate, ops
    from ditsep_b200.utils import shard
    from ditsep_b200.utils.synth import synth_mixtures

    c = CONFIGS["C5"]
    precision = args.precision or c["precision"]
    n_mix = args.sweep_mixtures
    B = args.batch or c["batch"]
    model, _, _ = build_model(precision, dev, N=N_STEPS_SDE)
    lengths = sweep_lengths(n_mix)
    base_mix, base_src = synth_mixtures(4, 2, max(lengths), fs=16000, seed=4242)
    mixtures = [base_mix[i % 4, :, :n].clone() for i, n in enumerate(lengths)]
    refs = [base_src[i % 4, :, :n].clone() for i, n in enumerate(lengths)]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def one_sweep(seed):
        t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
        t0.record()
        # the evaluation driver (reference row schema, SI-SDR / SIR / SAR with PIT on the device, summary)
        rows, summary, outs = evaluate.evaluate_sweep(model, mixtures, refs, fs=16000, N=N_STEPS_SDE, snr=0.5,
                                                      corrector_steps=1, denoise=True, max_batch=B, device=dev,
                                                      seed=seed, perceptual=False)
        rows = (rows, summary)
        t1.record()
        wav = shard.gather_waveforms(outs, lengths, dev) if world > 1 else None
        return rows, t0, t1, wav

    # warm-up: every bucket's launch plan is built and graph-captured once
    one_sweep(0)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n0 = ops.launch_count()
    e0.record()
    busy = 0.0
    for i in range(args.steps):
        rows, t0, t1, wav = one_sweep(100 + i)
        torch.cuda.synchronize()
        busy += t0.elapsed_time(t1)
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    busy_t = torch.tensor([busy / args.steps], device=dev)
    allbusy = [busy_t.clone() for _ in range(world)]
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_gather(allbusy, busy_t)
    launches = ops.launch_count() - n0
    if rank == 0:
        value = n_mix * args.steps / (float(ms.item()) * 1e-3)
        results, summary = rows
        assert sorted(results) == list(range(n_mix)) and all(r["nfe"] == NFE for r in results.values())
        line = {"metric": METRIC, "value": value, "unit": "mixtures/s", "n_gpus": world, "steps": args.steps,
                "warmup": 1, "ms_per_step": float(ms.item()) / args.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "bf16" if precision == "bf16" else "f32",
                "data": "synthetic",
                "config": {**workload_config("C5", B, precision), "n_mixtures": n_mix,
                           "mean_len_s": sum(lengths) / len(lengths) / 16000.0,
                           "buckets": sorted({shard.padded_frames(n) for n in lengths})},
                "e2e": {"value": value, "unit": "mixtures/s", "h2d_bytes_per_step": sum(lengths) * 4,
                        "d2h_bytes_per_step": 0},
                "gpu_launches": launches,
                "per_rank_busy_ms": [float(b.item()) for b in allbusy],
                "summary_random_weights": {k: summary[k] for k in ("si_sdr", "si_sir", "si_sar", "nfe", "runtime",
                                                                   "len_s", "number")}}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# ----------------------------------------------------------------------------------- rooflines
def _flush_l2(buf):
    buf.zero_()


def rooflines(model, name, B, dev, latent, value_per_gpu, hbm, tf_sust, peak_src):
    import torch

    from ditsep_b200 import ops

    c = CONFIGS[name]
    prof, plan = profile_score_step(model, name, B, dev, latent)
    conv_ms = sum(ms for k, ms, fl, by in prof if k.startswith("conv"))
    conv_fl = sum(fl for k, ms, fl, by in prof if k.startswith("conv"))
    total_ms = sum(ms for k, ms, fl, by in prof)
    achieved = conv_fl / (conv_ms * 1e-3) / 1e12 if conv_ms > 0 else 0.0
    roofline = {"bound": "tensor", "kernel": "conv class (tcgen05 implicit GEMM + CUDA-core narrow layers)",
                "achieved": achieved, "peak": tf_sust, "unit": "TFLOP/s", "frac": achieved / tf_sust,
                "flops": "algorithmic (nn.Conv2d 2*M*N*K, real Cin/Cout; identity shortcuts and operand padding = 0)",
                "peak_source": f"{peak_src} bf16_tflops_sustained",
                # DRAM bytes of ONE launch of the dominant kernel (conv_ts_kernel, 256x256 64->64, batch 32) from the committed
                # ncu --set full capture of this kernel: 321.2 MB read + 227.4 MB written against 536.9 MB algorithmic
                "traffic": 548.6e6 if (name == "C2" and B == 32) else None,
                "traffic_from_profile": "profiles/r2w_conv_ts.txt: ncu --set full capture of the dominant conv launch "
                                        "(dram__bytes_read.sum + dram__bytes_write.sum per launch); not re-measured by this run",
                "conv_share_of_score_step": conv_ms / total_ms if total_ms else None,
                "score_step_ms_event_sum": total_ms,
                "whole_path_frac": value_per_gpu * NFE * c["gflop"] * 1e9 / (tf_sust * 1e12)}
    # the single dominant kernel: the 3x3 64->64 convolutions at the top resolution (plain / with the identity tap)
    top = max((int(k.split(":")[1].split("x")[0]) * int(k.split(":")[1].split("x")[1]), k) for k, *_ in prof
              if k.startswith("conv_tc:"))[0] if any(k.startswith("conv_tc:") for k, *_ in prof) else 0
    dom = [(ms, fl) for k, ms, fl, by in prof if k.startswith("conv_tc:") and k.split(":")[2] == "64->64"
           and k.split(":")[3] in ("k3", "k3+x64")
           and int(k.split(":")[1].split("x")[0]) * int(k.split(":")[1].split("x")[1]) == top]
    if dom:
        tf = sum(f for _, f in dom) / sum(m for m, _ in dom) / 1e9
        roofline["dominant_kernel"] = {"name": "conv_tc 3x3 64->64 at the top resolution (with / without the fused "
                                               "identity-shortcut tap; algorithmic FLOPs)",
                                       "launches_per_score_step": len(dom),
                                       "avg_us": 1e3 * sum(m for m, _ in dom) / len(dom), "achieved_tflops": tf,
                                       "frac_of_sustained_peak": tf / tf_sust}
    by_kind, hb = {}, {}
    for k, ms, fl, by in prof:
        kind = k.split(":")[0]
        by_kind[kind] = by_kind.get(kind, 0.0) + ms
        if by:
            e = hb.setdefault(k, [0, 0.0, 0.0])
            e[0] += 1; e[1] += ms; e[2] += by
    # the kernels outside the launch plan: STFT + compress, expand + iSTFT, the two PC updates -- timed alone with the
    # L2 flushed between repetitions (in the real step a whole network evaluation separates two calls)
    extra = time_spectral_and_pc(model, name, B, dev, latent)
    for k, (ms, by) in extra.items():
        by_kind[k] = by_kind.get(k, 0.0) + ms
        hb[k] = [1, ms, by]
    roofline["ms_by_kernel_class"] = {k: round(v, 3) for k, v in sorted(by_kind.items(), key=lambda kv: -kv[1])}
    roofline_hbm = {"peak": hbm, "unit": "GB/s", "peak_source": f"{peak_src} hbm_gbs", "kernels": {}}
    for k, (cnt, ms, by) in sorted(hb.items(), key=lambda kv: -kv[1][1])[:14]:
        gbs = by / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
        roofline_hbm["kernels"][k] = {"launches": cnt, "us_avg": 1e3 * ms / cnt, "bytes_per_launch": by / cnt,
                                      "achieved": gbs, "frac": gbs / hbm}
    dump = os.environ.get("DITSEP_BENCH_DUMP")
    if dump:  # per-launch-shape table of the profiled score step (count, total ms, TFLOP/s or GB/s)
        agg = {}
        for k, ms, fl, by in prof:
            a = agg.setdefault(k, [0, 0.0, 0.0, 0.0])
            a[0] += 1; a[1] += ms; a[2] += fl; a[3] += by
        with open(dump, "w") as f:
            for k, (cnt, ms, fl, by) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
                f.write(f"{k:44s} n={cnt:3d} total {ms:8.3f} ms  avg {1e3 * ms / cnt:8.1f} us"
                        + (f"  {fl / ms / 1e9:7.1f} TFLOP/s" if fl else "")
                        + (f"  {by / ms / 1e6:7.0f} GB/s" if by else "") + "\n")
            for k, (ms, by) in extra.items():
                f.write(f"{k:44s} n=  1 total {ms:8.3f} ms  avg {1e3 * ms:8.1f} us  {by / ms / 1e6:7.0f} GB/s\n")
    return roofline, roofline_hbm


def profile_score_step(model, name, B, dev, latent):
    """(name, ms, algorithmic flops, algorithmic HBM bytes) for every launch of one score step's launch plan, CUDA
    events on the launching stream."""
    import torch

    c = CONFIGS[name]
    with torch.no_grad():
        if latent:
            xt = torch.randn(B, 2, 64, c["L"], device=dev)
            mix = torch.randn(B, 1, 64, c["L"], device=dev)
            t = torch.full((B,), 0.5, device=dev)
            model(xt, t, mix)
            eng = model.backbone.engine()
        else:
            xt = torch.randn(B, 2, c["T"], device=dev)
            mix = torch.randn(B, 1, c["T"], device=dev)
            t = torch.full((B,), 0.5, device=dev)
            model(xt, t, mix)
            eng = model.score_model.backbone.engine()
        plan = max(eng._plans.values(), key=lambda p: p.B * p.H * p.W)
        rows, evs = [], []
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
            nm = getattr(fn, "__name__", "op")
            fl, by = 0.0, 0.0
            if nm == "run":  # ConvTcPlan.run
                pl = fn.__self__
                Bq, H, W = pl.shape
                fl = getattr(pl, "algo_flops", 0.0)
                ks = 3 if pl.K >= 9 * pl.Cin else 1
                extra = pl.K - ks * ks * pl.Cin   # fused shortcut columns (identity: 0 algorithmic FLOPs; Conv_2: real)
                cin = pl.Cin if (extra or not fl) else int(round(fl / (2.0 * Bq * H * W * pl.Cout * ks * ks)))
                nm = f"conv_tc:{H}x{W}:{cin}->{pl.Cout}:k{ks}" + (f"+x{extra}" if extra else "")
            elif nm == "conv_direct":
                x, w = a[0], a[1]
                fl = 2.0 * x.shape[0] * x.shape[1] * x.shape[2] * w.shape[0] * w.shape[1] * w.shape[2]
                nm = f"conv_direct:{x.shape[1]}x{x.shape[2]}:{w.shape[2]}->{w.shape[0]}:k{int(w.shape[1] ** 0.5)}"
            elif a and hasattr(a[0], "shape") and len(a[0].shape) == 4:
                x = a[0]
                n_el = x.shape[0] * x.shape[1] * x.shape[2] * x.shape[3]
                e = x.element_size()
                if nm in ("gn_apply", "gn_apply_srcs"):                # (_srcs: scale / shift folded in the same launch)
                    by = 2.0 * n_el * e                                   # read x, write act(GN(x))
                    nm = "gn_apply"
                elif nm == "gn_partial_sums":
                    by = 1.0 * n_el * e
                elif nm == "fir_resample":
                    up = bool(a[1])
                    n_out = sum(1 for key in ("out_x", "out_h") if kw.get(key) is not None)
                    by = n_el * e * (1.0 + n_out * (4.0 if up else 0.25))  # read once, write every output once
                    nm = "fir_up" if up else "fir_down"
                elif nm == "split_hilo":
                    by = n_el * 4.0 + a[1].numel() * 2.0
                nm = f"{nm}:{x.shape[1]}x{x.shape[2]}x{x.shape[3]}"
            rows.append((nm, e0.elapsed_time(e1), fl, by))
    return rows, plan


def time_spectral_and_pc(model, name, B, dev, latent):
    """{kernel: (ms per launch, algorithmic bytes)} for STFT+compress, expand+iSTFT and the PC updates."""
    import torch

    from ditsep_b200 import ops
    from ditsep_b200._lib import SdeParams

    c = CONFIGS[name]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    out = {}

    def timeit(fn, reps=5):
        tot = 0.0
        for i in range(reps + 2):
            _flush_l2(flush)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            if i >= 2:
                tot += e0.elapsed_time(e1)
        return tot / reps

    if latent:
        L = 64 * c["L"]
        x = torch.randn(B, 2, L, device=dev); s = torch.randn(B, 2, L, device=dev)
        y = torch.randn(B, 1, L, device=dev); t = torch.full((B,), 0.5, device=dev)
        p = SdeParams(1, 2, 1.5, 0.96, 10.0, 30, 0.5)
    else:
        T = c["T"]
        sm = model.score_model
        x = torch.randn(B, 2, T, device=dev); s = torch.randn(B, 2, T, device=dev)
        y = torch.randn(B, 1, T, device=dev); t = torch.full((B,), 0.5, device=dev)
        p = SdeParams(0, 2, 2.0, 0.05, 0.5, 30, 0.5)
        plan = max(sm.backbone.engine()._plans.values(), key=lambda q: q.B * q.H * q.W)
        W = plan.W
        e_, f_ = sm._exp_factor()
        res = torch.empty(B, 2, T, device=dev)
        ms = timeit(lambda: ops.stft_compress(x, y, sm.stft.window, plan.feat, e_, f_, True))
        out[f"stft_compress:{T}"] = (ms, B * (3.0 * T * 4 + 6.0 * 256 * W * plan.feat.element_size()))
        ms = timeit(lambda: ops.expand_istft(plan.out, sm.stft_inv.window, res, e_, f_))
        out[f"expand_istft:{T}"] = (ms, B * (4.0 * 256 * W * 4 + 2.0 * T * 4))
    L_ = x[0, 0].numel()
    ms = timeit(lambda: ops.pc_corrector(p, x, s, t, seed=1, draw=1))
    out["pc_update:corrector"] = (ms, B * 3.0 * 2 * L_ * 4)
    ms = timeit(lambda: ops.pc_predictor(p, x, s, t, y, seed=1, draw=2))
    out["pc_update:predictor"] = (ms, B * 3.0 * 2 * L_ * 4)
    return out


# ----------------------------------------------------------------------------------- library bar (torch + cuDNN)
def library_baseline(name, B, dev):
    """The oracle's torch code (oracle/ncsnpp_ref.py: F.conv2d, group_norm, torch.stft ...) run on the SAME GPU: torch
    eager over cuDNN / cuBLAS / cuFFT, in TF32 (the reference's default arithmetic on a GPU) and under bf16 autocast,
    at batch 1 and at this run's batch.  ms per score step; nothing of it is on the product path."""
    import torch

    from oracle import ncsnpp_ref as R

    c = CONFIGS[name]
    torch.backends.cudnn.benchmark = True
    torch.backends.cudnn.allow_tf32 = True
    torch.backends.cuda.matmul.allow_tf32 = True
    res = {"kind": "torch eager + cuDNN/cuBLAS/cuFFT (the oracle's code on cuda), one score-network evaluation",
           "ms_per_score_step": {}}
    try:
        if c["kind"] == "latent":
            bcfg = R.NCSNppConfig(nf=128, ch_mult=(1, 2, 2), num_res_blocks=2, attn_resolutions=(16,), image_size=64,
                                  num_channels_in=3, num_channels_out=2, centered=True)
            sd = {k: v.to(dev) for k, v in R.make_state_dict(bcfg, seed=5, with_window=None).items()}
            fwd = lambda xt, t, mx: R.latent_score_forward_ref(sd, bcfg, xt, t, mx, max_latent_length=4)  # noqa: E731
            shape_x, shape_m = (2, 64, c["L"]), (1, 64, c["L"])
        else:
            cfg = R.ScoreConfig()
            sd = {k: v.to(dev) for k, v in R.make_state_dict(cfg.backbone, seed=0).items()}
            def fwd(xt, t, mx):  # score_forward_ref, with the STFT ends in fp32 (torch.complex has no bf16) and only
                feat, n_s, n_pad = R.pre_process_ref(torch.cat((xt, mx), dim=1), cfg, sd["stft.window"])  # the backbone
                out = R.ncsnpp_forward_ref(sd, cfg.backbone, feat, t, "backbone.")                        # autocast
                with torch.autocast("cuda", enabled=False):
                    return R.post_process_ref(out.float(), n_s, n_pad, cfg, sd["stft.window"])
            shape_x, shape_m = (2, c["T"]), (1, c["T"])
        for b in sorted({1, B}):
            xt = torch.randn(b, *shape_x, device=dev)
            mx = torch.randn(b, *shape_m, device=dev)
            t = torch.full((b,), 0.5, device=dev)
            for mode in ("tf32", "bf16_autocast"):
                def call():
                    with torch.no_grad():
                        if mode == "tf32":
                            return fwd(xt, t, mx)
                        with torch.autocast("cuda", dtype=torch.bfloat16):
                            return fwd(xt, t, mx)
                for _ in range(3):
                    call()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                reps = 5
                e0.record()
                for _ in range(reps):
                    call()
                e1.record()
                torch.cuda.synchronize()
                res["ms_per_score_step"][f"B{b}_{mode}"] = e0.elapsed_time(e1) / reps
            del xt, mx
            torch.cuda.empty_cache()
        best = min(v for k, v in res["ms_per_score_step"].items() if k.startswith(f"B{B}_"))
        res["value"] = B / (best * 1e-3 * NFE)
        res["unit"] = "mixtures/s (60 score evaluations per mixture; sampler updates not counted)"
    except Exception as ex:  # the bar is a report, never a reason to lose the bench line
        res["error"] = f"{type(ex).__name__}: {ex}"[:300]
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C2", choices=sorted(CONFIGS))
    ap.add_argument("--batch", type=int, default=0, help="mixtures per GPU per step (default: the config's)")
    ap.add_argument("--precision", default=None, choices=["bf16", "fp32"])
    ap.add_argument("--sweep-mixtures", type=int, default=256, help="C5: size of the fixed mixture set")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-library-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
