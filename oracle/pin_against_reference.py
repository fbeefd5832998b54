"""Pins the oracle against the REAL reference and writes the golden fixtures.

TEST INFRASTRUCTURE.  Runs only in the build container (needs /root/reference);
``python -m oracle.pin_against_reference [--full]``.

The reference's hot path is imported unmodified from /root/reference/src with three
import shims injected into ``sys.modules`` (no reference file is touched):
  1. ``hydra.utils.instantiate`` -- hydra is not installed; the only hot-path use is
     score_models.py:27 (build the backbone from ``_target_``).
  2. an empty ``utils`` package whose ``__path__`` is src/utils, so that
     ``from utils.registry import Registry`` (sdes.py:12) works without executing
     src/utils/__init__.py (which pulls Lightning + stable_audio_tools).
  3. ``torch.utils.cpp_extension.load`` returns a dummy during import, skipping the
     108 s JIT build of the two StyleGAN2 CUDA ops (op/upfirdn2d.py:10-16); on CPU
     tensors the reference takes ``upfirdn2d_native`` and never touches them.

What is pinned (max abs / rel-L2 printed, asserted):
  a. state_dict key names + shapes of ``make_state_dict`` == reference module's.
  b. one score evaluation (ScoreModelNCSNpp.forward) -- small and C1 shapes.
  c. the full PC sampler (reverse_diffusion + ald2, MixSDE) on taped noise.
  d. OUVESDE + "ald" sampler with a synthetic score function.
  e. latent wrapper (LatentScoreModelNCSNpp, centered=True).
Fixtures written to tests/golden/ are OUTPUTS OF THE REFERENCE (not of the oracle).
"""
from __future__ import annotations

import argparse
import importlib
import os
import sys
import time
import types

import numpy as np
import torch

REF_SRC = "/root/reference/src"
GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def import_reference():
    if not os.path.isdir(REF_SRC):
        raise RuntimeError("reference tree not present; this script only runs in the build container")

    def instantiate(cfg, *args, **kwargs):
        cfg = dict(cfg)
        cfg.pop("_recursive_", None)
        kwargs.pop("_recursive_", None)
        target = cfg.pop("_target_")
        mod, name = target.rsplit(".", 1)
        return getattr(importlib.import_module(mod), name)(*args, **cfg, **kwargs)

    hydra = types.ModuleType("hydra")
    hydra_utils = types.ModuleType("hydra.utils")
    hydra_utils.instantiate = instantiate
    hydra.utils = hydra_utils
    sys.modules.setdefault("hydra", hydra)
    sys.modules.setdefault("hydra.utils", hydra_utils)

    utils_pkg = types.ModuleType("utils")
    utils_pkg.__path__ = [os.path.join(REF_SRC, "utils")]
    sys.modules["utils"] = utils_pkg

    import torch.utils.cpp_extension as cpp_ext
    real_load = cpp_ext.load
    cpp_ext.load = lambda *a, **k: types.SimpleNamespace()
    sys.path.insert(0, REF_SRC)
    try:
        sdes = importlib.import_module("sdes")
        score_models = importlib.import_module("models.diffsep.score_models")
    finally:
        cpp_ext.load = real_load
    return sdes, score_models


def rel_l2(a, b):
    return float((a - b).norm() / b.norm().clamp(min=1e-30))


class TapedRandn:
    """Replaces torch.randn_like / torch.randn inside the reference sampler by tape draws."""

    def __init__(self, tape):
        self.tape, self.i = tape, 0

    def __enter__(self):
        self._rl, self._r = torch.randn_like, torch.randn
        torch.randn_like = lambda x, **k: self._next(x.shape)
        torch.randn = lambda *s, **k: self._next(s[0] if len(s) == 1 and not isinstance(s[0], int) else s)
        return self

    def _next(self, shape):
        z = self.tape.draw(self.i, shape)
        self.i += 1
        return z

    def __exit__(self, *a):
        torch.randn_like, torch.randn = self._rl, self._r


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--full", action="store_true", help="also run the full N=30 C1 sampler (~3 min)")
    ap.add_argument("--init-scale", type=float, default=1.0,
                    help="scale of the reference's init_scale=0 tensors (Conv_1, NIN_3, pyramid convs); 1.0 = worst case "
                         "for rounding-error growth, 0.01 = 'trained-like' set used by the bf16 parity tests")
    args = ap.parse_args()
    tag = "" if args.init_scale == 1.0 else f"_s{args.init_scale:g}".replace(".", "p")

    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from oracle import ncsnpp_ref as R
    from oracle import sampler_ref as S
    from ditsep_b200.utils.synth import synth_mixtures

    torch.manual_seed(0)
    torch.set_num_threads(os.cpu_count())
    sdes, score_models = import_reference()
    os.makedirs(GOLDEN, exist_ok=True)

    stft_args = dict(n_fft=510, hop_length=128, center=True, pad_mode="constant")

    def build_ref(nf, num_sources=2, **bk):
        backbone_args = dict(_target_="models.diffsep.ncsnpp.NCSNpp", nf=nf, ch_mult=[1, 1, 2, 2, 2, 2, 2],
                             num_res_blocks=2, attn_resolutions=[16], resamp_with_conv=True, image_size=256)
        backbone_args.update(bk)
        return score_models.ScoreModelNCSNpp(num_sources, dict(stft_args), backbone_args, transform="exponent",
                                             spec_abs_exponent=0.5, spec_factor=0.15).eval()

    # ---- a. key layout -------------------------------------------------------------
    for nf, nsrc in ((64, 2), (128, 2), (64, 3)):
        ref = build_ref(nf, nsrc)
        cfg = R.ScoreConfig(num_sources=nsrc, backbone=R.NCSNppConfig(nf=nf))
        sd = R.make_state_dict(cfg.backbone, seed=0)
        ref_sd = ref.state_dict()
        assert set(sd) == set(ref_sd), (sorted(set(sd) ^ set(ref_sd))[:10])
        for k in sd:
            assert tuple(sd[k].shape) == tuple(ref_sd[k].shape), (k, sd[k].shape, ref_sd[k].shape)
        assert torch.allclose(sd["stft.window"], ref_sd["stft.window"], atol=1e-7)
        print(f"[a] nf={nf} nsrc={nsrc}: {len(sd)} keys match (names + shapes)")

    # ---- b. one score evaluation ---------------------------------------------------
    cfg = R.ScoreConfig()
    sd = R.make_state_dict(cfg.backbone, seed=0, init_scale=args.init_scale)
    ref = build_ref(64)
    ref.load_state_dict(sd, strict=True)
    golden = {}
    for name, T, B in (("small", 7810, 2), ("c1", 32000, 1)):
        mix, src = synth_mixtures(B, 2, T, seed=1234)
        mixn, _, _ = S.normalize_batch_ref(mix)
        tape = S.NoiseTape(99)
        xt = 0.5 * mixn + 0.3 * tape.draw(0, (B, 2, T))
        t = torch.full((B,), 0.61)
        with torch.no_grad():
            t0 = time.perf_counter()
            out_ref = ref(xt, t, mixn)
            t_ref = time.perf_counter() - t0
            trace = {}
            out_orc = R.score_forward_ref(sd, cfg, xt, t, mixn, trace)
            feat_ref, _, _ = ref.pre_process(torch.cat((xt, mixn), dim=1))
        e = rel_l2(out_orc, out_ref)
        ef = rel_l2(trace["features"], feat_ref)
        print(f"[b] score step {name}: rel-L2 oracle vs reference = {e:.3e} (features {ef:.3e}); "
              f"reference {t_ref:.2f} s; |out| rms {float(out_ref.pow(2).mean().sqrt()):.3f}")
        assert e < 2e-5 and ef < 1e-6
        if name == "small":
            golden["score_small"] = out_ref.numpy()
            golden["feat_small_sub"] = feat_ref[:, :, ::16, ::4].numpy()
        else:
            golden["score_c1_sub"] = out_ref[..., ::16].numpy()
            golden["score_c1_norm"] = np.array(float(out_ref.norm()))

    # ---- b2. what bf16 arithmetic does to the REFERENCE itself ------------------------
    # The reference under torch.autocast(bfloat16) on the same inputs, against its own fp32 output: the size of the
    # error that is intrinsic to bf16 operands / activations for this weight set, independent of this repo's kernels.
    mix, src = synth_mixtures(2, 2, 7810, seed=1234)
    mixn, _, _ = S.normalize_batch_ref(mix)
    xt = 0.5 * mixn + 0.3 * S.NoiseTape(99).draw(0, (2, 2, 7810))
    t = torch.full((2,), 0.61)
    with torch.no_grad():
        out32 = ref(xt, t, mixn)
        with torch.autocast("cpu", dtype=torch.bfloat16):
            out16 = ref(xt, t, mixn)
    e_ac = rel_l2(out16.float(), out32)
    print(f"[b2] reference under torch.autocast(cpu, bfloat16) vs its own fp32 output (init_scale={args.init_scale:g}): "
          f"rel-L2 = {e_ac:.3e}")
    golden["ref_autocast_bf16_rel_l2"] = np.array(e_ac)

    # ---- b3. C3 at its maximum: 10 s @ 16 kHz, W = 1280 frames -------------------------
    if args.full:
        T3 = 160000
        mix, src = synth_mixtures(1, 2, T3, seed=1234)
        mixn, _, _ = S.normalize_batch_ref(mix)
        xt = 0.5 * mixn + 0.3 * S.NoiseTape(99).draw(0, (1, 2, T3))
        t = torch.full((1,), 0.61)
        with torch.no_grad():
            t0 = time.perf_counter()
            out_ref = ref(xt, t, mixn)
            t_ref = time.perf_counter() - t0
            out_orc = R.score_forward_ref(sd, cfg, xt, t, mixn)
        print(f"[b3] score step C3 max (T={T3}, W=1280): rel-L2 oracle vs reference = {rel_l2(out_orc, out_ref):.3e}; "
              f"reference {t_ref:.2f} s")
        assert rel_l2(out_orc, out_ref) < 2e-5
        golden["score_c3max_sub"] = out_ref[..., ::64].numpy()

    # ---- c. PC sampler, MixSDE + ald2, taped noise ---------------------------------
    def run_pair(T, B, N, tag):
        mix, src = synth_mixtures(B, 2, T, seed=1234)
        mixn, _, _ = S.normalize_batch_ref(mix)
        sde_ref = sdes.sdes.MixSDE(2, 2.0, 0.05, 0.5, N=N)
        tape = S.NoiseTape(4321)
        ref_scores_sub = []  # the REFERENCE's score at every evaluation, every 64th sample (per-step rel-L2 gate)

        def ref_recording(x, t, y):
            out = ref(x, t, y)
            ref_scores_sub.append(out[..., ::64].clone())
            return out

        with TapedRandn(tape) as tr:
            sampler = sdes.get_pc_sampler("reverse_diffusion", "ald2", sde=sde_ref, score_fn=ref_recording, y=mixn,
                                          eps=0.03, snr=0.5, corrector_steps=1, denoise=True)
            t0 = time.perf_counter()
            x_ref, nfe = sampler()
            t_ref = time.perf_counter() - t0
            ndraws = tr.i
        scores, orc_sub = [], []
        x_orc, nfe2 = S.pc_sampler_ref(lambda x, t, y: R.score_forward_ref(sd, cfg, x, t, y),
                                       S.MixSDERef(2, 2.0, 0.05, 0.5, N=N), mixn, S.NoiseTape(4321),
                                       corrector="ald2", eps=0.03, snr=0.5,
                                       on_score=lambda k, x, t, s: (scores.append(float(s.norm())),
                                                                    orc_sub.append(s[..., ::64].clone())))
        worst = max(rel_l2(a, b) for a, b in zip(orc_sub, ref_scores_sub))
        print(f"[c] sampler {tag}: worst per-step score rel-L2 oracle vs reference over {len(orc_sub)} evaluations = {worst:.3e}")
        assert len(orc_sub) == len(ref_scores_sub) == 2 * N and worst < 1e-3
        e = rel_l2(x_orc, x_ref)
        sis_ref, _ = S.si_sdr_pit_ref(src, x_ref)
        sis_orc, _ = S.si_sdr_pit_ref(src, x_orc)
        print(f"[c] sampler {tag}: N={N} nfe={nfe} draws={ndraws} rel-L2={e:.3e} reference {t_ref:.1f} s "
              f"SI-SDR ref {sis_ref.mean():.3f} dB oracle {sis_orc.mean():.3f} dB")
        assert nfe == nfe2 == 2 * N and ndraws == 1 + 2 * N
        assert e < 1e-3 and abs(float(sis_ref.mean() - sis_orc.mean())) < 0.01
        # the same reference sampler under torch.autocast(bfloat16), same tape: how far bf16 arithmetic moves the
        # REFERENCE's own trajectory (per-step score rel-L2 against its fp32 run, SI-SDR shift).  This is the yardstick
        # for the bf16 trajectory tests on weight sets where bf16 cannot meet 1e-2 (errors compound over 2N steps).
        fp32_sub = list(ref_scores_sub)
        ref_scores_sub.clear()
        with TapedRandn(S.NoiseTape(4321)), torch.autocast("cpu", dtype=torch.bfloat16):
            sampler = sdes.get_pc_sampler("reverse_diffusion", "ald2", sde=sdes.sdes.MixSDE(2, 2.0, 0.05, 0.5, N=N),
                                          score_fn=lambda x, t, y: ref_recording(x.float(), t, y).float(), y=mixn,
                                          eps=0.03, snr=0.5, corrector_steps=1, denoise=True)
            x_ac, _ = sampler()
        ac_steps = [rel_l2(a.float(), b) for a, b in zip(ref_scores_sub, fp32_sub)]
        sis_ac, _ = S.si_sdr_pit_ref(src, x_ac.float())
        ac = np.array([max(ac_steps), abs(float(sis_ac.mean() - sis_ref.mean())), rel_l2(x_ac.float(), x_ref)])
        print(f"[c] sampler {tag}: reference under bf16 autocast vs its own fp32 run: worst per-step score rel-L2 "
              f"{ac[0]:.3e}, |dSI-SDR| {ac[1]:.3f} dB, final rel-L2 {ac[2]:.3e}")
        return x_ref, scores, sis_ref, torch.stack(fp32_sub).numpy(), ac

    x_ref, scores, sis, sub, ac = run_pair(7810, 2, 3, "small")
    golden["sampler_small_out"] = x_ref.numpy()
    golden["sampler_small_scores_sub"] = sub
    golden["sampler_small_autocast"] = ac  # [worst per-step rel-L2, |dSI-SDR| dB, final rel-L2] of the reference in bf16
    golden["sampler_small_score_norms"] = np.array(scores)
    golden["sampler_small_sisdr"] = sis.numpy()
    if args.full:
        x_ref, scores, sis, sub, ac = run_pair(32000, 1, 30, "C1")
        np.savez_compressed(os.path.join(GOLDEN, f"sampler_c1{tag}.npz"), out=x_ref.numpy().astype(np.float32),
                            score_norms=np.array(scores), sisdr=sis.numpy(), scores_sub=sub.astype(np.float32),
                            autocast=ac)

    # ---- d. OUVESDE + ald with a synthetic score function ---------------------------
    B, T, N = 2, 4096, 5
    y = torch.randn(B, 1, T, generator=torch.Generator().manual_seed(5))

    def toy_score(x, t, yy):
        return -(x - 0.5 * yy) / (0.1 + t[:, None, None]) + 0.05 * torch.sin(3 * x)

    sde_ref = sdes.sdes.OUVESDE(1.5, 0.05, 0.5, N=N)
    with TapedRandn(S.NoiseTape(77)):
        x_ref, _ = sdes.get_pc_sampler("reverse_diffusion", "ald", sde=sde_ref, score_fn=toy_score, y=y, eps=0.03,
                                       snr=0.5, corrector_steps=1, denoise=True, n_spkrs=2)()
    x_orc, _ = S.pc_sampler_ref(toy_score, S.OUVESDERef(1.5, 0.05, 0.5, N=N), y, S.NoiseTape(77), corrector="ald")
    print(f"[d] OUVE/ald toy sampler: max abs err {float((x_orc - x_ref).abs().max()):.3e}")
    assert float((x_orc - x_ref).abs().max()) < 1e-4
    golden["ouve_toy_out"] = x_ref.numpy()
    # MixSDE + ald2 with the same toy score (pins the SDE algebra without the network)
    sde_ref = sdes.sdes.MixSDE(2, 2.0, 0.05, 0.5, N=N)
    with TapedRandn(S.NoiseTape(78)):
        x_ref, _ = sdes.get_pc_sampler("reverse_diffusion", "ald2", sde=sde_ref, score_fn=toy_score, y=y, eps=0.03,
                                       snr=0.5, corrector_steps=1, denoise=True)()
    x_orc, _ = S.pc_sampler_ref(toy_score, S.MixSDERef(2, 2.0, 0.05, 0.5, N=N), y, S.NoiseTape(78), corrector="ald2")
    print(f"[d] Mix/ald2 toy sampler: max abs err {float((x_orc - x_ref).abs().max()):.3e}")
    assert float((x_orc - x_ref).abs().max()) < 1e-4
    golden["mix_toy_out"] = x_ref.numpy()

    # ---- e. latent wrapper ------------------------------------------------------------
    bcfg = R.NCSNppConfig(nf=32, ch_mult=(1, 2, 2), num_res_blocks=2, attn_resolutions=(16,), image_size=64,
                          num_channels_in=3, num_channels_out=2, centered=True)
    sdl = R.make_state_dict(bcfg, seed=3, with_window=None)
    lat = score_models.LatentScoreModelNCSNpp(
        2, dict(_target_="models.diffsep.ncsnpp.NCSNpp", nf=32, ch_mult=[1, 2, 2], num_res_blocks=2,
                attn_resolutions=[16], resamp_with_conv=True, image_size=64, centered=True),
        max_latent_length=4).eval()
    lat.load_state_dict(sdl, strict=True)
    g = torch.Generator().manual_seed(11)
    xt = torch.randn(2, 2, 64, 14, generator=g)
    mx = torch.randn(2, 1, 64, 14, generator=g)
    t = torch.tensor([0.4, 0.9])
    with torch.no_grad():
        o_ref = lat(xt, t, mx)
        o_orc = R.latent_score_forward_ref(sdl, bcfg, xt, t, mx, max_latent_length=4)
    print(f"[e] latent wrapper: rel-L2 {rel_l2(o_orc, o_ref):.3e}")
    assert rel_l2(o_orc, o_ref) < 2e-5
    golden["latent_out"] = o_ref.numpy()

    # ---- e2. latent wrapper at the real C4 size (nf=128, ch_mult [1,2,2], L = 16 and 80, OUVE latent config) --------
    if args.full:
        bcfg4 = R.NCSNppConfig(nf=128, ch_mult=(1, 2, 2), num_res_blocks=2, attn_resolutions=(16,), image_size=64,
                               num_channels_in=3, num_channels_out=2, centered=True)
        sd4 = R.make_state_dict(bcfg4, seed=5, with_window=None, init_scale=args.init_scale)
        lat4 = score_models.LatentScoreModelNCSNpp(
            2, dict(_target_="models.diffsep.ncsnpp.NCSNpp", nf=128, ch_mult=[1, 2, 2], num_res_blocks=2,
                    attn_resolutions=[16], resamp_with_conv=True, image_size=64, centered=True),
            max_latent_length=4).eval()
        lat4.load_state_dict(sd4, strict=True)
        for L in (16, 80):
            g = torch.Generator().manual_seed(100 + L)
            xt = torch.randn(1, 2, 64, L, generator=g)
            mx = torch.randn(1, 1, 64, L, generator=g)
            t = torch.tensor([0.55])
            with torch.no_grad():
                o_ref = lat4(xt, t, mx)
                o_orc = R.latent_score_forward_ref(sd4, bcfg4, xt, t, mx, max_latent_length=4)
            print(f"[e2] latent wrapper nf=128 L={L}: rel-L2 oracle vs reference {rel_l2(o_orc, o_ref):.3e}")
            assert rel_l2(o_orc, o_ref) < 2e-5
            golden[f"latent_c4_L{L}"] = o_ref.numpy()

    np.savez_compressed(os.path.join(GOLDEN, f"reference_outputs{tag}.npz"), **golden)
    print("fixtures written to", GOLDEN)


if __name__ == "__main__":
    main()
