"""How chaotic is the bf16 trajectory of the worst-case weight set?  Measured on the REAL reference (TEST
INFRASTRUCTURE; build container only).

``python -m oracle.pin_bf16_spread [--c1]`` runs the unmodified reference PC sampler (N=3, T=7810, B=2, the ``small`` case of
pin_against_reference.py, init_scale 1.0) under ``torch.autocast(bfloat16)`` K times.  Run 0 is the plain autocast
run; in runs 1..K-1 the input of every score evaluation is multiplied by (1 + 2^-9 u), u uniform in [-1, 1] and fixed
per run: half a bf16 ulp at ONE tensor per evaluation, i.e. LESS than the difference between two bf16 implementations
of the network, which round differently at every layer.  For each run the final SI-SDR is compared with the
reference's own fp32 run (tests/golden/reference_outputs.npz).  The spread of that one number across runs is what a
bf16 trajectory test on this weight set can resolve; tests/test_gpu_parity.py::_traj_tol takes its SI-SDR gate from the
file written here (tests/golden/bf16_spread.npz) instead of from a single draw."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ncsnpp_ref as R  # noqa: E402
from oracle import sampler_ref as S  # noqa: E402
from oracle.pin_against_reference import GOLDEN, TapedRandn, import_reference, rel_l2  # noqa: E402


def spread(sdes, ref, T, B, N, K, x_fp32, sis_fp32, tag):
    from ditsep_b200.utils.synth import synth_mixtures
    mix, src = synth_mixtures(B, 2, T, seed=1234)
    mixn, _, _ = S.normalize_batch_ref(mix)
    d_db, e_fin = [], []
    for k in range(K):
        u = None if k == 0 else (torch.rand(B, 2, T, generator=torch.Generator().manual_seed(500 + k)) * 2 - 1)

        def score(x, t, y):
            x = x.float() if u is None else x.float() * (1.0 + 2.0 ** -9 * u)
            return ref(x, t, y).float()

        with torch.no_grad(), TapedRandn(S.NoiseTape(4321)), torch.autocast("cpu", dtype=torch.bfloat16):
            x, _ = sdes.get_pc_sampler("reverse_diffusion", "ald2", sde=sdes.sdes.MixSDE(2, 2.0, 0.05, 0.5, N=N),
                                       score_fn=score, y=mixn, eps=0.03, snr=0.5, corrector_steps=1, denoise=True)()
        sis, _ = S.si_sdr_pit_ref(src, x.float())
        d_db.append(abs(float(sis.mean()) - sis_fp32))
        e_fin.append(rel_l2(x.float(), x_fp32))
        print(f"{tag} run {k} ({'plain autocast' if k == 0 else 'input x (1 + 2^-9 u)'}): |dSI-SDR| vs the fp32 reference "
              f"{d_db[-1]:.4f} dB, final rel-L2 {e_fin[-1]:.3e}", flush=True)
    print(f"reference bf16, {tag}: |dSI-SDR| over {K} runs: min {min(d_db):.3f} max {max(d_db):.3f} "
          f"rms {float(np.sqrt(np.mean(np.square(d_db)))):.3f} dB", flush=True)
    return np.array(d_db), np.array(e_fin)


def main():
    torch.set_num_threads(os.cpu_count())
    sdes, score_models = import_reference()
    cfg = R.ScoreConfig()
    sd = R.make_state_dict(cfg.backbone, seed=0, init_scale=1.0)
    ref = score_models.ScoreModelNCSNpp(
        2, dict(n_fft=510, hop_length=128, center=True, pad_mode="constant"),
        dict(_target_="models.diffsep.ncsnpp.NCSNpp", nf=64, ch_mult=[1, 1, 2, 2, 2, 2, 2], num_res_blocks=2,
             attn_resolutions=[16], resamp_with_conv=True, image_size=256),
        transform="exponent", spec_abs_exponent=0.5, spec_factor=0.15).eval()
    ref.load_state_dict(sd, strict=True)
    path = os.path.join(GOLDEN, "bf16_spread.npz")
    out = dict(np.load(path)) if os.path.exists(path) else {}
    if "--c1" in sys.argv:  # the C1 trajectory (N=30, T=32000): ~10 minutes per run on 16 cores
        g = np.load(os.path.join(GOLDEN, "sampler_c1.npz"))
        d, e = spread(sdes, ref, 32000, 1, 30, 4, torch.tensor(g["out"]), float(torch.tensor(g["sisdr"]).mean()), "C1")
        out.update(c1_dsisdr_db=d, c1_final_rel_l2=e)
    else:
        g = np.load(os.path.join(GOLDEN, "reference_outputs.npz"))
        d, e = spread(sdes, ref, 7810, 2, 3, 6, torch.tensor(g["sampler_small_out"]),
                      float(torch.tensor(g["sampler_small_sisdr"]).mean()), "small")
        out.update(small_dsisdr_db=d, small_final_rel_l2=e)
    np.savez(path, **out)
    print("written", path)


if __name__ == "__main__":
    main()
