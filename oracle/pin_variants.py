"""Pins the sampler variants of SURVEY.md 8f-2 against the REAL reference (TEST INFRASTRUCTURE; build container only).

``python -m oracle.pin_variants`` runs the unmodified reference ``get_pc_scheduled_sampler`` (src/sdes/__init__.py:49-130)
for the three schedules with a closed-form score and a noise tape, asserts the oracle against it and writes the
REFERENCE's outputs to tests/golden/variants.npz."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import sampler_ref as S  # noqa: E402
from oracle.pin_against_reference import GOLDEN, TapedRandn, import_reference  # noqa: E402


def toy_score(x, t, yy):
    return -(x - 0.5 * yy) / (0.1 + t[:, None, None]) + 0.05 * torch.sin(3 * x)


def main():
    sdes, _ = import_reference()
    B, T, N = 2, 4096, 6
    y = torch.randn(B, 1, T, generator=torch.Generator().manual_seed(5))
    golden = {}
    for schedule in ("linear", "log", "revlog"):
        with TapedRandn(S.NoiseTape(80)):
            x_ref, nfe = sdes.get_pc_scheduled_sampler("reverse_diffusion", "ald2", sde=sdes.sdes.MixSDE(2, 2.0, 0.05, 0.5, N=N),
                                                       score_fn=toy_score, y=y, eps=0.03, snr=0.5, corrector_steps=1,
                                                       denoise=True, schedule=schedule)()
        x_orc, nfe2 = S.pc_sampler_ref(toy_score, S.MixSDERef(2, 2.0, 0.05, 0.5, N=N), y, S.NoiseTape(80), corrector="ald2",
                                       eps=0.03, snr=0.5, schedule=schedule)
        err = float((x_orc - x_ref).abs().max())
        print(f"scheduled sampler '{schedule}': nfe {nfe}, max abs err oracle vs reference {err:.3e}")
        assert nfe == nfe2 == 2 * N and err < 1e-4
        golden[f"scheduled_{schedule}"] = x_ref.numpy()
    # f2, second half: euler_maruyama predictor, langevin corrector, PriorMixSDE (registry name "priormix")
    cases = {
        "em_ald2_mix": ("euler_maruyama", "ald2", lambda: sdes.sdes.MixSDE(2, 2.0, 0.05, 0.5, N=N), lambda: S.MixSDERef(2, 2.0, 0.05, 0.5, N=N)),
        "rd_langevin_mix": ("reverse_diffusion", "langevin", lambda: sdes.sdes.MixSDE(2, 2.0, 0.05, 0.5, N=N), lambda: S.MixSDERef(2, 2.0, 0.05, 0.5, N=N)),
        "rd_ald2_priormix": ("reverse_diffusion", "ald2", lambda: sdes.sdes.PriorMixSDE(2, 2.0, 0.05, 0.5, N=N), lambda: S.PriorMixSDERef(2, 2.0, 0.05, 0.5, N=N)),
        "em_ald2_priormix_avg64": ("euler_maruyama", "ald2", lambda: sdes.sdes.PriorMixSDE(2, 2.0, 0.05, 0.5, N=N, avg_len=64), lambda: S.PriorMixSDERef(2, 2.0, 0.05, 0.5, N=N, avg_len=64)),
    }
    import warnings
    for name, (pred, corr, mk_ref, mk_orc) in cases.items():
        with TapedRandn(S.NoiseTape(81)), warnings.catch_warnings():
            warnings.simplefilter("ignore")
            x_ref, nfe = sdes.get_pc_sampler(pred, corr, sde=mk_ref(), score_fn=toy_score, y=y, eps=0.03, snr=0.5,
                                             corrector_steps=1, denoise=True)()
        x_orc, nfe2 = S.pc_sampler_ref(toy_score, mk_orc(), y, S.NoiseTape(81), corrector=corr, eps=0.03, snr=0.5, predictor=pred)
        err = float((x_orc - x_ref).abs().max())
        print(f"{name}: nfe {nfe}, max abs err oracle vs reference {err:.3e}, |x| max {float(x_ref.abs().max()):.3f}")
        assert nfe == nfe2 == 2 * N and err < 1e-4
        golden[name] = x_ref.numpy()
    golden["variants_y"] = y.numpy()
    np.savez_compressed(os.path.join(GOLDEN, "variants.npz"), **golden)
    print("written", os.path.join(GOLDEN, "variants.npz"))


if __name__ == "__main__":
    main()
