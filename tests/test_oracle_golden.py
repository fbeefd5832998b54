"""The oracle against the committed outputs of the REFERENCE (-m "not gpu").

tests/golden/*.npz were written by oracle/pin_against_reference.py from the unmodified
reference code (/root/reference, imported in the build container).  The reference ships no
tests or golden vectors for this path, so these fixtures are the pin."""
import numpy as np
import pytest
import torch

from common import rel_l2, sampler_inputs, score_inputs
from oracle import ncsnpp_ref as R
from oracle import sampler_ref as S


@pytest.fixture(scope="module")
def weights():
    cfg = R.ScoreConfig()
    return cfg, R.make_state_dict(cfg.backbone, seed=0)


def test_oracle_score_step_small(golden, weights):
    cfg, sd = weights
    xt, t, mixn = score_inputs(2, 7810)
    trace = {}
    with torch.no_grad():
        out = R.score_forward_ref(sd, cfg, xt, t, mixn, trace)
    assert rel_l2(out, torch.tensor(golden["ref"]["score_small"])) < 2e-5
    assert rel_l2(trace["features"][:, :, ::16, ::4], torch.tensor(golden["ref"]["feat_small_sub"])) < 1e-6


def test_oracle_score_step_c1(golden, weights):
    cfg, sd = weights
    xt, t, mixn = score_inputs(1, 32000)
    with torch.no_grad():
        out = R.score_forward_ref(sd, cfg, xt, t, mixn)
    assert rel_l2(out[..., ::16], torch.tensor(golden["ref"]["score_c1_sub"])) < 2e-5
    assert abs(float(out.norm()) / float(golden["ref"]["score_c1_norm"]) - 1) < 1e-5


def test_oracle_sampler_small(golden, weights):
    cfg, sd = weights
    mixn, src = sampler_inputs(2, 7810)
    norms = []
    x, nfe = S.pc_sampler_ref(lambda x, t, y: R.score_forward_ref(sd, cfg, x, t, y), S.MixSDERef(2, 2.0, 0.05, 0.5, N=3),
                              mixn, S.NoiseTape(4321), corrector="ald2", eps=0.03, snr=0.5,
                              on_score=lambda k, x, t, s: norms.append(float(s.norm())))
    assert nfe == 6
    assert rel_l2(x, torch.tensor(golden["ref"]["sampler_small_out"])) < 1e-4
    assert np.allclose(norms, golden["ref"]["sampler_small_score_norms"], rtol=1e-4)


def test_oracle_toy_samplers(golden):
    B, T, N = 2, 4096, 5
    y = torch.randn(B, 1, T, generator=torch.Generator().manual_seed(5))

    def toy_score(x, t, yy):
        return -(x - 0.5 * yy) / (0.1 + t[:, None, None]) + 0.05 * torch.sin(3 * x)

    x, _ = S.pc_sampler_ref(toy_score, S.OUVESDERef(1.5, 0.05, 0.5, N=N), y, S.NoiseTape(77), corrector="ald")
    assert float((x - torch.tensor(golden["ref"]["ouve_toy_out"])).abs().max()) < 1e-4
    x, _ = S.pc_sampler_ref(toy_score, S.MixSDERef(2, 2.0, 0.05, 0.5, N=N), y, S.NoiseTape(78), corrector="ald2")
    assert float((x - torch.tensor(golden["ref"]["mix_toy_out"])).abs().max()) < 1e-4


def test_oracle_latent(golden):
    bcfg = R.NCSNppConfig(nf=32, ch_mult=(1, 2, 2), num_res_blocks=2, attn_resolutions=(16,), image_size=64,
                          num_channels_in=3, num_channels_out=2, centered=True)
    sdl = R.make_state_dict(bcfg, seed=3, with_window=None)
    g = torch.Generator().manual_seed(11)
    xt = torch.randn(2, 2, 64, 14, generator=g)
    mx = torch.randn(2, 1, 64, 14, generator=g)
    with torch.no_grad():
        out = R.latent_score_forward_ref(sdl, bcfg, xt, torch.tensor([0.4, 0.9]), mx, max_latent_length=4)
    assert rel_l2(out, torch.tensor(golden["ref"]["latent_out"])) < 2e-5


def test_si_sdr_definition():
    g = torch.Generator().manual_seed(0)
    ref = torch.randn(1, 2, 4000, generator=g)
    est = torch.stack([3.0 * ref[0, 1] + 0.1 * torch.randn(4000, generator=g), 0.5 * ref[0, 0]])[None]
    v, perm = S.si_sdr_pit_ref(ref, est)
    assert perm[0] == (1, 0) and v[0, 0] > 60 and 25 < v[0, 1] < 35  # scale invariant, permutation found
