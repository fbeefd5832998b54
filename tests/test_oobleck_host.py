"""Oobleck VAE on the CPU: the oracle restatement against the reference's outputs (golden fixture), and the
product's host logic -- weight-norm folding, filter re-packing, strided views, polyphase transposed convolutions with their
shifted output views, in-place residual streams -- through the torch stand-ins of tests/emu_ops.py against the same fixture."""
import os

import numpy as np
import pytest
import torch

from ditsep_b200.diffsep_latent import LatentDiffSepModel, latent_default_config, vae_from_config
from ditsep_b200.models.oobleck import LatentCodec, OobleckVAE
from emu_ops import emulated_ops
from oracle import oobleck_ref as O


@pytest.fixture(scope="module")
def oob():
    return dict(np.load(os.path.join(os.path.dirname(__file__), "golden", "oobleck.npz")))


def _small(oob):
    c = [int(v) for v in oob["small_cfg"]]
    n = (len(c) - 3) // 2
    return c[0], c[1], c[2], c[3:3 + n], c[3 + n:]


def rel(a, b):
    return float((torch.as_tensor(a).double() - torch.as_tensor(b).double()).norm() / torch.as_tensor(b).double().norm())


def test_oracle_vs_reference_golden(oob):
    ch, lat, T, cm, st = _small(oob)
    sd = O.make_state_dict(ch, cm, st, lat, seed=7)
    mix, noise, est = (torch.tensor(oob[f"small_{k}"]) for k in ("mix", "noise", "est"))
    with torch.no_grad():
        assert rel(O.encoder_forward(sd, O.pad_to_hop(mix, int(np.prod(st))), st), oob["small_pre"]) < 1e-5
        assert rel(O.latent_encode(sd, mix, st, noise)[:, 0], oob["small_z"]) < 1e-5
        assert rel(O.latent_decode(sd, est, st), oob["small_wav"]) < 1e-5
    # pad() always pads, a whole hop when the length is already a multiple (src/utils/torch_utils.py:12-18)
    assert O.pad_to_hop(torch.zeros(1, 1, 128), 64).shape[-1] == 192


def test_vae_host_logic_vs_reference_golden(oob):
    ch, lat, T, cm, st = _small(oob)
    sd = O.make_state_dict(ch, cm, st, lat, seed=7)
    vae = OobleckVAE(ch, cm, st, lat, device="cpu").load_state_dict(sd)
    codec = LatentCodec(vae)
    with emulated_ops():
        z, zt = codec.encode(torch.tensor(oob["small_mix"]), noise=torch.tensor(oob["small_noise"]))
        wav = codec.decode(torch.tensor(oob["small_est"]), target_dim=T)
    assert zt is None and z.shape[1] == 1
    assert rel(z[:, 0], oob["small_z"]) < 2e-2                         # bf16 activation storage, as on the device
    assert wav.shape[-1] == T and rel(wav, oob["small_wav"][..., :T]) < 3e-2
    with pytest.raises(KeyError):
        OobleckVAE(ch, cm, st, lat, device="cpu").load_state_dict({**sd, "encoder.layers.99.weight_g": torch.zeros(1)})
    with pytest.raises(ValueError):
        OobleckVAE(ch, cm, [2, 3, 8], lat, device="cpu")


def test_latent_model_wiring(oob):
    """LatentDiffSepModel (src/diffsep_latent.py): config -> score model + OUVE SDE + VAE; separate = encode -> "ald" sampler
    in the latent space -> decode."""
    cfg = latent_default_config(nf=16, N=2)
    vae = vae_from_config(cfg["model"]["vae"], device="cpu")
    assert (vae.channels, vae.strides, vae.latent, vae.hop_length) == (128, [2, 4, 4, 8, 8], 64, 2048)
    ch, lat, T, cm, st = _small(oob)
    small = OobleckVAE(ch, cm, st, lat, device="cpu").load_state_dict(O.make_state_dict(ch, cm, st, lat, seed=7))
    model = LatentDiffSepModel(cfg, vae=small).eval()
    from oracle import sampler_ref as S
    mix = torch.tensor(oob["small_mix"])
    with emulated_ops(), torch.no_grad():
        est, nfe = model.separate(mix, target_dim=T, noise=torch.tensor(oob["small_noise"]), noise_tape=S.NoiseTape(3))
    assert est.shape == (2, 2, T) and nfe == 4 and torch.isfinite(est).all()
