"""Oobleck VAE around the latent sampler on the device against the REFERENCE's outputs
(tests/golden/oobleck.npz, written by oracle/pin_oobleck.py from the unmodified stable_audio_tools classes) and, layer by
layer, against the oracle restatement."""
import numpy as np
import pytest
import torch

from ditsep_b200 import ops
from ditsep_b200.models.oobleck import LatentCodec, OobleckVAE
from oracle import oobleck_ref as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def rel(a, b):
    return float((a.double().cpu() - b.double().cpu()).norm() / b.double().cpu().norm())


def _cfg(golden, name):
    c = [int(v) for v in golden[f"{name}_cfg"]]
    ch, lat, T = c[:3]
    n = (len(c) - 3) // 2
    return ch, lat, T, c[3:3 + n], c[3 + n:]


@pytest.fixture(scope="module")
def oob():
    import os
    return dict(np.load(os.path.join(os.path.dirname(__file__), "golden", "oobleck.npz")))


def test_conv1d_forms_vs_torch():
    """The three forms of ditsep_conv1d_tc (dilated, strided through the reshaped view, transposed in polyphase form with the
    shifted output view), bias / residual / ELU outputs, against torch's conv1d / conv_transpose1d on the same bf16 operands."""
    import torch.nn.functional as F
    g = torch.Generator().manual_seed(0)
    B, T, Cin, Cout = 2, 300, 128, 192
    x = torch.randn(B, Cin, T, generator=g).to(torch.bfloat16)
    xd = x.permute(0, 2, 1).contiguous().to(DEV)                                  # [B][T][Cin]
    for taps, dil in ((7, 1), (7, 3), (7, 9), (1, 1), (3, 1)):
        w = (torch.randn(Cout, Cin, taps, generator=g) / (Cin * taps) ** 0.5).to(torch.bfloat16)
        bias = 0.1 * torch.randn(Cout, generator=g)
        res = torch.randn(B, Cout, T, generator=g).to(torch.bfloat16)
        pad = dil * (taps - 1) // 2
        ref = F.conv1d(x.float(), w.float(), bias, padding=pad, dilation=dil) + res.float()
        out = torch.empty(B, T, Cout, device=DEV, dtype=torch.bfloat16)
        act = torch.empty_like(out)
        ops.conv1d_tc(xd, T, T * Cin, w.permute(0, 2, 1).reshape(Cout, -1).contiguous().to(DEV), Cout, taps, dil, pad, 1, T,
                      bias=bias.to(DEV), residual=res.permute(0, 2, 1).contiguous().to(DEV), ld_res=Cout, bs_res=T * Cout,
                      out=out, out_act=act, ld_out=Cout, bs_out=T * Cout)
        assert rel(out.float().permute(0, 2, 1), ref) < 4e-3, (taps, dil)
        assert rel(act.float(), F.elu(out.float())) < 4e-3
    for s in (2, 4, 8):
        Tm = 40 * s
        xs = torch.randn(B, Cin, Tm, generator=g).to(torch.bfloat16)
        w = (torch.randn(Cout, Cin, 2 * s, generator=g) / (Cin * 2 * s) ** 0.5).to(torch.bfloat16)
        ref = F.conv1d(xs.float(), w.float(), None, stride=s, padding=s // 2)
        out = torch.empty(B, Tm // s, Cout, device=DEV, dtype=torch.float32)
        ops.conv1d_tc(xs.permute(0, 2, 1).contiguous().to(DEV), Tm, Tm * Cin, w.permute(0, 2, 1).reshape(Cout, -1).contiguous().to(DEV),
                      Cout, 2 * s, 1, s // 2, s, Tm // s, out=out, ld_out=Cout, bs_out=(Tm // s) * Cout)
        assert ref.shape[-1] == Tm // s and rel(out.permute(0, 2, 1), ref) < 2e-5, s
        # transposed: (Cin, Co, 2s) -> polyphase filter [f * Co + co][m * Cin + ci]; output = buffer rows from s / 2 on
        Co, Tt = 64, 37
        xt = torch.randn(B, Cin, Tt, generator=g).to(torch.bfloat16)
        wt = (torch.randn(Cin, Co, 2 * s, generator=g) / (Cin * 2) ** 0.5).to(torch.bfloat16)
        bt = 0.1 * torch.randn(Co, generator=g)
        ref = F.conv_transpose1d(xt.float(), wt.float(), bt, stride=s, padding=s // 2)
        wp = wt.reshape(Cin, Co, 2, s).permute(3, 1, 2, 0).reshape(s * Co, 2 * Cin).contiguous()
        buf = torch.full((B, (Tt + 1) * s, Co), float("nan"), device=DEV, dtype=torch.float32)
        ops.conv1d_tc(xt.permute(0, 2, 1).contiguous().to(DEV), Tt, Tt * Cin, wp.to(DEV), s * Co, 2, -1, 0, 1, Tt + 1,
                      bias=bt.repeat(s).to(DEV), out=buf, ld_out=s * Co, bs_out=(Tt + 1) * s * Co)
        got = buf[:, s // 2: s // 2 + Tt * s].permute(0, 2, 1)
        assert ref.shape[-1] == Tt * s and rel(got, ref) < 2e-5, s


@pytest.mark.parametrize("name", ["small", "real"])
def test_oobleck_vae_vs_reference_golden(oob, name):
    ch, lat, T, cm, st = _cfg(oob, name)
    sd = O.make_state_dict(ch, cm, st, lat, seed=7)
    vae = OobleckVAE(ch, cm, st, lat, device=DEV).load_state_dict(sd)
    codec = LatentCodec(vae)
    mix = torch.tensor(oob[f"{name}_mix"]).to(DEV)
    noise = torch.tensor(oob[f"{name}_noise"]).to(DEV)
    n0 = ops.launch_count()
    pre = vae.encode(LatentCodec.pad(mix, vae.hop_length), skip_bottleneck=True)
    z, _ = codec.encode(mix, noise=noise)
    wav = codec.decode(torch.tensor(oob[f"{name}_est"]).to(DEV))
    torch.cuda.synchronize()
    e_pre, e_z, e_wav = rel(pre, torch.tensor(oob[f"{name}_pre"])), rel(z[:, 0], torch.tensor(oob[f"{name}_z"])), \
        rel(wav, torch.tensor(oob[f"{name}_wav"]))
    print(f"oobleck {name}: rel-L2 vs reference: encoder {e_pre:.2e}, latents {e_z:.2e}, decoder {e_wav:.2e} "
          f"({ops.launch_count() - n0} launches)")
    assert z.shape == (2, 1, lat // 2, oob[f"{name}_z"].shape[-1]) and wav.shape == oob[f"{name}_wav"].shape
    # bf16 activations through 30 + 36 convolutions with fp32 accumulation
    assert e_pre < 2e-2 and e_z < 2e-2 and e_wav < 3e-2
    # decode(target_dim) trims like diffsep_latent.py:126-127; the Philox bottleneck is reproducible per seed
    assert codec.decode(torch.tensor(oob[f"{name}_est"]).to(DEV), target_dim=T).shape[-1] == T
    z1, _ = codec.encode(mix, seed=5)
    z2, _ = codec.encode(mix, seed=5)
    assert torch.equal(z1, z2) and not torch.equal(z1, z)


def test_latent_separate_end_to_end_on_gpu(oob):
    """LatentDiffSepModel.separate (src/diffsep_latent.py:472-487) on the device: VAE encode -> latent "ald" sampler over the
    OUVE SDE -> VAE decode equals the same three calls made by hand; shapes and trimming as the reference's."""
    from ditsep_b200.diffsep_latent import LatentDiffSepModel, latent_default_config
    from oracle import ncsnpp_ref as R
    from oracle import sampler_ref as S
    ch, lat, T, cm, st = _cfg(oob, "small")
    vae = OobleckVAE(ch, cm, st, lat, device=DEV).load_state_dict(O.make_state_dict(ch, cm, st, lat, seed=7))
    model = LatentDiffSepModel(latent_default_config(nf=32, N=3, precision="bf16"), vae=vae)
    bcfg = R.NCSNppConfig(nf=32, ch_mult=(1, 2, 2), num_res_blocks=2, attn_resolutions=(16,), image_size=64,
                          num_channels_in=3, num_channels_out=2, centered=True)
    model.score_model.load_state_dict(R.make_state_dict(bcfg, seed=5, with_window=None), strict=True)
    model = model.to(DEV).eval()
    mix = torch.tensor(oob["small_mix"]).to(DEV)
    noise = torch.tensor(oob["small_noise"]).to(DEV)
    with torch.no_grad():
        est, nfe = model.separate(mix, target_dim=T, noise=noise, noise_tape=S.NoiseTape(3))
        z, _ = model.encode(mix, noise=noise)
        x, nfe2 = model.get_pc_sampler("reverse_diffusion", "ald", z, N=3, snr=0.5, corrector_steps=1,
                                       noise_tape=S.NoiseTape(3))()
        byhand = model.decode(x, T)
    assert nfe == nfe2 == 6 and est.shape == (2, 2, T) and torch.isfinite(est).all()
    assert torch.equal(est, byhand)
