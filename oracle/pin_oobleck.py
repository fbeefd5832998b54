"""Pins oracle/oobleck_ref.py against the REAL reference VAE classes (TEST INFRASTRUCTURE; build container only).

``python -m oracle.pin_oobleck`` imports the unmodified stable_audio_tools.models.autoencoders from /root/reference/src
(its package __init__ files and the heavy siblings it imports but the VAE never calls -- diffusion, transformer,
pretransforms, inference -- are replaced by empty stand-ins; `alias_free_torch` is absent here and only wraps
activations when antialias=True, which oobleck_finetune.json does not set), builds OobleckEncoder / OobleckDecoder /
VAEBottleneck with seeded weights, asserts the oracle against them and writes the REFERENCE's outputs to
tests/golden/oobleck.npz."""
import importlib
import os
import sys
import types

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oobleck_ref as O  # noqa: E402

REF_SRC = "/root/reference/src"
GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def import_reference_vae():
    class _Any:
        def __init__(self, *a, **k):
            pass

    def stub(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)

        def _placeholder(attr):  # any other public name the module imports: a placeholder class
            if attr.startswith("__"):
                raise AttributeError(attr)
            return _Any

        m.__getattr__ = _placeholder
        sys.modules[name] = m
        return m

    for pkg, sub in (("stable_audio_tools", ""), ("stable_audio_tools.models", "models"),
                     ("stable_audio_tools.inference", "inference")):
        m = types.ModuleType(pkg)
        m.__path__ = [os.path.join(REF_SRC, "stable_audio_tools", sub)]
        sys.modules[pkg] = m
    stub("alias_free_torch", Activation1d=_Any)
    for name in ("stable_audio_tools.inference.sampling", "stable_audio_tools.inference.utils",
                 "stable_audio_tools.models.diffusion", "stable_audio_tools.models.factory",
                 "stable_audio_tools.models.pretransforms", "stable_audio_tools.models.transformer"):
        stub(name)
    A = importlib.import_module("stable_audio_tools.models.autoencoders")
    Bn = importlib.import_module("stable_audio_tools.models.bottleneck")
    return A, Bn


class TapedRandnLike:
    def __init__(self, tensor):
        self.t = tensor

    def __enter__(self):
        self._orig = torch.randn_like
        torch.randn_like = lambda x, **kw: self.t.to(x.dtype)
        return self

    def __exit__(self, *a):
        torch.randn_like = self._orig


def main():
    A, Bn = import_reference_vae()
    golden = {}
    # (name, channels, c_mults, strides, encoder latent (mean | scale), samples): a small net with all five stages, and
    # the real oobleck_finetune.json widths (128 .. 2048 channels) on a short signal
    for name, ch, cm, st, lat, T in (("small", 64, [1, 2, 2], [2, 4, 8], 128, 3000),
                                     ("real", 128, [1, 2, 4, 8, 16], [2, 4, 4, 8, 8], 128, 4096)):
        torch.manual_seed(0)
        enc = A.OobleckEncoder(in_channels=1, channels=ch, c_mults=cm, strides=st, latent_dim=lat).eval()
        dec = A.OobleckDecoder(out_channels=1, channels=ch, c_mults=cm, strides=st, latent_dim=lat // 2).eval()
        sd = O.make_state_dict(ch, cm, st, lat, seed=7)
        # strict=True: the reference classes accept exactly these keys and shapes
        enc.load_state_dict({k[len("encoder."):]: v for k, v in sd.items() if k.startswith("encoder.")}, strict=True)
        dec.load_state_dict({k[len("decoder."):]: v for k, v in sd.items() if k.startswith("decoder.")}, strict=True)
        g = torch.Generator().manual_seed(3)
        hop = int(np.prod(st))
        mix = 0.3 * torch.randn(2, 1, T, generator=g)
        xin = O.pad_to_hop(mix, hop)
        with torch.no_grad():
            pre = enc(xin)
            noise = torch.randn(pre.shape[0], lat // 2, pre.shape[2], generator=g)
            with TapedRandnLike(noise):
                z = Bn.VAEBottleneck().encode(pre)
            est = torch.stack([z, 0.5 * z.flip(0)], dim=1)            # (B, 2 sources, latent, T')
            wav = dec(est.reshape(-1, *est.shape[2:])).reshape(2, 2, -1)
            o_pre = O.encoder_forward(sd, xin, st)
            o_z = O.latent_encode(sd, mix, st, noise)[:, 0]
            o_wav = O.latent_decode(sd, est, st)
        for what, a, b in (("encoder", o_pre, pre), ("bottleneck", o_z, z), ("decoder", o_wav, wav)):
            err = float((a - b).norm() / b.norm())
            print(f"{name} {what}: oracle vs reference rel-L2 {err:.2e}, |ref| rms {float(b.pow(2).mean().sqrt()):.3f}, shape {tuple(b.shape)}")
            assert err < 1e-5
        golden.update({f"{name}_mix": mix.numpy(), f"{name}_noise": noise.numpy(), f"{name}_pre": pre.numpy(),
                       f"{name}_z": z.numpy(), f"{name}_est": est.numpy(), f"{name}_wav": wav.numpy(),
                       f"{name}_cfg": np.array([ch, lat, T] + list(cm) + list(st))})
    np.savez_compressed(os.path.join(GOLDEN, "oobleck.npz"), **golden)
    print("written", os.path.join(GOLDEN, "oobleck.npz"), os.path.getsize(os.path.join(GOLDEN, "oobleck.npz")) // 1024, "KB")


if __name__ == "__main__":
    main()
