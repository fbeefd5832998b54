"""Pins the Schroedinger-bridge samplers of SURVEY.md 8f-2 against the REAL reference (TEST INFRASTRUCTURE; build
container only).

``python -m oracle.pin_sb`` runs the unmodified reference ``get_sb_sampler`` (src/sdes/__init__.py:284-389, "ode" and
"sde") with ``SBVESDE`` (src/sdes/sdes.py:701-779), once with a closed-form estimator and once through the reference's
own ``DiffSepOU.forward`` (src/diffsep.py:813-828: the EDM preconditioning around a closed-form "network"), asserts the
oracle against it and writes the REFERENCE's outputs to tests/golden/sb.npz.  ``DiffSepOU`` is a Lightning module whose
constructor needs hydra / omegaconf / torch_ema (absent here): its ``forward`` / ``_c_in`` / ``_c_out`` / ``_c_skip``
are called UNBOUND on a namespace carrying the attributes they read, so the code that runs is the reference's."""
import os
import sys
import types

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import sampler_ref as S  # noqa: E402
from oracle.pin_against_reference import GOLDEN, TapedRandn, import_reference  # noqa: E402


def toy_net(x, t, mix):
    """Stands in for the score network: smooth, non-linear, couples the sources and the mixture."""
    return 0.6 * mix - 0.35 * x + 0.1 * torch.tanh(2.0 * x.flip(1)) * (0.5 + t[:, None, None])


def toy_estimator(x, t, y):
    return 0.5 * y + 0.3 * torch.tanh(x) * (1.0 - 0.5 * t[:, None, None]) + 0.05 * x.flip(1)


def import_diffsep_ou():
    def stub(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules.setdefault(name, m)
        return sys.modules[name]

    stub("pytorch_lightning", LightningModule=torch.nn.Module)
    oc = stub("omegaconf", DictConfig=dict)
    oc.omegaconf = stub("omegaconf.omegaconf", open_dict=lambda c: c)
    stub("torch_ema", ExponentialMovingAverage=object)
    sys.modules["hydra.utils"].to_absolute_path = lambda p: p
    import utils  # the stub package installed by import_reference
    for name in ("normalize_batch", "denormalize_batch"):
        if not hasattr(utils, name):
            import importlib
            setattr(utils, name, getattr(importlib.import_module("utils.separate"), name))
    import importlib
    return importlib.import_module("diffsep").DiffSepOU


def reference_edm_model(DiffSepOU, sde, net, sigma_data=0.1):
    ns = types.SimpleNamespace(is_edm=True, sde=sde, score_model=net, network_scaling="1/sigma", sigma_data=sigma_data,
                               c_in="edm", c_out="edm", c_skip="edm")
    for name in ("_c_in", "_c_out", "_c_skip"):
        setattr(ns, name, types.MethodType(getattr(DiffSepOU, name), ns))
    return lambda xt, time, mix: DiffSepOU.forward(ns, xt, time, mix)


def main():
    sdes, _ = import_reference()
    DiffSepOU = import_diffsep_ou()
    B, T, N = 2, 4096, 6
    y = torch.randn(B, 1, T, generator=torch.Generator().manual_seed(11)) * 0.7
    golden = {"sb_y": y.numpy()}
    ref_sde = lambda: sdes.sdes.SBVESDE(2.6, 0.4, N=N)  # noqa: E731
    orc_sde = lambda: S.SBVESDERef(2.6, 0.4, N=N)  # noqa: E731

    # SBVESDE's closed forms on a time grid (sdes.py:741-770)
    tt = torch.tensor([1.0, 0.83, 0.5, 0.03, 1e-4])
    x0 = torch.randn(5, 2, 64, generator=torch.Generator().manual_seed(12))
    yy = torch.randn(5, 1, 64, generator=torch.Generator().manual_seed(13))
    mean_ref, std_ref = ref_sde().marginal_prob(x0, yy, tt)
    mean_orc, std_orc = orc_sde().marginal_prob(x0, yy, tt)
    assert torch.equal(mean_ref, mean_orc) and torch.equal(std_ref, std_orc)
    golden.update(sb_t=tt.numpy(), sb_x0=x0.numpy(), sb_yy=yy.numpy(), sb_mean=mean_ref.numpy(), sb_std=std_ref.numpy())

    # one preconditioned forward (diffsep.py:813-828)
    xt = torch.randn(B, 2, T, generator=torch.Generator().manual_seed(14))
    tv = torch.tensor([0.7, 0.2])
    f_ref = reference_edm_model(DiffSepOU, ref_sde(), toy_net)(xt, tv, y)
    f_orc = S.edm_forward_ref(toy_net, orc_sde(), xt, tv, y)
    err = float((f_ref - f_orc).abs().max())
    print(f"DiffSepOU.forward (EDM): max abs err oracle vs reference {err:.3e}")
    assert err < 1e-6
    golden.update(edm_xt=xt.numpy(), edm_t=tv.numpy(), edm_out=f_ref.numpy())

    for kind in ("ode", "sde"):
        for name, ref_model, orc_model in (
                ("plain", toy_estimator, toy_estimator),
                ("edm", reference_edm_model(DiffSepOU, ref_sde(), toy_net),
                 lambda a, b, c: S.edm_forward_ref(toy_net, orc_sde(), a, b, c))):
            with TapedRandn(S.NoiseTape(90)):
                x_ref, n_ref = sdes.get_sb_sampler(ref_sde(), ref_model, y=y, sampler_type=kind)()
            x_orc, n_orc = S.sb_sampler_ref(orc_sde(), orc_model, y, S.NoiseTape(90), sampler_type=kind)
            err = float((x_ref - x_orc).abs().max())
            print(f"sb {kind} / {name}: returned count {n_ref}, max abs err oracle vs reference {err:.3e}, "
                  f"|x| max {float(x_ref.abs().max()):.3f}, distance from the start {float((x_ref - y).abs().max()):.3f}")
            assert n_ref == n_orc == 50 and err < 1e-5
            golden[f"sb_{kind}_{name}"] = x_ref.numpy()
    np.savez_compressed(os.path.join(GOLDEN, "sb.npz"), **golden)
    print("written", os.path.join(GOLDEN, "sb.npz"))


if __name__ == "__main__":
    main()
