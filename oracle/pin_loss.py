"""Pins the score-matching loss forward of SURVEY.md 8f-4 against the REAL reference (TEST INFRASTRUCTURE; build
container only).

``python -m oracle.pin_loss`` calls the unmodified reference ``DiffSepModel.sample_prior`` (src/diffsep.py:113-181, every
init_hack branch), ``compute_score_loss`` (:344-358), ``compute_score_loss_init_hack_pit`` (:305-338) and the ``DiffSepOU`` counterparts
(:800-811, 832-869; OUVESDE and SBVESDE, the latter through the reference's EDM ``forward``) UNBOUND on a
namespace that carries the attributes they read (the Lightning constructor needs hydra / omegaconf / torch_ema, absent
here), with a closed-form score network, a fixed time vector and taped normal draws; asserts the oracle against them and
writes the REFERENCE's outputs to tests/golden/loss.npz."""
import os
import sys
import types

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import sampler_ref as S  # noqa: E402
from oracle.pin_against_reference import GOLDEN, TapedRandn, import_reference  # noqa: E402
from oracle.pin_sb import import_diffsep_ou  # noqa: E402


def toy_score(x, t, mix):
    return -(x - 0.5 * mix) / (0.2 + t[:, None, None]) + 0.1 * torch.tanh(x.flip(1))


def main():
    sdes, _ = import_reference()
    import_diffsep_ou()
    DiffSepModel = sys.modules["diffsep"].DiffSepModel
    B, T = 4, 1024
    g = torch.Generator().manual_seed(31)
    target = torch.randn(B, 2, T, generator=g) * torch.tensor([1.0, 0.4])[None, :, None]
    mix = target.sum(dim=1, keepdim=True)
    time = torch.tensor([0.05, 0.5, 0.985, 1.0])
    sel4 = torch.tensor([False, True, False, True])
    golden = {"loss_target": target.numpy(), "loss_time": time.numpy(), "loss_sel4": sel4.numpy()}

    def namespace(hack):
        sde = sdes.sdes.MixSDE(2, 2.0, 0.05, 0.5, N=30)
        ns = types.SimpleNamespace(sde=sde, init_hack=hack, t_rev_init=0.03, sample_time=lambda x: time.clone(),
                                   loss=torch.nn.MSELoss(reduction="none"))
        return ns

    class Callable(types.SimpleNamespace):
        def __call__(self, *a):
            return toy_score(*a)

    orc = S.MixSDERef(2, 2.0, 0.05, 0.5, N=30)
    for hack in (0, 1, 2, 3, 4, 5):
        ns = Callable(**namespace(hack).__dict__)
        ns.sample_prior = types.MethodType(DiffSepModel.sample_prior, ns)
        real_rand_like = torch.rand_like
        torch.rand_like = lambda t, **k: torch.where(sel4, 0.0, 1.0).to(t.dtype)  # init_hack 4: rand < 1/N  <=>  sel4
        try:
            with TapedRandn(S.NoiseTape(70)):
                x_t, t_out, L, z = DiffSepModel.sample_prior(ns, mix, target)
            with TapedRandn(S.NoiseTape(70)):
                loss = DiffSepModel.compute_score_loss(ns, mix, target)
        finally:
            torch.rand_like = real_rand_like
        z0 = S.NoiseTape(70).draw(0, target.shape)
        kw = dict(init_hack=hack if hack in (1, 2, 3, 4) else 0, t_rev_init=0.03, select4=sel4)
        x_o, t_o, L_o, z_o = S.sample_prior_ref(orc, mix, target, time, z0, **kw)
        loss_o = S.score_loss_ref(toy_score, orc, mix, target, time, z0, **kw)
        errs = [float((a - b).abs().max()) for a, b in ((x_t, x_o), (t_out, t_o), (L, L_o), (z, z_o), (loss, loss_o))]
        print(f"init_hack {hack}: oracle vs reference max abs err x_t {errs[0]:.2e} time {errs[1]:.2e} L {errs[2]:.2e} "
              f"z {errs[3]:.2e} loss {errs[4]:.2e}; loss {loss.tolist()}")
        assert max(errs) < 1e-5
        golden.update({f"loss_h{hack}_xt": x_t.numpy(), f"loss_h{hack}_time": t_out.numpy(), f"loss_h{hack}_L": L.numpy(),
                       f"loss_h{hack}_z": z.numpy(), f"loss_h{hack}_loss": loss.numpy()})
    ns = Callable(**namespace(5).__dict__)
    with TapedRandn(S.NoiseTape(71)):
        pit = DiffSepModel.compute_score_loss_init_hack_pit(ns, mix, target)
    pit_o = S.score_loss_init_hack_pit_ref(toy_score, orc, mix, target, S.NoiseTape(71).draw(0, target.shape))
    err = float((pit - pit_o).abs().max())
    print(f"init-hack PIT loss: oracle vs reference max abs err {err:.2e}; {pit.tolist()}")
    assert err < 1e-5
    golden["loss_pit"] = pit.numpy()

    # ---- the two remaining PIT variants (src/diffsep.py:183-303); their random choices are pinned: the order selection of
    # utils.select_elem_at_random by a fixed index vector, utils.shuffle_sources by the identity
    import utils as ref_utils
    sel = torch.tensor([1, 0, 1, 0])
    golden["loss_pit_select"] = sel.numpy()

    def fixed_select(x, dim=-1, batch_dim=0):
        x = x.moveaxis(dim, -1)
        idx = torch.broadcast_to(sel[(...,) + (None,) * (x.ndim - 1)], x.shape[:-1] + (1,))
        return torch.gather(x, dim=-1, index=idx).moveaxis(-1, dim)

    ref_utils.select_elem_at_random, ref_utils.shuffle_sources = fixed_select, (lambda x: x)
    ns = Callable(**namespace(6).__dict__)
    ns.config = types.SimpleNamespace(model=types.SimpleNamespace(mmnr_thresh_pit=-9.5))  # items 0-2 plain, item 3 PIT (-9.93 dB)
    with TapedRandn(S.NoiseTape(75)):
        wp = DiffSepModel.compute_score_loss_with_pit(ns, mix, target)
    with TapedRandn(S.NoiseTape(76)):
        wa = DiffSepModel.compute_score_loss_with_pit_allthetime(ns, mix, target)
    wp_o = S.score_loss_with_pit_ref(toy_score, orc, mix, target, time, S.NoiseTape(75).draw(0, target.shape), sel, -9.5)
    wa_o = S.score_loss_with_pit_allthetime_ref(toy_score, orc, mix, target, time, S.NoiseTape(76).draw(0, target.shape))
    e1, e2 = float((wp - wp_o).abs().max()), float((wa - wa_o).abs().max())
    print(f"with_pit: oracle vs reference max abs err {e1:.2e}; {wp.tolist()}\nwith_pit_allthetime: {e2:.2e}; {wa.tolist()}")
    assert e1 < 1e-5 and e2 < 1e-5 and wp.shape == wa.shape == (B,)
    golden.update(loss_with_pit=wp.numpy(), loss_with_pit_allthetime=wa.numpy())

    # ---- DiffSepOU (src/diffsep.py:796-869): scalar-std SDEs; the SBVESDE goes through the reference's own EDM forward
    DiffSepOU = sys.modules["diffsep"].DiffSepOU
    from oracle.pin_sb import toy_net

    def ou_namespace(sde, edm):
        class NS(types.SimpleNamespace):
            def __call__(self, xt, t, y):
                return DiffSepOU.forward(self, xt, t, y)

        ns = NS(sde=sde, is_edm=edm, sample_time=lambda x: time.clone(), loss=torch.nn.MSELoss(reduction="none"),
                score_model=toy_net if edm else toy_score, network_scaling="1/sigma", sigma_data=0.1, c_in="edm",
                c_out="edm", c_skip="edm")
        for name in ("_c_in", "_c_out", "_c_skip", "sample_prior", "_step"):
            setattr(ns, name, types.MethodType(getattr(DiffSepOU, name), ns))
        return ns

    for tag, mk_ref, orc_sde, edm in (
            ("ouve", lambda: sdes.sdes.OUVESDE(theta=1.5, sigma_min=0.05, sigma_max=0.5, N=30), S.OUVESDERef(1.5, 0.05, 0.5, N=30), False),
            ("sbve", lambda: sdes.sdes.SBVESDE(2.6, 0.4, N=30), S.SBVESDERef(2.6, 0.4, N=30), True)):
        ns = ou_namespace(mk_ref(), edm)
        with TapedRandn(S.NoiseTape(73)):
            x_t, t_out, sigma, z = DiffSepOU.sample_prior(ns, mix, target)
        with TapedRandn(S.NoiseTape(73)):
            loss = DiffSepOU.compute_score_loss(ns, mix, target)
        z0 = S.NoiseTape(73).draw(0, target.shape)
        model_o = (lambda a, b, c: S.edm_forward_ref(toy_net, orc_sde, a, b, c)) if edm else toy_score
        x_o, sigma_o = S.ou_sample_prior_ref(orc_sde, mix, target, time, z0)
        loss_o = S.ou_score_loss_ref(model_o, orc_sde, mix, target, time, z0)
        errs = [float((a - b).abs().max()) for a, b in ((x_t, x_o), (sigma, sigma_o), (loss, loss_o))]
        print(f"DiffSepOU / {tag}: oracle vs reference max abs err x_t {errs[0]:.2e} sigma {errs[1]:.2e} loss {errs[2]:.2e}; "
              f"loss {loss.tolist()}")
        assert max(errs) < 1e-5 * max(1.0, float(loss.abs().max())) and tuple(loss.shape) == (B, 2)
        golden.update({f"ou_{tag}_xt": x_t.numpy(), f"ou_{tag}_sigma": sigma.numpy(), f"ou_{tag}_loss": loss.numpy()})
    ns = ou_namespace(sdes.sdes.OUVESDE(theta=1.5, sigma_min=0.05, sigma_max=0.5, N=30), False)
    with TapedRandn(S.NoiseTape(74)):
        pit = DiffSepOU.compute_score_loss_init_hack_pit(ns, mix, target)
    pit_o = S.ou_score_loss_init_hack_pit_ref(toy_score, S.OUVESDERef(1.5, 0.05, 0.5, N=30), mix, target,
                                              S.NoiseTape(74).draw(0, target.shape))
    err = float((pit - pit_o).abs().max())
    print(f"DiffSepOU / ouve init-hack PIT loss: oracle vs reference max abs err {err:.2e}; {pit.tolist()}")
    assert err < 1e-5 * max(1.0, float(pit.abs().max())) and tuple(pit.shape) == (B, 2)
    golden["ou_ouve_pit"] = pit.numpy()
    np.savez_compressed(os.path.join(GOLDEN, "loss.npz"), **golden)
    print("written", os.path.join(GOLDEN, "loss.npz"))


if __name__ == "__main__":
    main()
