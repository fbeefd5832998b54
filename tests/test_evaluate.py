"""Evaluation sweep driver (ditsep_b200/evaluate.py): metrics against the oracle restatement and
known answers, the reference's row schema / summary / JSON dump, the sampler overrides.  CPU: kernels replaced by
the torch stand-ins of tests/emu_ops.py; the GPU run of the same driver is tests/test_gpu_parity.py."""
import json
import math

import numpy as np
import pytest
import torch

from common import build_model
from ditsep_b200 import evaluate as E
from emu_ops import emulated_ops
from oracle import metrics_ref as M
from oracle import sampler_ref as S


@pytest.mark.parametrize("n_src", [2, 3])
def test_si_bss_eval_matches_oracle(n_src):
    g = torch.Generator().manual_seed(11 + n_src)
    ref = torch.randn(3, n_src, 6000, generator=g)
    ref[:, 1] += 0.3 * ref[:, 0]                                   # correlated references: SIR != SDR
    mixm = torch.eye(n_src) + 0.1 * torch.randn(n_src, n_src, generator=g)
    est = torch.einsum("ij,bjt->bit", mixm, ref) + 0.05 * torch.randn(3, n_src, 6000, generator=g)
    order = list(range(1, n_src)) + [0]
    est = est[:, order]                                            # estimates come out rotated
    sdr, sir, sar, perm = E.si_bss_eval(ref, est)
    for b in range(3):
        o_sdr, o_sir, o_sar, o_perm = M.si_bss_eval_ref(ref[b].numpy(), est[b].numpy())
        assert perm[b].tolist() == o_perm.tolist()
        assert [order[p] for p in perm[b].tolist()] == list(range(n_src))
        for got, want in ((sdr, o_sdr), (sir, o_sir), (sar, o_sar)):
            np.testing.assert_allclose(got[b].numpy(), want, atol=2e-4, rtol=0)   # dB


def test_si_bss_eval_known_answers():
    g = torch.Generator().manual_seed(5)
    ref = torch.randn(1, 2, 8000, generator=g)
    ref[0, 1] -= (ref[0, 1] @ ref[0, 0]) / (ref[0, 0] @ ref[0, 0]) * ref[0, 0]    # orthogonal references
    noise = torch.randn(8000, generator=g)
    for r in ref[0]:
        noise -= (noise @ r) / (r @ r) * r
    noise *= ref[0, 0].norm() / noise.norm()
    est = torch.stack([3.0 * ref[0, 0] + 0.3 * ref[0, 1] * ref[0, 0].norm() / ref[0, 1].norm() + 0.03 * noise,
                       0.5 * ref[0, 1]])[None]
    sdr, sir, sar, perm = E.si_bss_eval(ref, est)
    assert perm.tolist() == [[0, 1]]
    assert abs(float(sir[0, 0]) - 20.0) < 1e-3                    # interference at 1/10 of the target amplitude
    assert abs(float(sar[0, 0]) - 10 * math.log10((9 + 0.09) / 0.0009)) < 1e-3
    assert abs(float(sdr[0, 0]) - 10 * math.log10(9 / (0.09 + 0.0009))) < 1e-3
    assert float(sdr[0, 1]) == 100.0 and float(sir[0, 1]) == 100.0   # exact scaled copy: clamp_db
    # same SI-SDR as the sweep's older single-metric helper
    from ditsep_b200.utils import shard
    old, _ = shard.si_sdr_pit(ref, est)
    assert abs(float(old[0, 0]) - float(sdr[0, 0])) < 1e-3


def test_summarize_reference_semantics():
    res = {0: {"batch_idx": 0, "si_sdr": [[10.0, 20.0]], "nfe": 60, "runtime": 1.0},
           1: {"batch_idx": 1, "si_sdr": [[float("inf"), 0.0]], "nfe": 60, "runtime": 3.0}}
    s = E.summarize(res)
    assert s["number"] == 2 and s["runtime"] == 2.0 and s["nfe"] == 60 and s["batch_idx"] == 0.5
    assert math.isinf(s["si_sdr"])                                # ignore_inf=True admits the infinite row (sic)
    assert E.summarize(res, ignore_inf=False)["si_sdr"] == 15.0


def test_evaluate_sweep_rows_summary_json(tmp_path):
    lengths = [2048, 1500, 2048, 1800]
    g = torch.Generator().manual_seed(99)
    srcs = [torch.randn(2, n, generator=g) for n in lengths]
    mixes = [s.sum(0, keepdim=True) for s in srcs]
    model, _, _ = build_model("fp32", N=2)
    tape = lambda i: S.NoiseTape(4321 + i).draw
    with emulated_ops(), torch.no_grad():
        res, summ, outs = E.evaluate_sweep(model, mixes, srcs, fs=8000, N=2, max_batch=2, output_dir=str(tmp_path),
                                           split="tt", noise_tape_factory=tape)
        res_b1, _, _ = E.evaluate_sweep(model, mixes, srcs, fs=8000, N=2, max_batch=1, noise_tape_factory=tape)
        res_s, _, _ = E.evaluate_sweep(model, mixes, srcs, fs=8000, N=3, snr=0.3, corrector_steps=2, denoise=False,
                                       schedule="log", max_batch=4, noise_tape_factory=tape)
    assert sorted(res) == [0, 1, 2, 3] and sorted(outs) == [0, 1, 2, 3]
    for i, row in res.items():
        assert tuple(row) == E.ROW_KEYS
        assert row["batch_idx"] == i and row["nfe"] == 4 and row["runtime"] > 0
        assert abs(row["len_s"] - lengths[i] / 8000) < 1e-6
        assert np.asarray(row["si_sdr"]).shape == (1, 2) and len(row["pesq"]) == 2
        # the row is what compute_metrics gives for the returned (already permuted) signals
        sdr, sir, sar, _, _, perm = E.compute_metrics(srcs[i][None], outs[i][None], 8000)
        assert perm.tolist() == [[0, 1]]
        np.testing.assert_allclose(row["si_sdr"], sdr.numpy(), atol=1e-4)
        np.testing.assert_allclose(row["si_sir"], sir.numpy(), atol=1e-4)
        np.testing.assert_allclose(row["si_sar"], sar.numpy(), atol=1e-4)
        np.testing.assert_allclose(row["si_sdr"], res_b1[i]["si_sdr"], atol=2e-3)   # batching does not change a row
    assert res_s[0]["nfe"] == 3 * (1 + 2) and res_s[0]["si_sdr"] != res[0]["si_sdr"]  # overrides reach the sampler
    assert summ["number"] == 4 and summ["nfe"] == 4
    assert abs(summ["si_sdr"] - np.mean([np.mean(r["si_sdr"]) for r in res.values()])) < 1e-5
    on_disk = json.load(open(tmp_path / "tt.json"))
    assert sorted(on_disk) == ["0", "1", "2", "3"] and on_disk["2"]["si_sir"] == res[2]["si_sir"]
    assert json.load(open(tmp_path / "tt_summary.json"))["number"] == 4
