"""TEST INFRASTRUCTURE ONLY (tests/ may import this; the product never does).

CPU restatement (numpy, explicit projections) of the separation metrics the reference's evaluation driver reports:
src/evaluate_mp.py:171-189 calls `fast_bss_eval.si_bss_eval_sources(ref, est, zero_mean=False,
compute_permutation=True, clamp_db=100)`.  fast_bss_eval (requirements.txt:21, unpinned there) is a third-party
package that is neither vendored in /root/reference nor installed in this image, so this file restates the published
definition -- BSS-eval v3 with a length-1 distortion filter (Le Roux et al. 2019 "SDR - half-baked or well done?";
Scheibler 2022 "SDR - medium rare with fast computations"):
    s_target = <est, ref_i> / <ref_i, ref_i> ref_i
    e_interf = P_span(all refs) est - s_target
    e_artif  = est - P_span(all refs) est
    SI-SDR = |s_target|^2 / |est - s_target|^2, SI-SIR = |s_target|^2 / |e_interf|^2,
    SI-SAR = |s_target + e_interf|^2 / |e_artif|^2
and the permutation that maximises the summed SI-SIR ("order according to SIR", evaluate_mp.py:177).
PARITY UNPINNED for these three numbers: no golden vector of fast_bss_eval exists in the reference; the anchor is the
published definition plus the known-answer cases of tests/test_evaluate.py.
"""
import itertools

import numpy as np


def si_bss_eval_ref(ref: np.ndarray, est: np.ndarray, clamp_db: float = 100.0):
    """ref, est: (n_src, T) float64 -> (si_sdr, si_sir, si_sar, perm), each (n_src,) in reference order."""
    ref, est = np.asarray(ref, np.float64), np.asarray(est, np.float64)
    n = ref.shape[0]
    sdr, sir, sar = (np.zeros((n, n)) for _ in range(3))
    with np.errstate(divide="ignore"):
        for j in range(n):
            coef, *_ = np.linalg.lstsq(ref.T, est[j], rcond=None)
            p_all = ref.T @ coef
            e_artif = est[j] - p_all
            for i in range(n):
                s_t = (est[j] @ ref[i]) / (ref[i] @ ref[i]) * ref[i]
                e_i = p_all - s_t
                db = lambda a, b: float(np.clip(10 * (np.log10(a @ a) - np.log10(b @ b)), -clamp_db, clamp_db))
                sdr[i, j], sir[i, j], sar[i, j] = db(s_t, est[j] - s_t), db(s_t, e_i), db(p_all, e_artif)
    best = max(itertools.permutations(range(n)), key=lambda p: sum(sir[i, p[i]] for i in range(n)))
    idx = (np.arange(n), np.array(best))
    return sdr[idx], sir[idx], sar[idx], np.array(best)
