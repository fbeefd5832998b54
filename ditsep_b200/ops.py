"""Tensor-level wrappers over the C ABI (torch is plumbing: device memory + streams).

Every function takes CUDA tensors, launches on torch's current stream and raises on any
error; CPU tensors are rejected (there is no CPU path).
Activations are NHWC views: a tensor of shape (B, H, W, C) whose last-dim stride is 1 and
whose pixel stride ``ld`` may exceed C (a channel slice of a wider concat buffer).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import torch

import os

from . import _lib
from ._lib import DitsepError, SdeParams, check  # noqa: F401  (re-exported)

F32, BF16 = 0, 1
NUM_SMS = 148


def tc_enabled() -> bool:
    """tcgen05 convolution path on (default) unless DITSEP_DISABLE_TC=1 (debug aid)."""
    return os.environ.get("DITSEP_DISABLE_TC", "0") != "1"


def graphs_enabled() -> bool:
    """CUDA-graph replay of a launch plan (default) unless DITSEP_GRAPH=0 (debugging, per-launch timing)."""
    return os.environ.get("DITSEP_GRAPH", "1") != "0"


def require_device(t: torch.Tensor, what: str):
    """The product path runs on sm_100a only: a CPU tensor is an error, never a fallback."""
    if not t.is_cuda:
        raise _lib.DitsepError(f"{what} runs on sm_100a only; there is no CPU fallback")

_DT = {torch.float32: F32, torch.bfloat16: BF16}


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise _lib.DitsepError("ditsep_b200 kernels need CUDA tensors; there is no CPU fallback")


def _dt(t: torch.Tensor) -> int:
    try:
        return _DT[t.dtype]
    except KeyError:
        raise _lib.DitsepError(f"unsupported activation dtype {t.dtype}") from None


def nhwc_ld(t: torch.Tensor) -> int:
    """Pixel stride of an NHWC view (B,H,W,C); checks that pixels are laid out densely in B,H,W."""
    b, h, w, c = t.shape
    sb, sh, sw, sc = t.stride()
    if c > 1 and sc != 1:
        raise _lib.DitsepError("NHWC view must have unit channel stride")
    ld = sw
    if (w > 1 and False) or (h > 1 and sh != w * ld) or (b > 1 and sb != h * w * ld):
        raise _lib.DitsepError(f"NHWC view is not pixel-dense: shape {tuple(t.shape)} strides {t.stride()}")
    return ld


def launch_count() -> int:
    return int(_lib.load().ditsep_launch_count())


def note_launches(n: int):
    """Account for kernel launches the library did not see one by one (a captured CUDA graph being replayed)."""
    _lib.load().ditsep_note_launches(n)


# ------------------------------------------------------------------ SDE updates
def _f32c(t: Optional[torch.Tensor], name: str):
    if t is None:
        return
    if t.dtype != torch.float32 or not t.is_contiguous():
        raise _lib.DitsepError(f"{name} must be contiguous float32")


def _state_dims(x: torch.Tensor, p: SdeParams) -> Tuple[int, int]:
    B = x.shape[0]
    if x.shape[1] != p.n_src:
        raise _lib.DitsepError(f"state has {x.shape[1]} sources, SDE params say {p.n_src}")
    L = x[0, 0].numel()
    return B, L


def _lengths(lengths, B: int):
    if lengths is None:
        return None
    if lengths.dtype != torch.int32 or not lengths.is_contiguous() or lengths.numel() != B:
        raise _lib.DitsepError("lengths must be a contiguous int32 tensor of B elements")
    _cuda(lengths)
    return lengths.data_ptr()


def pc_corrector(p: SdeParams, x, score, t, noise=None, seed=0, draw=0, x_mean=None, lengths=None):
    """lengths: optional (B,) int32 valid sample counts of a ragged batch (state kept at 0 past them)."""
    _cuda(x, score, t, noise, x_mean)
    for n, v in (("x", x), ("score", score), ("t", t), ("noise", noise), ("x_mean", x_mean)):
        _f32c(v, n)
    B, L = _state_dims(x, p)
    check(_lib.load().ditsep_pc_corrector(C.byref(p), _ptr(x), _ptr(score), _ptr(noise), seed, draw, _ptr(t),
                                          _ptr(x_mean), B, L, _lengths(lengths, B), _stream()), "pc_corrector")


def pc_predictor(p: SdeParams, x, score, t, y, noise=None, seed=0, draw=0, x_mean=None, lengths=None):
    _cuda(x, score, t, y, noise, x_mean)
    for n, v in (("x", x), ("score", score), ("t", t), ("y", y), ("noise", noise), ("x_mean", x_mean)):
        _f32c(v, n)
    B, L = _state_dims(x, p)
    check(_lib.load().ditsep_pc_predictor(C.byref(p), _ptr(x), _ptr(score), _ptr(noise), seed, draw, _ptr(t), _ptr(y),
                                          _ptr(x_mean), B, L, _lengths(lengths, B), _stream()), "pc_predictor")


def pc_prior(p: SdeParams, x, y, noise=None, seed=0, draw=0, lengths=None):
    _cuda(x, y, noise)
    for n, v in (("x", x), ("y", y), ("noise", noise)):
        _f32c(v, n)
    B, L = _state_dims(x, p)
    check(_lib.load().ditsep_pc_prior(C.byref(p), _ptr(x), _ptr(noise), seed, draw, _ptr(y), B, L,
                                      _lengths(lengths, B), _stream()), "pc_prior")


def pc_langevin(n_src: int, x, score, noise, step, x_mean=None, lengths=None):
    """One LangevinCorrector step with the batch-wide step size read from the 1-element device tensor `step`."""
    _cuda(x, score, noise, step, x_mean)
    for n, v in (("x", x), ("score", score), ("noise", noise), ("step", step), ("x_mean", x_mean)):
        _f32c(v, n)
    B = x.shape[0]
    L = x.numel() // (B * n_src)
    check(_lib.load().ditsep_pc_langevin(n_src, _ptr(x), _ptr(score), _ptr(noise), _ptr(step), _ptr(x_mean), B, L,
                                         _lengths(lengths, B), _stream()), "pc_langevin")


def energy_envelope(mix: torch.Tensor, out: torch.Tensor, avg_len: int):
    """PriorMixSDE's per-sample diffusion envelope: out (B,1,L) = 0.5 sqrt(max(box_avg_len(mix^2), 1e-4))."""
    _cuda(mix, out)
    _f32c(mix, "mix")
    _f32c(out, "out")
    B = mix.shape[0]
    check(_lib.load().ditsep_energy_envelope(_ptr(mix), _ptr(out), B, mix.numel() // B, int(avg_len), _stream()),
          "energy_envelope")


def pc_perturb(p: SdeParams, x, t, y=None, noise=None, seed=0, draw=0, x_mean=None):
    """Marginal sample in place: x holds the clean sources on entry, mean(x0, t) + L(t) z on return (ditsep_pc_perturb)."""
    _cuda(x, t, y, noise, x_mean)
    for n, v in (("x", x), ("t", t), ("y", y), ("noise", noise), ("x_mean", x_mean)):
        _f32c(v, n)
    B, L = _state_dims(x, p)
    check(_lib.load().ditsep_pc_perturb(C.byref(p), _ptr(x), _ptr(noise), seed, draw, _ptr(t), _ptr(y), _ptr(x_mean), B, L,
                                        None, _stream()), "pc_perturb")


def score_loss(p: SdeParams, pred, t, noise=None, seed=0, draw=0) -> torch.Tensor:
    """(B,) f32: mean over (source, sample) of (L(t) pred + z)^2 (ditsep_score_loss); z = `noise` or the Philox draw."""
    _cuda(pred, t, noise)
    for n, v in (("pred", pred), ("t", t), ("noise", noise)):
        _f32c(v, n)
    B, L = _state_dims(pred, p)
    if noise is not None and noise.numel() != pred.numel():
        raise _lib.DitsepError("score_loss: noise must have the prediction's shape")
    loss = torch.empty(B, device=pred.device, dtype=torch.float32)
    work = torch.empty(256 * B, device=pred.device, dtype=torch.float64)
    check(_lib.load().ditsep_score_loss(C.byref(p), _ptr(pred), _ptr(noise), seed, draw, _ptr(t), _ptr(loss), _ptr(work), B, L,
                                        _stream()), "score_loss")
    return loss


def score_loss_sigma(pred, sigma, noise=None, seed=0, draw=0) -> torch.Tensor:
    """(B,) f32: mean over everything but the batch axis of (sigma[b] pred + z)^2 (ditsep_score_loss_sigma)."""
    _cuda(pred, sigma, noise)
    for n, v in (("pred", pred), ("sigma", sigma), ("noise", noise)):
        _f32c(v, n)
    B, n_src = pred.shape[0], pred.shape[1]
    L = pred.numel() // (B * n_src)
    if sigma.numel() != B or (noise is not None and noise.numel() != pred.numel()):
        raise _lib.DitsepError("score_loss_sigma: sigma must be (B,), noise must have the prediction's shape")
    if not 1 <= n_src <= 4:
        raise _lib.DitsepError(f"score_loss_sigma: {n_src} sources (1..4)")
    loss = torch.empty(B, device=pred.device, dtype=torch.float32)
    work = torch.empty(256 * B, device=pred.device, dtype=torch.float64)
    check(_lib.load().ditsep_score_loss_sigma(_ptr(pred), _ptr(noise), seed, draw, _ptr(sigma), _ptr(loss), _ptr(work), B, n_src, L,
                                              _stream()), "score_loss_sigma")
    return loss


def sb_step(x, f, y, coef, *, x_out=None, x_in=None, y_in=None, est=None, noise=None, use_noise=False, seed=0, draw=0,
            lengths=None):
    """One Schroedinger-bridge sampler update with the EDM preconditioning folded in (ditsep_sb_step):
    coef (B, 8) f32 device rows [w_prev, w_est, w_y, w_z, c_skip, c_out, f_div, c_in_next]."""
    _cuda(x, f, y, coef, x_out, x_in, y_in, est, noise)
    for n, v in (("x", x), ("f", f), ("y", y), ("coef", coef), ("x_out", x_out), ("x_in", x_in), ("y_in", y_in),
                 ("est", est), ("noise", noise)):
        _f32c(v, n)
    B, n_src = x.shape[0], x.shape[1]
    L = x.numel() // (B * n_src)
    if coef.shape != (B, 8):
        raise _lib.DitsepError(f"sb_step: coef must be ({B}, 8), got {tuple(coef.shape)}")
    for n, v in (("f", f), ("x_out", x_out), ("x_in", x_in), ("est", est), ("noise", noise)):
        if v is not None and v.numel() != x.numel():
            raise _lib.DitsepError(f"sb_step: {n} has {v.numel()} elements, the state has {x.numel()}")
    for n, v in (("y", y), ("y_in", y_in)):
        if v is not None and v.numel() != B * L:
            raise _lib.DitsepError(f"sb_step: {n} has {v.numel()} elements, expected (B, 1, L) = {B * L}")
    check(_lib.load().ditsep_sb_step(_ptr(x), _ptr(f), _ptr(y), _ptr(noise), int(use_noise), seed, draw, _ptr(coef),
                                     _ptr(x_out), _ptr(x_in), _ptr(y_in), _ptr(est), B, n_src, L, _lengths(lengths, B),
                                     _stream()), "sb_step")


def philox_normal(out: torch.Tensor, seed: int, draw: int):
    _cuda(out)
    _f32c(out, "out")
    check(_lib.load().ditsep_philox_normal(_ptr(out), out.numel(), seed, draw, _stream()), "philox_normal")


def box_muller(bits: torch.Tensor, out: torch.Tensor):
    """The Philox stream's uniform -> normal transform on caller-supplied words (int32 view of 32-bit patterns)."""
    _cuda(bits, out)
    _f32c(out, "out")
    if bits.dtype != torch.int32 or not bits.is_contiguous() or bits.numel() != out.numel():
        raise _lib.DitsepError("box_muller: bits must be a contiguous int32 tensor of out's size")
    check(_lib.load().ditsep_box_muller(_ptr(bits), _ptr(out), out.numel(), _stream()), "box_muller")


# ------------------------------------------------------------------ spectral
def stft_compress(xt, mix, window, feat, exponent: float, factor: float, affine: bool, mix_cache=None, mix_mode: int = 0):
    """xt (B,n,T), mix (B,1,T) f32 -> feat NHWC (B,256,W,2(n+1)) f32|bf16 (dense).  mix_cache (B, W, 256, 2) f32 with
    mix_mode 1 (fill) / 2 (use): the mixture channel's transform kept across the evaluations of a sampler run."""
    _cuda(xt, mix, window, feat, mix_cache)
    for n, v in (("xt", xt), ("mix", mix), ("window", window)):
        _f32c(v, n)
    B, n_src, T = xt.shape
    if not feat.is_contiguous() or feat.shape[0] != B or feat.shape[1] != 256 or feat.shape[3] != 2 * (n_src + 1):
        raise _lib.DitsepError(f"feat must be dense (B,256,W,{2 * (n_src + 1)}), got {tuple(feat.shape)}")
    if mix_mode:
        _f32c(mix_cache, "mix_cache")
        if mix_cache is None or tuple(mix_cache.shape) != (B, feat.shape[2], 256, 2):
            raise _lib.DitsepError(f"mix_cache must be ({B}, {feat.shape[2]}, 256, 2) float32")
    check(_lib.load().ditsep_stft_compress_cached(_ptr(xt), _ptr(mix), _ptr(window), _ptr(feat), _dt(feat), B, n_src, T,
                                                  feat.shape[2], exponent, factor, int(affine), _ptr(mix_cache),
                                                  int(mix_mode), _stream()), "stft_compress")


def expand_istft(spec, window, out, exponent: float, factor: float):
    """spec NHWC (B,256,W,2n) f32 dense -> out (B,n,T) f32."""
    _cuda(spec, window, out)
    for n, v in (("spec", spec), ("window", window), ("out", out)):
        _f32c(v, n)
    B, n_src, T = out.shape
    if spec.shape[0] != B or spec.shape[1] != 256 or spec.shape[3] != 2 * n_src:
        raise _lib.DitsepError(f"spec must be (B,256,W,{2 * n_src}), got {tuple(spec.shape)}")
    check(_lib.load().ditsep_expand_istft(_ptr(spec), _ptr(window), _ptr(out), B, n_src, T, spec.shape[2], exponent,
                                          factor, _stream()), "expand_istft")


# ------------------------------------------------------------------ norm / resample
def new_sums(B: int, C: int, device) -> torch.Tensor:
    """Zeroed fixed-point statistics buffer [B][C][4] int64 (include/ditsep_b200.h: value = int + frac * 2^-32).
    Producers ADD to it: it must be zero again before the next score step (the engine zeroes one arena per step)."""
    return torch.zeros(B, C, 4, device=device, dtype=torch.int64)


def _sums_ok(sums, B: int, C: int, who: str):
    if sums.dtype != torch.int64 or not sums.is_contiguous() or sums.numel() != B * C * 4:
        raise _lib.DitsepError(f"{who}: sums must be a contiguous int64 [B={B}][C={C}][4] tensor")


def gn_sums(x, sums):
    """ADDS the per-channel {sum, sum of squares} of x NHWC (B,H,W,C) to the fixed-point sums [B][C][4]."""
    _cuda(x, sums)
    B, H, W, Cc = x.shape
    _sums_ok(sums, B, Cc, "gn_sums")
    check(_lib.load().ditsep_gn_sums(_ptr(x), _dt(x), B, H * W, Cc, nhwc_ld(x), _ptr(sums), _stream()), "gn_sums")


def make_gn_job(srcs, B: int, HW: int, G: int, gamma, beta, eps: float, scale_shift) -> "_lib.GnJob":
    """srcs: 1..4 sums tensors [B][C_i][4] covering the normalised tensor's channels end to end -> ditsep_gn_job."""
    _cuda(gamma, beta, scale_shift, *srcs)
    if not 1 <= len(srcs) <= 4:
        raise _lib.DitsepError("GroupNorm job: 1..4 statistics sources")
    job = _lib.GnJob()
    ctot = 0
    for i, t in enumerate(srcs):
        ci = t.shape[1]
        _sums_ok(t, B, ci, "GroupNorm job")
        job.src_sums[i] = t.data_ptr()
        job.src_C[i] = ci
        ctot += ci
    if scale_shift is not None and (scale_shift.dtype != torch.float32 or not scale_shift.is_contiguous()
                                    or scale_shift.numel() != B * 2 * ctot):
        raise _lib.DitsepError("GroupNorm job: scale_shift must be contiguous f32 (B, 2, C)")
    job.n_src, job.C, job.G, job.eps, job.HW = len(srcs), ctot, G, eps, HW
    job.gamma, job.beta = gamma.data_ptr(), beta.data_ptr()
    job.scale_shift = None if scale_shift is None else scale_shift.data_ptr()
    job._keep = (srcs, gamma, beta, scale_shift)
    return job


def gn_finalize(srcs, B: int, HW: int, G: int, gamma, beta, eps: float, scale_shift):
    """scale_shift (B, 2, C) of GroupNorm(G, C) from the fixed-point sums of its producers (1..4 sources, end to end
    along the channel axis)."""
    job = make_gn_job(srcs, B, HW, G, gamma, beta, eps, scale_shift)
    check(_lib.load().ditsep_gn_finalize(C.byref(job), B, _stream()), "gn_finalize")


def gn_apply_job(x, job, silu: bool, y):
    """y = act(GroupNorm(x)) with the scale / shift formed inside the kernel from the job's fixed-point sums
    (make_gn_job with scale_shift=None): gn_finalize + gn_apply in one launch."""
    _cuda(x, y)
    B, H, W, Cc = x.shape
    if y.dtype != x.dtype or job.C != Cc:
        raise _lib.DitsepError("gn_apply_job: dtype / channel mismatch")
    check(_lib.load().ditsep_gn_apply_job(_ptr(x), _dt(x), B, H * W, C.byref(job), nhwc_ld(x), int(silu), _ptr(y),
                                          nhwc_ld(y), _stream()), "gn_apply_job")


def gn_apply_srcs(x, srcs, G: int, gamma, beta, eps: float, silu: bool, y, cache: dict = None):
    """y = act(GroupNorm(G)(x)) straight from the producers' sums `srcs` (see gn_finalize): ONE launch, no scale / shift
    buffer.  cache: a dict owned by the caller's launch plan; the ditsep_gn_job struct is built on the first call."""
    job = None if cache is None else cache.get("job")
    if job is None:
        B, H, W, _ = x.shape
        job = make_gn_job(srcs, B, H * W, G, gamma, beta, eps, None)
        if cache is not None:
            cache["job"] = job
    gn_apply_job(x, job, silu, y)


def gn_stats(x, G: int, gamma, beta, eps: float, scale_shift):
    """x NHWC (B,H,W,C) -> scale_shift (B,2,C) f32: explicit statistics pass + finalize (allocates a sums buffer:
    not for use inside a captured launch plan)."""
    B, H, W, Cc = x.shape
    sums = new_sums(B, Cc, x.device)
    gn_sums(x, sums)
    gn_finalize([sums], B, H * W, G, gamma, beta, eps, scale_shift)


def gn_small(x, G: int, gamma, beta, eps: float, silu: bool, y):
    """act(GroupNorm(x)) of a small NHWC image batch in ONE launch (ditsep_gn_small); y may alias x."""
    _cuda(x, gamma, beta, y)
    B, H, W, Cc = x.shape
    if y.dtype != x.dtype:
        raise _lib.DitsepError("gn_small: dtype mismatch")
    check(_lib.load().ditsep_gn_small(_ptr(x), _dt(x), B, H * W, Cc, nhwc_ld(x), G, _ptr(gamma), _ptr(beta), eps,
                                      int(silu), _ptr(y), nhwc_ld(y), _stream()), "gn_small")


def gn_apply(x, scale_shift, silu: bool, y):
    _cuda(x, scale_shift, y)
    B, H, W, Cc = x.shape
    if y.dtype != x.dtype:
        raise _lib.DitsepError("gn_apply: dtype mismatch")
    check(_lib.load().ditsep_gn_apply(_ptr(x), _dt(x), B, H * W, Cc, nhwc_ld(x), _ptr(scale_shift), int(silu), _ptr(y),
                                      nhwc_ld(y), _stream()), "gn_apply")


def fir_resample(x, up: bool, out_x=None, scale_shift=None, out_h=None):
    _cuda(x, out_x, scale_shift, out_h)
    B, H, W, Cc = x.shape
    check(_lib.load().ditsep_fir_resample(_ptr(x), _dt(x), B, H, W, Cc, nhwc_ld(x), int(up), _ptr(scale_shift),
                                          _ptr(out_x), 0 if out_x is None else nhwc_ld(out_x), _ptr(out_h),
                                          0 if out_h is None else nhwc_ld(out_h), _stream()), "fir_resample")


# ------------------------------------------------------------------ convolution
def conv_direct(x, w, bias, out, ksize: int, bias_b=None, residual=None, out_scale: float = 1.0):
    """x NHWC, w (Cout, taps, Cin) f32, out NHWC.  out = out_scale*(conv + bias + bias_b[b] + residual)."""
    _cuda(x, w, bias, out, bias_b, residual)
    B, H, W, Cin = x.shape
    Cout = out.shape[3]
    if w.dtype != torch.float32 or not w.is_contiguous() or w.shape != (Cout, ksize * ksize, Cin):
        raise _lib.DitsepError(f"conv_direct: weights must be f32 (Cout,taps,Cin)={Cout, ksize * ksize, Cin}, got {tuple(w.shape)}")
    if residual is not None and residual.dtype != out.dtype:
        raise _lib.DitsepError("conv_direct: residual dtype must equal output dtype")
    check(_lib.load().ditsep_conv_direct(_ptr(x), _dt(x), B, H, W, Cin, nhwc_ld(x), _ptr(w), _ptr(bias), _ptr(bias_b),
                                         0 if bias_b is None else bias_b.stride(0), _ptr(residual),
                                         0 if residual is None else nhwc_ld(residual), out_scale, Cout, ksize,
                                         _ptr(out), _dt(out), nhwc_ld(out), _stream()),
          "conv_direct")


def split_hilo(x, out):
    """x (B,H,W,C) f32 contiguous -> out (B,H,W,ld) bf16 = [hi | lo | hi | 0] (ditsep_split_hilo)."""
    _cuda(x, out)
    if x.dtype != torch.float32 or out.dtype != torch.bfloat16 or not x.is_contiguous() or not out.is_contiguous():
        raise _lib.DitsepError("split_hilo: x must be contiguous f32, out contiguous bf16")
    B, H, W, Cc = x.shape
    check(_lib.load().ditsep_split_hilo(_ptr(x), B * H * W, Cc, _ptr(out), out.shape[3], _stream()), "split_hilo")


class ConvTcPlan:
    """Owns a ditsep_conv_plan (TMA descriptors for one input tensor + weight binding)."""

    def __init__(self, x: torch.Tensor, w_bf16: torch.Tensor, Cout: int, ksize: int, shortcut: torch.Tensor = None):
        """shortcut: optional second NHWC bf16 input (B,H,W,Cx) accumulated through the 1x1 filter packed behind the
        taps of w_bf16 ([CoutPad][ksize^2*Cin + Cx]) -- ditsep_conv_tc_plan_create_sc.  Raises DitsepError when the
        shape cannot be fused (filter not resident, image below one tile); callers fall back to the unfused path."""
        _cuda(x, w_bf16, shortcut)
        B, H, W, Cin = x.shape
        if x.dtype != torch.bfloat16 or w_bf16.dtype != torch.bfloat16 or not w_bf16.is_contiguous():
            raise _lib.DitsepError("conv_tc: bf16 activations and packed bf16 weights required")
        self._h = C.c_void_p()
        self._keep = (x, w_bf16, shortcut)
        self.Cout, self.Cin = Cout, Cin
        self.K = ksize * ksize * Cin + (0 if shortcut is None else shortcut.shape[3])
        self.shape = (B, H, W)
        if w_bf16.numel() % self.K or w_bf16.numel() // self.K < Cout:
            raise _lib.DitsepError(f"conv_tc: packed filter has {w_bf16.numel()} elements, K={self.K}, Cout={Cout}")
        if shortcut is None:
            check(_lib.load().ditsep_conv_tc_plan_create(C.byref(self._h), _ptr(x), B, H, W, Cin, nhwc_ld(x),
                                                         _ptr(w_bf16), Cout, ksize), "conv_tc_plan_create")
        else:
            if shortcut.dtype != torch.bfloat16 or tuple(shortcut.shape[:3]) != (B, H, W):
                raise _lib.DitsepError("conv_tc: shortcut must be bf16 NHWC with the input's B, H, W")
            check(_lib.load().ditsep_conv_tc_plan_create_sc(C.byref(self._h), _ptr(x), B, H, W, Cin, nhwc_ld(x),
                                                            _ptr(w_bf16), Cout, ksize, _ptr(shortcut), shortcut.shape[3],
                                                            nhwc_ld(shortcut)), "conv_tc_plan_create_sc")

    @property
    def tile(self) -> bool:
        """True when this plan runs the persistent tile kernel (whose epilogue can leave GroupNorm statistics)."""
        return bool(_lib.load().ditsep_conv_tc_has_stats(self._h))

    def run(self, bias, out, bias_b=None, residual=None, out_scale: float = 1.0, stats=None):
        """stats: fixed-point sums [B][Cout][4] int64 (new_sums) the epilogue ADDS the output's per-channel
        {sum, sum of squares} to (tile-kernel plans only)."""
        _cuda(bias, out, bias_b, residual, stats)
        if stats is not None:
            if not self.tile:
                raise _lib.DitsepError("conv_tc: this plan (image below one 8x16 tile) cannot produce statistics")
            _sums_ok(stats, self.shape[0], self.Cout, "conv_tc")
        check(_lib.load().ditsep_conv_tc_run(self._h, _ptr(bias), _ptr(bias_b),
                                             0 if bias_b is None else bias_b.stride(0), _ptr(residual),
                                             0 if residual is None else _dt(residual),
                                             0 if residual is None else nhwc_ld(residual), out_scale, _ptr(out),
                                             _dt(out), nhwc_ld(out), _ptr(stats), _stream()), "conv_tc_run")

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            try:
                _lib.load().ditsep_conv_tc_plan_destroy(h)
            except Exception:
                pass
            self._h = None


# ------------------------------------------------------------------ Oobleck VAE (1-D convolutions)
def conv1d_tc(x, T_in: int, bs_in: int, w_bf16, Cout: int, taps: int, dil: int, pad: int, s_view: int, T_out: int, bias=None,
              residual=None, ld_res: int = 0, bs_res: int = 0, out=None, out_act=None, ld_out: int = 0, bs_out: int = 0):
    """ditsep_conv1d_tc on explicit views: x / residual / out / out_act are bf16 (out may be fp32) tensors whose data_ptr is
    the view's first row; rows are ld apart (x: dense, ld = Cin), signals bs apart.  See include/ditsep_b200.h."""
    _cuda(x, w_bf16, bias, residual, out, out_act)
    B = x.shape[0]
    Cin = w_bf16.shape[1] // taps
    if x.dtype != torch.bfloat16 or w_bf16.dtype != torch.bfloat16 or not w_bf16.is_contiguous():
        raise _lib.DitsepError("conv1d_tc: bf16 activations and packed bf16 weights required")
    out_f32 = int(out is not None and out.dtype == torch.float32)
    check(_lib.load().ditsep_conv1d_tc(_ptr(x), B, T_in, Cin, bs_in, _ptr(w_bf16), Cout, taps, dil, pad, s_view, T_out,
                                       _ptr(bias), _ptr(residual), ld_res, bs_res, _ptr(out), _ptr(out_act), out_f32, ld_out,
                                       bs_out, _stream()), "conv1d_tc")


def conv1d_in(x, w, bias, out, out_act):
    """x (B, T) f32 -> conv7 (1 -> C) -> out / ELU(out), [B][T][C] bf16."""
    _cuda(x, w, bias, out, out_act)
    _f32c(x, "x")
    B, T = x.shape
    check(_lib.load().ditsep_conv1d_in(_ptr(x), _ptr(w), _ptr(bias), _ptr(out), _ptr(out_act), B, T, out.shape[-1], _stream()),
          "conv1d_in")


def conv1d_out(x, T: int, ld: int, bs: int, w, out, final_tanh: bool = True):
    """ELU'd activations (view: rows ld, signals bs apart) -> conv7 (C -> 1) -> tanh -> out (B, T) f32."""
    _cuda(x, w, out)
    _f32c(out, "out")
    check(_lib.load().ditsep_conv1d_out(_ptr(x), ld, bs, _ptr(w), _ptr(out), out.shape[0], T, w.shape[1], int(final_tanh),
                                        _stream()), "conv1d_out")


def vae_sample(pre, noise, z):
    """pre [B][T][2L] f32 (mean | scale), noise / z (B, L, T) f32: z = noise * (softplus(scale) + 1e-4) + mean."""
    _cuda(pre, noise, z)
    for n, v in (("pre", pre), ("noise", noise), ("z", z)):
        _f32c(v, n)
    B, L, T = z.shape
    check(_lib.load().ditsep_vae_sample(_ptr(pre), _ptr(noise), _ptr(z), B, T, L, _stream()), "vae_sample")


# ------------------------------------------------------------------ attention / temb / output
def attention(q, k, v, o):
    """q,k,v,o: NHWC views (B,H,W,C) sharing one pixel stride for q/k/v."""
    _cuda(q, k, v, o)
    B, H, W, Cc = q.shape
    ld = nhwc_ld(q)
    if nhwc_ld(k) != ld or nhwc_ld(v) != ld:
        raise _lib.DitsepError("attention: q, k, v must share the pixel stride")
    check(_lib.load().ditsep_attention(_ptr(q), _ptr(k), _ptr(v), _dt(q), B, H * W, Cc, ld, _ptr(o), nhwc_ld(o),
                                       _stream()), "attention")


def temb_mlp(t, gfp_w, w1, b1, w2, b2, act_temb):
    _cuda(t, gfp_w, w1, b1, w2, b2, act_temb)
    check(_lib.load().ditsep_temb_mlp(_ptr(t), _ptr(gfp_w), _ptr(w1), _ptr(b1), _ptr(w2), _ptr(b2), _ptr(act_temb),
                                      t.shape[0], gfp_w.shape[0], _stream()), "temb_mlp")


def temb_dense(act_temb, dense_w, dense_b, out):
    _cuda(act_temb, dense_w, dense_b, out)
    B, K = act_temb.shape
    check(_lib.load().ditsep_temb_dense(_ptr(act_temb), _ptr(dense_w), _ptr(dense_b), _ptr(out), B, K,
                                        dense_w.shape[0], _stream()), "temb_dense")


def output_layer(x, t, w, bias, out, scale_by_sigma: bool):
    """x NHWC (B,H,W,Cin) -> out dense (B,H,W,Cout) f32."""
    _cuda(x, t, w, bias, out)
    B, H, W, Cin = x.shape
    check(_lib.load().ditsep_output_layer(_ptr(x), _dt(x), B, H * W, Cin, nhwc_ld(x), _ptr(t), int(scale_by_sigma),
                                          _ptr(w), _ptr(bias), out.shape[3], _ptr(out), _stream()), "output_layer")
