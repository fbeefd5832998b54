"""CPU stand-ins for the C-ABI wrappers in ``ditsep_b200.ops`` -- TEST INFRASTRUCTURE.

The product path has no CPU fallback.  To check the HOST logic (launch plans, NHWC buffer
aliasing, concat-slot placement, weight packing, sampler loop, noise numbering) on a box
without a GPU, the ``-m "not gpu"`` tests swap every ``ops.*`` kernel wrapper for a small
torch restatement with the SAME calling convention (in-place writes into caller-allocated
NHWC views).  Each stand-in documents the kernel it mirrors; the ``-m gpu`` tests then check
the real kernels against the oracle.
"""
from __future__ import annotations

import contextlib
import math

import torch
import torch.nn.functional as F

from ditsep_b200 import ops

_NAMES = ["conv1d_tc", "conv1d_in", "conv1d_out", "vae_sample", "pc_corrector", "pc_predictor", "pc_prior", "pc_langevin", "pc_perturb", "score_loss", "score_loss_sigma", "energy_envelope", "sb_step", "philox_normal", "stft_compress", "expand_istft", "gn_stats",
          "gn_sums", "gn_finalize", "gn_apply", "gn_apply_srcs", "gn_small", "fir_resample", "conv_direct", "attention", "temb_mlp", "temb_dense", "output_layer"]


def _coef(p, mode, t):
    """make_coef of csrc/pc_step.cu, vectorised over the batch (t: (B,))."""
    ratio = p.sigma_max / p.sigma_min
    logsig = math.log(ratio)
    dt = 1.0 / p.N
    z = torch.zeros_like(t)
    c = dict(a0=z + 1, a1=z, a2=z, a3=z, a4=z, b0=z, b1=z)
    if p.sde == 0:
        lam = p.p0
        srp = ratio ** (2 * t)
        mult = p.sigma_min ** 2
        ev1 = mult * (srp - 1)
        ev2 = mult * (srp - torch.exp(-2 * lam * t)) / (1 + lam / logsig)
        r1, r2 = ev1.sqrt(), ev2.sqrt()
        if mode == "corrector":
            st = 2 * p.snr ** 2
            c.update(a2=st * ev2, a3=st * (ev1 - ev2), b0=2 * p.snr * r2, b1=2 * p.snr * (r1 - r2))
        elif mode == "predictor":
            G = p.sigma_min * ratio ** t * math.sqrt(2 * logsig) * math.sqrt(dt)
            c.update(a0=z + 1 + lam * dt, a1=z - lam * dt, a2=G * G, b0=G)
        elif mode == "perturb":
            e = torch.exp(-lam * t)
            c.update(a0=e, a1=1 - e, b0=r2, b1=r1 - r2)
        else:
            c.update(a0=z, a4=z + 0.5, b0=r2, b1=r1 - r2)
    else:
        th = p.p0
        std = torch.sqrt(p.sigma_min ** 2 * torch.exp(-2 * th * t) * (torch.exp(2 * (th + logsig) * t) - 1) * logsig
                         / (th + logsig))
        if mode == "corrector":
            eps = 2 * (p.snr * std) ** 2
            c.update(a2=eps, b0=torch.sqrt(2 * eps))
        elif mode == "predictor":
            G = p.sigma_min * ratio ** t * math.sqrt(2 * logsig) * math.sqrt(dt)
            c.update(a0=z + 1 + th * dt, a4=z - th * dt, a2=G * G, b0=G)
        elif mode == "perturb":
            e = torch.exp(-th * t)
            c.update(a0=e, a4=1 - e, b0=std)
        else:
            c.update(a0=z, a4=z + 1, b0=std)
    return c


def _pc(p, mode, x, score, t, y, noise, x_mean, lengths=None):
    if noise is None:
        raise AssertionError("the CPU stand-in needs a noise tape (Philox lives in the kernel)")
    B = x.shape[0]
    tt = torch.ones(B) if t is None else t
    c = _coef(p, mode, tt.double())
    v = lambda a: a.reshape(B, *([1] * (x.ndim - 1)))  # noqa: E731
    xs = torch.zeros_like(x, dtype=torch.float64) if mode == "prior" else x.double()
    s = torch.zeros_like(xs) if score is None else score.double()
    z = noise.double()
    xb, sb, zb = (a.mean(dim=1, keepdim=True) for a in (xs, s, z))
    yy = 0.0 if y is None else y.double().reshape(B, 1, *x.shape[2:])
    env = getattr(p, "_keep", None)  # PriorMixSDE: the tensor behind ditsep_sde_params.elem_scale
    ev = 1.0 if env is None else env.double().reshape(B, 1, *x.shape[2:])
    m = v(c["a0"]) * xs + v(c["a1"]) * xb + ev * ev * (v(c["a2"]) * s + v(c["a3"]) * sb) + v(c["a4"]) * yy
    xo = m + ev * (v(c["b0"]) * z + v(c["b1"]) * zb)
    if lengths is not None:  # ragged batch: state kept at exactly 0 past each row's length (csrc/pc_step.cu)
        L = x[0, 0].numel()
        live = (torch.arange(L)[None] < lengths.reshape(B, 1).long()).reshape(B, 1, *x.shape[2:])
        xo, m = xo * live, m * live
    x.copy_(xo.float())
    if x_mean is not None:
        x_mean.copy_(m.float())


def pc_corrector(p, x, score, t, noise=None, seed=0, draw=0, x_mean=None, lengths=None):
    _pc(p, "corrector", x, score, t, None, noise, x_mean, lengths)


def pc_predictor(p, x, score, t, y, noise=None, seed=0, draw=0, x_mean=None, lengths=None):
    _pc(p, "predictor", x, score, t, y if p.sde == 1 else None, noise, x_mean, lengths)


def pc_prior(p, x, y, noise=None, seed=0, draw=0, lengths=None):
    _pc(p, "prior", x, None, None, y, noise, None, lengths)


def pc_perturb(p, x, t, y=None, noise=None, seed=0, draw=0, x_mean=None):
    _pc(p, "perturb", x, None, t, y if p.sde == 1 else None, noise, x_mean)


def score_loss(p, pred, t, noise=None, seed=0, draw=0):
    """csrc/pc_step.cu::score_loss_*_kernel: mean over (source, sample) of (b0 pred + b1 mean_src(pred) + z)^2."""
    if noise is None:
        raise AssertionError("the CPU stand-in needs the noise tensor (Philox lives in the kernel)")
    B = pred.shape[0]
    c = _coef(p, "perturb", t.double())
    v = lambda a: a.reshape(B, *([1] * (pred.ndim - 1)))  # noqa: E731
    pr = pred.double()
    r = v(c["b0"]) * pr + v(c["b1"]) * pr.mean(dim=1, keepdim=True) + noise.double()
    return r.square().flatten(1).mean(dim=1).float()


def score_loss_sigma(pred, sigma, noise=None, seed=0, draw=0):
    if noise is None:
        raise AssertionError("the CPU stand-in needs the noise tensor (Philox lives in the kernel)")
    B = pred.shape[0]
    r = sigma.double().reshape(B, *([1] * (pred.ndim - 1))) * pred.double() + noise.double()
    return r.square().flatten(1).mean(dim=1).float()


def pc_langevin(n_src, x, score, noise, step, x_mean=None, lengths=None):
    eps = step.double().reshape(())
    m = x.double() + eps * score.double()
    xo = m + torch.sqrt(2 * eps) * noise.double()
    x.copy_(xo.float())
    if x_mean is not None:
        x_mean.copy_(m.float())


def energy_envelope(mix, out, avg_len):
    """csrc/pc_step.cu::energy_envelope_kernel: zero-padded box sum of mix^2 over [l - k/2, l - k/2 + k), in double."""
    B, L = mix.shape[0], mix.numel() // mix.shape[0]
    sq = torch.nn.functional.pad(mix.double().reshape(B, L) ** 2, (avg_len // 2, avg_len - 1 - avg_len // 2))
    cs = torch.cat([torch.zeros(B, 1, dtype=torch.float64), sq.cumsum(dim=1)], dim=1)
    box = (cs[:, avg_len:] - cs[:, :-avg_len]) / avg_len
    out.copy_((0.5 * box.clamp(min=1e-4).sqrt()).float().reshape(out.shape))


def sb_step(x, f, y, coef, *, x_out=None, x_in=None, y_in=None, est=None, noise=None, use_noise=False, seed=0, draw=0,
            lengths=None):
    """csrc/pc_step.cu::sb_step_kernel: est = c_skip x + c_out (f / f_div); x' = w_prev x + w_est est + (w_z z | w_y y);
    x_in = c_in x', y_in = c_in y -- fp32, individually rounded products like the kernel."""
    c = [coef[:, i].reshape(-1, 1, 1) for i in range(8)]
    e = c[4] * x + c[5] * (f / c[6]) if f is not None else torch.zeros_like(x)
    v = c[0] * x + c[1] * e
    if int(use_noise) != 1:
        v = v + c[2] * y
    if int(use_noise) != 0:
        if noise is None:
            raise NotImplementedError("the Philox stream only exists on the device: CPU tests pass a noise tape")
        v = v + c[3] * noise
    if lengths is not None:
        live = torch.arange(x.shape[-1])[None, None, :] < lengths.reshape(-1, 1, 1)
        v, e = v * live, e * live
    u = c[7] * v
    yi = c[7] * y if y is not None else None
    if lengths is not None and yi is not None:
        yi = yi * live
    for dst, src in ((x_in, u), (y_in, yi), (est, e), (x_out, v)):  # x_out last: it may alias x
        if dst is not None:
            dst.copy_(src.reshape(dst.shape))


def _view(t, B, rows, C, ld, bs):
    """The (B, rows, C) view the C ABI describes by (first row pointer, row stride ld, signal stride bs)."""
    return torch.as_strided(t, (B, rows, C), (bs, ld, 1), t.storage_offset())


def conv1d_tc(x, T_in, bs_in, w_bf16, Cout, taps, dil, pad, s_view, T_out, bias=None, residual=None, ld_res=0, bs_res=0,
              out=None, out_act=None, ld_out=0, bs_out=0):
    """ditsep_conv1d_tc: out[t] = sum_k W_k x[t * s + k * dil - pad] (zero outside [0, T_in)), fp32 accumulation of the bf16
    operands, + bias + residual; raw and / or ELU output through the views."""
    B = x.shape[0]
    Cin = w_bf16.shape[1] // taps
    xv = _view(x, B, T_in, Cin, Cin, bs_in).float()
    w = w_bf16.float().reshape(Cout, taps, Cin)
    t = torch.arange(T_out)
    acc = torch.zeros(B, T_out, Cout)
    for k in range(taps):
        idx = t * s_view + k * dil - pad
        ok = (idx >= 0) & (idx < T_in)
        g = torch.zeros(B, T_out, Cin)
        g[:, ok] = xv[:, idx[ok]]
        acc += g @ w[:, k].t()
    if bias is not None:
        acc += bias
    if residual is not None:
        acc += _view(residual, B, T_out, Cout, ld_res, bs_res).float()
    stored = acc
    if out is not None:
        _view(out, B, T_out, Cout, ld_out, bs_out).copy_(acc)
        stored = _view(out, B, T_out, Cout, ld_out, bs_out).float()
    if out_act is not None:
        _view(out_act, B, T_out, Cout, ld_out, bs_out).copy_(F.elu(stored))


def conv1d_in(x, w, bias, out, out_act):
    y = F.conv1d(x[:, None].float(), w[:, None].float(), bias, padding=3).permute(0, 2, 1)
    out.copy_(y)
    out_act.copy_(F.elu(out.float()))


def conv1d_out(x, T, ld, bs, w, out, final_tanh=True):
    B, C = out.shape[0], w.shape[1]
    xv = _view(x, B, T, C, ld, bs).float().permute(0, 2, 1)
    y = F.conv1d(xv, w.t()[None].float(), None, padding=3)[:, 0]
    out.copy_(torch.tanh(y) if final_tanh else y)


def vae_sample(pre, noise, z):
    L = z.shape[1]
    mean, scale = pre[..., :L].permute(0, 2, 1), pre[..., L:].permute(0, 2, 1)
    z.copy_(noise * (F.softplus(scale) + 1e-4) + mean)


def philox_normal(out, seed, draw):
    out.normal_(generator=torch.Generator().manual_seed(seed * 1000003 + draw))


def stft_compress(xt, mix, window, feat, exponent, factor, affine, mix_cache=None, mix_mode=0):
    """ditsep_stft_compress: right pad 382, STFT 510/128 centre zero pad, |X|^e e^{j arg} f,
    [Re.., Im..], frames zero-padded to W, optional 2x-1 over the whole tile."""
    x = torch.cat((xt, mix), dim=1)
    B, C, T = x.shape
    z = F.pad(x.reshape(B * C, T), (255, 255 + 382))
    s = torch.stft(z, 510, hop_length=128, win_length=510, window=window, center=False, onesided=True,
                   return_complex=True).reshape(B, C, 256, -1)
    if exponent == 0:
        s = torch.log1p(s.abs()) * torch.exp(1j * s.angle())
    elif exponent != 1:
        s = s.abs() ** exponent * torch.exp(1j * s.angle())
    s = s * factor
    f = torch.cat((s.real, s.imag), dim=1)  # (B, 2C, 256, frames)
    f = F.pad(f, (0, feat.shape[2] - f.shape[-1]))
    if affine:
        f = 2 * f - 1
    feat.copy_(f.permute(0, 2, 3, 1).to(feat.dtype))


def expand_istft(spec, window, out, exponent, factor):
    B, n, T = out.shape
    frames = 1 + (T + 382) // 128
    s = spec[:, :, :frames, :].permute(0, 3, 1, 2)  # (B, 2n, 256, frames)
    c = torch.complex(s[:, :n].contiguous(), s[:, n:].contiguous()) / abs(factor)
    if exponent == 0:
        c = (torch.exp(c.abs()) - 1) * torch.exp(1j * c.angle())
    elif exponent != 1:
        c = c.abs() ** (1 / exponent) * torch.exp(1j * c.angle())
    y = torch.istft(c.reshape(B * n, 256, frames), 510, hop_length=128, win_length=510, window=window, center=True,
                    onesided=True).reshape(B, n, -1)
    out.copy_(y[..., :T])


def _gn_from_channel_sums(s, HW, G, gamma, beta, eps, scale_shift):
    """s: (B, C, 2) double {sum, sum of squares} -> scale / shift, as gn_job_image of csrc/norm_fir.cu."""
    B, C = s.shape[:2]
    cpg = C // G
    g = s.reshape(B, G, cpg, 2).sum(dim=2)
    n = HW * cpg
    mean = g[..., 0] / n
    var = (g[..., 1] / n - mean * mean).clamp(min=0)
    rstd = 1 / torch.sqrt(var + eps)
    sc = rstd.repeat_interleave(cpg, dim=1) * gamma.double()
    sh = beta.double() - mean.repeat_interleave(cpg, dim=1) * sc
    scale_shift[:, 0].copy_(sc.float())
    scale_shift[:, 1].copy_(sh.float())


def _fx_encode(v):
    """double -> (int, frac * 2^32) pair of the fixed-point sums (common.cuh fx_add)."""
    fl = torch.floor(v)
    return fl.to(torch.int64), ((v - fl) * 4294967296.0).to(torch.int64)


def _fx_decode(sums):
    """[B][C][4] int64 -> (B, C, 2) double."""
    d = sums.double()
    return torch.stack((d[..., 0] + d[..., 1] / 4294967296.0, d[..., 2] + d[..., 3] / 4294967296.0), dim=-1)


def gn_sums(x, sums):
    """ditsep_gn_sums: ADDS per-channel {sum, sum of squares} to the fixed-point sums [B][C][4]."""
    B, H, W, C = x.shape
    xf = x.double().reshape(B, H * W, C)
    for k, v in ((0, xf.sum(dim=1)), (2, (xf * xf).sum(dim=1))):
        hi, lo = _fx_encode(v)
        sums[..., k] += hi
        sums[..., k + 1] += lo


def gn_finalize(srcs, B, HW, G, gamma, beta, eps, scale_shift):
    """ditsep_gn_finalize: sources laid end to end along the channel axis."""
    s = torch.cat([_fx_decode(t) for t in srcs], dim=1)  # (B, C, 2)
    _gn_from_channel_sums(s, HW, G, gamma, beta, eps, scale_shift)


def gn_apply_srcs(x, srcs, G, gamma, beta, eps, silu, y, cache=None):
    B, H, W, C = x.shape
    ss = torch.empty(B, 2, C)
    gn_finalize(srcs, B, H * W, G, gamma, beta, eps, ss)
    gn_apply(x, ss, silu, y)


def gn_stats(x, G, gamma, beta, eps, scale_shift):
    B, H, W, C = x.shape
    xf = x.double().reshape(B, H * W, C)
    _gn_from_channel_sums(torch.stack((xf.sum(dim=1), (xf * xf).sum(dim=1)), dim=-1), H * W, G, gamma, beta, eps,
                          scale_shift)


def gn_small(x, G, gamma, beta, eps, silu, y):
    B, H, W, C = x.shape
    ss = torch.empty(B, 2, C)
    gn_stats(x, G, gamma, beta, eps, ss)
    gn_apply(x, ss, silu, y)


def gn_apply(x, scale_shift, silu, y):
    v = x.float() * scale_shift[:, 0][:, None, None, :] + scale_shift[:, 1][:, None, None, :]
    y.copy_((F.silu(v) if silu else v).to(y.dtype))


def _fir(x_nhwc, up):
    """upfirdn2d with k = outer([1,3,3,1])/64 (x4 when upsampling), pads (2,1) up / (1,1) down."""
    x = x_nhwc.float().permute(0, 3, 1, 2)
    n, c, h, w = x.shape
    k1 = torch.tensor([1.0, 3.0, 3.0, 1.0])
    k = torch.outer(k1, k1) / 64.0
    z = x.reshape(n * c, 1, h, w)
    if up:
        buf = z.new_zeros(n * c, 1, 2 * h, 2 * w)
        buf[:, :, ::2, ::2] = z
        z = F.conv2d(F.pad(buf, [2, 1, 2, 1]), (k * 4).flip(0, 1)[None, None])
    else:
        z = F.conv2d(F.pad(z, [1, 1, 1, 1]), k.flip(0, 1)[None, None])[:, :, ::2, ::2]
    return z.reshape(n, c, z.shape[-2], z.shape[-1]).permute(0, 2, 3, 1)


def fir_resample(x, up, out_x=None, scale_shift=None, out_h=None):
    if out_x is not None:
        out_x.copy_(_fir(x, up).to(out_x.dtype))
    if out_h is not None:
        v = F.silu(x.float() * scale_shift[:, 0][:, None, None, :] + scale_shift[:, 1][:, None, None, :])
        out_h.copy_(_fir(v, up).to(out_h.dtype))


def conv_direct(x, w, bias, out, ksize, bias_b=None, residual=None, out_scale=1.0):
    cout, taps, cin = w.shape
    wt = w.reshape(cout, ksize, ksize, cin).permute(0, 3, 1, 2)
    y = F.conv2d(x.float().permute(0, 3, 1, 2), wt, bias, padding=ksize // 2).permute(0, 2, 3, 1)
    if bias_b is not None:
        y = y + bias_b[:, None, None, :]
    if residual is not None:
        y = y + residual.float()
    out.copy_((y * out_scale).to(out.dtype))


def attention(q, k, v, o):
    B, H, W, C = q.shape
    qf, kf, vf = (a.float().reshape(B, H * W, C) for a in (q, k, v))
    p = torch.softmax(qf @ kf.transpose(1, 2) * C ** -0.5, dim=-1)
    o.copy_((p @ vf).reshape(B, H, W, C).to(o.dtype))


def temb_mlp(t, gfp_w, w1, b1, w2, b2, act_temb):
    xp = torch.log(t)[:, None] * gfp_w[None, :] * 2 * math.pi
    e = torch.cat([torch.sin(xp), torch.cos(xp)], dim=-1)
    h = F.silu(F.linear(e, w1, b1))
    act_temb.copy_(F.silu(F.linear(h, w2, b2)))


def temb_dense(act_temb, dense_w, dense_b, out):
    out.copy_(F.linear(act_temb, dense_w, dense_b))


def output_layer(x, t, w, bias, out, scale_by_sigma):
    v = x.float()
    if scale_by_sigma:
        v = v / t[:, None, None, None]
    out.copy_(v @ w.t() + bias)


@contextlib.contextmanager
def emulated_ops():
    """Swap the kernel wrappers for the stand-ins above (and back)."""
    # the three policy hooks of ops: no tcgen05 plans, no CUDA graphs, CPU tensors accepted -- patched HERE, so the
    # package itself carries no emulation switch
    policy = {"tc_enabled": lambda: False, "graphs_enabled": lambda: False, "require_device": lambda t, what: None}
    saved = {n: getattr(ops, n) for n in list(_NAMES) + list(policy)}
    try:
        for n in _NAMES:
            setattr(ops, n, globals()[n])
        for n, f in policy.items():
            setattr(ops, n, f)
        yield
    finally:
        for n, f in saved.items():
            setattr(ops, n, f)
