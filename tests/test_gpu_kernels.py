"""Per-kernel parity (-m gpu): every C-ABI entry point against a torch restatement of the
reference op (tests/emu_ops.py, fp64/fp32 on the CPU) on the same seeded inputs.
Tolerances are written next to each check: fp32 kernels ~1e-5 relative, bf16 storage 2^-8."""
import math

import pytest
import torch

import emu_ops as E
from ditsep_b200 import ops
from ditsep_b200._lib import SdeParams

pytestmark = pytest.mark.gpu
DEV = "cuda"


def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / b.norm().clamp(min=1e-30))


def g(seed):
    return torch.Generator().manual_seed(seed)


# ------------------------------------------------------------------ fused SDE updates
@pytest.mark.parametrize("sde,n_src", [(0, 2), (1, 2), (1, 3), (0, 3)])
@pytest.mark.parametrize("L", [32000, 4099, 7])
def test_pc_updates(sde, n_src, L):
    B = 3
    p = SdeParams(sde, n_src, 2.0 if sde == 0 else 1.5, 0.05, 0.5, 30, 0.5)
    x = torch.randn(B, n_src, L, generator=g(1))
    s = torch.randn(B, n_src, L, generator=g(2))
    z = torch.randn(B, n_src, L, generator=g(3))
    y = torch.randn(B, 1, L, generator=g(4))
    t = torch.tensor([1.0, 0.5, 0.03])
    for mode in ("corrector", "predictor", "prior"):
        if mode == "prior" and sde == 0 and n_src != 2:
            continue
        xr, mr = x.clone(), torch.empty_like(x)
        xd, md = x.clone().to(DEV), torch.empty_like(x).to(DEV)
        if mode == "corrector":
            E.pc_corrector(p, xr, s, t, noise=z, x_mean=mr)
            ops.pc_corrector(p, xd, s.to(DEV), t.to(DEV), noise=z.to(DEV), x_mean=md)
        elif mode == "predictor":
            E.pc_predictor(p, xr, s, t, y, noise=z, x_mean=mr)
            ops.pc_predictor(p, xd, s.to(DEV), t.to(DEV), y.to(DEV), noise=z.to(DEV), x_mean=md)
        else:
            E.pc_prior(p, xr, y, noise=z)
            ops.pc_prior(p, xd, y.to(DEV), noise=z.to(DEV))
        assert rel(xd, xr) < 2e-6, (mode, rel(xd, xr))  # fp32 elementwise, coefficients via powf/expf
        if mode != "prior":
            assert rel(md, mr) < 2e-6, mode


def test_pc_philox_noise_is_standard_normal_and_keyed():
    n = 1 << 22
    a = torch.empty(n, device=DEV)
    b = torch.empty(n, device=DEV)
    ops.philox_normal(a, seed=4321, draw=0)
    ops.philox_normal(b, seed=4321, draw=0)
    assert torch.equal(a, b)  # deterministic in (seed, draw)
    ops.philox_normal(b, seed=4321, draw=1)
    assert abs(float((a * b).mean())) < 5e-3  # draws are independent
    m, v = float(a.mean()), float(a.var())
    k = float(((a - m) ** 4).mean() / v ** 2)
    assert abs(m) < 3e-3 and abs(v - 1) < 5e-3 and abs(k - 3) < 3e-2
    # the fused update with noise=None consumes the same stream: x = 0 + 1*z for OUVE prior with y = 0
    p = SdeParams(1, 2, 1.5, 0.05, 0.5, 30, 0.5)
    x = torch.empty(2, 2, 1024, device=DEV)
    ops.pc_prior(p, x, torch.zeros(2, 1, 1024, device=DEV), noise=None, seed=7, draw=3)
    z = torch.empty(2 * 2 * 1024, device=DEV)
    ops.philox_normal(z, seed=7, draw=3)
    std1 = math.sqrt(0.05 ** 2 * math.exp(-3.0) * (math.exp(2 * (1.5 + math.log(10)) * 1.0) - 1) * math.log(10) / (1.5 + math.log(10)))
    assert rel(x.flatten(), z * std1) < 1e-5


def test_box_muller_endpoint_uniforms_are_finite():
    """Uniform words at and next to both ends of the range (u -> 1 makes __logf(u) ~ -6e-8, which the fast-math
    intrinsic may return with the wrong sign: sqrt of a negative = NaN without the clamp) and a dense sweep of the top
    2^16 words: every output finite, radius 0 at u = 1."""
    ends = [0xFFFFFFFF, 0xFFFFFF00, 0xFFFFFE00, 0xFFFFFF7F, 0xFFFFFF80, 0, 1, 0x80000000]
    top = torch.arange(0x100000000 - (1 << 16), 0x100000000, dtype=torch.int64)
    words = torch.cat([torch.tensor(ends, dtype=torch.int64), top])
    words = words[:words.numel() // 2 * 2]
    # radii words (4q, 4q+2) take the test values; angle words a fixed pattern
    bits = torch.empty(words.numel() * 2, dtype=torch.int64)
    bits[0::2] = words
    bits[1::2] = 0x12345678
    bits32 = (bits & 0xFFFFFFFF).to(torch.int64)
    bits32 = torch.where(bits32 >= 2 ** 31, bits32 - 2 ** 32, bits32).to(torch.int32).to(DEV)
    out = torch.empty(bits32.numel(), device=DEV)
    ops.box_muller(bits32, out)
    assert torch.isfinite(out).all()
    r = torch.hypot(out[0::2], out[1::2]).cpu()
    assert float(r[0]) < 1e-3  # u = 1 -> radius ~ 0
    assert float(r[5]) > 6.0    # u = 2^-33 -> radius sqrt(-2 ln u) ~ 6.76


# ------------------------------------------------------------------ spectral
@pytest.mark.parametrize("B,n_src,T", [(1, 2, 32000), (2, 2, 7810), (3, 3, 5001), (1, 1, 300)])
@pytest.mark.parametrize("affine", [False, True])
def test_stft_compress(B, n_src, T, affine):
    xt = torch.randn(B, n_src, T, generator=g(5))
    mix = torch.randn(B, 1, T, generator=g(6))
    win = torch.hann_window(510)
    frames = 1 + (T + 382) // 128
    W = (frames + 63) // 64 * 64
    C2 = 2 * (n_src + 1)
    ref = torch.empty(B, 256, W, C2)
    E.stft_compress(xt, mix, win, ref, 0.5, 0.15, affine)
    out = torch.empty(B, 256, W, C2, device=DEV)
    ops.stft_compress(xt.to(DEV), mix.to(DEV), win.to(DEV), out, 0.5, 0.15, affine)
    # sqrt compression amplifies the relative error of small bins: 1e-4 in rel-L2 over the tile
    assert rel(out, ref) < 1e-4, rel(out, ref)
    assert float((out.cpu() - ref).abs().max()) < 5e-3
    outb = torch.empty(B, 256, W, C2, device=DEV, dtype=torch.bfloat16)
    ops.stft_compress(xt.to(DEV), mix.to(DEV), win.to(DEV), outb, 0.5, 0.15, affine)
    assert rel(outb.float(), ref) < 6e-3  # bf16 storage
    # exponent 1 path = plain STFT
    E.stft_compress(xt, mix, win, ref, 1.0, 1.0, False)
    ops.stft_compress(xt.to(DEV), mix.to(DEV), win.to(DEV), out, 1.0, 1.0, False)
    assert rel(out, ref) < 2e-5


@pytest.mark.parametrize("B,n_src,T", [(1, 2, 32000), (2, 2, 7810), (2, 3, 5001)])
def test_expand_istft(B, n_src, T):
    frames = 1 + (T + 382) // 128
    W = (frames + 63) // 64 * 64
    spec = 0.1 * torch.randn(B, 256, W, 2 * n_src, generator=g(7))
    win = torch.hann_window(510)
    ref = torch.empty(B, n_src, T)
    E.expand_istft(spec, win, ref, 0.5, 0.15)
    out = torch.empty(B, n_src, T, device=DEV)
    ops.expand_istft(spec.to(DEV), win.to(DEV), out, 0.5, 0.15)
    assert rel(out, ref) < 2e-5, rel(out, ref)


def test_spectral_log_transform():
    """transform="log" (score_models.py:49-51,66-68): exponent 0 in the C ABI = log1p(|X|) e^{j arg X} |factor| and back."""
    B, n_src, T = 2, 2, 7810
    xt, mix = torch.randn(B, n_src, T, generator=g(8)), torch.randn(B, 1, T, generator=g(9))
    win = torch.hann_window(510)
    frames = 1 + (T + 382) // 128
    W = (frames + 63) // 64 * 64
    ref = torch.empty(B, 256, W, 2 * (n_src + 1))
    E.stft_compress(xt, mix, win, ref, 0.0, 0.15, False)
    out = torch.empty(B, 256, W, 2 * (n_src + 1), device=DEV)
    ops.stft_compress(xt.to(DEV), mix.to(DEV), win.to(DEV), out, 0.0, 0.15, False)
    assert rel(out, ref) < 2e-5, rel(out, ref)
    spec = 0.1 * torch.randn(B, 256, W, 2 * n_src, generator=g(10))
    back_ref = torch.empty(B, n_src, T)
    E.expand_istft(spec, win, back_ref, 0.0, 0.15)
    back = torch.empty(B, n_src, T, device=DEV)
    ops.expand_istft(spec.to(DEV), win.to(DEV), back, 0.0, 0.15)
    assert rel(back, back_ref) < 2e-5, rel(back, back_ref)
    # round trip through both kernels: expand(compress(x)) == x
    rt = torch.empty(B, n_src, T, device=DEV)
    ops.expand_istft(out[..., [0, 1, 3, 4]].contiguous(), win.to(DEV), rt, 0.0, 0.15)
    assert rel(rt, xt) < 1e-4


def test_stft_roundtrip_full_size():
    """size-independent property at the benchmark shape: iSTFT(expand(compress(STFT(x)))) == x."""
    B, n, T = 32, 2, 32000
    xt = torch.randn(B, n + 1, T, generator=g(8)).to(DEV)
    win = torch.hann_window(510, device=DEV)
    feat = torch.empty(B, 256, 256, 2 * (n + 1), device=DEV)
    ops.stft_compress(xt[:, :n].contiguous(), xt[:, n:].contiguous(), win, feat, 0.5, 0.15, False)
    out = torch.empty(B, n + 1, T, device=DEV)
    ops.expand_istft(feat, win, out, 0.5, 0.15)
    assert rel(out, xt) < 1e-5


# ------------------------------------------------------------------ GroupNorm / FIR
@pytest.mark.parametrize("C,ld,H,W,B", [(64, 64, 32, 32, 2), (128, 192, 16, 8, 3), (192, 192, 8, 8, 1), (256, 256, 4, 4, 2),
                                        (64, 128, 256, 256, 1)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_groupnorm(C, ld, H, W, B, dtype):
    buf = (torch.randn(B, H, W, ld, generator=g(9)) * 2 + 0.5).to(dtype)
    x = buf[..., ld - C:]
    gamma = 1 + 0.1 * torch.randn(C, generator=g(10))
    beta = 0.1 * torch.randn(C, generator=g(11))
    G = min(C // 4, 32)
    ss_ref = torch.empty(B, 2, C)
    E.gn_stats(x, G, gamma, beta, 1e-6, ss_ref)
    xd = buf.to(DEV)[..., ld - C:]
    ss = torch.empty(B, 2, C, device=DEV)
    ops.gn_stats(xd, G, gamma.to(DEV), beta.to(DEV), 1e-6, ss)
    assert rel(ss, ss_ref) < 1e-5
    # the fixed-point sums themselves: exact integers, so two passes are bit-identical and image b alone == in the batch
    s1, s2 = ops.new_sums(B, C, DEV), ops.new_sums(B, C, DEV)
    ops.gn_sums(xd, s1)
    ops.gn_sums(xd, s2)
    assert torch.equal(s1, s2)
    sb = ops.new_sums(1, C, DEV)
    ops.gn_sums(xd[B - 1:], sb)
    assert torch.equal(sb[0], s1[B - 1])
    tot = E._fx_decode(s1.cpu())
    xf = x.double().reshape(B, H * W, C)
    assert rel(tot[..., 1], (xf * xf).sum(dim=1)) < 1e-5
    for silu in (False, True):
        y_ref = torch.empty(B, H, W, C, dtype=dtype)
        E.gn_apply(x, ss_ref, silu, y_ref)
        y = torch.empty(B, H, W, C, device=DEV, dtype=dtype)
        ops.gn_apply(xd, ss, silu, y)
        assert rel(y.float(), y_ref.float()) < (1e-5 if dtype == torch.float32 else 6e-3)
    # in place
    ybuf = buf.to(DEV).clone()
    yv = ybuf[..., ld - C:]
    ops.gn_apply(yv, ss, True, yv)
    assert rel(yv.float(), y_ref.float()) < (1e-5 if dtype == torch.float32 else 6e-3)


@pytest.mark.parametrize("B,H,W,C,ld,dtype,silu", [(3, 16, 16, 128, 128, torch.bfloat16, True),
                                                   (2, 4, 4, 256, 256, torch.bfloat16, True),
                                                   (2, 8, 8, 64, 192, torch.float32, False),
                                                   (1, 16, 12, 64, 64, torch.float32, True)])
def test_gn_small_one_launch(B, H, W, C, ld, dtype, silu):
    """ditsep_gn_small = statistics + normalise (+ SiLU) of a small image in one launch, against the two-step
    definition (emulated gn_stats + gn_apply); also in place and on a channel slice of a wider buffer."""
    buf = (torch.randn(B, H, W, ld, generator=g(40)) * 1.7 + 0.3).to(dtype)
    x = buf[..., ld - C:]
    G = min(C // 4, 32)
    gamma, beta = 1 + 0.1 * torch.randn(C, generator=g(41)), 0.1 * torch.randn(C, generator=g(42))
    ref = torch.empty(B, H, W, C, dtype=dtype)
    E.gn_small(x, G, gamma, beta, 1e-6, silu, ref)
    xd = buf.to(DEV)[..., ld - C:]
    out = torch.empty(B, H, W, C, device=DEV, dtype=dtype)
    ops.gn_small(xd, G, gamma.to(DEV), beta.to(DEV), 1e-6, silu, out)
    tol = 6e-3 if dtype == torch.bfloat16 else 2e-5
    assert rel(out.float(), ref.float()) < tol
    ops.gn_small(xd, G, gamma.to(DEV), beta.to(DEV), 1e-6, silu, xd)  # in place
    assert torch.equal(xd, out)


@pytest.mark.parametrize("up", [False, True])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_fir_resample(up, dtype):
    B, H, W, C = 2, 16, 8, 64
    x = torch.randn(B, H, W, C, generator=g(12)).to(dtype)
    ss = torch.stack([1 + 0.1 * torch.randn(B, C, generator=g(13)), 0.1 * torch.randn(B, C, generator=g(14))], dim=1)
    Ho, Wo = (2 * H, 2 * W) if up else (H // 2, W // 2)
    rx, rh = torch.empty(B, Ho, Wo, C, dtype=dtype), torch.empty(B, Ho, Wo, C, dtype=dtype)
    E.fir_resample(x, up, out_x=rx, scale_shift=ss, out_h=rh)
    ox = torch.empty(B, Ho, Wo, C, device=DEV, dtype=dtype)
    oh = torch.empty(B, Ho, Wo, C, device=DEV, dtype=dtype)
    ops.fir_resample(x.to(DEV), up, out_x=ox, scale_shift=ss.to(DEV), out_h=oh)
    tol = 1e-5 if dtype == torch.float32 else 6e-3
    assert rel(ox.float(), rx.float()) < tol and rel(oh.float(), rh.float()) < tol
    ox.zero_()
    ops.fir_resample(x.to(DEV), up, out_x=ox)
    assert rel(ox.float(), rx.float()) < tol
    # narrow (6-channel pyramid) path, fp32
    xn = torch.randn(B, H, W, 6, generator=g(15))
    rn = torch.empty(B, Ho, Wo, 6)
    E.fir_resample(xn, up, out_x=rn)
    on = torch.empty(B, Ho, Wo, 6, device=DEV)
    ops.fir_resample(xn.to(DEV), up, out_x=on)
    assert rel(on, rn) < 1e-5


# ------------------------------------------------------------------ convolutions
def _conv_case(B, H, W, Cin, Cout, k, seed, ld_extra=0):
    buf = torch.randn(B, H, W, Cin + ld_extra, generator=g(seed))
    w = torch.randn(Cout, k * k, Cin, generator=g(seed + 1)) / math.sqrt(k * k * Cin)
    bias = 0.1 * torch.randn(Cout, generator=g(seed + 2))
    bias_b = 0.1 * torch.randn(B, Cout + 5, generator=g(seed + 3))
    res = torch.randn(B, H, W, Cout, generator=g(seed + 4))
    return buf, w, bias, bias_b, res


@pytest.mark.parametrize("B,H,W,Cin,Cout,k", [(2, 16, 16, 64, 64, 3), (1, 9, 13, 6, 64, 3), (2, 8, 8, 6, 128, 1),
                                              (1, 32, 32, 128, 6, 3), (3, 4, 4, 256, 128, 1)])
def test_conv_direct(B, H, W, Cin, Cout, k):
    buf, w, bias, bias_b, res = _conv_case(B, H, W, Cin, Cout, k, 20, ld_extra=8)
    x = buf[..., 8:]
    ref = torch.empty(B, H, W, Cout)
    E.conv_direct(x, w, bias, ref, k, bias_b=bias_b[:, 5:], residual=res, out_scale=0.7)
    out = torch.empty(B, H, W, Cout, device=DEV)
    ops.conv_direct(buf.to(DEV)[..., 8:], w.to(DEV), bias.to(DEV), out, k, bias_b=bias_b.to(DEV)[:, 5:],
                    residual=res.to(DEV), out_scale=0.7)
    assert rel(out, ref) < 1e-5
    # in-place residual + plain
    out2 = res.to(DEV).clone()
    ops.conv_direct(buf.to(DEV)[..., 8:], w.to(DEV), bias.to(DEV), out2, k, residual=out2)
    E.conv_direct(x, w, bias, ref, k, residual=res)
    assert rel(out2, ref) < 1e-5


@pytest.mark.parametrize("B,H,W,Cin,Cout,ld_extra", [(2, 16, 24, 6, 64, 0), (3, 5, 7, 6, 128, 64), (1, 8, 8, 8, 64, 0)])
def test_conv1x1_narrow_combine_pattern(B, H, W, Cin, Cout, ld_extra):
    """Combine "sum" (layerspp.py:57-62): dst = Conv1x1(fp32 input pyramid, 6 channels) + dst, bf16 dst inside a wider
    concat buffer -- the streaming fast path of ditsep_conv_direct."""
    x = torch.randn(B, H, W, Cin, generator=g(60))
    w = torch.randn(Cout, 1, Cin, generator=g(61)) / math.sqrt(Cin)
    bias = 0.1 * torch.randn(Cout, generator=g(62))
    buf = torch.randn(B, H, W, Cout + ld_extra, generator=g(63)).to(torch.bfloat16)
    dst = buf.to(DEV)[..., ld_extra:]
    ref = torch.empty(B, H, W, Cout)
    E.conv_direct(x, w, bias, ref, 1, residual=buf[..., ld_extra:].float())
    ops.conv_direct(x.to(DEV), w.to(DEV), bias.to(DEV), dst, 1, residual=dst)
    assert rel(dst.float(), ref) < 4e-3          # bf16 rounding of the stored sum
    f32 = torch.empty(B, H, W, Cout, device=DEV)
    ops.conv_direct(x.to(DEV), w.to(DEV), bias.to(DEV), f32, 1, out_scale=0.5)
    E.conv_direct(x, w, bias, ref, 1, out_scale=0.5)
    assert rel(f32, ref) < 1e-6


TC_CASES = [
    # B, H, W, Cin, ld_extra, Cout, k
    (2, 16, 16, 64, 0, 64, 3),
    (1, 64, 64, 128, 64, 128, 3),     # channel slice of a concat buffer
    (1, 256, 256, 64, 0, 64, 3),      # the dominant shape of the score net
    (2, 32, 32, 192, 0, 64, 3),
    (3, 4, 4, 256, 0, 128, 3),        # deepest level, batch not a multiple of the batch box
    (1, 8, 24, 128, 0, 128, 3),       # non power-of-two width (variable-length utterances)
    (2, 16, 16, 128, 0, 384, 1),      # fused q/k/v projection (three N tiles)
    (1, 32, 32, 64, 0, 6, 3),         # progressive-output conv (N padded to 16)
    (2, 16, 16, 128, 0, 64, 1),
    (1, 40, 20, 64, 0, 64, 3),        # halo tiles with ragged right / bottom edges
    (2, 32, 16, 256, 0, 128, 3),      # four 64-channel blocks, filter streamed
]


def _conv_sums(buf):
    """Fixed-point sums [B][C][4] int64 -> (B, C, 2) double {sum, sum of squares}."""
    return E._fx_decode(buf.cpu())


@pytest.mark.parametrize("B,H,W,Cin,ld_extra,Cout,k", TC_CASES)
def test_conv_tc(B, H, W, Cin, ld_extra, Cout, k):
    buf, w, bias, bias_b, res = _conv_case(B, H, W, Cin, Cout, k, 30, ld_extra=ld_extra)
    xb = buf.to(torch.bfloat16)
    wb = w.to(torch.bfloat16)
    x = xb[..., ld_extra:]
    ref = torch.empty(B, H, W, Cout)
    E.conv_direct(x.float(), wb.float(), bias, ref, k, bias_b=bias_b[:, 5:], residual=res, out_scale=0.7)
    cp = (Cout + 15) // 16 * 16
    wpad = torch.zeros(cp, k * k, Cin, dtype=torch.bfloat16)
    wpad[:Cout] = wb
    xd = xb.to(DEV)[..., ld_extra:]
    plan = ops.ConvTcPlan(xd, wpad.to(DEV), Cout, k)
    out = torch.empty(B, H, W, Cout, device=DEV)
    tile_kernel = H >= 16 and W >= 8
    assert plan.tile == tile_kernel
    stats_buf = ops.new_sums(B, Cout, DEV) if tile_kernel else None
    plan.run(bias.to(DEV), out, bias_b=bias_b.to(DEV)[:, 5:], residual=res.to(DEV), out_scale=0.7, stats=stats_buf)
    torch.cuda.synchronize()
    # same bf16 operands, fp32 accumulation on both sides: only summation order differs
    assert rel(out, ref) < 2e-5, rel(out, ref)
    if tile_kernel:
        # fused GroupNorm statistics of the stored output: per-image, per-channel sum and sum of squares
        sums = _conv_sums(stats_buf)
        o = out.double().cpu()
        assert rel(sums[..., 1], (o * o).sum(dim=(1, 2))) < 1e-5
        assert float(((sums[..., 0] - o.sum(dim=(1, 2))).abs() / o.abs().sum(dim=(1, 2))).max()) < 1e-5
        G = min(max(Cout // 4, 1), 32) if Cout % 4 == 0 else 1
        gamma, beta = 1 + 0.1 * torch.randn(Cout, generator=g(1)), 0.1 * torch.randn(Cout, generator=g(2))
        ss_ref = torch.empty(B, 2, Cout)
        E.gn_stats(out.cpu(), G, gamma, beta, 1e-6, ss_ref)
        ss = torch.empty(B, 2, Cout, device=DEV)
        ops.gn_finalize([stats_buf], B, H * W, G, gamma.to(DEV), beta.to(DEV), 1e-6, ss)
        assert rel(ss, ss_ref) < 1e-4, rel(ss, ss_ref)
        # bit-reproducible although accumulated with atomics (integer adds commute): a second run gives identical sums
        again = ops.new_sums(B, Cout, DEV)
        plan.run(bias.to(DEV), out, bias_b=bias_b.to(DEV)[:, 5:], residual=res.to(DEV), out_scale=0.7, stats=again)
        assert torch.equal(stats_buf, again)
        if B > 1:  # batch-invariant: image B-1 alone leaves the same sums as inside the batch
            plan1 = ops.ConvTcPlan(xd[B - 1:], wpad.to(DEV), Cout, k)
            st1 = ops.new_sums(1, Cout, DEV)
            out1 = torch.empty(1, H, W, Cout, device=DEV)
            plan1.run(bias.to(DEV), out1, bias_b=bias_b.to(DEV)[B - 1:, 5:], residual=res.to(DEV)[B - 1:], out_scale=0.7,
                      stats=st1)
            assert torch.equal(out1, out[B - 1:])
            assert torch.equal(st1[0], stats_buf[B - 1])
    else:
        with pytest.raises(ops.DitsepError):
            plan.run(bias.to(DEV), out, stats=torch.zeros(16, device=DEV))
    outb = torch.empty(B, H, W, Cout, device=DEV, dtype=torch.bfloat16)
    resb = res.to(torch.bfloat16)
    plan.run(bias.to(DEV), outb, residual=resb.to(DEV))
    E.conv_direct(x.float(), wb.float(), bias, ref, k, residual=resb.float())
    assert rel(outb.float(), ref) < 4e-3  # bf16 output rounding
    # in-place residual (Combine / res-block shortcut pattern)
    out2 = resb.to(DEV).clone()
    plan.run(bias.to(DEV), out2, residual=out2)
    assert rel(out2.float(), ref) < 4e-3


@pytest.mark.parametrize("B,H,W,Cin,Cout,Cx,k,ld_extra", [
    (2, 32, 24, 64, 64, 64, 3, 0),      # identity-shaped shortcut (Cx == Cout)
    (3, 48, 40, 64, 64, 128, 3, 64),    # up-path block: x = concat of 128 channels inside a wider buffer
    (1, 16, 16, 128, 128, 128, 3, 0),   # two N tiles, two main chunks, two shortcut chunks
    (2, 17, 21, 64, 48, 64, 1, 0),      # ragged image, 1x1 main conv, Cout not a multiple of 64
])
def test_conv_tc_fused_shortcut(B, H, W, Cin, Cout, Cx, k, ld_extra):
    """ditsep_conv_tc_plan_create_sc: out = (conv_k(h) + conv_1x1(x) + bias + bias_b) * scale in ONE launch, the
    shortcut accumulated in TMEM as an extra tap; statistics partials are those of the stored output."""
    h = torch.randn(B, H, W, Cin, generator=g(80)).to(torch.bfloat16)
    xbuf = torch.randn(B, H, W, Cx + ld_extra, generator=g(81)).to(torch.bfloat16)
    x = xbuf[..., ld_extra:]
    wm = (torch.randn(Cout, k * k, Cin, generator=g(82)) / math.sqrt(k * k * Cin)).to(torch.bfloat16)
    wsc = (torch.randn(Cout, 1, Cx, generator=g(83)) / math.sqrt(Cx)).to(torch.bfloat16)
    if Cx == Cout and k == 3 and B == 2:
        wsc = torch.eye(Cout).reshape(Cout, 1, Cx).to(torch.bfloat16)  # plain residual as an identity filter
    bias = 0.1 * torch.randn(Cout, generator=g(84))
    bias_b = 0.1 * torch.randn(B, Cout, generator=g(85))
    cp = (Cout + 15) // 16 * 16
    wpad = torch.zeros(cp, k * k * Cin + Cx, dtype=torch.bfloat16)
    wpad[:Cout, :k * k * Cin] = wm.reshape(Cout, -1)
    wpad[:Cout, k * k * Cin:] = wsc.reshape(Cout, -1)
    hd, xd = h.to(DEV), xbuf.to(DEV)[..., ld_extra:]
    plan = ops.ConvTcPlan(hd, wpad.to(DEV), Cout, k, shortcut=xd)
    stats_buf = ops.new_sums(B, Cout, DEV)
    out = torch.empty(B, H, W, Cout, device=DEV)
    plan.run(bias.to(DEV), out, bias_b=bias_b.to(DEV), out_scale=0.7, stats=stats_buf)
    sc = torch.empty(B, H, W, Cout)
    E.conv_direct(x.float(), wsc.float(), torch.zeros(Cout), sc, 1)
    ref = torch.empty(B, H, W, Cout)
    E.conv_direct(h.float(), wm.float(), bias, ref, k, bias_b=bias_b, residual=sc, out_scale=0.7)
    assert rel(out, ref) < 2e-5, rel(out, ref)
    sums = _conv_sums(stats_buf)
    o = out.double().cpu()
    assert rel(sums[..., 1], (o * o).sum(dim=(1, 2))) < 1e-5
    # bf16 output + second run bit-identical
    outb = torch.empty(B, H, W, Cout, device=DEV, dtype=torch.bfloat16)
    plan.run(bias.to(DEV), outb, bias_b=bias_b.to(DEV), out_scale=0.7)
    outb2 = torch.empty_like(outb)
    plan.run(bias.to(DEV), outb2, bias_b=bias_b.to(DEV), out_scale=0.7)
    assert torch.equal(outb, outb2) and rel(outb.float(), ref) < 4e-3


@pytest.mark.parametrize("B,H,W,Cx,ld_out", [
    (2, 64, 64, 0, 128),       # 5 row tiles of 15 (last: 4 rows -> N = 48); output = a slice of a concat buffer
    (1, 128, 72, 64, 192),     # identity-shaped shortcut chunk, last tile 8 rows (N = 80)
    (3, 100, 64, 128, 80),     # two shortcut chunks (all 512 TMEM columns), ragged height, three images, run-time stride
    (1, 256, 256, 0, 64),      # the dominant shape: 18 row tiles, the last with ONE valid row (N = 16)
])
def test_conv_ts_filter_in_tensor_memory(B, H, W, Cx, ld_out, monkeypatch):
    """conv_ts_kernel (64 -> 64 3x3, filter as the TMEM A operand, two taps stacked on M, pixels on N) against the fp32
    reference and against conv_tc_tile_kernel on the same operands; fused statistics, batch invariance, repeatability."""
    Cin = Cout = 64
    h = torch.randn(B, H, W, Cin, generator=g(90)).to(torch.bfloat16)
    wm = (torch.randn(Cout, 9, Cin, generator=g(91)) / math.sqrt(9 * Cin)).to(torch.bfloat16)
    bias = 0.1 * torch.randn(Cout, generator=g(92))
    bias_b = 0.1 * torch.randn(B, Cout, generator=g(93))
    wpad = torch.zeros(Cout, 9 * Cin + Cx, dtype=torch.bfloat16)
    wpad[:, :9 * Cin] = wm.reshape(Cout, -1)
    sc = torch.zeros(B, H, W, Cout)
    xd = None
    if Cx:
        x = torch.randn(B, H, W, Cx, generator=g(94)).to(torch.bfloat16)
        wsc = (torch.randn(Cout, 1, Cx, generator=g(95)) / math.sqrt(Cx)).to(torch.bfloat16)
        wpad[:, 9 * Cin:] = wsc.reshape(Cout, -1)
        E.conv_direct(x.float(), wsc.float(), torch.zeros(Cout), sc, 1)
        xd = x.to(DEV)
    ref = torch.empty(B, H, W, Cout)
    E.conv_direct(h.float(), wm.float(), bias, ref, 3, bias_b=bias_b, residual=sc, out_scale=0.7)
    hd, wd = h.to(DEV), wpad.to(DEV)

    def run(ts, hd_=hd, xd_=xd, bb=bias_b):
        monkeypatch.setenv("DITSEP_CONV_TS", "1" if ts else "0")   # read when the plan is created
        plan = ops.ConvTcPlan(hd_, wd, Cout, 3, shortcut=xd_)
        st = ops.new_sums(hd_.shape[0], Cout, DEV)
        buf = torch.full((hd_.shape[0], H, W, ld_out), float("nan"), device=DEV, dtype=torch.bfloat16)
        out = buf[..., ld_out - Cout:]                          # compile-time strides 64 / 128 / 192, or the run-time path
        plan.run(bias.to(DEV), out, bias_b=bb.to(DEV), out_scale=0.7, stats=st)
        torch.cuda.synchronize()
        assert ld_out == Cout or torch.isnan(buf[..., :ld_out - Cout].float()).all()   # nothing written beside the slice
        return out.contiguous(), st

    n0 = ops.launch_count()
    out, st = run(True)
    old, st_old = run(False)
    assert not torch.isnan(out.float()).any()
    e_ref, e_old = rel(out.float(), ref), rel(out.float(), old.float())
    print(f"conv_ts {B}x{H}x{W} Cx={Cx}: rel-L2 vs fp32 reference {e_ref:.2e} (tile kernel {rel(old.float(), ref):.2e}), "
          f"vs tile kernel {e_old:.2e}, bf16 outputs that differ: {float((out != old).float().mean()):.2e}")
    assert e_ref < 4e-3 and e_old < 1e-3             # bf16 output rounding; the kernels differ in summation order only
    assert float((out != old).float().mean()) < 0.02
    sums = _conv_sums(st)
    of = out.double().cpu()
    # statistics are those of the fp32 values BEFORE the bf16 rounding of the store (zero-mean noise of 2^-9 relative)
    assert rel(sums[..., 1], (of * of).sum(dim=(1, 2))) < 3e-4
    assert float(((sums[..., 0] - of.sum(dim=(1, 2))).abs() / of.abs().sum(dim=(1, 2))).max()) < 3e-4
    assert rel(sums, _conv_sums(st_old)) < 2e-5                      # and equal to the tile kernel's to fp32 rounding
    out2, st2 = run(True)
    assert torch.equal(out, out2) and torch.equal(st, st2)
    if B > 1:  # batch-invariant: the last image alone gives the same bits
        o1, s1 = run(True, hd[B - 1:], None if xd is None else xd[B - 1:], bias_b[B - 1:])
        assert torch.equal(o1, out[B - 1:]) and torch.equal(s1[0], st[B - 1])


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_gn_apply_srcs_equals_finalize_then_apply(dtype):
    """ditsep_gn_apply_job (scale / shift folded inside the apply kernel from two producers' sums) is bit-identical to
    gn_finalize followed by gn_apply."""
    B, H, W, C1, C2 = 3, 24, 16, 128, 64
    x = torch.randn(B, H, W, C1 + C2, generator=g(70)).to(DEV).to(dtype)
    s1, s2 = ops.new_sums(B, C1, DEV), ops.new_sums(B, C2, DEV)
    ops.gn_sums(x[..., :C1], s1)
    ops.gn_sums(x[..., C1:], s2)
    C = C1 + C2
    gamma, beta = (1 + 0.1 * torch.randn(C, generator=g(71))).to(DEV), (0.1 * torch.randn(C, generator=g(72))).to(DEV)
    ss = torch.empty(B, 2, C, device=DEV)
    ops.gn_finalize([s1, s2], B, H * W, 32, gamma, beta, 1e-6, ss)
    want = torch.empty_like(x)
    ops.gn_apply(x, ss, True, want)
    got, cache = torch.empty_like(x), {}
    for _ in range(2):  # second call reuses the cached job struct
        got.fill_(float("nan"))
        ops.gn_apply_srcs(x, [s1, s2], 32, gamma, beta, 1e-6, True, got, cache=cache)
    assert torch.equal(got, want)
    ref = torch.nn.functional.silu(torch.nn.functional.group_norm(x.float().permute(0, 3, 1, 2), 32, gamma, beta, 1e-6))
    assert rel(got.float().permute(0, 3, 1, 2), ref) < (2e-5 if dtype == torch.float32 else 5e-3)


def test_gn_finalize_concat_sources():
    """GroupNorm over torch.cat([h, skip]) (ncsnpp.py:413) from two producers' statistics: a conv
    epilogue's sums for h (128 ch) and an explicit statistics pass for skip (64 ch); 6-channel groups straddle the
    boundary."""
    B, H, W = 2, 32, 16
    buf, w, bias, _, _ = _conv_case(B, H, W, 64, 128, 3, 61)
    cat = torch.zeros(B, H, W, 192, device=DEV, dtype=torch.bfloat16)
    plan = ops.ConvTcPlan(buf.to(DEV).to(torch.bfloat16), w.to(DEV).to(torch.bfloat16), 128, 3)
    st = ops.new_sums(B, 128, DEV)
    plan.run(bias.to(DEV), cat[..., :128], stats=st)
    skip = torch.randn(B, H, W, 64, generator=g(62)).to(DEV).to(torch.bfloat16)
    cat[..., 128:] = skip
    part = ops.new_sums(B, 64, DEV)
    ops.gn_sums(cat[..., 128:], part)
    gamma, beta = 1 + 0.1 * torch.randn(192, generator=g(1)), 0.1 * torch.randn(192, generator=g(2))
    ss = torch.empty(B, 2, 192, device=DEV)
    ops.gn_finalize([st, part], B, H * W, 32, gamma.to(DEV), beta.to(DEV), 1e-6, ss)
    ss_ref = torch.empty(B, 2, 192)
    E.gn_stats(cat.float().cpu(), 32, gamma, beta, 1e-6, ss_ref)
    assert rel(ss, ss_ref) < 1e-4, rel(ss, ss_ref)


def test_conv_tc_shifted_descriptor_views():
    """The tile kernel reads the nine taps as row-shifted views of ONE 128B-swizzled TMA tile with the UMMA
    descriptor's base offset left at 0 (the swizzle is keyed on absolute shared-memory address bits).  A probe
    on the B200 measured rel-L2 5e-7 that way and 1.3 with base_offset = (start >> 7) & 7."""
    B, H, W, Cin, Cout, k = 1, 32, 32, 64, 64, 3
    buf, w, bias, bias_b, res = _conv_case(B, H, W, Cin, Cout, k, 77)
    xb, wb = buf.to(torch.bfloat16), w.to(torch.bfloat16)
    ref = torch.empty(B, H, W, Cout)
    E.conv_direct(xb.float(), wb.float(), bias, ref, k)
    plan = ops.ConvTcPlan(xb.to(DEV), wb.to(DEV), Cout, k)
    out = torch.empty(B, H, W, Cout, device=DEV)
    plan.run(bias.to(DEV), out)
    torch.cuda.synchronize()
    assert rel(out, ref) < 2e-5


# ------------------------------------------------------------------ attention / temb / output layer
@pytest.mark.parametrize("B,H,W,C", [(2, 16, 16, 128), (1, 4, 4, 128), (1, 16, 80, 128), (2, 4, 5, 64)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_attention(B, H, W, C, dtype):
    qkv = torch.randn(B, H, W, 3 * C, generator=g(40)).to(dtype)
    ref = torch.empty(B, H, W, C, dtype=dtype)
    E.attention(qkv[..., :C], qkv[..., C:2 * C], qkv[..., 2 * C:], ref)
    d = qkv.to(DEV)
    out = torch.empty(B, H, W, C, device=DEV, dtype=dtype)
    ops.attention(d[..., :C], d[..., C:2 * C], d[..., 2 * C:], out)
    assert rel(out.float(), ref.float()) < (2e-5 if dtype == torch.float32 else 6e-3)


@pytest.mark.parametrize("B,H,W,C", [(2, 16, 16, 128), (8, 16, 80, 128), (3, 16, 20, 256), (2, 12, 9, 128), (1, 16, 5, 256),
                                     (2, 8, 8, 128)])
def test_attention_tensor_core_kernel(B, H, W, C, monkeypatch):
    """attention_tc_kernel (tcgen05: Q K^T and P V as UMMA tiles, V as an MN-major operand, two-pass softmax) against the
    fp32 torch statement and against the CUDA-core kernel; q / k / v are slices of one fused projection buffer, S is
    256 (C2), 1280 (C3 at its maximum), 320 (C4, C = 256) and sequences that are not multiples of the key tile."""
    qkv = torch.randn(B, H, W, 3 * C, generator=g(41)).to(torch.bfloat16)
    qkv[..., :C] *= 3.0                                  # peaky softmax rows as well as flat ones
    ref = torch.empty(B, H, W, C)
    E.attention(qkv[..., :C].float(), qkv[..., C:2 * C].float(), qkv[..., 2 * C:].float(), ref)
    d = qkv.to(DEV)
    n0 = ops.launch_count()
    out = torch.full((B, H, W, C), float("nan"), device=DEV, dtype=torch.bfloat16)
    ops.attention(d[..., :C], d[..., C:2 * C], d[..., 2 * C:], out)
    monkeypatch.setenv("DITSEP_ATTN_TC", "0")
    old = torch.empty_like(out)
    ops.attention(d[..., :C], d[..., C:2 * C], d[..., 2 * C:], old)
    torch.cuda.synchronize()
    e, e_old = rel(out.float(), ref), rel(old.float(), ref)
    print(f"attention_tc B={B} S={H * W} C={C}: rel-L2 vs fp32 {e:.2e} (CUDA-core kernel {e_old:.2e})")
    assert not torch.isnan(out.float()).any()
    assert e < 6e-3 and e < 2.0 * e_old + 1e-3


def test_temb_and_output_layer():
    B, nf = 3, 64
    t = torch.tensor([1.0, 0.4, 0.03])
    gfp = 16 * torch.randn(nf, generator=g(50))
    w1, b1 = torch.randn(4 * nf, 2 * nf, generator=g(51)) / 12, 0.1 * torch.randn(4 * nf, generator=g(52))
    w2, b2 = torch.randn(4 * nf, 4 * nf, generator=g(53)) / 16, 0.1 * torch.randn(4 * nf, generator=g(54))
    ref = torch.empty(B, 4 * nf)
    E.temb_mlp(t, gfp, w1, b1, w2, b2, ref)
    out = torch.empty(B, 4 * nf, device=DEV)
    ops.temb_mlp(*(a.to(DEV) for a in (t, gfp, w1, b1, w2, b2)), out)
    # sin/cos of arguments up to ~16*2pi*|log t|*3 ~ 1e3 rad in fp32: 1e-4
    assert rel(out, ref) < 2e-4, rel(out, ref)
    dw, db = torch.randn(777, 4 * nf, generator=g(55)) / 16, torch.randn(777, generator=g(56))
    dref = torch.empty(B, 777)
    E.temb_dense(ref, dw, db, dref)
    dout = torch.empty(B, 777, device=DEV)
    ops.temb_dense(ref.to(DEV), dw.to(DEV), db.to(DEV), dout)
    assert rel(dout, dref) < 1e-5
    x = torch.randn(B, 8, 8, 6, generator=g(57))
    w, b = torch.randn(4, 6, generator=g(58)), torch.randn(4, generator=g(59))
    oref = torch.empty(B, 8, 8, 4)
    E.output_layer(x, t, w, b, oref, True)
    o = torch.empty(B, 8, 8, 4, device=DEV)
    ops.output_layer(x.to(DEV), t.to(DEV), w.to(DEV), b.to(DEV), o, True)
    assert rel(o, oref) < 1e-5


def test_cpu_tensors_are_rejected():
    with pytest.raises(ops.DitsepError):
        ops.gn_apply(torch.zeros(1, 2, 2, 8), torch.zeros(1, 2, 8), True, torch.zeros(1, 2, 2, 8))


def test_stft_mixture_cache_modes_and_hook():
    """ditsep_stft_compress_cached: mode 1 fills the mixture channel's cache and writes the same features as the plain
    call, mode 2 reads it back bit-for-bit while the sources change; the score model only uses it for a tensor a
    sampler announced (note_constant_mixture) and refills after an in-place change of that tensor."""
    import torch
    from common import build_model, score_inputs
    from ditsep_b200 import ops
    dev = "cuda"
    B, T, W = 3, 9000, 128
    g = torch.Generator().manual_seed(8)
    xt, xt2 = (torch.randn(B, 2, T, generator=g).to(dev) for _ in range(2))
    mix = torch.randn(B, 1, T, generator=g).to(dev)
    win = torch.hann_window(510).to(dev)
    for dt in (torch.float32, torch.bfloat16):
        plain, filled, used, plain2 = (torch.empty(B, 256, W, 6, device=dev, dtype=dt) for _ in range(4))
        cache = torch.full((B, W, 256, 2), float("nan"), device=dev)
        ops.stft_compress(xt, mix, win, plain, 0.5, 0.15, True)
        ops.stft_compress(xt, mix, win, filled, 0.5, 0.15, True, mix_cache=cache, mix_mode=1)
        assert torch.equal(plain, filled) and torch.isfinite(cache).all()
        ops.stft_compress(xt2, mix, win, plain2, 0.5, 0.15, True)
        ops.stft_compress(xt2, torch.zeros_like(mix), win, used, 0.5, 0.15, True, mix_cache=cache, mix_mode=2)  # mixture NOT read
        assert torch.equal(plain2.float(), used.float())
    model, _, _ = build_model("bf16", dev, N=2)
    xs, t, mixn = (a.to(dev) for a in score_inputs(2, 7810))
    sm = model.score_model
    with torch.no_grad():
        ref = model(xs, t, mixn)
        assert getattr(sm, "_mixc", None) is None          # nothing announced: no cache
        model.note_constant_mixture(mixn)
        a = model(xs, t, mixn)                              # fills
        assert sm._mixc is not None and sm._mixc["ref"] is mixn
        b = model(xs, t, mixn)                              # served from the cache
        assert torch.equal(a, ref) and torch.equal(b, ref)
        mixn.mul_(0.5)                                      # in-place change: version bump -> refill
        c = model(xs, t, mixn)
        model.note_constant_mixture(None)
        d = model(xs, t, mixn)
        assert torch.equal(c, d) and not torch.equal(c, ref) and sm._mixc is None
