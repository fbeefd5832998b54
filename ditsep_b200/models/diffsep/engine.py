"""Execution engine of the NCSN++ score network: NHWC activations, one plan per input shape.

``NcsnppEngine`` turns ``NCSNpp.forward`` (src/models/diffsep/ncsnpp.py:320-478) into a
fixed list of kernel launches through the C ABI (``ops``):

* weights are re-packed once per parameter version: conv filters to ``[Cout][tap][Cin]``
  (fp32 for the CUDA-core path, bf16 for the tcgen05 path), the three attention
  projections q/k/v to one ``[3C][C]`` 1x1 filter, all res-blocks' ``Dense_0`` to one
  ``[R][4nf]`` matrix evaluated by a single launch per step;
* activations are NHWC (pixel-major) so that channel vectors are contiguous for the
  implicit-GEMM K dimension and for the 8-channel vector packs of the HBM-bound kernels;
* the U-Net skip concatenations (``torch.cat([h, hs.pop()], dim=1)``, ncsnpp.py:413) cost
  nothing: every down-path tensor is produced directly inside the channel slice of the
  concat buffer that its up-path consumer reads (pixel stride ``ld`` > C);
* the res-block tail ``(shortcut + h) / sqrt(2)`` (layerspp.py:322-328), the temb bias add
  (:316-317) and the Combine / progressive-pyramid sums are conv epilogues.

A plan is a list of ``(fn, args, kwargs)`` built once per (B, H, W); executing it issues
only kernel launches (no allocation, no host sync), which makes it CUDA-graph capturable.
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import torch

from ... import ops


@dataclass
class ModuleSpec:
    """One entry of ``NCSNpp.all_modules`` (ncsnpp.py:107-309)."""

    kind: str  # gfp | linear | conv3x3 | resblock | attn | combine | gn
    in_ch: int = 0
    out_ch: int = 0
    up: bool = False
    down: bool = False


def layout_of(m) -> List[ModuleSpec]:
    """Order and shapes of ``all_modules`` for the supported configuration family."""
    nf, R = m.nf, m.num_resolutions
    channels = m.num_channels_in
    mods = [ModuleSpec("gfp", out_ch=nf), ModuleSpec("linear", 2 * nf, 4 * nf), ModuleSpec("linear", 4 * nf, 4 * nf),
            ModuleSpec("conv3x3", channels, nf)]
    hs_c = [nf]
    in_ch = nf
    for lvl in range(R):
        for _ in range(m.num_res_blocks):
            out_ch = nf * m.ch_mult[lvl]
            mods.append(ModuleSpec("resblock", in_ch, out_ch))
            in_ch = out_ch
            if m.all_resolutions[lvl] in m.attn_resolutions:
                mods.append(ModuleSpec("attn", in_ch, in_ch))
            hs_c.append(in_ch)
        if lvl != R - 1:
            mods.append(ModuleSpec("resblock", in_ch, in_ch, down=True))
            mods.append(ModuleSpec("combine", channels, in_ch))
            hs_c.append(in_ch)
    in_ch = hs_c[-1]
    mods += [ModuleSpec("resblock", in_ch, in_ch), ModuleSpec("attn", in_ch, in_ch), ModuleSpec("resblock", in_ch, in_ch)]
    for lvl in reversed(range(R)):
        for _ in range(m.num_res_blocks + 1):
            out_ch = nf * m.ch_mult[lvl]
            mods.append(ModuleSpec("resblock", in_ch + hs_c.pop(), out_ch))
            in_ch = out_ch
        if m.all_resolutions[lvl] in m.attn_resolutions:
            mods.append(ModuleSpec("attn", in_ch, in_ch))
        mods.append(ModuleSpec("gn", in_ch, in_ch))
        mods.append(ModuleSpec("conv3x3", in_ch, channels))
        if lvl != 0:
            mods.append(ModuleSpec("resblock", in_ch, in_ch, up=True))
    assert not hs_c
    return mods


class _Pool:
    """Shape-keyed pool of temporaries used while a plan is being built."""

    def __init__(self, device):
        self.device = device
        self.free: Dict[Tuple, List[torch.Tensor]] = {}
        self.total_bytes = 0

    def get(self, shape, dtype) -> torch.Tensor:
        key = (tuple(shape), dtype)
        lst = self.free.get(key)
        if lst:
            return lst.pop()
        t = torch.empty(shape, device=self.device, dtype=dtype)
        self.total_bytes += t.numel() * t.element_size()
        return t

    def put(self, t: torch.Tensor):
        self.free.setdefault((tuple(t.shape), t.dtype), []).append(t)


class _SumsArena:
    """Fixed-point GroupNorm statistics buffers [B][C][4] int64 of one plan, carved from a few large chunks so that
    ONE memset per chunk at the head of the launch list re-zeroes all of them (producers ADD with integer atomics)."""

    CHUNK = 1 << 20  # int64 elements (8 MiB)

    def __init__(self, device):
        self.device, self.chunks, self.used = device, [], 0

    def new(self, B: int, C: int) -> torch.Tensor:
        n = B * C * 4
        if not self.chunks or self.used + n > self.chunks[-1].numel():
            self.chunks.append(torch.zeros(max(self.CHUNK, n), device=self.device, dtype=torch.int64))
            self.used = 0
        t = self.chunks[-1][self.used:self.used + n].view(B, C, 4)
        self.used += n
        return t

    def zero(self):
        for c in self.chunks:
            c.zero_()


@dataclass
class _GnSrcs:
    """A GroupNorm whose scale / shift have not been materialised: the producers' fixed-point sums (1..4 sources laid
    end to end along the channel axis) + the layer's parameters.  gn_apply folds them itself (ops.gn_apply_srcs)."""
    srcs: list
    G: int
    gamma: torch.Tensor
    beta: torch.Tensor


class _ConvW:
    """Packed filter of one convolution."""

    def __init__(self, w_f32, bias, ksize, w_bf16=None):
        self.w, self.bias, self.ksize, self.w_bf16 = w_f32, bias, ksize, w_bf16
        self.cout, self.cin = w_f32.shape[0], w_f32.shape[2]
        # ALGORITHMIC multiply-accumulates per output pixel (2*M*N*K convention of SURVEY.md 8d, K = taps * real Cin):
        # what nn.Conv2d does in the reference, not what the packed operand executes (the hi/lo-widened first layer
        # executes K = 576 for 54 real; a fused identity shortcut executes 64 extra columns for a plain add)
        self.algo_macs = self.cout * w_f32.shape[1] * self.cin


def _pack_conv(weight: torch.Tensor, bias: torch.Tensor, want_bf16: bool) -> _ConvW:
    """(Cout, Cin, kh, kw) -> [Cout][tap][Cin] fp32 (+ bf16 copy with Cout padded to 16)."""
    cout, cin, kh, kw = weight.shape
    w = weight.detach().float().permute(0, 2, 3, 1).reshape(cout, kh * kw, cin).contiguous()
    wb = None
    if want_bf16 and cin % 64 == 0:
        cp = (cout + 15) // 16 * 16
        wb = torch.zeros(cp, kh * kw, cin, device=w.device, dtype=torch.bfloat16)
        wb[:cout] = w.to(torch.bfloat16)
    return _ConvW(w, bias.detach().float().contiguous(), kh, wb)


def _fuse_shortcut(main: _ConvW, sc: Optional[_ConvW], cx: int) -> Optional[_ConvW]:
    """Filter of a res-block's second conv with its skip path riding behind the taps along K
    (ditsep_conv_tc_plan_create_sc): [CoutPad][taps*Cin + Cx] bf16, the last Cx columns being Conv_2's 1x1 filter
    (layerspp.py:322-324) or the identity (plain `x + h`, :325-327); bias = b_conv1 (+ b_conv2)."""
    if main.w_bf16 is None or cx % 64 or (sc is None and cx != main.cout):
        return None
    cp, taps, cin = main.w_bf16.shape
    wb = torch.zeros(cp, taps * cin + cx, device=main.w_bf16.device, dtype=torch.bfloat16)
    wb[:, :taps * cin] = main.w_bf16.reshape(cp, taps * cin)
    if sc is None:
        wb[:main.cout, taps * cin:] = torch.eye(main.cout, device=wb.device, dtype=torch.bfloat16)
        bias = main.bias
    else:
        wb[:main.cout, taps * cin:] = sc.w.reshape(main.cout, cx).to(torch.bfloat16)
        bias = (main.bias + sc.bias).contiguous()
    out = _ConvW(main.w, bias, main.ksize, wb)
    out.cx = cx
    out.algo_macs = main.algo_macs + (0 if sc is None else main.cout * cx)  # Conv_2 is a real 1x1 conv; x + h is not
    return out


def _pack_conv_hilo(weight: torch.Tensor, bias: torch.Tensor, width: int = 64) -> Optional[_ConvW]:
    """First-layer filter for the tensor-core path (ops.split_hilo): input channels [wh | wh | wl | 0] with
    w = wh + wl in bf16, matching activations [xh | xl | xh | 0]; (Cout, Cin, kh, kw) -> [CoutPad][tap][width]."""
    cout, cin, kh, kw = weight.shape
    if 3 * cin > width:
        return None
    w = weight.detach().float().permute(0, 2, 3, 1).reshape(cout, kh * kw, cin)
    wh = w.to(torch.bfloat16)
    wl = (w - wh.float()).to(torch.bfloat16)
    cp = (cout + 15) // 16 * 16
    wb = torch.zeros(cp, kh * kw, width, device=w.device, dtype=torch.bfloat16)
    wb[:cout, :, 0:cin] = wh
    wb[:cout, :, cin:2 * cin] = wh
    wb[:cout, :, 2 * cin:3 * cin] = wl
    out = _ConvW(wb[:cout].float(), bias.detach().float().contiguous(), kh, wb)
    out.algo_macs = cout * kh * kw * cin
    return out


class NcsnppEngine:
    def __init__(self, model):
        self.m = model
        self._wkey = None
        self._plans: Dict[Tuple, "_Plan"] = {}
        self.w: Dict[str, object] = {}

    # ------------------------------------------------------------------ weights
    def _weights_key(self):
        ps = list(self.m.parameters())
        return (self.m.precision, tuple((p.data_ptr(), p._version) for p in ps))

    def _pack(self):
        key = self._weights_key()
        if key == self._wkey:
            return
        m = self.m
        bf = m.precision == "bf16"
        w: Dict[str, object] = {}
        am = m.all_modules
        dense_w, dense_b, off = [], [], 0
        for i, spec in enumerate(m.layout):
            mod = am[i]
            if spec.kind == "gfp":
                w["gfp"] = mod.W.detach().float().contiguous()
            elif spec.kind == "linear":
                w[f"lin{i}"] = (mod.weight.detach().float().contiguous(), mod.bias.detach().float().contiguous())
            elif spec.kind == "conv3x3":
                w[f"conv{i}"] = _pack_conv(mod.weight, mod.bias, bf)
                if bf and i == 3:  # the input conv: tensor-core variant over the hi/lo-split features
                    w["conv3_hilo"] = _pack_conv_hilo(mod.weight, mod.bias)
            elif spec.kind == "gn":
                w[f"gn{i}"] = (mod.weight.detach().float().contiguous(), mod.bias.detach().float().contiguous())
            elif spec.kind == "combine":
                w[f"conv{i}"] = _pack_conv(mod.Conv_0.weight, mod.Conv_0.bias, bf)
            elif spec.kind == "attn":
                w[f"gn{i}"] = (mod.GroupNorm_0.weight.detach().float().contiguous(),
                               mod.GroupNorm_0.bias.detach().float().contiguous())
                # NIN: y = x W + b with W (in, out) -> 1x1 filter [out][in]
                wq = torch.cat([mod.NIN_0.W.detach().t(), mod.NIN_1.W.detach().t(), mod.NIN_2.W.detach().t()], dim=0)
                bq = torch.cat([mod.NIN_0.b.detach(), mod.NIN_1.b.detach(), mod.NIN_2.b.detach()], dim=0)
                w[f"qkv{i}"] = _pack_conv(wq[:, :, None, None], bq, bf)
                w[f"proj{i}"] = _pack_conv(mod.NIN_3.W.detach().t()[:, :, None, None], mod.NIN_3.b, bf)
            elif spec.kind == "resblock":
                w[f"gn{i}a"] = (mod.GroupNorm_0.weight.detach().float().contiguous(),
                                mod.GroupNorm_0.bias.detach().float().contiguous())
                w[f"gn{i}b"] = (mod.GroupNorm_1.weight.detach().float().contiguous(),
                                mod.GroupNorm_1.bias.detach().float().contiguous())
                w[f"conv{i}a"] = _pack_conv(mod.Conv_0.weight, mod.Conv_0.bias, bf)
                w[f"conv{i}b"] = _pack_conv(mod.Conv_1.weight, mod.Conv_1.bias, bf)
                if hasattr(mod, "Conv_2"):
                    w[f"conv{i}c"] = _pack_conv(mod.Conv_2.weight, mod.Conv_2.bias, bf)
                if bf:  # second conv with the skip path fused in (used where the kernel can keep the filter resident)
                    w[f"conv{i}b_sc"] = _fuse_shortcut(w[f"conv{i}b"], w.get(f"conv{i}c"), spec.in_ch)
                dense_w.append(mod.Dense_0.weight.detach().float())
                dense_b.append(mod.Dense_0.bias.detach().float())
                w[f"dense_off{i}"] = off
                off += spec.out_ch
        w["dense_w"] = torch.cat(dense_w, dim=0).contiguous()
        w["dense_b"] = torch.cat(dense_b, dim=0).contiguous()
        w["dense_rows"] = off
        ol = m.output_layer
        w["out_w"] = ol.weight.detach().float().reshape(ol.weight.shape[0], ol.weight.shape[1]).contiguous()
        w["out_b"] = ol.bias.detach().float().contiguous()
        self.w = w
        self._wkey = key
        self._plans.clear()  # plans hold pointers to the packed tensors

    # ------------------------------------------------------------------ entry points
    def plan(self, B: int, H: int, W: int, device) -> "_Plan":
        self._pack()
        key = (B, H, W, str(device))
        p = self._plans.get(key)
        if p is None:
            p = _Plan(self, B, H, W, device)
            self._plans[key] = p
        return p

    def forward_nhwc(self, feat: torch.Tensor, time_cond: torch.Tensor, affine_done: bool = True) -> torch.Tensor:
        """feat: (B, H, W, Cin) fp32 NHWC, with NCSN++'s ``2x-1`` (ncsnpp.py:348-350) already
        applied unless the model is ``centered``.  Returns the plan's output buffer
        (B, H, W, Cout) fp32 -- valid until the next call."""
        ops.require_device(feat, "NCSNpp")
        if not affine_done and not self.m.centered:
            feat = 2 * feat - 1.0
        B, H, W, C = feat.shape
        p = self.plan(B, H, W, feat.device)
        if feat.data_ptr() != p.feat.data_ptr():
            p.feat.copy_(feat)
        p.t.copy_(time_cond)
        p.run()
        return p.out


class _Plan:
    """Buffers + launch list for one (B, H, W)."""

    def __init__(self, eng: NcsnppEngine, B: int, H: int, W: int, device):
        m, w = eng.m, eng.w
        self.eng, self.B, self.H, self.W, self.device = eng, B, H, W, device
        self.steps: List[Tuple] = []
        adt = torch.bfloat16 if m.precision == "bf16" else torch.float32
        self.adt = adt
        f32 = torch.float32
        nf, R = m.nf, m.num_resolutions
        if H % (1 << (R - 1)) or W % (1 << (R - 1)):
            raise ops.DitsepError(f"NCSNpp input {H}x{W} must be divisible by {1 << (R - 1)}")
        pool = _Pool(device)
        self.pool = pool
        new = lambda *shape, dtype=adt: torch.empty(shape, device=device, dtype=dtype)  # noqa: E731
        Cin0 = m.num_channels_in
        self.feat = new(B, H, W, Cin0, dtype=f32)
        self.t = new(B, dtype=f32)
        self.out = new(B, H, W, m.num_channels_out, dtype=f32)
        self.act_temb = new(B, 4 * nf, dtype=f32)
        self.dense = new(B, w["dense_rows"], dtype=f32)
        self.sums = _SumsArena(device)
        self._tc_plans = []
        self._graph, self._graph_launches, self._runs = None, 0, 0
        # (data_ptr, channels) of a produced NHWC view -> (statistics buffer, layout): where its producer
        # left the per-channel sums GroupNorm needs (conv epilogue partials, or an explicit pass)
        self.stat_src: Dict[Tuple[int, int], Tuple[torch.Tensor, object]] = {}
        add = self.steps.append

        add((self.sums.zero, (), {}))  # statistics are accumulated with atomics: start every evaluation from zero
        # ---- time embedding (ncsnpp.py:325-346) + every res-block's Dense_0 (layerspp.py:316-317)
        add((ops.temb_mlp, (self.t, w["gfp"], *w["lin1"], *w["lin2"], self.act_temb), {}))
        add((ops.temb_dense, (self.act_temb, w["dense_w"], w["dense_b"], self.dense), {}))

        # ---- where does every U-Net tensor live?  (static replay of the hs push / pop order)
        layout = m.layout
        res = [(H >> l, W >> l) for l in range(R)]
        # down-path pushes: (level, channels)
        pushes = [(0, nf)]
        i = 4
        in_ch = nf
        for lvl in range(R):
            for _ in range(m.num_res_blocks):
                in_ch = layout[i].out_ch
                i += 1
                if layout[i].kind == "attn":
                    i += 1
                pushes.append((lvl, in_ch))
            if lvl != R - 1:
                i += 2
                pushes.append((lvl + 1, in_ch))
        # up-path consumers pop in reverse order; consumer j reads cat[h_j, push[n-1-j]]
        n_up = len(pushes)
        i_mid = i
        up_specs = [s for s in layout[i_mid + 3:] if s.kind == "resblock" and not s.up]
        assert len(up_specs) == n_up
        self.concat: List[torch.Tensor] = []
        for j, spec in enumerate(up_specs):
            lvl, cs = pushes[n_up - 1 - j]
            hh, ww = res[lvl]
            self.concat.append(new(B, hh, ww, spec.in_ch))
        def skip_slot(k):  # destination of push k  # noqa: E306
            j = n_up - 1 - k
            cs = pushes[k][1]
            return self.concat[j][..., self.concat[j].shape[3] - cs:]
        def h_slot(j):  # destination of the main-path tensor feeding up-block j  # noqa: E306
            cs = pushes[n_up - 1 - j][1]
            return self.concat[j][..., :self.concat[j].shape[3] - cs]

        # ---- down path (ncsnpp.py:352-397)
        pyr_in = self.feat
        k = 0
        cur = skip_slot(k)
        hilo = w.get("conv3_hilo")
        if hilo is not None and adt == torch.bfloat16 and ops.tc_enabled():
            # first layer on the tensor cores: 6 fp32 channels widened to [hi | lo | hi | 0] bf16 (ops.split_hilo);
            # channels >= 3*Cin0 are zeroed once here and never written again
            self.feat_tc = torch.zeros(B, H, W, hilo.cin, device=device, dtype=torch.bfloat16)
            add((ops.split_hilo, (self.feat, self.feat_tc), {}))
            self._conv(self.feat_tc, hilo, cur, want_stats=True)
        else:
            self._conv(self.feat, w["conv3"], cur, want_stats=True)
        k += 1
        i = 4
        for lvl in range(R):
            for _ in range(m.num_res_blocks):
                spec = layout[i]
                if layout[i + 1].kind == "attn":
                    hh, ww = res[lvl]
                    tmp = pool.get((B, hh, ww, spec.out_ch), adt)
                    self._resblock(i, spec, cur, tmp)
                    dst = skip_slot(k)
                    self._attn(i + 1, layout[i + 1], tmp, dst)
                    pool.put(tmp)
                    i += 2
                else:
                    dst = skip_slot(k)
                    self._resblock(i, spec, cur, dst)
                    i += 1
                cur = dst
                k += 1
            if lvl != R - 1:
                spec = layout[i]
                dst = skip_slot(k)
                self._resblock(i, spec, cur, dst)
                hh, ww = res[lvl + 1]
                nxt = new(B, hh, ww, Cin0, dtype=f32)  # FIR-downsampled input pyramid (ncsnpp.py:387)
                add((ops.fir_resample, (pyr_in, False), {"out_x": nxt}))
                pyr_in = nxt
                # Combine "sum" (layerspp.py:57-62): dst = Conv_0(pyramid) + dst
                self._conv(pyr_in, w[f"conv{i + 1}"], dst, residual=dst, want_stats=True)
                i += 2
                cur = dst
                k += 1
        assert k == n_up and i == i_mid

        # ---- middle (ncsnpp.py:399-405)
        hh, ww = res[R - 1]
        c = layout[i].out_ch
        t1, t2 = pool.get((B, hh, ww, c), adt), pool.get((B, hh, ww, c), adt)
        self._resblock(i, layout[i], cur, t1)
        self._attn(i + 1, layout[i + 1], t1, t2)
        self._resblock(i + 2, layout[i + 2], t2, h_slot(0))
        pool.put(t1)
        pool.put(t2)
        i += 3

        # ---- up path (ncsnpp.py:410-460)
        pyramid = None
        j = 0
        for lvl in reversed(range(R)):
            hh, ww = res[lvl]
            for b in range(m.num_res_blocks + 1):
                spec = layout[i]
                last_of_level = b == m.num_res_blocks
                if last_of_level:
                    dst = pool.get((B, hh, ww, spec.out_ch), adt)
                else:
                    dst = h_slot(j + 1)
                self._resblock(i, spec, self.concat[j], dst)
                cur = dst
                i += 1
                j += 1
            if layout[i].kind == "attn":
                tmp = pool.get((B, hh, ww, layout[i].in_ch), adt)
                self._attn(i, layout[i], cur, tmp)
                pool.put(cur)
                cur = tmp
                i += 1
            # progressive output skip: GN -> SiLU -> conv3x3(C -> Cin0), summed over levels
            cch = layout[i].in_ch
            if self._small(cur):
                hact = pool.get(tuple(cur.shape), adt)
                self._gn_act(cur, w[f"gn{i}"], True, hact)
                src, gnk = hact, {}
            else:
                hact = None
                src, gnk = cur, {"in_ss": self._scale_shift(cur, w[f"gn{i}"])}
            i += 1
            newpyr = new(B, hh, ww, Cin0, dtype=f32)
            if pyramid is None:
                self._conv(src, w[f"conv{i}"], newpyr, **gnk)
            else:
                add((ops.fir_resample, (pyramid, True), {"out_x": newpyr}))
                self._conv(src, w[f"conv{i}"], newpyr, residual=newpyr, **gnk)
            if hact is not None:
                pool.put(hact)
            pyramid = newpyr
            i += 1
            if lvl != 0:
                spec = layout[i]
                self._resblock(i, spec, cur, h_slot(j))
                pool.put(cur)
                i += 1
            else:
                pool.put(cur)
        assert i == len(layout) and j == n_up
        add((ops.output_layer, (pyramid, self.t, w["out_w"], w["out_b"], self.out, bool(m.scale_by_sigma)), {}))

    # ------------------------------------------------------------------ building blocks
    def _conv(self, x, cw: _ConvW, out, bias_b=None, residual=None, out_scale: float = 1.0, want_stats: bool = False,
              shortcut=None, in_ss=None, in_silu: bool = True, gn_inplace: bool = False) -> bool:
        """3x3 / 1x1 conv with the fused epilogue; tcgen05 path when the shape allows it.  want_stats:
        GroupNorm will normalise `out` (alone or as part of a concat buffer) -- its per-channel sums come
        from the conv epilogue when the kernel can leave them, else from one explicit pass over `out`."""
        use_tc = (cw.w_bf16 is not None and x.dtype == torch.bfloat16 and ops.tc_enabled())
        kw = {"bias_b": bias_b, "residual": residual, "out_scale": out_scale}
        plan = None
        if shortcut is not None:
            # fused skip path (shortcut: the block input x, added through the 1x1 filter packed behind cw's taps):
            # returns False -- nothing appended -- when this shape cannot be fused; the caller keeps the unfused steps
            if not use_tc or shortcut.dtype != torch.bfloat16 or os.environ.get("DITSEP_FUSE_SHORTCUT", "1") == "0":
                return False
            try:
                plan = ops.ConvTcPlan(x, cw.w_bf16, cw.cout, cw.ksize, shortcut=shortcut)
            except ops.DitsepError:
                return False
        if in_ss is not None:
            # conv(act(GroupNorm(x))) (layerspp.py:297-298,318-320): one normalise + SiLU pass, then the convolution.
            # (Applying it to the landed tiles inside the conv kernel was measured slower in round 1 and removed.)
            act = x if gn_inplace else self.pool.get(tuple(x.shape), x.dtype)
            if isinstance(in_ss, _GnSrcs):  # scale / shift formed inside the apply kernel from the producers' sums
                self.steps.append((ops.gn_apply_srcs, (x, in_ss.srcs, in_ss.G, in_ss.gamma, in_ss.beta, 1e-6, in_silu, act),
                                   {"cache": {}}))
            else:
                self.steps.append((ops.gn_apply, (x, in_ss, in_silu, act), {}))
            if act is not x:
                self.pool.put(act)  # free again once this conv has run (the launch list is sequential)
            x = act
            plan = None if shortcut is None else ops.ConvTcPlan(x, cw.w_bf16, cw.cout, cw.ksize, shortcut=shortcut)
        key = (out.data_ptr(), out.shape[3])
        self.stat_src.pop(key, None)  # whatever was known about this view is stale now
        fused = False
        if use_tc:
            if plan is None:
                plan = ops.ConvTcPlan(x, cw.w_bf16, cw.cout, cw.ksize)
            self._tc_plans.append(plan)
            plan.algo_flops = 2.0 * x.shape[0] * x.shape[1] * x.shape[2] * cw.algo_macs
            if want_stats and plan.tile:
                buf = self.sums.new(x.shape[0], cw.cout)
                kw["stats"] = buf
                self.stat_src[key] = buf
                fused = True
            self.steps.append((plan.run, (cw.bias, out), kw))
        else:
            self.steps.append((ops.conv_direct, (x, cw.w, cw.bias, out, cw.ksize), kw))
        if want_stats and not fused:
            # LAZY: the explicit statistics pass is only appended when a consumer asks for scale / shift
            # (_scale_shift); small-image consumers use the one-launch gn_small instead and never do
            self.stat_src[key] = ("lazy", out)
        return True

    def _scale_shift(self, x, gnw, want_tensor: bool = False):
        """GroupNorm(min(C/4,32), C, eps=1e-6) of the NHWC view x folded to per-(b,c) scale/shift
        (layerspp.py:269-271,297,318).  Uses the producers' recorded sums when every channel range of
        x has one (a concat buffer = [main path | skip]); otherwise one statistics pass over x.
        want_tensor: the consumer reads a (B, 2, C) buffer (the FIR kernels): one gn_finalize launch writes it.
        Otherwise the sums themselves are returned (_GnSrcs) and the consumer -- gn_apply -- folds them itself."""
        B, H, W, C = x.shape
        G = min(C // 4, 32)
        srcs, ptr, left = [], x.data_ptr(), C
        while left > 0:
            hit = None
            for (p0, c0), v in self.stat_src.items():
                if p0 == ptr and c0 <= left and (hit is None or c0 > hit[0]):
                    hit = (c0, v)
            if hit is None:
                srcs = None
                break
            if isinstance(hit[1], tuple):  # "lazy": the producer left no statistics: one explicit pass over its output, now
                t = hit[1][1]
                buf = self.sums.new(t.shape[0], t.shape[3])
                self.steps.append((ops.gn_sums, (t, buf), {}))
                self.stat_src[(t.data_ptr(), t.shape[3])] = buf
                hit = (hit[0], buf)
            srcs.append(hit[1])
            ptr += hit[0] * x.element_size()
            left -= hit[0]
        if not srcs or len(srcs) > 4:  # unknown producers: one statistics pass over the whole view
            buf = self.sums.new(B, C)
            self.steps.append((ops.gn_sums, (x, buf), {}))
            srcs = [buf]
        # Folding the finalize arithmetic into gn_apply (ops.gn_apply_srcs) was MEASURED SLOWER (r2zd: 48.6 vs 50.3
        # mixtures/s; every CTA of the apply kernel pays the ~5 us prologue, and a 2-wave grid pays it twice, where the
        # separate launch hides under its predecessor's tail): opt-in only.
        if not want_tensor and C <= 512 and os.environ.get("DITSEP_GN_FOLD", "0") == "1":
            return _GnSrcs(srcs, G, gnw[0], gnw[1])
        ss = torch.empty(B, 2, C, device=self.device, dtype=torch.float32)
        self.steps.append((ops.gn_finalize, (srcs, B, H * W, G, gnw[0], gnw[1], 1e-6, ss), {}))
        return ss

    @staticmethod
    def _small(x) -> bool:
        """Images of at most 16 x 16 pixels: statistics + normalise + activation in ONE launch (ops.gn_small) instead
        of partial sums -> finalize -> apply (or the conv's fused operand path)."""
        B, H, W, C = x.shape
        return H * W <= 256 and C <= 256 and C % 8 == 0 and 256 % (C // 8) == 0 and \
            os.environ.get("DITSEP_GN_SMALL", "1") != "0"

    def _gn_act(self, x, gnw, silu: bool, out):
        """out = act?(GroupNorm(min(C/4, 32), C, eps=1e-6)(x)) for a small image (out may be x)."""
        C = x.shape[3]
        self.steps.append((ops.gn_small, (x, min(C // 4, 32), gnw[0], gnw[1], 1e-6, silu, out), {}))

    def _resblock(self, i: int, spec: ModuleSpec, x, out):
        """ResnetBlockBigGANpp.forward (layerspp.py:296-328)."""
        w, pool, add = self.eng.w, self.pool, self.steps.append
        B, H, W, cin = x.shape
        cout = spec.out_ch
        small0 = self._small(x) and not (spec.up or spec.down)
        ss0 = None if small0 else self._scale_shift(x, w[f"gn{i}a"], want_tensor=bool(spec.up or spec.down))
        if spec.up or spec.down:
            Ho, Wo = (2 * H, 2 * W) if spec.up else (H // 2, W // 2)
            xr = pool.get((B, Ho, Wo, cin), self.adt)
            hA = pool.get((B, Ho, Wo, cin), self.adt)
            add((ops.fir_resample, (x, spec.up), {"out_x": xr, "scale_shift": ss0, "out_h": hA}))
        else:
            Ho, Wo = H, W
            xr = x
            hA = None
        h1 = pool.get((B, Ho, Wo, cout), self.adt)
        off = w[f"dense_off{i}"]
        if small0:
            hA = pool.get((B, H, W, cin), self.adt)
            self._gn_act(x, w[f"gn{i}a"], True, hA)
        if hA is None:  # act(GroupNorm_0(x)) rides in conv0's operand path
            self._conv(x, w[f"conv{i}a"], h1, bias_b=self.dense[:, off:off + cout], want_stats=True, in_ss=ss0)
        else:
            self._conv(hA, w[f"conv{i}a"], h1, bias_b=self.dense[:, off:off + cout], want_stats=True)
            pool.put(hA)
        if self._small(h1):
            self._gn_act(h1, w[f"gn{i}b"], True, h1)  # in place, one launch
            gn1 = {}
        else:
            ss1 = self._scale_shift(h1, w[f"gn{i}b"])
            # act(GroupNorm_1(h)) rides in conv1's operand path (or, unfused, is applied to h1 in place)
            gn1 = {"in_ss": ss1, "in_silu": True, "gn_inplace": True}
        fused = w.get(f"conv{i}b_sc")
        if not (fused is not None and self._conv(h1, fused, out, shortcut=xr, out_scale=1.0 / math.sqrt(2.0),
                                                 want_stats=True, **gn1)):
            if f"conv{i}c" in w:
                self._conv(xr, w[f"conv{i}c"], out)
                res = out
            else:
                res = xr
            self._conv(h1, w[f"conv{i}b"], out, residual=res, out_scale=1.0 / math.sqrt(2.0), want_stats=True, **gn1)
        self.stat_src.pop((h1.data_ptr(), cout), None)
        pool.put(h1)
        if xr is not x:
            pool.put(xr)

    def _attn(self, i: int, spec: ModuleSpec, x, out):
        """AttnBlockpp.forward (layerspp.py:81-97)."""
        w, pool, add = self.eng.w, self.pool, self.steps.append
        B, H, W, C = x.shape
        hn = pool.get((B, H, W, C), self.adt)
        qkv = pool.get((B, H, W, 3 * C), self.adt)
        if self._small(x):
            self._gn_act(x, w[f"gn{i}"], False, hn)
            self._conv(hn, w[f"qkv{i}"], qkv)
        else:
            ss = self._scale_shift(x, w[f"gn{i}"])
            self._conv(x, w[f"qkv{i}"], qkv, in_ss=ss, in_silu=False)  # GroupNorm_0 (no activation) in the operand path
        add((ops.attention, (qkv[..., :C], qkv[..., C:2 * C], qkv[..., 2 * C:], hn), {}))
        self._conv(hn, w[f"proj{i}"], out, residual=x, out_scale=1.0 / math.sqrt(2.0), want_stats=True)
        pool.put(hn)
        pool.put(qkv)

    def run_eager(self):
        for fn, args, kwargs in self.steps:
            fn(*args, **kwargs)

    def run(self):
        """One score-network evaluation.  The launch list touches only the plan's own buffers, so after one
        eager pass it is captured into a CUDA graph and replayed: ~430 launches per evaluation become one
        graph launch (at batch 1 the host launch path, not the GPU, bounds an eager step).
        DITSEP_GRAPH=0 keeps it eager (debugging, per-launch timing)."""
        if not ops.graphs_enabled():
            return self.run_eager()
        if self._graph is None:
            if self._runs == 0:  # first call: plain launches (loads modules, surfaces launch errors eagerly)
                self._runs = 1
                return self.run_eager()
            n0 = ops.launch_count()
            g = torch.cuda.CUDAGraph()
            cur = torch.cuda.current_stream()
            side = torch.cuda.Stream(device=self.device)
            side.wait_stream(cur)
            with torch.cuda.stream(side):
                with torch.cuda.graph(g, stream=side, capture_error_mode="thread_local"):
                    self.run_eager()
            cur.wait_stream(side)
            self._graph = g
            self._graph_launches = ops.launch_count() - n0
            ops.note_launches(-self._graph_launches)  # capture issued no work; replays are counted below
        self._graph.replay()
        ops.note_launches(self._graph_launches)
