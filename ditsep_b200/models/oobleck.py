"""Oobleck VAE around the latent sampler: LatentDiffSep.encode / decode
(src/diffsep_latent.py:107-128) over OobleckEncoder / OobleckDecoder / VAEBottleneck
(stable_audio_tools/models/autoencoders.py:59-85,228-357, bottleneck.py:57-86) in the `oobleck_finetune.json`
configuration (ELU activations, weight-normalised convolutions, final tanh).

`OobleckVAE.load_state_dict` takes the reference's key layout (encoder.layers.N..., decoder.layers.N..., weight_g /
weight_v of torch.nn.utils.weight_norm, or plain .weight), folds the weight norm and re-packs every filter once for the
tensor-core kernel of csrc/conv1d.cu.  Activations live as [B][T][C] bf16; every ELU is produced by the epilogue of the
convolution in front of it, strided convolutions read a reshaped view instead of gathering, transposed convolutions run
in polyphase form and hand a shifted view of their output buffer to the next layer -- no pass over the data that is not
a convolution.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Sequence

import torch

from .. import ops


def _wn(sd: Dict[str, torch.Tensor], key: str) -> torch.Tensor:
    """Effective weight of a weight-normalised layer: g * v / ||v|| over every axis but the first."""
    if key + ".weight" in sd:
        return sd[key + ".weight"].float()
    v, g = sd[key + ".weight_v"].float(), sd[key + ".weight_g"].float()
    return v * (g / v.flatten(1).norm(dim=1).reshape(-1, *([1] * (v.dim() - 1))))


class _Conv:
    """One tensor-core convolution: packed bf16 filter [Cout][taps * Cin] + fp32 bias + its tap geometry."""

    def __init__(self, w, bias, taps, dil, pad, s_view, cout, device):
        self.w = w.to(device=device, dtype=torch.bfloat16).contiguous()
        self.bias = None if bias is None else bias.to(device=device, dtype=torch.float32).contiguous()
        self.taps, self.dil, self.pad, self.s_view, self.cout = taps, dil, pad, s_view, cout

    @classmethod
    def plain(cls, sd, key, dil, device, stride=1):
        w = _wn(sd, key)                                     # (Cout, Cin, K)
        cout, cin, k = w.shape
        pad = dil * (k - 1) // 2 if stride == 1 else math.ceil(stride / 2)
        return cls(w.permute(0, 2, 1).reshape(cout, k * cin), sd.get(key + ".bias"), k, dil, pad, stride, cout, device)

    @classmethod
    def transposed(cls, sd, key, s, device):
        w = _wn(sd, key)                                     # (Cin, Cout, 2s); tap k = m * s + f
        cin, cout, k = w.shape
        assert k == 2 * s
        wp = w.reshape(cin, cout, 2, s).permute(3, 1, 2, 0).reshape(s * cout, 2 * cin)   # [f * Cout + co][m * Cin + ci]
        b = sd.get(key + ".bias")
        return cls(wp, None if b is None else b.float().repeat(s), 2, -1, 0, 1, s * cout, device)


class OobleckVAE:
    """encode: (B, 1, T) waveform -> (B, latent, T / hop) latents; decode: latents -> (B, 1, T' * hop) waveform."""

    def __init__(self, channels: int = 128, c_mults: Sequence[int] = (1, 2, 4, 8, 16), strides: Sequence[int] = (2, 4, 4, 8, 8),
                 enc_latent: int = 128, final_tanh: bool = True, device="cuda"):
        if any(s % 2 for s in strides):
            raise ValueError("OobleckVAE: even strides only (kernel 2s, padding s/2)")
        if channels % 64 or (enc_latent // 2) % 64:
            raise ValueError("OobleckVAE: channels and the latent width must be multiples of 64 (tensor-core K blocks)")
        self.channels, self.c_mults, self.strides = channels, [1] + list(c_mults), list(strides)
        self.enc_latent, self.latent, self.final_tanh = enc_latent, enc_latent // 2, final_tanh
        self.hop_length = int(math.prod(strides))
        self.device = torch.device(device)
        self._w = None
        self._plans = {}

    # ------------------------------------------------------------------ weights
    def load_state_dict(self, sd: Dict[str, torch.Tensor], strict: bool = True):
        dev, n = self.device, len(self.strides)
        used = set()

        def take(cons, key, *a):
            for suf in (".weight", ".weight_g", ".weight_v", ".bias"):
                if key + suf in sd:
                    used.add(key + suf)
            return cons(sd, key, *a)

        w = {}
        first = _wn(sd, "encoder.layers.0")                  # (C, 1, 7)
        used.update(k for k in sd if k.startswith("encoder.layers.0."))
        w["enc_in"] = (first[:, 0, :].contiguous().to(dev), sd["encoder.layers.0.bias"].float().to(dev))
        for i, s in enumerate(self.strides):
            b = f"encoder.layers.{i + 1}.layers."
            for j, d in enumerate((1, 3, 9)):
                w[f"enc{i}_ru{j}a"] = take(_Conv.plain, f"{b}{j}.layers.1", d, dev)
                w[f"enc{i}_ru{j}b"] = take(_Conv.plain, f"{b}{j}.layers.3", 1, dev)
            w[f"enc{i}_down"] = take(_Conv.plain, b + "4", 1, dev, s)
        w["enc_out"] = take(_Conv.plain, f"encoder.layers.{n + 2}", 1, dev)
        w["dec_in"] = take(_Conv.plain, "decoder.layers.0", 1, dev)
        for idx, i in enumerate(range(n, 0, -1)):
            b, s = f"decoder.layers.{idx + 1}.layers.", self.strides[i - 1]
            w[f"dec{idx}_up"] = take(_Conv.transposed, b + "1", s, dev)
            for j, d in enumerate((1, 3, 9)):
                w[f"dec{idx}_ru{j}a"] = take(_Conv.plain, f"{b}{j + 2}.layers.1", d, dev)
                w[f"dec{idx}_ru{j}b"] = take(_Conv.plain, f"{b}{j + 2}.layers.3", 1, dev)
        last_key = f"decoder.layers.{n + 2}"
        last = _wn(sd, last_key)                             # (1, C, 7), no bias
        used.update(k for k in sd if k.startswith(last_key + "."))
        w["dec_out"] = last[0].t().contiguous().to(dev)      # [7][C]
        if strict:
            extra = sorted(k for k in sd if (k.startswith("encoder.") or k.startswith("decoder.")) and k not in used)
            if extra:
                raise KeyError(f"OobleckVAE.load_state_dict: unexpected keys {extra[:4]} ...")
        self._w = w
        self._plans = {}
        return self

    def _weights(self):
        if self._w is None:
            raise ops.DitsepError("OobleckVAE: load_state_dict first")
        return self._w

    # ------------------------------------------------------------------ building blocks ([B][T][C] bf16, views explicit)
    @staticmethod
    def _conv(plan, c: _Conv, x, T_in, bs_in, T_out, residual=None, res_geom=(0, 0), out=None, out_act=None, ld_out=0, bs_out=0):
        plan.add(ops.conv1d_tc, x, T_in, bs_in, c.w, c.cout, c.taps, c.dil, c.pad, c.s_view, T_out, bias=c.bias,
                 residual=residual, ld_res=res_geom[0], bs_res=res_geom[1], out=out, out_act=out_act, ld_out=ld_out,
                 bs_out=bs_out)

    def _res_units(self, plan, w, tag, h, a, T, C, bs):
        """Three residual units (dilations 1, 3, 9) on the stream h with a = ELU(h), both views of rows C / signals bs
        apart: t = ELU(conv7_d(a)); h += conv1(t); a = ELU(h) -- all in the two convolutions' epilogues, in place."""
        B = h.shape[0]
        t = torch.empty(B, T, C, device=h.device, dtype=torch.bfloat16)
        for j in range(3):
            self._conv(plan, w[f"{tag}_ru{j}a"], a, T, bs, T, out_act=t, ld_out=C, bs_out=T * C)
            self._conv(plan, w[f"{tag}_ru{j}b"], t, T, T * C, T, residual=h, res_geom=(C, bs), out=h, out_act=a, ld_out=C,
                       bs_out=bs)

    # ------------------------------------------------------------------ launch plans (one per shape, CUDA-graph replayed)
    def _encode_plan(self, B: int, T: int, dev):
        w, plan = self._weights(), _Plan(dev)
        C = self.c_mults[0] * self.channels
        plan.x = torch.empty(B, T, device=dev, dtype=torch.float32)
        h = torch.empty(B, T, C, device=dev, dtype=torch.bfloat16)
        a = torch.empty_like(h)
        plan.add(ops.conv1d_in, plan.x, w["enc_in"][0], w["enc_in"][1], h, a)
        for i, s in enumerate(self.strides):
            self._res_units(plan, w, f"enc{i}", h, a, T, C, T * C)
            Cn, Tn = self.c_mults[i + 1] * self.channels, T // s
            hn = torch.empty(B, Tn, Cn, device=dev, dtype=torch.bfloat16)
            an = torch.empty_like(hn)
            self._conv(plan, w[f"enc{i}_down"], a, T, T * C, Tn, out=hn, out_act=an, ld_out=Cn, bs_out=Tn * Cn)
            h, a, T, C = hn, an, Tn, Cn
        plan.pre = torch.empty(B, T, self.enc_latent, device=dev, dtype=torch.float32)
        self._conv(plan, w["enc_out"], a, T, T * C, T, out=plan.pre, ld_out=self.enc_latent, bs_out=T * self.enc_latent)
        plan.noise = torch.empty(B, self.latent, T, device=dev, dtype=torch.float32)
        plan.z = torch.empty(B, self.latent, T, device=dev, dtype=torch.float32)
        plan.add(ops.vae_sample, plan.pre, plan.noise, plan.z)
        return plan

    def _decode_plan(self, B: int, T: int, dev):
        w, plan = self._weights(), _Plan(dev)
        L, C = self.latent, self.c_mults[-1] * self.channels
        plan.zin = torch.empty(B, T, L, device=dev, dtype=torch.bfloat16)            # [B][T'][latent]
        h = torch.empty(B, T, C, device=dev, dtype=torch.bfloat16)
        a = torch.empty_like(h)
        self._conv(plan, w["dec_in"], plan.zin, T, T * L, T, out=h, out_act=a, ld_out=C, bs_out=T * C)
        bs = T * C
        for idx, i in enumerate(range(len(self.strides), 0, -1)):
            s, Cn = self.strides[i - 1], self.c_mults[i - 1] * self.channels
            # polyphase transposed convolution: rows (j, f) of a [B][(T + 1) * s][Cn] buffer; the layer's output is the view
            # that starts s / 2 rows in (signals (T + 1) * s * Cn apart)
            hb = torch.empty(B, (T + 1) * s, Cn, device=dev, dtype=torch.bfloat16)
            ab = torch.empty_like(hb)
            self._conv(plan, w[f"dec{idx}_up"], a, T, bs, T + 1, out=hb, out_act=ab, ld_out=s * Cn, bs_out=(T + 1) * s * Cn)
            h, a = hb[:, s // 2:], ab[:, s // 2:]
            T, C, bs = T * s, Cn, (T + 1) * s * Cn
            self._res_units(plan, w, f"dec{idx}", h, a, T, C, bs)
        plan.out = torch.empty(B, T, device=dev, dtype=torch.float32)
        plan.add(ops.conv1d_out, a, T, C, bs, w["dec_out"], plan.out, self.final_tanh)
        return plan

    def _plan(self, kind: str, B: int, T: int, dev):
        key = (kind, B, T, str(dev))
        if key not in self._plans:
            if len(self._plans) >= 8:  # a handful of shapes per run; drop the oldest rather than grow without bound
                self._plans.pop(next(iter(self._plans)))
            self._plans[key] = (self._encode_plan if kind == "enc" else self._decode_plan)(B, T, dev)
        return self._plans[key]

    # ------------------------------------------------------------------ encoder / bottleneck / decoder
    @torch.no_grad()
    def encode(self, audio: torch.Tensor, noise: Optional[torch.Tensor] = None, seed: int = 0,
               skip_bottleneck: bool = False) -> torch.Tensor:
        """AudioAutoencoder.encode (autoencoders.py:440-485) without pretransform: audio (B, 1, T), T a multiple of the hop.
        noise: the bottleneck's randn draw (B, latent, T / hop) (parity runs); default: the library's Philox stream."""
        ops.require_device(audio, "OobleckVAE.encode")
        B, one, T = audio.shape
        if one != 1 or T % self.hop_length:
            raise ValueError(f"OobleckVAE.encode: (B, 1, T) with T a multiple of {self.hop_length}, got {tuple(audio.shape)}")
        plan = self._plan("enc", B, T, audio.device)
        plan.x.copy_(audio.reshape(B, T))
        if noise is None:
            ops.philox_normal(plan.noise, seed, 0)
        else:
            plan.noise.copy_(noise)
        plan.run()
        return plan.pre.permute(0, 2, 1).contiguous() if skip_bottleneck else plan.z.clone()

    @torch.no_grad()
    def decode(self, latents: torch.Tensor) -> torch.Tensor:
        """AudioAutoencoder.decode (autoencoders.py:487-528): latents (B, latent, T') -> (B, 1, T' * hop)."""
        ops.require_device(latents, "OobleckVAE.decode")
        B, L, T = latents.shape
        if L != self.latent:
            raise ValueError(f"OobleckVAE.decode: {self.latent} latent channels expected, got {L}")
        plan = self._plan("dec", B, T, latents.device)
        plan.zin.copy_(latents.permute(0, 2, 1))
        plan.run()
        return plan.out.clone()[:, None, :]


class _Plan:
    """A fixed launch list over buffers it owns: eager on the first call, then captured into a CUDA graph and replayed
    (tensor maps are kernel parameters, so they are baked in; DITSEP_GRAPH=0 keeps it eager) -- the launch-plan idiom of
    models/diffsep/engine.py.  ~70 - 110 launches of 10 - 100 us each are host-bound otherwise."""

    def __init__(self, device):
        self.steps, self.device, self._graph, self._runs, self._graph_launches = [], device, None, 0, 0

    def add(self, fn, *args, **kwargs):
        self.steps.append((fn, args, kwargs))

    def run_eager(self):
        for fn, args, kwargs in self.steps:
            fn(*args, **kwargs)

    def run(self):
        if not ops.graphs_enabled():
            return self.run_eager()
        if self._graph is None:
            if self._runs == 0:
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
            ops.note_launches(-self._graph_launches)
        self._graph.replay()
        ops.note_launches(self._graph_launches)


class LatentCodec:
    """LatentDiffSep.encode / decode (src/diffsep_latent.py:107-128) around an OobleckVAE."""

    def __init__(self, vae: OobleckVAE):
        self.vae = vae

    @staticmethod
    def pad(x: torch.Tensor, hop: int) -> torch.Tensor:
        """src/utils/torch_utils.py:12-18: ALWAYS pads (a whole hop when the length is already a multiple)."""
        return torch.nn.functional.pad(x, (0, hop - x.shape[-1] % hop))

    def encode(self, mix, target=None, noise=None, noise_target=None, seed: int = 0):
        hop = self.vae.hop_length
        z = self.vae.encode(self.pad(mix, hop), noise=noise, seed=seed).unsqueeze(1)
        if target is not None:
            t = self.pad(target, hop)
            bsz, n_src, seq = t.shape
            zt = self.vae.encode(t.reshape(bsz * n_src, 1, seq), noise=noise_target, seed=seed + 1)
            target = zt.reshape(bsz, n_src, *zt.shape[1:])
        return z, target

    def decode(self, est, target_dim=None):
        bsz, n_src, latent, seq = est.shape
        out = self.vae.decode(est.reshape(bsz * n_src, latent, seq)).reshape(bsz, n_src, -1)
        return out[..., :target_dim] if target_dim is not None else out
