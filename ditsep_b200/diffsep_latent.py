"""``LatentDiffSepModel``: the latent separation model (mirror of the inference side of src/diffsep_latent.py).

``encode`` / ``decode`` (:107-128) wrap the Oobleck VAE around the latent PC sampler; ``separate`` (:472-487) is
encode -> sampler ("reverse_diffusion" + "ald" over the OUVE SDE, LatentScoreModelNCSNpp) -> decode.  Training losses, EMA
and the Lightning plumbing are out of scope."""
from __future__ import annotations

import json
from typing import Optional

import torch

from .diffsep import DiffSepModel, _to_plain
from .models.oobleck import LatentCodec, OobleckVAE


def latent_default_config(nf=128, n_src=2, N=30, precision=None, max_latent_length=4):
    """src/config/latent_diffsep_ouve/model/default.yaml:16-46 as a plain dict."""
    backbone = {"_target_": "models.diffsep.ncsnpp.NCSNpp", "nf": nf, "ch_mult": [1, 2, 2], "num_res_blocks": 2,
                "attn_resolutions": [16], "resamp_with_conv": True, "image_size": 64, "centered": True}
    if precision is not None:
        backbone["precision"] = precision
    return {"model": {
        "fs": 8000, "n_speakers": n_src, "t_eps": 0.03, "t_rev_init": 0.03,
        "score_model": {"_target_": "models.diffsep.score_models.LatentScoreModelNCSNpp", "num_sources": n_src,
                        "backbone_args": backbone, "max_latent_length": max_latent_length},
        "sde": {"_target_": "sdes.OUVESDE", "theta": 1.5, "sigma_min": 0.96, "sigma_max": 10.0, "N": N},
        "sampler": {"N": N, "snr": 0.5, "corrector_steps": 1},
        "vae": {"config_path": None, "ckpt_path": None, "trainable_vae": False,
                "config": {"encoder": {"type": "oobleck", "config": {"in_channels": 1, "channels": 128,
                                                                      "c_mults": [1, 2, 4, 8, 16], "strides": [2, 4, 4, 8, 8],
                                                                      "latent_dim": 128}},
                           "decoder": {"type": "oobleck", "config": {"out_channels": 1, "channels": 128,
                                                                      "c_mults": [1, 2, 4, 8, 16], "strides": [2, 4, 4, 8, 8],
                                                                      "latent_dim": 64}},
                           "bottleneck": {"type": "vae"}}},
    }}


def vae_from_config(vae_cfg, device="cuda") -> OobleckVAE:
    """utils.load_stable_model (diffsep_latent.py:41-45): the stable-audio-tools JSON at `config_path` (or its `model`
    section given inline as `config`), `oobleck` encoder / decoder with a `vae` bottleneck; weights from `ckpt_path` when
    given (a state_dict, possibly under "state_dict")."""
    cfg = _to_plain(vae_cfg)
    model = cfg.get("config")
    if cfg.get("config_path"):
        with open(cfg["config_path"]) as f:
            model = json.load(f)["model"]
    enc, dec = model["encoder"], model["decoder"]
    if enc.get("type") != "oobleck" or dec.get("type") != "oobleck" or model.get("bottleneck", {}).get("type") != "vae":
        raise NotImplementedError("only the oobleck encoder / decoder with a vae bottleneck has sm_100a kernels")
    ec, dc = enc["config"], dec["config"]
    if ec.get("use_snake") or dc.get("use_snake") or ec.get("in_channels", 1) != 1 or dc.get("out_channels", 1) != 1:
        raise NotImplementedError("oobleck_finetune.json shape only: ELU activations, one audio channel")
    vae = OobleckVAE(ec.get("channels", 128), ec["c_mults"], ec["strides"], ec["latent_dim"], dc.get("final_tanh", True),
                     device=device)
    if cfg.get("ckpt_path"):
        sd = torch.load(cfg["ckpt_path"], map_location="cpu")
        vae.load_state_dict(sd.get("state_dict", sd))
    return vae


class LatentDiffSepModel(DiffSepModel):
    def __init__(self, config, vae: Optional[OobleckVAE] = None, device="cuda"):
        super().__init__(config)
        self.vae = vae if vae is not None else vae_from_config(self.config.model.vae, device=device)
        self.codec = LatentCodec(self.vae)
        self.max_len_lat = 0

    @torch.no_grad()
    def encode(self, mix, target=None, **noise_kw):
        """diffsep_latent.py:107-118: pad to the hop, VAE-encode; mix (B, 1, T) -> (B, 1, latent, T'), target (B, n, T) ->
        (B, n, latent, T')."""
        z, zt = self.codec.encode(mix, target, **noise_kw)
        self.max_len_lat = max(self.max_len_lat, z.shape[-1])
        return z, zt

    @torch.no_grad()
    def decode(self, est, target_dim=None):
        """diffsep_latent.py:120-128."""
        return self.codec.decode(est, target_dim)

    def separate(self, mix, target_dim=None, latent=False, **kwargs):
        """diffsep_latent.py:472-487: (B, 1, T) waveform (or latents when latent=True) -> (B, n_src, T) estimates, nfe."""
        if not latent:
            mix, _ = self.encode(mix, None, **{k: kwargs.pop(k) for k in ("noise",) if k in kwargs})
        sampler_kwargs = dict(_to_plain(self.config.model.sampler))
        sampler_kwargs.update(kwargs)
        sampler = self.get_pc_sampler("reverse_diffusion", "ald", mix, **sampler_kwargs)
        est, *others = sampler()
        return (self.decode(est, target_dim), *others)
