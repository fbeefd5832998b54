"""Batch sharding of independent mixtures over ranks (one process per GPU) and the final gather.

The sampler has no cross-mixture coupling (sdes/__init__.py:133-193 of the reference works per batch row
for ald2 / ald), so the path shards by mixture with NO data-path collective.  The reference splits a
dataset into contiguous index ranges per device (src/evaluate_mp.py:496-504), which is length-imbalanced;
here mixtures are bucketed by padded STFT frame count and dealt round-robin inside each bucket, so every
rank sees the same mix of lengths.  The only collective is one all-gather of the per-mixture metric rows
(and, optionally, of the padded separated waveforms) at the end of a sweep (src/evaluate_mp.py:506-514
collects the per-process results the same way, through a multiprocessing queue).
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

import torch
import torch.distributed as dist

HOP, PAD_SAMPLES, FRAME_QUANTUM = 128, 382, 64  # score_models.py:83-91,107-116 (hop 128, right pad 382, frames -> x64)


def padded_frames(n_samples: int) -> int:
    """Number of STFT frames the score network sees for a mixture of n_samples (multiple of 64)."""
    frames = 1 + (n_samples + PAD_SAMPLES) // HOP
    return (frames + FRAME_QUANTUM - 1) // FRAME_QUANTUM * FRAME_QUANTUM


def bucket_round_robin(lengths: Sequence[int], world_size: int) -> List[List[int]]:
    """Indices of the mixtures each rank processes: sort by (padded frame count, index), deal `idx % world_size`
    inside each bucket, carrying the deal position across buckets so ranks stay balanced to +-1 mixture."""
    if world_size < 1:
        raise ValueError(f"world_size={world_size}")
    order = sorted(range(len(lengths)), key=lambda i: (padded_frames(int(lengths[i])), i))
    parts: List[List[int]] = [[] for _ in range(world_size)]
    for pos, i in enumerate(order):
        parts[pos % world_size].append(i)
    return parts


def batches_for_rank(lengths: Sequence[int], world_size: int, rank: int, max_batch: int) -> List[List[int]]:
    """This rank's mixtures grouped into batches of equal padded frame count (one launch plan per group)."""
    mine = bucket_round_robin(lengths, world_size)[rank]
    by_frames: Dict[int, List[int]] = {}
    for i in mine:
        by_frames.setdefault(padded_frames(int(lengths[i])), []).append(i)
    out: List[List[int]] = []
    for fr in sorted(by_frames):
        idx = by_frames[fr]
        out += [idx[k:k + max_batch] for k in range(0, len(idx), max_batch)]
    return out


def si_sdr_pit(ref: torch.Tensor, est: torch.Tensor, eps: float = 1e-8) -> Tuple[torch.Tensor, torch.Tensor]:
    """Scale-invariant SDR with the best source permutation, per mixture.  ref, est: (B, n_src, T).
    Restates fast_bss_eval.si_bss_eval_sources(ref, est, zero_mean=False, compute_permutation=True) as the
    reference calls it (src/evaluate_mp.py:173-175): alpha = <est, ref> / <ref, ref>,
    10 log10(|alpha ref|^2 / |est - alpha ref|^2).  Returns (si_sdr (B, n_src) in reference order, perm (B, n_src))."""
    import itertools

    B, n, _ = ref.shape
    dot = torch.einsum("bit,bjt->bij", ref, est)            # <ref_i, est_j>
    rr = (ref * ref).sum(-1)                                 # (B, n)
    ee = (est * est).sum(-1)
    proj = dot * dot / (rr[:, :, None] + eps)                # |alpha ref_i|^2 for pair (i, j)
    noise = ee[:, None, :] - proj
    val = 10.0 * torch.log10((proj + eps) / (noise.clamp_min(0.0) + eps))  # (B, n ref, n est)
    perms = list(itertools.permutations(range(n)))
    scores = torch.stack([sum(val[:, i, p[i]] for i in range(n)) for p in perms], dim=1)  # (B, n!)
    best = scores.argmax(dim=1)
    perm = torch.tensor(perms, device=ref.device)[best]      # (B, n): est index matched to ref i
    sdr = torch.gather(val, 2, perm[:, :, None]).squeeze(2)
    return sdr, perm


def gather_rows(rows: torch.Tensor, group=None) -> torch.Tensor:
    """All-gather of per-mixture metric rows.  rows: (n_local, k) float32, column 0 = global mixture index.
    Ranks may hold different n_local: rows are padded to the maximum (pad index -1), gathered with ONE
    all_gather_into_tensor (NCCL) or all_gather (gloo), the padding dropped and the result sorted by index."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return rows[torch.argsort(rows[:, 0])] if rows.numel() else rows
    world = dist.get_world_size(group)
    n = torch.tensor([rows.shape[0]], device=rows.device, dtype=torch.int64)
    dist.all_reduce(n, op=dist.ReduceOp.MAX, group=group)
    n_max = int(n.item())
    k = rows.shape[1]
    padded = torch.full((n_max, k), -1.0, device=rows.device, dtype=torch.float32)
    padded[:rows.shape[0]] = rows
    if rows.is_cuda:
        out = torch.empty(world * n_max, k, device=rows.device, dtype=torch.float32)
        dist.all_gather_into_tensor(out, padded, group=group)
    else:
        parts = [torch.empty_like(padded) for _ in range(world)]
        dist.all_gather(parts, padded, group=group)
        out = torch.cat(parts, dim=0)
    out = out[out[:, 0] >= 0]
    return out[torch.argsort(out[:, 0])]


def ragged_tape(tape_for_mixture, indices: Sequence[int], lens: Sequence[int]):
    """Noise tape of a ragged batch built from per-mixture tapes (parity runs): draw i of the batch is the stack of
    every mixture's own draw i at its OWN length, zero-padded to the batch length -- so a mixture consumes the same
    noise whether it runs alone or inside any batch.  tape_for_mixture(idx) -> callable (i, shape) -> tensor."""
    tapes = [tape_for_mixture(i) for i in indices]

    def tape(i, like):
        z = torch.zeros(like.shape, dtype=torch.float32)
        for r, (tp, n) in enumerate(zip(tapes, lens)):
            z[r, :, :n] = tp(i, (1, like.shape[1], n))[0]
        return z

    return tape


def normalize_ragged(mixtures: Sequence[torch.Tensor], batch: Sequence[int], device=None):
    """Zero-padded (B, 1, T_max) batch of the mixtures `batch`, each standardised over ITS OWN samples (mean, unbiased
    std clamped at 1e-5: src/utils/separate.py:56-62 applied per mixture as the reference's batch-1 evaluation does,
    src/evaluate_mp.py:289,311); the padding stays exactly zero.  Returns (y, lengths int32 (B,), mean (B,1,1), std)."""
    lens = [int(mixtures[i].shape[-1]) for i in batch]
    t_run = max(lens)
    y = torch.zeros(len(batch), 1, t_run)
    mean, std = torch.zeros(len(batch), 1, 1), torch.ones(len(batch), 1, 1)
    for r, i in enumerate(batch):
        m = mixtures[i].reshape(-1).float().cpu()
        mean[r], std[r] = m.mean(), m.std().clamp(min=1e-5)
        y[r, 0, :lens[r]] = (m - mean[r, 0, 0]) / std[r, 0, 0]
    lengths = torch.tensor(lens, dtype=torch.int32)
    if device is not None:
        y, lengths, mean, std = y.to(device), lengths.to(device), mean.to(device), std.to(device)
    return y, lengths, mean, std


def gather_waveforms(outs: Dict[int, torch.Tensor], lengths: Sequence[int], device, group=None) -> Dict[int, torch.Tensor]:
    """All-gather of the separated waveforms (the optional second collective of SURVEY.md 8e).  outs: this rank's
    {idx: (n_src, T_idx)}; every rank gets every mixture back.  ONE all_gather_into_tensor of a zero-padded
    (n_local_max, 1 + n_src * T_max) block per rank (slot 0 carries the global index, -1 = padding)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return dict(outs)
    world = dist.get_world_size(group)
    t_max = max(int(n) for n in lengths)
    n_src = next(iter(outs.values())).shape[0] if outs else 2
    n = torch.tensor([len(outs), n_src], device=device, dtype=torch.int64)
    dist.all_reduce(n, op=dist.ReduceOp.MAX, group=group)
    n_max, n_src = int(n[0].item()), int(n[1].item())
    block = torch.zeros(n_max, 1 + n_src * t_max, device=device, dtype=torch.float32)
    block[:, 0] = -1.0
    for r, (i, est) in enumerate(sorted(outs.items())):
        block[r, 0] = float(i)
        block[r, 1:].view(n_src, t_max)[:, :est.shape[-1]] = est.to(device)
    if block.is_cuda:
        full = torch.empty(world * n_max, block.shape[1], device=device, dtype=torch.float32)
        dist.all_gather_into_tensor(full, block, group=group)
    else:
        parts = [torch.empty_like(block) for _ in range(world)]
        dist.all_gather(parts, block, group=group)
        full = torch.cat(parts, dim=0)
    res = {}
    for row in full[full[:, 0] >= 0]:
        i = int(row[0].item())
        res[i] = row[1:].view(n_src, t_max)[:, :int(lengths[i])]
    return res


def separate_sharded(model, mixtures: Sequence[torch.Tensor], references=None, max_batch: int = 32, device=None,
                     sampler_kwargs=None, seed: int = 0, rank: int = None, world_size: int = None,
                     noise_tape_factory=None, predictor: str = "reverse_diffusion", corrector: str = "ald2",
                     row_fn=None):
    """Sweep over independent mixtures, batch-sharded over the ranks of the default process group.

    mixtures: list of (1, T_i) tensors (any lengths); references: optional list of (n_src, T_i).
    Every rank runs `model.get_pc_sampler` on its own length-bucketed batches (no collective inside), then the
    rows [idx, n_samples, si_sdr_0..n-1] are gathered once.  Returns (rows (n_total, 2 + n_src) on every rank,
    {idx: separated (n_src, T_i)} for this rank's mixtures).

    Mixtures of one bucket (same padded STFT frame count) share a batch although their lengths differ by up to 8191
    samples: each is standardised over its own samples, zero-padded, and the sampler is told the valid lengths
    (`lengths=`), which keeps the state at zero past each mixture's end.  In exact arithmetic every mixture then sees
    the frames, the GroupNorm statistics and the score of a batch-1 run of itself -- the reference's evaluation
    (src/evaluate_mp.py:289,311); with a per-mixture noise tape (`ragged_tape`) the results are independent of
    max_batch and of the world size (tests/test_sharding_gloo.py, tests/test_gpu_parity.py).
    noise_tape_factory(idx) -> callable (i, shape) -> tensor: the tape of mixture idx (parity runs); default:
    in-kernel Philox keyed by the global index of the batch's first mixture.
    row_fn(idx, est (n_src, T_i) de-normalised, nfe, runtime_s) -> list of floats: replaces the default row (the
    evaluation driver, ditsep_b200/evaluate.py, uses it for the reference's row schema); runtime_s = the batch's
    sampler wall time (device synchronised) divided by the batch size.  sampler_kwargs may carry `schedule`
    ("linear" / "log" / "revlog": the scheduled sampler, diffsep.py:100-106 here, src/diffsep.py:404-441 there)."""
    import time

    if world_size is None:
        world_size = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
    if rank is None:
        rank = dist.get_rank() if dist.is_available() and dist.is_initialized() else 0
    kw = {"N": 30, "snr": 0.5, "corrector_steps": 1, "denoise": True}
    kw.update(sampler_kwargs or {})
    lengths = [int(m.shape[-1]) for m in mixtures]
    n_src = int(kw.get("n_spkrs", getattr(model.sde, "ndim", 2)))
    rows, outs = [], {}
    for batch in batches_for_rank(lengths, world_size, rank, max_batch):
        yn, lens, mean, std = normalize_ragged(mixtures, batch, device)
        noise = ({"noise_tape": ragged_tape(noise_tape_factory, batch, [lengths[i] for i in batch])}
                 if noise_tape_factory is not None else {"seed": seed + 7919 * batch[0]})
        ragged = len(set(lengths[i] for i in batch)) > 1
        sampler = model.get_pc_sampler(predictor, corrector, yn, lengths=lens if ragged else None, **noise, **kw)
        if yn.is_cuda:
            torch.cuda.synchronize(yn.device)
        t0 = time.perf_counter()
        x, nfe = sampler()
        if yn.is_cuda:
            torch.cuda.synchronize(yn.device)
        t_run = (time.perf_counter() - t0) / len(batch)
        x = x * std + mean
        for r, i in enumerate(batch):
            est = x[r, :, :lengths[i]]
            outs[i] = est
            if row_fn is not None:
                rows.append([float(v) for v in row_fn(i, est, nfe, t_run)])
                continue
            row = [float(i), float(lengths[i])]
            if references is not None:
                ref = references[i].to(est.device)[None]
                sdr, _ = si_sdr_pit(ref, est[None])
                row += [float(v) for v in sdr[0]]
            else:
                row += [float("nan")] * n_src
            rows.append(row)
    dev = device if device is not None else "cpu"
    width = len(rows[0]) if rows else 2 + n_src
    if row_fn is not None and dist.is_available() and dist.is_initialized() and world_size > 1:
        w = torch.tensor([width], device=dev, dtype=torch.int64)   # a rank with no mixtures learns the width
        dist.all_reduce(w, op=dist.ReduceOp.MAX)
        width = int(w.item())
    table = torch.tensor(rows, dtype=torch.float32, device=dev).reshape(-1, width)
    return gather_rows(table), outs
