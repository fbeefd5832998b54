"""N>1 path on CPU: world_size-2 gloo processes, kernels replaced by the torch stand-ins of tests/emu_ops.py.
What is exercised is the product's own host logic: length-bucketed round-robin partition, per-rank sampler runs
with no data-path collective, and the single final gather of metric rows (ditsep_b200/utils/shard.py)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ditsep_b200.utils import shard

LENGTHS = [2048, 1500, 2048, 3000, 1500, 1800, 2500, 1700]  # 1500..2500 share one frame bucket (64 frames)


def _mixtures():
    g = torch.Generator().manual_seed(99)
    srcs = [torch.randn(2, n, generator=g) for n in LENGTHS]
    return [s.sum(0, keepdim=True) for s in srcs], srcs


def _run(rank, world, port, q, max_batch=1):
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path[:0] = [here, os.path.dirname(here)]
    from common import build_model
    from emu_ops import emulated_ops
    from oracle import sampler_ref as S
    torch.set_num_threads(2)
    if world > 1:
        dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    model, _, _ = build_model("fp32", N=2)
    mixes, srcs = _mixtures()
    with emulated_ops(), torch.no_grad():
        rows, outs = shard.separate_sharded(model, mixes, references=srcs, max_batch=max_batch,
                                            sampler_kwargs={"N": 2}, rank=rank, world_size=world,
                                            noise_tape_factory=lambda i: S.NoiseTape(4321 + i).draw)
    q.put((rank, rows.numpy().copy(), {k: v.numpy().copy() for k, v in outs.items()}))  # by value: the child exits
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _launch(world, max_batch=1):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_run, args=(r, world, port, q, max_batch)) for r in range(world)]
    for p in procs:
        p.start()
    got = []
    import queue as _q
    import time
    t_end = time.time() + 600
    while len(got) < world and time.time() < t_end:
        try:
            got.append(q.get(timeout=2))
        except _q.Empty:
            assert all(p.is_alive() or p.exitcode == 0 for p in procs), "a rank died"
    assert len(got) == world
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    got = [(r, torch.from_numpy(rows), {k: torch.from_numpy(v) for k, v in outs.items()}) for r, rows, outs in got]
    return sorted(got, key=lambda t: t[0])


def test_partition_is_balanced_and_complete():
    lengths = [32000] * 10 + [64000] * 7 + [20000] * 5
    for world in (1, 2, 4, 8):
        parts = shard.bucket_round_robin(lengths, world)
        assert sorted(i for p in parts for i in p) == list(range(len(lengths)))
        sizes = [len(p) for p in parts]
        assert max(sizes) - min(sizes) <= 1
        # every rank gets the same number of long mixtures to +-1 (length balance, not just count balance)
        longs = [sum(1 for i in p if lengths[i] == 64000) for p in parts]
        assert max(longs) - min(longs) <= 1
    for b in shard.batches_for_rank(lengths, 2, 0, max_batch=4):
        assert len(b) <= 4 and len({shard.padded_frames(lengths[i]) for i in b}) == 1
    assert shard.padded_frames(32000) == 256 and shard.padded_frames(160000) == 1280


def test_si_sdr_pit_known_values():
    g = torch.Generator().manual_seed(3)
    ref = torch.randn(2, 2, 4000, generator=g)
    est = torch.stack([ref[:, 1] * 0.5, ref[:, 0] * 2.0 + 0.1 * torch.randn(2, 4000, generator=g)], dim=1)  # swapped
    sdr, perm = shard.si_sdr_pit(ref, est)
    assert perm.tolist() == [[1, 0], [1, 0]]
    assert torch.all(sdr[:, 1] > 60)                         # exact scaled copy
    assert torch.all((sdr[:, 0] > 23) & (sdr[:, 0] < 29))    # 20 log10(2 / 0.1) = 26 dB


def test_world_size_2_gloo_matches_single_process():
    one = _launch(1)
    two = _launch(2)
    rows1 = one[0][1]
    assert rows1.shape == (len(LENGTHS), 4) and rows1[:, 0].tolist() == list(range(len(LENGTHS)))
    for rank, rows, outs in two:
        assert torch.equal(rows, rows1)                       # every rank holds the full, identical table
    parts = shard.bucket_round_robin(LENGTHS, 2)
    for rank, rows, outs in two:
        assert sorted(outs) == sorted(parts[rank])            # each rank separated exactly its shard
        for i, est in outs.items():
            assert est.shape == (2, LENGTHS[i])
            assert torch.equal(est, one[0][2][i])


def test_ragged_batches_match_batch_1_runs():
    """Mixtures of DIFFERENT lengths sharing a batch (one frame bucket, zero-padded, per-mixture statistics, state
    masked past each length) give what each mixture gives alone at batch 1 with the same noise tape -- the
    reference evaluates at batch 1 (src/evaluate_mp.py:289,311).  Round 1 normalised over the padding and let the
    prior noise into it."""
    one = _launch(1, max_batch=1)
    rag = _launch(2, max_batch=4)
    assert any(len({LENGTHS[i] for i in b}) > 1 for r in range(2) for b in shard.batches_for_rank(LENGTHS, 2, r, 4))
    ref_outs = one[0][2]
    for rank, rows, outs in rag:
        assert torch.allclose(rows, one[0][1], atol=2e-3, rtol=0)       # SI-SDR rows, dB
        for i, est in outs.items():
            d = float((est - ref_outs[i]).norm() / ref_outs[i].norm())
            assert est.shape == (2, LENGTHS[i]) and d < 2e-5, (i, d)    # fp32 emulation: summation order only
