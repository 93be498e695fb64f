"""Host-side multi-GPU logic (utterance sharding + waveform gather) on CPU: world_size 2, gloo backend."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mlx_audio_b200 import sharding


def test_lpt_assign_balances_and_is_deterministic():
    rng = np.random.default_rng(4001)
    frames = rng.integers(240, 401, size=4096).tolist()
    for world in (1, 2, 4, 8):
        a = sharding.lpt_assign(frames, world)
        assert sorted(i for r in a for i in r) == list(range(4096))
        loads = [sum(frames[i] for i in r) for r in a]
        assert max(loads) - min(loads) <= max(frames)  # LPT bound
        assert a == sharding.lpt_assign(frames, world)
    assert sharding.lpt_assign([], 2) == [[], []]
    assert sharding.lpt_assign([5], 2) == [[0], []]
    with pytest.raises(ValueError):
        sharding.lpt_assign([1], 0)


def test_bucket_by_length():
    frames = [3, 5, 3, 3, 5, 7]
    b = sharding.bucket_by_length([0, 1, 2, 3, 4, 5], frames, 2)
    assert b == [(7, [5]), (5, [1, 4]), (3, [0, 2]), (3, [3])]
    with pytest.raises(ValueError):
        sharding.bucket_by_length([0], frames, 0)


def test_assign_buckets_partitions_batches_and_balances():
    rng = np.random.default_rng(4001)
    frames = rng.integers(240, 401, size=4096).tolist()
    for world in (1, 2, 4, 8):
        plan = sharding.assign_buckets(frames, world, 64)
        assert plan == sharding.assign_buckets(frames, world, 64)
        assert sorted(i for r in plan for _, ids in r for i in ids) == list(range(4096))
        for r in plan:
            assert all(1 <= len(ids) <= 64 and all(frames[i] == t for i in ids) for t, ids in r)
            assert [t for t, _ in r] == sorted((t for t, _ in r), reverse=True)
        loads = [sum(t * len(ids) for t, ids in r) for r in plan]
        assert max(loads) <= 1.005 * (sum(frames) / world)        # within 0.5 % of perfect balance on this corpus
        # batches stay large: a rank sees every length it owns in one (or, above max_batch, few) batches
        assert sum(len(r) for r in plan) <= 161 + 4096 // 64 + world
    # equal lengths: whole batches of 64
    plan = sharding.assign_buckets([320] * 4096, 8, 64)
    assert all(len(r) == 8 and all(len(ids) == 64 for _, ids in r) for r in plan)
    assert sharding.assign_buckets([], 2, 4) == [[], []]
    with pytest.raises(ValueError):
        sharding.assign_buckets([1], 0, 4)
    with pytest.raises(ValueError):
        sharding.assign_buckets([1], 1, 0)


def _fake_wave(i, t):
    """Deterministic stand-in for the decoder output of utterance i (T = t)."""
    n = t * sharding.SAMPLES_PER_FRAME
    return (torch.arange(n, dtype=torch.float32) * 1e-3 + float(i)).reshape(1, 1, n)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _padded_wave(i, t_true, t_pad):
    """Row of a ragged batch: the utterance's waveform, then padding the receiver must never hand out."""
    w = torch.full((1, 1, t_pad * sharding.SAMPLES_PER_FRAME), -777.0)
    w[0, 0, :t_true * sharding.SAMPLES_PER_FRAME] = _fake_wave(i, t_true).reshape(-1)
    return w


def _worker(rank, world, port, frames, q, ragged=False):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        calls = []

        def decode_batch(t, ids):
            calls.append((t, list(ids)))
            if ragged:
                return torch.cat([_padded_wave(i, frames[i], t) for i in ids], 0)
            return torch.cat([_fake_wave(i, t) for i in ids], 0)

        out = sharding.decode_sharded(decode_batch, frames, max_batch=2, ragged=ragged, bucket=4)
        if rank == 0:
            ok = all(torch.equal(out[i].reshape(-1), _fake_wave(i, frames[i]).reshape(-1)) for i in range(len(frames)))
            q.put(("gather", ok, len(out)))
        else:
            assert out is None
        q.put(("calls", rank, calls))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_decode_sharded_world2_gloo():
    frames = [4, 2, 7, 2, 4, 3, 1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, frames, q)) for r in range(2)]
    for p in procs:
        p.start()
    msgs = [q.get(timeout=100) for _ in range(3)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    gather = [m for m in msgs if m[0] == "gather"][0]
    assert gather[1] is True and gather[2] == len(frames)
    calls = {m[1]: m[2] for m in msgs if m[0] == "calls"}
    plan = sharding.assign_buckets(frames, 2, 2)
    for r in (0, 1):
        assert calls[r] == [(t, list(ids)) for t, ids in plan[r]]   # every rank decodes exactly its batches, in order
        assert all(len(ids) <= 2 for _, ids in calls[r])
        assert all(frames[i] == t for t, ids in calls[r] for i in ids)


@pytest.mark.timeout(120)
def test_decode_sharded_ragged_world2_gloo():
    """Mixed-length batches (lengths= inside one launch): every rank decodes its padded batches, rank 0 receives the
    padded rows and hands out exactly each utterance's own samples."""
    frames = [4, 2, 7, 2, 4, 3, 1, 6, 5]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, frames, q, True)) for r in range(2)]
    for p in procs:
        p.start()
    msgs = [q.get(timeout=100) for _ in range(3)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    gather = [m for m in msgs if m[0] == "gather"][0]
    assert gather[1] is True and gather[2] == len(frames)
    calls = {m[1]: m[2] for m in msgs if m[0] == "calls"}
    plan = sharding.assign_ragged(frames, 2, 2, bucket=4)
    for r in (0, 1):
        assert calls[r] == [(t, list(ids)) for t, ids in plan[r]]
        assert all(t % 4 == 0 and all(frames[i] <= t for i in ids) for t, ids in calls[r])


def test_assign_ragged_balance_plans_and_padding():
    rng = np.random.default_rng(4001)
    frames = rng.integers(240, 401, size=4096).tolist()
    for world in (1, 2, 4, 8):
        plan = sharding.assign_ragged(frames, world, 64)
        assert plan == sharding.assign_ragged(frames, world, 64)
        assert sorted(i for r in plan for _, ids in r for i in ids) == list(range(4096))
        loads = [sum(frames[i] for _, ids in r for i in ids) for r in plan]
        assert max(loads) <= 1.002 * (sum(frames) / world)        # snake dealing: within 0.2 % of perfect balance
        for r in plan:
            assert all(len(ids) == 64 for _, ids in r)            # 4096 / world / 64: whole batches
            assert all(t % 32 == 0 and t - 32 < max(frames[i] for i in ids) <= t for t, ids in r)
            assert len({(len(ids), t) for t, ids in r}) <= 8      # launch plans per rank: (batch, padded length)
            spread = max(max(frames[i] for i in ids) - min(frames[i] for i in ids) for _, ids in r)
            assert spread <= 8 * world + 8                        # a batch holds neighbouring lengths
    assert sharding.assign_ragged([], 2, 4) == [[], []]
    assert sharding.assign_ragged([5, 9], 1, 8, bucket=4) == [[(12, [1, 0])]]
    with pytest.raises(ValueError):
        sharding.assign_ragged([1], 0, 4)


def test_decode_sharded_single_process():
    frames = [2, 3, 2]
    out = sharding.decode_sharded(lambda t, ids: torch.cat([_fake_wave(i, t) for i in ids], 0), frames, max_batch=8)
    assert [o.numel() for o in out] == [1200, 1800, 1200]
    assert torch.equal(out[1], _fake_wave(1, 3).reshape(-1))
