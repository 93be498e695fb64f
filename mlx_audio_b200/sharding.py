"""Utterance sharding across the GPUs of one box and the final waveform gather.

The decoder has no cross-utterance dependence (instance norm reduces over time only, istftnet.py:221-230; the
style is per utterance, :334-337), so the path is partitioned by utterance: every rank owns the full decoder
and a subset of the utterances; the ONLY exchange is the gather of the finished waveforms on rank 0.

Host logic only (torch.distributed plumbing); works with the ``nccl`` backend on GPUs and with ``gloo`` on CPU
tensors (the world-size-2 tests).  Nothing here computes audio.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

SAMPLES_PER_FRAME = 600


def lpt_assign(frames: Sequence[int], world: int) -> List[List[int]]:
    """Longest-processing-time-first assignment of utterances (cost = frame count T) to ``world`` ranks.
    Deterministic: ties broken by utterance index, then by rank.  Returns per-rank utterance indices sorted
    ascending (so that a rank's outputs keep the caller's relative order)."""
    if world < 1:
        raise ValueError("world must be >= 1")
    load = [0] * world
    out: List[List[int]] = [[] for _ in range(world)]
    for i in sorted(range(len(frames)), key=lambda i: (-int(frames[i]), i)):
        r = min(range(world), key=lambda r: (load[r], r))
        out[r].append(i)
        load[r] += int(frames[i])
    return [sorted(a) for a in out]


def bucket_by_length(indices: Sequence[int], frames: Sequence[int], max_batch: int) -> List[Tuple[int, List[int]]]:
    """Group a rank's utterances into same-T batches of at most ``max_batch`` (one launch plan / CUDA graph per
    (B, T) shape).  Returns [(T, [utterance indices])] ordered by decreasing T."""
    if max_batch < 1:
        raise ValueError("max_batch must be >= 1")
    by_t: Dict[int, List[int]] = {}
    for i in indices:
        by_t.setdefault(int(frames[i]), []).append(i)
    out = []
    for t in sorted(by_t, reverse=True):
        ids = by_t[t]
        for k in range(0, len(ids), max_batch):
            out.append((t, ids[k:k + max_batch]))
    return out


def gather_waveforms(local: Dict[int, torch.Tensor], frames: Sequence[int], assignment: List[List[int]],
                     dst: int = 0, group=None, packed: Optional[torch.Tensor] = None) -> Optional[List[torch.Tensor]]:
    """gather(v) of ragged waveforms to ``dst``.  ``local``: {utterance index: (600*T,) float32 tensor} for the
    utterances this rank decoded.  Every rank packs its waveforms into ONE flat buffer (order = its assignment
    list) and rank ``dst`` receives each peer's buffer with a single point-to-point transfer (sizes are known
    everywhere from ``frames`` + ``assignment``: no size exchange, no padding).  Returns the utterance-ordered
    list on ``dst``, None elsewhere."""
    rank = dist.get_rank(group)
    world = dist.get_world_size(group)
    mine = assignment[rank]
    dev = next(iter(local.values())).device if local else torch.device("cpu")
    sizes = [[int(frames[i]) * SAMPLES_PER_FRAME for i in assignment[r]] for r in range(world)]
    if packed is not None:      # the caller already holds its waveforms packed in assignment order
        flat = packed
        dev = packed.device
    else:
        flat = torch.cat([local[i].reshape(-1) for i in mine]) if mine else torch.empty(0, dtype=torch.float32, device=dev)
    if flat.numel() != sum(sizes[rank]):
        raise ValueError(f"rank {rank}: packed {flat.numel()} samples, expected {sum(sizes[rank])}")
    if rank != dst:
        if flat.numel():
            dist.send(flat, dst=dst, group=group)
        return None
    out: List[Optional[torch.Tensor]] = [None] * len(frames)
    for r in range(world):
        if r == dst:
            buf = flat
        else:
            buf = torch.empty(sum(sizes[r]), dtype=torch.float32, device=dev)
            if buf.numel():
                dist.recv(buf, src=r, group=group)
        off = 0
        for i, n in zip(assignment[r], sizes[r]):
            out[i] = buf[off:off + n]
            off += n
    return out  # type: ignore[return-value]


def assign_buckets(frames: Sequence[int], world: int, max_batch: int) -> List[List[Tuple[int, List[int]]]]:
    """Bucket first, then balance: utterances of equal length T are split into near-equal batches of at most
    ``max_batch`` and the BATCHES (cost = size * T) are assigned longest-processing-time-first to ``world`` ranks,
    so that a rank decodes few, large same-T batches instead of many small ones (a ragged corpus dealt utterance by
    utterance leaves every rank with a handful of utterances per length).  Deterministic on every rank.  Returns
    per rank [(T, [utterance indices])] ordered by decreasing T."""
    if world < 1:
        raise ValueError("world must be >= 1")
    if max_batch < 1:
        raise ValueError("max_batch must be >= 1")
    by_t: Dict[int, List[int]] = {}
    for i, t in enumerate(frames):
        by_t.setdefault(int(t), []).append(i)
    batches: List[Tuple[int, List[int]]] = []
    for t, ids in by_t.items():
        n_b = -(-len(ids) // max_batch)
        base, rem = divmod(len(ids), n_b)
        k = 0
        for j in range(n_b):
            n = base + (1 if j < rem else 0)
            batches.append((t, ids[k:k + n]))
            k += n
    load = [0] * world
    out: List[List[Tuple[int, List[int]]]] = [[] for _ in range(world)]
    for t, ids in sorted(batches, key=lambda b: (-len(b[1]) * b[0], -b[0], b[1][0])):
        r = min(range(world), key=lambda r: (load[r], r))
        out[r].append((t, ids))
        load[r] += len(ids) * t
    # refinement: whole batches leave the ranks up to one batch apart; move the tail of one batch from the most to the
    # least loaded rank (at most `world` moves), which brings the spread down to about one utterance
    for _ in range(world):
        hi = max(range(world), key=lambda r: (load[r], -r))
        lo = min(range(world), key=lambda r: (load[r], r))
        gap = load[hi] - load[lo]
        cand = [(k, b) for k, b in enumerate(out[hi]) if len(b[1]) > 1 and b[0] <= gap // 2]
        if not cand:
            break
        k, (t, ids) = max(cand, key=lambda kb: (len(kb[1][1]), kb[1][0]))
        n = min(len(ids) - 1, gap // (2 * t))
        out[hi][k] = (t, ids[:-n])
        out[lo].append((t, ids[-n:]))
        load[hi] -= n * t
        load[lo] += n * t
    return [sorted(a, key=lambda b: (-b[0], b[1][0])) for a in out]


def assign_ragged(frames: Sequence[int], world: int, max_batch: int, bucket: int = 32) -> List[List[Tuple[int, List[int]]]]:
    """Mixed-length batches for ``Decoder.__call__(..., lengths=)``: the utterances, sorted by length, are dealt to the
    ranks in snake order (every rank sees the same length distribution: near-equal load without moving single
    utterances afterwards), and each rank cuts its sorted share into batches of ``max_batch`` (the short remainder
    last).  A batch is padded to its longest utterance rounded up to a multiple of ``bucket`` frames, so that a corpus
    needs a handful of launch plans (one per (batch size, padded length)) instead of one per distinct length; the
    padding costs no convolution work (the kernels build their tile lists from the lengths).  Deterministic on every
    rank.  Returns per rank [(T_padded, [utterance indices])], longest first."""
    if world < 1 or max_batch < 1 or bucket < 1:
        raise ValueError("world, max_batch and bucket must be >= 1")
    order = sorted(range(len(frames)), key=lambda i: (-int(frames[i]), i))
    share: List[List[int]] = [[] for _ in range(world)]
    for k, i in enumerate(order):
        rnd, pos = divmod(k, world)
        share[pos if rnd % 2 == 0 else world - 1 - pos].append(i)
    out: List[List[Tuple[int, List[int]]]] = []
    for ids in share:
        batches = []
        for k in range(0, len(ids), max_batch):
            sel = ids[k:k + max_batch]
            t = max(int(frames[i]) for i in sel)
            batches.append((-(-t // bucket) * bucket, sel))
        out.append(batches)
    return out


def decode_sharded(decode_batch: Callable[[int, List[int]], torch.Tensor], frames: Sequence[int], max_batch: int = 64,
                   dst: int = 0, group=None, ragged: bool = False, bucket: int = 32) -> Optional[List[torch.Tensor]]:
    """Decode ``len(frames)`` utterances over all ranks.  ``decode_batch(T, ids)`` must return the (len(ids), 1,
    600*T) waveforms of utterances ``ids`` - on the GPU path it is ``Decoder.__call__`` on the rank's own inputs.
    ``ragged=False``: batches hold utterances of ONE length T (``assign_buckets``); ``ragged=True``: mixed lengths padded
    to T (``assign_ragged``; ``decode_batch`` passes ``lengths=`` and the waveform of utterance i is the first
    600*frames[i] samples of its row).  Each rank writes its batches straight into ONE flat buffer, which is what the
    gather sends (ragged: the padded rows travel, <= bucket/2 frames of padding per utterance on average).  Returns the
    utterance-ordered waveform list on ``dst`` (views of the received buffers), None elsewhere."""
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    plan = assign_ragged(frames, world, max_batch, bucket) if ragged else assign_buckets(frames, world, max_batch)
    assignment = [[i for _, ids in plan[r] for i in ids] for r in range(world)]
    # samples each utterance occupies in its rank's flat buffer (its row's padded length), and the samples that count
    pitch = [[t * SAMPLES_PER_FRAME for t, ids in plan[r] for _ in ids] for r in range(world)]
    flat: Optional[torch.Tensor] = None
    off = 0
    total = sum(pitch[rank])
    for t, ids in plan[rank]:
        y = decode_batch(t, ids)
        if y.shape[0] != len(ids) or y.shape[-1] != t * SAMPLES_PER_FRAME:
            raise ValueError(f"decode_batch returned {tuple(y.shape)} for {len(ids)} utterances of T={t}")
        if flat is None:
            flat = torch.empty(total, dtype=torch.float32, device=y.device)
        n = len(ids) * t * SAMPLES_PER_FRAME
        flat[off:off + n].copy_(y.reshape(-1))        # the decoder reuses its output buffer
        off += n
    if flat is None:   # this rank owns no utterances (fewer batches than ranks)
        on_gpu = dist.is_initialized() and dist.get_backend(group) == "nccl"
        flat = torch.empty(0, dtype=torch.float32, device=torch.device("cuda", torch.cuda.current_device()) if on_gpu else "cpu")

    def views(buf: torch.Tensor, r: int) -> Dict[int, torch.Tensor]:
        o, res = 0, {}
        for i, p in zip(assignment[r], pitch[r]):
            res[i] = buf[o:o + int(frames[i]) * SAMPLES_PER_FRAME]
            o += p
        return res

    if world == 1:
        local = views(flat, 0)
        return [local[i] for i in range(len(frames))]
    if rank != dst:
        if flat.numel():
            dist.send(flat, dst=dst, group=group)
        return None
    out: List[Optional[torch.Tensor]] = [None] * len(frames)
    for r in range(world):
        if r == dst:
            buf = flat
        else:
            buf = torch.empty(sum(pitch[r]), dtype=torch.float32, device=flat.device)
            if buf.numel():
                dist.recv(buf, src=r, group=group)
        for i, v in views(buf, r).items():
            out[i] = v
    return out  # type: ignore[return-value]


class OverlappedGather:
    """The path's one exchange - gather of a step's (B, 1, S) waveforms on rank ``dst`` - issued on a SIDE stream so
    that it overlaps the next forward ("issue per bucket so it overlaps compute").  The decoder reuses its
    output buffer, so each call first copies ``y`` into one of ``depth`` staging buffers on the caller's stream (15 us for
    49 MB); the gather of step i then runs under the forward of step i+1, and the staging buffer is reused once the
    gather that read it has finished (event, no host synchronisation).  ``wait()`` joins the side stream."""

    def __init__(self, shape, device, dst: int = 0, depth: int = 2, group=None):
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.dst, self.depth, self.group = dst, depth, group
        self.side = torch.cuda.Stream(device)
        self.stage = [torch.empty(shape, dtype=torch.float32, device=device) for _ in range(depth)]
        self.lists = ([[torch.empty(shape, dtype=torch.float32, device=device) for _ in range(self.world)]
                       for _ in range(depth)] if self.rank == dst else None)
        self.done: List[Optional[torch.cuda.Event]] = [None] * depth
        self.i = 0

    def __call__(self, y: torch.Tensor):
        k = self.i % self.depth
        self.i += 1
        cur = torch.cuda.current_stream(y.device)
        if self.done[k] is not None:
            cur.wait_event(self.done[k])                 # the gather that last read this staging buffer has finished
        self.stage[k].copy_(y, non_blocking=True)
        ready = cur.record_event()
        with torch.cuda.stream(self.side):
            self.side.wait_event(ready)
            dist.gather(self.stage[k], self.lists[k] if self.lists is not None else None, dst=self.dst, group=self.group)
            self.done[k] = self.side.record_event()
        return self.lists[k] if self.lists is not None else None

    def wait(self):
        torch.cuda.current_stream(self.stage[0].device).wait_stream(self.side)
