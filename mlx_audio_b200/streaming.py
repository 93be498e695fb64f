"""Streaming driver for long-form synthesis: batches of HOST inputs go through a
``Decoder`` with the host->device copy of batch i+1 and the device->host copy of batch i-1 overlapping the forward
of batch i (three CUDA streams, ``depth`` staging slots, events between them).  The reference has no counterpart - it
decodes one chunk per call and materialises each result before starting the next (kokoro.py:135-168, pipeline.py:381-402).

Host logic only: every byte of audio is computed by ``Decoder.__call__``.
"""
from __future__ import annotations

import math
from typing import Callable, Dict, Iterable, Iterator, List, Optional

import torch

_KEYS = ("asr", "F0_curve", "N", "s")


def split_batches(n: int, max_batch: int) -> List[int]:
    """Batch sizes for ``n`` equal-length chunks: full batches of ``max_batch`` with the remainder FIRST, so that the
    pipeline's un-overlapped head (the first host->device copy + the first forward's ramp) is the short batch and every
    later copy hides under a full forward."""
    if n <= 0:
        return []
    full, rem = divmod(n, max_batch)
    return ([rem] if rem else []) + [max_batch] * full


class StreamingDecoder:
    """``for wav in StreamingDecoder(dec).run(batches): ...`` - ``batches`` yields dicts with (preferably pinned) host
    tensors ``asr (B,512,T), F0_curve (B,2T), N (B,2T), s (B,128)``; ``wav`` is a pinned host tensor (B, 1, 600T) that
    stays valid until ``depth`` further batches have been requested.  Shapes may differ from batch to batch."""

    def __init__(self, decoder, depth: int = 2, post: Optional[Callable[[torch.Tensor], None]] = None):
        if depth < 2:
            raise ValueError("depth must be >= 2 (one slot in flight, one being filled)")
        self.dec, self.depth, self.post = decoder, depth, post
        self.dev = torch.device(decoder.device)
        self.h2d = torch.cuda.Stream(self.dev)
        self.d2h = torch.cuda.Stream(self.dev)
        self.slots: List[Dict[str, object]] = [dict(shape=None) for _ in range(depth)]

    def _slot(self, k: int, shapes) -> Dict[str, object]:
        """Staging buffers of slot ``k`` viewed at ``shapes``.  The backing stores only ever grow (pinned-host and device
        allocations cost milliseconds: a stream whose batch shapes vary must not re-allocate per batch)."""
        sl = self.slots[k]
        if sl["shape"] != shapes:
            B, T = shapes[0][0], shapes[0][2]
            need = {n: math.prod(sh) for n, sh in zip(_KEYS, shapes)}
            need["out"] = B * 600 * T
            store = sl.setdefault("store", {})
            for n, cnt in need.items():
                if n not in store or store[n].numel() < cnt:
                    store[n] = torch.empty(cnt, dtype=torch.float32, device=self.dev)
                    if n == "out":
                        store["out_host"] = torch.empty(cnt, dtype=torch.float32).pin_memory()
            sl["in"] = {n: store[n][:need[n]].view(sh) for n, sh in zip(_KEYS, shapes)}
            sl["out_dev"] = store["out"][:need["out"]].view(B, 1, 600 * T)
            sl["out_host"] = store["out_host"][:need["out"]].view(B, 1, 600 * T)
            sl["shape"] = shapes
        return sl

    def run(self, batches: Iterable[Dict[str, torch.Tensor]]) -> Iterator[torch.Tensor]:
        compute = torch.cuda.current_stream(self.dev)
        pending: List[Dict[str, object]] = []
        for i, b in enumerate(batches):
            host = {n: torch.as_tensor(b[n]) for n in _KEYS}
            sl = self._slot(i % self.depth, tuple(tuple(host[n].shape) for n in _KEYS))
            with torch.cuda.stream(self.h2d):
                if "in_free" in sl:
                    self.h2d.wait_event(sl["in_free"])      # the forward that read this slot's inputs has finished
                for n in _KEYS:
                    sl["in"][n].copy_(host[n], non_blocking=True)
                sl["in_ready"] = self.h2d.record_event()
            compute.wait_event(sl["in_ready"])
            x = sl["in"]
            y = self.dec(x["asr"], x["F0_curve"], x["N"], x["s"], borrow=True)   # copied out below
            if self.post is not None:
                self.post(y)
            if "out_done" in sl:
                compute.wait_event(sl["out_done"])          # the previous user of this slot has left the device
            sl["out_dev"].copy_(y, non_blocking=True)       # the decoder reuses its output buffer
            sl["out_ready"] = compute.record_event()
            sl["in_free"] = sl["out_ready"]
            with torch.cuda.stream(self.d2h):
                self.d2h.wait_event(sl["out_ready"])
                sl["out_host"].copy_(sl["out_dev"], non_blocking=True)
                sl["out_done"] = self.d2h.record_event()
            pending.append(sl)
            if len(pending) >= self.depth:                  # hand out the oldest result once the pipe is full
                done = pending.pop(0)
                done["out_done"].synchronize()
                yield done["out_host"]
        for done in pending:
            done["out_done"].synchronize()
            yield done["out_host"]
