"""Host -> device input pipeline of the training step: the batch of step i+1 is copied from pinned host memory on a side stream
while step i computes, and the loss of a step is read back one step late, so neither PCIe direction sits on the step's critical
path.  Plumbing only (torch streams / events / pinned memory); replaces what Lightning's DataLoader(pin_memory=True) +
``batch.to(device, non_blocking=True)`` + ``self.log(loss)`` do around ``GraphBinsLM.training_step`` (``modules/GraphBinsLM.py:121-139``)."""
from __future__ import annotations

from typing import Any, Callable, Iterable, Iterator, Optional

import torch


def _map(obj: Any, fn: Callable[[torch.Tensor], torch.Tensor]) -> Any:
    if torch.is_tensor(obj):
        return fn(obj)
    if isinstance(obj, (list, tuple)):
        return type(obj)(_map(o, fn) for o in obj)
    if isinstance(obj, dict):
        return {k: _map(v, fn) for k, v in obj.items()}
    return obj


class DevicePrefetcher:
    """Iterate device copies of pinned host batches (any nesting of tensors / lists / tuples / dicts / None); the copy of the
    next batch is in flight on ``self.stream`` while the caller works on the current one."""

    def __init__(self, batches: Iterable[Any], device, depth: int = 1):
        self.it: Iterator[Any] = iter(batches)
        self.device = torch.device(device)
        self.stream = torch.cuda.Stream(self.device)
        self.queue = []
        for _ in range(max(1, depth)):
            self._enqueue()

    def _enqueue(self) -> None:
        try:
            host = next(self.it)
        except StopIteration:
            return
        with torch.cuda.stream(self.stream):
            dev = _map(host, lambda t: t.to(self.device, non_blocking=True))
            ev = torch.cuda.Event()
            ev.record(self.stream)
        self.queue.append((dev, ev))

    def __iter__(self):
        return self

    def __next__(self):
        if not self.queue:
            raise StopIteration
        dev, ev = self.queue.pop(0)
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(ev)
        _map(dev, lambda t: (t.record_stream(cur), t)[1])      # allocated on the copy stream, consumed on the current one
        self._enqueue()
        return dev


class LaggedScalars:
    """Device scalars -> host floats without stalling the stream: ``push`` starts an async copy into pinned memory and returns
    the values pushed ``lag`` calls earlier (None until then); ``flush`` returns everything still in flight."""

    def __init__(self, n: int, lag: int = 1):
        self.n, self.lag = n, lag
        self.slots = [(torch.empty(n, dtype=torch.float32).pin_memory(), torch.cuda.Event()) for _ in range(lag + 1)]
        self.pending = []
        self.i = 0

    def push(self, *scalars: torch.Tensor) -> Optional[list]:
        buf, ev = self.slots[self.i % len(self.slots)]
        self.i += 1
        for j, s in enumerate(scalars):
            buf[j:j + 1].copy_(s.detach().reshape(1), non_blocking=True)
        ev.record()
        self.pending.append((buf, ev))
        if len(self.pending) > self.lag:
            b, e = self.pending.pop(0)
            e.synchronize()
            return b.tolist()
        return None

    def flush(self) -> list:
        out = []
        for b, e in self.pending:
            e.synchronize()
            out.append(b.tolist())
        self.pending = []
        return out
