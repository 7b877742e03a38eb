"""Double-buffered host -> device feeder for the training step (the caller side of ``train_step``: upstream's ``get_batch`` hands
numpy/CPU batches to the model every step).

While step *i* runs on the compute stream, step *i+1*'s batch is copied from pinned host memory on a copy stream into the other
device slot, so the PCIe transfer (c2: 55 MB/step with uint8 frames) hides behind the kernels.  Each slot owns its device tensors
and, optionally, a CUDA graph of the step captured over them.  ``step()`` returns the step's loss as a Python float (8-byte
device -> host read + stream sync every step), exactly what ``SequenceTrainer.train_step()`` returns.
"""
from __future__ import annotations

from typing import Callable, Dict

import torch


class HostBatchPipeline:
    def __init__(self, make_step: Callable[[Dict[str, torch.Tensor]], Callable[[], torch.Tensor]], example: Dict[str, torch.Tensor],
                 device: torch.device, use_graph: bool = True, warmup: int = 3, deferred_loss: bool = False,
                 indexed: Dict[str, tuple] | None = None):
        """make_step(device_batch) -> step() returning a device tensor whose element 0 is the loss; `example` fixes shapes/dtypes.
        indexed: {name: (index_key, table)} — input `name` of the step is a gather ``table[index]`` of a small device-resident table
        (table [n, *row_shape] on `device`, index int64 with the leading dims of the input): host batches then carry `index_key` instead
        of `name`, only the indices cross PCIe, and the rows are expanded into the slot's device tensor on the copy stream (lossless;
        the step sees the same tensor).  The reference's Minecraft2d audio channel is such an input: every timestep's spectrogram is one
        of three precomputed MFCC images (REF environments/Minecraft/Minecraft.py:145-159, 222-225).
        deferred_loss: ``step()`` returns the loss of the PREVIOUS step (nan on the first call; ``flush()`` returns the last one): every
        step's loss is still copied to the host, but the host no longer waits for step i before enqueueing step i + 1, so the GPU never
        idles for a host round trip.  Default False = upstream ``train_step()`` semantics (the step's own loss, one sync per step)."""
        self.device = device
        self.copy_stream = torch.cuda.Stream(device=device)
        self.slots = []
        self.indexed = dict(indexed or {})
        self._index_of = {ik: (name, table) for name, (ik, table) in self.indexed.items()}
        for name, (ik, table) in self.indexed.items():
            if name not in example or table.device != torch.device(device) or tuple(example[name].shape[-(table.dim() - 1):]) != tuple(table.shape[1:]):
                raise ValueError(f"indexed input {name!r}: table must live on {device} and its rows must match the input's trailing dims")
        for _ in range(2):
            dev = {k: torch.empty(v.shape, dtype=v.dtype, device=device) for k, v in example.items()}
            idx = {ik: torch.empty(example[name].shape[:example[name].dim() - (table.dim() - 1)], dtype=torch.long, device=device)
                   for name, (ik, table) in self.indexed.items()}
            self.slots.append(dict(dev=dev, idx=idx, step=make_step(dev), graph=None, out=None, copied=torch.cuda.Event(), consumed=torch.cuda.Event()))
        self.loss_host = torch.empty(2, dtype=torch.float32).pin_memory()
        self.deferred_loss = bool(deferred_loss)
        self._loss_hosts = [torch.empty(2, dtype=torch.float32).pin_memory() for _ in range(2)]
        self._loss_events = [torch.cuda.Event() for _ in range(2)]
        self._pending, self._parity = None, 0
        # bytes that cross PCIe per batch: the full tensors, except indexed inputs (8 bytes per gathered row)
        self.bytes_per_batch = sum(v.numel() * v.element_size() for k, v in example.items() if k not in self.indexed) + \
            sum(8 * s_idx.numel() for s_idx in self.slots[0]["idx"].values())
        self._cur, self._filled = 0, False
        self._copy_first = self.bytes_per_batch > (32 << 20)
        # warm-up on real data, then capture one graph per slot
        for s in self.slots:
            for k, v in example.items():
                s["dev"][k].copy_(v, non_blocking=True)
            for _ in range(warmup):
                s["out"] = s["step"]()
        torch.cuda.synchronize(device)
        if use_graph:
            for s in self.slots:
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    s["out"] = s["step"]()
                s["graph"] = g
            torch.cuda.synchronize(device)
        for s in self.slots:
            s["consumed"].record(torch.cuda.current_stream(device))

    def prefetch(self, host_batch: Dict[str, torch.Tensor]):
        """Enqueue the H2D copy of the NEXT batch (pinned host tensors) into the idle slot on the copy stream."""
        s = self.slots[self._cur if not self._filled else 1 - self._cur]
        self.copy_stream.wait_event(s["consumed"])  # the step that last read this slot has finished
        with torch.cuda.stream(self.copy_stream):
            for k, v in host_batch.items():
                if k in self._index_of:  # indices cross PCIe, the rows are gathered from the resident table on the copy stream
                    name, table = self._index_of[k]
                    s["idx"][k].copy_(v, non_blocking=True)
                    torch.index_select(table, 0, s["idx"][k].view(-1), out=s["dev"][name].view(-1, *table.shape[1:]))
                else:
                    s["dev"][k].copy_(v, non_blocking=True)
            s["copied"].record(self.copy_stream)
        self._filled = True

    def step(self, next_host_batch: Dict[str, torch.Tensor] | None = None) -> float:
        """Run the step on the batch prefetched last; meanwhile copy `next_host_batch` into the other slot."""
        if not self._filled:
            raise RuntimeError("HostBatchPipeline.step() before prefetch()")
        s = self.slots[self._cur]
        main = torch.cuda.current_stream(self.device)
        main.wait_event(s["copied"])
        # Large batches (the copy takes about as long as the step): start the next copy first, it is the longer pole.  Small ones (indices,
        # uint8 frames): launch the step first — the host time of enqueueing the copies (tens of microseconds) would otherwise sit between
        # the previous step's sync and this step's first kernel.
        if next_host_batch is not None and self._copy_first:
            self.prefetch(next_host_batch)  # overlaps with the kernels enqueued next
        if s["graph"] is not None:
            s["graph"].replay()
            out = s["out"]
        else:
            out = s["step"]()
        s["consumed"].record(main)
        if next_host_batch is not None and not self._copy_first:
            self.prefetch(next_host_batch)  # overlaps with the kernels just enqueued
        if next_host_batch is not None:
            self._cur = 1 - self._cur
        else:
            self._filled = False
        if not self.deferred_loss:
            self.loss_host.copy_(out[:2] if out.numel() >= 2 else out, non_blocking=True)
            main.synchronize()
            return float(self.loss_host[0])
        buf, ev = self._loss_hosts[self._parity], self._loss_events[self._parity]
        buf.copy_(out[:2] if out.numel() >= 2 else out, non_blocking=True)
        ev.record(main)
        prev, self._pending, self._parity = self._pending, (buf, ev), 1 - self._parity
        if prev is None:
            return float("nan")
        prev[1].synchronize()  # the PREVIOUS step's loss: normally long finished, while this step is already running
        return float(prev[0][0])

    def flush(self) -> float:
        """deferred_loss mode: wait for and return the loss of the last step enqueued."""
        if self._pending is None:
            return float("nan")
        buf, ev = self._pending
        self._pending = None
        ev.synchronize()
        return float(buf[0])
