"""Data-parallel gradient exchange: one process per GPU, batch sharded, NCCL all-reduce of the flat
gradient buffer overlapped with the rest of backward.

Reference idiom: ``cleanrl/ppo_atari_multigpu.py:345-359`` flattens all grads, ``all_reduce(SUM)``, divides by world size and
copies back, after backward and on gloo by default.  Here the gradients already live in one flat buffer, ``1/world_size`` is
folded into the loss gradient (``grad_scale``), and every finished backward stage is reduced on a side stream while earlier
layers are still computing.  Works on any ``torch.distributed`` backend (nccl on GPUs, gloo in the CPU tests).
"""
from __future__ import annotations

import torch
import torch.distributed as dist


class GradAllReducer:
    """Coalesces finished gradient ranges of one flat buffer into buckets and all-reduces them on a communication stream."""

    def __init__(self, bucket_bytes: int = 2 << 20, group=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.bucket_elems = max(1, bucket_bytes // 4)
        self._pending = {}  # id(flat) -> [flat, lo, hi]
        self._comm = None
        self.n_collectives = 0

    def _stream(self, device):
        if device.type != "cuda":
            return None
        if self._comm is None:
            self._comm = torch.cuda.Stream(device=device)
        return self._comm

    def _flush(self, flat, lo, hi, wait=None):
        if self.world == 1 or hi <= lo:
            return
        chunk = flat[lo:hi]
        comm = self._stream(flat.device)
        self.n_collectives += 1
        if comm is None:
            dist.all_reduce(chunk, op=dist.ReduceOp.SUM, group=self.group)
            return
        if wait is None:
            comm.wait_stream(torch.cuda.current_stream(flat.device))  # the stage's kernels are enqueued before this point
        else:
            wait(comm)  # a narrower dependency than "everything enqueued so far" (e.g. the encoder's FC-gradient event)
        with torch.cuda.stream(comm):
            dist.all_reduce(chunk, op=dist.ReduceOp.SUM, group=self.group)

    def flush(self, flat: torch.Tensor = None):
        """Send what is pending (of ``flat``, or of every buffer) now instead of waiting for a full bucket."""
        for key, cur in list(self._pending.items()):
            if flat is None or cur[0] is flat:
                self._flush(*cur)
                del self._pending[key]

    def ready_now(self, flat: torch.Tensor, begin: int, end: int, wait=None):
        """All-reduce flat[begin:end] immediately; ``wait(comm_stream)`` replaces the default dependency on the current stream."""
        self._flush(flat, begin, end, wait)

    def callback_for(self, model):
        """``on_stage_done`` callback for ``model.loss_and_backward`` (DecisionTransformer or MultimodalDecisionTransformer).

        Trunk stages are coalesced into buckets and the last one flushes the rest, so the trunk's gradients travel while the towers'
        backward runs; the encoder's FC range goes out as soon as the library's FC event fires (while the conv layers still run); only
        the conv range (a few hundred KB) is exchanged after the last kernel."""
        dt = getattr(model, "dt", model)
        enc = getattr(model, "encoder", None)
        last = dt.n_stages - 1

        def cb(stage, begin, end):
            if stage == "encoder_fc":
                wait = enc.wait_fc_grads if enc.flat_grads.is_cuda else None
                self.ready_now(enc.flat_grads, begin, end, wait)
            elif stage == "encoder":
                self.ready_now(enc.flat_grads, begin, end)
            else:
                self.ready(dt.flat_grads, begin, end)
                if stage == last:
                    self.flush(dt.flat_grads)

        return cb

    def ready(self, flat: torch.Tensor, begin: int, end: int):
        """Mark flat[begin:end] as final for this step (its producing kernels are already enqueued on the current stream)."""
        key = id(flat)
        cur = self._pending.get(key)
        if cur is None:
            self._pending[key] = [flat, begin, end]
        elif end == cur[1]:
            cur[1] = begin
        elif begin == cur[2]:
            cur[2] = end
        else:  # not adjacent: send what we have, start a new range
            self._flush(*cur)
            self._pending[key] = [flat, begin, end]
        cur = self._pending[key]
        if cur[2] - cur[1] >= self.bucket_elems:
            self._flush(*cur)
            del self._pending[key]

    def finish(self):
        """Flush the remaining ranges and make the current stream wait for every collective of this step."""
        for cur in list(self._pending.values()):
            self._flush(*cur)
        self._pending.clear()
        if self._comm is not None:
            torch.cuda.current_stream(self._comm.device).wait_stream(self._comm)


def broadcast_parameters(flat_params, src: int = 0, group=None):
    """Identical initial replicas (the reference relies on a shared torch seed, ppo_atari_multigpu.py:197-200; a broadcast is stricter)."""
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        for f in flat_params:
            dist.broadcast(f, src=src, group=group)


def replicas_consistent(flat_params, group=None) -> bool:
    """Replica-consistency check the reference only prints (ppo_atari_multigpu.py:269-271): bitwise-equal parameter checksums."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return True
    ok = True
    for f in flat_params:
        s = f.view(torch.int32).to(torch.int64).sum().reshape(1)
        lo, hi = s.clone(), s.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=group)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=group)
        ok = ok and bool((lo == hi).item())
    return ok


def shard_batch(tensors, rank: int, world: int):
    """Rank r takes windows [r*B_local, (r+1)*B_local) of the global batch (ppo_atari_multigpu.py:159-161)."""
    out = []
    for t in tensors:
        B = t.shape[0]
        assert B % world == 0, f"global batch {B} not divisible by world size {world}"
        b = B // world
        out.append(t[rank * b:(rank + 1) * b])
    return out
