"""HostBatchPipeline (pinned host batches -> double-buffered H2D -> graph-captured step -> loss read-back): the losses it returns are those
of running the same steps directly on device-resident copies of the same batches, in the same order; deferred_loss=True returns the same
sequence one call later (nan first, flush() for the last)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(dev):
    from util import build_pair, make_inputs

    _, model, K = build_pair("c1", dev)
    batches = []
    for seed in range(5):
        s, a, r, t, m = make_inputs(8, K, 11, 3, 1000, seed=10 + seed)
        batches.append({"states": s.pin_memory(), "actions": a.pin_memory(), "rtg": r.pin_memory(), "timesteps": t.pin_memory(), "mask": m.pin_memory()})

    def make_step(d):
        return lambda: model.loss_and_backward(d["states"], d["actions"], d["rtg"], d["timesteps"], d["mask"])

    return model, batches, make_step


@pytest.mark.parametrize("use_graph", [True, False])
def test_pipeline_losses_match_direct_steps(cuda_device, use_graph):
    from mtrl_b200.pipeline import HostBatchPipeline

    dev = cuda_device
    model, batches, make_step = _setup(dev)
    direct = []
    for b in batches:
        d = {k: v.to(dev) for k, v in b.items()}
        direct.append(float(make_step(d)()[0]))
    pipe = HostBatchPipeline(make_step, batches[0], dev, use_graph=use_graph)
    pipe.prefetch(batches[0])
    got = [pipe.step(batches[i + 1] if i + 1 < len(batches) else None) for i in range(len(batches))]
    assert got == direct  # same kernels, same inputs, no optimizer step in between: bit-identical losses

    pipe = HostBatchPipeline(make_step, batches[0], dev, use_graph=use_graph, deferred_loss=True)
    pipe.prefetch(batches[0])
    lagged = [pipe.step(batches[i + 1] if i + 1 < len(batches) else None) for i in range(len(batches))]
    assert math.isnan(lagged[0]) and lagged[1:] == direct[:-1]
    assert pipe.flush() == direct[-1]
    assert math.isnan(pipe.flush())  # nothing pending any more
