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


def test_indexed_input_equals_full_tensor_input(cuda_device):
    """An input drawn from a small table (the reference's Minecraft2d audio channel: one of three MFCC images per timestep) may travel as
    its indices: the rows are gathered on the copy stream and the step sees the same tensor, so the losses are bit-identical and only
    8 bytes per row cross PCIe; a batch that carries the full tensor instead still works on the same pipeline."""
    from mtrl_b200.pipeline import HostBatchPipeline
    from util import make_inputs

    dev = cuda_device
    _, model, batches, make_step = (None, *_setup(dev))
    K, sd = batches[0]["states"].shape[1], batches[0]["states"].shape[2]
    g = torch.Generator().manual_seed(3)
    table = torch.randn(3, sd, generator=g)
    full, feed = [], []
    for b in batches:
        idx = torch.randint(0, 3, b["states"].shape[:2], generator=g)
        fb = dict(b, states=table[idx].contiguous().pin_memory())
        full.append(fb)
        feed.append({**{k: v for k, v in fb.items() if k != "states"}, "state_index": idx.pin_memory()})
    direct = [float(make_step({k: v.to(dev) for k, v in b.items()})()[0]) for b in full]
    pipe = HostBatchPipeline(make_step, full[0], dev, indexed={"states": ("state_index", table.to(dev))})
    assert pipe.bytes_per_batch == sum(v.numel() * v.element_size() for k, v in full[0].items() if k != "states") + 8 * full[0]["states"].shape[0] * K
    pipe.prefetch(feed[0])
    got = [pipe.step(feed[i + 1] if i + 1 < len(feed) else None) for i in range(len(feed))]
    assert got == direct
    pipe.prefetch(full[1])  # the full tensor is still accepted
    assert pipe.step(None) == direct[1]
    with pytest.raises(ValueError):
        HostBatchPipeline(make_step, full[0], dev, indexed={"states": ("state_index", table[:, :4].to(dev))})
