"""dtb_colsum_bf16_static: the allocation-free variant of dtb_colsum_bf16 (scratch in a file-scope static __device__ array inside
libdtb200.so).  It has no counterpart in the reference; the checker is the existing entry
point, and torch's fp32 column sum bounds both.  Collected last on purpose."""
import pytest
import torch

from mtrl_b200 import _lib

pytestmark = pytest.mark.gpu

STATIC_MAX_N = 1024  # DTB_COLSUM_STATIC_MAX_N in include/dtb200.h


def _case(dev, M, N, ld, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    full = torch.randn(M, ld, generator=g).to(dev).to(torch.bfloat16)
    return full[:, :N]


@pytest.mark.parametrize("M,N,ld", [(5760, 384, 384), (1920, 512, 1024), (37, 8, 8), (1, 128, 128), (30720, 32, 64), (4096, 1024, 1024)])
def test_static_colsum_matches_existing_entry_point_twice_in_a_row(cuda_device, M, N, ld):
    dev = cuda_device
    for call, seed in enumerate((11, 12)):  # two calls in a row: the second one finds the first one's partials in the static array
        x = _case(dev, M, N, ld, seed + M)
        want = torch.empty(N, device=dev)
        _lib.colsum_bf16(x, want)  # existing entry point, caller-provided scratch
        got = torch.full((N,), float("nan"), device=dev)
        before = torch.cuda.memory_allocated(dev)
        _lib.colsum_bf16_static(x, got)
        assert torch.cuda.memory_allocated(dev) == before
        torch.cuda.synchronize()
        assert torch.equal(got, want), (call, (got - want).abs().max().item())  # same kernels, same slab count: bit-identical
        ref = x.float().sum(0)
        assert (got - ref).abs().max().item() <= 2e-3 * max(1.0, ref.abs().max().item()) + 1e-3 * M ** 0.5
        # accumulate = 1 adds onto what is there
        _lib.colsum_bf16(x, want, accumulate=True)
        _lib.colsum_bf16_static(x, got, accumulate=True)
        torch.cuda.synchronize()
        assert torch.equal(got, want)


def test_static_colsum_smaller_call_after_larger_one(cuda_device):
    """The array keeps the last call's contents; a following smaller problem must not read stale partials."""
    dev = cuda_device
    big = _case(dev, 8192, 1024, 1024, 5)
    small = _case(dev, 100, 64, 64, 6)
    out_big = torch.empty(1024, device=dev)
    _lib.colsum_bf16_static(big, out_big)
    got, want = torch.empty(64, device=dev), torch.empty(64, device=dev)
    _lib.colsum_bf16_static(small, got)
    _lib.colsum_bf16(small, want)
    torch.cuda.synchronize()
    assert torch.equal(got, want)


def test_static_colsum_rejects_sizes_beyond_the_limit(cuda_device):
    dev = cuda_device
    x = torch.zeros(16, STATIC_MAX_N + 8, device=dev, dtype=torch.bfloat16)
    out = torch.zeros(STATIC_MAX_N + 8, device=dev)
    with pytest.raises(_lib.DtbError, match="exceeds the static scratch limit"):
        _lib.colsum_bf16_static(x, out)
    with pytest.raises(_lib.DtbError):
        _lib.colsum_bf16_static(torch.zeros(16, 12, device=dev, dtype=torch.bfloat16), out)  # N not a multiple of 8
