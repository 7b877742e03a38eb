"""tcgen05 GEMM (dtb_gemm_bf16) vs a torch fp32 matmul of the same bf16-rounded operands.

Tolerance: operands are identical bf16 values, accumulation is fp32 on both sides, so the only
differences are summation order and the final bf16 rounding of the output: rel-err (L2) <= 2e-3
for fp32 outputs and <= 6e-3 for bf16 outputs.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def rel_err(a, b):
    a = a.float()
    b = b.float()
    return ((a - b).norm() / b.norm().clamp_min(1e-20)).item()


def _rand(shape, dev, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return torch.randn(shape, generator=g).to(dev).to(torch.bfloat16)


@pytest.mark.parametrize("M,N,K,bn", [(128, 128, 64, 0), (256, 256, 128, 256), (384, 384, 768, 128), (3840, 384, 128, 0),
                                     (1000, 96, 200, 32), (130, 64, 3136, 64), (4096, 2304, 768, 256)])
def test_gemm_kmajor_plain(cuda_device, M, N, K, bn):
    from mtrl_b200 import _lib

    A = _rand((M, K), cuda_device, 1)
    Bt = _rand((N, K), cuda_device, 2)  # B stored [N][K]
    out = torch.full((M, N), float("nan"), device=cuda_device, dtype=torch.float32)
    outb = torch.zeros((M, N), device=cuda_device, dtype=torch.bfloat16)
    _lib.gemm_bf16(A, Bt, M, N, K, out_f32=out, out_bf16=outb, bn=bn)
    torch.cuda.synchronize()
    ref = A.float() @ Bt.float().t()
    assert rel_err(out, ref) < 2e-3
    assert rel_err(outb, ref) < 6e-3


@pytest.mark.parametrize("bn", [0, 1000])  # 1000 = CTA-pair kernel with the staged (warp-coalesced) epilogue
@pytest.mark.parametrize("act", [0, 1, 2])
def test_gemm_epilogue_bias_act_residual(cuda_device, act, bn):
    from mtrl_b200 import _lib

    M, N, K = (640, 512, 256) if bn == 0 else (700, 520, 256)  # ragged M and N tails for the pair kernel
    A = _rand((M, K), cuda_device, 3)
    Bt = _rand((N, K), cuda_device, 4)
    bias = torch.randn(N, device=cuda_device)
    res = torch.randn(M, N, device=cuda_device)
    out = torch.empty((M, N), device=cuda_device, dtype=torch.float32)
    pre = torch.empty((M, N), device=cuda_device, dtype=torch.bfloat16)
    _lib.gemm_bf16(A, Bt, M, N, K, bias=bias, residual=res, act=act, out_f32=out, out_preact=pre, bn=bn)
    torch.cuda.synchronize()
    z = A.float() @ Bt.float().t() + bias
    y = z if act == 0 else (torch.relu(z) if act == 1 else torch.nn.functional.gelu(z, approximate="tanh"))
    assert rel_err(out, y + res) < 2e-3
    assert rel_err(pre, z) < 6e-3


@pytest.mark.parametrize("bn", [0, 1000])
@pytest.mark.parametrize("dact", [1, 2])
def test_gemm_dgrad_dact(cuda_device, dact, bn):
    """dX = (dY @ W^T) * act'(aux): B = W stored [K_out][N_red] is K-major."""
    from mtrl_b200 import _lib

    M, Nred, Kout = (512, 384, 256) if bn == 0 else (777, 384, 264)
    dY = _rand((M, Nred), cuda_device, 5)
    W = _rand((Kout, Nred), cuda_device, 6)
    aux = _rand((M, Kout), cuda_device, 7)
    out = torch.empty((M, Kout), device=cuda_device, dtype=torch.bfloat16)
    _lib.gemm_bf16(dY, W, M, Kout, Nred, aux=aux, dact=dact, out_bf16=out, bn=bn)
    torch.cuda.synchronize()
    g = dY.float() @ W.float().t()
    if dact == 1:
        ref = g * (aux.float() > 0)
    else:
        x = aux.float().requires_grad_(True)
        torch.nn.functional.gelu(x, approximate="tanh").backward(g)
        ref = x.grad
    assert rel_err(out, ref) < 6e-3


@pytest.mark.parametrize("Mtok,Kin,Nout,split", [(512, 128, 128, 1), (4096, 128, 384, 4), (3000, 256, 512, 3), (8192, 768, 256, 8),
                                                (640, 32, 64, 2)])
def test_gemm_wgrad_mn_major_splitk(cuda_device, Mtok, Kin, Nout, split):
    """dW[Kin, Nout] = X^T dY with X [Mtok, Kin], dY [Mtok, Nout] read in place (both MN-major)."""
    from mtrl_b200 import _lib

    X = _rand((Mtok, Kin), cuda_device, 8)
    dY = _rand((Mtok, Nout), cuda_device, 9)
    ref = X.float().t() @ dY.float()
    out = torch.ones((Kin, Nout), device=cuda_device, dtype=torch.float32)
    part = torch.empty((max(split, 1), Kin, Nout), device=cuda_device, dtype=torch.float32)
    if split == 1:
        _lib.gemm_bf16(X, dY, Kin, Nout, Mtok, a_mn_major=True, b_mn_major=True, out_f32=out)
        torch.cuda.synchronize()
        assert rel_err(out, ref) < 2e-3
    else:
        _lib.gemm_bf16(X, dY, Kin, Nout, Mtok, a_mn_major=True, b_mn_major=True, out_f32=out, partial=part, split_k=split,
                       accumulate=True)
        torch.cuda.synchronize()
        assert rel_err(out, ref + 1.0) < 2e-3


def test_gemm_mixed_major(cuda_device):
    """Forward with the Conv1D weight read in place: A K-major, B = W[K][N] MN-major."""
    from mtrl_b200 import _lib

    M, N, K = 768, 256, 192
    A = _rand((M, K), cuda_device, 10)
    W = _rand((K, N), cuda_device, 11)
    out = torch.empty((M, N), device=cuda_device, dtype=torch.float32)
    _lib.gemm_bf16(A, W, M, N, K, b_mn_major=True, out_f32=out)
    torch.cuda.synchronize()
    assert rel_err(out, A.float() @ W.float()) < 2e-3


def test_gemm_rejects_bad_args(cuda_device):
    from mtrl_b200 import _lib

    A = _rand((128, 64), cuda_device, 1)
    with pytest.raises(_lib.DtbError):
        _lib.gemm_bf16(A, A, 128, 60, 64, out_f32=torch.empty((128, 60), device=cuda_device))


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("M,N,K,split", [(512, 256, 128, 1), (1000, 768, 320, 1), (2048, 512, 1024, 4), (600, 384, 4096, 7)])
def test_gemm_cta_pair(cuda_device, a_mn, b_mn, M, N, K, split):
    """cta_group::2 kernel (bn=1000 forces it): 256 x 256 tiles over two SMs, every operand major, ragged M / N tails, split-K."""
    from mtrl_b200 import _lib

    A = _rand((K, M) if a_mn else (M, K), cuda_device, 21)
    B = _rand((K, N) if b_mn else (N, K), cuda_device, 22)
    ref = (A.float().t() if a_mn else A.float()) @ (B.float() if b_mn else B.float().t())
    out = torch.full((M, N), float("nan"), device=cuda_device, dtype=torch.float32)
    if split == 1:
        bias = torch.randn(N, device=cuda_device)
        outb = torch.zeros((M, N), device=cuda_device, dtype=torch.bfloat16)
        _lib.gemm_bf16(A, B, M, N, K, a_mn_major=a_mn, b_mn_major=b_mn, bias=bias, out_f32=out, out_bf16=outb, bn=1000)
        torch.cuda.synchronize()
        assert rel_err(out, ref + bias) < 2e-3
        assert rel_err(outb, ref + bias) < 6e-3
    else:
        part = torch.empty((split, M, N), device=cuda_device, dtype=torch.float32)
        _lib.gemm_bf16(A, B, M, N, K, a_mn_major=a_mn, b_mn_major=b_mn, out_f32=out, partial=part, split_k=split, bn=1000)
        torch.cuda.synchronize()
        assert rel_err(out, ref) < 2e-3


@pytest.mark.parametrize("M,N,K,b_mn", [(5760, 128, 128, True), (5760, 128, 512, True), (90, 128, 128, True), (9600, 256, 1024, True),
                                        (333, 64, 192, False), (1, 256, 256, True)])
def test_gemm_fused_layernorm_epilogue(cuda_device, M, N, K, b_mn):
    """c_proj + bias + residual + LayerNorm in one kernel (dtb_gemm_bf16_ln) vs addmm + layer_norm in fp32 on the same bf16 operands."""
    from mtrl_b200 import _lib

    dev = cuda_device
    g = torch.Generator().manual_seed(M + N + K)
    A = (torch.randn(M, K, generator=g) * 0.5).to(dev).to(torch.bfloat16)
    W = (torch.randn(K, N, generator=g) * 0.1).to(dev).to(torch.bfloat16)  # Conv1D layout [in, out]
    Wop = W if b_mn else W.t().contiguous()
    bias = torch.randn(N, generator=g).to(dev)
    res = (torch.randn(M, N, generator=g) * 2 + 0.3).to(dev)
    gamma, beta = (torch.randn(N, generator=g) * 0.5 + 1).to(dev), torch.randn(N, generator=g).to(dev)
    out = torch.full((M, N), float("nan"), device=dev)
    y = torch.full((M, N), float("nan"), device=dev, dtype=torch.bfloat16)
    mean, rstd = torch.empty(M, device=dev), torch.empty(M, device=dev)
    _lib.gemm_bf16_ln(A, Wop, M, N, K, gamma, beta, 1e-5, y, b_mn_major=b_mn, bias=bias, residual=res, out_f32=out, mean=mean, rstd=rstd)
    torch.cuda.synchronize()
    ref = A.float() @ W.float() + bias + res
    assert rel_err(out, ref) < 2e-3
    ref_y = torch.nn.functional.layer_norm(out, (N,), gamma, beta, 1e-5)  # LN of the kernel's own fp32 rows: isolates the fused part
    assert rel_err(y, ref_y) < 4e-3
    assert rel_err(mean, out.mean(1)) < 1e-5
    assert rel_err(rstd, 1.0 / torch.sqrt(out.var(1, unbiased=False) + 1e-5)) < 1e-5


@pytest.mark.parametrize("Mtok,Kin,Nout,split", [(5760, 128, 384, 4), (5760, 512, 128, 8), (1920, 1024, 128, 3), (38400, 768, 768, 2)])
def test_gemm_splitk_reduced_in_kernel_is_deterministic(cuda_device, Mtok, Kin, Nout, split):
    """Split-K weight gradient with the reduction done by the last-arriving CTA of every tile: same bits on every run (fixed slice
    order), equal to the fp32 reference, and the ticket counters return to zero (second call correct)."""
    from mtrl_b200 import _lib

    dev = cuda_device
    g = torch.Generator().manual_seed(Mtok + Kin)
    X = torch.randn(Mtok, Kin, generator=g).to(dev).to(torch.bfloat16)
    dY = torch.randn(Mtok, Nout, generator=g).to(dev).to(torch.bfloat16)
    ref = X.float().t() @ dY.float()
    partial = torch.empty(split, Kin, Nout, device=dev)
    outs = []
    _lib.set_option("splitk_inkernel", 1)
    try:
        for _ in range(3):
            out = torch.full((Kin, Nout), float("nan"), device=dev)
            _lib.gemm_bf16(X, dY, Kin, Nout, Mtok, a_mn_major=True, b_mn_major=True, out_f32=out, partial=partial, split_k=split)
            torch.cuda.synchronize()
            assert rel_err(out, ref) < 2e-3
            outs.append(out)
    finally:
        _lib.set_option("splitk_inkernel", 0)
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[1], outs[2])
    out = torch.full((Kin, Nout), float("nan"), device=dev)  # default: separate reduction launch (its own fixed summation order)
    _lib.gemm_bf16(X, dY, Kin, Nout, Mtok, a_mn_major=True, b_mn_major=True, out_f32=out, partial=partial, split_k=split)
    torch.cuda.synchronize()
    assert rel_err(out, outs[0]) < 1e-6
