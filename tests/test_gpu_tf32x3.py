"""GPU tests of the fp32-faithful tensor-core GEMM path (``ops.precision = "tf32x3"``, the default
and the mode bench.py times): tcgen05 kind::tf32 with every operand split x = hi + lo in shared
memory and hi*hi + hi*lo + lo*hi accumulated in fp32.

Tolerance: the same rel 2e-5 of an fp64 reference (``max|got-ref| / max|ref|``) that the exact FMA
kernel is held to, at the shapes of the c3 workload (5e-5 for reductions longer than 10 000 terms:
the tensor core adds its products into the fp32 accumulator in a different order and with a different
rounding than a chain of FMAs; observed <= 2.1e-5); decisions (support set ``y > 0`` and per-group
argmax, what the reference's consumers read) must equal the fp64 decisions wherever the fp64 value
is not within fp32 rounding of a tie."""
import numpy as np
import pytest
import torch

from gsc_b200 import ops

pytestmark = pytest.mark.gpu
DEV = "cuda"
TOL = 2e-5


def rel_err(got, ref):
    got, ref = got.double(), ref.double()
    return float((got - ref).abs().max() / ref.abs().max().clamp_min(1e-30))


def _pad4(t):
    ld = (t.shape[1] + 3) // 4 * 4
    buf = torch.zeros(t.shape[0], ld, dtype=torch.float32, device=DEV)
    buf[:, :t.shape[1]] = t
    return buf[:, :t.shape[1]]


@pytest.mark.parametrize("M,N,Ks", [(512, 128, (64,)), (300, 256, (256,)), (4096, 1728, (256,)),
                                    (100, 64, (24, 1728, 1728)), (1024, 64, (24, 30000)),
                                    (1024, 30000, (256,)), (257, 136, (40,))])
def test_x3_linear_fwd(M, N, Ks):
    gen = torch.Generator(device=DEV).manual_seed(M + N)
    xs = [torch.randn(M, K, generator=gen, device=DEV) for K in Ks]
    W = torch.randn(N, sum(Ks), generator=gen, device=DEV) / np.sqrt(sum(Ks))
    b = torch.randn(N, generator=gen, device=DEV)
    mul = (torch.rand(M, N, generator=gen, device=DEV) > 0.3).float()
    xd, Wd = [_pad4(x) for x in xs], _pad4(W)
    ref = torch.relu(torch.cat(xs, 1).double() @ W.double().t() + b.double()) * mul.double()
    with ops.use_precision("tf32x3"):
        got = ops.linear_fwd(xd, Wd, b, relu=True, mul=mul)
    with ops.use_precision("fp32"):
        exact = ops.linear_fwd(xd, Wd, b, relu=True, mul=mul)
    e3, e1 = rel_err(got, ref), rel_err(exact, ref)
    print(f"fwd M={M} N={N} K={Ks}: tf32x3 err {e3:.2e}, fp32 FMA err {e1:.2e}")
    assert e3 < TOL and e1 < TOL
    if M * N * sum(Ks) >= 4e6:      # smaller problems stay on the FMA kernel by design
        assert not torch.equal(got, exact), "bit-identical to the FMA kernel: did the tensor-core path run?"


@pytest.mark.parametrize("M,N,K", [(256, 128, 128), (100, 64, 1728), (1024, 64, 30000), (1024, 30000, 256), (333, 256, 4000)])
def test_x3_dgrad(M, N, K):
    """dX[M,K] = dY[M,N] @ W[N,K]  (B operand MN-major), gated."""
    gen = torch.Generator(device=DEV).manual_seed(M + K)
    dY = torch.randn(M, N, generator=gen, device=DEV)
    W = torch.randn(N, K + 8, generator=gen, device=DEV)
    gate = torch.randn(M, K, generator=gen, device=DEV)
    ref = (dY.double() @ W[:, 8:8 + K].double()) * (gate > 0)
    with ops.use_precision("tf32x3"):
        got = ops.linear_dgrad(dY, W, 8, K, gate=gate)
    e = rel_err(got, ref)
    print(f"dgrad M={M} N={N} K={K}: tf32x3 err {e:.2e}")
    assert e < (TOL if N <= 10000 else 5e-5)


@pytest.mark.parametrize("M,N,K", [(256, 128, 128), (100, 256, 30000), (1024, 64, 30000), (1024, 64, 3456), (4096, 256, 1000)])
def test_x3_wgrad(M, N, K):
    """dW[N,K] = dY[M,N]^T @ X[M,K]  (both operands MN-major, reduction over the batch)."""
    gen = torch.Generator(device=DEV).manual_seed(M + K + 1)
    dY = torch.randn(M, N, generator=gen, device=DEV)
    X = torch.randn(M, K, generator=gen, device=DEV)
    dW = torch.zeros(N, K + 8, device=DEV)
    with ops.use_precision("tf32x3"):
        ops.linear_wgrad(dY, [X], dW[:, 4:4 + K])
    e = rel_err(dW[:, 4:4 + K], dY.double().t() @ X.double())
    print(f"wgrad M={M} N={N} K={K}: tf32x3 err {e:.2e}")
    assert e < (TOL if M <= 10000 else 5e-5)
    assert float(dW[:, :4].abs().max()) == 0 and float(dW[:, 4 + K:].abs().max()) == 0


def test_x3_decisions_match_fp64_at_c3_size():
    """Actor layer 2 at the c3 shape (1024 states, A = 30 000): the support set and the per-group
    argmax over the N_max destinations equal the fp64 decisions except where fp64 itself is within
    fp32 rounding of a tie (|y| or the top-2 gap below 4e-6 of the output scale)."""
    M, K, n_max, groups = 1024, 256, 100, 3
    A = n_max * groups * n_max
    gen = torch.Generator(device=DEV).manual_seed(11)
    h = torch.relu(torch.randn(M, K, generator=gen, device=DEV))
    W = torch.randn(A, K, generator=gen, device=DEV) / 16
    b = torch.randn(A, generator=gen, device=DEV) / 4
    n = torch.full((M,), n_max, dtype=torch.int32, device=DEV)
    with ops.use_precision("tf32x3"):
        y = ops.linear_fwd([h], W, b, smask=(n, n_max, groups))
    margin = 4e-6
    flips = near_zero = 0
    worst = 0.0
    for r0 in range(0, M, 128):                      # fp64 reference in row blocks (memory)
        ref = h[r0:r0 + 128].double() @ W.double().t() + b.double()
        got = y[r0:r0 + 128]
        scale = float(ref.abs().max())
        worst = max(worst, float((got.double() - ref).abs().max()) / scale)
        diff = (got > 0) != (ref > 0)
        near = ref.abs() < margin * scale
        flips += int((diff & ~near).sum())
        near_zero += int((diff & near).sum())
        g_ref, g_got = ref.view(-1, n_max), got.reshape(-1, n_max)
        top2 = g_ref.topk(2, dim=1).values
        clear = (top2[:, 0] - top2[:, 1]) > margin * scale
        assert torch.equal(g_got.argmax(1)[clear], g_ref.argmax(1)[clear]), "per-group argmax differs from fp64"
    print(f"c3-size layer 2: err {worst:.2e}; support-set differences {flips} (+{near_zero} within rounding of zero) of {M * A}")
    assert worst < TOL
    assert flips == 0
