"""GPU tests of the tensor-core (tcgen05, TF32 operands, fp32 accumulate) GEMM path.

Tolerance is stated separately from the fp32 path: TF32 keeps 10 mantissa bits, so results are
compared to the fp64 reference at rel 2e-3 (max|got-ref| / max|ref|); the exact-fp32 kernel stays
the parity path (rel 1e-4)."""
import numpy as np
import pytest
import torch

from gsc_b200 import ops

pytestmark = pytest.mark.gpu
DEV = "cuda"
TF32_TOL = 2e-3


def rel_err(got, ref):
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    return float(np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-30))


def _padded(t, ld):
    """Copy of ``t`` [r, c] with row stride ``ld`` floats (TMA needs strides % 4 floats == 0)."""
    buf = torch.zeros(t.shape[0], ld, dtype=torch.float32, device=DEV)
    buf[:, :t.shape[1]] = t
    return buf[:, :t.shape[1]]


@pytest.mark.parametrize("M,N,Ks", [(512, 128, (64,)), (300, 256, (256,)), (4096, 1728, (256,)),
                                    (100, 64, (24, 1728, 1728)), (1000, 256, (24, 30000)),
                                    (100, 30000, (256,)), (257, 136, (40,))])
def test_tf32_linear_fwd(M, N, Ks):
    gen = torch.Generator().manual_seed(M + N)
    xs = [torch.randn(M, K, generator=gen, dtype=torch.float64) for K in Ks]
    W = torch.randn(N, sum(Ks), generator=gen, dtype=torch.float64) / np.sqrt(sum(Ks))
    b = torch.randn(N, generator=gen, dtype=torch.float64)
    mul = (torch.rand(M, N, generator=gen) > 0.3).double()
    ref = torch.relu(torch.cat(xs, 1) @ W.t() + b) * mul
    xd = [_padded(x.float().to(DEV), (x.shape[1] + 3) // 4 * 4) for x in xs]
    Wd = _padded(W.float().to(DEV), (W.shape[1] + 3) // 4 * 4)
    with ops.use_precision("tf32"):
        got = ops.linear_fwd(xd, Wd, b.float().to(DEV), relu=True, mul=mul.float().to(DEV))
    exact = ops.linear_fwd(xd, Wd, b.float().to(DEV), relu=True, mul=mul.float().to(DEV))
    e_tc, e_fp32 = rel_err(got.cpu().numpy(), ref.numpy()), rel_err(exact.cpu().numpy(), ref.numpy())
    print(f"fwd M={M} N={N} K={Ks}: tf32 err {e_tc:.2e}, fp32 err {e_fp32:.2e}")
    assert e_fp32 < 2e-5
    assert e_tc < TF32_TOL
    assert e_tc > 1e-6 or sum(Ks) < 64, "suspiciously exact: did the tensor-core path run?"


@pytest.mark.parametrize("M,N,K", [(256, 128, 128), (100, 64, 1728), (1024, 30000, 256), (333, 256, 4000)])
def test_tf32_dgrad(M, N, K):
    """dX[M,K] = dY[M,N] @ W[N,K]  (B operand MN-major)."""
    gen = torch.Generator().manual_seed(M + K)
    dY = torch.randn(M, N, generator=gen, dtype=torch.float64)
    W = torch.randn(N, K + 8, generator=gen, dtype=torch.float64)
    gate = torch.randn(M, K, generator=gen, dtype=torch.float64)
    ref = (dY @ W[:, 8:8 + K]) * (gate > 0)
    with ops.use_precision("tf32"):
        got = ops.linear_dgrad(dY.float().to(DEV), W.float().to(DEV), 8, K, gate=gate.float().to(DEV))
    e = rel_err(got.cpu().numpy(), ref.numpy())
    print(f"dgrad M={M} N={N} K={K}: tf32 err {e:.2e}")
    assert 1e-6 < e < TF32_TOL


@pytest.mark.parametrize("M,N,K", [(256, 128, 128), (100, 256, 30000), (1024, 64, 3456), (4096, 256, 1000)])
def test_tf32_wgrad(M, N, K):
    """dW[N,K] = dY[M,N]^T @ X[M,K]  (both operands MN-major, reduction over the batch)."""
    gen = torch.Generator().manual_seed(M + K + 1)
    dY = torch.randn(M, N, generator=gen, dtype=torch.float64)
    X = torch.randn(M, K, generator=gen, dtype=torch.float64)
    dW = torch.zeros(N, K + 8, device=DEV)
    with ops.use_precision("tf32"):
        ops.linear_wgrad(dY.float().to(DEV), [X.float().to(DEV)], dW[:, 4:4 + K])
    e = rel_err(dW[:, 4:4 + K].cpu().numpy(), (dY.t() @ X).numpy())
    print(f"wgrad M={M} N={N} K={K}: tf32 err {e:.2e}")
    assert 1e-6 < e < TF32_TOL
    assert float(dW[:, :4].abs().max()) == 0 and float(dW[:, 4 + K:].abs().max()) == 0


def test_tf32_falls_back_to_fp32_when_not_tma_mappable():
    gen = torch.Generator().manual_seed(1)
    x = torch.randn(256, 22, generator=gen).to(DEV)            # 88-byte rows: not a multiple of 16
    W = torch.randn(128, 22, generator=gen).to(DEV)
    with ops.use_precision("tf32"):
        got = ops.linear_fwd([x], W, None)
    assert torch.equal(got, ops.linear_fwd([x], W, None))


# ---------------------------------------------------------------- whole networks in TF32 mode
def test_tf32_networks_against_golden(golden_ddpg):
    """Forward, gradients and two fused DDPG steps with the tensor-core GEMMs, against the fp64
    golden fixtures at the TF32 tolerance; reports how many decisions differ from the fp32 oracle."""
    from gsc_b200 import ddpg
    from test_gpu_models import _batches, _gscale, _nets, grad_err
    g = golden_ddpg
    spec, (actor, qf1, target_actor, qf1_target) = _nets(g)
    cur, nxt = _batches(g)
    with ops.use_precision("tf32"):
        y = ddpg.predict(actor, cur)
        upd = ddpg.DDPGUpdater(actor, qf1, target_actor, qf1_target, lr=1e-3, gamma=0.99, tau=1e-2)
        actions = torch.from_numpy(g["actions"]).to(DEV)
        rewards, dones = torch.from_numpy(g["rewards"]).to(DEV), torch.from_numpy(g["dones"]).to(DEV)
        qv, ql, al = upd.update(cur, nxt, actions, rewards, dones)
    assert rel_err(y.cpu().numpy(), g["y_f64"]) < TF32_TOL
    flips = int(((y.cpu().numpy() > 0) != (g["y_f32"] > 0)).sum())
    print(f"tf32 actor forward err {rel_err(y.cpu().numpy(), g['y_f64']):.2e}; support-set flips {flips} of {y.numel()}")
    assert flips <= y.numel() // 1000
    names_q = [k for k, _ in qf1.named_parameters()]
    for k, v in zip(names_q, qf1.param_views(upd.g_critic)):
        assert grad_err(v.cpu().numpy(), g["gcritic_f64." + k], _gscale(g, "gcritic_f64.")) < 5 * TF32_TOL, k
    names_a = [k for k, _ in actor.named_parameters()]
    for k, v in zip(names_a, actor.param_views(upd.g_actor)):
        assert grad_err(v.cpu().numpy(), g["gactor_f64." + k], _gscale(g, "gactor_f64.")) < 5 * TF32_TOL, k
