"""The closed forms behind the centralized edge-MLP kernels (csrc/central.cu, csrc/nnops.cu segmax_bn_*), checked in fp64 on the CPU
against torch autograd of the plain sequence  Linear -> BatchNorm (batch statistics over the E = N*k edge rows) -> LeakyReLU  (x2)
-> max over each point's k edges  (deltaconv/nn/deltaconv.py:50-52 with mlp_depth = 2):

  * first block: mean_c = w_c . m, var_c = w_c^T (S - m m^T) w_c from the moments m = E[d], S = E[d d^T] of the edge offsets
    (dc_central_stats); backward from the dense gradient of its activation with the three sums
        S1_c = sum dz, S2_c = sum dz*xhat, A_c = sum dz * d      (one pass over the E rows: central_dense_bwd_kernel)
        dW_c = gamma_c invstd_c ( A_c - S1_c m - S2_c invstd_c (S w_c - mean_c m) ),  dgamma = S2, dbeta = S1
    (central_bwd_finalize_kernel);
  * last block folded into the max: the gradient of the activation is non-zero at the arg-max edge only, so S1, S2 run over the N
    points, and  dy = gamma invstd (dz [edge is the arg-max] - S1/E - xhat S2/E)  for every edge row (segmax_bn_bwd_vec4_kernel).

The GPU tests compare the kernels with the plain sequence on the device; this file pins the algebra where no GPU exists."""
import torch

SLOPE = 0.2


def _setup(seed=0, N=40, k=7, C1=6, C2=5):
    g = torch.Generator().manual_seed(seed)
    pos = torch.randn(N, 3, generator=g, dtype=torch.float64)
    idx = torch.stack([torch.randperm(N, generator=g)[:k] for _ in range(N)])
    d = (pos[idx] - pos[:, None, :]).reshape(N * k, 3)
    W1 = torch.randn(C1, 3, generator=g, dtype=torch.float64, requires_grad=True)
    W2 = torch.randn(C2, C1, generator=g, dtype=torch.float64, requires_grad=True)
    bn = [(torch.randn(C, generator=g, dtype=torch.float64, requires_grad=True), torch.randn(C, generator=g, dtype=torch.float64, requires_grad=True))
          for C in (C1, C2)]
    return N, k, d, W1, W2, bn, g


def _bn_act(t, gamma, beta, eps=1e-5):
    mu, var = t.mean(0), t.var(0, unbiased=False)
    xh = (t - mu) / torch.sqrt(var + eps)
    return torch.nn.functional.leaky_relu(gamma * xh + beta, SLOPE), xh, mu, 1.0 / torch.sqrt(var + eps)


def test_first_block_statistics_from_the_moments_of_the_offsets():
    N, k, d, W1, _, _, _ = _setup()
    t = d @ W1.T
    m, S = d.mean(0), (d[:, :, None] * d[:, None, :]).mean(0)
    mean = W1 @ m
    var = torch.einsum('cq,qp,cp->c', W1, S - torch.outer(m, m), W1)
    assert torch.allclose(mean, t.mean(0), atol=1e-12) and torch.allclose(var, t.var(0, unbiased=False), atol=1e-12)


def test_first_block_backward_from_the_dense_gradient():
    N, k, d, W1, _, bn, g = _setup(1)
    gamma, beta = bn[0]
    h, xh, mu, invstd = _bn_act(d @ W1.T, gamma, beta)
    dh = torch.randn(h.shape, generator=g, dtype=torch.float64)
    h.backward(dh)
    z = gamma * xh + beta
    dz = dh * torch.where(z > 0, 1.0, SLOPE)
    S1, S2, A = dz.sum(0), (dz * xh).sum(0), dz.T @ d                       # [C], [C], [C,3]
    m, S = d.mean(0), (d[:, :, None] * d[:, None, :]).mean(0)
    Sw = W1.detach() @ S                                                     # row c = S w_c
    dW = (gamma * invstd)[:, None] * (A - S1[:, None] * m - (S2 * invstd)[:, None] * (Sw - mu[:, None] * m))
    assert torch.allclose(dW.detach(), W1.grad, atol=1e-10)
    assert torch.allclose(S2.detach(), gamma.grad, atol=1e-10) and torch.allclose(S1.detach(), beta.grad, atol=1e-10)


def test_last_block_folded_into_the_max():
    N, k, d, W1, W2, bn, g = _setup(2)
    h1, _, _, _ = _bn_act(d @ W1.T, *bn[0])
    y = (h1.detach() @ W2.detach().T).requires_grad_(True)                            # pre-activation of the last block, [E, C2]
    gamma, beta = bn[1]
    a, xh, mu, invstd = _bn_act(y, gamma, beta)
    out, arg = a.reshape(N, k, -1).max(1)
    dout = torch.randn(out.shape, generator=g, dtype=torch.float64)
    out.backward(dout)
    # forward: the maximum of the activation is the activation of the extreme of (+-)y — increasing for gamma > 0, decreasing otherwise
    y3 = y.detach().reshape(N, k, -1)
    sel = torch.where(gamma.detach() >= 0, y3.max(1).values, y3.min(1).values)
    z_sel = gamma.detach() * (sel - mu.detach()) * invstd.detach() + beta.detach()
    assert torch.allclose(torch.nn.functional.leaky_relu(z_sel, SLOPE), out.detach(), atol=1e-12)
    # backward: sums over the N arg-max rows, then every edge row
    E = N * k
    xh3 = xh.detach().reshape(N, k, -1)
    xh_arg = torch.gather(xh3, 1, arg[:, None, :]).squeeze(1)                # [N, C2]
    z_arg = gamma.detach() * xh_arg + beta.detach()
    dz_arg = dout * torch.where(z_arg > 0, 1.0, SLOPE)
    S1, S2 = dz_arg.sum(0), (dz_arg * xh_arg).sum(0)
    hit = torch.zeros(N, k, out.size(1), dtype=torch.float64).scatter_(1, arg[:, None, :], dz_arg[:, None, :])
    dy = (gamma.detach() * invstd.detach()) * (hit - S1 / E - xh3 * (S2 / E))
    assert torch.allclose(dy.reshape(E, -1), y.grad, atol=1e-10)
    assert torch.allclose(S2, gamma.grad, atol=1e-10) and torch.allclose(S1, beta.grad, atol=1e-10)


def test_head_linear_without_the_broadcast():
    """ops.linear_bcast_bn_act's algebra (deltanet_segmentation.py:61-66): cat([g[batch], x], 1) @ W.T = x @ W_x.T + (g @ W_g.T)[batch]
    with W = [W_g | W_x]; the gradients of g and of W_g come from the per-cloud sums of dy."""
    gen = torch.Generator().manual_seed(3)
    N, B, Kg, Kx, C = 50, 3, 7, 4, 6
    batch = torch.sort(torch.randint(0, B, (N,), generator=gen)).values
    x = torch.randn(N, Kx, generator=gen, dtype=torch.float64, requires_grad=True)
    g = torch.randn(B, Kg, generator=gen, dtype=torch.float64, requires_grad=True)
    W = torch.randn(C, Kg + Kx, generator=gen, dtype=torch.float64, requires_grad=True)
    y = torch.cat([g[batch], x], 1) @ W.T
    dy = torch.randn(N, C, generator=gen, dtype=torch.float64)
    y.backward(dy)
    Wg, Wx = W.detach()[:, :Kg], W.detach()[:, Kg:]
    assert torch.allclose(x.detach() @ Wx.T + (g.detach() @ Wg.T)[batch], y.detach(), atol=1e-12)
    dyg = torch.zeros(B, C, dtype=torch.float64).index_add_(0, batch, dy)
    assert torch.allclose(dy @ Wx, x.grad, atol=1e-12) and torch.allclose(dyg @ Wg, g.grad, atol=1e-12)
    assert torch.allclose(torch.cat([dyg.T @ g.detach(), dy.T @ x.detach()], 1), W.grad, atol=1e-12)
