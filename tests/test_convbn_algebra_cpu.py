"""The algebra of the folded 1x1 convolution + batch-norm (+ shortcut + ReLU) restated in float64 and
checked against autograd of the plain computation, on the CPU.  These are the formulas
csrc/kernels/convbn_aux.cu evaluates (forward statistics from the Gram matrix and column sums of the
INPUT; weight gradient as a closed-form correction of P = g^T X; input gradient as one GEMM over two
K sources plus a bias row) - the GPU tests check the kernels, this checks that the formulas are the
derivative.  (Workload: conv*_block*_3_conv/_3_bn/_add/_out of ResNet-50,
core/tests/examples/call_run_within_script_with_keras_fit.py:79-83.)"""
import pytest
import torch


def _plain(X, W, gamma, beta, R, eps):
    y = X @ W.t()
    mean = y.mean(0)
    var = y.var(0, unbiased=False)
    z = (y - mean) / torch.sqrt(var + eps) * gamma + beta
    if R is not None:
        z = z + R
    return torch.relu(z)


def _folded_forward(X, W, gamma, beta, R, eps):
    M = X.shape[0]
    S = X.sum(0)                                  # column sums of the input
    G = X.t() @ X                                 # Gram matrix of the input
    U = W @ G                                     # saved for the backward pass
    mean = (W @ S) / M                            # E[y_c] = w_c . S / M
    ey2 = (U * W).sum(1) / M                      # E[y_c^2] = w_c^T G w_c / M
    rstd = 1.0 / torch.sqrt(ey2 - mean * mean + eps)
    scale = gamma * rstd
    shift = beta - mean * scale
    z = (X @ W.t()) * scale + shift               # the fused GEMM's epilogue: y never stored
    if R is not None:
        z = z + R
    z = torch.relu(z)
    return z, (S, U, mean, rstd, z > 0)


def _folded_backward(dz, X, W, gamma, saved):
    S, U, mean, rstd, mask = saved
    M = X.shape[0]
    g = dz * mask                                 # masked upstream gradient (= gradient of the shortcut)
    P = g.t() @ X                                 # [N, K]: the one big reduction of the backward pass
    sum_g = g.sum(0)
    sum_gx = rstd * ((P * W).sum(1) - mean * sum_g)          # sum of g * xhat, without y
    a, b = sum_g / M, sum_gx / M
    sp = gamma * rstd
    t = sp * b * rstd
    dW = sp[:, None] * (P - a[:, None] * S[None, :] - (b * rstd)[:, None] * (U - mean[:, None] * S[None, :]))
    Ws = sp[:, None] * W
    Wt = -t[:, None] * W
    Qn = Wt.t() @ W                               # -(W^T diag(t) W)
    nbias = -((sp * a - t * mean)[:, None] * W).sum(0)
    dX = g @ Ws + X @ Qn + nbias                  # one GEMM over two K sources + a bias row
    return dX, dW, sum_gx, sum_g, g               # ..., dgamma, dbeta, dR


@pytest.mark.parametrize("M,K,N,residual", [(97, 16, 64, True), (64, 32, 128, False), (301, 8, 32, True)])
def test_folded_forward_and_backward_are_the_plain_computation_and_its_derivative(M, K, N, residual):
    torch.manual_seed(M + K)
    dd = torch.float64
    X = torch.randn(M, K, dtype=dd).relu()                  # the input of a 1x1 convolution follows a ReLU
    W = torch.randn(N, K, dtype=dd) / K ** 0.5
    gamma = torch.rand(N, dtype=dd) + 0.5
    beta = torch.randn(N, dtype=dd) * 0.3
    R = torch.randn(M, N, dtype=dd) if residual else None
    dz = torch.randn(M, N, dtype=dd)
    eps = 1e-3

    leaves = [t.clone().requires_grad_(True) for t in (X, W, gamma, beta)] + \
        ([R.clone().requires_grad_(True)] if residual else [])
    want = _plain(*leaves[:4], leaves[4] if residual else None, eps)
    grads = torch.autograd.grad(want, leaves, dz)

    z, saved = _folded_forward(X, W, gamma, beta, R, eps)
    torch.testing.assert_close(z, want.detach(), rtol=1e-9, atol=1e-9)
    dX, dW, dgamma, dbeta, dR = _folded_backward(dz, X, W, gamma, saved)
    torch.testing.assert_close(dX, grads[0], rtol=1e-8, atol=1e-9)
    torch.testing.assert_close(dW, grads[1], rtol=1e-8, atol=1e-9)
    torch.testing.assert_close(dgamma, grads[2], rtol=1e-8, atol=1e-9)
    torch.testing.assert_close(dbeta, grads[3], rtol=1e-8, atol=1e-9)
    if residual:
        torch.testing.assert_close(dR, grads[4], rtol=1e-9, atol=1e-9)


def test_hi_lo_split_of_the_gram_matrix_keeps_the_variance():
    """U = W Gram goes through bf16 tensor cores as two products over Gram = hi + lo (both bf16); the
    variance is a difference of two such numbers, so the split must keep ~16 bits."""
    torch.manual_seed(0)
    M, K, N = 4096, 64, 32
    X = torch.randn(M, K).relu().to(torch.bfloat16).float()
    W = (torch.randn(N, K) / K ** 0.5).to(torch.bfloat16).float()
    G = X.t() @ X
    hi = G.to(torch.bfloat16).float()
    lo = (G - hi).to(torch.bfloat16).float()
    U2 = W @ hi + W @ lo
    U = W.double() @ G.double()
    mean = (W.double() @ X.double().sum(0)) / M
    var = (U * W.double()).sum(1) / M - mean ** 2
    var2 = (U2.double() * W.double()).sum(1) / M - mean ** 2
    assert ((var2 - var).abs() / var).max().item() < 2e-3
    one = (W @ hi).double()                        # hi alone is not enough
    var1 = (one * W.double()).sum(1) / M - mean ** 2
    assert ((var1 - var).abs() / var).max().item() > 5 * ((var2 - var).abs() / var).max().item()
