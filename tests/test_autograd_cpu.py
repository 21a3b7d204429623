"""Host-side gradient plumbing (pot_b200/_autograd.py) on CPU: the backward formulas of the differentiable `dist` and
of the lazy loss are checked against torch autograd on a torch restatement; the kernels are replaced by torch stand-ins
here (the GPU tests in test_gradients_gpu.py run the real ones against goldens from the reference)."""

import torch

from pot_b200 import _autograd


def _dist_torch(X, Y, metric):
    same = Y is None
    Y = X if same else Y
    C = (X * X).sum(1)[:, None] + (Y * Y).sum(1)[None, :] - 2 * X @ Y.t()
    C = torch.clamp(C, min=0)
    if metric == "euclidean":
        C = torch.sqrt(C)
    if same:
        C = C * (1 - torch.eye(X.shape[0], dtype=X.dtype))
    return C


def test_dist_function_backward_matches_autograd():
    g = torch.Generator().manual_seed(0)
    for metric in ("sqeuclidean", "euclidean"):
        for same in (False, True):
            X = torch.randn(7, 3, dtype=torch.float64, generator=g, requires_grad=True)
            Y = X if same else torch.randn(5, 3, dtype=torch.float64, generator=g, requires_grad=True)
            W = torch.randn(7, 7 if same else 5, dtype=torch.float64, generator=g)
            C = _autograd.DistFunction.apply(X, Y, metric, same, lambda x, y, m: _dist_torch(x, y, m))
            (C * W).sum().backward()
            got = (X.grad.clone(), None if same else Y.grad.clone())
            X.grad = None
            if not same:
                Y.grad = None
            if metric == "euclidean" and same:
                # autograd through sqrt(0) on the diagonal gives nan (in the reference too); here C = 0 contributes 0
                assert torch.isfinite(got[0]).all()
                continue
            (_dist_torch(X, None if same else Y, metric) * W).sum().backward()
            assert torch.allclose(got[0], X.grad, rtol=1e-10, atol=1e-12), (metric, same)
            if not same:
                assert torch.allclose(got[1], Y.grad, rtol=1e-10, atol=1e-12), (metric, same)


def test_lazy_loss_backward_is_the_envelope_gradient(monkeypatch):
    from pot_b200 import empirical

    monkeypatch.setattr(empirical, "_dist_kernel", lambda x, y, metric: _dist_torch(x, y, metric))  # no GPU here
    g = torch.Generator().manual_seed(1)
    n, m, d, reg = 9, 6, 4, 0.7
    for metric in ("sqeuclidean", "euclidean"):
        X = torch.randn(n, d, dtype=torch.float64, generator=g, requires_grad=True)
        Y = torch.randn(m, d, dtype=torch.float64, generator=g, requires_grad=True)
        a = torch.full((n,), 1.0 / n, dtype=torch.float64, requires_grad=True)
        b = torch.full((m,), 1.0 / m, dtype=torch.float64, requires_grad=True)
        f = torch.randn(n, dtype=torch.float64, generator=g) * 0.1
        gg = torch.randn(m, dtype=torch.float64, generator=g) * 0.1

        def plan_block(xa, xb, fi, gj, reg_, metric_):
            return torch.exp(fi[:, None] + gj[None, :] - _dist_torch(xa, xb, metric_) / reg_)

        P = plan_block(X.detach(), Y.detach(), f, gg, reg, metric)
        loss = (P * _dist_torch(X.detach(), Y.detach(), metric)).sum()
        out = _autograd.LazyLossFunction.apply(loss, X, Y, a, b, f, gg, reg, metric, plan_block, 4)
        (3.0 * out).backward()
        gX, gY, ga, gb = X.grad.clone(), Y.grad.clone(), a.grad.clone(), b.grad.clone()
        X.grad = Y.grad = None
        (3.0 * (P * _dist_torch(X, Y, metric)).sum()).backward()  # dL/dC = P held fixed
        assert torch.allclose(gX, X.grad, rtol=1e-10, atol=1e-12)
        assert torch.allclose(gY, Y.grad, rtol=1e-10, atol=1e-12)
        assert torch.allclose(ga, 3.0 * reg * (f - f.mean()))
        assert torch.allclose(gb, 3.0 * reg * (gg - gg.mean()))


def test_set_gradients_casts_to_the_input_dtype():
    M = torch.ones(2, 2, dtype=torch.float32, requires_grad=True)
    a = torch.ones(2, dtype=torch.float64, requires_grad=True)
    v = _autograd.set_gradients(torch.tensor(1.0), (M, a, None), (torch.full((2, 2), 2.0), torch.full((2,), 3.0), None))
    (2 * v).backward()
    assert M.grad.dtype == torch.float32 and torch.all(M.grad == 4)
    assert a.grad.dtype == torch.float64 and torch.all(a.grad == 6)
