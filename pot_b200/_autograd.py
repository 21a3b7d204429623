"""Gradients around the forward-only kernels.

The CUDA kernels do not record an autograd graph.  The reference is differentiable on this path in two ways: by
unrolling the Python loop (`ot.sinkhorn2`, test/test_bregman.py:203-225) and, in `ot.solve`, by prescribing the
gradient of the value at convergence, `nx.set_gradients(value, (M, a, b), (plan, reg (f - mean f), reg (g - mean g)))`
(ot/solvers.py:485-494, ot/backend.py:2140-2145).  Here every loss-returning entry point attaches that second, envelope
gradient to its value with a custom `autograd.Function`; the plan-returning ones warn that their result is detached.

    set_gradients(val, inputs, grads)   value with prescribed gradients (the role of nx.set_gradients)
    DistFunction                        `dist` differentiable wrt the point clouds (sqeuclidean / euclidean)
    LazyLossFunction                    lazy `empirical_sinkhorn2` loss differentiable wrt (X_s, X_t, a, b): the plan is
                                        rebuilt block by block by the lazy-plan kernel in backward, never stored
"""

import warnings

import torch

ENVELOPE_NOTE = ("pot_b200: the CUDA kernels are forward-only; the returned loss carries the envelope gradient "
                 "(plan, reg (f - mean f), reg (g - mean g)) of ot.solve(grad='envelope') (ot/solvers.py:485-494), which "
                 "equals the gradient of the regularised value at convergence, not the unrolled autodiff gradient")


def any_requires_grad(*ts):
    return torch.is_grad_enabled() and any(isinstance(t, torch.Tensor) and t.requires_grad for t in ts)


class _ValFunction(torch.autograd.Function):
    """value with prescribed gradients (the role of nx.set_gradients, ot/backend.py:2140-2145)."""

    @staticmethod
    def forward(ctx, val, grads, *inputs):
        ctx.grads = grads
        ctx.needs = [isinstance(t, torch.Tensor) and t.requires_grad for t in inputs]
        ctx.dtypes = [t.dtype if isinstance(t, torch.Tensor) else None for t in inputs]
        return val.clone()

    @staticmethod
    def backward(ctx, grad_output):
        out = tuple((g * grad_output).to(dt) if need else None
                    for g, need, dt in zip(ctx.grads, ctx.needs, ctx.dtypes))
        return (None, None) + out


def set_gradients(val, inputs, grads):
    return _ValFunction.apply(val, grads, *inputs)


def warn_detached(fn_name, *tensors):
    """Plan-returning entry points: say so instead of silently returning a constant to a training loop."""
    if any_requires_grad(*tensors):
        warnings.warn(f"pot_b200.{fn_name}: an input requires grad but the CUDA kernels are forward-only; the returned "
                      "plan / potentials are detached.  Use sinkhorn2 / empirical_sinkhorn2 / solve(...).value for a "
                      "differentiable loss (envelope gradient)")


def envelope_grads(plan, log_u, log_v, reg):
    """(d/dM, d/da, d/db) of ot/solvers.py:485-494 from the log-scalings."""
    return plan, reg * (log_u - log_u.mean()), reg * (log_v - log_v.mean())


class DistFunction(torch.autograd.Function):
    """C = dist(X, Y) from the dist kernel, differentiable wrt X and Y like the reference's torch backend
    (ot/utils.py:270-307).  With G = dL/dC:
        sqeuclidean: dL/dX = 2 (rowsum(G) * X - G Y),       dL/dY = 2 (colsum(G) * Y - G^T X)
        euclidean  : the same with G / (2 C) in place of G (entries with C = 0 contribute nothing).
    The clamp at 0 and the zeroed diagonal (`same`) have zero gradient there, as in the reference."""

    @staticmethod
    def forward(ctx, X, Y, metric, same, kernel):
        C = kernel(X.detach(), None if same else Y.detach(), metric)
        ctx.metric, ctx.same = metric, same
        ctx.save_for_backward(X.detach(), Y.detach(), C if metric == "euclidean" else None)
        return C

    @staticmethod
    def backward(ctx, G):
        X, Y, C = ctx.saved_tensors
        if ctx.metric == "euclidean":
            G = torch.where(C > 0, G / (2 * C), torch.zeros_like(G))
        elif ctx.same:
            G = G.clone().fill_diagonal_(0.0)
        gX = 2 * (G.sum(1, keepdim=True) * X - G @ Y) if ctx.needs_input_grad[0] else None
        gY = 2 * (G.sum(0)[:, None] * Y - G.t() @ X) if ctx.needs_input_grad[1] else None
        return gX, gY, None, None, None


class LazyLossFunction(torch.autograd.Function):
    """loss = sum_ij C_ij P_ij of the lazy solve with the envelope gradient: dL/dC = P, i.e. for the squared euclidean
    cost dL/dX_s = 2 (rowsum(P) * X_s - P X_t), dL/dX_t = 2 (colsum(P) * X_t - P^T X_s), dL/da = reg (f - mean f),
    dL/db = reg (g - mean g) (f, g the log-scalings).  P is produced block by block (`block` rows at a time) by the
    lazy-plan kernel inside backward and never stored whole."""

    @staticmethod
    def forward(ctx, loss, X_s, X_t, a, b, f, g, reg, metric, plan_block, block):
        ctx.reg, ctx.metric, ctx.plan_block, ctx.block = reg, metric, plan_block, block
        ctx.save_for_backward(X_s.detach().contiguous(), X_t.detach().contiguous(), f, g)
        return loss.clone()

    @staticmethod
    def backward(ctx, grad_output):
        X_s, X_t, f, g = ctx.saved_tensors
        need_s, need_t, need_a, need_b = ctx.needs_input_grad[1:5]
        gXs = gXt = ga = gb = None
        if need_s or need_t:
            gXs = torch.zeros_like(X_s)
            gXt = torch.zeros_like(X_t)
            for i0 in range(0, X_s.shape[0], ctx.block):
                xs = X_s[i0:i0 + ctx.block]
                P = ctx.plan_block(xs, X_t, f[i0:i0 + ctx.block], g, ctx.reg, ctx.metric)
                if ctx.metric == "euclidean":
                    from .empirical import _dist_kernel

                    C = _dist_kernel(xs, X_t, "euclidean")
                    P = torch.where(C > 0, P / (2 * C), torch.zeros_like(P))
                gXs[i0:i0 + ctx.block] = 2 * (P.sum(1, keepdim=True) * xs - P @ X_t)
                gXt += 2 * (P.sum(0)[:, None] * X_t - P.t() @ xs)
            gXs, gXt = gXs * grad_output, gXt * grad_output
        if need_a:
            ga = ctx.reg * (f - f.mean()) * grad_output
        if need_b:
            gb = ctx.reg * (g - g.mean()) * grad_output
        return (None, gXs if need_s else None, gXt if need_t else None, ga, gb, None, None, None, None, None, None)
