"""paddle.incubate.optimizer.functional (reference: python/paddle/incubate/optimizer/functional/{bfgs,lbfgs,
line_search}.py): quasi-Newton minimisers of a scalar function of one tensor.

Both minimisers share a strong-Wolfe line search (bracketing, then zoom with cubic interpolation; c1 = 1e-4, c2 = 0.9).
BFGS keeps the dense inverse-Hessian estimate H and updates it with
``H <- (I - rho s y^T) H (I - rho y s^T) + rho s s^T``; L-BFGS keeps the last ``history_size`` (s, y) pairs and applies
the two-loop recursion.
"""
from __future__ import annotations

import torch

__all__ = ["minimize_bfgs", "minimize_lbfgs"]


def _value_and_grad(f, x):
    x = x.detach().clone().requires_grad_(True)
    with torch.enable_grad():
        v = f(x)
        (g,) = torch.autograd.grad(v, x)
    return v.detach(), g.detach()


def _cubic_min(a, fa, ga, b, fb, gb):
    """Minimiser of the cubic through (a, fa, ga), (b, fb, gb), clamped to the bracket; bisection when it degenerates."""
    d1 = ga + gb - 3 * (fa - fb) / (a - b)
    sq = d1 * d1 - ga * gb
    if sq < 0:
        return (a + b) / 2
    d2 = sq ** 0.5
    t = b - (b - a) * (gb + d2 - d1) / (gb - ga + 2 * d2) if a <= b else a - (a - b) * (ga + d2 - d1) / (ga - gb + 2 * d2)
    lo, hi = (a, b) if a <= b else (b, a)
    return min(max(t, lo), hi)


def _strong_wolfe(f, x, p, f0, g0, step, max_iters, c1=1e-4, c2=0.9):
    """-> (step, value, gradient, function calls).  Falls back to the best point seen when the budget runs out."""
    d0 = float(torch.dot(g0.reshape(-1), p.reshape(-1)))
    calls = 0

    def phi(a):
        nonlocal calls
        calls += 1
        v, g = _value_and_grad(f, x + a * p)
        return float(v), g, float(torch.dot(g.reshape(-1), p.reshape(-1)))
    f0 = float(f0)
    a_prev, f_prev, d_prev, g_prev = 0.0, f0, d0, g0
    a = float(step)
    bracket = None
    for i in range(max_iters):
        fa, ga, da = phi(a)
        if fa > f0 + c1 * a * d0 or (i > 0 and fa >= f_prev):
            bracket = (a_prev, f_prev, d_prev, g_prev, a, fa, da, ga)
            break
        if abs(da) <= -c2 * d0:
            return a, fa, ga, calls
        if da >= 0:
            bracket = (a, fa, da, ga, a_prev, f_prev, d_prev, g_prev)
            break
        a_prev, f_prev, d_prev, g_prev = a, fa, da, ga
        a = a * 2.0
    if bracket is None:
        return a_prev, f_prev, g_prev, calls
    lo, f_lo, d_lo, g_lo, hi, f_hi, d_hi, g_hi = bracket
    for _ in range(max_iters):
        if abs(hi - lo) < 1e-12:
            break
        a = _cubic_min(lo, f_lo, d_lo, hi, f_hi, d_hi)
        span = abs(hi - lo)
        if min(abs(a - lo), abs(a - hi)) < 0.1 * span:          # keep away from the ends, or the zoom stalls
            a = (lo + hi) / 2
        fa, ga, da = phi(a)
        if fa > f0 + c1 * a * d0 or fa >= f_lo:
            hi, f_hi, d_hi, g_hi = a, fa, da, ga
        else:
            if abs(da) <= -c2 * d0:
                return a, fa, ga, calls
            if da * (hi - lo) >= 0:
                hi, f_hi, d_hi, g_hi = lo, f_lo, d_lo, g_lo
            lo, f_lo, d_lo, g_lo = a, fa, da, ga
    return lo, f_lo, g_lo, calls


def _setup(objective_func, initial_position, dtype, line_search_fn, initial_inverse_hessian_estimate):
    if dtype not in ("float32", "float64"):
        raise ValueError(f"The dtype must be 'float32' or 'float64', but the specified is {dtype}.")
    if line_search_fn != "strong_wolfe":
        raise NotImplementedError(f"Currently only support line_search_fn = 'strong_wolfe', but the specified is '{line_search_fn}'")
    td = torch.float32 if dtype == "float32" else torch.float64
    x = torch.as_tensor(initial_position).detach().to(td).clone()
    n = x.numel()
    if initial_inverse_hessian_estimate is None:
        h0 = torch.eye(n, dtype=td)
    else:
        h0 = torch.as_tensor(initial_inverse_hessian_estimate).to(td)
        if not torch.allclose(h0, h0.t()) or bool((torch.linalg.eigvalsh(h0.double()) <= 0).any()):
            raise ValueError("The initial_inverse_hessian_estimate should be symmetric and positive definite.")
    return x, h0, td


def minimize_bfgs(objective_func, initial_position, max_iters=50, tolerance_grad=1e-7, tolerance_change=1e-9,
                  initial_inverse_hessian_estimate=None, line_search_fn="strong_wolfe", max_line_search_iters=50,
                  initial_step_length=1.0, dtype="float32", name=None):
    """-> (is_converge, num_func_calls, position, objective_value, objective_gradient, inverse_hessian_estimate)."""
    x, h, td = _setup(objective_func, initial_position, dtype, line_search_fn, initial_inverse_hessian_estimate)
    shape = x.shape
    value, grad = _value_and_grad(objective_func, x)
    calls, converged = 1, False
    eye = torch.eye(x.numel(), dtype=td)
    for _ in range(max_iters):
        if float(grad.abs().max()) < tolerance_grad:
            converged = True
            break
        p = -(h @ grad.reshape(-1)).reshape(shape)
        a, new_value, new_grad, used = _strong_wolfe(objective_func, x, p, value, grad, initial_step_length, max_line_search_iters)
        calls += used
        s = (a * p).reshape(-1)
        y = (new_grad - grad).reshape(-1)
        x, change = x + a * p, abs(float(new_value) - float(value))
        value, grad = torch.as_tensor(new_value, dtype=td), new_grad
        sy = float(torch.dot(s, y))
        if sy > 1e-10:
            rho = 1.0 / sy
            left = eye - rho * torch.outer(s, y)
            h = left @ h @ left.t() + rho * torch.outer(s, s)
        if float(grad.abs().max()) < tolerance_grad or change < tolerance_change or float(s.abs().max()) < tolerance_change:
            converged = True
            break
    return (torch.tensor(converged), torch.tensor(calls, dtype=torch.int64), x, torch.as_tensor(value, dtype=td), grad, h)


def minimize_lbfgs(objective_func, initial_position, history_size=100, max_iters=50, tolerance_grad=1e-8, tolerance_change=1e-8,
                   initial_inverse_hessian_estimate=None, line_search_fn="strong_wolfe", max_line_search_iters=50,
                   initial_step_length=1.0, dtype="float32", name=None):
    """-> (is_converge, num_func_calls, position, objective_value, objective_gradient)."""
    x, h0, td = _setup(objective_func, initial_position, dtype, line_search_fn, initial_inverse_hessian_estimate)
    shape = x.shape
    value, grad = _value_and_grad(objective_func, x)
    calls, converged = 1, False
    hist_s, hist_y, hist_rho = [], [], []
    for _ in range(max_iters):
        if float(grad.abs().max()) < tolerance_grad:
            converged = True
            break
        q = grad.reshape(-1).clone()
        alphas = []
        for s, y, rho in zip(reversed(hist_s), reversed(hist_y), reversed(hist_rho)):
            al = rho * torch.dot(s, q)
            alphas.append(al)
            q = q - al * y
        r = h0 @ q
        if hist_s:                                                     # scale by the latest curvature, as usual
            r = r * (torch.dot(hist_s[-1], hist_y[-1]) / torch.dot(hist_y[-1], hist_y[-1])) if initial_inverse_hessian_estimate is None else r
        for (s, y, rho), al in zip(zip(hist_s, hist_y, hist_rho), reversed(alphas)):
            r = r + s * (al - rho * torch.dot(y, r))
        p = -r.reshape(shape)
        a, new_value, new_grad, used = _strong_wolfe(objective_func, x, p, value, grad, initial_step_length, max_line_search_iters)
        calls += used
        s = (a * p).reshape(-1)
        y = (new_grad - grad).reshape(-1)
        x, change = x + a * p, abs(float(new_value) - float(value))
        value, grad = torch.as_tensor(new_value, dtype=td), new_grad
        sy = float(torch.dot(s, y))
        if sy > 1e-10:
            hist_s.append(s)
            hist_y.append(y)
            hist_rho.append(1.0 / sy)
            if len(hist_s) > history_size:
                hist_s.pop(0), hist_y.pop(0), hist_rho.pop(0)
        if float(grad.abs().max()) < tolerance_grad or change < tolerance_change or float(s.abs().max()) < tolerance_change:
            converged = True
            break
    return (torch.tensor(converged), torch.tensor(calls, dtype=torch.int64), x, torch.as_tensor(value, dtype=td), grad)
