"""Shim of `torchdiffeq.odeint` (unpinned third-party dependency, not installed).  TEST INFRASTRUCTURE.
Fixed-grid solvers restated from the published algorithm: the grid is `t` itself, every grid point is an
output point (reference call sites: model/cfm.py:212, 335, 471)."""
import torch


def odeint(func, y0, t, *, method="euler", **kwargs):
    ys = [y0]
    y = y0
    for i in range(len(t) - 1):
        t0, t1 = t[i], t[i + 1]
        dt = t1 - t0
        if method == "euler":
            dy = dt * func(t0, y)
        elif method == "midpoint":
            half = 0.5 * dt
            dy = dt * func(t0 + half, y + func(t0, y) * half)
        else:
            raise ValueError(method)
        y = y + dy
        ys.append(y)
    return torch.stack(ys)
