"""TEST INFRASTRUCTURE: numpy restatement of lcsim's Algo::fitBSpline (external/lcsim/lcsim/src/algo.cpp:449-589,
linspace :377-385), the curve plan_m1 draws through its sampled points (lc/light_curtain.py:596-598).

Written independently of the library's C++ (csrc/lc_spline.cu), scalar by scalar in numpy float32 / Python float
(= C double) following the reference's expression types.  Pinned: bit-identical to the reference's own two function
definitions compiled from where they lie (oracle/ref_spline.py -> oracle/_ref/libspline_ref.so) on the cases of
tests/golden/spline_ref.npz (tests/test_oracle_golden.py), which tests/golden/make_plan_m1.py generated with it.
"""
import numpy as np

f32 = np.float32


def _linspace(a, b, N):
    """algo.cpp:377-385: float h = (b - a) / float(N - 1); values accumulated by repeated float addition."""
    a, b = f32(a), f32(b)
    with np.errstate(divide="ignore", invalid="ignore"):
        h = f32(f32(b - a) / f32(N - 1))
    if N <= 0:
        return np.zeros(0, np.float32)
    steps = np.full(N - 1, h, np.float32)
    return np.concatenate([[a], f32(a) + np.add.accumulate(steps, dtype=np.float32)]).astype(np.float32) if a == 0 else \
        np.array([_acc(a, h, k) for k in range(N)], np.float32)


def _acc(a, h, k):
    v = f32(a)
    for _ in range(k):
        v = f32(v + h)
    return v


def fit_bspline(pts):
    """pts [n,3] (x, y, weight) -> curve [m,2] float32, like pylc.fitBSpline(pts, params, True)."""
    pts = np.asarray(pts, np.float32).reshape(-1, 3)
    n_pts = pts.shape[0]
    assert n_pts >= 2          # the reference reads row 1 of its input
    inp = pts[::-1].copy()                                            # :462
    c1 = [f32(float(f32(inp[0, k] + inp[1, k])) / 2.0) for k in range(3)]            # float sum, double halving (:465)
    c2 = [f32(float(f32(inp[-1, k] + inp[-2, k])) / 2.0) for k in range(3)]
    inp = np.vstack([inp, np.array(c1, np.float32), np.array(c2, np.float32)]).astype(np.float32)
    n = inp.shape[0] - 2
    dx, dy = np.zeros(n, np.float32), np.zeros(n, np.float32)
    Ax, Ay, Bi = np.zeros(n, np.float32), np.zeros(n, np.float32), np.zeros(n, np.float32)
    DIV = f32(3.0)
    dx[0] = f32(f32(inp[1, 0] - inp[0, 0]) / DIV)
    dy[0] = f32(f32(inp[1, 1] - inp[0, 1]) / DIV)
    dx[n - 1] = f32(f32(inp[n - 1, 0] - inp[n - 2, 0]) / DIV)
    dy[n - 1] = f32(f32(inp[n - 1, 1] - inp[n - 2, 1]) / DIV)
    Bi[1] = f32(-1.0 / float(inp[1, 2]))                              # double division, stored as float
    Ax[1] = f32(f32(-f32(f32(inp[2, 0] - inp[0, 0]) - dx[0])) * Bi[1])
    Ay[1] = f32(f32(-f32(f32(inp[2, 1] - inp[0, 1]) - dy[0])) * Bi[1])
    for i in range(2, n - 1):
        Bi[i] = f32(f32(-1.0) / f32(inp[i, 2] + Bi[i - 1]))           # int / float -> float division
        Ax[i] = f32(f32(-f32(f32(inp[i + 1, 0] - inp[i - 1, 0]) - Ax[i - 1])) * Bi[i])
        Ay[i] = f32(f32(-f32(f32(inp[i + 1, 1] - inp[i - 1, 1]) - Ay[i - 1])) * Bi[i])
    for i in range(n - 2, 0, -1):
        dx[i] = f32(Ax[i] + f32(dx[i + 1] * Bi[i]))
        dy[i] = f32(Ay[i] + f32(dy[i + 1] * Bi[i]))
    xs, ys = [np.array([inp[0, 0]], np.float32)], [np.array([inp[0, 1]], np.float32)]
    three, two = f32(3.0), f32(2.0)
    for i in range(n - 1):
        ddx, ddy = float(f32(inp[i, 0] - inp[i + 1, 0])), float(f32(inp[i, 1] - inp[i + 1, 1]))
        dist = f32(np.sqrt(ddx ** 2 + ddy ** 2))                      # pow / sqrt in double, stored as float
        count = int(float(dist) / 0.01)
        p0 = (inp[i, 0], inp[i, 1])
        p1 = (f32(inp[i, 0] + dx[i]), f32(inp[i, 1] + dy[i]))
        p2 = (f32(inp[i + 1, 0] - dx[i + 1]), f32(inp[i + 1, 1] - dy[i + 1]))
        p3 = (inp[i + 1, 0], inp[i + 1, 1])
        co = []
        for a in (0, 1):
            A = f32(f32(f32(p3[a] - f32(three * p2[a])) + f32(three * p1[a])) - p0[a])
            B = f32(three * f32(f32(p2[a] - f32(two * p1[a])) + p0[a]))
            C = f32(three * f32(p1[a] - p0[a]))
            co.append((A, B, C, p0[a]))
        if count < 1:
            continue
        t = _linspace(0.0, f32(1 + 0.0 - (1.0 / float(f32(count)))), count)
        for (A, B, C, Dd), acc in zip(co, (xs, ys)):
            t2 = (t * t).astype(np.float32)
            t3 = (t2 * t).astype(np.float32)
            v = ((((t3 * A).astype(np.float32) + (t2 * B).astype(np.float32)).astype(np.float32)
                  + (t * C).astype(np.float32)).astype(np.float32) + Dd).astype(np.float32)
            acc.append(v)
    x, y = np.concatenate(xs), np.concatenate(ys)
    if x.size == 1:
        x = np.array([x[0], f32(float(x[0]) + 0.2), f32(float(x[0]) - 0.2)], np.float32)
        y = np.array([y[0], y[0], y[0]], np.float32)
    return np.stack([x[::-1], y[::-1]], 1).astype(np.float32)
