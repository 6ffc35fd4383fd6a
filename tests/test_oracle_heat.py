"""Pin the heat-conduction restatement (oracle/heat_oracle.c) with analytic answers: the reference prints no
numbers to compare with, so eigenmode decay of the discrete operator and a dense solve are the anchors."""
import numpy as np


def test_thomas_against_dense(oracle):
    rng = np.random.default_rng(0)
    n = 37
    a, c = -rng.random(n), -rng.random(n)
    b = 2.5 + rng.random(n)
    a[0] = 0.0
    c[-1] = 0.0
    d = rng.random(n)
    A = np.diag(b) + np.diag(a[1:], -1) + np.diag(c[:-1], 1)
    assert np.allclose(oracle.thomas(a, b, c, d), np.linalg.solve(A, d), rtol=1e-13, atol=1e-14)


def test_ftcs_eigenmode_decay(oracle):
    """sin(pi x) sin(2 pi y) is an eigenvector of the 5-point operator: one FTCS step multiplies it by
    1 - 4 Fo_x sin^2(pi dx/2) - 4 Fo_y sin^2(2 pi dy/2) exactly (up to rounding)."""
    nx, ny, nsteps = 41, 33, 7
    x = np.linspace(0.0, 1.0, nx)
    y = np.linspace(0.0, 1.0, ny)
    t0 = np.asfortranarray(np.sin(np.pi * x)[:, None] * np.sin(2 * np.pi * y)[None, :])
    t0[0, :] = t0[-1, :] = 0.0
    t0[:, 0] = t0[:, -1] = 0.0
    k, fo = 5.0, 0.2
    dx, dy = 1.0 / (nx - 1), 1.0 / (ny - 1)
    dt = fo * dx * dx / k
    t = oracle.heat2d_ftcs(t0, nsteps, k=k, dt=dt, qval=0.0)
    lam = 1.0 - 4 * k * dt / dx**2 * np.sin(np.pi * dx / 2) ** 2 - 4 * k * dt / dy**2 * np.sin(np.pi * dy) ** 2
    assert np.abs(t - lam**nsteps * t0).max() < 1e-13


def test_ftcs_reduces_to_the_1d_reference_formula(oracle):
    """With no y-variation in the interior (rows identical, ya/yb far apart) the 2-D update equals the 1-D
    update of fdm1d_explicit_heat_cond.f90:133-135 along x, source included."""
    nx, ny = 21, 5
    k, rho, cp = 5.0, 1.0, 1.0
    dx = 1.0 / (nx - 1)
    dt = 0.2 * dx * dx / k
    rng = np.random.default_rng(3)
    row = rng.random(nx)
    t0 = np.asfortranarray(np.repeat(row[:, None], ny, axis=1))
    t = oracle.heat2d_ftcs(t0, 1, k=k, dt=dt, lim=(0.0, 1.0, 0.0, 1.0))
    x = np.arange(nx) * dx
    q = np.where(x < oracle.QTHR_REF, 1.0e3, 0.0)
    want = row + (k / rho / cp * (np.roll(row, -1) - 2 * row + np.roll(row, 1)) / dx**2 + q / rho / cp) * dt
    j = ny // 2      # middle row: y-neighbours equal -> y-term is exactly zero
    assert np.abs(t[1:-1, j] - want[1:-1]).max() <= 4e-16 * np.abs(want).max()
    assert t[0, j] == 0.0 and t[-1, j] == 0.0


def test_lod_implicit_steady_state(oracle):
    """Implicit LOD steps are unconditionally stable and converge to the steady solution of the same operator."""
    nx = ny = 17
    t0 = np.zeros((nx, ny), order="F")
    t = oracle.heat2d_lod(t0, 400, dt=5e-3, tbc=(1.0, 1.0, 1.0, 1.0), qval=0.0)
    assert np.abs(t - 1.0).max() < 1e-6
