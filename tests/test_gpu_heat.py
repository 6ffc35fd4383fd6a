"""Parity of the CUDA FTCS path with the CPU oracle: bit-exact in `exact` mode, 1e-12 relative per step in the
production (pre-multiplied, FMA) mode -- the tolerance BASELINE.json:north_star states for explicit FD fields."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def noise(nx, ny, seed=20260001):
    return np.asfortranarray(np.random.default_rng(seed).random((nx, ny)))


@pytest.mark.parametrize("shape", [(3, 3), (5, 4), (66, 34), (130, 257), (1026, 1026)])
def test_ftcs_exact_mode_bitwise(cfd, oracle, shape):
    nx, ny = shape
    t0 = noise(nx, ny)
    h = cfd.Heat2D(nx, ny, tleft=1.0, tright=2.0, tbottom=3.0, ttop=4.0)
    h.upload(t0)
    h.step(3, exact=True)
    want = oracle.heat2d_ftcs(t0, 3, dt=h.dt, tbc=(1.0, 2.0, 3.0, 4.0))
    assert np.array_equal(h.download(), want)
    h.close()


@pytest.mark.parametrize("shape", [(66, 34), (1026, 1026), (2050, 514)])
def test_ftcs_fast_mode_1e12_per_step(cfd, oracle, shape):
    nx, ny = shape
    t0 = noise(nx, ny)
    h = cfd.Heat2D(nx, ny)
    h.upload(t0)
    cur = t0
    for _ in range(4):                 # per-step comparison, restarting the oracle from the GPU state
        h.step(1)
        want = oracle.heat2d_ftcs(cur, 1, dt=h.dt)
        got = h.download()
        assert np.abs(got - want).max() <= 1e-12 * np.abs(want).max()
        cur = got
    h.close()


def test_ftcs_reference_setup_source_and_bc(cfd, oracle):
    """SURVEY 8d input 2 at reduced size: T0 = 0, q = 1e3 where x < 0.1, Dirichlet 0, Fo = 0.2, 100 steps."""
    nx = ny = 258
    t0 = np.zeros((nx, ny), order="F")
    h = cfd.Heat2D(nx, ny)
    h.upload(t0)
    h.step(100)
    want = oracle.heat2d_ftcs(t0, 100, dt=h.dt)
    got = h.download()
    assert np.abs(got - want).max() <= 1e-11 * np.abs(want).max()
    assert got[:, 0].max() == 0.0 and got[0, :].max() == 0.0
    assert got[2, ny // 2] > 0.0 and got[nx // 2, ny // 2] >= 0.0
    h.close()


def test_ftcs_full_size_eigenmode(cfd):
    """BASELINE.json config 2 size (16384^2): analytic decay of a discrete eigenmode, 10 steps."""
    n = 16384
    x = np.linspace(0.0, 1.0, n)
    s = np.sin(np.pi * x)
    s[0] = s[-1] = 0.0
    t0 = np.asfortranarray(s[:, None] * s[None, :])
    h = cfd.Heat2D(n, n, qval=0.0)
    h.upload(t0)
    h.step(10)
    got = h.download()
    dx = 1.0 / (n - 1)
    lam = 1.0 - 8 * (5.0 * h.dt / dx**2) * np.sin(np.pi * dx / 2) ** 2
    assert np.abs(got - lam**10 * t0).max() < 1e-12
    h.close()


@pytest.mark.parametrize("shape", [(5, 4), (66, 34), (130, 257), (515, 300)])
def test_implicit_lod_bitwise(cfd, oracle, shape):
    """Implicit Euler, locally one-dimensional (tridiagonal assembly + thomas of the implicit 1-D program applied
    along x then y): same operation order as the oracle -> same bits."""
    nx, ny = shape
    t0 = noise(nx, ny, 7)
    h = cfd.Heat2D(nx, ny, dt=2.0e-4, tleft=1.0, tright=2.0, tbottom=3.0, ttop=4.0)
    h.upload(t0)
    h.implicit_step(3, exact=True)
    want = oracle.heat2d_lod(t0, 3, dt=2.0e-4, tbc=(1.0, 2.0, 3.0, 4.0))
    assert np.array_equal(h.download(), want)
    h.close()
    # production path (reciprocal pivots, FMAs): 1e-12 relative per step, restarting the oracle from the GPU state
    h = cfd.Heat2D(nx, ny, dt=2.0e-4, tleft=1.0, tright=2.0, tbottom=3.0, ttop=4.0)
    h.upload(t0)
    cur = t0
    for _ in range(3):
        h.implicit_step(1)
        want = oracle.heat2d_lod(cur, 1, dt=2.0e-4, tbc=(1.0, 2.0, 3.0, 4.0))
        cur = h.download()
        assert np.abs(cur - want).max() <= 1e-12 * np.abs(want).max()
    h.close()


@pytest.mark.parametrize("shape,fo", [((2100, 1300), 5.0), ((1030, 4100), 0.7), ((4100, 1030), 20.0)])
def test_implicit_lod_chunked_lines_match_the_serial_solve(cfd, oracle, shape, fo):
    """Grids large enough for the production sweeps to cut the lines into chunks (each warming up over the
    contraction length of the Thomas recurrences): same result as the one-thread-per-line exact mode, which is
    bit-identical with the oracle, to rounding."""
    nx, ny = shape
    t0 = noise(nx, ny, 11) * 100.0
    kw = dict(fo=fo, tleft=1.0, tright=2.0, tbottom=3.0, ttop=4.0)
    a, b = cfd.Heat2D(nx, ny, **kw), cfd.Heat2D(nx, ny, **kw)
    a.upload(t0)
    b.upload(t0)
    for k in range(3):
        a.implicit_step(1, exact=True)
        b.implicit_step(1)
        ta, tb = a.download(), b.download()
        assert np.abs(ta - tb).max() <= 1e-12 * np.abs(ta).max(), k
        b.upload(ta)
    want = oracle.heat2d_lod(t0, 1, dt=a.dt, tbc=(1.0, 2.0, 3.0, 4.0)) if nx * ny < 3_000_000 else None
    if want is not None:
        a.upload(t0)
        a.implicit_step(1, exact=True)
        assert np.array_equal(a.download(), want)
    a.close()
    b.close()


def test_implicit_lod_large_time_step_is_stable(cfd):
    """Fo = 50 (250 x the explicit limit): bounded, monotone approach to the steady state."""
    n = 1026
    h = cfd.Heat2D(n, n, fo=50.0, qval=0.0, tleft=1.0, tright=1.0, tbottom=1.0, ttop=1.0)
    h.upload(np.zeros((n, n), order="F"))
    h.implicit_step(40)
    t = h.download()
    assert t.min() >= -1e-12 and t.max() <= 1.0 + 1e-12
    h.close()
