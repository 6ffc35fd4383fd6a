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


# ---- variable conductivity: the operator of steady-heat-conduction-2d/fdm2d_steady_heat_cond.f90:162-190 ---------------
def _lam(x, y):
    return 20.0 * (1.0 + 3.0 * x / 0.03 + 0.5 * np.sin(40.0 * y))


def _qd(x, y):
    return 5.0e7 * (1.0 + 10.0 * x) + 0.0 * y


@pytest.mark.parametrize("shape", [(7, 5), (66, 34), (130, 257), (1030, 70)])
def test_variable_conductivity_operator_and_explicit_steps_bitwise(cfd, oracle, shape):
    """Four face conductivities per node (LAMBDA at the face mid-points, :171-174): the residual of the steady
    equations and explicit steps with that operator, bit for bit against oracle/heat_oracle.c."""
    nx, ny = shape
    lim = (0.0, 0.03, 0.0, 0.02)
    o = oracle.HeatVar(nx, ny, _lam, _qd, lim=lim)
    dx, dy = o.dx, o.dy
    dt = 0.2 * min(dx, dy) ** 2 / 80.0 * 7.8e3 * 460.0          # Fo = 0.2 at the largest conductivity
    h = cfd.Heat2D(nx, ny, rho=7.8e3, cp=460.0, xa=lim[0], xb=lim[1], ya=lim[2], yb=lim[3], tleft=300.0, tright=310.0,
                   tbottom=320.0, ttop=330.0, dt=dt)
    h.set_operator(_lam, _qd)
    t0 = np.asfortranarray(300.0 + 50.0 * np.random.default_rng(11).random((nx, ny)))
    h.upload(t0)
    assert np.array_equal(h.apply(), o.apply(t0))
    h.step_var(5)
    want = o.ftcs(t0, 5, dt, rho=7.8e3, cp=460.0, tbc=(300.0, 310.0, 320.0, 330.0))
    assert np.array_equal(h.download(), want)
    h.close()


def test_matrix_free_gauss_seidel_reference_case(cfd, oracle):
    """The program as shipped (7 x 5 nodes, LAMBDA = 20, QDOT = 5e7, walls at 300, rf 0.5, tol 1e-7, from zero): the
    two-colour device passes against the same passes in the oracle bit for bit, and -- run to a tight tolerance --
    against the reference's lexicographic order and a direct solve of the 35 x 35 system."""
    nx, ny, lim = 7, 5, (0.0, 0.03, 0.0, 0.02)
    lam, qd = (lambda x, y: 20.0 + 0.0 * x), (lambda x, y: 5.0e7 + 0.0 * x)
    o = oracle.HeatVar(nx, ny, lam, qd, lim=lim)
    h = cfd.Heat2D(nx, ny, xa=lim[0], xb=lim[1], ya=lim[2], yb=lim[3], tleft=300.0, tright=300.0, tbottom=300.0, ttop=300.0)
    h.set_operator(lam, qd)
    h.upload(np.zeros((nx, ny), order="F"))
    n_g, d_g = h.steady_gs()
    x_o, n_o, d_o = o.gauss_seidel(order=1)
    assert n_g == n_o and abs(d_g - d_o) <= 1e-12 * d_o
    assert np.array_equal(h.download(), x_o)
    x_ref, n_ref, _ = o.gauss_seidel(order=0)                     # the reference's own order: 151 passes, same plate
    assert abs(n_ref - n_g) <= 3 and np.abs(x_ref - x_o).max() < 1e-5
    h.upload(np.zeros((nx, ny), order="F"))
    h.steady_gs(maxit=5000, tol=1e-12)
    x_seq, _, _ = o.gauss_seidel(order=0, tol=1e-12, imax=5000)
    got = h.download()
    assert np.abs(got - x_seq).max() <= 1e-8 * np.abs(x_seq).max()
    # direct solve of the rows interior() / boundary() define
    A, b = np.zeros((nx * ny, nx * ny)), np.zeros(nx * ny)
    for j in range(ny):
        for i in range(nx):
            k = j * nx + i
            if i in (0, nx - 1) or j in (0, ny - 1):
                A[k, k], b[k] = 1.0, 300.0
            else:
                A[k, k - 1], A[k, k + 1], A[k, k - nx], A[k, k + nx] = o.aw[i, j], o.ae[i, j], o.as_[i, j], o.an[i, j]
                A[k, k], b[k] = o.ap[i, j], -o.q[i, j]
    exact = np.linalg.solve(A, b).reshape(ny, nx).T
    assert np.abs(got - exact).max() <= 1e-8 * np.abs(exact).max()
    assert abs(exact[3, 2] - 398.0) < 0.1                         # the plate's centre temperature
    h.close()


def test_matrix_free_gauss_seidel_variable_conductivity(cfd, oracle):
    nx, ny, lim = 41, 29, (0.0, 0.03, 0.0, 0.02)
    o = oracle.HeatVar(nx, ny, _lam, _qd, lim=lim)
    h = cfd.Heat2D(nx, ny, xa=lim[0], xb=lim[1], ya=lim[2], yb=lim[3], tleft=300.0, tright=310.0, tbottom=320.0, ttop=330.0)
    h.set_operator(_lam, _qd)
    x0 = np.asfortranarray(300.0 + np.random.default_rng(2).random((nx, ny)))
    h.upload(x0)
    n_g, d_g = h.steady_gs(maxit=40, tol=0.0)
    x_o, n_o, d_o = o.gauss_seidel(x0=x0, tbc=(300.0, 310.0, 320.0, 330.0), tol=0.0, imax=40, order=1)
    assert n_g == n_o == 40 and abs(d_g - d_o) <= 1e-11 * d_o
    assert np.array_equal(h.download(), x_o)
    h.close()
