"""Scalar-transport demos on the device: line sweeps in both directions, the convergence loops and the
per-step time terms against oracle/lin_oracle.c."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cfd():
    import cfd_b200
    return cfd_b200


@pytest.fixture(scope="module")
def oracle():
    from oracle import oracle as o
    return o


@pytest.mark.parametrize("isch", [1, 2])
@pytest.mark.parametrize("isol", [1, 2])
@pytest.mark.parametrize("n", [20, 37])
def test_stag_point_sequential_bitwise(cfd, oracle, isch, isol, n):
    o = oracle.StagPoint(n, n, isch=isch, isol=isol)
    g = cfd.StagnationPoint(n, n, isch=isch, isol=isol, order="sequential")
    for k in range(3):                                    # single sweeps: same bits
        (o.wesweep if isol == 1 else o.snsweep)()
        (g.sys.wesweep if isol == 1 else g.sys.snsweep)()
        assert np.array_equal(g.sys.phi(), o.phi), k
        assert abs(g.sys.residual() - o.residual()) <= 1e-12 * o.residual()
    it_o, it_g = o.solve(), g.solve()                     # the whole loop: same iteration count, same field
    assert it_g == it_o
    assert np.array_equal(g.sys.phi(), o.phi)
    phi, fwall = g.finish()
    assert fwall == o.finish() and np.array_equal(phi, o.phi)
    g.close()


@pytest.mark.parametrize("isol", [1, 2])
def test_stag_point_red_black(cfd, oracle, isol):
    n = 40
    o = oracle.StagPoint(n, n, isch=1, isol=isol, order=1)
    g = cfd.StagnationPoint(n, n, isch=1, isol=isol, order="redblack")
    for k in range(5):
        (o.wesweep if isol == 1 else o.snsweep)()
        (g.sys.wesweep if isol == 1 else g.sys.snsweep)()
    assert np.abs(g.sys.phi() - o.phi).max() < 1e-13
    it_o, it_g = o.solve(), g.solve()
    assert abs(it_g - it_o) <= 1
    assert np.abs(g.sys.phi() - o.phi).max() < 1e-6       # both stop at rsm < 1e-4
    g.close()


@pytest.mark.parametrize("itsch", [1, 2, 3])
@pytest.mark.parametrize("isol", [1, 2])
def test_unsteady_stag_point_bitwise(cfd, oracle, itsch, isol):
    o = oracle.StagPoint(20, 20, isch=2, isol=isol)
    g = cfd.StagnationPointUnsteady(20, 20, isch=2, itsch=itsch, isol=isol, order="sequential")
    for k in range(6):
        it_o = o.step(itsch)
        it_g = g.step(1)[0]
        assert it_g == it_o, k
        assert np.array_equal(g.sys.phi(), o.phi), k
    g.close()


@pytest.mark.parametrize("bc", [1, 2])
@pytest.mark.parametrize("order", ["sequential", "redblack"])
def test_oblique_convection(cfd, oracle, bc, order):
    o = oracle.Oblique(24, 24, east_north_bc=bc, order=0 if order == "sequential" else 1)
    g = cfd.ObliqueConvection(24, 24, east_north_bc=bc, order=order)
    it_o, it_g = o.solve(), g.solve()
    assert it_g == it_o
    if order == "sequential":
        assert np.array_equal(g.finish(), o.finish())
    else:
        assert np.abs(g.finish() - o.finish()).max() < 1e-13
    g.close()


def test_lin_calls_need_raw_coefficients(cfd):
    from cfd_b200._lib import CfdError
    c = cfd.CavityFVM(10, 10, order="redblack")
    with pytest.raises(CfdError):
        cfd.LinearSystem5.wesweep(type("S", (), {"lib": c.lib, "_h": c._h})())
    c.close()
