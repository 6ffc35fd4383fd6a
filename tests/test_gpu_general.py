"""Parity of the general-family CUDA path (calcu/calcv/calcp/calct of buoyancy_cav.f90 through the C ABI) with the
CPU oracle: bit-exact coefficients and fields in sequential-order mode, 1e-8 on converged fields in red/black."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
KAT = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "cylinder_kat.json")))
COEF = (("ae", "ae"), ("aw", "aw"), ("an", "an"), ("as", "as_"), ("ap", "ap"), ("su", "su"))


def linf_rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


def randomise_buoyancy(o, g, seed):
    rng = np.random.default_rng(seed)
    ni, nj = o.ni, o.nj
    for name, scale, base in (("u", 2e-3, 0.0), ("v", 2e-3, 0.0), ("p", 1e-6, 0.0), ("t", 0.4, 0.5), ("den", 0.2, 1.0), ("cp", 0.2, 1.0)):
        a = getattr(o, name)
        a[:, :] = base + scale * (rng.random((ni, nj)) - 0.5)
        g.upload(name, a)


@pytest.mark.parametrize("n", [12, 34, 67])
@pytest.mark.parametrize("iadv", [1, 2, 3, 4])
def test_buoyancy_assembly_bitwise(cfd, oracle, n, iadv):
    o = oracle.General("buoyancy", ni=n, nj=n, iadv=iadv)
    g = cfd.BuoyantCavity(n, n, iadv=iadv, order="sequential")
    randomise_buoyancy(o, g, 3 + iadv)
    sl_u = (slice(2, n - 1), slice(1, n - 1))
    sl_v = (slice(1, n - 1), slice(2, n - 1))
    sl_p = (slice(1, n - 1), slice(1, n - 1))
    for eq, name, sl, dname in ((0, "u", sl_u, "du"), (1, "v", sl_v, "dv"), (2, "p", sl_p, None), (3, "t", sl_p, None)):
        o.call("assemble_" + name)
        getattr(g, "assemble_" + name)()
        for gn, on in COEF:
            assert np.array_equal(g.download(gn)[sl], getattr(o, on)[sl]), (name, gn)
        if dname:
            assert np.array_equal(g.download(dname)[sl], getattr(o, dname)[sl]), dname
        assert abs(g.residuals()[eq] - o.resor[eq]) <= 1e-11 * abs(o.resor[eq]) + 1e-300
        if name == "t":          # the oracle scaled den by cp in place (buoyancy_cav.f90:372); undo for the next round
            o.den[:, :] = o.den / o.cp
    g.close()


def test_cylinder_case_sequential_bitwise_and_kat(cfd, oracle):
    """cylin_cond.f90 exactly as shipped: the known-answer history of SURVEY 8(c)-3 and, after 3000 passes (where
    the reference iteration is bitwise stationary), the same bits as the oracle in every node."""
    g = cfd.CylinderConduction(order="sequential")
    done = 0
    for it in (1, 2, 10, 100, 1000, 2000, 3000):
        res = g.iterate(it - done)
        done = it
        tres, tmon = KAT["history"][str(it)]
        assert abs(res[3] - tres) <= (1e-6 if it in (1000, 2000) else 1e-9) * tres
        assert g.probe("t", 26, 26) == pytest.approx(tmon, rel=1e-13)
    assert g.probe("t", 26, 26) == KAT["final_t_mon"]
    o = oracle.General("cylinder")
    for _ in range(3000):
        o.call("calct")
    assert np.array_equal(g.download("t"), o.t)
    t = g.finish()                                          # boundary() of the last pass, host side
    qw = sum(g.gam[0, j] * (g.y[j] - g.y[j - 1]) * (t[0, j] - t[1, j]) / (g.x[1] - g.x[0]) for j in range(1, g.nj - 1))
    assert qw == pytest.approx(KAT["qwall_west"], rel=1e-12)
    assert np.abs(t).sum() == pytest.approx(KAT["sum_abs_t"], rel=1e-12)
    g.close()


def test_cylinder_redblack_reaches_the_reference_solution(cfd, oracle):
    o = oracle.General("cylinder")
    o.run()
    g = cfd.CylinderConduction(order="redblack")
    g.iterate(4000)
    t = g.finish()
    assert linf_rel(t[1:-1, 1:-1], o.t[1:-1, 1:-1]) < 1e-8
    assert abs(t.max() - KAT["t_max"]) < 1e-8 * KAT["t_max"]
    g.close()


@pytest.mark.parametrize("n", [22, 34])
def test_buoyancy_trajectory_sequential_bitwise(cfd, oracle, n):
    o = oracle.General("buoyancy", ni=n, nj=n)
    g = cfd.BuoyantCavity(n, n, order="sequential")
    for _ in range(25):
        o.call("calcu"); o.call("calcv"); o.call("calcp"); o.call("calct")
    res = g.iterate(25)
    for name in ("u", "v", "p", "t"):
        assert np.array_equal(g.download(name), getattr(o, name)), name
    for a, b in zip(res, o.resor):
        assert abs(a - b) <= 1e-10 * abs(b) + 1e-300
    g.close()


@pytest.mark.parametrize("n", [12, 34, 67])
@pytest.mark.parametrize("iadv", [1, 2, 3, 4])
def test_buoyancy_production_assembly_vs_oracle(cfd, oracle, n, iadv):
    """Red/black contexts use general_fast_kernels.cu and store the equations divided by ap: every coefficient must
    equal the oracle's coefficient / ap to rounding (same randomised fields as the bitwise test)."""
    check_production_assembly(cfd, oracle, n, iadv)


@pytest.mark.parametrize("n,tdma_m", [(2100, 16), (1790, 4), (5200, 13)])
@pytest.mark.parametrize("iadv", [3, 4])
def test_buoyancy_tma_assembly_vs_oracle(cfd, oracle, n, tdma_m, iadv):
    """Long lines take the TMA-staged kernels of general_tma_kernels.cu: T = 160 layout threads (one block per line),
    T = 448 (a line split over two blocks of 224 consumers, padding nodes in the last warp, as at 8192^2); T = 416 has
    no aligned split and stays on the register-window kernels.  Same check as above."""
    check_production_assembly(cfd, oracle, n, iadv, tdma_m=tdma_m)


def check_production_assembly(cfd, oracle, n, iadv, **kw):
    o = oracle.General("buoyancy", ni=n, nj=n, iadv=iadv)
    g = cfd.BuoyantCavity(n, n, iadv=iadv, order="redblack", **kw)
    randomise_buoyancy(o, g, 3 + iadv)
    sl_u = (slice(2, n - 1), slice(1, n - 1))
    sl_v = (slice(1, n - 1), slice(2, n - 1))
    sl_p = (slice(1, n - 1), slice(1, n - 1))
    for eq, name, sl, dname in ((0, "u", sl_u, "du"), (1, "v", sl_v, "dv"), (2, "p", sl_p, None), (3, "t", sl_p, None)):
        o.call("assemble_" + name)
        getattr(g, "assemble_" + name)()
        ap = o.ap[sl]
        for gn, on in COEF:
            if gn == "ap":
                continue
            want = getattr(o, on)[sl] / ap
            got = g.download(gn)[sl]
            assert np.abs(got - want).max() <= 1e-11 * max(1e-300, np.abs(want).max()), (name, gn, np.abs(got - want).max())
        if dname:
            want = getattr(o, dname)[sl]
            assert np.abs(g.download(dname)[sl] - want).max() <= 1e-11 * np.abs(want).max(), dname
        assert abs(g.residuals()[eq] - o.resor[eq]) <= 1e-10 * abs(o.resor[eq]) + 1e-300
        if name == "t":
            o.den[:, :] = o.den / o.cp
    g.close()


def test_buoyancy_redblack_converged_vs_sequential_oracle(cfd, oracle):
    n = 34
    o = oracle.General("buoyancy", ni=n, nj=n)
    o.run(max_iter=6000, tol=1e-17)
    g = cfd.BuoyantCavity(n, n, order="redblack")
    g.iterate(6000)
    for name in ("u", "v", "t"):
        assert linf_rel(g.download(name)[1:-1, 1:-1], getattr(o, name)[1:-1, 1:-1]) < 1e-8, name
    g.close()


def test_buoyancy_130_trajectory_redblack(cfd, oracle):
    n = 130
    o = oracle.General("buoyancy", ni=n, nj=n, order=1)
    for _ in range(10):
        o.call("calcu"); o.call("calcv"); o.call("calcp"); o.call("calct")
    g = cfd.BuoyantCavity(n, n, order="redblack")
    g.iterate(10)
    for name in ("u", "v", "p", "t"):
        assert linf_rel(g.download(name), getattr(o, name)) < 1e-10, name
    g.close()


def test_full_size_8194_buoyancy_properties(cfd):
    """BASELINE.json configs[3] size (8192^2 CV): no CPU answer exists here, so check size-independent properties:
    finite fields, Dirichlet T walls and no-slip walls untouched, p(ipref,jpref) stays the reference 0, the hot wall
    heats its neighbours and nothing leaves [0,1] by more than rounding of the hybrid scheme, and the C-ABI
    subroutine path (calcu+calcv+calcp+calct) gives the same bits as cfd_iterate."""
    n = 8194
    g = cfd.BuoyantCavity(n, n, order="redblack")
    r1 = g.iterate(1)
    r4 = g.iterate(3)
    assert np.isfinite(r1).all() and np.isfinite(r4).all()
    t = g.download("t")
    assert np.isfinite(t).all()
    assert np.all(t[0, :] == 1.0) and np.all(t[-1, :] == 0.0)
    assert t[1, 1:-1].min() > 0.5 and t[-2, 1:-1].max() < 0.5          # conduction has started at both walls
    assert t.min() >= -1e-9 and t.max() <= 1.0 + 1e-9
    for name in ("u", "v"):
        a = g.download(name)
        assert np.isfinite(a).all()
        assert not a[0, :].any() and not a[-1, :].any() and not a[:, 0].any() and not a[:, -1].any(), name
    v = g.download("v")
    assert v[2, n // 2] > 0.0 and v[-2, n // 2] < 0.0                   # up the hot wall, down the cold one
    assert g.probe("p", g.ipref, g.jpref) == 0.0
    del a, v
    h = cfd.BuoyantCavity(n, n, order="redblack")
    for _ in range(4):
        h.calcu(); h.calcv(); h.calcp(); h.calct()
    assert np.array_equal(h.download("t"), t)
    assert np.array_equal(h.download("u"), g.download("u"))
    g.close(); h.close()


def test_full_size_assembly_launches_are_reproducible(cfd):
    """The TMA-staged assembly kernels at the BASELINE configs[3] size (lines split over two blocks, a last warp that
    holds padding nodes): the same launch on the same rough fields must give the same bits every time.  A ring
    slot read too early or too late shows here as a large difference in a few nodes of one line."""
    n = 8194
    g = cfd.BuoyantCavity(n, n, order="redblack")
    rng = np.random.default_rng(5)
    for name, scale, base in (("u", 2e-3, 0.0), ("v", 2e-3, 0.0), ("p", 1e-6, 0.0), ("t", 0.4, 0.5), ("den", 0.2, 1.0), ("cp", 0.2, 1.0)):
        g.upload(name, np.asfortranarray(base + scale * (rng.random((n, n)) - 0.5)))
    for name in ("u", "v", "p", "t"):
        fn = getattr(g, "assemble_" + name)
        fn()
        ref = g.download("su")
        assert np.isfinite(ref[2:-1, 2:-1]).all()
        for rep in range(5):
            fn()
            assert np.array_equal(g.download("su"), ref), (name, rep)
    g.close()


def test_general_rejects_what_the_reference_cannot_do(cfd):
    with pytest.raises(cfd.CfdError, match="ni == nj"):
        cfd.GeneralFVM(40, 30, lsolve_flow=True)
    g = cfd.GeneralFVM(40, 30, lsolve_flow=False)     # T only: fine
    g.close()
