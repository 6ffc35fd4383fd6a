"""Stencil passes of the explicit FD cavity on the GPU vs the oracle: bit-exact fields (the SIP solve of p' runs
in the oracle on the host between the two device calls -- device SIP is SURVEY 8f-1 'next')."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape", [(11, 11), (33, 33), (40, 24), (24, 40), (257, 257)])
def test_fd_cavity_steps_bitwise(cfd, oracle, shape):
    ni, nj = shape
    o = oracle.FDCavity(ni, nj, dt=1.0e-4 if ni < 100 else 2.0e-6)
    g = cfd.FDCavity(ni, nj, dt=1.0e-4 if ni < 100 else 2.0e-6)

    def solve_pp(su):
        assert np.array_equal(su, o.su)           # same source -> the oracle's sipsol result is the answer
        o.solve_pp()
        return o.pp

    for it in range(25):
        o.predict()
        err_g = g.step(solve_pp)
        err_o = o.correct()
        assert abs(err_g - err_o) <= 1e-12 * abs(err_o)
    for n in ("u", "v", "p", "uo", "vo", "po"):
        assert np.array_equal(g.download(n), getattr(o, n)), n
    g.close()


def test_fd_cavity_converges_to_reference_default(cfd, oracle):
    """ni = nj = 11, Re = 10, dt = 1e-4 (the shipped literals): both reach the same steady state."""
    o = oracle.FDCavity()
    g = cfd.FDCavity()

    def solve_pp(su):
        o.su[:, :] = su
        o.pp[:, :] = 0.0
        o.solve_pp()
        return o.pp

    for it in range(3000):
        e = g.step(solve_pp)
    for it in range(3000):
        o.step()
    assert np.abs(g.download("u") - o.u).max() < 1e-12
    g.close()


@pytest.mark.parametrize("shape", [(11, 11), (33, 33), (40, 24), (24, 40), (70, 130), (257, 257)])
def test_device_sip_matches_sipsol(cfd, oracle, shape):
    """Stone's SIP on the device (band-pipelined wavefront) against the restated sipsol: same iteration count,
    same bits in pp."""
    ni, nj = shape
    dt = 1.0e-4 if ni < 100 else 2.0e-6
    o = oracle.FDCavity(ni, nj, dt=dt)
    g = cfd.FDCavity(ni, nj, dt=dt)
    for it in range(4):
        o.predict()
        g.predict()
        n_o = o.solve_pp()
        n_g = g.solve_pp()
        assert n_g == n_o, (it, n_g, n_o)
        assert np.array_equal(g.download("pp"), o.pp)
        e_o, e_g = o.correct(), g.correct()
        assert abs(e_g - e_o) <= 1e-12 * abs(e_o)
    g.close()


def test_fd_cavity_fully_on_device_converges(cfd, oracle):
    o = oracle.FDCavity()
    g = cfd.FDCavity()
    for it in range(3000):
        eg = g.step()
        eo = o.step()
    assert np.array_equal(g.download("u"), o.u) or np.abs(g.download("u") - o.u).max() < 1e-13
    assert eg < 1e-6
    g.close()


@pytest.mark.parametrize("shape,steady", [((11, 11), False), ((33, 33), False), ((40, 24), False), ((21, 21), True), ((130, 70), False)])
def test_implicit_fd_cavity_bitwise(cfd, oracle, shape, steady):
    """fdm_implicit_cavity.f90: linearised coefficient assembly, three SIP solves with array coefficients,
    correction, boundary conditions, continuity norm -- all on the device, same bits as the restatement."""
    ni, nj = shape
    kw = dict(re=10.0, dt=1.0e-4)
    urf = (0.5, 0.5, 0.1) if steady else (1.0, 1.0, 0.5)    # the steady form needs under-relaxation (source :8-16)
    o = oracle.FDCavityImplicit(ni, nj, steady=steady, urf=urf, **kw)
    g = cfd.FDCavity(ni, nj, implicit=True, steady=steady, urfu=urf[0], urfv=urf[1], urfp=urf[2], **kw)
    for it in range(6):
        for w, name in ((0, "assemble_u"), (1, "assemble_v"), (2, "assemble_p")):
            o.call(name)
            g.implicit_assemble(w)
            for f, of in (("ae", "ae"), ("aw", "aw"), ("an", "an"), ("as", "as_"), ("ap", "ap"), ("su", "su")):
                assert np.array_equal(g.download(f), getattr(o, of), equal_nan=True), (it, name, f)
            assert g.implicit_solve(w) == o.solve(w), (it, name)
            phi = ("u", "v", "pp")[w]
            assert np.array_equal(g.download(phi), getattr(o, phi), equal_nan=True), (it, phi)
        e_o, e_g = o.call("correct"), g.implicit_correct()
        assert abs(e_g - e_o) <= 1e-12 * abs(e_o) + 1e-300
    for n in ("u", "v", "p"):
        assert np.array_equal(g.download(n), getattr(o, n), equal_nan=True), n
    assert np.isfinite(g.download("u")).all()
    g.close()


def test_implicit_steady_divergence_is_reproduced_too(cfd, oracle):
    """With urf = 1 the steady form blows up after three passes (the source warns about it, :8-16); the device path
    goes non-finite in exactly the same nodes as the restatement."""
    o = oracle.FDCavityImplicit(21, 21, steady=True)
    g = cfd.FDCavity(21, 21, implicit=True, steady=True, urfp=0.5)
    for _ in range(4):
        o.step()
        g.implicit_step()
    assert np.array_equal(g.download("v"), o.v, equal_nan=True) and np.isnan(o.v).any()
    g.close()


@pytest.mark.parametrize("shape,negate", [((40, 56), True), ((130, 97), True), ((65, 300), False)])
def test_sipsol_drop_in_on_host_arrays(cfd, oracle, shape, negate):
    """cfd_sipsol = `call sipsol(1, i2, 1, j2, -aw, -ae, -as, -an, ap, su, phi)` on host arrays (what the Fortran shim's
    sipsol forwards to): same iteration count and the same bits as the restated sipsol on a random diagonally dominant
    system, with the program's arrays (negated inside) and with arrays passed as they are."""
    from cfd_b200.fdcavity import sipsol
    ld, ncols = shape
    rng = np.random.default_rng(ld * 1000 + ncols)
    aw, ae, as_, an = (np.asfortranarray(0.1 + rng.random(shape)) for _ in range(4))
    ap = np.asfortranarray(aw + ae + as_ + an + 0.05 + 0.5 * rng.random(shape))
    su = np.asfortranarray(rng.random(shape) - 0.5)
    phi0 = np.asfortranarray(rng.random(shape))
    want = phi0.copy(order="F")
    it_want = oracle.sipsol_arr(aw, ae, as_, an, ap, su, want)
    got = phi0.copy(order="F")
    if negate:
        it, rsm = sipsol(aw, ae, as_, an, ap, su, got)
    else:
        it, rsm = sipsol(-aw, -ae, -as_, -an, ap, su, got, negate=False)
    assert it == it_want and 0 < it < 3000
    assert rsm < 1.0e-6
    assert np.array_equal(got, want)
    assert np.array_equal(got[0, :], phi0[0, :]) and np.array_equal(got[:, -1], phi0[:, -1])     # boundary values untouched
