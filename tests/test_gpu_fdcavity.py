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
