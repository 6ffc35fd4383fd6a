"""oracle/lin_oracle.c (scalar-transport demos): the reference ships no outputs for these programs, so the
restatement is checked for internal consistency and for the behaviour the programs' own comments describe."""
import numpy as np
import pytest

from oracle import oracle


@pytest.mark.parametrize("isch", [1, 2])
def test_stag_point_sweep_directions_agree(isch):
    a, b = oracle.StagPoint(20, 20, isch=isch, isol=1), oracle.StagPoint(20, 20, isch=isch, isol=2)
    ia, ib = a.solve(), b.solve()
    fa, fb = a.finish(), b.finish()
    assert 100 < ia < 1000 and 100 < ib < 1000 and a.rsm < 1e-4 and b.rsm < 1e-4
    assert abs(fa - fb) < 2e-4 * abs(fa)                 # both loops stop at rsm < 1e-4 of the same system
    assert 0.0 <= a.phi.min() and a.phi.max() <= 1.0     # bounded (UDS) / nearly bounded (CDS at this Peclet number)
    assert -0.16 < fa < -0.13


def test_stag_point_red_black_order_reaches_the_same_solution():
    a, b = oracle.StagPoint(16, 16, isch=1, isol=1, order=0), oracle.StagPoint(16, 16, isch=1, isol=1, order=1)
    for o in (a, b):
        for _ in range(600):
            o.wesweep()
    assert np.abs(a.phi - b.phi).max() < 1e-12


def test_unsteady_schemes_agree_and_approach_the_steady_state():
    out = {}
    for itsch in (1, 2, 3):
        o = oracle.StagPoint(20, 20, isch=2, isol=2)
        for _ in range(50):
            o.step(itsch)
        out[itsch] = o.phi.copy()
    assert np.abs(out[2] - out[3]).max() < 5e-3 and np.abs(out[1] - out[3]).max() < 5e-3
    assert np.abs(out[2] - out[3]).max() > 0.0


@pytest.mark.parametrize("bc", [1, 2])
def test_oblique_step_profile(bc):
    o = oracle.Oblique(20, 20, east_north_bc=bc)
    it = o.solve()
    phi = o.finish()
    assert it == 1                                        # pure upwinding + a W->E sweep: exact in one pass
    d = np.array([phi[i, o.ni - 2 - i] for i in range(1, o.ni - 2)])
    assert np.all(np.diff(d) <= 1e-15) and d[0] > 0.99 and d[-1] < 0.01      # smeared, monotone step (numerical diffusion)
    assert np.abs(d + d[::-1] - 1.0).max() < 1e-12                           # antisymmetric about the flow diagonal
