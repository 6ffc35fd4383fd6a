"""Pin oracle/general_oracle.c (grid, DAPEC/ARIMEAN/HARMEAN, calct, tdma, the shipped `user` module of
cylin_cond.f90) to the known-answer table of SURVEY.md 8(c)-3, and sanity-check the authored buoyancy module."""
import json
import os

import numpy as np

KAT = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "cylinder_kat.json")))


def rel(a, b):
    return abs(a - b) / max(abs(b), 1e-300)


def test_cylinder_case_as_shipped(oracle):
    g = oracle.General("cylinder")
    n, h = g.run(history=True)
    assert n == KAT["iterations"]          # never meets its own 1e-10 stop rule (SURVEY 8c finding)
    for it, (tres, tmon) in KAT["history"].items():
        row = h[int(it) - 1]
        assert rel(row[3], tres) < (1e-6 if int(it) in (1000, 2000) else 1e-12), (it, row[3], tres)
        assert rel(row[4], tmon) < 1e-13, (it, row[4], tmon)
    assert h[-1][3] == KAT["final_tres"] and h[-1][4] == KAT["final_t_mon"]     # bitwise stationary
    assert rel(np.abs(g.t).sum(), KAT["sum_abs_t"]) < 1e-13
    assert rel(g.t.max(), KAT["t_max"]) < 1e-13 and rel(g.t.min(), KAT["t_min"]) < 1e-13
    qw = sum(g.gam[0, j] * (g.y[j] - g.y[j - 1]) * (g.t[0, j] - g.t[1, j]) / (g.x[1] - g.x[0]) for j in range(1, g.nj - 1))
    assert rel(qw, KAT["qwall_west"]) < 1e-12


def test_host_grid_matches_oracle_bitwise(cfd, oracle):
    for user, kw, icrd, lim in (("cylinder", {}, 2, (0.0, 2.0, 0.0, 1.0, 1.0)),
                                ("buoyancy", dict(ni=34, nj=40), 1, (0.0, 1.0, 0.0, 1.0, 1.0))):
        o = oracle.General(user, **kw)
        g = cfd.grid(o.ni, o.nj, *lim, icrd)
        for k, v in g.items():
            assert np.array_equal(v, getattr(o, k)), (user, k)


def test_buoyancy_module_converges_and_is_physical(oracle):
    g = oracle.General("buoyancy", ni=34, nj=34)
    n = g.run(max_iter=5000, tol=1e-10)
    assert 100 < n < 5000
    assert g.t.min() >= -1e-12 and g.t.max() <= 1.0 + 1e-12
    mid = g.nj // 2
    assert g.v[3, mid] > 0.0 > g.v[g.ni - 4, mid]        # up along the hot west wall, down along the cold one
    # red/black lines reach the same fixed point (tight tolerance: the 1e-10 rule is loose for these magnitudes)
    a = oracle.General("buoyancy", ni=22, nj=22)
    b = oracle.General("buoyancy", ni=22, nj=22, order=1)
    a.run(max_iter=3000, tol=1e-17)
    b.run(max_iter=3000, tol=1e-17)
    assert np.abs(a.t - b.t).max() < 1e-11 and np.abs(a.u - b.u).max() < 1e-11 * max(np.abs(a.u).max(), 1e-30) + 1e-15
