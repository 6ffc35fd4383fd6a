"""Pin the CPU oracle (oracle/cavity_oracle.c) to the known-answer table of SURVEY.md 8(c) (tests/golden/
cavity_kat.json).  The reference itself has no tests or golden outputs and cannot be compiled here (Fortran,
no compiler), so this table -- produced by an independent transliteration -- is the only external pin."""
import json
import os

import numpy as np
import pytest

KAT = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "cavity_kat.json")))


def rel(a, b):
    return abs(a - b) / max(abs(b), 1e-300)


def test_cv10_history_and_count(oracle):
    k = KAT["cv10_re100"]
    c = oracle.Cavity(10, 10, 100.0)
    assert (c.imon, c.jmon) == (k["imon"], k["jmon"])
    n, h = c.run(history=True)
    assert n == k["iters"]
    for it, row in k["history"].items():
        got = h[int(it) - 1]
        # residual sums are cancellation-prone near convergence: 1e-3 there, 1e-11 early; monitors 1e-12
        rtol = 1e-11 if int(it) <= 10 else (1e-9 if int(it) == 100 else 2e-3)
        for g, w in zip(got[:3], row[:3]):
            assert rel(g, w) < rtol, (it, g, w)
        for g, w in zip(got[3:], row[3:]):
            assert rel(g, w) < 1e-12, (it, g, w)
    uc, vc = c.centred_vel()
    assert rel(uc[5, 5], k["uc_mon"]) < 1e-12
    assert rel(np.abs(c.u).sum(), k["sum_abs_u"]) < 1e-12
    assert rel(np.abs(c.v).sum(), k["sum_abs_v"]) < 1e-12
    assert rel(np.abs(c.p).sum(), k["sum_abs_p"]) < 1e-12
    assert c.p[5, 5] == 0.0        # the reference cell never moves (SURVEY 8c note)


def test_iter1_analytic_resoru(oracle):
    # 9 lid-adjacent u-cells x an = mu*dx/(dy/2) = 10*0.1/0.05 = 20 (SURVEY 8c sanity)
    c = oracle.Cavity(10, 10, 100.0)
    assert abs(c.assemble_u() - 180.0) < 1e-10


def test_cv32_counts_and_sums(oracle):
    k = KAT["cv32_re100"]
    c = oracle.Cavity(32, 32, 100.0)
    assert c.run() == k["iters"]
    uc, _ = c.centred_vel()
    assert rel(uc[c.imon - 1, :].min(), k["min_centreline_uc"]) < 1e-11
    assert rel(np.abs(c.u).sum(), k["sum_abs_u"]) < 1e-12
    assert rel(np.abs(c.v).sum(), k["sum_abs_v"]) < 1e-12
    assert rel(np.abs(c.p).sum(), k["sum_abs_p"]) < 1e-12
    seq_u = c.u.copy()
    c2 = oracle.Cavity(32, 32, 100.0, order=1)
    assert c2.run() == k["iters_redblack"]
    # red/black lines: different iterates, same fixed point
    assert np.abs(c2.u - seq_u).max() / np.abs(seq_u).max() < 1e-11


@pytest.mark.slow
def test_cv64_count(oracle):
    k = KAT["cv64_re100"]
    c = oracle.Cavity(64, 64, 100.0)
    n, h = c.run(history=True)
    assert n == k["iters"]
    assert rel(h[-1][3], k["final"][3]) < 1e-11 and rel(h[-1][4], k["final"][4]) < 1e-11
    uc, _ = c.centred_vel()
    assert rel(uc[32, 32], k["uc_mon"]) < 1e-11


def test_schemes_and_rectangular(oracle):
    # UDS / CDS / hybrid all run on a non-square grid and keep the lid row untouched
    for iadv in (1, 2, 3):
        c = oracle.Cavity(16, 8, 10.0, iadv=iadv)
        c.run(maxit=50)
        assert np.all(c.u[1:, -1] == 1.0) and np.isfinite(c.u).all() and np.isfinite(c.p).all()
