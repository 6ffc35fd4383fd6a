"""The one piece of published data the reference validates against: its figures overlay the cavity's centre-line u on
Ghia, Ghia & Shin (1982), Table I (lid-driven-cavity-2d/figures/fvm_59_59_re_400.jpeg, re_100_51_51.jpeg; SURVEY.md
section 6).  The reference holds no numbers, so this table is the only pin that is neither the C oracle nor the second
restatement: the converged 256^2 solutions (committed fixtures, written by tools/gen_cavity_converged.py) must pass
through its 17 points.  The hybrid scheme on 256^2 is within 0.0027 (Re=100) / 0.0036 (Re=1000) of the table; the
bound below leaves a little room and would catch any wrong term in the discretisation (a first-order-only convection
scheme is off by > 0.02 at Re=1000)."""
import json
import os

import numpy as np
import pytest

HERE = os.path.dirname(__file__)
GHIA = json.load(open(os.path.join(HERE, "golden", "ghia1982_u_centreline.json")))
TOL = {100: 0.004, 400: 0.006, 1000: 0.005}


def ghia_error(u_line, re):
    """max |u(y_k) - table| with u_line = u(i_centre, 1..nj) on the staggered grid (wall values at both ends, cell
    centres in between: cavity_fvm.f90:387-398), linear interpolation in y."""
    ncv = len(u_line) - 2
    y = np.concatenate([[0.0], (np.arange(ncv) + 0.5) / ncv, [1.0]])
    yg, ug = np.array(GHIA["y"])[::-1], np.array(GHIA["u"][str(re)])[::-1]
    return float(np.abs(np.interp(yg, y, u_line) - ug).max())


@pytest.mark.parametrize("re", [100, 1000])
def test_converged_fixture_passes_through_ghia_table(re):
    f = np.load(os.path.join(HERE, "golden", f"cavity_cv256_re{re}.npz"))
    assert ghia_error(f["u_centre"], re) < TOL[re]


def test_table_is_what_the_literature_prints():
    """Guards the transcription: lid value 1, wall value 0, a single minimum per Reynolds number at the published height."""
    for re, ymin, umin in ((100, 0.4531, -0.21090), (400, 0.2813, -0.32726), (1000, 0.1719, -0.38289)):
        u = np.array(GHIA["u"][str(re)])
        k = int(u.argmin())
        assert u[0] == 1.0 and u[-1] == 0.0 and GHIA["y"][k] == ymin and u[k] == umin


@pytest.mark.gpu
def test_device_cavity_re400_59x59_matches_ghia_like_the_reference_figure():
    """The very case of the reference's figure (FVM SIMPLE, 59 x 59 CV, Re = 400) run on the device to the literal stop
    rule: the centre-line profile lies on Ghia's points as the figure shows (59^2 with the hybrid scheme: within 0.03)."""
    import cfd_b200
    g = cfd_b200.CavityFVM(59, 59, order="redblack")
    g.init_lid_driven(reynolds=400.0)
    n = g.run(maxit=200000, check_every=20)
    assert n < 200000
    u = g.download("u")
    ic = 59 // 2 + 2                       # xu(ic) = (ic - 2) / 59: the u-line nearest x = 0.5
    e0, e1 = ghia_error(u[ic - 1, :], 400), ghia_error(u[ic, :], 400)
    assert min(e0, e1) < 0.03, (e0, e1)
    g.close()


@pytest.mark.gpu
@pytest.mark.parametrize("re", [100, 1000])
def test_device_cavity_256_matches_ghia(re):
    """256^2 on the device (red/black order, literal stop rule): the centre line through cfd_extract_line, against the table."""
    import cfd_b200
    f = np.load(os.path.join(HERE, "golden", f"cavity_cv256_re{re}.npz"))
    g = cfd_b200.CavityFVM(256, 256, order="redblack")
    g.init_lid_driven(reynolds=float(re))
    g.run(maxit=int(2 * f["iters"]), check_every=50)
    u_line = g.extract_line("u", int(f["imon"]))
    assert ghia_error(u_line, re) < TOL[re]
    g.close()


def test_oracle_59x59_re400_lies_on_ghia_like_the_reference_figure(oracle):
    """The C oracle on the reference figure's case (sequential line order, literal stop rule): the only check of the
    oracle against data that was not produced in this repository."""
    o = oracle.Cavity(59, 59, 400.0, fast=True)
    n = o.run(maxit=200000)
    assert n < 200000
    ic = 59 // 2 + 2
    e0, e1 = ghia_error(o.u[ic - 1, :], 400), ghia_error(o.u[ic, :], 400)
    assert min(e0, e1) < 0.03, (e0, e1)
