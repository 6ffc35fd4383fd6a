"""The TEACH-T (backward step) restatement: no reference output exists to pin it (SURVEY 8c has no table for this
program), so these are consistency checks -- FP64 vs the reference's own FP32 arithmetic, mass conservation at
the outlet, red/black vs sequential fixed point -- plus bit-parity of the host-side grid/init mirror."""
import numpy as np


def test_f64_and_reference_f32_agree_to_single_precision(oracle):
    a = oracle.Teach(40, 30)
    b = oracle.Teach(40, 30, dtype=np.float32)
    ra, rb = a.iterate(150), b.iterate(150)
    assert np.abs(a.u - b.u).max() < 2e-6 * np.abs(a.u).max()
    assert np.abs(a.v - b.v).max() < 1e-5 * max(np.abs(a.v).max(), 1e-3)
    assert np.abs(a.p - b.p).max() < 1e-3 * np.abs(a.p).max()
    assert ra[2] < 1e-4 and rb[2] < 1e-4


def test_outlet_mass_balance_and_symmetry(oracle):
    a = oracle.Teach(50, 40)
    a.iterate(300)
    ni, nj = a.ni, a.nj
    flow = sum(0.5 * (a.den[ni - 1, j] + a.den[ni - 2, j]) * a.r[j] * a.sns[j] * a.u[ni - 1, j] for j in range(1, nj - 1))
    assert abs(flow - a.s.flowin) < 1e-10 * a.s.flowin          # modu: global mass balance (promod.f90:79-94)
    assert np.abs(a.u[:ni - 1, 0] - a.u[:ni - 1, 1]).max() < 1e-5  # symmetry axis u(i,1) = u(i,2) (set before the sweeps)
    assert a.u[2, a.jstep + 3] < 0.0 or a.u[3, a.jstep + 3] < 0.0  # recirculation behind the step


def test_redblack_same_fixed_point(oracle):
    a, b = oracle.Teach(40, 30), oracle.Teach(40, 30, order=1)
    a.iterate(1500)
    b.iterate(1500)
    assert np.abs(a.u - b.u).max() < 1e-12


def test_host_grid_and_init_match_oracle_bitwise(cfd, oracle):
    for ni, nj, epsx, indcos in ((40, 30, 1.05, 2), (64, 20, 1.0, 1), (100, 100, 1.05, 2)):
        o = oracle.Teach(ni, nj, epsx=epsx, indcos=indcos)
        x, y = cfd.teach_grid(ni, nj, 0.0124, epsx)
        assert np.array_equal(x, o.x[:ni]) and np.array_equal(y, o.y[:nj])
        g = cfd.teach_init(x, y, indcos)
        for k, v in g.items():
            assert np.array_equal(v, getattr(o, k)[:len(v)]), k


def test_turbulent_restatement_precisions_agree_and_converge(oracle):
    """teach_turb.inc (k-epsilon + energy + wall functions): the FP64 oracle and the FP32 build (the reference's own
    arithmetic) follow the same trajectory to single precision; residuals fall; wall y+ leaves the laminar branch."""
    import numpy as np
    o64 = oracle.TeachTurb(40, 30)
    o32 = oracle.TeachTurb(40, 30, dtype=np.float32)
    # initial effective viscosity = laminar + turbulent part from the inlet k, eps (props, general.f90:165-190)
    te0, ed0 = 0.03 * 0.5 ** 2, (0.03 * 0.5 ** 2) ** 1.5 / (0.005 * 0.0124)
    v0 = 0.7 * (1000.0 * te0 ** 2 * 0.09 / ed0 + 0.8e-3) + 0.3 * 0.8e-3
    assert abs(o64.vis[5, 5] - v0) < 1e-15
    r_first = o64.iterate_turb(20)
    o32.iterate_turb(20)
    r_last = o64.iterate_turb(80)
    o32.iterate_turb(80)
    assert all(np.isfinite(r_last)) and r_last[0] < r_first[0] and r_last[3] < r_first[3]
    for name in ("u", "te", "t", "vis"):
        a, b = getattr(o64, name), getattr(o32, name)
        assert np.abs(a - b).max() <= 2e-5 * np.abs(a).max(), name
    assert (o64.t >= -1e-9).all() and o64.t.max() <= 100.0 + 1e-9          # between inlet and wall temperature
    assert o64.te.min() >= 0.0 and (o64.ed[1:-1, 1:-1] > 0).all()
    assert o64.xplusw[:30].max() > 11.63                                     # step wall: log-law branch already active
    assert o64.yplusn[1:39].max() < 11.63                                    # top wall: still the laminar branch ...
    lo = oracle.TeachTurb(40, 30, viscos=2.0e-5)
    r = lo.iterate_turb(20)
    assert all(np.isfinite(r)) and lo.yplusn[1:39].max() > 11.63             # ... the log law needs a lower viscosity
