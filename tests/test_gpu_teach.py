"""Parity of the TEACH-family CUDA path (calcu/calcv/calcp + modu/modv of backward-step-flow-2d through the C ABI)
with the FP64 oracle: bit-exact in sequential-order mode; red/black mode to the same fixed point; and agreement
with the reference's own FP32 arithmetic to single precision."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
COEF = (("ae", "ae"), ("aw", "aw"), ("an", "an"), ("as", "as_"), ("ap", "ap"), ("su", "su"))


def linf_rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


@pytest.mark.parametrize("shape,epsx,indcos", [((40, 30), 1.05, 2), ((64, 20), 1.0, 1), ((100, 100), 1.05, 2), ((24, 300), 1.02, 2)])
def test_assembly_bitwise(cfd, oracle, shape, epsx, indcos):
    ni, nj = shape
    o = oracle.Teach(ni, nj, epsx=epsx, indcos=indcos)
    g = cfd.TeachFVM(ni, nj, epsx=epsx, indcos=indcos, order="sequential")
    assert g.flowin == o.s.flowin
    o.iterate(3)                       # get non-trivial fields
    for name in ("u", "v", "p"):
        g.upload(name, getattr(o, name)[:ni, :])
    o.call("assemble_u")
    g.assemble_u()
    g.synchronize()
    sl = (slice(2, ni - 1), slice(1, nj - 1))
    for gn, on in COEF:
        assert np.array_equal(g.download(gn)[sl], getattr(o, on)[:ni][sl]), ("u", gn)
    assert np.array_equal(g.download("du")[sl], o.du[:ni][sl])
    assert np.array_equal(g.download("u"), o.u[:ni])          # symmetry row and outlet column written by modu
    assert abs(g.residuals()[0] - o.residuals[0]) <= 1e-11 * o.residuals[0]
    o.call("assemble_v")
    g.assemble_v()
    sl = (slice(1, ni - 1), slice(2, nj - 1))
    for gn, on in COEF:
        assert np.array_equal(g.download(gn)[sl], getattr(o, on)[:ni][sl]), ("v", gn)
    assert np.array_equal(g.download("dv")[sl], o.dv[:ni][sl])
    o.call("assemble_p")
    g.assemble_p()
    sl = (slice(1, ni - 1), slice(1, nj - 1))
    for gn, on in COEF:
        assert np.array_equal(g.download(gn)[sl], getattr(o, on)[:ni][sl]), ("p", gn)
    g.close()


@pytest.mark.parametrize("shape", [(40, 30), (100, 100)])
def test_trajectory_sequential_bitwise(cfd, oracle, shape):
    ni, nj = shape
    o = oracle.Teach(ni, nj)
    g = cfd.TeachFVM(ni, nj, order="sequential")
    ro = o.iterate(30)
    rg = g.iterate(30)
    for name in ("u", "v", "p"):
        assert np.array_equal(g.download(name), getattr(o, name)[:ni]), name
    for a, b in zip(rg, ro):
        assert abs(a - b) <= 1e-10 * abs(b)
    g.close()


def test_redblack_converged_and_reference_fp32(cfd, oracle):
    ni, nj = 40, 30
    o = oracle.Teach(ni, nj)
    o.iterate(1500)
    g = cfd.TeachFVM(ni, nj, order="redblack")
    g.iterate(1500)
    for name in ("u", "v", "p"):
        assert linf_rel(g.download(name), getattr(o, name)[:ni]) < 1e-8, name
    f32 = oracle.Teach(ni, nj, dtype=np.float32)     # the reference as written
    f32.iterate(400)
    assert linf_rel(g.download("u"), f32.u[:ni].astype(np.float64)) < 1e-4
    g.close()


def test_leading_dimension_it_larger_than_ni(cfd, oracle):
    """The reference allocates (it,jt) >= (ni,nj) (control.f90:22-25,49-55): uploads/downloads take ld = it."""
    ni, nj, it = 40, 30, 64
    o = oracle.Teach(ni, nj, it=it)
    g = cfd.TeachFVM(ni, nj, order="sequential")
    o.iterate(5)
    g.iterate(5)
    big = np.zeros((it, nj), order="F")
    g.download_ptr("u", big.ctypes.data, it)
    assert np.array_equal(big[:ni], o.u[:ni]) and not big[ni:].any()
    g.close()


def test_config5_shape_properties(cfd):
    """BASELINE.json configs[4] at a reduced but same-aspect size (4096 x 1024, uniform x: the shipped epsx = 1.05
    overflows at these sizes): finite fields, outlet mass balance, symmetry row."""
    ni, nj = 4096, 1024
    g = cfd.TeachFVM(ni, nj, epsx=1.0, order="redblack")
    r = g.iterate(10)
    u = g.download("u")
    assert np.isfinite(u).all() and np.isfinite(r).all()
    flow = (0.5 * (1000.0 + 1000.0) * g.r[1:nj - 1] * g.sns[1:nj - 1] * u[ni - 1, 1:nj - 1]).sum()
    assert abs(flow - g.flowin) < 1e-9 * g.flowin
    g.close()


# ---- k-epsilon / energy: calcte, calced, calct, props, wall functions ---------------------------
def _rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


@pytest.mark.parametrize("shape,order,viscos", [((24, 20), "sequential", 0.8e-3), ((40, 30), "sequential", 0.8e-3),
                                                ((40, 30), "redblack", 0.8e-3), ((40, 30), "sequential", 2.0e-5),
                                                ((40, 30), "redblack", 2.0e-5)])
def test_turbulent_trajectory_vs_oracle(cfd, oracle, shape, order, viscos):
    """The shipped case (k-epsilon + energy, all incal* on): every field and wall array follows the restatement to
    rounding (log / pow / sqrt of the wall laws come from different math libraries, so not bit for bit)."""
    ni, nj = shape
    # viscos = 2e-5 puts the top-wall y+ above 11.63: the log-law branches of modu / modte / modt are exercised too
    o = oracle.TeachTurb(ni, nj, order=0 if order == "sequential" else 1, viscos=viscos)
    g = cfd.TeachFVM(ni, nj, order=order, turbulent=True, viscos=viscos)
    for name in ("vis", "gamh", "te", "ed", "t"):
        assert _rel(g.download(name), getattr(o, name)) < 1e-14, ("initial", name)
    for it in range(4):
        ro = o.iterate_turb(5)
        rg = g.iterate_turb(5)
        for a, b in zip(rg, ro):
            assert abs(a - b) <= 1e-8 * abs(b) + 1e-300, (it, rg, ro)
    tol = 1e-10 if order == "sequential" else 1e-9
    for name in ("u", "v", "p", "te", "ed", "t", "vis", "gen", "gamh"):
        assert _rel(g.download(name), getattr(o, name)) < tol, name
    for name in ("taun", "yplusn"):
        assert _rel(g.wall_download(name)[:ni], getattr(o, name)[:ni]) < tol, name
    for name in ("tauw", "xplusw"):
        assert _rel(g.wall_download(name)[:nj], getattr(o, name)[:nj]) < tol, name
    if viscos < 1e-4 and order == "sequential":
        assert g.wall_download("yplusn").max() > 11.63 and o.yplusn[:ni].max() > 11.63
    g.close()


def test_turbulent_stagewise_vs_oracle(cfd, oracle):
    """One pass, stage by stage, from the initial fields: coefficient arrays of the three scalar equations."""
    ni, nj = 30, 26
    o = oracle.TeachTurb(ni, nj)
    g = cfd.TeachFVM(ni, nj, order="sequential", turbulent=True)
    o.call("calcu"); o.call("calcv"); o.call("calcp")
    g.calcu(); g.calcv(); g.calcp()
    sl = (slice(1, ni - 1), slice(1, nj - 1))
    coef = (("ae", "ae"), ("aw", "aw"), ("an", "an"), ("as", "as_"), ("ap", "ap"), ("su", "su"))
    for stage, solve, phi in (("assemble_te", "calcte", "te"), ("assemble_ed", "calced", "ed"), ("assemble_t", "calct", "t")):
        o.tcall(stage)
        if stage != "assemble_t":            # (T has no assemble-only entry point: compared after its solve)
            getattr(g, stage)()
            for gn, on in coef:
                assert _rel(g.download(gn)[sl], getattr(o, on)[sl]) < 1e-12, (stage, gn)
        for n in range({"te": o.q.nswpk, "ed": o.q.nswpd, "t": o.q.nswpt}[phi]):
            o.lisolv(2, 2, getattr(o, phi))
        getattr(g, solve)()                  # assembles again from the unchanged state, then sweeps
        assert _rel(g.download(phi), getattr(o, phi)) < 1e-11, phi
    o.tcall("props")
    g.props()
    assert _rel(g.download("vis"), o.vis) < 1e-12
    g.close()


@pytest.mark.parametrize("viscos", [0.8e-3, 2.0e-5])
def test_wall_coefficients_postprocessing(cfd, oracle, viscos):
    """Shear-stress coefficient and Nusselt number along the large-duct wall (control.f90:254-272) from the device's wall
    arrays and the wall-adjacent row of te, t, den -- against the same formulas evaluated on the oracle's fields after the
    same 30 passes.  viscos = 2e-5 takes the log-law branch (y+ > 11.63), 0.8e-3 the laminar one."""
    ni, nj = 40, 30
    o = oracle.TeachTurb(ni, nj, order=0, viscos=viscos)
    g = cfd.TeachFVM(ni, nj, order="sequential", turbulent=True, viscos=viscos)
    o.iterate_turb(30)
    g.iterate_turb(30)
    rows = np.array(g.wall_coefficients())
    assert len(rows) == ni - 1 - g.istep + 1
    m = g.model
    rlarge, uin, densit, tin, cpp, tk = 0.0124, 0.5, 1000.0, 0.0, 4.190e3, 6.0e-1
    rsmall = 0.5 * (o.y[g.jstep - 1] + o.y[g.jstep])
    ularge = uin * (rsmall / rlarge) ** 2
    xlref, delt, dyn = rlarge - rsmall, abs(tin - m.twalln), o.yv[nj - 1] - o.y[nj - 2]
    j = nj - 2                                                  # njm1, 0-based
    branches = set()
    for r in rows:
        i = int(r[0])
        ssc = o.taun[i - 1] / (densit * ularge * ularge)
        if o.yplusn[i - 1] <= 11.63:
            hf = tk * abs(m.twalln - o.t[i - 1, j]) / dyn
            branches.add("laminar")
        else:
            uplusn = np.log(m.elog * o.yplusn[i - 1]) / m.cappa
            hf = o.den[i - 1, j] * cpp * np.sqrt(o.te[i - 1, j]) * (tin - o.t[i - 1, j]) / m.prandt / (uplusn + m.pfun)
            branches.add("log")
        want = (i, o.xu[i - 1] / xlref, ssc, abs(hf) * xlref / (tk * delt), o.x[i - 1] / xlref)
        for a, b in zip(r, want):
            assert abs(a - b) <= 1e-9 * abs(b) + 1e-300, (i, r, want)
    assert ("log" in branches) == (viscos < 1e-4)
    # the extracted row itself is the downloaded field's row
    assert np.array_equal(g.extract_row("t", nj - 1), g.download("t")[:, nj - 2])
    g.close()
