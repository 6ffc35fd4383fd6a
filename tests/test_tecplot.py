"""Host-side result writers (cavity_fvm.f90:567-607): Fortran es14.7 formatting and file structure."""
import numpy as np

from cfd_b200 import tecplot


def test_es14_7_matches_fortran_edit_descriptor():
    assert tecplot.es14_7(-0.15782975450135306) == "-1.5782975E-01"
    assert tecplot.es14_7(0.0) == " 0.0000000E+00"
    assert tecplot.es14_7(1.0) == " 1.0000000E+00"
    assert tecplot.es14_7(2793.942303605858) == " 2.7939423E+03"
    assert tecplot.es14_7(9.99999999e-11) == " 1.0000000E-10"
    assert all(len(tecplot.es14_7(v)) == 14 for v in (1e-300, -1e-300, -3.5e7, 12345.678))
    assert tecplot.es14_7(-1.2345678e-100) == "-1.2345678-100"      # three-digit exponent: the letter is dropped
    assert tecplot.i3(12) == " 12" and tecplot.i3(999) == "999" and tecplot.i3(4098) == "***"


def test_result_files_structure(tmp_path, oracle):
    c = oracle.Cavity(10, 10, 100.0)
    c.run()
    uc, vc = c.centred_vel()
    p = tecplot.corner_pressure(c.p, c.ipref, c.jpref)
    assert p[c.ipref - 1, c.jpref - 1] == 0.0
    assert p[0, 0] == p[1, 0] + p[0, 1] - p[1, 1]
    tecplot.write_result_dat(tmp_path / "result.dat", c.x, c.y, uc, vc, p)
    lines = open(tmp_path / "result.dat").read().splitlines()
    assert lines[0] == ' title="Lid Driven Cavity Flow"' and lines[2] == "zone  i= 12  j= 12  f=point"
    assert len(lines) == 3 + 12 * 12 and all(len(l) == 75 for l in lines[3:])
    rec = lines[3 + 5 * 12 + 5].split()            # node (6,6): i fastest
    assert float(rec[2]) == float(tecplot.es14_7(uc[5, 5])) and abs(float(rec[2]) + 0.15782975) < 1e-8
    tecplot.write_centrelines(tmp_path / "uvel.txt", tmp_path / "vvel.txt", c.x, c.y, uc, vc, c.imon, c.jmon)
    assert len(open(tmp_path / "uvel.txt").read().splitlines()) == 12


def test_general_postproc_and_teach_writers(tmp_path):
    import numpy as np
    from cfd_b200 import tecplot as tp
    rng = np.random.default_rng(5)
    ni, nj = 7, 6
    u, v, p, t = (np.asfortranarray(rng.random((ni, nj))) for _ in range(4))
    xdif, ydif = 0.1 + rng.random(ni), 0.1 + rng.random(nj)
    uc, vc, pp, tt = tp.general_postproc(u, v, p, t, xdif, ydif, 3, 3)
    assert uc[0, 2] == u[1, 2] and uc[3, 2] == 0.5 * (u[3, 2] + u[4, 2]) and vc[2, 0] == v[2, 1]
    assert pp[2, 2] == 0.0                                           # referenced to (ipref, jpref)
    assert tt[0, 0] == t[1, 0] + t[0, 1] - t[1, 1] and np.array_equal(tt[1:-1, 1:-1], t[1:-1, 1:-1])
    # linear pressure field is extrapolated exactly on a uniform grid
    xi, yj = np.arange(ni)[:, None] * 1.0, np.arange(nj)[None, :] * 1.0
    lin = np.asfortranarray(2.0 * xi + 3.0 * yj)
    lin_in = lin.copy(); lin_in[0, :] = lin_in[-1, :] = 0; lin_in[:, 0] = lin_in[:, -1] = 0
    _, _, pl, _ = tp.general_postproc(u, v, lin_in, t, np.ones(ni), np.ones(nj), 3, 3)
    assert np.allclose(pl, lin - lin[2, 2], atol=1e-12)
    uc2, vc2 = tp.teach_centred_vels(u, v)
    assert uc2[2, 1] == u[2, 1] + 0.5 * (u[3, 1] - u[2, 1]) and vc2[1, 2] == v[1, 2] + 0.5 * (v[1, 2] - v[1, 3])
    f = tmp_path / "vel.dat"
    tp.teach_velprint(f, 2, 2, np.arange(ni) * 0.5, np.arange(nj) * 0.25, u, v)
    lines = f.read_text().split("\n")
    assert lines[0] == "" and lines[1] == "ZONE   I=  6   j=  5   F=POINT"
    assert len([x for x in lines[2:] if x]) == (ni - 1) * (nj - 1)
    assert lines[2].startswith("  5.000E-01    2.500E-01   ") and len(lines[2]) == 1 + 10 + 3 * 13
    g = tmp_path / "t.dat"
    tp.teach_tecprint(g, 2, 2, np.arange(ni) * 0.5, np.arange(nj) * 0.25, t)
    assert len(g.read_text().split("\n")[2]) == 1 + 10 + 2 * 13
