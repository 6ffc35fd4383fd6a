"""Host-side result writers (cavity_fvm.f90:567-607): Fortran es14.7 formatting and file structure."""
import numpy as np

from cfd_b200 import tecplot


def test_es14_7_matches_fortran_edit_descriptor():
    assert tecplot.es14_7(-0.15782975450135306) == "-1.5782975E-01"
    assert tecplot.es14_7(0.0) == " 0.0000000E+00"
    assert tecplot.es14_7(1.0) == " 1.0000000E+00"
    assert tecplot.es14_7(2793.942303605858) == " 2.7939423E+03"
    assert tecplot.es14_7(9.99999999e-11) == " 1.0000000E-10"
    assert all(len(tecplot.es14_7(v)) == 14 for v in (1e-300, -3.5e7, 12345.678))


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
