"""The C side of the boundary: include/cfd_b200.h compiles as plain C, its structs have the layout the ctypes mirror
assumes, and examples/cavity_driver.c (`program cavity_fvm` call for call over the C ABI) builds against the library.
Without a GPU the driver must stop with the library's "no CUDA device" error; with one it must reproduce the
reference's 232 outer iterations of the default 10 x 10 case."""
import ctypes as C
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "cfd_b200", "lib")


def _build_driver(tmp_path):
    exe = str(tmp_path / "cavity_driver")
    subprocess.run(["gcc", "-O2", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "examples", "cavity_driver.c"), "-L" + LIBDIR, "-lcfd_b200",
                    "-Wl,-rpath," + LIBDIR, "-lm", "-o", exe], check=True)
    return exe


def test_header_struct_layout_matches_ctypes(cfd, tmp_path):
    from cfd_b200 import _lib
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "cfd_b200.h"\nint main(void) {\n'
                   'printf("%zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(cfd_config), sizeof(cfd_heat_config), sizeof(cfd_fdcav_config),'
                   ' sizeof(cfd_turbulence), sizeof(cfd_hook_desc), offsetof(cfd_config, sew_nim1), offsetof(cfd_config, joff),'
                   ' offsetof(cfd_config, raw_coefficients));\nreturn 0; }\n')
    exe = str(tmp_path / "sz")
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), str(src), "-o", exe], check=True)
    got = [int(x) for x in subprocess.run([exe], check=True, capture_output=True, text=True).stdout.split()]
    want = [C.sizeof(_lib.CfdConfig), C.sizeof(_lib.CfdHeatConfig), C.sizeof(_lib.CfdFdcavConfig), C.sizeof(_lib.CfdTurbulence),
            C.sizeof(_lib.CfdHookDesc), _lib.CfdConfig.sew_nim1.offset, _lib.CfdConfig.joff.offset,
            _lib.CfdConfig.raw_coefficients.offset]
    assert got == want


def test_c_driver_builds_and_fails_loudly_without_a_gpu(cfd, tmp_path):
    import torch
    exe = _build_driver(tmp_path)
    if torch.cuda.is_available():
        return
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode != 0 and "no CUDA device" in r.stderr


@pytest.mark.gpu
def test_c_driver_reproduces_the_reference_default_case(cfd, tmp_path):
    exe = _build_driver(tmp_path)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "converged after 232 outer iterations" in r.stdout, r.stdout[-500:]
