"""cfd_b200 -- B200-native (sm_100a, FP64) hot path of brli3/CFD behind a C ABI (include/cfd_b200.h).

The package holds the CUDA sources (csrc/), the built library (lib/libcfd_b200.so, git-ignored) and thin
host-side mirrors of the reference's program interfaces.  No CPU fallback exists anywhere in the package.
"""
from ._lib import CfdError, LIB_PATH, load  # noqa: F401
from .cavity import CavityFVM, centred_vel, setgrid  # noqa: F401
from .heat import Heat2D  # noqa: F401
from .fdcavity import FDCavity  # noqa: F401
from .teach import TeachFVM, teach_grid, teach_init  # noqa: F401
from .general import BuoyantCavity, CylinderConduction, GeneralFVM, grid  # noqa: F401
from .scalar import LinearSystem5, ObliqueConvection, StagnationPoint, StagnationPointUnsteady, lin_grid  # noqa: F401

__version__ = "0.1.0"
