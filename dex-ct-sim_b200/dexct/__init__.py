"""dexct -- B200-native (sm_100a) implementation of the data-parallel hot path of gjadick/dex-ct-sim.

Layout of this directory (put it on ``sys.path``, exactly where the reference's own modules would be found):
    xtomosim/      drop-in ``system`` / ``forward_project`` / ``back_project`` / ``xcompy`` modules
    matdecomp.py   drop-in ``get_basismat_sinos``
    dexct/         ctypes binding + device operators + view sharding
    csrc/          the CUDA kernels and the C ABI (``libdexct.so``)
"""
from . import config  # noqa: F401
