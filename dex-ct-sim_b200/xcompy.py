"""Top-level alias: the reference does ``import xcompy as xc`` (/root/reference/matdecomp.py:7) after putting the
``xtomosim`` directory on ``sys.path`` (main.py:17)."""
from xtomosim.xcompy import *  # noqa: F401,F403
from xtomosim.xcompy import mixatten, parse_matcomp, register_element_table, set_provider  # noqa: F401
