"""Drop-in ``xtomosim.back_project``: ``get_recon(sino, ct, spec, N_matrix, FOV, ramp) -> (recon_raw, recon_HU)``.

Call sites: /root/reference/main.py:134,168; described at README.md:30-31 ("fan-beam filtered back projection with a
sinc window filter").  Row f1 of SURVEY.md section 8 -- scheduled after the projector and the decomposition.
"""
from __future__ import annotations


def get_recon(sino, ct, spec, N_matrix, FOV, ramp):
    raise NotImplementedError('fan-beam FBP (SURVEY.md section 8 row f1) is not implemented yet')
