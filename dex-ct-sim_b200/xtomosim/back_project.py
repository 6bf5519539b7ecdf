"""Drop-in ``xtomosim.back_project``: ``get_recon(sino, ct, spec, N_matrix, FOV, ramp) -> (recon_raw, recon_HU)``.

Call sites: /root/reference/main.py:134,168; described at README.md:30-31 ("fan-beam filtered back projection with a
sinc window filter"); parameters from input/params.txt:33-35.  The reference's implementation is in an absent submodule,
so the algorithm is the textbook equiangular fan-beam FBP, running in csrc/fbp.cu as a gather over
views.  With a process group every rank back-projects its own view range and the partial images are summed with one
all-reduce.
"""
from __future__ import annotations

import numpy as np
import torch

from dexct import config, dist
from dexct.ops import dtype_code, fbp_fan, hounsfield, make_scan
from . import xcompy as xc

WATER = 'H(11.2)O(88.8)'          # /root/reference/plots.py:140
WATER_DENSITY = 1.0


def effective_mu_water(ct, spec):
    """Attenuation of water [1/cm] averaged over the detected spectrum W(E) = I0 eta (E if eid) dE -- the level a
    reconstruction of ``sino_log`` assigns to water, used for the Hounsfield scale."""
    from .forward_project import spectral_weights
    W = spectral_weights(ct, spec)
    mu = WATER_DENSITY * xc.mixatten(WATER, spec.E)
    return float(np.sum(W * mu) / np.sum(W))


def recon_device(sino, ct, N_matrix, FOV, ramp, dtype=None):
    """Device-resident core: ``sino`` (host array or device tensor, ``[N_proj, (N_rows,) N_channels]``, the WHOLE scan)
    -> image tensor ``[N_rows, N, N]``, summed over ranks when views are sharded."""
    code = dtype_code(config.dtype if dtype is None else dtype)
    td = torch.float32 if code == 0 else torch.float64
    # rotation_angle_total (input/params.txt:24): one full turn, or a short scan of at least pi + fan angle, which takes
    # Parker's weights inside the filter kernel; anything else is refused by the library with the numbers in the message
    n_views = ct.N_proj
    rank, ws = dist.world() if config.shard_views else (0, 1)
    v0, nv = dist.shard_range(n_views, rank, ws)
    local_only = ws > 1 and not config.gather_views and sino.shape[0] == nv and nv != n_views
    if (sino.shape[0] != n_views and not local_only) or sino.shape[-1] != ct.N_channels:
        raise ValueError(f'sinogram shape {tuple(sino.shape)} does not match the scanner ({n_views} views x {ct.N_channels} channels)')
    part = sino if local_only else sino[v0:v0 + nv]      # local_only: the caller holds just this rank's view range
    dev = torch.device('cuda', torch.cuda.current_device())
    if isinstance(part, torch.Tensor):
        local = part.to(device=dev, dtype=td)
    else:
        part = np.asarray(part)
        if part.dtype not in (np.float32, np.float64) or not part.flags.c_contiguous:
            part = np.ascontiguousarray(part, dtype=np.float64)
        local = torch.empty(part.shape, dtype=td, device=dev)
        local.copy_(torch.from_numpy(part), non_blocking=True)
    scan = make_scan(ct.N_proj, ct.N_channels, ct.gamma_fan, ct.SID, ct.SDD, n_rows=ct.N_rows, h_iso=ct.h_iso,
                     rotation=ct.rotation, beta0=ct.beta0)
    img = fbp_fan(scan, local, N_matrix, FOV, ramp, view_begin=v0, dtype=code)
    if ws > 1:
        torch.distributed.all_reduce(img)
    return img


def get_recon(sino, ct, spec, N_matrix, FOV, ramp):
    """Reconstruct a fan-beam sinogram.

    ``sino`` ``[N_proj, N_channels]`` line integrals (``sino_log`` of get_sino, or a basis-material sinogram);
    ``N_matrix`` pixels per side, ``FOV`` side of the square [cm], ``ramp`` filter cut-off as a fraction of Nyquist.
    Returns ``(recon_raw, recon_HU)``: float64 arrays ``[N_matrix, N_matrix]`` (``[N_rows, N, N]`` for multi-row
    scanners), attenuation per cm and Hounsfield units relative to water at the effective energy of ``spec``.
    """
    code = dtype_code(config.dtype)
    img = recon_device(sino, ct, N_matrix, FOV, ramp, dtype=code)
    hu = hounsfield(img, effective_mu_water(ct, spec), dtype=code)
    both = torch.stack([img, hu])
    h = torch.empty(both.shape, dtype=torch.float64, pin_memory=True)
    h.copy_(both, non_blocking=True)
    torch.cuda.current_stream().synchronize()
    a = h.numpy()
    return (a[0, 0], a[1, 0]) if ct.N_rows == 1 else (a[0], a[1])
