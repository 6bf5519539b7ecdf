"""ORACLE (test infrastructure only) -- float64 numpy restatement of equiangular fan-beam filtered back-projection.

Stands in for ``xtomosim.back_project.get_recon`` (/root/reference/main.py:134,168).  PARITY UNPINNED: the reference's
own implementation is in the absent submodule ``gjadick/x-tomo-sim@f21a7e7`` (/root/reference/.gitmodules:1-3) and the
reference holds no test or golden image for it; README.md:30-31 only says "fan-beam filtered back projection with a
sinc window filter".  What is restated here is the textbook algorithm (Kak & Slaney, Principles of Computerized
Tomographic Imaging, ch. 3.4.1, equiangular rays) with the project's geometry conventions (DESIGN.md sections 3 and 7),
anchored by the analytic known-answer tests in tests/test_oracle_fbp.py (a uniform disc reconstructs to its
attenuation).  Only tests/ may import this module.
"""
from __future__ import annotations

import numpy as np


def _s(u):
    """sin(u)^2 / u, continuous at 0."""
    u = np.asarray(u, dtype=np.float64)
    out = np.zeros_like(u)
    nz = u != 0
    out[nz] = np.sin(u[nz]) ** 2 / u[nz]
    return out


def ramp_sinc_response(t, fc):
    """Impulse response at t of H(f) = |f| sinc(f / (2 fc)) for |f| <= fc (0 beyond): the ramp filter with a sinc
    window cut at fc.  Closed form of  2 int_0^fc f sinc(f / 2fc) cos(2 pi f t) df.  fc = Nyquist gives the
    Shepp-Logan kernel  2 / (pi^2 tau^2 (1 - 4 n^2))  at t = n tau."""
    t = np.asarray(t, dtype=np.float64)
    return (2.0 * fc * fc / np.pi) * (_s(0.25 * np.pi + np.pi * fc * t) + _s(0.25 * np.pi - np.pi * fc * t))


def filter_taps(n_ch, dgamma, ramp):
    """g[n], n = -(n_ch-1) .. n_ch-1:  1/2 (gamma / sin gamma)^2 h(gamma) at gamma = n dgamma (Kak & Slaney eq. 3.4-96)."""
    n = np.arange(-(n_ch - 1), n_ch)
    g = n * float(dgamma)
    h = ramp_sinc_response(g, ramp / (2.0 * dgamma))
    w = np.ones_like(g)
    w[n != 0] = (g[n != 0] / np.sin(g[n != 0])) ** 2
    return 0.5 * w * h


def filter_lines(sino, gamma0, dgamma, SID, ramp):
    """Pre-weight (SID cos gamma) and convolve every line [..., n_ch] with the taps, times dgamma."""
    sino = np.asarray(sino, dtype=np.float64)
    n_ch = sino.shape[-1]
    gam = gamma0 + np.arange(n_ch) * dgamma
    pre = sino * (SID * np.cos(gam))
    g = filter_taps(n_ch, dgamma, ramp)
    flat = pre.reshape(-1, n_ch)
    out = np.stack([np.convolve(line, g)[n_ch - 1:2 * n_ch - 1] for line in flat])
    return (out * dgamma).reshape(sino.shape)


def parker_weights(n_views, dbeta, gamma):
    """Short-scan weights [n_views, n_ch] for a scan of n_views views dbeta apart covering less than a full turn
    (Parker 1982, in the form that takes any angular range pi + 2 delta with delta >= the half fan angle -- the "virtual fan"
    of Silver 2000): every line through the object is measured once or twice over such a scan, the conjugate of ray
    (beta, gamma) being (beta + pi + 2 gamma, -gamma) for a source moving towards +beta, and the weights of a conjugate
    pair add up to 1, going smoothly to 0 at both ends of the scan:
        w = sin^2(pi/4 * b / (delta - gamma))                   0 <= b <= 2 delta - 2 gamma
        w = 1                                                   2 delta - 2 gamma <= b <= pi - 2 gamma
        w = sin^2(pi/4 * (pi + 2 delta - b) / (delta + gamma))  pi - 2 gamma <= b <= pi + 2 delta
    with b = v |dbeta| the angle from the start of the scan.  A scan towards -beta is the mirror image: gamma -> -gamma.
    The full-turn formula carries a factor 1/2 (every line measured twice), so these weights are used TWICE over."""
    gamma = np.asarray(gamma, dtype=np.float64)
    total = n_views * abs(float(dbeta))
    delta = 0.5 * (total - np.pi)
    gmax = float(np.max(np.abs(gamma)))
    if not (delta >= gmax - 1e-12 and total < 2.0 * np.pi):
        raise ValueError(f'short-scan weights need pi + fan <= rotation < 2 pi (rotation {total}, fan {2 * gmax})')
    g = (gamma if dbeta >= 0 else -gamma)[None, :]
    b = (np.arange(n_views) * abs(float(dbeta)))[:, None]
    return parker(b, g, delta)


def parker(b, g, delta):
    """The weight function itself: b angle from the start of the scan, g fan angle (scan towards +beta), range pi + 2 delta;
    0 outside the scan."""
    b, g = np.broadcast_arrays(np.asarray(b, dtype=np.float64), np.asarray(g, dtype=np.float64))
    tiny = 1e-300
    w = np.ones(b.shape)
    w1 = np.sin(0.25 * np.pi * b / np.maximum(delta - g, tiny)) ** 2
    w3 = np.sin(0.25 * np.pi * (np.pi + 2.0 * delta - b) / np.maximum(delta + g, tiny)) ** 2
    w = np.where(b < 2.0 * delta - 2.0 * g, w1, w)
    w = np.where(b > np.pi - 2.0 * g, w3, w)
    return np.where((b < 0) | (b > np.pi + 2.0 * delta), 0.0, w)


def fbp_fan(sino, beta, gamma0, dgamma, dbeta, SID, n_matrix, fov, ramp, short_scan=False):
    """sino [n_views, n_ch] (or [n_views, n_rows, n_ch]) -> image [n, n] (or [n_rows, n, n]), float64.

    Pixel (j, i) is centred at x = (i - (n-1)/2) fov/n, y = (j - (n-1)/2) fov/n.  For the source at
    SID (cos beta, sin beta) and rays along -(cos(beta+gamma), sin(beta+gamma)): A = SID - (x cos b + y sin b),
    B = y cos b - x sin b, gamma' = atan2(-B, A), L^2 = A^2 + B^2.  The filtered line is interpolated linearly between
    channel centres and continued by zero beyond the first/last channel (so it tapers to 0 over one channel).
    """
    sino = np.asarray(sino, dtype=np.float64)
    squeeze = sino.ndim == 2
    if squeeze:
        sino = sino[:, None, :]
    n_views, n_rows, n_ch = sino.shape
    if short_scan:
        sino = sino * (2.0 * parker_weights(n_views, dbeta, gamma0 + np.arange(n_ch) * dgamma))[:, None, :]
    q = filter_lines(sino, gamma0, dgamma, SID, ramp)
    qp = np.zeros((n_views, n_rows, n_ch + 2))
    qp[..., 1:-1] = q
    c = (np.arange(n_matrix) - 0.5 * (n_matrix - 1)) * (fov / n_matrix)
    X, Y = np.meshgrid(c, c)                       # X varies along axis 1 (i), Y along axis 0 (j)
    img = np.zeros((n_rows, n_matrix, n_matrix))
    for v in range(n_views):
        cb, sb = np.cos(beta[v]), np.sin(beta[v])
        A = SID - (X * cb + Y * sb)
        B = Y * cb - X * sb
        u = (np.arctan2(-B, A) - gamma0) / dgamma
        u = np.clip(u, -1.0, float(n_ch))
        fl = np.minimum(np.floor(u), n_ch - 1.0)
        i0 = fl.astype(np.int64)
        f = u - fl
        w = 1.0 / (A * A + B * B)
        for r in range(n_rows):
            line = qp[v, r]
            img[r] += (line[i0 + 1] + f * (line[i0 + 2] - line[i0 + 1])) * w
    img *= abs(dbeta)                              # the sum over views is a quadrature in |d beta|, whichever way the source turns
    return img[0] if squeeze else img


def hounsfield(image, mu_water):
    return 1000.0 * (np.asarray(image, dtype=np.float64) - mu_water) / mu_water


def disc_sinogram(beta, gamma, SID, centre, radius, mu):
    """Exact line integrals of a uniform disc for the project's ray geometry: [n_views, n_ch]."""
    b = np.asarray(beta)[:, None]
    g = np.asarray(gamma)[None, :]
    Sx, Sy = SID * np.cos(b), SID * np.sin(b)
    ex, ey = -np.cos(b + g), -np.sin(b + g)
    d = (centre[0] - Sx) * ey - (centre[1] - Sy) * ex          # signed distance of the centre from the ray
    return 2.0 * mu * np.sqrt(np.maximum(radius * radius - d * d, 0.0))
