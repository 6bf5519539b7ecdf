"""Synthetic workloads of SURVEY.md section 8(d): phantoms, spectra and attenuation tables of the named benchmark configs.

Everything is generated from ``numpy.random.default_rng(20240)``; nothing here touches the GPU.  The reference ships
no phantom (its XCAT data is licensed), so these stand in for it in tests and benchmarks.
"""
from __future__ import annotations

import numpy as np

from xtomosim import xcompy as xc
from xtomosim.system import FanBeamGeometry, VoxelPhantom, xRaySpectrum

SEED = 20240

# label -> name, density [g/cm3], closed-form coefficients of mu/rho(E) = c_C f_KN(E)/f_KN(60) + c_P (E/60)^-3 [cm2/g],
# and a composition string for code paths that go through xcompy instead.
MATERIALS = [
    dict(name='air',         density=0.0012, c_C=0.1870, c_P=0.0010, composition='N(75.5)O(23.2)Ar(1.3)'),
    dict(name='soft tissue', density=1.06,   c_C=0.1960, c_P=0.0086, composition='H(10.2)C(14.3)N(3.4)O(70.8)Na(0.2)P(0.3)S(0.3)Cl(0.2)K(0.3)'),
    dict(name='bone',        density=1.92,   c_C=0.1800, c_P=0.1340, composition='H(3.4)C(15.5)N(4.2)O(43.5)Na(0.1)Mg(0.2)P(10.3)S(0.3)Ca(22.5)'),
    dict(name='fat',         density=0.95,   c_C=0.1990, c_P=0.0040, composition='H(11.4)C(59.8)N(0.7)O(27.8)Na(0.1)S(0.1)Cl(0.1)'),
    dict(name='lung',        density=0.26,   c_C=0.1950, c_P=0.0090, composition='H(10.3)C(10.5)N(3.1)O(74.9)Na(0.2)P(0.2)S(0.3)Cl(0.3)K(0.2)'),
    dict(name='muscle',      density=1.05,   c_C=0.1950, c_P=0.0090, composition='H(10.2)C(14.3)N(3.4)O(71.0)Na(0.1)P(0.2)S(0.3)Cl(0.1)K(0.4)'),
    dict(name='titanium',    density=4.51,   c_C=0.1650, c_P=0.6000, composition='Ti(100)'),
    dict(name='steel',       density=7.80,   c_C=0.1680, c_P=1.0300, composition='Fe(100)'),
]
K = len(MATERIALS)


def mu_table(E_keV, materials=MATERIALS):
    """Closed-form linear attenuation [1/cm], [K, nE] (SURVEY.md section 8d 'mu tables')."""
    E = np.asarray(E_keV, dtype=np.float64)
    kn = xc.klein_nishina(E) / xc.klein_nishina(60.0)
    return np.array([m['density'] * (m['c_C'] * kn + m['c_P'] * (E / 60.0) ** -3.0) for m in materials])


def phantom_labels(n, nz=1, n_inserts=24, seed=SEED, device=None):
    """uint8 labels [nz, n, n]: air, a soft-tissue body ellipse/ellipsoid (semi-axes 0.42 n, 0.30 n voxels; 0.45 nz in z),
    and ``n_inserts`` random ellipsoidal inserts of labels 2..7 inside the body.  With ``device`` the volume is rasterised
    there with torch (same arithmetic, same result) and returned as a tensor: the 1024^3 phantom of config C5 takes minutes
    in numpy."""
    rng = np.random.default_rng(seed)
    c = 0.5 * (n - 1)
    cz = 0.5 * (nz - 1)
    az = max(0.45 * nz, 0.5)
    y, x = np.mgrid[0:n, 0:n].astype(np.float32)
    ins = []
    for _ in range(n_inserts):
        lab = int(rng.integers(2, 8))
        rx, ry = rng.uniform(0.015 * n, 0.07 * n, size=2)
        rz = rng.uniform(0.05, 0.4) * max(nz, 1)
        th = rng.uniform(0, 2 * np.pi)
        rad = np.sqrt(rng.uniform(0, 0.75))
        cx = c + 0.42 * n * rad * np.cos(th)
        cy = c + 0.30 * n * rad * np.sin(th)
        czi = cz + rng.uniform(-0.3, 0.3) * nz
        rot = rng.uniform(0, np.pi)
        ins.append((lab, cx, cy, czi, rx, ry, rz, rot))
    if device is not None:
        return _labels_on_device(n, nz, ins, c, cz, az, device)
    out = np.zeros((nz, n, n), dtype=np.uint8)
    for k in range(nz):
        zz = ((k - cz) / az) ** 2 if nz > 1 else 0.0
        sl = np.zeros((n, n), dtype=np.uint8)
        body = ((x - c) / (0.42 * n)) ** 2 + ((y - c) / (0.30 * n)) ** 2 + zz < 1.0
        sl[body] = 1
        for lab, cx, cy, czi, rx, ry, rz, rot in ins:
            dz2 = ((k - czi) / rz) ** 2 if nz > 1 else 0.0
            if dz2 >= 1.0:
                continue
            xr = (x - cx) * np.cos(rot) + (y - cy) * np.sin(rot)
            yr = -(x - cx) * np.sin(rot) + (y - cy) * np.cos(rot)
            sl[body & ((xr / rx) ** 2 + (yr / ry) ** 2 + dz2 < 1.0)] = lab
        out[k] = sl
    return out


def _labels_on_device(n, nz, ins, c, cz, az, device):
    import torch
    dev = torch.device(device)
    f32, f64 = torch.float32, torch.float64
    x = torch.arange(n, dtype=f32, device=dev)[None, None, :]
    y = torch.arange(n, dtype=f32, device=dev)[None, :, None]
    x64, y64 = x.to(f64), y.to(f64)
    out = torch.empty((nz, n, n), dtype=torch.uint8, device=dev)
    plane = ((x - c) / (0.42 * n)) ** 2 + ((y - c) / (0.30 * n)) ** 2          # float32 like the numpy path
    step = max(1, min(nz, (1 << 26) // (n * n)))
    for k0 in range(0, nz, step):
        k = torch.arange(k0, min(k0 + step, nz), dtype=f64, device=dev)[:, None, None]
        zz = (((k - cz) / az) ** 2).to(f32) if nz > 1 else torch.zeros_like(k, dtype=f32)
        body = plane + zz < 1.0
        sl = body.to(torch.uint8)
        for lab, cx, cy, czi, rx, ry, rz, rot in ins:
            dz2 = ((k - czi) / rz) ** 2 if nz > 1 else torch.zeros_like(k)
            if not bool((dz2 < 1.0).any()):
                continue
            xr = (x64 - cx) * np.cos(rot) + (y64 - cy) * np.sin(rot)
            yr = -(x64 - cx) * np.sin(rot) + (y64 - cy) * np.cos(rot)
            hit = body & ((xr / rx) ** 2 + (yr / ry) ** 2 + dz2 < 1.0) & (dz2 < 1.0)
            sl = torch.where(hit, torch.full_like(sl, lab), sl)
        out[k0:k0 + sl.shape[0]] = sl
    return out


def make_phantom(n=512, nz=1, voxel=0.1, seed=SEED, closed_form=True, device=None):
    """``device``: rasterise the labels there and keep them as a tensor (``VoxelPhantom.M`` is then a CUDA tensor, accepted
    by dexct.ops.Projector; the numpy-facing drop-in entry points need the default host array)."""
    ph = VoxelPhantom(f'synthetic_{n}x{n}x{nz}', None, None, n, n, nz, z_index=nz // 2 if nz > 1 else 0,
                      dx=voxel, dy=voxel, dz=voxel, labels=phantom_labels(n, nz, seed=seed, device=device), materials=MATERIALS)
    if closed_form:
        ph.mu_table = lambda E, _m=MATERIALS: mu_table(E, _m)      # bench configs use the closed-form tables
    return ph


def gaussian_spectrum(name, mean, sigma, E, total=1.0e5):
    I0 = np.exp(-0.5 * ((E - mean) / sigma) ** 2)
    return xRaySpectrum(None, name, E=E, I0=I0 * (total / I0.sum()))


def make_spectra(n_bins=120, e0=20.0):
    """Two Gaussian-shaped spectra (mean 55 / sigma 18 and mean 95 / sigma 25 keV), 1e5 counts each, 1 keV bins."""
    E = e0 + np.arange(n_bins, dtype=np.float64)
    return gaussian_spectrum('low', 55.0, 18.0, E), gaussian_spectrum('high', 95.0, 25.0, E)


def make_scanner(n_ch=900, n_views=1000, n_rows=1, h_iso=0.1, eid=False):
    """Geometry constants of /root/reference/input/params.txt:18-27; ideal photon-counting detector (eta = 1)."""
    return FanBeamGeometry(n_ch, n_views, 0.8230337, 60.0, 100.0, h_iso=h_iso, eid=eid, detector_file=None,
                           N_rows=n_rows, rotation=6.283185)


CONFIGS = {
    # name: phantom n, nz, voxel, channels, rows, views, energy bins, row pitch at iso
    'C2': dict(n=512, nz=1, voxel=0.1, n_ch=900, n_rows=1, n_views=1000, n_bins=120, h_iso=0.1),
    'C4': dict(n=512, nz=512, voxel=0.1, n_ch=900, n_rows=64, n_views=1200, n_bins=120, h_iso=0.1),
    'C5': dict(n=1024, nz=1024, voxel=0.05, n_ch=1800, n_rows=128, n_views=2400, n_bins=150, h_iso=0.05),
}


def make_config(name, n_views=None, nz=None, n=None, device=None):
    """(ct, phantom, spec_low, spec_high) of a named configuration; sizes can be overridden for small test cases."""
    c = dict(CONFIGS[name])
    if n_views is not None:
        c['n_views'] = n_views
    if nz is not None:
        c['nz'] = nz
    if n is not None:
        c['voxel'] = c['voxel'] * c['n'] / n
        c['n'] = n
    ct = make_scanner(c['n_ch'], c['n_views'], c['n_rows'], c['h_iso'])
    ph = make_phantom(c['n'], c['nz'], c['voxel'], device=device)
    lo, hi = make_spectra(c['n_bins'])
    return ct, ph, lo, hi


def gn_problem(n_views=2048, n_bins=2048, n_energy=120, seed=SEED):
    """Config C3: a_true uniform on [0,30] x [0,5] g/cm2, noiseless counts from the decomposition's own forward model,
    water/bone-like bases.  Returns (g [2, n_views, n_bins], i0 [2, nE], mus [2, nE], a_true [n_views, n_bins, 2])."""
    rng = np.random.default_rng(seed)
    E = 20.0 + np.arange(n_energy, dtype=np.float64)
    lo, hi = make_spectra(n_energy)
    i0 = np.stack([lo.I0, hi.I0])
    mus = mu_table(E)[1:3] / np.array([[MATERIALS[1]['density']], [MATERIALS[2]['density']]])   # mass attenuation
    a = np.empty((n_views, n_bins, 2))
    a[..., 0] = rng.uniform(0.0, 30.0, size=(n_views, n_bins))
    a[..., 1] = rng.uniform(0.0, 5.0, size=(n_views, n_bins))
    g = np.empty((2, n_views, n_bins))
    for v in range(n_views):                                      # chunked to bound memory
        att = np.exp(-(a[v, :, 0, None] * mus[0] + a[v, :, 1, None] * mus[1]))
        g[0, v] = att @ i0[0]
        g[1, v] = att @ i0[1]
    return g, i0, mus, a
