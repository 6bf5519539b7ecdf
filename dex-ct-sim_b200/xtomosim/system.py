"""Scanner / phantom / spectrum objects with the names and attributes the reference's callers use.

The reference's own ``xtomosim/system.py`` is in a submodule that is absent from the mounted tree, so
these classes are re-created from their call sites:
  read_parameter_file      /root/reference/main.py:89-94, keys /root/reference/input/params.txt:2-35
  xRaySpectrum             main.py:66-68, matdecomp.py:140,149
  FanBeamGeometry          plots.py:109-111; attributes read at main.py:68, matdecomp.py:146-147
  VoxelPhantom             plots.py:124-126,251,290
They are thin data carriers; all heavy work happens in libdexct.so.
"""
from __future__ import annotations

import csv
import json
import os

import numpy as np

from . import xcompy as xc

WATER = 'H(11.2)O(88.8)'          # plots.py:140


def _read_pairs(filename):
    """float32 file [x_0..x_{n-1}, y_0..y_{n-1}] (spectra and detector efficiencies)."""
    a = np.fromfile(filename, dtype=np.float32)
    if a.size == 0 or a.size % 2:
        raise ValueError(f'{filename}: expected an even number of float32 values, got {a.size}')
    n = a.size // 2
    return a[:n].astype(np.float64), a[n:].astype(np.float64)


class xRaySpectrum:
    """Polychromatic spectrum: ``E`` [keV] bin energies, ``I0`` photons per cm^2 (per mGy until rescaled)."""

    def __init__(self, filename, name, E=None, I0=None):
        self.filename = filename
        self.name = name
        if filename is not None:
            self.E, self.I0 = _read_pairs(filename)
        else:
            self.E = np.asarray(E, dtype=np.float64)
            self.I0 = np.asarray(I0, dtype=np.float64).copy()
        if self.E.shape != self.I0.shape or self.E.ndim != 1:
            raise ValueError('E and I0 must be 1-D arrays of equal length')

    def rescale_counts(self, scale):
        """main.py:68 -- scale to photons per detector pixel per view: A_iso * dose / N_proj."""
        self.I0 = self.I0 * scale
        return self

    def bin_widths(self):
        """matdecomp.py:142 convention: the first bin runs from 0 to E[0]."""
        return np.append([self.E[0]], self.E[1:] - self.E[:-1])


class FanBeamGeometry:
    """Third-generation equiangular fan-beam scanner (cone beam when N_rows > 1).

    Positional signature as used by the reference (plots.py:109-111).  Keyword extensions (not in the reference):
    ``N_rows`` detector rows, ``rotation`` total gantry angle, ``beta0`` start angle; ``noise`` (None = follow
    ``dexct.config.noise``) switches Poisson counting statistics on for this scanner, ``noise_seed`` keys its generator.
    """

    def __init__(self, N_channels, N_proj, gamma_fan, SID, SDD, h_iso=1.0, eid=True, detector_file=None,
                 N_rows=1, rotation=2.0 * np.pi, beta0=0.0, det_E=None, det_eta_E=None, noise=None, noise_seed=None):
        self.noise = noise
        self.noise_seed = noise_seed
        self.noise_stream = 0                                # acquisitions simulated so far (one random stream each)
        self.N_channels = int(N_channels)
        self.N_proj = int(N_proj)
        self.N_rows = int(N_rows)
        self.gamma_fan = float(gamma_fan)
        self.SID = float(SID)
        self.SDD = float(SDD)
        self.h_iso = float(h_iso)
        self.eid = bool(eid)
        self.rotation = float(rotation)
        # "rotation_angle_total": 6.283185 (input/params.txt:27) is 2 pi written with seven digits: a total within 1e-5 of
        # a full turn IS a full turn (project-defined), which keeps view v + N/4 exactly a quarter turn
        # from view v -- the symmetry the back-projection uses
        if abs(abs(self.rotation) - 2.0 * np.pi) < 1e-5:
            self.rotation = float(np.copysign(2.0 * np.pi, self.rotation))
        self.beta0 = float(beta0)
        self.dgamma = self.gamma_fan / self.N_channels
        self.dbeta = self.rotation / self.N_proj
        self.s_iso = self.SID * self.dgamma                  # channel pitch at the isocentre [cm]
        self.A_iso = self.s_iso * self.h_iso                 # pixel area at the isocentre [cm^2] (main.py:68)
        self.gammas = (np.arange(self.N_channels) - 0.5 * (self.N_channels - 1)) * self.dgamma
        self.betas = self.beta0 + np.arange(self.N_proj) * self.dbeta
        self.detector_file = detector_file
        if detector_file is not None:
            self.det_E, self.det_eta_E = _read_pairs(detector_file)
        elif det_E is not None:
            self.det_E = np.asarray(det_E, dtype=np.float64)
            self.det_eta_E = np.asarray(det_eta_E, dtype=np.float64)
        else:                                                # ideal detector
            self.det_E = np.array([1.0, 1.0e4])
            self.det_eta_E = np.array([1.0, 1.0])

    def det_response(self, E):
        """Detector weighting of matdecomp.py:146-148: efficiency, times energy for an energy-integrating detector."""
        r = np.interp(E, self.det_E, self.det_eta_E)
        return r * E if self.eid else r


ScannerGeometry = FanBeamGeometry


class _DeviceArray:
    """Minimal stand-in for the cupy array ``M_mono`` returns in the reference (plots.py:251 calls ``.get()``)."""

    def __init__(self, tensor):
        self.tensor = tensor

    def get(self):
        return self.tensor.detach().cpu().numpy()

    @property
    def shape(self):
        return tuple(self.tensor.shape)


class VoxelPhantom:
    """uint8 material-label volume plus its material table.

    ``filename``: raw uint8, Nz*Ny*Nx values, x fastest.  ``matcomp_filename``: CSV ``id,name,density,composition``
    (the reference's file format is unknown -- its phantom and CSV are licensed XCAT data that are not shipped; this
    format is project-defined), one row per label value, composition in xcompy syntax.
    """

    def __init__(self, name, filename, matcomp_filename, Nx, Ny, Nz, z_index=0, dx=0.1, dy=0.1, dz=0.1,
                 labels=None, materials=None):
        self.name = name
        self.filename = filename
        self.matcomp_filename = matcomp_filename
        self.Nx, self.Ny, self.Nz = int(Nx), int(Ny), int(Nz)
        self.dx, self.dy, self.dz = float(dx), float(dy), float(dz)
        self.z_index = int(z_index)
        if labels is None:
            raw = np.fromfile(filename, dtype=np.uint8)
            if raw.size != self.Nx * self.Ny * self.Nz:
                raise ValueError(f'{filename}: {raw.size} bytes, expected {self.Nx}x{self.Ny}x{self.Nz}')
            labels = raw.reshape(self.Nz, self.Ny, self.Nx)
        if type(labels).__module__.startswith('torch'):      # device-resident labels (large synthetic volumes)
            if str(labels.dtype) != 'torch.uint8':
                raise ValueError('label tensors must be uint8')
            labels = labels.contiguous()
        else:
            labels = np.ascontiguousarray(labels, dtype=np.uint8)
        if labels.ndim == 2:
            labels = labels[None]
        if tuple(labels.shape) != (self.Nz, self.Ny, self.Nx):
            raise ValueError(f'labels have shape {tuple(labels.shape)}, expected {(self.Nz, self.Ny, self.Nx)}')
        self.M = labels
        self.materials = read_materials(matcomp_filename) if materials is None else list(materials)
        if int(self.M.max()) >= len(self.materials):
            raise ValueError(f'label {int(self.M.max())} has no row in the material table ({len(self.materials)} rows)')

    @property
    def K(self):
        return len(self.materials)

    def mu_table(self, E_keV):
        """Linear attenuation [1/cm] of every label at energies E: [K, nE] = density * mixatten(composition)."""
        E = np.asarray(E_keV, dtype=np.float64)
        return np.array([m['density'] * xc.mixatten(m['composition'], E) if m['density'] > 0 else np.zeros_like(E)
                         for m in self.materials])

    def M_mono(self, E0):
        """Ground-truth monoenergetic image of slice z_index in Hounsfield units (plots.py:251,290)."""
        import torch
        mu = self.mu_table(np.array([float(E0)]))[:, 0]
        mu_w = xc.mixatten(WATER, np.array([float(E0)]))[0]
        sl = self.M[self.z_index]
        if isinstance(sl, torch.Tensor):                     # device-resident labels: one slice comes to the host
            sl = sl.cpu().numpy()
        img = 1000.0 * (mu[sl] - mu_w) / mu_w
        dev = 'cuda' if torch.cuda.is_available() else 'cpu'
        return _DeviceArray(torch.as_tensor(img.astype(np.float32), device=dev))


def read_materials(path):
    rows = []
    with open(path, newline='') as f:
        for rec in csv.reader(f):
            if not rec or rec[0].strip().startswith('#') or rec[0].strip().lower() == 'id':
                continue
            rows.append((int(rec[0]), dict(name=rec[1].strip(), density=float(rec[2]), composition=rec[3].strip())))
    rows.sort(key=lambda r: r[0])
    if [r[0] for r in rows] != list(range(len(rows))):
        raise ValueError(f'{path}: material ids must be 0..K-1 without gaps')
    return [r[1] for r in rows]


def write_materials(path, materials):
    with open(path, 'w', newline='') as f:
        w = csv.writer(f)
        w.writerow(['id', 'name', 'density', 'composition'])
        for i, m in enumerate(materials):
            w.writerow([i, m['name'], repr(float(m['density'])), m['composition']])


def read_parameter_file(path):
    """JSON parameter file -> list of ``[run_id, do_fp, do_bp, ct, phantom, spec_or_None, N_matrix, FOV, ramp]``
    (unpacked at main.py:90-94).  The file holds one JSON object or a list of them."""
    with open(path) as f:
        data = json.load(f)
    out = []
    for p in (data if isinstance(data, list) else [data]):
        if p.get('scanner_geometry', 'fan_beam') != 'fan_beam':
            raise ValueError(f"scanner_geometry {p['scanner_geometry']!r} is not supported (fan_beam only)")
        if p.get('phantom_type', 'voxel') != 'voxel':
            raise ValueError(f"phantom_type {p['phantom_type']!r} is not supported (voxel only)")
        ct = FanBeamGeometry(N_channels=p['N_channels'], N_proj=p['N_projections'], gamma_fan=p['fan_angle_total'],
                             SID=p['SID'], SDD=p['SDD'], h_iso=p.get('detector_px_height', 1.0),
                             eid=(p.get('detector_mode', 'eid') == 'eid'), detector_file=p.get('detector_filename'),
                             N_rows=p.get('N_rows', 1), rotation=p.get('rotation_angle_total', 2.0 * np.pi),
                             noise=p.get('noise'), noise_seed=p.get('noise_seed'))
        phantom = VoxelPhantom(p.get('phantom_id', 'phantom'), p['phantom_filename'], p['matcomp_filename'],
                               p['Nx'], p['Ny'], p['Nz'], z_index=p.get('z_index', 0),
                               dx=p.get('dx', 0.1), dy=p.get('dy', 0.1), dz=p.get('dz', 0.1))
        spec = None
        if p.get('spectrum_filename', 'NA') not in ('NA', None, ''):
            spec = xRaySpectrum(p['spectrum_filename'], p.get('spectrum_id', 'spectrum'))
            n = p.get('N_photons_per_cm2_per_scan', 'NA')
            if n not in ('NA', None, ''):
                spec.rescale_counts(float(n) * ct.A_iso / ct.N_proj / max(float(np.sum(spec.I0)), 1e-300))
        out.append([p.get('RUN_ID', os.path.splitext(os.path.basename(path))[0]), bool(p.get('forward_project', True)),
                    bool(p.get('back_project', True)), ct, phantom, spec,
                    int(p.get('N_recon_matrix', 512)), float(p.get('FOV_recon', 50.0)),
                    float(p.get('ramp_filter_percent_Nyquist', 1.0))])
    return out
