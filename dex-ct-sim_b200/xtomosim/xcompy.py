"""Stand-in for ``xcompy`` (NIST XCOM mass-attenuation lookup), which the reference imports as ``xcompy``
(/root/reference/matdecomp.py:7,158) and ``xtomosim.xcompy`` (/root/reference/plots.py:16,138-140).

The real package and its XCOM tables are absent from the reference tree and from this machine, and this
is a host-side table lookup far from the hot path.  ``mixatten`` therefore evaluates a smooth closed-form model per
element -- Klein-Nishina incoherent scattering per electron, a Z- and E-power-law photoelectric term with a K-edge jump,
and above 1.022 MeV unscreened Bethe-Heitler pair production (nuclear field Z^2 plus an electron-field share rising with
energy) -- mixed by weight fraction.  It reproduces water and ICRU bone within 1 % over 20-150 keV (the fit range of the
photoelectric term) and water within 0.3 % of the NIST totals from 1.25 to 10 MeV (ICRU bone -3 % at 6 MeV; the reference's
shipped configuration pairs a 0.1-6 MeV spectrum with an 80 kV one, main.py:101), but it is NOT reference data.  Real tables
can be plugged in with
:func:`register_element_table` or by replacing the whole provider with :func:`set_provider`; everything downstream
takes mu(E) as arrays.
"""
from __future__ import annotations

import re

import numpy as np

# symbol -> (Z, A [g/mol], K-edge [keV])
ELEMENTS = {
    'H': (1, 1.008, 0.0136), 'He': (2, 4.0026, 0.0246), 'Li': (3, 6.94, 0.055), 'Be': (4, 9.0122, 0.111),
    'B': (5, 10.81, 0.188), 'C': (6, 12.011, 0.284), 'N': (7, 14.007, 0.410), 'O': (8, 15.999, 0.543),
    'F': (9, 18.998, 0.697), 'Ne': (10, 20.180, 0.870), 'Na': (11, 22.990, 1.071), 'Mg': (12, 24.305, 1.303),
    'Al': (13, 26.982, 1.560), 'Si': (14, 28.085, 1.839), 'P': (15, 30.974, 2.146), 'S': (16, 32.06, 2.472),
    'Cl': (17, 35.45, 2.822), 'Ar': (18, 39.948, 3.206), 'K': (19, 39.098, 3.608), 'Ca': (20, 40.078, 4.038),
    'Sc': (21, 44.956, 4.492), 'Ti': (22, 47.867, 4.966), 'V': (23, 50.942, 5.465), 'Cr': (24, 51.996, 5.989),
    'Mn': (25, 54.938, 6.539), 'Fe': (26, 55.845, 7.112), 'Co': (27, 58.933, 7.709), 'Ni': (28, 58.693, 8.333),
    'Cu': (29, 63.546, 8.979), 'Zn': (30, 65.38, 9.659), 'Mo': (42, 95.95, 20.0), 'Ag': (47, 107.87, 25.514),
    'I': (53, 126.90, 33.169), 'Ba': (56, 137.33, 37.441), 'Gd': (64, 157.25, 50.239), 'W': (74, 183.84, 69.525),
    'Au': (79, 196.97, 80.725), 'Pb': (82, 207.2, 88.005),
}

_N_A = 6.02214076e23
_R_E2 = 7.940787e-26          # classical electron radius squared [cm^2]
_MEC2 = 510.99895             # keV

_tables = {}                  # symbol -> (E [keV], mu/rho [cm2/g]) registered real data
_provider = None
generation = 0                # bumped whenever the data behind mixatten changes (memoising callers key on it)


def klein_nishina(E_keV):
    """Total Klein-Nishina cross-section per electron [cm^2]."""
    k = np.asarray(E_keV, dtype=np.float64) / _MEC2
    l1 = np.log1p(2.0 * k)
    return 2.0 * np.pi * _R_E2 * ((1.0 + k) / k ** 2 * (2.0 * (1.0 + k) / (1.0 + 2.0 * k) - l1 / k)
                                  + l1 / (2.0 * k) - (1.0 + 3.0 * k) / (1.0 + 2.0 * k) ** 2)


_ALPHA_RE2 = _R_E2 / 137.035999     # fine-structure constant x r_e^2 [cm^2]
_ZETA3 = 1.2020569031595942


def pair_production(E_keV):
    """Unscreened Bethe-Heitler pair cross-section in the nuclear field per Z^2 [cm^2], closed forms of Maximon (J. Res.
    NBS 72B, 1968): the threshold expansion in eps = (2k - 4) / (2 + k + 2 sqrt(2k)) below k = E / mc^2 = 4 and the
    high-energy series in 2/k above; zero below the 1.022 MeV threshold.  Screening is ignored (a few per cent for light
    elements below 10 MeV)."""
    k = np.asarray(E_keV, dtype=np.float64) / _MEC2
    out = np.zeros_like(k)
    lo = (k > 2.0) & (k < 4.0)
    if lo.any():
        kk = k[lo]
        e = (2.0 * kk - 4.0) / (2.0 + kk + 2.0 * np.sqrt(2.0 * kk))
        out[lo] = (2.0 * np.pi / 3.0) * ((kk - 2.0) / kk) ** 3 * (1.0 + e / 2.0 + 23.0 * e ** 2 / 40.0 + 11.0 * e ** 3 / 60.0
                                                                   + 29.0 * e ** 4 / 960.0)
    hi = k >= 4.0
    if hi.any():
        kk = k[hi]
        L = np.log(2.0 * kk)
        r = 2.0 / kk
        out[hi] = (28.0 / 9.0 * L - 218.0 / 27.0
                   + r ** 2 * (6.0 * L - 3.5 + 2.0 / 3.0 * L ** 3 - L ** 2 - np.pi ** 2 / 3.0 * L + 2.0 * _ZETA3 + np.pi ** 2 / 6.0)
                   - r ** 4 * (3.0 / 16.0 * L + 1.0 / 8.0)
                   - r ** 6 * (29.0 / (9.0 * 256.0) * L - 77.0 / (27.0 * 512.0)))
    return _ALPHA_RE2 * out


def _element_model(sym, E):
    Z, A, edge = ELEMENTS[sym]
    compton = _N_A * Z / A * klein_nishina(E)
    # pair production above 1.022 MeV (the reference's shipped configuration uses a 0.1-6 MeV spectrum, main.py:101): nuclear
    # field Z^2, plus the electron field (triplets, threshold 2.044 MeV) as eta(E) Z with eta rising from 0 towards 1
    k = E / _MEC2
    eta = np.where(k > 4.0, 1.0 - np.exp(-(np.maximum(k, 4.0) - 4.0) / 12.0), 0.0)
    compton = compton + _N_A / A * Z * (Z + eta) * pair_production(E)
    # photoelectric (+ coherent) term: c Z^n / A / E^k above the K edge, an eighth of it below.  (c, n, k) were
    # least-squares fitted to seven published NIST values each of water and ICRU cortical bone, 20-150 keV (<1 %).
    pe = 11.8094 * Z ** 4.2336 / A / np.maximum(E, 1e-3) ** 2.9697
    pe = np.where(E >= edge, pe, pe / 8.0)
    return compton + pe


def element_atten(sym, E_keV):
    E = np.asarray(E_keV, dtype=np.float64)
    if sym in _tables:
        te, tm = _tables[sym]
        return np.exp(np.interp(np.log(E), np.log(te), np.log(tm)))
    if sym not in ELEMENTS:
        raise KeyError(f'element {sym!r} unknown to the xcompy stand-in; register a table for it')
    return _element_model(sym, E)


def parse_matcomp(matcomp):
    """'H(11.2)O(88.8)' -> [('H', 0.112), ('O', 0.888)] (weight fractions, normalised)."""
    items = re.findall(r'([A-Z][a-z]?)\(([-+0-9.eE]+)\)', matcomp)
    if not items or ''.join(f'{s}({w})' for s, w in items) != matcomp.replace(' ', ''):
        raise ValueError(f'cannot parse material composition {matcomp!r}')
    w = np.array([float(x) for _, x in items])
    if w.sum() <= 0:
        raise ValueError(f'composition {matcomp!r} has no weight')
    return [(s, wi) for (s, _), wi in zip(items, w / w.sum())]


def mixatten(matcomp, E_keV):
    """Mass attenuation coefficient [cm^2/g] of a mixture given as ``'El(wt%)El(wt%)...'`` at energies E [keV]."""
    if _provider is not None:
        return np.asarray(_provider(matcomp, E_keV), dtype=np.float64)
    E = np.asarray(E_keV, dtype=np.float64)
    out = np.zeros_like(E)
    for sym, w in parse_matcomp(matcomp):
        out = out + w * element_atten(sym, E)
    return out


def register_element_table(sym, E_keV, mu_rho):
    """Use tabulated (e.g. XCOM) data for one element; log-log interpolated."""
    global generation
    _tables[sym] = (np.asarray(E_keV, dtype=np.float64), np.asarray(mu_rho, dtype=np.float64))
    generation += 1


def set_provider(fn):
    """Replace ``mixatten`` wholesale (e.g. with the real xcompy): fn(matcomp, E_keV) -> mu/rho.  None restores the model."""
    global _provider, generation
    _provider = fn
    generation += 1
