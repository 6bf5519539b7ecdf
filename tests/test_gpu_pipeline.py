"""The call sequence of /root/reference/main.py:89-178 through the drop-in modules, on the GPU box.

main.py itself cannot run there (the reference tree is not shipped to the GPU box, and it needs matplotlib + LaTeX), so
this test replays its calls one for one -- same functions, same argument order, same use of the results -- on a small
phantom written in the reference's file formats, with the shipped spectrum / detector files (tests/golden/input).
"""
import json
import os

import numpy as np
import pytest

from dexct import synthetic as S

pytestmark = pytest.mark.gpu


def _write_inputs(tmp_path, golden_dir, n=128):
    from xtomosim.system import write_materials
    lab = S.phantom_labels(n)
    (tmp_path / 'ph.bin').write_bytes(lab.tobytes())
    mats = [dict(m) for m in S.MATERIALS]
    write_materials(tmp_path / 'mats.csv', mats)
    params = {"RUN_ID": "t", "forward_project": True, "back_project": True, "phantom_type": "voxel", "phantom_id": "p",
              "phantom_filename": str(tmp_path / 'ph.bin'), "matcomp_filename": str(tmp_path / 'mats.csv'),
              "Nx": n, "Ny": n, "Nz": 1, "dx": 51.2 / n, "dy": 51.2 / n, "dz": 0.1, "z_index": 0,
              "scanner_geometry": "fan_beam", "SID": 60.0, "SDD": 100.0, "N_channels": 200, "N_projections": 180,
              "fan_angle_total": 0.8230337, "rotation_angle_total": 6.283185, "detector_px_height": 1.0,
              "detector_mode": "eid", "detector_filename": os.path.join(golden_dir, 'input/detector/eta_eid_mv.bin'),
              "spectrum_id": "NA", "spectrum_filename": "NA", "N_photons_per_cm2_per_scan": "NA",
              "N_recon_matrix": 128, "FOV_recon": 50.0, "ramp_filter_percent_Nyquist": 0.8}
    (tmp_path / 'params.txt').write_text(json.dumps(params))
    return str(tmp_path / 'params.txt')


def test_main_py_call_sequence(tmp_path, golden_dir):
    from xtomosim.system import read_parameter_file, xRaySpectrum               # main.py:19
    from xtomosim.forward_project import get_sino                                # main.py:20
    from xtomosim.back_project import get_recon                                  # main.py:21
    from matdecomp import get_basismat_sinos                                     # main.py:22

    def load_spectrum(spec_id, ct, dose):                                        # main.py:64-69
        spec = xRaySpectrum(os.path.join(golden_dir, f'input/spectrum/{spec_id}_1mGy_float32.bin'), spec_id)
        spec.rescale_counts(ct.A_iso * dose / ct.N_proj)
        return spec

    for params in read_parameter_file(_write_inputs(tmp_path, golden_dir)):      # main.py:89-90
        run_id, do_fp, do_bp, ct, phantom, _spec, N_matrix, FOV, ramp = params   # main.py:91-94
        assert do_fp and do_bp
        for spec_id1, spec_id2, D1, D2 in [['detunedMV', '80kV', 9, 1]]:         # main.py:101
            dect_specs, dect_sinos = [], []
            for spec_id, dose in ((spec_id1, D1), (spec_id2, D2)):               # main.py:109
                spec = load_spectrum(spec_id, ct, dose)
                dect_specs.append(spec)
                sino_raw, sino_log = get_sino(ct, phantom, spec)                 # main.py:120
                sino_raw.astype(np.float32).tofile(str(tmp_path / f'{spec_id}_raw.bin'))   # main.py:121
                assert sino_raw.shape == sino_log.shape == (ct.N_proj, ct.N_channels)
                assert np.isfinite(sino_raw).all() and (sino_raw > 0).all() and (sino_log > -1e-5).all()
                dect_sinos.append(sino_raw)
                recon_raw, recon_HU = get_recon(sino_log, ct, spec, N_matrix, FOV, ramp)    # main.py:134
                assert recon_raw.shape == recon_HU.shape == (N_matrix, N_matrix)
                assert np.isfinite(recon_raw).all()
                # air above the body but inside the scanned field (the fan covers a disc of 24 cm radius: the corners of
                # the 50 cm square are outside it), tissue inside the body
                air = recon_raw[110:116, 58:70].mean()
                tissue = recon_raw[N_matrix // 2 - 4:N_matrix // 2 + 4, 30:34].mean()
                assert tissue > 0.01 and abs(air) < 0.1 * tissue
                from xtomosim.back_project import effective_mu_water
                mw = effective_mu_water(ct, spec)
                assert np.allclose(recon_HU, 1000.0 * (recon_raw - mw) / mw, rtol=1e-4, atol=0.5)
            spec1, spec2 = dect_specs
            sino1, sino2 = dect_sinos
            matsino1, matsino2 = get_basismat_sinos(ct, sino1, sino2, spec1, spec2, n_iters=50)   # main.py:153
            assert matsino1.shape == matsino2.shape == sino1.shape and matsino1.dtype == np.float64
            assert np.isfinite(matsino1).all() and np.isfinite(matsino2).all()
            air = sino1 >= 0.95 * sino1.max()                                    # matdecomp.py:195-196
            assert (matsino1[air] == 0).all() and (matsino2[air] == 0).all() and matsino1[~air].max() > 1.0
            for matsino in (matsino1, matsino2):                                 # main.py:167-168
                recon_raw, _ = get_recon(matsino, ct, spec1, N_matrix, FOV, ramp)
                assert recon_raw.shape == (N_matrix, N_matrix) and np.isfinite(recon_raw).all()
            # (no level check on the basis images: the rays through the steel and titanium inserts are photon-starved at
            # 80 kV and decompose to huge values of either sign -- in the reference as well -- which streak the image)


@pytest.mark.parametrize('cfg,n_spec', [('C2', 2), ('C4', 2), ('C4', 1)])
def test_chunked_host_paths_equal_the_one_shot_paths(cfg, n_spec, monkeypatch):
    """Large scans take get_sino(s) / get_basismat_sinos in view chunks, with the host copies of finished chunks overlapping the
    kernels of the next ones (forward_project._project_to_host, matdecomp._decompose_host_pipelined).  Forced here on a small
    scan: bit-identical arrays, including a second get_sino call served from the stored lengths and the air mask, whose
    threshold comes from the maximum over ALL chunks."""
    import matdecomp
    from dexct import synthetic as S
    from xtomosim import forward_project as FP
    ct, ph, lo, hi = S.make_config(cfg, n=64, nz=48 if cfg == 'C4' else None, n_views=37)
    specs = [lo, hi][:n_spec]

    def run():
        FP.clear_cache()
        if n_spec == 1:
            first = FP.get_sino(ct, ph, lo)
            second = FP.get_sino(ct, ph, hi)                       # same energy grid: served from the stored lengths
            return [first, second]
        return FP.get_sinos(ct, ph, specs)
    want = run()
    monkeypatch.setattr(FP, 'PIPELINE_MIN_BYTES', 0)
    got = run()
    for (r0, l0), (r1, l1) in zip(want, got):
        assert r1.dtype == np.float64 and r1.shape == r0.shape
        assert np.array_equal(r0, r1) and np.array_equal(l0, l1)
    raw_lo, raw_hi = want[0][0], want[1][0]
    m_want = matdecomp.get_basismat_sinos(ct, raw_lo, raw_hi, lo, hi, n_iters=8)
    monkeypatch.setattr(matdecomp, 'PIPELINE_MIN_BYTES', 0)
    m_got = matdecomp.get_basismat_sinos(ct, raw_lo, raw_hi, lo, hi, n_iters=8)
    for a, b in zip(m_want, m_got):
        assert b.dtype == np.float64 and np.array_equal(a, b)
    assert (m_got[0] == 0).any() and (m_got[0] != 0).any()         # the air mask acted, and not everywhere
    FP.clear_cache()
