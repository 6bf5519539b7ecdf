"""get_sino / get_sinos / M_mono on device-resident phantoms (the form the large synthetic volumes take), and the
invalidation of the prepared workspace when a phantom is edited in place."""
import numpy as np
import pytest
import torch

from dexct import synthetic as S
from xtomosim import forward_project as FP

pytestmark = pytest.mark.gpu


def test_get_sino_with_device_phantom_equals_host_phantom():
    ct, ph, lo, hi = S.make_config('C4', n=64, nz=64, n_views=12)
    assert isinstance(ph.M, np.ndarray)
    FP.clear_cache()
    raw_h, log_h = FP.get_sino(ct, ph, lo)
    ct2, ph2, _, _ = S.make_config('C4', n=64, nz=64, n_views=12, device=torch.device('cuda', 0))
    assert isinstance(ph2.M, torch.Tensor) and ph2.M.is_cuda
    assert np.array_equal(ph2.M.cpu().numpy(), ph.M)
    raw_d, log_d = FP.get_sino(ct2, ph2, lo)
    assert np.array_equal(raw_d, raw_h) and np.array_equal(log_d, log_h)
    both = FP.get_sinos(ct2, ph2, [lo, hi])
    # the fused two-spectrum epilogue and the single-spectrum one sum the bins in a different order (float32): 1e-5
    assert np.allclose(both[0][0], raw_h, rtol=1e-5)
    hu_h, hu_d = ph.M_mono(70.0).get(), ph2.M_mono(70.0).get()
    assert np.array_equal(hu_h, hu_d)


@pytest.mark.parametrize('on_device', [False, True])
def test_in_place_edit_of_a_large_phantom_invalidates_the_workspace(on_device):
    """> 64 MiB of labels: an edit of a few voxels between two get_sino calls must change the sinogram (round 1 keyed the
    cache on every 4099th byte and served stale layouts)."""
    n, nz = 512, 272                                                      # 68 MiB
    lab = np.zeros((nz, n, n), dtype=np.uint8)
    lab[:, 150:360, 120:400] = 1
    ct, ph, lo, hi = S.make_config('C4', n=64, nz=64, n_views=4)
    from xtomosim.system import VoxelPhantom
    M = torch.as_tensor(lab, device='cuda') if on_device else lab
    big = VoxelPhantom('big', None, None, n, n, nz, dx=0.1, dy=0.1, dz=0.1, labels=M, materials=ph.materials)
    FP.clear_cache()
    raw0, _ = FP.get_sino(ct, big, lo)
    raw0b, _ = FP.get_sino(ct, big, lo)
    # the second call is served from the stored lengths (poly_detect): same values, another summation order in float32
    assert np.allclose(raw0, raw0b, rtol=1e-5)
    big.M[nz // 2 - 3:nz // 2 + 3, 250:262, 251:263] = 7                   # a small dense insert near the centre, in place
    raw1, _ = FP.get_sino(ct, big, lo)
    assert not np.allclose(raw0, raw1, rtol=1e-3)
    assert (raw1 <= raw0 * (1 + 1e-5)).all() and raw1.min() < raw0.min()
    FP.clear_cache()
    raw2, _ = FP.get_sino(ct, big, lo)                                     # a fresh workspace gives the same as the invalidated one
    assert np.array_equal(raw1, raw2)
    FP.clear_cache()
