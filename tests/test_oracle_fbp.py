"""The FBP oracle against analytic known answers (CPU).  The reference has no golden image for get_recon (its source is
absent), so the oracle is anchored on closed forms: the Shepp-Logan kernel, the zero DC gain of a ramp filter, and the
reconstruction of a uniform disc from its exact fan-beam line integrals."""
import numpy as np

from oracle import fbp_oracle as F

SID, FAN = 60.0, 0.8230337        # /root/reference/input/params.txt:19,23


def test_ramp_1_is_the_shepp_logan_kernel():
    tau = 0.01
    n = np.arange(-50, 51)
    h = F.ramp_sinc_response(n * tau, 1.0 / (2 * tau))
    assert np.allclose(h, 2.0 / (np.pi ** 2 * tau ** 2 * (1.0 - 4.0 * n ** 2)), rtol=1e-12)


def test_kernel_is_smooth_through_its_removable_singularity():
    # ramp = 0.5 puts t = 1/(4 fc) exactly on tap n = 1
    tau = 0.02
    fc = 0.5 / (2 * tau)
    t = 1.0 / (4 * fc)
    eps = 1e-7
    around = F.ramp_sinc_response(np.array([t - eps, t, t + eps]), fc)
    assert abs(around[1] - 0.5 * (around[0] + around[2])) < 1e-6 * abs(around[1])


def test_taps_have_no_dc_gain_and_are_even():
    for ramp in (1.0, 0.8, 0.37):
        g = F.filter_taps(900, FAN / 900, ramp)
        assert np.array_equal(g, g[::-1])
        # the ramp filter removes the mean; what is left is the truncated (oscillating, 1/n) tail and the (gamma/sin gamma)^2 weight
        assert abs(g.sum()) < 1e-2 * g.max()


def _disc_case(n_ch, n_views, n, fov, centre, radius, ramp):
    dg = FAN / n_ch
    gamma0 = -0.5 * (n_ch - 1) * dg
    dbeta = 2 * np.pi / n_views
    beta = np.arange(n_views) * dbeta
    sino = F.disc_sinogram(beta, gamma0 + np.arange(n_ch) * dg, SID, centre, radius, 0.2)
    img = F.fbp_fan(sino, beta, gamma0, dg, dbeta, SID, n, fov, ramp)
    c = (np.arange(n) - 0.5 * (n - 1)) * (fov / n)
    X, Y = np.meshgrid(c, c)
    r = np.hypot(X - centre[0], Y - centre[1])
    return img, r


def test_uniform_disc_reconstructs_to_its_attenuation():
    img, r = _disc_case(n_ch=400, n_views=360, n=96, fov=30.0, centre=(3.0, -2.0), radius=8.0, ramp=1.0)
    inside, outside = img[r < 6.5], img[(r > 9.5) & (r < 14.0)]
    assert abs(inside.mean() - 0.2) < 0.002 and inside.std() < 0.004
    assert abs(outside.mean()) < 0.002
    # the orientation convention: the disc sits at positive x (axis 1) and negative y (axis 0)
    j, i = np.unravel_index(np.argmax(img > 0.1), img.shape)
    assert img[:, 60:].sum() > img[:, :36].sum() and img[:40].sum() > img[56:].sum()


def test_lower_cutoff_smooths_but_keeps_the_level():
    sharp, r = _disc_case(240, 180, 64, 30.0, (0.0, 0.0), 7.0, 1.0)
    soft, _ = _disc_case(240, 180, 64, 30.0, (0.0, 0.0), 7.0, 0.4)
    assert abs(soft[r < 4.0].mean() - 0.2) < 0.004
    edge = (r > 6.0) & (r < 8.0)
    assert np.abs(np.diff(soft, axis=1))[edge[:, 1:]].max() < np.abs(np.diff(sharp, axis=1))[edge[:, 1:]].max()


def test_multi_row_is_row_by_row_and_hu():
    dg = FAN / 120
    gamma0 = -0.5 * 119 * dg
    beta = np.arange(90) * (2 * np.pi / 90)
    s = F.disc_sinogram(beta, gamma0 + np.arange(120) * dg, SID, (0.0, 0.0), 6.0, 0.2)
    both = F.fbp_fan(np.stack([s, 2 * s], axis=1), beta, gamma0, dg, 2 * np.pi / 90, SID, 32, 25.0, 0.8)
    one = F.fbp_fan(s, beta, gamma0, dg, 2 * np.pi / 90, SID, 32, 25.0, 0.8)
    assert np.allclose(both[0], one, rtol=0, atol=1e-15) and np.allclose(both[1], 2 * one, rtol=1e-13, atol=1e-15)
    assert np.allclose(F.hounsfield(np.array([0.0, 0.2, 0.4]), 0.2), [-1000.0, 0.0, 1000.0])


def test_parker_weights_of_conjugate_rays_add_up_to_one():
    n_ch = 121
    dg = FAN / n_ch
    gam = (np.arange(n_ch) - 0.5 * (n_ch - 1)) * dg
    for total, sign in ((np.pi + FAN, 1), (np.pi + FAN + 0.7, 1), (4.5, -1), (5.9, 1)):
        n_views = 3000
        db = sign * total / n_views
        w = F.parker_weights(n_views, db, gam)
        assert w.min() >= 0.0 and w.max() <= 1.0 and np.all(w[0] < 1e-6)
        delta = 0.5 * (total - np.pi)
        g = (gam if sign > 0 else -gam)[None, :]
        b = (np.arange(n_views) * abs(db))[:, None]
        assert np.array_equal(w, F.parker(b, g, delta))
        # the same line is also seen from (b + pi + 2 g, -g) and from (b - pi + 2 g, -g), whichever lie inside the scan:
        # the weights of all sightings of a line add up to 1
        total_w = w + F.parker(b + np.pi + 2.0 * g, -g, delta) + F.parker(b - np.pi + 2.0 * g, -g, delta)
        assert np.max(np.abs(total_w - 1.0)) < 1e-12, (total, sign)


def test_short_scan_reconstructs_the_disc_in_both_directions():
    n_ch, n, fov = 400, 96, 30.0
    dg = FAN / n_ch
    gamma0 = -0.5 * (n_ch - 1) * dg
    for total, sign, beta0 in ((np.pi + FAN + 0.05, 1, 0.3), (4.4, -1, -1.1), (5.6, 1, 2.0)):
        n_views = 300
        db = sign * total / n_views
        beta = beta0 + np.arange(n_views) * db
        sino = F.disc_sinogram(beta, gamma0 + np.arange(n_ch) * dg, SID, (3.0, -2.0), 8.0, 0.2)
        img = F.fbp_fan(sino, beta, gamma0, dg, db, SID, n, fov, 1.0, short_scan=True)
        c = (np.arange(n) - 0.5 * (n - 1)) * (fov / n)
        X, Y = np.meshgrid(c, c)
        r = np.hypot(X - 3.0, Y + 2.0)
        inside, outside = img[r < 6.5], img[(r > 9.5) & (r < 14.0)]
        assert abs(inside.mean() - 0.2) < 0.003 and inside.std() < 0.006, (total, sign, inside.mean(), inside.std())
        assert abs(outside.mean()) < 0.003
    with np.testing.assert_raises(ValueError):
        F.parker_weights(100, (np.pi + 0.5 * FAN) / 100, gamma0 + np.arange(n_ch) * dg)
