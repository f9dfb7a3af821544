"""TEST INFRASTRUCTURE (oracle) -- numpy restatement of the reference's pinhole / PSF characterisation helpers; the checker
for the device implementation (csrc/calib.cu, sr459.calibration).  Only tests/ may import it.

Mirror of the functions the reference keeps in `data_collection/psf_mtf_utils.py` (`find_peak` :48-51,
`extract_psf` :54-71, `subpixel_centre` :74-78, `radial_average` :81-102, `gauss2d` :105-112,
`fit_gaussian_psf` :115-131, `compute_mtf` :134-170, `mtf_at_fraction` :173-186, `zoom_crop` :189-197) and of
`find_pinhole_center` in `calibration_beam_shift/calibrate_shift_grid.py:66-96`.
Same names, arguments and return values.  These run on 30-pixel ROIs once per calibration, far from the
throughput path, so they are host numpy like the reference's; what differs is that nothing here needs scipy or
OpenCV: the Gaussian smoothing is `analysis.gaussian_filter`, the centre of mass and the image moments are
plain sums, and the bounded 7-parameter Gaussian fit is a projected Levenberg-Marquardt iteration with the
analytic Jacobian (the reference calls scipy's `curve_fit`; both descend to the same least-squares minimum, the
tests pin centre, widths and fit image against the reference's own output).
"""
from __future__ import annotations

import numpy as np

from oracle.analysis_oracle import gaussian_filter, mtf_at_fraction  # noqa: F401  (mtf_at_fraction: identical definition, :173-186)

PSF_CROP_RADIUS = 15            # calibrate_shift_grid.py: radius handed to extract_psf


def find_peak(img, smooth_sigma: float = 2.0):
    """(row, col) of the first maximum of the lightly smoothed image -- psf_mtf_utils.py:48-51."""
    sm = gaussian_filter(img, sigma=smooth_sigma)
    return np.unravel_index(sm.argmax(), sm.shape)


def extract_psf(img, center, radius, bg_percentile: float = 50.0):
    """Background-subtracted ROI around `center`; the background is a percentile of the outer ring
    (pixels at least 0.6 * radius from the ROI centre) -- psf_mtf_utils.py:54-71."""
    r, c = int(center[0]), int(center[1])
    h, w = img.shape
    roi = np.array(img[max(r - radius, 0):min(r + radius + 1, h), max(c - radius, 0):min(c + radius + 1, w)], dtype=np.float64)
    cy, cx = roi.shape[0] // 2, roi.shape[1] // 2
    yy, xx = np.ogrid[:roi.shape[0], :roi.shape[1]]
    ring = ((yy - cy) ** 2 + (xx - cx) ** 2) >= (radius * 0.6) ** 2
    roi -= np.percentile(roi[ring], bg_percentile)
    np.maximum(roi, 0, out=roi)
    return roi


def subpixel_centre(psf):
    """Centre of mass (row, col) of the PSF above 10 % of its peak -- psf_mtf_utils.py:74-78."""
    psf = np.asarray(psf, dtype=np.float64)
    m = np.where(psf > psf.max() * 0.1, psf, 0.0)
    tot = m.sum()
    rows, cols = np.arange(psf.shape[0]), np.arange(psf.shape[1])
    return float((m.sum(axis=1) * rows).sum() / tot), float((m.sum(axis=0) * cols).sum() / tot)


def radial_average(data_2d, center=None, max_radius=None):
    """Azimuthal mean over integer-radius rings. Returns (radii, profile) -- psf_mtf_utils.py:81-102."""
    data_2d = np.asarray(data_2d, dtype=np.float64)
    h, w = data_2d.shape
    cy, cx = (h / 2.0, w / 2.0) if center is None else center
    yy, xx = np.mgrid[:h, :w]
    ring = np.sqrt((xx - cx) ** 2 + (yy - cy) ** 2).astype(int)
    if max_radius is None:
        max_radius = int(min(cy, cx, h - cy, w - cx))
    radii = np.arange(0, max_radius)
    keep = ring < max_radius
    cnt = np.bincount(ring[keep], minlength=max_radius)[:max_radius]
    tot = np.bincount(ring[keep], weights=data_2d[keep], minlength=max_radius)[:max_radius]
    profile = np.where(cnt > 0, tot / np.maximum(cnt, 1), 0.0)
    return radii, profile


def _quad(theta, sx, sy):
    ct, st, s2 = np.cos(theta) ** 2, np.sin(theta) ** 2, np.sin(2 * theta)
    return ct / (2 * sx ** 2) + st / (2 * sy ** 2), -s2 / (4 * sx ** 2) + s2 / (4 * sy ** 2), st / (2 * sx ** 2) + ct / (2 * sy ** 2)


def gauss2d(xy, amp, x0, y0, sigma_x, sigma_y, theta, offset):
    """Rotated 2-D Gaussian, flattened -- psf_mtf_utils.py:105-112."""
    x, y = xy
    a, b, c = _quad(theta, sigma_x, sigma_y)
    dx, dy = x - x0, y - y0
    return (offset + amp * np.exp(-(a * dx ** 2 + 2 * b * dx * dy + c * dy ** 2))).ravel()


def _model_and_jacobian(p, x, y):
    amp, x0, y0, sx, sy, th, off = p
    a, b, c = _quad(th, sx, sy)
    dx, dy = x - x0, y - y0
    e = np.exp(-(a * dx ** 2 + 2 * b * dx * dy + c * dy ** 2))
    g = amp * e
    ct, st, s2, c2 = np.cos(th) ** 2, np.sin(th) ** 2, np.sin(2 * th), np.cos(2 * th)
    da_sx, db_sx, dc_sx = -ct / sx ** 3, s2 / (2 * sx ** 3), -st / sx ** 3
    da_sy, db_sy, dc_sy = -st / sy ** 3, -s2 / (2 * sy ** 3), -ct / sy ** 3
    k = 1 / (2 * sy ** 2) - 1 / (2 * sx ** 2)
    da_th, db_th, dc_th = s2 * k, c2 * k, -s2 * k
    q = lambda da, db, dc: -g * (da * dx ** 2 + 2 * db * dx * dy + dc * dy ** 2)   # noqa: E731
    J = np.stack([e, g * (2 * a * dx + 2 * b * dy), g * (2 * b * dx + 2 * c * dy),
                  q(da_sx, db_sx, dc_sx), q(da_sy, db_sy, dc_sy), q(da_th, db_th, dc_th), np.ones_like(e)], axis=1)
    return off + g, J


def fit_gaussian_psf(psf, max_iter: int = 400):
    """Bounded least-squares fit of gauss2d to the ROI. Returns (params, fit_image) or (None, None)
    -- psf_mtf_utils.py:115-131 (same start point and bounds)."""
    psf = np.asarray(psf, dtype=np.float64)
    h, w = psf.shape
    yy, xx = np.mgrid[:h, :w]
    x, y, z = xx.ravel().astype(np.float64), yy.ravel().astype(np.float64), psf.ravel()
    cy, cx = subpixel_centre(psf)
    pk = psf.max()
    p = np.array([pk, cx, cy, 2.0, 2.0, 0.0, 0.0])
    lo = np.array([0, 0, 0, 0.3, 0.3, -np.pi, -np.inf])
    hi = np.array([pk * 2, w, h, w / 2, h / 2, np.pi, pk * 0.5])
    if not (np.isfinite(p).all() and pk > 0):
        return None, None
    p = np.clip(p, lo, hi)
    f, J = _model_and_jacobian(p, x, y)
    r = f - z
    cost, lam = float(r @ r), 1e-3
    for _ in range(max_iter):
        A, gvec = J.T @ J, J.T @ r
        step_ok = False
        for _ in range(30):
            try:
                dp = np.linalg.solve(A + lam * np.diag(np.maximum(np.diag(A), 1e-12)), -gvec)
            except np.linalg.LinAlgError:
                lam *= 10
                continue
            pn = np.clip(p + dp, lo, hi)
            fn, Jn = _model_and_jacobian(pn, x, y)
            rn = fn - z
            cn = float(rn @ rn)
            if cn < cost:
                step_ok = True
                break
            lam *= 10
        if not step_ok:
            break
        done = (cost - cn) <= 1e-15 * max(cost, 1e-300) and np.max(np.abs(pn - p)) <= 1e-10
        p, r, J, cost, lam = pn, rn, Jn, cn, max(lam / 10, 1e-12)
        if done:
            break
    else:
        return None, None
    return p, gauss2d((xx, yy), *p).reshape(h, w)


def compute_mtf(psf, pixel_pitch_um=None):
    """MTF from the PSF: zero-pad to >= 256, 2-D FFT magnitude, radial average.
    Returns (freq, mtf_radial, mtf_2d, freq_label, nyquist) -- psf_mtf_utils.py:134-170."""
    psf = np.asarray(psf, dtype=np.float64)
    pad = max(256, psf.shape[0], psf.shape[1])
    big = np.zeros((pad, pad))
    r0, c0 = (pad - psf.shape[0]) // 2, (pad - psf.shape[1]) // 2
    big[r0:r0 + psf.shape[0], c0:c0 + psf.shape[1]] = psf
    tot = big.sum()
    if tot > 0:
        big /= tot
    mtf_2d = np.abs(np.fft.fftshift(np.fft.fft2(np.fft.ifftshift(big))))
    if mtf_2d.max() > 0:
        mtf_2d /= mtf_2d.max()
    radii, prof = radial_average(mtf_2d, (pad / 2.0, pad / 2.0), pad // 2)
    cpp = radii.astype(float) / pad
    if pixel_pitch_um is not None:
        return cpp / (pixel_pitch_um * 1e-3), prof, mtf_2d, "cycles/mm", 1.0 / (2.0 * pixel_pitch_um * 1e-3)
    return cpp, prof, mtf_2d, "cycles/pixel", 0.5


def zoom_crop(img, center, half_width):
    """Square crop of img around (row, col) -- psf_mtf_utils.py:189-197."""
    cy, cx = int(round(center[0])), int(round(center[1]))
    h, w = img.shape
    return img[max(cy - half_width, 0):min(cy + half_width + 1, h), max(cx - half_width, 0):min(cx + half_width + 1, w)]


def find_pinhole_center(img, crop_radius: int = PSF_CROP_RADIUS):
    """Sub-pixel (cx, cy) of a bright pinhole: Gaussian fit of the background-subtracted ROI around the peak,
    centre of mass above 10 % of the ROI maximum if the fit fails -- calibrate_shift_grid.py:66-96."""
    gray = np.asarray(img, dtype=np.float64)
    pr, pc = find_peak(gray)
    pr, pc = int(pr), int(pc)
    roi = extract_psf(gray, (pr, pc), crop_radius)
    r0, c0 = max(pr - crop_radius, 0), max(pc - crop_radius, 0)
    popt, _ = fit_gaussian_psf(roi)
    if popt is not None:
        return popt[1] + c0, popt[2] + r0
    t = np.clip(roi - np.median(roi), 0, None)
    t[t < t.max() * 0.1] = 0
    m00 = t.sum()
    if m00 == 0:
        return float(pc), float(pr)
    rows, cols = np.arange(t.shape[0]), np.arange(t.shape[1])
    return float((t.sum(axis=0) * cols).sum() / m00 + c0), float((t.sum(axis=1) * rows).sum() / m00 + r0)
