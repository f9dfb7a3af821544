"""CPU baseline: the reference's pipeline on the reference's own numerical library (scipy).

TEST / BENCH INFRASTRUCTURE ONLY (see oracle/sr_oracle.py's header for who may import this).

The reference is pure Python whose arithmetic is three scipy calls
(mono_cal_target/run_sr.py:43-44): ndimage.zoom, ndimage.shift and
signal.fftconvolve.  /root/reference does not exist on the GPU box, but scipy does
(same image), so the honest "reference CPU path" there is the same composition of the
same scipy calls.  This port states that composition once, in float64, single threaded
like the reference; bench.py times it (cpu_baseline.kind == "port") and runs one copy
per host core for the all-core figure (the only parallelism the reference admits is
across independent frame sets, mono_barcodes/run_sr.py:301).  tests/test_oracle_pinning.py
checks it against oracle/sr_oracle.py and the reference-generated golden vectors.
"""
from __future__ import annotations

import time

import numpy as np
from scipy import ndimage as _ndi
from scipy.signal import fftconvolve as _fftconvolve


class ScipyPipeline:
    def __init__(self, shifts_yx, psf, factor=2, step=0.5, clip=(0.0, 255.0)):
        self.shifts = [(float(a), float(b)) for a, b in shifts_yx]
        self.psf = np.asarray(psf, dtype=np.float64)
        self.psf_t = self.psf[::-1, ::-1]
        self.f = int(factor)
        self.step = float(step)
        self.clip = clip

    # run_sr.py:279
    def native(self, mean_lr):
        return _ndi.zoom(mean_lr, self.f, order=3)

    # run_sr.py:181-187
    def saa(self, frames):
        total = 0.0
        for frame, (dy, dx) in zip(frames, self.shifts):
            total = total + _ndi.shift(_ndi.zoom(frame, self.f, order=3), (dy * self.f, dx * self.f), order=3, mode="nearest")
        return total / len(frames)

    # run_sr.py:161-165
    def simulate(self, hr, s):
        b = _fftconvolve(hr, self.psf, mode="same")
        return _ndi.shift(b, (s[0] * self.f, s[1] * self.f), order=3, mode="nearest")[:: self.f, :: self.f]

    # run_sr.py:168-178
    def project_back(self, err, s, shape):
        up = np.zeros(shape)
        up[:: self.f, :: self.f] = err
        moved = _ndi.shift(up, (-s[0] * self.f, -s[1] * self.f), order=3, mode="nearest")
        return _fftconvolve(moved, self.psf_t, mode="same")

    # run_sr.py:190-209
    def iterate(self, frames, hr):
        corr = np.zeros_like(hr)
        mse = 0.0
        for frame, s in zip(frames, self.shifts):
            err = frame - self.simulate(hr, s)
            mse += float(np.mean(err * err))
            corr += self.project_back(err, s, hr.shape)
        hr = np.clip(hr + self.step * corr / len(frames), *self.clip)
        return hr, mse / len(frames)

    def ibp(self, frames, hr_init, n_iter):
        hr = np.array(hr_init, dtype=np.float64, copy=True)
        errs = []
        for _ in range(n_iter):
            hr, e = self.iterate(frames, hr)
            errs.append(e)
        return hr, errs


def timed_sample(args):
    """Worker: SAA + k_iter IBP iterations on one synthetic frame set; returns (t_saa, t_iter_mean).

    args = (seed, n, h, w, shifts, psf, k_iter)."""
    seed, n, h, w, shifts, psf, k_iter = args
    rng = np.random.default_rng(seed)
    frames = [rng.uniform(0, 255, (h, w)) for _ in range(n)]
    pipe = ScipyPipeline(shifts, psf)
    t0 = time.perf_counter()
    hr = pipe.saa(frames)
    t1 = time.perf_counter()
    for _ in range(k_iter):
        hr, _ = pipe.iterate(frames, hr)
    t2 = time.perf_counter()
    return (t1 - t0, (t2 - t1) / max(k_iter, 1))


def full_size_frame_sample(args):
    """Worker for bench.py: the reference's work for ONE FRAME of a job at the job's FULL size, timed piecewise.

    A job is n * [zoom + shift] (SAA, run_sr.py:181-187) followed by K * (n * [forward_model + residual +
    back_project] + update/clip) (run_sr.py:197-208); every frame and every iteration repeats the same scipy calls on
    arrays of the same shape, so   t_job = n * t_saa_frame + K * (n * t_frame_iter + t_update)   with nothing
    extrapolated across image sizes (scipy's cost per pixel grows with the working set, which is why the size must be
    the real one).

    args = (seed, h, w, shift_yx, psf, do_saa); returns (t_saa_frame or None, t_frame_iter, t_update)."""
    seed, h, w, s, psf, do_saa = args
    rng = np.random.default_rng(seed)
    frame = rng.uniform(0, 255, (h, w))
    pipe = ScipyPipeline([s], psf)
    t_saa = None
    if do_saa:
        t0 = time.perf_counter()
        hr = _ndi.shift(_ndi.zoom(frame, pipe.f, order=3), (s[0] * pipe.f, s[1] * pipe.f), order=3, mode="nearest")
        t_saa = time.perf_counter() - t0
    else:
        hr = rng.uniform(0, 255, (pipe.f * h, pipe.f * w))
    t0 = time.perf_counter()
    err = frame - pipe.simulate(hr, s)
    mse = float(np.mean(err * err))
    corr = pipe.project_back(err, s, hr.shape)
    t1 = time.perf_counter()
    hr = np.clip(hr + pipe.step * corr / 4.0, *pipe.clip)
    t2 = time.perf_counter()
    return (t_saa, t1 - t0, t2 - t1, mse)
