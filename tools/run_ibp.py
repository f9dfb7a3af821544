#!/usr/bin/env python3
"""Dev helper: run SAA + K IBP iterations of one config on cuda:0 (target command for ncu).

    python tools/run_ibp.py [--workload c1|c2|c4] [--iters K] [--reps R]
"""
import argparse, os, sys, time
import numpy as np
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import torch
import sr459

NOMINAL = [(+0.5, -0.5), (+0.5, +0.5), (-0.5, -0.5), (-0.5, +0.5)]
CALIBRATED = [(0.47225, -0.43385), (0.4897, 0.4641), (-0.4798, -0.4553), (-0.4856, 0.4369)]

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="c1")
ap.add_argument("--iters", type=int, default=10)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--path", type=int, default=0)
a = ap.parse_args()
rng = np.random.default_rng(459)
if a.workload == "c1":
    h, w, shifts, psf = 1536, 2048, NOMINAL, sr459.make_gaussian_psf()
elif a.workload == "c2":
    h, w, shifts = 768, 1024, CALIBRATED
    psf = sr459.make_gaussian_psf(); psf[0, :] = 0; psf[:, 0] = 0; psf /= psf.sum()
elif a.workload == "c4":
    h, w, shifts = 1536, 2048, NOMINAL
    psf = sr459.make_gaussian_psf(31, 4.0)
elif a.workload == "c1five":
    h, w, shifts, psf = 1536, 2048, [(0.0, 0.0)] + NOMINAL, sr459.make_gaussian_psf()
else:
    raise SystemExit("unknown workload")
lrs = [torch.from_numpy(rng.uniform(0, 255, (h, w)).astype(np.float32)).cuda() for _ in shifts]
plan = sr459.Plan(len(shifts), h, w, shifts, psf)
if a.path:
    plan.set_path(a.path)
hr = torch.empty((2 * h, 2 * w), dtype=torch.float32, device="cuda")
for r in range(a.reps):
    torch.cuda.synchronize()
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    e0.record(); plan.saa(lrs, out=hr); e1.record(); plan.ibp_(lrs, hr, a.iters, want_errors=False); e2.record()
    torch.cuda.synchronize()
    print(f"rep {r}: saa {e0.elapsed_time(e1)*1e3:.1f} us, ibp {e1.elapsed_time(e2)*1e3/a.iters:.1f} us/iter", flush=True)
print("errors", plan.errors(a.iters)[-1])
