import sys, numpy as np
sys.path.insert(0, '/root/repo')
import torch, sr459
from oracle import sr_oracle as O
sys.path.insert(0, '/root/repo/tools')
from time_configs import measured7, NOMINAL
rng = np.random.default_rng(5)
h, w = 64, 96
lrs = [rng.uniform(0, 255, (h, w)) for _ in NOMINAL]
for name, psf in (("meas7", measured7()), ("gauss", O.make_gaussian_psf())):
    saa = O.shift_and_add_integer(lrs, NOMINAL)
    ref, eref = O.ibp_integer_shifts(lrs, NOMINAL, psf, saa, n_iter=50)
    for path in (2, 3, 8, 0):
        plan = sr459.Plan(4, h, w, NOMINAL, psf)
        try: plan.set_path(path)
        except Exception as e: print(name, path, "n/a"); continue
        dl = [torch.from_numpy(l.astype(np.float32)).cuda() for l in lrs]
        hr = torch.from_numpy(saa.astype(np.float32)).cuda()
        errs = plan.ibp_(dl, hr, 50)
        d = np.abs(hr.cpu().numpy() - ref).max()
        print(name, "path", path, "max|d| after 50 it", d, "mse rel", abs(errs[-1] / eref[-1] - 1), "mse", errs[-1], eref[-1])
        plan.close()
