"""Dev probe: fractional-shift IBP, default path (table FIR + polyphase interior) vs the spline stages (path 1)."""
import sys, numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import torch, sr459
from time_configs import measured7
CAL = [(0.47225, -0.43385), (0.4897, 0.4641), (-0.4798, -0.4553), (-0.4856, 0.4369)]
lib = sr459._lib.load()
for (h, w, it) in ((64, 96, 5), (176, 336, 5), (768, 1024, 50)):
    rng = np.random.default_rng(459)
    lrs = [torch.from_numpy(rng.uniform(0, 255, (h, w)).astype(np.float32)).cuda() for _ in CAL]
    plan = sr459.Plan(4, h, w, CAL, measured7())
    saa = plan.saa(lrs)
    out = {}
    for path in (0, 1):
        hr = saa.clone()
        c0 = lib.sr459_launch_count()
        e = plan.set_path(path).ibp_(lrs, hr, it)
        out[path] = (hr.cpu().numpy(), e, lib.sr459_launch_count() - c0)
    d = np.abs(out[0][0] - out[1][0])
    iy, ix = np.unravel_index(np.argmax(d), d.shape)
    print(f"{h}x{w} it={it}: launches {out[0][2]} vs {out[1][2]}; max|d|={d.max():.3e} at ({iy},{ix}); mse {out[0][1][-1]:.4f} vs {out[1][1][-1]:.4f}; "
          f"interior max|d|={d[80:-80, 140:-140].max() if h > 200 else float('nan'):.3e}")
    plan.close()
