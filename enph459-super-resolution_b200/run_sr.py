"""run_sr.py-compatible command line on top of the CUDA engine.

    python -m sr459.run_sr --experiment mono_cal_target|rgb_cal_target|mono_barcodes|rgb_barcodes \
           [--psf gaussian|measured] [--psf-dir DIR] [--data-dir DIR] [--output-dir DIR] [--iterations K]

The reference ships four near-identical scripts (mono_cal_target/run_sr.py:320-366 and the same
`main()` in the other three); their flags, directory walks, per-session outputs
(`native_2x.png`, `SAA.png`, `SAA_IBP.png`, `LR_mean.png` | `LR_red_mean.png`, `shifts.json`,
`done.flag`) and skip-if-done behaviour are reproduced here; `--experiment` selects which script's
loader and constants apply.  Frames go to the GPU as uint8 and every numeric step (Bayer-red
extraction, repetition mean, Native-2x, SAA, IBP, the truncating uint8 cast) runs in libsr459.
The matplotlib plots (`comparison.png`, `convergence.png`) are not produced; the convergence
curve is written as `errors.json` instead.
"""
from __future__ import annotations

import argparse
import json
import os
import re
import time
from typing import List, Sequence, Tuple

import numpy as np

from . import ops

# per-experiment constants (reference file:line)
EXPERIMENTS = {
    # name: (IBP_ITERATIONS, red Bayer channel?, LR-mean file name)
    "mono_cal_target": (80, False, "LR_mean.png"),       # mono_cal_target/run_sr.py:56, :309
    "rgb_cal_target": (50, True, "LR_red_mean.png"),     # rgb_cal_target/run_sr.py:59, :323
    "mono_barcodes": (80, False, "LR_mean.png"),         # mono_barcodes/run_sr.py:66
    "rgb_barcodes": (80, True, "LR_red_mean.png"),       # rgb_barcodes/run_sr.py:73, :358
}
IMAGE_SHIFTS = [("center.png", (0.0, 0.0)), ("shift_0.png", (+0.5, -0.5)), ("shift_1.png", (+0.5, +0.5)),
                ("shift_2.png", (-0.5, -0.5)), ("shift_3.png", (-0.5, +0.5))]          # mono_cal_target/run_sr.py:60-66
CORNER_SHIFTS = [(+0.5, -0.5), (+0.5, +0.5), (-0.5, -0.5), (-0.5, +0.5)]               # mono_barcodes/run_sr.py:71-76
CORNER_ORDER = ["(-x,+y)", "(+x,+y)", "(-x,-y)", "(+x,-y)"]                            # rgb_cal_target/run_sr.py:61


def _load_png(path: str):
    """PNG -> device tensor: uint8 for 8-bit grey files (converted / gathered by the ingest kernels),
    float32 channel mean for RGB files exactly as load_gray does on the host (run_sr.py:73-75)."""
    from PIL import Image
    torch = ops._torch()
    img = np.array(Image.open(path))
    if img.ndim == 3 or img.dtype != np.uint8:
        a = np.array(img, dtype=np.float64)
        a = a.mean(axis=2) if a.ndim == 3 else a
        return torch.from_numpy(a.astype(np.float32)).cuda()
    return torch.from_numpy(np.ascontiguousarray(img)).cuda()


def _ingest(groups, red: bool):
    """groups[k] = the repetitions of frame k as device tensors -> (frames, mean_lr, mean_u8 | None)."""
    torch = ops._torch()
    if all(t.dtype == torch.uint8 for g in groups for t in g):
        return ops.ingest_groups(groups, red=red)                      # one fused kernel, float64 means
    frames = []                                                        # RGB / 16-bit files: stage by stage
    for g in groups:
        fs = [t if t.dtype == torch.float32 else ops.u8_to_f32(t) for t in g]
        fs = [ops.extract_red(f) for f in fs] if red else fs
        frames.append(ops.mean_frames(fs) if len(fs) > 1 else fs[0])
    return frames, ops.mean_frames(frames), None


def load_session(session_dir: str):
    """mono_cal_target loader (run_sr.py:78-99): centre + four shifted frames, nominal shifts."""
    groups, shifts = [], []
    for fname, s in IMAGE_SHIFTS:
        path = os.path.join(session_dir, fname)
        if not os.path.exists(path):
            print(f"  WARNING: {fname} not found, skipping")
            continue
        groups.append([_load_png(path)])
        shifts.append(s)
    if len(groups) < 2:
        raise FileNotFoundError(f"Need at least 2 images in {session_dir}")
    return groups, shifts


def load_combo(combo_dir: str):
    """rgb_cal_target loader (run_sr.py:78-113): red Bayer sites, mean over the repetitions of each
    corner, calibrated shifts = metadata pixels / 2."""
    with open(os.path.join(combo_dir, "metadata.json")) as fp:
        meta = json.load(fp)

    def get_shift(label):
        if "expected_shifts" in meta:
            s = meta["expected_shifts"][label]
            return s["dy_px"] / 2.0, s["dx_px"] / 2.0
        if "corners" in meta:
            c = meta["corners"][label]
            return c["expected_dy_px"] / 2.0, c["expected_dx_px"] / 2.0
        raise KeyError(f"Cannot find shift for {label} in metadata")

    groups, shifts = [], []
    for idx, label in enumerate(CORNER_ORDER):
        reps = sorted(f for f in os.listdir(combo_dir) if f.startswith(f"corner{idx}_rep") and f.endswith(".png"))
        if not reps:
            raise FileNotFoundError(f"No images for corner{idx} in {combo_dir}")
        groups.append([_load_png(os.path.join(combo_dir, r)) for r in reps])
        shifts.append(get_shift(label))
    return groups, shifts


def load_reps(session_dir: str):
    """*_barcodes loader (mono_barcodes/run_sr.py:89-130, rgb_barcodes/run_sr.py:102-145): every
    repetition index is an independent four-frame job."""
    rep_indices = sorted({int(m.group(1)) for f in os.listdir(session_dir)
                          for m in [re.match(r"corner\d+_rep(\d+)\.png", f)] if m})
    if not rep_indices:
        raise FileNotFoundError(f"No corner*_rep*.png files in {session_dir}")
    all_reps = []
    for ri in rep_indices:
        groups = []
        for ci in range(4):
            path = os.path.join(session_dir, f"corner{ci}_rep{ri:02d}.png")
            if not os.path.exists(path):
                raise FileNotFoundError(f"Missing {path}")
            groups.append([_load_png(path)])
        all_reps.append(groups)
    return all_reps, list(CORNER_SHIFTS)


def _save_u8(img, path: str):
    from PIL import Image
    Image.fromarray(ops.to_uint8(img).cpu().numpy()).save(path)        # clip + truncating cast on the device (:303)


def run_job(groups: Sequence, red: bool, shifts: Sequence[Tuple[float, float]], psf, n_iter: int, out_dir: str,
            lr_name: str, extra_json=None) -> dict:
    """process_session's numeric body and outputs (run_sr.py:272-314)."""
    torch = ops._torch()
    os.makedirs(out_dir, exist_ok=True)
    t0 = time.time()
    frames, mean_lr, mean_u8 = _ingest(groups, red)
    n, (h, w) = len(frames), frames[0].shape
    plan = ops.get_plan(n, h, w, shifts, psf)
    native = plan.native_2x(mean_lr)
    saa = plan.saa(list(frames))
    hr = saa.clone()
    errors = plan.ibp_(list(frames), hr, n_iter)
    torch.cuda.synchronize()
    dt = time.time() - t0
    for name, img in (("native_2x", native), ("SAA", saa), ("SAA_IBP", hr)):
        _save_u8(img, os.path.join(out_dir, f"{name}.png"))
    if mean_u8 is not None:
        from PIL import Image
        Image.fromarray(mean_u8.cpu().numpy()).save(os.path.join(out_dir, lr_name))
    else:
        _save_u8(mean_lr, os.path.join(out_dir, lr_name))
    if extra_json:
        for fname, obj in extra_json.items():
            with open(os.path.join(out_dir, fname), "w") as fp:
                json.dump(obj, fp, indent=2)
    with open(os.path.join(out_dir, "errors.json"), "w") as fp:
        json.dump({"mse_per_iteration": errors}, fp)
    open(os.path.join(out_dir, "done.flag"), "w").close()
    print(f"  {n} frames {h}x{w} -> {2*h}x{2*w}: Native-2x, SAA, SAA+IBP({n_iter}) in {dt*1e3:.1f} ms; output {out_dir}")
    return {"errors": errors, "seconds": dt}


def process_session(experiment: str, session_dir: str, psf, output_base: str, n_iter: int) -> int:
    """One session / combo directory; returns the number of SR jobs run (skips finished ones)."""
    _, red, lr_name = EXPERIMENTS[experiment]
    name = os.path.basename(session_dir)
    jobs = 0
    if experiment in ("mono_cal_target", "rgb_cal_target"):
        out_dir = os.path.join(output_base, name)
        if os.path.exists(os.path.join(out_dir, "done.flag")):
            print(f"  [skip] {name} - already done")
            return 0
        if experiment == "mono_cal_target":
            groups, shifts = load_session(session_dir)
            extra = None
        else:
            groups, shifts = load_combo(session_dir)
            extra = {"shifts.json": {"shifts_lr_yx": [list(s) for s in shifts], "corner_labels": CORNER_ORDER}}
        run_job(groups, red, shifts, psf, n_iter, out_dir, lr_name, extra)
        return 1
    all_reps, shifts = load_reps(session_dir)
    for rep_idx, groups in enumerate(all_reps):
        out_dir = os.path.join(output_base, name, f"rep{rep_idx}")
        if os.path.exists(os.path.join(out_dir, "done.flag")):
            print(f"  [skip] rep {rep_idx} - already done")
            continue
        run_job(groups, red, shifts, psf, n_iter, out_dir, lr_name)
        jobs += 1
    return jobs


def find_sessions(experiment: str, data_dir: str) -> List[str]:
    dirs = [os.path.join(data_dir, d) for d in sorted(os.listdir(data_dir)) if os.path.isdir(os.path.join(data_dir, d))]
    if experiment == "mono_cal_target":                                   # run_sr.py:338-343
        dirs = [d for d in dirs if os.path.exists(os.path.join(d, "center.png"))]
        what = "session folders with center.png"
    elif experiment == "rgb_cal_target":                                  # rgb_cal_target/run_sr.py:352-356
        dirs = [d for d in dirs if os.path.exists(os.path.join(d, "metadata.json"))]
        what = "directories with metadata.json"
    else:
        what = "session folders"
    if not dirs:
        raise FileNotFoundError(f"No {what} found in {data_dir}")
    return dirs


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description="Multi-frame super-resolution (Native-2x, SAA, SAA+IBP) on B200")
    ap.add_argument("--experiment", choices=sorted(EXPERIMENTS), default="mono_cal_target")
    ap.add_argument("--psf", choices=["gaussian", "measured"], default="gaussian",
                    help="PSF type: gaussian (default) or measured from calibration data")
    ap.add_argument("--psf-dir", default=None, help="Path to PSF calibration images")
    ap.add_argument("--data-dir", default=None, help="Path to data directory (default: <experiment>/data)")
    ap.add_argument("--output-dir", default=None, help="Output directory (default: <experiment>/results)")
    ap.add_argument("--iterations", type=int, default=None, help="IBP iterations (default: the experiment's own constant)")
    args = ap.parse_args(argv)
    data_dir = args.data_dir or os.path.join(args.experiment, "data")
    output_dir = args.output_dir or os.path.join(args.experiment, "results")
    psf_dir = args.psf_dir or os.path.join(args.experiment, "..", "calibration_beam_shift", "data")
    n_iter = args.iterations if args.iterations is not None else EXPERIMENTS[args.experiment][0]

    sessions = find_sessions(args.experiment, data_dir)
    print(f"Found {len(sessions)} session(s):\n" + "\n".join(f"  {os.path.basename(s)}" for s in sessions))
    psf = ops.load_measured_psf(psf_dir) if args.psf == "measured" else ops.make_gaussian_psf()
    t_total, jobs = time.time(), 0
    for i, s in enumerate(sessions, 1):
        print(f"\n[{i}/{len(sessions)}] {os.path.basename(s)}")
        jobs += process_session(args.experiment, s, psf, output_dir, n_iter)
    print(f"\nAll sessions done ({jobs} job(s)) in {time.time() - t_total:.1f} s")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
