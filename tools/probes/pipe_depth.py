"""Dev probe: FrameSetPipeline throughput vs depth (uint8 in / uint8 out, CUDA-graph jobs, config 1)."""
import sys, time, numpy as np
sys.path.insert(0, '/root/repo')
import torch, sr459
from sr459.driver import FrameSetPipeline
NOM = [(+0.5, -0.5), (+0.5, +0.5), (-0.5, -0.5), (-0.5, +0.5)]
rng = np.random.default_rng(1)
sets = [torch.from_numpy(rng.integers(0, 256, (4, 1536, 2048), dtype=np.uint8)).pin_memory() for _ in range(4)]
for depth in (2, 3, 4, 6):
    p = FrameSetPipeline(4, 1536, 2048, NOM, sr459.make_gaussian_psf(), 50, depth=depth, in_dtype="u8", out_dtype="u8", use_graph=True)
    p.run([sets[i % 4] for i in range(6)])
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    p.run([sets[i % 4] for i in range(50)])
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"depth {depth}: {dt / 50 * 1e3:.4f} ms/job  {50 * 12.582912 / dt:.0f} HR Mpix/s", flush=True)
    p.close()
