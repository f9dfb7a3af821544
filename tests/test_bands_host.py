"""Host logic of the row-band multi-GPU path on CPU: band layout arithmetic, and the per-iteration
halo exchange + error all-reduce at world_size 2 and 3 over gloo, with the oracle standing in for the
CUDA iteration kernel (the oracle is the checker here, never the product)."""
import os
import socket

import numpy as np
import pytest

import sr459
from sr459.bands import BandLayout, band_loop, exchange_halos, reduce_errors
from oracle import sr_oracle as O

NOMINAL = [(+0.5, -0.5), (+0.5, +0.5), (-0.5, -0.5), (-0.5, +0.5)]


def test_band_layout_tiles_the_image_and_keeps_halos():
    for h, world, reach in [(1536, 8, 15), (1536, 4, 3), (100, 3, 3), (64, 2, 15), (17, 1, 3)]:
        lays = [BandLayout(h, 40, reach, r, world) for r in range(world)]
        assert lays[0].own_y0 == 0 and lays[-1].own_y1 == 2 * h
        for a, b in zip(lays, lays[1:]):
            assert a.own_y1 == b.own_y0                       # bands tile the rows, even boundaries
            assert a.own_y1 % 2 == 0
        for l in lays:
            assert l.halo == 2 * reach + 2
            assert l.hr_lo == max(0, l.own_y0 - l.halo) and l.hr_hi == min(2 * h, l.own_y1 + l.halo)
            # every LR row a site within reach+1 of the owned rows can touch is present
            assert l.lr_lo <= max(0, (l.own_y0 - reach - 2) // 2) and l.lr_hi >= min(h, (l.own_y1 + reach + 2) // 2 + 1)
    with pytest.raises(ValueError):
        BandLayout(64, 40, 15, 0, 8)                          # 16-row bands are thinner than the 32-row halo
    with pytest.raises(ValueError):
        BandLayout(4, 40, 3, 0, 8)


def _oracle_band_iterate(lay, lr_full, ints, psf, cur_local, step=0.5):
    """One IBP iteration for the rows a band owns, from its local buffer only (zeros elsewhere)."""
    H, W, n = lay.H, lay.W, len(lr_full)
    full = np.zeros((H, W))
    full[lay.hr_lo:lay.hr_hi] = cur_local
    B = O.conv2_same_zero(full, psf)
    acc = np.zeros((H, W))
    sq = 0.0
    for lr, (sy, sx) in zip(lr_full, ints):
        err = lr - O.clamp_gather(B, sy, sx)[::2, ::2]
        sq += float((err[lay.own_i0:lay.own_i1] ** 2).sum())
        up = np.zeros((H, W))
        up[::2, ::2] = err
        acc += O.clamp_gather(up, -sy, -sx)
    new = np.clip(full + (step / n) * O.conv2_same_zero(acc, psf[::-1, ::-1]), 0.0, 255.0)
    return new[lay.own_y0:lay.own_y1], sq


def _worker(rank, world, port, h, w, psf_size, n_iter, out_dir):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(77)
        lrs = [rng.uniform(0, 255, (h, w)) for _ in NOMINAL]
        ints = [(int(2 * dy), int(2 * dx)) for dy, dx in NOMINAL]
        psf = O.make_gaussian_psf(psf_size, 1.0 if psf_size == 7 else 2.5)
        hr0 = O.shift_and_add_integer(lrs, NOMINAL)
        lay = BandLayout(h, w, psf_size // 2, rank, world)
        cur = torch.from_numpy(np.ascontiguousarray(lay.split_hr(hr0)))
        scratch = cur.clone()
        sums = []

        def iterate(c, nx, it):
            rows, sq = _oracle_band_iterate(lay, lrs, ints, psf, c.numpy())
            nx[lay.owned] = torch.from_numpy(rows)
            sums.append(sq)

        res = band_loop(lay, cur, scratch, n_iter, iterate, exchange_halos)
        errors = reduce_errors(sums, len(lrs), lay)
        np.savez(os.path.join(out_dir, f"band{rank}.npz"), rows=res[lay.owned].numpy(), errors=np.array(errors),
                 y0=lay.own_y0, y1=lay.own_y1)
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world,h,w,psf_size", [(2, 40, 24, 7), (3, 48, 16, 7), (2, 36, 16, 15)])
def test_band_loop_over_gloo_matches_full_image_oracle(tmp_path, world, h, w, psf_size):
    import torch.multiprocessing as mp
    n_iter = 3
    mp.spawn(_worker, args=(world, _free_port(), h, w, psf_size, n_iter, str(tmp_path)), nprocs=world, join=True)
    rng = np.random.default_rng(77)
    lrs = [rng.uniform(0, 255, (h, w)) for _ in NOMINAL]
    psf = O.make_gaussian_psf(psf_size, 1.0 if psf_size == 7 else 2.5)
    hr0 = O.shift_and_add_integer(lrs, NOMINAL)
    hr_ref, e_ref = O.ibp_integer_shifts(lrs, NOMINAL, psf, hr0, n_iter=n_iter)
    got = np.zeros_like(hr_ref)
    for r in range(world):
        z = np.load(tmp_path / f"band{r}.npz")
        got[int(z["y0"]):int(z["y1"])] = z["rows"]
        np.testing.assert_allclose(z["errors"], e_ref, rtol=1e-12)      # every rank holds the all-reduced errors
    assert np.max(np.abs(got - hr_ref)) < 1e-9


def test_single_rank_exchange_is_a_no_op():
    import torch
    lay = BandLayout(16, 8, 3, 0, 1)
    t = torch.arange(32 * 16, dtype=torch.float64).reshape(32, 16)
    u = t.clone()
    exchange_halos(u, lay)
    assert torch.equal(t, u)
    assert reduce_errors([float(16 * 8 * 4 * 2)], 4, lay) == [2.0]


def _handle_worker(rank, world, port, out_dir):
    import torch.distributed as dist
    from sr459.bands import exchange_handles
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mine = bytes([rank]) * 64                             # stands in for a cudaIpcMemHandle_t
        up, down = exchange_handles(mine, rank, world)
        with open(os.path.join(out_dir, f"h{rank}.txt"), "w") as f:
            f.write(f"{up[0] if up else -1} {down[0] if down else -1} {len(up or b'x' * 64)} {len(down or b'x' * 64)}")
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_arena_handles_reach_the_right_neighbours_over_gloo(tmp_path, world):
    """The only collective of the peer-memory band path besides the final error all-reduce: every rank must end up with the
    handle of rank-1 (its upper neighbour) and rank+1 (its lower one), None at the image ends."""
    import torch.multiprocessing as mp
    mp.spawn(_handle_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        up, down, lu, ld = map(int, (tmp_path / f"h{r}.txt").read_text().split())
        assert up == (r - 1 if r > 0 else -1) and down == (r + 1 if r < world - 1 else -1)
        assert lu == 64 and ld == 64


def test_band_layout_extended_lr_rows_and_fractional_halo():
    from sr459.bands import LR_MARGIN, band_halo
    assert band_halo(3, True) == 8 and band_halo(15, True) == 32
    assert band_halo(3, False) == 48 and band_halo(15, False) == 80     # 2 * (reach + 19), rounded up to 16
    for h, world, reach, halo in [(1536, 8, 15, None), (768, 4, 3, 48), (100, 3, 3, None)]:
        for r in range(world):
            l = BandLayout(h, 40, reach, r, world, halo=halo)
            assert l.lr_ext_lo == max(0, l.hr_lo // 2 - LR_MARGIN) and l.lr_ext_hi == min(h, l.hr_hi // 2 + LR_MARGIN)
            assert l.lr_ext_lo <= l.lr_lo and l.lr_hi <= l.lr_ext_hi
            # zoom's column prefilter: 16-sample tail + 3 taps beyond the rows the band's HR rows map to
            assert l.lr_ext_lo == 0 or l.hr_lo // 2 - l.lr_ext_lo >= 19
            assert l.lr_ext_hi == h or l.lr_ext_hi - l.hr_hi // 2 >= 19
