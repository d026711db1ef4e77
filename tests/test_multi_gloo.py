"""N>1 host logic on CPU: world_size-2 gloo run of the tile partition + gather + unpack used by the multi-GPU frame."""
import os
import socket
import sys

import numpy as np
import pytest
from conftest import ROOT


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, W, H, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch
    import torch.distributed as dist
    import rtp_b200 as rt
    from rtp_b200_pkg import multi_gpu as mg
    dist.init_process_group("gloo", rank=rank, world_size=world)
    yy, xx = np.mgrid[0:H, 0:W]
    full = np.stack([xx + 1000 * yy, xx * 0.5, yy * 0.25], -1).astype(np.float32)   # the frame every rank would agree on
    tx, ty = mg.tile_grid(W, H)
    mine = np.zeros_like(full)                                                       # this rank renders only its tiles
    for t in range(tx * ty):
        if rt.tile_owner(t, world)[0] != rank:      # the library's partition (rt_tile_owner, host arithmetic)
            continue
        x0, y0 = (t % tx) * 16, (t // tx) * 16
        mine[y0:y0 + 16, x0:x0 + 16] = full[y0:y0 + 16, x0:x0 + 16]
    packed = torch.from_numpy(mg.pack_tiles_np(mine, rank, world))
    assert packed.numel() == rt.tiles_packed_floats(W, H, world) == mg.packed_floats(W, H, world)
    allp = mg.gather_packed(packed, rank, world, dist)
    ok = True
    if rank == 0:
        out = mg.unpack_tiles_np(allp.numpy(), W, H, world)
        ok = out.tobytes() == full.tobytes()
    # max-over-ranks timing reduction used by bench.py
    t = torch.tensor([float(rank + 1)])
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ok = ok and t.item() == world
    q.put((rank, ok))
    dist.destroy_process_group()


@pytest.mark.parametrize("W,H", [(64, 48), (200, 100), (33, 17)])
def test_tile_gather_world2(W, H):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, W, H, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(60)
    assert sorted(res) == [(0, True), (1, True)]


def test_pack_unpack_numpy_roundtrip():
    sys.path.insert(0, ROOT)
    import rtp_b200  # noqa: F401
    from rtp_b200_pkg import multi_gpu as mg
    rng = np.random.default_rng(0)
    for W, H, world in [(1920, 1080, 8), (50, 50, 3), (16, 16, 4)]:
        fb = rng.random((H, W, 3)).astype(np.float32)
        allp = np.concatenate([mg.pack_tiles_np(fb, r, world) for r in range(world)])
        assert mg.unpack_tiles_np(allp, W, H, world).tobytes() == fb.tobytes()


def test_tile_partition_is_a_balanced_bijection():
    """rt_tile_owner (C) == multi_gpu.tile_owner (numpy mirror); every tile has one owner, every rank gets one tile of every
    group of `world` consecutive ids, and a rank does not own the same tile columns in every row (the failure of t % world
    whenever the tiles per row are a multiple of world: 120 at 1080p, 480 at 8K)."""
    sys.path.insert(0, ROOT)
    import rtp_b200 as rt
    from rtp_b200_pkg import multi_gpu as mg
    for W, H, world in [(1920, 1080, 8), (1920, 1080, 2), (7680, 4320, 8), (200, 100, 3), (16, 16, 4), (64, 48, 1)]:
        tx, ty = mg.tile_grid(W, H)
        owners = np.array([rt.tile_owner(t, world) for t in range(tx * ty)])
        assert [tuple(o) for o in owners] == [mg.tile_owner(t, world) for t in range(tx * ty)]
        assert all(mg.tile_of_rank(k, r, world) == t for t, (r, k) in enumerate(owners))
        counts = np.bincount(owners[:, 0], minlength=world)
        assert counts.max() - counts.min() <= 1
        for g in range(tx * ty // world):
            assert sorted(owners[g * world:(g + 1) * world, 0]) == list(range(world))
        if world > 1 and tx % world == 0 and ty >= 8:
            cols_row0 = set(np.nonzero(owners[:tx, 0] == 0)[0])
            assert any(set(np.nonzero(owners[r * tx:(r + 1) * tx, 0] == 0)[0]) != cols_row0 for r in range(1, ty))
