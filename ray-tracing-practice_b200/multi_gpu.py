"""Multi-GPU frame: interleaved 16x16 tiles, one process per GPU, gather of the packed tiles to rank 0.

The reference is single-GPU; this is the B200-native extension BASELINE.json asks for.  Pixels
are independent and the scene is read-only, so every rank holds the whole scene and renders its share of the tiles
(rt_render_tiles; tiles are dealt in groups of `world` consecutive ids, the position inside a group rotated by a hash
of the group -- include/rt_b200.h, rt_tile_owner).  The only exchange is the end-of-frame gather of W*H*12/world
bytes per rank to rank 0 (grouped ncclSend / ncclRecv inside rt_render_tiles), followed by an unpack kernel.  The numpy
pack / unpack below state the same tile layout for the CPU (gloo) tests; TiledFrame is the same frame assembled with
torch.distributed instead of the library's own communicator.
"""
import numpy as np

TILE = 16


def tile_grid(W, H):
    return (W + TILE - 1) // TILE, (H + TILE - 1) // TILE


def packed_floats(W, H, world):
    tx, ty = tile_grid(W, H)
    per = (tx * ty + world - 1) // world
    return per * TILE * TILE * 3


def tile_rot(g, world):
    """position rotation of tile group g (kernels.h: tile_rot)"""
    return (((g * 0x9E3779B9) & 0xFFFFFFFF) * world) >> 32 if world > 1 else 0


def tile_of_rank(k, rank, world):
    return k * world + (rank - tile_rot(k, world)) % world


def tile_owner(t, world):
    k = t // world
    return (t % world + tile_rot(k, world)) % world, k


def pack_tiles_np(fb, rank, world):
    """numpy statement of rt_tiles_pack: rank's tiles in increasing k, 16x16x3 floats each, zero padded."""
    H, W = fb.shape[:2]
    tx, ty = tile_grid(W, H)
    out = np.zeros(packed_floats(W, H, world), np.float32).reshape(-1, TILE, TILE, 3)
    for k in range((tx * ty + world - 1) // world):
        t = tile_of_rank(k, rank, world)
        if t >= tx * ty:
            continue
        x0, y0 = (t % tx) * TILE, (t // tx) * TILE
        blk = fb[y0:y0 + TILE, x0:x0 + TILE]
        out[k, :blk.shape[0], :blk.shape[1]] = blk
    return out.reshape(-1)


def unpack_tiles_np(packed_all, W, H, world):
    """numpy statement of rt_tiles_unpack: packed_all = concatenation of every rank's packed buffer."""
    tx, ty = tile_grid(W, H)
    per = packed_floats(W, H, world)
    fb = np.zeros((H, W, 3), np.float32)
    for t in range(tx * ty):
        r, k = tile_owner(t, world)
        blk = packed_all[r * per:(r + 1) * per].reshape(-1, TILE, TILE, 3)[k]
        x0, y0 = (t % tx) * TILE, (t // tx) * TILE
        h, w = min(TILE, H - y0), min(TILE, W - x0)
        fb[y0:y0 + h, x0:x0 + w] = blk[:h, :w]
    return fb


def gather_packed(packed, rank, world, dist, dst=0):
    """gather every rank's packed tile buffer on `dst` (torch tensors; NCCL on GPUs, gloo in the CPU tests)."""
    import torch
    if world == 1:
        return packed
    bufs = [torch.empty_like(packed) for _ in range(world)] if rank == dst else None
    dist.gather(packed, bufs, dst=dst)
    return torch.cat(bufs) if rank == dst else None


class TiledFrame:
    """Device buffers + calls for one rank's share of a frame (GPU path)."""

    def __init__(self, scene, W, H, rank, world, device):
        import torch
        self.scene, self.W, self.H, self.rank, self.world = scene, W, H, rank, world
        self.fb = torch.zeros(H * W * 3, dtype=torch.float32, device=device)
        n = packed_floats(W, H, world)
        self.packed = torch.zeros(n, dtype=torch.float32, device=device)
        self.full = torch.zeros(H * W * 3, dtype=torch.float32, device=device) if rank == 0 else None

    def render(self, cam, stream, **opts):
        self.scene.render_device(cam, self.fb.data_ptr(), stream, tile_rank=self.rank, tile_nranks=self.world, **opts)

    def gather(self, dist, stream):
        """pack own tiles, gather to rank 0, unpack there. Returns the full frame tensor on rank 0."""
        if self.world == 1:
            return self.fb
        self.scene.tiles_pack(self.fb.data_ptr(), self.W, self.H, self.rank, self.world, self.packed.data_ptr(), stream)
        allp = gather_packed(self.packed, self.rank, self.world, dist)
        if self.rank == 0:
            self.scene.tiles_unpack(allp.data_ptr(), self.W, self.H, self.world, self.full.data_ptr(), stream)
            return self.full
        return None
