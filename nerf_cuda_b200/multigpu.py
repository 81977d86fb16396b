"""One-process-per-GPU frame sharding and the framebuffer gather (torch.distributed: NCCL over NVLink on GPUs, gloo in
the CPU tests).

The reference's multi-GPU is a compile-time `#define NGPU 2` (/root/reference/include/nerf-cuda/common.h:91): pixel
p = NGPU*tid + gpu goes to device `gpu` (render_utils.h:37), every device holds a full model replica
(nerf_render.cu:173-183,458-462), and the frame is re-interleaved on the host after two D2H copies per device
(nerf_render.cu:345-358).  Here the shard count is the world size, shards are round-robin 32x32-pixel image TILES
(balance + locality), rays never exchange data while rendering, and the only collective is one gather
of packed u8 tiles to rank 0 followed by a tile transpose into pixel order.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist


TILE = 32  # NRF_TILE_PX (csrc/nrf_device.cuh)


class FrameSharding:
    """Tile <-> shard map shared with the kernels (ShardMap in csrc/nrf_device.cuh): the image is cut into 32x32-pixel
    tiles numbered row-major; tile t belongs to shard t % world and is that shard's local tile t // world.  A shard's
    packed output is [local tiles, 32, 32, C], pixels row-major inside the tile; slots of edge tiles that fall outside
    the image are present (and ignored by `assemble`)."""

    def __init__(self, width: int, height: int, world: int):
        if world < 1:
            raise ValueError("world must be >= 1")
        self.W, self.H, self.world = width, height, world
        self.tiles_x, self.tiles_y = -(-width // TILE), -(-height // TILE)
        self.n_tiles = self.tiles_x * self.tiles_y
        self.tiles_per_shard = -(-self.n_tiles // world)                  # max over shards (gather stride)

    def local_tiles(self, shard: int) -> int:
        return self.n_tiles // self.world + (1 if shard < self.n_tiles % self.world else 0)

    def local_count(self, shard: int) -> int:
        return self.local_tiles(shard) * TILE * TILE

    def packed_pixels(self, shard: int) -> np.ndarray:
        """Global pixel index of every slot of the shard's packed output (-1 outside the image), [tiles, 32, 32]."""
        t = np.arange(self.local_tiles(shard), dtype=np.int64) * self.world + shard
        ty, tx = t // self.tiles_x, t % self.tiles_x
        y = ty[:, None, None] * TILE + np.arange(TILE)[None, :, None]
        x = tx[:, None, None] * TILE + np.arange(TILE)[None, None, :]
        return np.where((x < self.W) & (y < self.H), y * self.W + x, -1)

    def assemble(self, gathered: torch.Tensor) -> torch.Tensor:
        """gathered: [world, tiles_per_shard, 32, 32, C] packed tiles of every shard -> [H, W, C] frame.
        Pure view permutations + one copy: tile t lives in shard t % world at local index t // world."""
        C = gathered.shape[-1]
        g = gathered.permute(1, 0, 2, 3, 4).reshape(self.tiles_per_shard * self.world, TILE, TILE, C)[: self.n_tiles]
        g = g.reshape(self.tiles_y, self.tiles_x, TILE, TILE, C).permute(0, 2, 1, 3, 4)
        return g.reshape(self.tiles_y * TILE, self.tiles_x * TILE, C)[: self.H, : self.W]


class FrameGather:
    """Gathers every rank's packed u8 tiles to `dst` and returns the assembled frame there (None elsewhere)."""

    def __init__(self, sharding: FrameSharding, rank: int, device, channels=3, dst=0, group=None):
        self.s, self.rank, self.dst, self.group, self.C = sharding, rank, dst, group, channels
        self.n_tiles = sharding.local_tiles(rank)
        self.send = torch.zeros(sharding.tiles_per_shard, TILE, TILE, channels, dtype=torch.uint8, device=device)
        self.recv = (torch.empty(sharding.world, sharding.tiles_per_shard, TILE, TILE, channels, dtype=torch.uint8,
                                 device=device) if rank == dst else None)

    def __call__(self, local: torch.Tensor) -> torch.Tensor | None:
        """local: [local tiles, 32, 32, C] u8 (this rank's packed tiles)."""
        local = local.reshape(-1, TILE, TILE, self.C)
        if local.shape[0] != self.n_tiles:
            raise ValueError(f"rank {self.rank}: expected {self.n_tiles} local tiles, got {local.shape[0]}")
        self.send[: self.n_tiles].copy_(local, non_blocking=True)
        if self.s.world == 1:
            return self.s.assemble(self.send.unsqueeze(0))
        out = list(self.recv.unbind(0)) if self.rank == self.dst else None
        dist.gather(self.send, out, dst=self.dst, group=self.group)
        return self.s.assemble(self.recv) if self.rank == self.dst else None
