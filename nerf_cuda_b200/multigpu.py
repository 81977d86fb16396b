"""One-process-per-GPU frame sharding and the framebuffer gather (torch.distributed: NCCL over NVLink on GPUs, gloo in
the CPU tests).

The reference's multi-GPU is a compile-time `#define NGPU 2` (/root/reference/include/nerf-cuda/common.h:91): pixel
p = NGPU*tid + gpu goes to device `gpu` (render_utils.h:37), every device holds a full model replica
(nerf_render.cu:173-183,458-462), and the frame is re-interleaved on the host after two D2H copies per device
(nerf_render.cu:345-358).  Here the shard count is the world size, shards are round-robin BLOCKS of `block` consecutive
pixels (block = 1 reproduces the reference's interleave), rays never exchange data while rendering, and the only
collective is one gather of packed u8 pixels to rank 0 followed by a block transpose into pixel order.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist


class FrameSharding:
    """Pixel <-> (shard, local ray) map shared with the kernels (ShardMap in csrc/nrf_device.cuh)."""

    def __init__(self, width: int, height: int, world: int, block: int):
        if world < 1 or block < 1:
            raise ValueError("world and block must be >= 1")
        self.W, self.H, self.world, self.block = width, height, world, block
        self.n_pixels = width * height
        self.n_blocks = (self.n_pixels + block - 1) // block
        self.blocks_per_shard = (self.n_blocks + world - 1) // world      # max over shards
        self.n_max = self.blocks_per_shard * block                        # padded local ray count (gather stride)

    def local_count(self, shard: int) -> int:
        mine = self.n_blocks // self.world + (1 if shard < self.n_blocks % self.world else 0)
        n = mine * self.block
        last = self.n_blocks - 1
        if last >= 0 and last % self.world == shard:
            n -= self.n_blocks * self.block - self.n_pixels
        return n

    def local_pixels(self, shard: int) -> np.ndarray:
        q = np.arange(self.local_count(shard), dtype=np.int64)
        return ((q // self.block) * self.world + shard) * self.block + q % self.block

    def assemble(self, gathered: torch.Tensor) -> torch.Tensor:
        """gathered: [world, n_max, C] packed local rays of every shard -> [H, W, C] frame in pixel order.
        A pure view permutation + one copy: pixel block b lives in shard b % world at local block b // world."""
        C = gathered.shape[-1]
        g = gathered.view(self.world, self.blocks_per_shard, self.block, C).permute(1, 0, 2, 3)
        flat = g.reshape(self.blocks_per_shard * self.world * self.block, C)
        return flat[: self.n_pixels].reshape(self.H, self.W, C)


class FrameGather:
    """Gathers every rank's packed u8 pixels to `dst` and returns the assembled frame there (None elsewhere)."""

    def __init__(self, sharding: FrameSharding, rank: int, device, channels=3, dst=0, group=None):
        self.s, self.rank, self.dst, self.group, self.C = sharding, rank, dst, group, channels
        self.n_local = sharding.local_count(rank)
        self.send = torch.zeros(sharding.n_max, channels, dtype=torch.uint8, device=device)
        self.recv = (torch.empty(sharding.world, sharding.n_max, channels, dtype=torch.uint8, device=device)
                     if rank == dst else None)

    def __call__(self, local: torch.Tensor) -> torch.Tensor | None:
        """local: [n_local, C] u8 (this rank's packed pixels, local ray order)."""
        if local.shape[0] != self.n_local:
            raise ValueError(f"rank {self.rank}: expected {self.n_local} local rays, got {local.shape[0]}")
        self.send[: self.n_local].copy_(local, non_blocking=True)
        if self.s.world == 1:
            return self.s.assemble(self.send.unsqueeze(0))
        out = list(self.recv.unbind(0)) if self.rank == self.dst else None
        dist.gather(self.send, out, dst=self.dst, group=self.group)
        return self.s.assemble(self.recv) if self.rank == self.dst else None
