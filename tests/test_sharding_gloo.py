"""CPU test of the N>1 path (world_size 2 and 3, gloo): frame sharding + framebuffer gather + assembly.  Each rank
"renders" its shard with the CPU oracle standing in for the device (the oracle is the checker here, allowed in tests/),
packs u8 pixels in local ray order exactly like the CUDA path does, and rank 0 must end up with the same frame as a
single-shard render."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from nerf_cuda_b200 import synthetic as syn  # noqa: E402
from nerf_cuda_b200.multigpu import FrameGather, FrameSharding  # noqa: E402


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _render_packed(O, intr, pose, W, H, packed):
    """Oracle render of a shard's packed tile slots ([tiles,32,32] global pixel or -1), row by row inside each tile."""
    out = np.zeros(packed.shape + (3,), np.float32)
    for t in range(packed.shape[0]):
        for y in range(32):
            row = packed[t, y]
            n = int((row >= 0).sum())     # valid pixels of a tile row are a prefix (x grows along the row)
            if n:
                rgb, _, _ = O.render(intr, pose, W, H, int(row[0]), n)
                out[t, y, :n] = rgb
    return out


def _worker(rank, world, port, W, H, q):
    from oracle import orc
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sc = syn.make_scene("tiny")
        O = orc.Oracle(sc.desc, sc.params, sc.grid, sc.aabb)
        intr, pose = syn.intrinsics(W, H), syn.look_at_pose(4.0, 20.0, 40.0)
        sh = FrameSharding(W, H, world)
        local = O.to_u8(_render_packed(O, intr, pose, W, H, sh.packed_pixels(rank)))
        frame = FrameGather(sh, rank, "cpu")(torch.from_numpy(local))
        if rank == 0:
            full, _, _ = O.render(intr, pose, W, H)
            q.put(bool(np.array_equal(frame.numpy(), O.to_u8(full).reshape(H, W, 3))))
        else:
            assert frame is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,W,H", [(2, 72, 40), (2, 33, 17), (3, 100, 64)])
def test_sharded_frame_equals_single_shard(world, W, H):
    from oracle import orc
    orc.lib()   # build once before forking
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, W, H, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True


def test_assemble_inverts_the_packing():
    for W, H, world in [(100, 70, 3), (1920, 1080, 8), (5, 8, 4), (64, 32, 2)]:
        sh = FrameSharding(W, H, world)
        g = torch.full((world, sh.tiles_per_shard, 32, 32, 1), -7, dtype=torch.int64)
        for r in range(world):
            pk = sh.packed_pixels(r)
            g[r, : pk.shape[0], :, :, 0] = torch.from_numpy(pk)
        assert torch.equal(sh.assemble(g).reshape(-1), torch.arange(W * H))
