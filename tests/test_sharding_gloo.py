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


def _render_pixels(O, intr, pose, W, H, pixels, block):
    """Oracle render of an arbitrary block-structured pixel list (contiguous runs of <= block pixels)."""
    out = np.zeros((pixels.size, 3), np.float32)
    i = 0
    while i < pixels.size:
        j = i
        while j + 1 < pixels.size and pixels[j + 1] == pixels[j] + 1:
            j += 1
        rgb, _, _ = O.render(intr, pose, W, H, int(pixels[i]), j - i + 1)
        out[i:j + 1] = rgb
        i = j + 1
    return out


def _worker(rank, world, port, W, H, block, q):
    from oracle import orc
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sc = syn.make_scene("tiny")
        O = orc.Oracle(sc.desc, sc.params, sc.grid, sc.aabb)
        intr, pose = syn.intrinsics(W, H), syn.look_at_pose(4.0, 20.0, 40.0)
        sh = FrameSharding(W, H, world, block)
        pix = sh.local_pixels(rank)
        local = O.to_u8(_render_pixels(O, intr, pose, W, H, pix, block))
        frame = FrameGather(sh, rank, "cpu")(torch.from_numpy(local))
        if rank == 0:
            full, _, _ = O.render(intr, pose, W, H)
            q.put(bool(np.array_equal(frame.numpy(), O.to_u8(full).reshape(H, W, 3))))
        else:
            assert frame is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,W,H,block", [(2, 40, 24, 128), (2, 33, 17, 1), (3, 40, 24, 64)])
def test_sharded_frame_equals_single_shard(world, W, H, block):
    from oracle import orc
    orc.lib()   # build once before forking
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, W, H, block, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True


def test_sharding_counts_match_c_abi_rule():
    for n_pix, world, block in [(1000, 3, 128), (1920 * 1080, 8, 128), (5, 8, 4), (64, 2, 1)]:
        sh = FrameSharding(n_pix, 1, world, block)
        tot = 0
        for r in range(world):
            px = sh.local_pixels(r)
            assert px.size == sh.local_count(r) <= sh.n_max
            tot += px.size
        assert tot == n_pix
        # assemble() inverts the packing
        g = torch.zeros(world, sh.n_max, 1, dtype=torch.int64)
        for r in range(world):
            px = sh.local_pixels(r)
            g[r, : px.size, 0] = torch.from_numpy(px)
        assert torch.equal(sh.assemble(g).reshape(-1), torch.arange(n_pix))
