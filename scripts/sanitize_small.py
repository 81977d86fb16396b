"""Small end-to-end exercise for compute-sanitizer: frames (whole, sharded, non-multiple-of-4 width), batch, density grid, post-process."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from nerf_cuda_b200 import NerfRender, synthetic as syn
for name in ("tiny", "tiny3"):
    sc = syn.make_scene(name)
    r = NerfRender(1); r.load_raw(sc.desc, sc.params, sc.grid, sc.aabb)
    for W, H in ((64, 48), (50, 37)):
        r.set_shard(0, 1); r.set_resolution((W, H))
        intr = syn.intrinsics(W, H)
        poses = np.stack([syn.look_at_pose(syn.camera_radius(name), 20.0, az) for az in (10.0, 130.0, 250.0)])
        r.render_frame(intr, poses[0]); r.render_frame_device(intr, poses[1]); r.synchronize()
        r.render_batch(intr, poses)
        r.set_shard(1, 3); r.set_resolution((W, H)); r.render_frame(intr, poses[2])
    r.set_shard(0, 1); r.set_resolution((32, 32))
    r.postprocess(np.random.rand(1000, 4).astype(np.float32), np.random.rand(1000, 4).astype(np.float32), 1.0, 1, 0.5,
                  [0.1, 0.2, 0.3, 1.0], 1, 2, True)
    r.close()
# round 2: termination-aware rounds on an opaque scene (all schedules), the peer frame with its fence kernels (one rank),
# the stage probes, the wrapped-dense level mode (log2_hashmap_size 22: finest level, reference's 32-bit stride overflow)
sc = syn.make_scene("tiny", opaque=True)
r = NerfRender(1); r.load_raw(sc.desc, sc.params, sc.grid, sc.aabb)
r.set_resolution((48, 40)); intr = syn.intrinsics(48, 40); pose = syn.look_at_pose(4.0, 20.0, 30.0)
for mode, bounds in ((1, [8, 32, 128]), (1, [1]), (1, [63, 64, 65]), (0, None), (2, None)):
    r.set_termination(mode, bounds); r.render_frame(intr, pose); r.render_frame(intr, pose)
r.set_shard(0, 1); r.set_resolution((48, 40))
h = r.peer_frame_export(); r.peer_frame_copy_out(True)
for k in range(4):
    r.render_frame_device(intr, pose)
    if k: r.peer_frame_collect(1)
r.peer_frame_collect(0); r.frame_join(); r.synchronize(); r.peer_frame_detach()
pos = np.random.rand(300, 3).astype(np.float32); d01 = np.random.rand(300, 3).astype(np.float32)
r.probe_encode(pos, want_indices=True); r.probe_sh(d01); r.probe_samples(pos, d01); r.level_scales()
r.close()
d = syn.make_desc(22, 1.0, 1)
r = NerfRender(1); r.load_raw(d, syn.random_params(d, 7, 1.0), syn.occupancy_grid(d, "sphere"), syn.aabb_of(d))
r.probe_encode(pos, want_indices=True); r.set_resolution((32, 24)); r.render_frame(syn.intrinsics(32, 24), pose)
r.close()
d = syn.make_desc(14, 1.0, 1, grid_size=32)
r = NerfRender(1); r.load_raw(d, syn.random_params(d, 1337, 1.0), syn.occupancy_grid(d, "sphere"), syn.aabb_of(d))
r.generate_density_grid(0.95, reset=False, seed=3, jitter=True)
r.close()
print("sanitize exercise done")
