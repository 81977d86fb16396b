"""nrf_render_batch throughput (pipelined) vs one-at-a-time nrf_render_frame + copy, host frames out in both cases."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from nerf_cuda_b200 import NerfRender, synthetic as syn
wl = sys.argv[1] if len(sys.argv) > 1 else "C3"
cfg, W, H, _ = bench.WORKLOADS[wl]
sc = syn.make_scene(cfg)
r = NerfRender(1); r.load_raw(sc.desc, sc.params, sc.grid, sc.aabb); r.set_resolution((W, H))
intr, poses = syn.intrinsics(W, H), np.stack(bench.poses_for(wl) * 2)
r.render_batch(intr, poses)   # sizes every buffer
t0 = time.perf_counter(); out = r.render_batch(intr, poses); t1 = time.perf_counter()
keep = np.empty_like(out)
t2 = time.perf_counter()
for k, p in enumerate(poses):
    keep[k] = r.render_frame(intr, p)[0]
t3 = time.perf_counter()
n = len(poses)
print(f"{wl}: batch {1e3 * (t1 - t0) / n:.3f} ms/frame ({W * H * n / (t1 - t0) / 1e6:.1f} Mrays/s), "
      f"single frames + copy {1e3 * (t3 - t2) / n:.3f} ms/frame ({W * H * n / (t3 - t2) / 1e6:.1f} Mrays/s), identical: {np.array_equal(out, keep)}")
