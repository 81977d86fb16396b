"""Per-frame wall time of the synchronous host call (render_frame: enqueue + D2H + sync) against the device time the same
frame reports, two cycles over the workload's poses: python scripts/e2e_frames.py C4"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from nerf_cuda_b200 import NerfRender, synthetic as syn
wl = sys.argv[1] if len(sys.argv) > 1 else "C3"
cfg, W, H, _ = bench.WORKLOADS[wl]
sc = syn.make_scene(cfg)
r = NerfRender(1); r.load_raw(sc.desc, sc.params, sc.grid, sc.aabb); r.set_resolution((W, H))
intr = syn.intrinsics(W, H)
for cycle in range(3):
    for i, p in enumerate(bench.poses_for(wl)):
        t0 = time.perf_counter(); r.render_frame(intr, p); wall = (time.perf_counter() - t0) * 1e3
        st = r.stats()
        print(cycle, i, int(st.n_samples), "wall %.2f" % wall, "device %.2f" % st.ms_total, flush=True)
