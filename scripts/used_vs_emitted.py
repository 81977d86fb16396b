"""How many emitted samples lie behind the point where their ray's transmittance fell below 1e-4 (composite early termination)?"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from nerf_cuda_b200 import NerfRender, synthetic as syn
wl = sys.argv[1] if len(sys.argv) > 1 else "C3"
cfg, W, H, _ = bench.WORKLOADS[wl]
sc = syn.make_scene(cfg)
r = NerfRender(1); r.load_raw(sc.desc, sc.params, sc.grid, sc.aabb); r.set_resolution((W, H))
for i, p in enumerate(bench.poses_for(wl)):
    r.render_frame_device(syn.intrinsics(W, H), p)
    _, _, used = r.float_frame()
    st = r.stats()
    print(i, "emitted", st.n_samples, "composited", int(used.sum()), f"{used.sum() / st.n_samples:.3f}", "max used", int(used.max()),
          "rays>256 used", int((used > 256).sum()))
