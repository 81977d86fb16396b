"""Samples per pose of a bench workload (to attribute an ncu capture of one launch)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from nerf_cuda_b200 import NerfRender, synthetic as syn
wl = sys.argv[1] if len(sys.argv) > 1 else "C3"
cfg, W, H, _ = bench.WORKLOADS[wl]
sc = syn.make_scene(cfg)
r = NerfRender(1); r.load_raw(sc.desc, sc.params, sc.grid, sc.aabb); r.set_resolution((W, H))
out = []
for p in bench.poses_for(wl):
    r.render_frame_device(syn.intrinsics(W, H), p); st = r.stats()
    out.append({"samples": int(st.n_samples), "ms_total": st.ms_total, "ms_setup": st.ms_setup, "ms_march": st.ms_march,
                "ms_network": st.ms_network, "ms_composite": st.ms_composite})
print(json.dumps(out))
