"""Small driver for ncu: one C2-scene frame at reduced resolution, then each network stage once."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_cuda_b200 import NerfRender, synthetic as syn
cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
W = int(sys.argv[2]) if len(sys.argv) > 2 else 400
H = int(sys.argv[3]) if len(sys.argv) > 3 else W
sc = syn.make_scene(cfg)
r = NerfRender(1)
r.load_raw(sc.desc, sc.params, sc.grid, sc.aabb)
r.set_resolution((W, H))
intr, pose = syn.intrinsics(W, H), syn.look_at_pose(syn.camera_radius(cfg), 20.0, 45.0)
for _ in range(2):
    r.render_frame_device(intr, pose)
r.synchronize()
st = r.stats()
print("samples", st.n_samples, "ms", st.ms_total, st.ms_setup, st.ms_march, st.ms_network, st.ms_composite)
for what in (0, 1, 2):
    print("stage", what, r.bench_stage(what, 0, 2), "ms")
