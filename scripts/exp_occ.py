"""Experiment: encode-only gather vs resident warps per SM."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_cuda_b200 import NerfRender, synthetic as syn
for cfg, W in (("C2", 800), ("C3", 1000)):
    sc = syn.make_scene(cfg)
    r = NerfRender(1); r.load_raw(sc.desc, sc.params, sc.grid, sc.aabb)
    r.set_resolution((W, W))
    r.render_frame_device(syn.intrinsics(W, W), syn.look_at_pose(syn.camera_radius(cfg), 20.0, 45.0)); r.synchronize()
    print(cfg, "samples", r.stats().n_samples)
    for v, name in ((3, "16 warps"), (4, "24 warps"), (0, "32 warps/64 regs"), (1, "48 warps/40 regs"), (2, "64 warps/32 regs")):
        ms = C.c_float()
        assert r._l.nrf_exp_encode(r._ctx, v, 3, C.byref(ms)) == 0
        print(f"  {name:20s} {ms.value:7.3f} ms")
    r.close()
