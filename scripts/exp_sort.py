"""Experiment: how much does spatial ordering of samples speed up the hash-grid gather?"""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from nerf_cuda_b200 import NerfRender, synthetic as syn
from oracle import orc
cfg = "C2"
sc = syn.make_scene(cfg)
r = NerfRender(1); r.load_raw(sc.desc, sc.params, sc.grid, sc.aabb)
O = orc.Oracle(sc.desc, sc.params, sc.grid, sc.aabb)
W = 300
intr, pose = syn.intrinsics(W, W), syn.look_at_pose(4.0, 20.0, 45.0)
o, d, nr, fr = O.rays(intr, pose, W, W)
cnt, smp = O.march(o, d, nr, fr)
pos = (smp[:, :3] * 0.5 + 0.5).astype(np.float32)
n = pos.shape[0]
print("samples", n)
def bench(p, name):
    p = np.ascontiguousarray(p, np.float32)
    ms = C.c_float()
    rc = r._l.nrf_bench_encode_positions(r._ctx, p.shape[0], p.ctypes.data_as(C.POINTER(C.c_float)), 3, C.byref(ms))
    assert rc == 0
    print(f"{name:40s} {ms.value:8.3f} ms  {p.shape[0]/ms.value/1e6:8.2f} Gsamples/s")
bench(pos, "ray order (as rendered)")
rng = np.random.default_rng(0)
bench(pos[rng.permutation(n)], "random order")
for res in (2048, 1024, 512):
    g = np.floor(pos * res).astype(np.int64)
    key = (g[:, 2] * res + g[:, 1]) * res + g[:, 0]
    bench(pos[np.argsort(key, kind="stable")], f"sorted by (z,y,x) cell at res {res}")
g = np.floor(pos * 2048).astype(np.int64)
key = (g[:, 2] * 2048 + g[:, 1]) * 64 + (g[:, 0] >> 5)
bench(pos[np.argsort(key, kind="stable")], "sorted by (z,y,x>>5) at res 2048")
# coarse bucket then ray order inside (cheap partial sort)
for res in (64, 128, 256):
    g = np.floor(pos * res).astype(np.int64)
    key = (g[:, 2] * res + g[:, 1]) * res + g[:, 0]
    bench(pos[np.argsort(key, kind="stable")], f"bucketed by cell at res {res}, ray order inside")
