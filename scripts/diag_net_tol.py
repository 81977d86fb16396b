import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from nerf_cuda_b200 import NerfRender, synthetic as syn
from oracle import orc
for name in ("tiny", "tiny16", "tiny3"):
    sc = syn.make_scene(name)
    r = NerfRender(1); r.load_raw(sc.desc, sc.params, sc.grid, sc.aabb)
    O = orc.Oracle(sc.desc, sc.params, sc.grid, sc.aabb)
    rng = np.random.default_rng(7)
    n = 5000
    pos = rng.random((n, 3), dtype=np.float32)
    dirs = rng.normal(size=(n, 3)).astype(np.float32)
    dirs = (dirs / np.linalg.norm(dirs, axis=1, keepdims=True) * 0.5 + 0.5).astype(np.float32)
    enc = O.encode(pos)
    ref = O.network(enc, dirs)
    for impl in (1, 0):
        out = r.probe_network(enc, dirs, impl=impl)
        err = np.abs(out - ref) / np.maximum(np.abs(ref), 1e-2)
        i, j = np.unravel_index(err.argmax(), err.shape)
        srt = np.sort(err.reshape(-1))[::-1][:5]
        print(name, "impl", impl, "max", err.max(), "at", (i, j), "out", out[i, j], "ref", ref[i, j], "top5", srt,
              "frac>5e-3", (err > 5e-3).mean())
    r.close()
