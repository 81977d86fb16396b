"""Roofline denominators measured on the box (nrf_bench_memory): L2 / HBM streaming bandwidth and the random
8-byte-slot gather rate over windows the size of one hashed level (2 MB), the C2/C3 table (24 MB) and the C5 table (512 MB)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_cuda_b200 import NerfRender, synthetic as syn
sc = syn.make_scene("tiny")
r = NerfRender(1)
r.load_raw(sc.desc, sc.params, sc.grid, sc.aabb)
out = {}
for name, mode, win in (("stream_split_24MB(L1-contaminated)", 0, 24 << 20), ("stream_split_64MB", 0, 64 << 20),
                        ("stream_l2_whole_8MB", 2, 8 << 20), ("stream_l2_whole_24MB", 2, 24 << 20), ("stream_l2_whole_28MB", 2, 28 << 20),
                        ("stream_l2_whole_64MB", 2, 64 << 20), ("stream_hbm_4GB", 0, 4 << 30),
                        ("gather_2MB", 1, 2 << 20), ("gather_24MB", 1, 24 << 20), ("gather_512MB", 1, 512 << 20)):
    g, l = r.bench_memory(mode, win, 3)
    out[name] = {"GB/s_requested": round(g, 1), "Gloads/s": round(l, 2),
                 **({"GB/s_sectors(32B)": round(l * 32, 1), "lanes_per_clk_per_SM@1965MHz": round(l / 148 / 1.965, 2)} if mode == 1 else {})}
    print(name, out[name], flush=True)
r.close()
print(json.dumps(out))
