#!/usr/bin/env python
"""Generates tests/golden/*.npz by EXECUTING THE REFERENCE'S OWN KERNELS on the host (oracle/_ref/libnerf_ref.so,
built by oracle/build_ref.sh from /root/reference — dev container only).  The reference ships no golden vectors
(SURVEY.md §4), so these fixtures are what pins oracle/nerf_oracle.cpp; tests/test_oracle_golden.py replays them
anywhere (no /root/reference needed).

Inputs are regenerated from seeds by nerf_cuda_b200.synthetic (numpy PCG64 streams are stable by policy); each file
stores a sha256 of the parameter / grid arrays it was made from so that drift is detected instead of mis-compared.

Usage (dev container):  make -C oracle ref && python tests/golden/make_golden.py
"""
import ctypes as C
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from nerf_cuda_b200 import synthetic as syn  # noqa: E402
from oracle import orc  # noqa: E402

F, U16, U32, I32 = C.c_float, C.c_uint16, C.c_uint32, C.c_int


def p(a, ty):
    return a.ctypes.data_as(C.POINTER(ty))


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def golden_inputs(name):
    """The seeded inputs every golden case uses (shared with tests/test_oracle_golden.py)."""
    sc = syn.make_scene(name)
    W, H = 24, 16
    intr = syn.intrinsics(W, H)
    pose = syn.look_at_pose(syn.camera_radius(name), 20.0, 33.0)
    rng = np.random.default_rng(2024)
    pos = rng.random((1024, 3), dtype=np.float32)
    pos[:8] = [[0, 0, 0], [1, 1, 1], [0.5, 0.5, 0.5], [1, 0, 0], [0, 1, 0], [0, 0, 1], [0.999999, 0.5, 0.25], [1e-7, 0.3, 0.7]]
    dirs = rng.normal(size=(1024, 3)).astype(np.float32)
    dirs = (dirs / np.linalg.norm(dirs, axis=1, keepdims=True) * 0.5 + 0.5).astype(np.float32)
    return sc, W, H, intr, pose, pos, dirs


def main():
    if not orc.ref_available():
        raise SystemExit("oracle/_ref/libnerf_ref.so missing: run `make -C oracle ref` in the dev container")
    R = orc.ref()
    for name in ("tiny", "tiny16"):
        sc, W, H, intr, pose, pos, dirs = golden_inputs(name)
        d = sc.desc
        O = orc.Oracle(d, sc.params, sc.grid, sc.aabb)
        n = W * H
        # --- rays: set_rays_o/set_rays_d + kernel_near_far_from_aabb (reference NGPU=2 interleave, re-ordered to pixels)
        ro, rd = np.zeros(3, np.float32), np.zeros((n, 3), np.float32)
        nr, fr = np.zeros(n, np.float32), np.zeros(n, np.float32)
        assert R.ref_rays(F(d.scale), p(intr, F), p(pose, F), W, n, p(sc.aabb, F), F(0.2), p(ro, F), p(rd, F), p(nr, F), p(fr, F)) == 0
        # --- march: kernel_march_rays over whole rays (n_step = 1024, deltas pre-zeroed = D1 clean)
        S = 1024
        xyz = np.zeros((n, S, 3), np.float32)
        dl = np.zeros((n, S, 2), np.float32)
        cnt = np.zeros(n, np.int32)
        assert R.ref_march(n, p(ro, F), p(rd, F), p(nr, F), p(fr, F), F(d.bound), F(1.0 / 128), int(d.cascades),
                           int(d.grid_size), p(sc.grid, F), F(d.mean_density), S, p(xyz, F), p(dl, F), p(cnt, I32)) == 0
        m = np.arange(S)[None, :] < cnt[:, None]
        march_xyz, march_dl = xyz[m], dl[m]      # ray-major compaction
        # --- hash grid: kernel_grid<__half,3,2>
        offs, scales, res, pls = O.level_table()
        enc = np.zeros((pos.shape[0], 32), np.uint16)
        assert R.ref_encode(pos.shape[0], p(pos, F), p(O.table, U16), p(offs, U32), 16, U32(d.base_resolution),
                            F(np.log2(np.float32(pls), dtype=np.float32)), p(enc, U16)) == 0
        # --- SH: kernel_sh<__half>, degree 4
        sh = np.zeros((dirs.shape[0], 16), np.uint16)
        assert R.ref_sh(dirs.shape[0], p(dirs, F), p(sh, U16)) == 0
        # --- whole frame: the reference's host loop + its own march/composite kernels; the network (wmma, cannot run
        # on the host) is the oracle's scalar MLP + encode through a callback.
        NETFN = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_int, C.POINTER(F), C.POINTER(U16))

        def net(_user, n_cols, B, net_in, net_out):
            a = np.ctypeslib.as_array(net_in, (6, B))
            o = np.ctypeslib.as_array(net_out, (16, B))
            out = O.samples(np.ascontiguousarray(a[0:3, :n_cols].T), np.ascontiguousarray(a[3:6, :n_cols].T))
            o[0:4, :n_cols] = out.astype(np.float16).view(np.uint16).T   # values are exact halves already

        img, dep, ws = np.zeros((n, 3), np.float32), np.zeros(n, np.float32), np.zeros(n, np.float32)
        spr = np.zeros(n, np.int32)
        R.ref_render_loop.restype = C.c_int
        iters = R.ref_render_loop(n, p(ro, F), p(rd, F), p(nr, F), p(fr, F), F(d.bound), F(1.0 / 128), int(d.cascades),
                                  int(d.grid_size), p(sc.grid, F), F(d.mean_density), NETFN(net), None, p(img, F),
                                  p(dep, F), p(ws, F), p(spr, I32))
        out = os.path.join(HERE, f"ref_{name}.npz")
        np.savez_compressed(
            out, params_sha=sha(sc.params), grid_sha=sha(sc.grid), W=W, H=H, intr=intr, pose=pose,
            rays_o=ro, rays_d=rd, nears=nr, fars=fr, march_counts=cnt, march_xyz=march_xyz, march_deltas=march_dl,
            positions=pos, enc=enc, dirs=dirs, sh=sh, offsets=offs, scales=scales, resolutions=res,
            frame_rgb=img, frame_depth=dep, frame_wsum=ws, frame_samples=spr, frame_iters=iters)
        print(f"{out}: {os.path.getsize(out) / 1024:.0f} KB, {int(cnt.sum())} march samples, {iters} loop iterations, "
              f"{int(spr.sum())} composited samples")


if __name__ == "__main__":
    main()
