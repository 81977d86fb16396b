"""Dev-container only (-m "not gpu", skipped where oracle/_ref/libnerf_ref.so is absent, e.g. on the GPU box): the oracle
against the REFERENCE'S OWN kernels compiled for the host, live, on larger seeded inputs than the committed fixtures
(tests/golden) cover: many poses, all cascades, wide fields of view with missing rays, 16k positions."""
import ctypes as C

import numpy as np
import pytest

from nerf_cuda_b200 import synthetic as syn
from oracle import orc

pytestmark = pytest.mark.skipif(not orc.ref_available(), reason="oracle/_ref not built (needs /root/reference)")
F, U16, U32, I32 = C.c_float, C.c_uint16, C.c_uint32, C.c_int


def p(a, ty):
    return a.ctypes.data_as(C.POINTER(ty))


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.fixture(scope="module", params=["tiny", "tiny16"])
def case(request):
    sc = syn.make_scene(request.param)
    return request.param, sc, orc.Oracle(sc.desc, sc.params, sc.grid, sc.aabb), orc.ref()


def ref_rays(R, sc, intr, pose, W, H):
    n = W * H
    ro, rd = np.zeros(3, np.float32), np.zeros((n, 3), np.float32)
    nr, fr = np.zeros(n, np.float32), np.zeros(n, np.float32)
    assert R.ref_rays(F(sc.desc.scale), p(intr, F), p(pose, F), W, n, p(sc.aabb, F), F(0.2), p(ro, F), p(rd, F), p(nr, F), p(fr, F)) == 0
    return ro, rd, nr, fr


@pytest.mark.parametrize("az,el,fov,rad_mul", [(0, 20, 50, 1.0), (123, -35, 90, 1.0), (271, 60, 120, 3.0), (45, 0, 30, 0.1)])
def test_rays_and_march(case, az, el, fov, rad_mul):
    name, sc, O, R = case
    W, H = 32, 20
    intr = syn.intrinsics(W, H, fov_deg=fov)
    pose = syn.look_at_pose(syn.camera_radius(name) * rad_mul, el, az)   # rad_mul 0.1: camera INSIDE the box
    ro, rd, nr, fr = ref_rays(R, sc, intr, pose, W, H)
    o2, d2, n2, f2 = O.rays(intr, pose, W, H)
    assert np.array_equal(bits(ro), bits(o2)) and np.array_equal(bits(rd), bits(d2))
    assert np.array_equal(bits(nr), bits(n2)) and np.array_equal(bits(fr), bits(f2))
    n, S, d = W * H, 1024, sc.desc
    xyz, dl, cnt = np.zeros((n, S, 3), np.float32), np.zeros((n, S, 2), np.float32), np.zeros(n, np.int32)
    assert R.ref_march(n, p(ro, F), p(rd, F), p(nr, F), p(fr, F), F(d.bound), F(1.0 / 128), int(d.cascades),
                       int(d.grid_size), p(sc.grid, F), F(d.mean_density), S, p(xyz, F), p(dl, F), p(cnt, I32)) == 0
    c2, smp = O.march(ro, rd, nr, fr)
    assert np.array_equal(cnt, c2)
    m = np.arange(S)[None, :] < cnt[:, None]
    assert np.array_equal(bits(smp[:, 0:3]), bits(xyz[m])) and np.array_equal(bits(smp[:, 3]), bits(dl[m][:, 0]))


def test_encode_and_sh(case):
    name, sc, O, R = case
    rng = np.random.default_rng(99)
    n = 16384
    pos = rng.random((n, 3), dtype=np.float32)
    pos[:3] = [[0, 0, 0], [1, 1, 1], [np.nextafter(np.float32(1), np.float32(0))] * 3]
    offs, _, _, pls = O.level_table()
    enc = np.zeros((n, 32), np.uint16)
    assert R.ref_encode(n, p(pos, F), p(O.table, U16), p(offs, U32), 16, U32(sc.desc.base_resolution),
                        F(np.log2(np.float32(pls), dtype=np.float32)), p(enc, U16)) == 0
    assert np.array_equal(O.encode(pos), enc)
    dirs = rng.random((n, 3), dtype=np.float32)
    sh = np.zeros((n, 16), np.uint16)
    assert R.ref_sh(n, p(dirs, F), p(sh, U16)) == 0
    assert np.array_equal(O.sh(dirs), sh)
