"""Parity at the configurations of BASELINE.json that the small scenes of test_gpu_parity.py do not reach: C5 (log2_hashmap_size
24: every fine level hashed into a 16 M-entry window, 1 GB table in HBM — tcnn grid.h:100-117 grid_index, :899-931 offset table),
tcnn's initial table range U(-1e-4, 1e-4), C2 and C4 at full size.  The oracle renders bands of full pixel rows (seconds on the
CPU); the rest of each frame is covered by the size-independent properties of test_full_size_c3_frame."""
import os

import numpy as np
import pytest

from nerf_cuda_b200 import NerfRender, synthetic as syn
from oracle import orc

pytestmark = pytest.mark.gpu
IMG_TOL = 2.0 / 255.0


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def check_bands(r, O, intr, pose, W, H, rows, n_rows=2):
    u8 = r.render_frame(intr, pose)[0].copy()
    rgb, dep, cnt = (x.copy() for x in r.float_frame())
    st = r.stats()
    assert u8.shape == (H, W, 3) and cnt.min() >= 0 and cnt.max() <= 1024
    assert st.n_samples_marched >= st.n_samples >= st.n_samples_used == int(cnt.sum()) > 0
    for row in rows:
        pix0, npix = row * W, n_rows * W
        rgb2, dep2, cnt2 = O.render(intr, pose, W, H, pix0, npix)
        sl = slice(pix0, pix0 + npix)
        assert np.array_equal(cnt[sl], cnt2), row                      # composited counts bit-exact
        assert np.abs(rgb[sl] - rgb2).max() <= IMG_TOL and np.abs(dep[sl] - dep2).max() <= 1e-3, row
    assert np.array_equal(r.render_frame(intr, pose)[0], u8)           # idempotent
    return u8, cnt


@pytest.fixture(scope="module")
def c5(built_lib):
    sc = syn.make_scene("C5")
    r = NerfRender(1)
    r.load_raw(sc.desc, sc.params, sc.grid, sc.aabb)
    O = orc.cuda_oracle(r, sc.desc, sc.params, sc.grid, sc.aabb)
    yield sc, r, O
    r.close()


def test_c5_level_table_and_encode_bit_exact(c5):
    sc, r, O = c5
    from nerf_cuda_b200 import level_table
    offs = level_table(sc.desc)[0]
    assert np.array_equal(offs, O.level_table()[0])
    per_level = np.diff(offs.astype(np.int64))
    assert per_level.max() == 1 << 24 and (per_level == 1 << 24).sum() == 7      # the seven finest levels fill the whole 2^24 window
    rng = np.random.default_rng(5)
    pos = rng.uniform(0, 1, size=(20000, 3)).astype(np.float32)
    pos[:8] = [[0, 0, 0], [1, 1, 1], [0.5, 0.5, 0.5], [1, 0, 0], [0, 1, 0], [0, 0, 1], [0.999999, 0.999999, 0.999999], [1e-7, 1e-7, 1e-7]]
    enc, idx = r.probe_encode(pos, want_indices=True)
    enc2, idx2 = O.encode(pos, want_indices=True)
    assert np.array_equal(idx, idx2)                                   # hashed and dense indices bit-exact
    assert np.array_equal(enc, enc2)                                   # fp16 features bit-exact
    assert idx.max() >= (1 << 23)                                      # the upper half of the window is really addressed


def test_c5_full_size_frame(c5):
    sc, r, O = c5
    W = H = 800
    r.set_resolution((W, H))
    intr, pose = syn.intrinsics(W, H), syn.look_at_pose(syn.camera_radius("C5"), 20.0, 45.0)
    check_bands(r, O, intr, pose, W, H, (5, H // 2 - 1, H - 9))


def test_init_range_table(built_lib):
    """tcnn initialises the table U(-1e-4, 1e-4) (grid.h:952-958): features are fp16 subnormal-adjacent, the encoding's output
    is ~1e-4 and the MLP sees nearly zero input — the interpolation must still be bit-exact and the frame within the bars."""
    sc = syn.make_scene("tiny", table_range=1e-4)
    r = NerfRender(1)
    r.load_raw(sc.desc, sc.params, sc.grid, sc.aabb)
    O = orc.cuda_oracle(r, sc.desc, sc.params, sc.grid, sc.aabb)
    pos = np.random.default_rng(11).uniform(0, 1, size=(5000, 3)).astype(np.float32)
    enc = r.probe_encode(pos)
    assert np.array_equal(enc, O.encode(pos))
    f = orc.half_to_float(enc)
    assert 0 < np.abs(f).max() <= 1.0001e-4
    W, H = 96, 72
    r.set_resolution((W, H))
    check_bands(r, O, syn.intrinsics(W, H), syn.look_at_pose(4.0, 20.0, 130.0), W, H, (0, 35, 70))
    r.close()


@pytest.mark.parametrize("wl,pi,rows", [("C2", 3, (4, 399, 790)), ("C4", 5, (7, 1079, 2150))])
def test_full_size_c2_and_c4_frames(built_lib, wl, pi, rows):
    import bench
    cfg, W, H, _ = bench.WORKLOADS[wl]
    sc = syn.make_scene(cfg)
    r = NerfRender(1)
    r.load_raw(sc.desc, sc.params, sc.grid, sc.aabb)
    r.set_resolution((W, H))
    O = orc.cuda_oracle(r, sc.desc, sc.params, sc.grid, sc.aabb)
    u8, cnt = check_bands(r, O, syn.intrinsics(W, H), bench.poses_for(wl)[pi], W, H, rows, n_rows=1 if wl == "C4" else 2)
    assert (cnt > 0).mean() > 0.2
    r.close()


def test_snapshot_written_by_the_reference_library_renders_like_its_arrays(built_lib):
    """tests/golden/ref_nlohmann.msgpack (bytes produced by the reference's vendored nlohmann::json::to_msgpack: float64 and
    integer-typed numbers, uint16/int16 encodings, alphabetical keys; 2 cascades, 32^3 grid, Sigmoid rgb, T = 2^8) loaded
    through reload_network_from_file renders the same bytes as the same arrays handed over with load_raw, and the frame
    agrees with the oracle built from those arrays (counts bit-exact)."""
    from test_snapshot_fixtures import GOLD, expected_arrays, inspect   # (tests/ is on sys.path: rootdir conftest)
    path = os.path.join(GOLD, "ref_nlohmann.msgpack")
    desc, n_p, n_g, aabb, _, _ = inspect(path)
    params, grid = expected_arrays()
    assert n_p == params.size and n_g == grid.size
    W, H = 80, 60
    intr, pose = syn.intrinsics(W, H), syn.look_at_pose(6.0, 25.0, 40.0)
    os.environ["NRF_SNAPSHOT_CACHE"] = "0"
    try:
        a = NerfRender(1)
        a.reload_network_from_file(path)
    finally:
        os.environ.pop("NRF_SNAPSHOT_CACHE", None)
    a.set_resolution((W, H))
    rgb_a, dep_a = (x.copy() for x in a.render_frame(intr, pose))
    fa, da, ca = (x.copy() for x in a.float_frame())
    a.close()
    b = NerfRender(1)
    b.load_raw(desc, params, grid, aabb.astype(np.float32))
    b.set_resolution((W, H))
    rgb_b, dep_b = (x.copy() for x in b.render_frame(intr, pose))
    O = orc.cuda_oracle(b, desc, params, grid, aabb.astype(np.float32))
    b.close()
    assert np.array_equal(rgb_a, rgb_b) and np.array_equal(dep_a, dep_b)
    rgb_o, dep_o, cnt_o = O.render(intr, pose, W, H)
    assert np.array_equal(ca, cnt_o) and ca.sum() > 0
    assert np.abs(fa - rgb_o).max() <= IMG_TOL and np.abs(da - dep_o).max() <= 1e-3


def test_c1_whole_frame_against_the_oracle(built_lib):
    """C1 (BASELINE.json configs[0], the CPU-runnable case: 200x200, C2's model, SURVEY.md §8d pose): the WHOLE frame against
    the oracle — every ray's composited count bit-exact, every pixel within the image bars, the u8 frame within 2 LSB of the
    cast of the oracle's floats (truncation at a .0 boundary may flip one LSB, never more)."""
    import bench
    cfg, W, H, _ = bench.WORKLOADS["C1"]
    sc = syn.make_scene(cfg)
    r = NerfRender(1)
    r.load_raw(sc.desc, sc.params, sc.grid, sc.aabb)
    r.set_resolution((W, H))
    intr, pose = syn.intrinsics(W, H), bench.poses_for("C1")[0]
    u8, d8 = (x.copy() for x in r.render_frame(intr, pose))
    rgb, dep, cnt = (x.copy() for x in r.float_frame())
    st = r.stats()
    O = orc.cuda_oracle(r, sc.desc, sc.params, sc.grid, sc.aabb)
    rgb2, dep2, cnt2 = O.render(intr, pose, W, H)
    assert np.array_equal(cnt, cnt2) and st.n_samples_used == int(cnt2.sum()) > 0
    assert np.abs(rgb - rgb2).max() <= IMG_TOL and np.abs(dep - dep2).max() <= 1e-3
    mse = float(np.mean((np.clip(rgb, 0, 1).astype(np.float64) - np.clip(rgb2, 0, 1)) ** 2))
    assert mse == 0 or 10 * np.log10(1.0 / mse) >= 45.0
    assert np.abs(u8.astype(int) - O.to_u8(rgb2).reshape(H, W, 3).astype(int)).max() <= 2
    r.close()
