"""CPU tests (-m "not gpu"): the oracle (oracle/nerf_oracle.cpp) against the committed golden fixtures that were produced
by running the REFERENCE'S OWN kernels on the host (tests/golden/make_golden.py).  This is what pins the oracle; the GPU
parity tests (tests/test_gpu_parity.py) then compare the CUDA path with the pinned oracle.

Bars: rays, traversal, sample counts, hash-grid features (fp16 bit patterns) and SH coefficients BIT-EXACT; the whole
frame (reference host loop + reference march/composite kernels, network = oracle MLP through a callback) bit-exact in
sample counts and within 1e-6 absolute in colour/depth (the reference's 8-samples-per-iteration schedule and the oracle's
per-ray loop execute the same float operations in the same order, so this is in practice exact as well)."""
import hashlib
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))

import make_golden  # noqa: E402  (shares golden_inputs with the generator)
from oracle import orc  # noqa: E402
from nerf_cuda_b200 import synthetic as syn  # noqa: E402


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.fixture(scope="module", params=["tiny", "tiny16"])
def case(request):
    g = np.load(os.path.join(HERE, "golden", f"ref_{request.param}.npz"))
    sc, W, H, intr, pose, pos, dirs = make_golden.golden_inputs(request.param)
    assert sha(sc.params) == str(g["params_sha"]) and sha(sc.grid) == str(g["grid_sha"]), \
        "seeded inputs drifted from the ones the golden file was generated with"
    assert np.array_equal(pos, g["positions"]) and np.array_equal(dirs, g["dirs"])
    O = orc.Oracle(sc.desc, sc.params, sc.grid, sc.aabb)
    return g, sc, O, W, H, intr, pose


def test_rays_match_reference_kernels(case):
    g, sc, O, W, H, intr, pose = case
    o, d, nr, fr = O.rays(intr, pose, W, H)
    assert np.array_equal(bits(o), bits(g["rays_o"]))
    assert np.array_equal(bits(d), bits(g["rays_d"]))
    assert np.array_equal(bits(nr), bits(g["nears"])) and np.array_equal(bits(fr), bits(g["fars"]))


def test_march_matches_reference_kernel(case):
    g, sc, O, W, H, intr, pose = case
    cnt, smp = O.march(g["rays_o"], g["rays_d"], g["nears"], g["fars"])
    assert np.array_equal(cnt, g["march_counts"])
    assert cnt.sum() > 1000 and cnt.min() != cnt.max()     # ragged: rays differ in sample count (tiny also has empty rays)
    assert np.array_equal(bits(smp[:, 0:3]), bits(g["march_xyz"]))
    assert np.array_equal(bits(smp[:, 3]), bits(g["march_deltas"][:, 0]))   # deltas(0) = dt
    # deltas(1) = t - last_t (render_utils.h:629-631): first sample of a ray measures from near
    ray = smp[:, 7].view(np.int32)
    first = np.r_[True, ray[1:] != ray[:-1]]
    prev = np.where(first, g["nears"][ray], np.r_[np.float32(0), smp[:-1, 4]]).astype(np.float32)
    assert np.array_equal(bits(smp[:, 4] - prev), bits(g["march_deltas"][:, 1]))


def test_level_table_matches_golden(case):
    g, sc, O, *_ = case
    offs, scl, res, _ = O.level_table()
    assert np.array_equal(offs, g["offsets"]) and np.array_equal(bits(scl), bits(g["scales"]))
    assert np.array_equal(res, g["resolutions"])


def test_encode_matches_reference_kernel(case):
    g, sc, O, *_ = case
    assert np.array_equal(O.encode(g["positions"]), g["enc"])


def test_sh_matches_reference_kernel(case):
    g, sc, O, *_ = case
    assert np.array_equal(O.sh(g["dirs"]), g["sh"])


def test_frame_matches_reference_loop(case):
    g, sc, O, W, H, intr, pose = case
    rgb, dep, cnt = O.render(intr, pose, W, H)
    assert np.array_equal(cnt, g["frame_samples"])
    assert np.abs(rgb - g["frame_rgb"]).max() <= 1e-6
    assert np.abs(dep - g["frame_depth"]).max() <= 1e-6
    assert int(g["frame_iters"]) > 1


def test_u8_conversion():
    # nerf_render.cu:355-358: (unsigned char)(255.0 * v), truncation; clamped where the reference is UB (defect D4)
    O = orc.lib()
    v = np.array([0.0, 0.5, 0.999, 1.0, 1.5, -0.2, np.nan, 0.00392157], np.float32)
    out = np.zeros(v.size, np.uint8)
    import ctypes as C
    O.orc_to_u8(v.ctypes.data_as(C.POINTER(C.c_float)), out.ctypes.data_as(C.POINTER(C.c_uint8)), C.c_size_t(v.size))
    assert out.tolist() == [0, 127, 254, 255, 255, 0, 0, 1]


def test_pcg32_known_answers():
    """pcg32 (tcnn default_rng_t, used by the density-grid jitter): the published demo stream of the PCG reference
    implementation (seed 42, sequence 54) and skip-ahead == stepping."""
    import ctypes as C
    l = orc.lib()
    out = np.zeros(6, np.uint32)
    l.orc_pcg32(C.c_uint64(42), C.c_uint64(54), C.c_uint64(0), 6, out.ctypes.data_as(C.POINTER(C.c_uint32)))
    assert [hex(v) for v in out] == ["0xa15c02b7", "0x7b47f409", "0xba1d3330", "0x83d2f293", "0xbfa4784b", "0xcbed606e"]
    a, b = np.zeros(40, np.uint32), np.zeros(8, np.uint32)
    l.orc_pcg32(C.c_uint64(1337), C.c_uint64(1), C.c_uint64(0), 40, a.ctypes.data_as(C.POINTER(C.c_uint32)))
    l.orc_pcg32(C.c_uint64(1337), C.c_uint64(1), C.c_uint64(32), 8, b.ctypes.data_as(C.POINTER(C.c_uint32)))
    assert np.array_equal(a[32:], b)


def test_density_grid_oracle_properties():
    """generate_density_grid restatement (nerf_render.cu:388-429): reset starts from 1/64, the update is an EMA-max,
    cells marked negative are left alone, and cell centres are what the march's cell index addresses."""
    d = syn.make_desc(14, 1.0, 1, grid_size=16)
    sc = syn.Scene(d, syn.random_params(d, 1337, 1.0), syn.occupancy_grid(d, "sphere"), syn.aabb_of(d))
    O = orc.Oracle(sc.desc, sc.params, sc.grid, sc.aabb)
    g1, m1 = O.density_grid(0.95, reset=True)
    assert np.all(g1 >= np.float32(1.0 / 64) * np.float32(0.95)) and abs(m1 - g1.mean()) < 1e-6
    g2, _ = O.density_grid(0.5, reset=False, grid=g1)
    assert np.all(g2 <= g1) and np.all(g2 >= g1 * np.float32(0.5))
    neg = sc.grid.copy(); neg[:100] = -1.0
    g3, _ = O.density_grid(0.95, reset=False, grid=neg)
    assert np.all(g3[:100] == -1.0)
    # sigma itself: the floor of the EMA with decay 0 is 0.001691 * sigma(cell centre)
    g4, _ = O.density_grid(0.0, reset=False, grid=np.zeros_like(sc.grid))
    H = d.grid_size
    idx = np.array([0, 5, H * H * 3 + H * 2 + 1, H ** 3 - 1])
    ids = np.stack([idx // (H * H), idx % (H * H) // H, idx % H], 1).astype(np.float32)
    centres = (np.float32(-1.0) + np.float32(2.0) / np.float32(H - 1) * ids) * np.float32(1.0 - 1.0 / H)
    sig = O.samples((centres * np.float32(0.5) + np.float32(0.5)).astype(np.float32), np.full((4, 3), 0.5, np.float32))[:, 3]
    assert np.allclose(g4[idx], np.float32(0.001691) * sig, rtol=1e-6)
    gj, _ = O.density_grid(0.0, reset=False, seed=5, jitter=True, grid=np.zeros_like(sc.grid))
    assert not np.array_equal(gj, g4)
