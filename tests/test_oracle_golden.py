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
