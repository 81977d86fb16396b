"""Termination-aware evaluation rounds (-m gpu): a frame evaluated in rounds of samples per ray — rays whose
transmittance fell below 1e-4 are neither marched, evaluated nor composited in the next round, as the reference drops
them from its next wavefront (render_utils.h:727 rays_t = -1, kernel_compact_rays :394-415) — must give exactly the
pixels, depths and composited counts of the whole-ray frame and of the oracle, for every round schedule, while the
network evaluates far fewer samples on a scene where rays do terminate."""
import numpy as np
import pytest

from nerf_cuda_b200 import NerfRender, synthetic as syn
from oracle import orc

pytestmark = pytest.mark.gpu
IMG_TOL = 2.0 / 255.0


def _render(r, intr, pose):
    u8, d8 = (a.copy() for a in r.render_frame(intr, pose))
    rgb, dep, cnt = (a.copy() for a in r.float_frame())
    return u8, d8, rgb, dep, cnt, r.stats()


@pytest.mark.parametrize("name,opaque", [("tiny", True), ("tiny16", True), ("tiny3", True), ("tiny", False)])
def test_round_schedules_give_identical_frames(built_lib, name, opaque):
    sc = syn.make_scene(name, opaque=opaque)
    W, H = 96, 64
    intr, pose = syn.intrinsics(W, H), syn.look_at_pose(syn.camera_radius(name), 20.0, 40.0)
    r = NerfRender(1)
    r.load_raw(sc.desc, sc.params, sc.grid, sc.aabb)
    r.set_resolution((W, H))
    r.set_termination(0)
    ref = _render(r, intr, pose)
    assert ref[5].n_rounds == 1 and ref[5].n_samples == ref[5].n_samples_marched
    O = orc.cuda_oracle(r, sc.desc, sc.params, sc.grid, sc.aabb)
    rgb_o, dep_o, cnt_o = O.render(intr, pose, W, H)
    assert np.array_equal(ref[4], cnt_o)                       # composited counts bit-exact vs oracle
    assert np.abs(ref[2] - rgb_o).max() <= IMG_TOL
    assert ref[5].n_samples_used == int(cnt_o.sum())
    o, d, nr, fr = O.rays(intr, pose, W, H)
    c_all, _ = O.march(o, d, nr, fr, want_samples=False)
    assert ref[5].n_samples_marched == int(c_all.sum())
    if opaque:
        assert cnt_o.sum() < 0.5 * c_all.sum()                 # the scene really terminates rays
    for bounds in ([16, 64], [1], [8, 70, 200], [63, 64, 65], [1000]):
        r.set_termination(1, bounds)
        got = _render(r, intr, pose)
        assert got[5].n_rounds == len(bounds) + 1
        for a, b in zip(ref[:5], got[:5]):
            assert np.array_equal(a, b), bounds               # bytes, floats and counts identical to the whole-ray frame
        assert got[5].n_samples_marched == ref[5].n_samples_marched and got[5].n_samples_used == ref[5].n_samples_used
        assert got[5].n_samples_used <= got[5].n_samples <= got[5].n_samples_marched
    if opaque:
        r.set_termination(1, [8, 32, 128])                     # the default schedule
        st = _render(r, intr, pose)[5]
        assert st.n_samples < 0.9 * st.n_samples_marched
        if name == "tiny":                                     # up to 128 samples per ray through the sphere
            assert st.n_samples * 2 <= st.n_samples_marched    # >= 2x fewer network samples
    r.close()


def test_rounds_beyond_the_t_cache_resume_the_march(built_lib):
    """Every cell occupied: hundreds of samples per ray, far beyond the 64 cached per ray, so later rounds resume the
    march from the state the previous round left (t_resume); opaque and transparent weights."""
    for opaque in (False, True):
        sc = syn.make_scene("tiny16", opaque=opaque)
        grid = np.ones_like(sc.grid)
        W, H = 48, 32
        intr, pose = syn.intrinsics(W, H), syn.look_at_pose(12.0, 20.0, 33.0)
        r = NerfRender(1)
        r.load_raw(sc.desc, sc.params, grid, sc.aabb)
        r.set_resolution((W, H))
        r.set_termination(0)
        ref = _render(r, intr, pose)
        assert ref[5].n_samples_marched > 100 * W * H
        for bounds in ([16, 64], [100, 300, 700], [64]):
            r.set_termination(1, bounds)
            got = _render(r, intr, pose)
            for a, b in zip(ref[:5], got[:5]):
                assert np.array_equal(a, b), (opaque, bounds)
        O = orc.cuda_oracle(r, sc.desc, sc.params, grid, sc.aabb)
        rgb_o, dep_o, cnt_o = O.render(intr, pose, W, H)
        assert np.array_equal(ref[4], cnt_o) and np.abs(ref[2] - rgb_o).max() <= IMG_TOL
        r.close()


def test_automatic_mode_follows_the_scene(built_lib):
    """mode 2 (default): a context starts with whole rays; once a validated frame shows that less than 3/4 of the marched
    samples were composited, the following frames run in rounds — and never on a scene where nothing terminates."""
    W, H = 64, 48
    intr = syn.intrinsics(W, H)
    for opaque, want_rounds in ((True, 4), (False, 1)):
        sc = syn.make_scene("tiny", opaque=opaque)
        r = NerfRender(1)
        r.load_raw(sc.desc, sc.params, sc.grid, sc.aabb)
        r.set_resolution((W, H))
        frames = []
        for k in range(3):
            pose = syn.look_at_pose(4.0, 20.0, 30.0)
            frames.append(_render(r, intr, pose))
        assert frames[0][5].n_rounds == 1      # (also when the first frame had to be rendered again to size the buffers)
        assert frames[1][5].n_rounds == want_rounds and frames[2][5].n_rounds == want_rounds
        for f in frames[1:]:
            for a, b in zip(frames[0][:5], f[:5]):
                assert np.array_equal(a, b)
        if opaque:
            assert frames[2][5].n_samples * 2 <= frames[2][5].n_samples_marched == frames[0][5].n_samples_marched
        # asynchronous frames + batch call take the same decision and stay identical
        r.render_frame_device(intr, syn.look_at_pose(4.0, 20.0, 30.0))
        r.synchronize()
        assert np.array_equal(r.float_frame()[0], frames[0][2])
        out = r.render_batch(intr, np.stack([syn.look_at_pose(4.0, 20.0, 30.0)] * 3))
        assert all(np.array_equal(out[k].reshape(H, W, 3), frames[0][0]) for k in range(3))
        r.close()


def test_set_termination_rejects_bad_schedules(built_lib):
    sc = syn.make_scene("tiny")
    r = NerfRender(1)
    r.load_raw(sc.desc, sc.params, sc.grid, sc.aabb)
    for bad in ([0], [1024], [5, 5], [9, 3], [1, 2, 3, 4]):
        with pytest.raises(Exception):
            r.set_termination(1, bad)
    with pytest.raises(Exception):
        r.set_termination(3)
    r.close()
