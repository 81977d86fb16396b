"""Parity against an EXECUTION of the reference's own CUDA renderer on a B200.

tests/golden/ref_cuda_{tiny,tiny16,tiny_init}.npz were produced by baseline/run_ref_cuda.py on a B200 from
baseline/_ref/ref_cuda_driver_d1 — the reference's src/nerf_render.cu + tiny-cuda-nn compiled for sm_100 by
baseline/build_ref_cuda.sh (D1-clean variant) — on synthetic models handed over as .msgpack snapshots:
  net_*    NerfNetwork<__half>::inference_mixed_precision_impl (nerf_network.h:148-196) on 4096 seeded samples, plus the
           hash encoding, density-network output and SH encoding it computes on the way (tcnn kernel_grid, kernel_mlp_fused
           with wmma fp16 accumulators, kernel_sh);
  march_*  set_rays_o/d, kernel_near_far_from_aabb, init_step0, kernel_march_rays over whole rays (32x32 frame);
  frame_*  NerfRender::render_frame, 64x64, two poses (float image / depth before the u8 cast);
  level_scales  exp2f(level*log2(per_level_scale))*base-1 as kernel_grid evaluates it on the device (MUFU.EX2).

CPU half (-m "not gpu"): the oracle in its "cuda" numerics (a*b+c contracted exactly where the reference's SASS has an
FFMA, device level scales) reproduces the reference BIT FOR BIT in every stage except the MLP, whose wmma fp16 accumulation
no fp32-accumulating restatement can match bit for bit — there the bars are north_star's.  The "host" numerics (no
contraction, host exp2f) are counted against the same data: that is the FMA-contraction difference SURVEY.md §8(c) asks
to be reported, not hidden.
GPU half (-m gpu): the product's kernels against the same vectors, through the C ABI."""
import os

import numpy as np
import pytest

from nerf_cuda_b200 import synthetic as syn
from oracle import orc

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["tiny", "tiny16", "tiny_init"]
IMG_TOL, PSNR_MIN = 2.0 / 255.0, 45.0
# network bars against the reference's fp16-accumulating MLP: sigma and every rgb value with |ref| >= RGB_FLOOR within
# 1e-2 relative (north_star); below the floor (a sum of +- terms that cancels) the error is the fp16 accumulation noise of
# the hidden layers, bounded absolutely: 1e-3 = two fp16 ulps of an activation in [0.5, 1).
REL_TOL, RGB_FLOOR, ABS_TOL = 1e-2, 0.05, 1e-3


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def load(name):
    d = np.load(os.path.join(GOLD, f"ref_cuda_{name}.npz"))
    sc = syn.make_scene(str(d["scene"]), rgb_activation=int(d["rgb_activation"]), table_range=float(d["table_range"]))
    return d, sc


def psnr(a, b):
    mse = float(np.mean((np.clip(a, 0, 1).astype(np.float64) - np.clip(b, 0, 1).astype(np.float64)) ** 2))
    return 99.0 if mse == 0 else 10.0 * np.log10(1.0 / mse)


def check_network(out, ref, what):
    """out, ref [n,4] float {r,g,b,sigma}; returns the report line."""
    err = np.abs(out - ref)
    assert np.isfinite(out).all()
    assert err.max() <= ABS_TOL, f"{what}: max abs err {err.max()}"
    sig_rel = (err[:, 3] / np.maximum(np.abs(ref[:, 3]), 1e-30)).max()
    assert sig_rel <= REL_TOL, f"{what}: sigma rel err {sig_rel}"
    rgb_ref, rgb_err = ref[:, :3], err[:, :3]
    big = np.abs(rgb_ref) >= RGB_FLOOR
    rgb_rel = (rgb_err[big] / np.abs(rgb_ref[big])).max() if big.any() else 0.0
    assert rgb_rel <= REL_TOL, f"{what}: rgb rel err {rgb_rel}"
    return (f"{what}: max abs {err.max():.2e}, sigma rel {sig_rel:.2e}, rgb rel {rgb_rel:.2e} over {int(big.sum())} values "
            f">= {RGB_FLOOR}, {int((~big).sum())} values under the floor (abs err there <= {rgb_err[~big].max() if (~big).any() else 0:.2e})")


# ------------------------------------------------------------------------------------------------ CPU: oracle vs reference
@pytest.mark.parametrize("name", CASES)
def test_oracle_cuda_numerics_reproduce_the_reference_bit_for_bit(name, capsys):
    d, sc = load(name)
    O = orc.Oracle(sc.desc, sc.params, sc.grid, sc.aabb, numerics="cuda", scales=d["level_scales"])
    # hash encoding + SH: fp16 features bit-exact
    assert np.array_equal(O.encode(d["net_pos"]), d["net_enc_f16"])
    assert np.array_equal(O.sh(d["net_dirs01"]), d["net_sh_f16"])
    # rays, near/far, traversal
    W = H = int(d["march_size"])
    o, rd, nr, fr = O.rays(syn.intrinsics(W, H), d["march_pose"], W, H)
    assert np.array_equal(bits(rd), bits(d["march_rays_d"]))
    assert np.array_equal(bits(nr), bits(d["march_nears"])) and np.array_equal(bits(fr), bits(d["march_fars"]))
    cnt, smp = O.march(o, rd, nr, fr)
    assert np.array_equal(cnt, d["march_counts"])            # sample counts bit-exact on every ray
    off = np.concatenate([[0], np.cumsum(cnt)])
    for r in np.nonzero(cnt)[0]:
        s = smp[off[r]:off[r + 1]]
        assert np.array_equal(bits(s[:, :3]), bits(d["march_xyz"][r, :cnt[r]])), r        # positions
        assert np.array_equal(bits(s[:, 3]), bits(d["march_deltas"][r, :cnt[r], 0])), r    # dt
    assert cnt.sum() > 10000
    # network: MLP wiring (param order, den16 | sh16, sigma = exp(den[0])) against kernel_mlp_fused's outputs
    ref_out = orc.half_to_float(d["net_out_f16"])
    line = check_network(O.network(d["net_enc_f16"], d["net_dirs01"]), ref_out, f"{name} oracle MLP on the reference's encodings")
    line2 = check_network(O.samples(d["net_pos"], d["net_dirs01"]), ref_out, f"{name} oracle encode+MLP")
    # density network output (16 values per sample), hidden wiring included
    with capsys.disabled():
        print("\n  " + line + "\n  " + line2)


@pytest.mark.parametrize("name", CASES)
def test_oracle_frames_match_the_reference_renderer(name, capsys):
    d, sc = load(name)
    O = orc.Oracle(sc.desc, sc.params, sc.grid, sc.aabb, numerics="cuda", scales=d["level_scales"])
    W = int(d["frame_size"])
    for k in range(2):
        rgb, dep, cnt = O.render(d["frame_intr"], d["frame_poses"][k], W, W)
        ref, dref = d["frame_rgb"][k], d["frame_depth"][k]
        ok = ~np.isnan(dref)                                   # defect D3: missed rays give 0/0 in the reference
        assert np.abs(rgb - ref).max() <= IMG_TOL and psnr(rgb, ref) >= PSNR_MIN
        assert np.abs(dep[ok] - dref[ok]).max() <= 1e-3
        with capsys.disabled():
            print(f"\n  {name} frame {k}: oracle vs reference-CUDA render_frame max abs {np.abs(rgb - ref).max():.2e}, PSNR {psnr(rgb, ref):.1f} dB")


@pytest.mark.parametrize("name", CASES)
def test_fma_contraction_differences_are_counted(name, capsys):
    """The same vectors against the oracle's "host" numerics (expression trees as written, host exp2f): what differs
    between the reference compiled for the host and the reference compiled by nvcc.  Sample counts do not change."""
    d, sc = load(name)
    O = orc.Oracle(sc.desc, sc.params, sc.grid, sc.aabb)       # host numerics
    enc = O.encode(d["net_pos"])
    n_enc = int((enc != d["net_enc_f16"]).sum())
    host_scales = O.level_table()[1]
    n_scale = int((bits(host_scales) != bits(d["level_scales"])).sum())
    n_sh = int((O.sh(d["net_dirs01"]) != d["net_sh_f16"]).sum())
    W = H = int(d["march_size"])
    o, rd, nr, fr = O.rays(syn.intrinsics(W, H), d["march_pose"], W, H)
    cnt, smp = O.march(o, rd, nr, fr)
    assert np.array_equal(cnt, d["march_counts"])
    off = np.concatenate([[0], np.cumsum(cnt)])
    n_pos = sum(int((bits(smp[off[r]:off[r + 1], :3]) != bits(d["march_xyz"][r, :cnt[r]])).any(axis=1).sum()) for r in np.nonzero(cnt)[0])
    max_pos = max(float(np.abs(smp[off[r]:off[r + 1], :3] - d["march_xyz"][r, :cnt[r]]).max()) for r in np.nonzero(cnt)[0])
    n_far = int((bits(fr) != bits(d["march_fars"])).sum())
    assert n_enc > 0 and n_pos > 0 and max_pos < 1e-5 and n_scale > 0   # the two numerics really differ, by ulps
    with capsys.disabled():
        print(f"\n  {name} host-vs-nvcc numerics: {n_scale}/16 level scales differ (exp2f), {n_enc}/{enc.size} fp16 features, "
              f"{n_sh}/{d['net_sh_f16'].size} SH values, {n_pos}/{int(cnt.sum())} sample positions (max {max_pos:.1e}), "
              f"{n_far}/{fr.size} far values; sample counts identical")


# ------------------------------------------------------------------------------------------------ GPU: product vs reference
@pytest.fixture(scope="module", params=CASES)
def loaded(request, built_lib):
    from nerf_cuda_b200 import NerfRender
    d, sc = load(request.param)
    r = NerfRender(1)
    r.load_raw(sc.desc, sc.params, sc.grid, sc.aabb)
    yield request.param, d, sc, r
    r.close()


@pytest.mark.gpu
def test_product_stages_are_bit_exact_against_the_reference_cuda_run(loaded):
    name, d, sc, r = loaded
    assert np.array_equal(bits(r.level_scales()), bits(d["level_scales"]))     # same MUFU.EX2 scale on every level
    enc, idx = r.probe_encode(d["net_pos"], want_indices=True)
    assert np.array_equal(enc, d["net_enc_f16"])                               # hash-grid features bit-exact
    O = orc.cuda_oracle(r, sc.desc, sc.params, sc.grid, sc.aabb)
    assert np.array_equal(idx, O.encode(d["net_pos"], want_indices=True)[1])   # and the indices behind them
    assert np.array_equal(r.probe_sh(d["net_dirs01"]), d["net_sh_f16"])        # SH coefficients bit-exact
    W = H = int(d["march_size"])
    o, rd, nr, fr = r.probe_rays(syn.intrinsics(W, H), d["march_pose"], W, H)
    assert np.array_equal(bits(rd), bits(d["march_rays_d"]))
    assert np.array_equal(bits(nr), bits(d["march_nears"])) and np.array_equal(bits(fr), bits(d["march_fars"]))
    cnt, smp = r.probe_march(o, rd, nr, fr)
    assert np.array_equal(cnt, d["march_counts"])                              # occupancy traversal: counts bit-exact
    off = np.concatenate([[0], np.cumsum(cnt)])
    for q in np.nonzero(cnt)[0]:
        s = smp[off[q]:off[q + 1]]
        assert np.array_equal(bits(s[:, :3]), bits(d["march_xyz"][q, :cnt[q]])), q
        assert np.array_equal(bits(s[:, 3]), bits(d["march_deltas"][q, :cnt[q], 0])), q


@pytest.mark.gpu
@pytest.mark.parametrize("impl", [0, 1])
def test_product_network_against_the_reference_mlp(loaded, impl, capsys):
    name, d, sc, r = loaded
    ref_out = orc.half_to_float(d["net_out_f16"])
    what = "tcgen05" if impl == 0 else "SIMT check kernel"
    line = check_network(r.probe_network(d["net_enc_f16"], d["net_dirs01"], impl=impl), ref_out, f"{name} {what} on the reference's encodings")
    with capsys.disabled():
        print("\n  " + line)
    if impl == 0:
        line = check_network(r.probe_samples(d["net_pos"], d["net_dirs01"]), ref_out, f"{name} fused k_network")
        with capsys.disabled():
            print("  " + line)


@pytest.mark.gpu
def test_product_frames_against_the_reference_renderer(loaded, capsys):
    name, d, sc, r = loaded
    W = int(d["frame_size"])
    r.set_resolution((W, W))
    for k in range(2):
        r.render_frame(d["frame_intr"], d["frame_poses"][k])
        rgb, dep, cnt = r.float_frame()
        ref, dref = d["frame_rgb"][k], d["frame_depth"][k]
        ok = ~np.isnan(dref)
        assert np.abs(rgb - ref).max() <= IMG_TOL and psnr(rgb, ref) >= PSNR_MIN
        assert np.abs(dep[ok] - dref[ok]).max() <= 1e-3
        with capsys.disabled():
            print(f"\n  {name} frame {k}: product vs reference-CUDA render_frame max abs {np.abs(rgb - ref).max():.2e}, PSNR {psnr(rgb, ref):.1f} dB")
