"""CPU tests (-m "not gpu"): host logic of the product library and the drop-in boundary.  No compute entry point is
called (there is no GPU here and the library has no CPU path) — these check that the C-ABI library loads, exports every
symbol include/*.h declares, fails loudly without a device, and that the host-side arithmetic (level table, parameter
count, shard map, snapshot reader) matches the oracle / the reference's rules."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from nerf_cuda_b200 import NerfError, capi, synthetic as syn
from nerf_cuda_b200.snapshot import write_snapshot
from oracle import orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = []
    for h in ("nerf_b200.h", "nerf_b200_ext.h"):
        src = open(os.path.join(ROOT, "include", h)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names += re.findall(r"\b(nrf_[a-z0-9_]+)\s*\(", src)
    return sorted(set(names))


def test_library_exports_every_declared_symbol(built_lib):
    lib = capi.lib()
    syms = declared_symbols()
    assert len(syms) >= 25
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/*.h but not exported by libnerf_b200.so"
    assert set(capi.EXPORTS) <= set(syms)
    assert b"sm_100a" in lib.nrf_version()
    # ... and nothing else: kernel stubs, cross-TU launch helpers and experiment hooks are local (linker version script)
    import subprocess
    out = subprocess.run(["nm", "-D", "--defined-only", built_lib], capture_output=True, text=True, check=True).stdout
    exported = sorted(ln.split()[-1].split("@")[0] for ln in out.splitlines() if len(ln.split()) == 3 and ln.split()[1] in "TDBR")
    assert exported == syms, sorted(set(exported) ^ set(syms))


def test_no_cpu_fallback(built_lib):
    """Without a CUDA device every path to compute fails with NRF_ENODEV; nothing silently runs on the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(NerfError) as e:
        capi.NerfRender(1)
    assert e.value.code == -2 and "no CPU path" in str(e.value)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "nerf_cuda_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f)).read()
                for pat in (r"#\s*include[^\n]*oracle", r"^\s*(from|import)\s+oracle", r"import\s+orc\b", r"libnerf_oracle",
                            r"libnerf_ref", r"dlopen"):
                    assert not re.search(pat, src, flags=re.M), f"{f} references the oracle ({pat})"


@pytest.mark.parametrize("name", ["C2", "C3", "C5", "tiny", "tiny16"])
def test_level_table_and_param_count(built_lib, name):
    d = {"C2": syn.make_desc(19, 1.0, 1), "C3": syn.make_desc(19, 16.0, 5), "C5": syn.make_desc(24, 1.0, 1),
         "tiny": syn.make_desc(14, 1.0, 1), "tiny16": syn.make_desc(14, 16.0, 5)}[name]
    offs, scl, res = capi.level_table(d)
    o2, s2, r2, _ = syn.level_table(d)                       # numpy restatement
    # numpy's exp2 may differ from glibc exp2f by an ulp: sizes must agree exactly, scales to 2 ulp
    assert np.array_equal(offs, o2) and np.array_equal(res, r2)
    assert np.abs(scl.view(np.int32).astype(np.int64) - s2.view(np.int32)).max() <= 2
    o3, s3, r3 = np.zeros(17, np.uint32), np.zeros(16, np.float32), np.zeros(16, np.uint32)
    pls = C.c_float()
    assert orc.lib().orc_level_table(C.byref(d), o3.ctypes.data_as(C.POINTER(C.c_uint32)), s3.ctypes.data_as(C.POINTER(C.c_float)),
                                     r3.ctypes.data_as(C.POINTER(C.c_uint32)), C.byref(pls)) == 0
    assert np.array_equal(offs, o3) and np.array_equal(scl.view(np.uint32), s3.view(np.uint32)) and np.array_equal(res, r3)
    assert capi.param_count(d) == 10240 + 2 * int(offs[-1])
    if name == "C2":   # verified by running the reference ctor math on the host
        assert [int(offs[i + 1] - offs[i]) for i in range(6)] == [4096, 12168, 29792, 79512, 205384, 524288]
        assert capi.param_count(d) == 10240 + 12196240


def test_bad_descriptions_are_rejected(built_lib):
    d = syn.make_desc()
    d.n_levels = 0
    with pytest.raises(NerfError):
        capi.level_table(d)
    d = syn.make_desc(log2_hashmap_size=40)
    with pytest.raises(NerfError):
        capi.param_count(d)


def _ray_pixel(W, H, shard, n_shards, q):
    """Independent restatement of ShardMap (csrc/nrf_device.cuh): q = local_tile*1024 + patch*32 + lane; tiles of 32x32
    pixels round-robin over shards; 4x8 patches of 8x4 pixels per tile; lane = 8*row + col inside the patch."""
    tiles_x = -(-W // 32)
    lt, w = divmod(q, 1024)
    patch, lane = divmod(w, 32)
    tile = lt * n_shards + shard
    ty, tx = divmod(tile, tiles_x)
    px = tx * 32 + (patch % 4) * 8 + lane % 8
    py = ty * 32 + (patch // 4) * 4 + lane // 8
    return py * W + px if px < W and py < H else -1


@pytest.mark.parametrize("W,H,world", [(100, 80, 1), (100, 80, 3), (1920, 1080, 8), (33, 17, 2), (64, 64, 4), (31, 5, 8)])
def test_tile_sharding_partitions_the_frame(W, H, world):
    """FrameSharding (host side of the N>1 path) against the ray order the kernels use: shards are disjoint, cover every
    pixel exactly once, and a 32-lane warp is one 8x4-pixel patch."""
    from nerf_cuda_b200.multigpu import FrameSharding
    sh = FrameSharding(W, H, world)
    seen = np.zeros(W * H, np.int32)
    for s in range(world):
        pk = sh.packed_pixels(s)
        assert pk.shape == (sh.local_tiles(s), 32, 32) and sh.local_count(s) == pk.size
        valid = pk[pk >= 0]
        seen[valid] += 1
        # ray order -> pixel, checked on a sample of slots (all of them for small frames)
        n = sh.local_count(s)
        qs = np.arange(n) if n <= 8192 else np.random.default_rng(s).integers(0, n, 4096)
        for q in qs:
            p = _ray_pixel(W, H, s, world, int(q))
            lt, w = divmod(int(q), 1024)
            patch, lane = divmod(w, 32)
            yin, xin = (patch // 4) * 4 + lane // 8, (patch % 4) * 8 + lane % 8
            assert pk[lt, yin, xin] == p
    assert (seen == 1).all()
    assert sum(sh.local_tiles(s) for s in range(world)) == sh.n_tiles


def _inspect(path):
    lib = capi.lib()
    d = syn.ModelDesc()
    n_p, n_g = C.c_uint64(), C.c_uint64()
    aabb = (C.c_float * 6)()
    ps, gs = C.c_double(), C.c_double()
    rc = lib.nrf_snapshot_inspect(os.fsencode(path), C.byref(d), C.byref(n_p), C.byref(n_g), aabb, C.byref(ps), C.byref(gs))
    if rc:
        raise NerfError(rc, lib.nrf_last_error().decode())
    return d, n_p.value, n_g.value, list(aabb), ps.value, gs.value


def test_snapshot_reader_roundtrip(built_lib, tmp_path):
    sc = syn.make_scene("tiny16", rgb_activation=syn.NRF_ACT_SIGMOID)
    path = str(tmp_path / "synthetic.msgpack")
    write_snapshot(path, sc.desc, sc.params, sc.grid, sc.aabb)
    d, n_p, n_g, aabb, ps, gs = _inspect(path)
    for f, _ in syn.ModelDesc._fields_:
        if f == "per_level_scale":
            continue   # derived at load time (nerf_render.cu:154-165) when absent from the file
        assert getattr(d, f) == getattr(sc.desc, f), f
    assert n_p == sc.params.size == capi.param_count(sc.desc) and n_g == sc.grid.size
    assert aabb == [float(x) for x in sc.aabb]
    assert ps == float(sc.params.astype(np.float64).sum()) and gs == float(sc.grid.astype(np.float64).sum())


def test_snapshot_reader_accepts_generic_msgpack_numbers(built_lib, tmp_path):
    """nlohmann's from_msgpack accepts any numeric encoding (float64, ints); so does the reader."""
    import msgpack
    H = 4
    top = {"encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 14,
                        "base_resolution": 16},
           "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64,
                       "n_hidden_layers": 1},
           "rgb_network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "Sigmoid",
                           "n_neurons": 64, "n_hidden_layers": 2},
           "dir_encoding": {"otype": "Composite", "nested": [{"otype": "SphericalHarmonics", "degree": 4}, {"otype": "Identity"}]},
           "loss": {"otype": "Huber"}, "unused": [1, 2, {"a": None}],
           "snapshot": {"aabb": [-1, -1, -1, 1, 1, 1], "bound": 1, "scale": 0.33, "cascade": 1, "density_grid_size": H,
                        "mean_density": 0.5, "density_grid": [0, 1] * (H ** 3 // 2), "params": [0.25, -1, 2.5]}}
    path = str(tmp_path / "generic.msgpack")
    open(path, "wb").write(msgpack.packb(top))   # float64 + small ints
    d, n_p, n_g, aabb, ps, gs = _inspect(path)
    assert (d.log2_hashmap_size, d.rgb_activation, d.sigma_activation, d.grid_size) == (14, syn.NRF_ACT_SIGMOID, syn.NRF_ACT_EXPONENTIAL, H)
    assert (n_p, n_g, ps, gs) == (3, H ** 3, 1.75, H ** 3 / 2)
    assert abs(d.scale - 0.33) < 1e-7 and d.mean_density == 0.5


def test_snapshot_reader_errors_match_reference(built_lib, tmp_path):
    import msgpack
    # wrong extension / missing file: nerf_render.cu:73-76,98,105-107
    with pytest.raises(NerfError) as e:
        _inspect(str(tmp_path / "model.bin"))
    assert e.value.code == -4 and ".msgpack" in str(e.value)
    with pytest.raises(NerfError) as e:
        _inspect(str(tmp_path / "missing.msgpack"))
    assert e.value.code == -4 and "does not exist" in str(e.value)
    # no "snapshot" key: nerf_render.cu:434-436
    p = str(tmp_path / "nosnap.msgpack")
    open(p, "wb").write(msgpack.packb({"encoding": {"otype": "HashGrid"}}))
    with pytest.raises(NerfError) as e:
        _inspect(p)
    assert e.value.code == -5 and "does not contain a snapshot" in str(e.value)
    # cascade / grid size mismatch: nerf_render.cu:467-469
    p = str(tmp_path / "badgrid.msgpack")
    open(p, "wb").write(msgpack.packb({"snapshot": {"aabb": [0] * 6, "cascade": 2, "density_grid_size": 2,
                                                   "density_grid": [0.0] * 8, "params": []}}))
    with pytest.raises(NerfError) as e:
        _inspect(p)
    assert "Incompatible number of grid cascades" in str(e.value)
    # truncated file
    p = str(tmp_path / "trunc.msgpack")
    sc = syn.make_scene("tiny")
    write_snapshot(p, sc.desc, sc.params[:100], sc.grid, sc.aabb)
    data = open(p, "rb").read()
    open(p, "wb").write(data[: len(data) // 2])
    with pytest.raises(NerfError) as e:
        _inspect(p)
    assert e.value.code == -5


def test_bench_reference_arm_contract_on_cpu():
    """bench.py --impl reference: under torchrun only rank 0 works (the other ranks exit 0 silently); without the reference's
    CUDA binary or a GPU the arm is the CPU port, and its JSON line carries the keys the driver reads."""
    import json
    import subprocess
    import sys
    bench = os.path.join(ROOT, "bench.py")
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, bench, "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0 and r.stdout.strip() == ""
    import torch
    if torch.cuda.is_available():
        return
    env["RANK"] = "0"
    env["LOCAL_RANK"] = "0"
    r = subprocess.run([sys.executable, bench, "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0", "--workload", "C1"],
                       capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stderr[-1000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "rays/s" and line["value"] > 0 and line["n_gpus"] == 2
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
