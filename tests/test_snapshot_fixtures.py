"""Snapshot / config files as the REFERENCE's library writes them (-m "not gpu": host-only parsing through the C ABI).

tests/golden/ref_nlohmann.msgpack was written by nlohmann::json::to_msgpack — the vendored library whose from_msgpack
decodes snapshots in NerfRender::load_network_config (/root/reference/src/nerf_render.cu:85-86) — from the DOM built in
oracle/ref_harness/make_snapshot.cpp (committed generator: tests/golden/make_snapshot_fixtures.sh).  It contains every
number encoding that library chooses on its own (float64, float32, fixints, uint16/32, int16, str8 keys, alphabetical
key order).  ref_child.json / ref_parent.json exercise load_network_config's .json branch: comments + the "parent"
chain merged with merge_patch (nerf_render.cu:30-44,79-82)."""
import ctypes as C
import os
import shutil

import numpy as np
import pytest

from nerf_cuda_b200 import capi
from nerf_cuda_b200.capi import NerfError
from nerf_cuda_b200.synthetic import ModelDesc

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
L, LOG2T, H, CAS = 16, 8, 32, 2
N_TABLE = L * (1 << LOG2T) * 2
N_PARAMS, N_GRID = 10240 + N_TABLE, CAS * H ** 3


def expected_arrays():
    """The closed formulas of make_snapshot.cpp, restated; conversion to float32 = data["params"].at(i) into a float
    (nerf_network.h:430-432)."""
    i = np.arange(N_GRID, dtype=np.int64)
    g = np.zeros(N_GRID, np.float64)
    m = i % 8
    g[m == 1] = 1
    g[m == 2] = 0.5
    g[m == 3] = 0.1 * (i[m == 3] % 13)
    g[(m == 4) & (i % 64 == 4)] = 300
    g[m == 5] = np.where(i[m == 5] % 128 == 5, 70000, 1)
    g[(m == 6) & (i % 256 == 6)] = -3
    g[m == 7] = np.where(i[m == 7] % 512 == 7, -200, 0.25)
    j = np.arange(N_PARAMS, dtype=np.uint64)
    v = ((j * np.uint64(2654435761)) % np.uint64(2001)).astype(np.float64)
    v = (v - 1000.0) / 4096.0
    v = np.where(j % np.uint64(5) == 0, v * 1.1, v)
    return v.astype(np.float32), g.astype(np.float32)


def inspect(path):
    lib = capi.lib()
    d, n_p, n_g = ModelDesc(), C.c_uint64(), C.c_uint64()
    aabb = (C.c_float * 6)()
    ps, gs = C.c_double(), C.c_double()
    rc = lib.nrf_snapshot_inspect(os.fsencode(path), C.byref(d), C.byref(n_p), C.byref(n_g), aabb, C.byref(ps), C.byref(gs))
    if rc:
        raise NerfError(rc, lib.nrf_last_error().decode())
    return d, n_p.value, n_g.value, np.array(aabb[:]), ps.value, gs.value


def test_fixture_contains_the_encodings_a_handwritten_reader_gets_wrong():
    raw = open(os.path.join(GOLD, "ref_nlohmann.msgpack"), "rb").read()
    import msgpack
    d = msgpack.unpackb(raw, raw=False, strict_map_key=False)
    assert list(d.keys()) == sorted(d.keys())                       # nlohmann objects are ordered maps
    assert isinstance(d["snapshot"]["bound"], int) and isinstance(d["snapshot"]["scale"], float)
    assert b"\xcb" in raw and b"\xca" in raw and b"\xcd\x01\x2c" in raw and b"\xce\x00\x01\x11\x70" in raw   # f64, f32, u16 300, u32 70000
    assert b"\xd1\xff\x38" in raw and b"\xfd" in raw                # int16 -200, negative fixint -3
    assert b"\xd9" in raw                                           # str8 key (> 31 bytes)


def test_reader_matches_the_reference_serialiser(built_lib, monkeypatch):
    monkeypatch.setenv("NRF_SNAPSHOT_CACHE", "0")
    d, n_p, n_g, aabb, ps, gs = inspect(os.path.join(GOLD, "ref_nlohmann.msgpack"))
    assert (d.n_levels, d.n_features_per_level, d.log2_hashmap_size, d.base_resolution) == (L, 2, LOG2T, 16)
    assert d.per_level_scale == 1.5 and d.n_neurons == 64 and d.density_hidden_layers == 1 and d.rgb_hidden_layers == 2
    assert d.sh_degree == 4 and d.rgb_activation == 3 and d.sigma_activation == 2
    assert d.bound == 2.0 and d.scale == np.float32(0.33) and d.cascades == CAS and d.grid_size == H
    assert d.mean_density == 0.0078125
    assert np.array_equal(aabb, [-2, -2, -2, 2, 2, 2])
    assert (n_p, n_g) == (N_PARAMS, N_GRID) and capi.param_count(d) == N_PARAMS
    p, g = expected_arrays()
    assert ps == float(p.astype(np.float64).sum()) and gs == float(g.astype(np.float64).sum())   # every value, exactly


def test_json_config_with_parent_chain_and_comments(built_lib):
    d, n_p, n_g, _, _, _ = inspect(os.path.join(GOLD, "ref_child.json"))
    assert (n_p, n_g) == (0, 0)
    assert d.log2_hashmap_size == LOG2T          # child overrides the parent's 19
    assert d.rgb_activation == 3                 # child: Sigmoid over the parent's None
    assert d.n_levels == L and d.base_resolution == 16 and d.per_level_scale == 1.5 and d.n_neurons == 64   # inherited
    dp = inspect(os.path.join(GOLD, "ref_parent.json"))[0]
    assert dp.log2_hashmap_size == 19 and dp.rgb_activation == 0
    with pytest.raises(NerfError) as e:
        inspect(os.path.join(GOLD, "no_such_config.json"))
    assert "does not exist" in str(e.value)


def test_json_parent_missing_and_n_features(built_lib, tmp_path):
    (tmp_path / "c.json").write_text('{"parent": "missing.json"}')
    with pytest.raises(NerfError) as e:
        inspect(str(tmp_path / "c.json"))
    assert "missing.json" in str(e.value) and "does not exist" in str(e.value)
    # nerf_render.cu:131-137: n_features / n_features_per_level replaces n_levels when present and > 0
    (tmp_path / "f.json").write_text('{"encoding": {"otype": "HashGrid", "n_features": 32, "n_features_per_level": 2, "n_levels": 3}}')
    assert inspect(str(tmp_path / "f.json"))[0].n_levels == 16
    # settings the renderer does not implement are rejected, not ignored
    (tmp_path / "a.json").write_text('{"network": {"output_activation": "ReLU"}}')
    with pytest.raises(NerfError) as e:
        inspect(str(tmp_path / "a.json"))
    assert "output_activation" in str(e.value)


def test_snapshot_without_aabb_is_an_error(built_lib, tmp_path, monkeypatch):
    import msgpack
    monkeypatch.setenv("NRF_SNAPSHOT_CACHE", "0")
    p = str(tmp_path / "noaabb.msgpack")
    open(p, "wb").write(msgpack.packb({"snapshot": {"cascade": 1, "density_grid_size": 2, "density_grid": [0.0] * 8, "params": [0.0]}}))
    with pytest.raises(NerfError) as e:
        inspect(p)
    assert "aabb" in str(e.value)


def test_binary_side_cache_is_used_and_stale_caches_are_rejected(built_lib, tmp_path, monkeypatch):
    monkeypatch.setenv("NRF_SNAPSHOT_CACHE", "1")
    src = str(tmp_path / "snap.msgpack")
    shutil.copy(os.path.join(GOLD, "ref_nlohmann.msgpack"), src)
    ref = inspect(src)
    cache = src + ".nrfcache"
    assert os.path.exists(cache) and os.path.getsize(cache) >= 4 * (N_PARAMS + N_GRID)
    # the cache really is what gets read: corrupt one float in it and the sums move
    raw = bytearray(open(cache, "rb").read())
    raw[-4:] = np.float32(1234.5).tobytes()
    open(cache, "wb").write(raw)
    got = inspect(src)
    assert got[5] != ref[5] and got[4] == ref[4]
    # a rewritten source (same size, other content and mtime) invalidates it: parsed again, cache refreshed
    data = bytearray(open(src, "rb").read())
    k = data.rindex(b"\xca")                       # last float32 of "params"
    data[k + 1:k + 5] = np.array(0.5, ">f4").tobytes()
    open(src, "wb").write(data)
    os.utime(src, ns=(os.stat(src).st_atime_ns, os.stat(src).st_mtime_ns + 5_000_000_000))
    fresh = inspect(src)
    assert fresh[5] == ref[5] and fresh[4] != ref[4]
    # truncated cache: rejected
    open(cache, "wb").write(open(cache, "rb").read()[:-100])
    assert inspect(src)[4] == fresh[4] and os.path.getsize(cache) >= 4 * (N_PARAMS + N_GRID)
    # disabled: no cache file is written
    monkeypatch.setenv("NRF_SNAPSHOT_CACHE", "0")
    src2 = str(tmp_path / "snap2.msgpack")
    shutil.copy(os.path.join(GOLD, "ref_nlohmann.msgpack"), src2)
    inspect(src2)
    assert not os.path.exists(src2 + ".nrfcache")
