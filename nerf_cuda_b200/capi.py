"""ctypes binding of the C ABI (include/nerf_b200.h, include/nerf_b200_ext.h) and a Python mirror of the reference's
`ngp::NerfRender` interface (/root/reference/include/nerf-cuda/nerf_render.h:29-50): same method names, argument
meaning and error behaviour (exceptions where the reference throws std::runtime_error).

There is no CPU path: loading fails loudly if the CUDA library is missing, and every compute call raises
`NerfError` when no sm_100 device is present.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .synthetic import ModelDesc

HERE = os.path.dirname(os.path.abspath(__file__))
# NRF_LIB: an experiment build of the same library (nerf_cuda_b200/build.py --variant), for A/B measurements only
LIB_PATH = os.environ.get("NRF_LIB") or os.path.join(HERE, "libnerf_b200.so")


class NerfError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"[nrf {code}] {msg}")
        self.code = code


class Stats(C.Structure):
    _fields_ = [("n_rays", C.c_uint64), ("n_samples", C.c_uint64), ("n_hit_rays", C.c_uint64),
                ("ms_total", C.c_float), ("ms_setup", C.c_float), ("ms_march", C.c_float),
                ("ms_network", C.c_float), ("ms_composite", C.c_float), ("n_launches", C.c_uint32),
                ("n_gpus", C.c_uint32), ("n_samples_marched", C.c_uint64), ("n_samples_used", C.c_uint64),
                ("n_rounds", C.c_uint32), ("reserved", C.c_uint32)]


EXPORTS = [
    "nrf_last_error", "nrf_version", "nrf_level_table", "nrf_param_count", "nrf_create", "nrf_destroy",
    "nrf_load_snapshot", "nrf_load_raw", "nrf_get_model_desc", "nrf_set_resolution", "nrf_render_frame",
    "nrf_render_batch", "nrf_render_frame_device", "nrf_get_float_frame", "nrf_get_stats", "nrf_synchronize",
    "nrf_probe_rays", "nrf_probe_march", "nrf_probe_encode", "nrf_probe_network", "nrf_probe_samples",
    "nrf_bench_stage", "nrf_set_shard", "nrf_local_rays", "nrf_device_outputs", "nrf_stream", "nrf_shard_pixel",
    "nrf_snapshot_inspect",
]

_lib = None


def lib():
    """Loads libnerf_b200.so; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: run `python -m nerf_cuda_b200.build` "
                              "(there is no CPU or PyTorch fallback for the hot path)")
        _lib = C.CDLL(LIB_PATH)
        _lib.nrf_last_error.restype = C.c_char_p
        _lib.nrf_version.restype = C.c_char_p
    return _lib


def _p(a, ty):
    return a.ctypes.data_as(C.POINTER(ty)) if a is not None else None


def _f32(a):
    return np.ascontiguousarray(a, np.float32)


def _check(rc):
    if rc != 0:
        raise NerfError(rc, lib().nrf_last_error().decode())


def level_table(desc: ModelDesc):
    L = desc.n_levels
    offs, sc, res = np.zeros(L + 1, np.uint32), np.zeros(L, np.float32), np.zeros(L, np.uint32)
    _check(lib().nrf_level_table(C.byref(desc), _p(offs, C.c_uint32), _p(sc, C.c_float), _p(res, C.c_uint32)))
    return offs, sc, res


def param_count(desc: ModelDesc) -> int:
    n = C.c_uint64()
    _check(lib().nrf_param_count(C.byref(desc), C.byref(n)))
    return n.value


class NerfRender:
    """Mirror of ngp::NerfRender.  `n_gpus` replaces the reference's compile-time NGPU (common.h:91)."""

    def __init__(self, n_gpus=1, device_ids=None):
        self._l = lib()
        self._ctx = C.c_void_p()
        ids = (C.c_int * n_gpus)(*device_ids) if device_ids is not None else None
        _check(self._l.nrf_create(n_gpus, ids, C.byref(self._ctx)))
        self.W = self.H = 0
        self._shard = (0, 1)   # (shard, n_shards)

    def close(self):
        if getattr(self, "_ctx", None) and self._ctx.value:
            self._l.nrf_destroy(self._ctx)
            self._ctx = C.c_void_p()

    __del__ = close

    # ---- reference interface --------------------------------------------------------------------------------------
    def reload_network_from_file(self, path: str):
        _check(self._l.nrf_load_snapshot(self._ctx, os.fsencode(path)))

    load_snapshot = reload_network_from_file

    def load_raw(self, desc: ModelDesc, params, grid, aabb):
        params, grid, aabb = _f32(params), _f32(grid), _f32(aabb)
        _check(self._l.nrf_load_raw(self._ctx, C.byref(desc), _p(params, C.c_float), C.c_size_t(params.size),
                                    _p(grid, C.c_float), C.c_size_t(grid.size), _p(aabb, C.c_float)))

    def set_resolution(self, resolution):
        W, H = int(resolution[0]), int(resolution[1])
        _check(self._l.nrf_set_resolution(self._ctx, W, H))
        self.W, self.H = W, H

    def render_frame(self, cam, pos):
        """cam = (fl_x, fl_y, cx, cy) (struct Camera); pos = 4x4 camera-to-world (NeRF convention).
        Returns (rgb u8 [H,W,3], depth u8 [H,W]) views of context-owned pinned memory (struct Image semantics),
        or this shard's packed tiles [n_tiles,32,32,3] / [n_tiles,32,32] for a sharded context."""
        intr, pose = _f32(cam).reshape(4), _f32(pos).reshape(16)
        rgb, dep = C.POINTER(C.c_uint8)(), C.POINTER(C.c_uint8)()
        _check(self._l.nrf_render_frame(self._ctx, _p(intr, C.c_float), _p(pose, C.c_float), C.byref(rgb), C.byref(dep)))
        if self._shard[1] == 1:
            return (np.ctypeslib.as_array(rgb, (self.H, self.W, 3)), np.ctypeslib.as_array(dep, (self.H, self.W)))
        t = self.local_rays() // 1024
        return np.ctypeslib.as_array(rgb, (t, 32, 32, 3)), np.ctypeslib.as_array(dep, (t, 32, 32))

    # ---- extensions -----------------------------------------------------------------------------------------------
    def render_frame_device(self, cam, pos):
        intr, pose = _f32(cam).reshape(4), _f32(pos).reshape(16)
        _check(self._l.nrf_render_frame_device(self._ctx, _p(intr, C.c_float), _p(pose, C.c_float)))

    def render_batch(self, cam, poses, want_depth=False):
        intr, poses = _f32(cam).reshape(4), _f32(poses).reshape(-1, 16)
        whole = self._shard[1] == 1
        n = self.W * self.H if whole else self.local_rays()
        out = np.zeros((poses.shape[0], n, 3), np.uint8)
        dep = np.zeros((poses.shape[0], n), np.uint8) if want_depth else None
        _check(self._l.nrf_render_batch_rgbd(self._ctx, _p(intr, C.c_float), _p(poses, C.c_float), poses.shape[0],
                                             _p(out, C.c_uint8), _p(dep, C.c_uint8)))
        out = out.reshape(poses.shape[0], self.H, self.W, 3) if whole else out.reshape(poses.shape[0], -1, 32, 32, 3)
        if not want_depth:
            return out
        return out, (dep.reshape(poses.shape[0], self.H, self.W) if whole else dep.reshape(poses.shape[0], -1, 32, 32))

    def float_frame(self):
        """Float image of the last frame + composited samples per ray: global pixel order [W*H] for a whole-frame
        context, packed tile order [n_local] for a sharded one."""
        n = self.W * self.H if self._shard[1] == 1 else self.local_rays()
        rgb, dep, cnt = np.zeros((n, 3), np.float32), np.zeros(n, np.float32), np.zeros(n, np.int32)
        _check(self._l.nrf_get_float_frame(self._ctx, _p(rgb, C.c_float), _p(dep, C.c_float), _p(cnt, C.c_int32)))
        return rgb, dep, cnt

    def synchronize(self):
        _check(self._l.nrf_synchronize(self._ctx))

    def stats(self) -> Stats:
        s = Stats()
        _check(self._l.nrf_get_stats(self._ctx, C.byref(s)))
        return s

    def set_shard(self, shard, n_shards):
        _check(self._l.nrf_set_shard(self._ctx, shard, n_shards))
        self._shard = (shard, n_shards)
        self.W = self.H = 0

    def local_rays(self) -> int:
        n = C.c_int()
        _check(self._l.nrf_local_rays(self._ctx, C.byref(n)))
        return n.value

    def shard_pixel(self, q: int) -> int:
        p = C.c_int()
        _check(self._l.nrf_shard_pixel(self._ctx, q, C.byref(p)))
        return p.value

    def device_outputs(self):
        a, b, c = C.c_void_p(), C.c_void_p(), C.c_void_p()
        _check(self._l.nrf_device_outputs(self._ctx, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def stream(self) -> int:
        s = C.c_void_p()
        _check(self._l.nrf_stream(self._ctx, C.byref(s)))
        return s.value or 0

    # ---- peer frame (nerf_b200_ext.h): the gather fused into the composite kernel, one process per GPU ---------
    def peer_frame_export(self) -> bytes:
        h = (C.c_ubyte * 64)()
        _check(self._l.nrf_peer_frame_export(self._ctx, h))
        return bytes(h)

    def peer_frame_attach(self, handle: bytes):
        h = (C.c_ubyte * 64).from_buffer_copy(handle)
        _check(self._l.nrf_peer_frame_attach(self._ctx, h))

    def peer_frame_detach(self):
        _check(self._l.nrf_peer_frame_detach(self._ctx))

    def peer_frame_device(self) -> int:
        p = C.c_void_p()
        _check(self._l.nrf_peer_frame_device(self._ctx, C.byref(p)))
        return p.value

    def peer_frame_to_host(self):
        """Owner: the latest frame on the host, complete when the call returns; (rgb[H,W,3], depth[H,W]) views."""
        return self.peer_frame_collect(0)

    def peer_frame_copy_out(self, on=True):
        _check(self._l.nrf_peer_frame_copy_out(self._ctx, int(bool(on))))

    def peer_frame_collect(self, back=0):
        """Owner: frame (latest - back) as (rgb[H,W,3], depth[H,W]) views of pinned memory; other ranks: waits for the rank's
        own latest frame and returns None."""
        rgb, dep = C.POINTER(C.c_ubyte)(), C.POINTER(C.c_ubyte)()
        _check(self._l.nrf_peer_frame_collect(self._ctx, int(back), C.byref(rgb), C.byref(dep)))
        if not rgb:
            return None
        W, H = self.W, self.H
        return (np.ctypeslib.as_array(rgb, shape=(H, W, 3)), np.ctypeslib.as_array(dep, shape=(H, W)))

    def frame_join(self):
        _check(self._l.nrf_frame_join(self._ctx))

    def bench_stage(self, what, n=0, iters=5) -> float:
        ms = C.c_float()
        _check(self._l.nrf_bench_stage(self._ctx, what, n, iters, C.byref(ms)))
        return ms.value

    def generate_density_grid(self, decay=0.95, reset=True, seed=0, jitter=False) -> float:
        """NerfRender::generate_density_grid (nerf_render.cu:388-429); returns the new mean density."""
        m = C.c_float()
        _check(self._l.nrf_generate_density_grid(self._ctx, C.c_float(decay), int(bool(reset)), C.c_ulonglong(seed),
                                                 int(bool(jitter)), C.byref(m)))
        return m.value

    def density_grid(self, n_cells: int):
        """(float grid [n_cells], occupancy bitfield [n_cells/32]) resident on device 0."""
        g, b = np.zeros(n_cells, np.float32), np.zeros(n_cells // 32, np.uint32)
        _check(self._l.nrf_get_density_grid(self._ctx, _p(g, C.c_float), _p(b, C.c_uint32)))
        return g, b

    def postprocess(self, accum, frame=None, sample_count=0.0, color_space=0, exposure=0.0, background=None,
                    output_color_space=1, tonemap_curve=0, clamp=True, want_output=True):
        """accumulate + tonemap (render_buffer.cu:224-258,529-555) on RGBA float arrays [n,4]; returns (accum, out)."""
        accum = np.ascontiguousarray(accum, np.float32).copy()
        n = accum.shape[0]
        fr = _f32(frame) if frame is not None else None
        bg = _f32(background if background is not None else [0, 0, 0, 0])
        out = np.zeros_like(accum) if want_output else None
        _check(self._l.nrf_postprocess(self._ctx, n, _p(fr, C.c_float), _p(accum, C.c_float), C.c_float(sample_count),
                                       color_space, C.c_float(exposure), _p(bg, C.c_float), output_color_space,
                                       tonemap_curve, int(bool(clamp)), _p(out, C.c_float)))
        return accum, out

    def level_scales(self):
        """Per-level grid scales as the device evaluates them (nerf_b200_ext.h nrf_get_level_scales): what the oracle's
        "cuda" numerics need, since MUFU.EX2 cannot be reproduced on the host."""
        sc = np.zeros(16, np.float32)
        _check(self._l.nrf_get_level_scales(self._ctx, _p(sc, C.c_float), None))
        return sc

    def set_termination(self, mode=2, bounds=None):
        """Termination-aware evaluation rounds (nerf_b200_ext.h nrf_set_termination): mode 0 whole rays, 1 rounds,
        2 automatic; bounds = cumulative per-ray sample boundaries of all rounds but the last (default 8, 32, 128)."""
        b = (C.c_int * len(bounds))(*bounds) if bounds else None
        _check(self._l.nrf_set_termination(self._ctx, int(mode), b, len(bounds) if bounds else 0))

    def overflowed_frames(self) -> int:
        """Asynchronous frames that did not fit the sample buffers so far (nerf_b200_ext.h nrf_overflowed_frames)."""
        n = C.c_uint64(0)
        _check(self._l.nrf_overflowed_frames(self._ctx, C.byref(n)))
        return int(n.value)

    def bench_memory(self, mode, window_bytes, iters=3):
        """Memory micro-benchmark (nerf_b200_ext.h nrf_bench_memory): returns (GB/s requested, loads per ns)."""
        g, l = C.c_float(), C.c_float()
        _check(self._l.nrf_bench_memory(self._ctx, mode, C.c_size_t(window_bytes), iters, C.byref(g), C.byref(l)))
        return g.value, l.value

    # ---- stage probes ---------------------------------------------------------------------------------------------
    def probe_rays(self, cam, pos, W, H, pix0=0, n=None):
        n = W * H - pix0 if n is None else n
        intr, pose = _f32(cam).reshape(4), _f32(pos).reshape(16)
        o, d = np.zeros(3, np.float32), np.zeros((n, 3), np.float32)
        nr, fr = np.zeros(n, np.float32), np.zeros(n, np.float32)
        _check(self._l.nrf_probe_rays(self._ctx, _p(intr, C.c_float), _p(pose, C.c_float), W, H, pix0, n,
                                      _p(o, C.c_float), _p(d, C.c_float), _p(nr, C.c_float), _p(fr, C.c_float)))
        return o, d, nr, fr

    def probe_march(self, o, d, nears, fars, want_samples=True):
        o, d, nears, fars = _f32(o), _f32(d), _f32(nears), _f32(fars)
        n = d.shape[0]
        counts = np.zeros(n, np.int32)
        _check(self._l.nrf_probe_march(self._ctx, n, _p(o, C.c_float), _p(d, C.c_float), _p(nears, C.c_float),
                                       _p(fars, C.c_float), _p(counts, C.c_int32), None, C.c_size_t(0)))
        if not want_samples:
            return counts, None
        total = int(counts.sum())
        samples = np.zeros((max(total, 1), 8), np.float32)
        _check(self._l.nrf_probe_march(self._ctx, n, _p(o, C.c_float), _p(d, C.c_float), _p(nears, C.c_float),
                                       _p(fars, C.c_float), _p(counts, C.c_int32), _p(samples, C.c_float),
                                       C.c_size_t(samples.shape[0])))
        return counts, samples[:total]

    def probe_encode(self, positions, want_indices=False):
        positions = _f32(positions)
        n = positions.shape[0]
        enc = np.zeros((n, 32), np.uint16)
        idx = np.zeros((n, 16, 8), np.uint32) if want_indices else None
        _check(self._l.nrf_probe_encode(self._ctx, n, _p(positions, C.c_float), _p(enc, C.c_uint16), _p(idx, C.c_uint32)))
        return (enc, idx) if want_indices else enc

    def probe_sh(self, dirs01):
        dirs01 = _f32(dirs01)
        n = dirs01.shape[0]
        out = np.zeros((n, 16), np.uint16)
        _check(self._l.nrf_probe_sh(self._ctx, n, _p(dirs01, C.c_float), _p(out, C.c_uint16)))
        return out

    def probe_network(self, enc_u16, dirs01, impl=0):
        enc_u16, dirs01 = np.ascontiguousarray(enc_u16, np.uint16), _f32(dirs01)
        n = enc_u16.shape[0]
        out = np.zeros((n, 4), np.float32)
        _check(self._l.nrf_probe_network(self._ctx, n, _p(enc_u16, C.c_uint16), _p(dirs01, C.c_float), _p(out, C.c_float), impl))
        return out

    def eval_samples_noalloc(self, n, positions_dev, dirs_dev, out_dev):
        """nrf_eval_samples_noalloc: DEVICE pointers (ints): n x float4 positions, n x float4 directions (both normalised to
        [0,1]), n x 4 fp16 out {r,g,b,sigma}.  Allocates nothing; asynchronous on the context's stream (synchronize() after)."""
        _check(self._l.nrf_eval_samples_noalloc(self._ctx, int(n), C.c_void_p(int(positions_dev)), C.c_void_p(int(dirs_dev)),
                                                C.c_void_p(int(out_dev))))

    def probe_samples(self, positions01, dirs01):
        positions01, dirs01 = _f32(positions01), _f32(dirs01)
        n = positions01.shape[0]
        out = np.zeros((n, 4), np.float32)
        _check(self._l.nrf_probe_samples(self._ctx, n, _p(positions01, C.c_float), _p(dirs01, C.c_float), _p(out, C.c_float)))
        return out
