"""ctypes loader for the CPU oracle (oracle/libnerf_oracle.so) and, in the dev container, the host build of the
reference's own kernels (oracle/_ref/libnerf_ref.so).

TEST INFRASTRUCTURE: importable only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  The product package never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_u16p = np.ctypeslib.ndpointer(np.uint16, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")


def build(ref=True):
    subprocess.run(["make", "-s", "-C", HERE, "libnerf_oracle.so"], check=True)
    if ref and os.path.isdir("/root/reference/include/nerf-cuda"):
        subprocess.run([os.path.join(HERE, "build_ref.sh")], check=True, stdout=subprocess.DEVNULL)


_lib = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(HERE, "libnerf_oracle.so")
        if not os.path.exists(path):
            build(ref=False)
        _lib = C.CDLL(path)
        _lib.orc_march.restype = C.c_longlong
        _lib.orc_render_frame.restype = C.c_longlong
    return _lib


def _p(a, ty):
    return a.ctypes.data_as(C.POINTER(ty)) if a is not None else None


class Oracle:
    """Holds a model (fp16 weights + table, float grid) and exposes each stage of the hot path.

    numerics: "host" = the reference's expression trees as written (what its kernels give compiled for the host; pinned by
    oracle/_ref); "cuda" = as nvcc compiles them for the GPU — a*b+c contracted where the reference's sm_100 SASS has an
    FFMA — with `scales` = the per-level grid scales as the device's exp2f gives them (NerfRender.level_scales(), or the
    `level_scales` array of tests/golden/ref_cuda_*.npz); pinned by an execution of the reference's CUDA build."""

    def __init__(self, desc, params_f32, grid, aabb, numerics="host", scales=None):
        self.l = lib()
        assert numerics in ("host", "cuda")
        self.fma = numerics == "cuda"
        self.scales = None if scales is None else np.ascontiguousarray(scales, np.float32)
        self.desc = desc
        self.grid = np.ascontiguousarray(grid, np.float32)
        self.aabb = np.ascontiguousarray(aabb, np.float32)
        p16 = np.empty(params_f32.size, np.uint16)
        self.l.orc_cast_f16(_p(np.ascontiguousarray(params_f32, np.float32), C.c_float), _p(p16, C.c_uint16),
                            C.c_size_t(params_f32.size))
        self.mlp = np.ascontiguousarray(p16[:10240])
        self.table = np.ascontiguousarray(p16[10240:])

    def _mode(self):
        n = 0 if self.scales is None else int(self.scales.size)
        self.l.orc_set_numerics(int(self.fma), _p(self.scales, C.c_float), n)

    def level_table(self):
        self._mode()
        L = self.desc.n_levels
        offs, sc, res = np.zeros(L + 1, np.uint32), np.zeros(L, np.float32), np.zeros(L, np.uint32)
        pls = C.c_float()
        rc = self.l.orc_level_table(C.byref(self.desc), _p(offs, C.c_uint32), _p(sc, C.c_float), _p(res, C.c_uint32),
                                    C.byref(pls))
        assert rc == 0
        return offs, sc, res, pls.value

    def rays(self, intr, pose, W, H, pix0=0, n=None):
        self._mode()
        n = W * H - pix0 if n is None else n
        o, d = np.zeros(3, np.float32), np.zeros((n, 3), np.float32)
        nr, fr = np.zeros(n, np.float32), np.zeros(n, np.float32)
        self.l.orc_rays(C.byref(self.desc), _p(self.aabb, C.c_float), _p(intr, C.c_float), _p(pose, C.c_float), W, H,
                        pix0, n, _p(o, C.c_float), _p(d, C.c_float), _p(nr, C.c_float), _p(fr, C.c_float))
        return o, d, nr, fr

    def march(self, o, d, nears, fars, want_samples=True):
        self._mode()
        n = d.shape[0]
        counts = np.zeros(n, np.int32)
        args = (C.byref(self.desc), _p(self.grid, C.c_float), n, _p(o, C.c_float), _p(d, C.c_float),
                _p(nears, C.c_float), _p(fars, C.c_float), _p(counts, C.c_int32))
        total = self.l.orc_march(*args, None, C.c_size_t(0))
        if not want_samples:
            return counts, None
        samples = np.zeros((max(total, 1), 8), np.float32)
        got = self.l.orc_march(*args, _p(samples, C.c_float), C.c_size_t(samples.shape[0]))
        assert got == total
        return counts, samples[:total]

    def encode(self, positions, want_indices=False):
        self._mode()
        n = positions.shape[0]
        enc = np.zeros((n, 32), np.uint16)
        idx = np.zeros((n, self.desc.n_levels, 8), np.uint32) if want_indices else None
        rc = self.l.orc_encode(C.byref(self.desc), _p(self.table, C.c_uint16), n,
                               _p(np.ascontiguousarray(positions, np.float32), C.c_float), _p(enc, C.c_uint16),
                               _p(idx, C.c_uint32))
        assert rc == 0
        return (enc, idx) if want_indices else enc

    def sh(self, dirs01):
        self._mode()
        n = dirs01.shape[0]
        out = np.zeros((n, 16), np.uint16)
        self.l.orc_sh(n, _p(np.ascontiguousarray(dirs01, np.float32), C.c_float), _p(out, C.c_uint16))
        return out

    def network(self, enc_u16, dirs01):
        self._mode()
        n = enc_u16.shape[0]
        out = np.zeros((n, 4), np.float32)
        rc = self.l.orc_network(C.byref(self.desc), _p(self.mlp, C.c_uint16), n,
                                _p(np.ascontiguousarray(enc_u16, np.uint16), C.c_uint16),
                                _p(np.ascontiguousarray(dirs01, np.float32), C.c_float), _p(out, C.c_float))
        assert rc == 0
        return out

    def samples(self, positions01, dirs01):
        self._mode()
        n = positions01.shape[0]
        out = np.zeros((n, 4), np.float32)
        rc = self.l.orc_samples(C.byref(self.desc), _p(self.mlp, C.c_uint16), _p(self.table, C.c_uint16), n,
                                _p(np.ascontiguousarray(positions01, np.float32), C.c_float),
                                _p(np.ascontiguousarray(dirs01, np.float32), C.c_float), _p(out, C.c_float))
        assert rc == 0
        return out

    def density_grid(self, decay=0.95, reset=True, seed=0, jitter=False, grid=None):
        """generate_density_grid restatement: returns (new float grid [C*H^3], mean density)."""
        self._mode()
        g = np.ascontiguousarray(self.grid if grid is None else grid, np.float32).copy()
        self.l.orc_density_grid.restype = C.c_float
        mean = self.l.orc_density_grid(C.byref(self.desc), _p(self.mlp, C.c_uint16), _p(self.table, C.c_uint16),
                                       C.c_float(decay), int(bool(reset)), C.c_uint64(seed), int(bool(jitter)),
                                       _p(g, C.c_float))
        assert mean >= 0
        return g, float(mean)

    def render(self, intr, pose, W, H, pix0=0, npix=None):
        self._mode()
        npix = W * H - pix0 if npix is None else npix
        rgb, depth = np.zeros((npix, 3), np.float32), np.zeros(npix, np.float32)
        counts = np.zeros(npix, np.int32)
        total = self.l.orc_render_frame(C.byref(self.desc), _p(self.mlp, C.c_uint16), _p(self.table, C.c_uint16),
                                        _p(self.grid, C.c_float), _p(self.aabb, C.c_float), _p(intr, C.c_float),
                                        _p(pose, C.c_float), W, H, pix0, npix, _p(rgb, C.c_float),
                                        _p(depth, C.c_float), _p(counts, C.c_int32))
        assert total >= 0
        return rgb, depth, counts

    def to_u8(self, img):
        a = np.ascontiguousarray(img, np.float32)
        out = np.zeros(a.shape, np.uint8)
        self.l.orc_to_u8(_p(a, C.c_float), _p(out, C.c_uint8), C.c_size_t(a.size))
        return out

    def num_threads(self):
        return self.l.orc_num_threads()

    def set_num_threads(self, n):
        """OpenMP threads of every later call (overrides an inherited OMP_NUM_THREADS); returns the count in effect."""
        return self.l.orc_set_num_threads(int(n))


def cuda_oracle(renderer, desc, params_f32, grid, aabb):
    """Oracle in the reference-CUDA numerics for comparison with a renderer: the per-level grid scales are the ones the
    renderer's device evaluated (the same MUFU.EX2 the reference's kernel_grid uses); `renderer` must hold a model with
    the same encoding parameters."""
    return Oracle(desc, params_f32, grid, aabb, numerics="cuda", scales=renderer.level_scales())


def half_to_float(u16):
    return np.ascontiguousarray(u16, np.uint16).view(np.float16).astype(np.float32)


# ---- host build of the reference's own kernels (dev container only) -------------------------------------------------
_ref = None


def ref_available():
    return os.path.exists(os.path.join(HERE, "_ref", "libnerf_ref.so"))


def ref():
    global _ref
    if _ref is None:
        _ref = C.CDLL(os.path.join(HERE, "_ref", "libnerf_ref.so"))
    return _ref
