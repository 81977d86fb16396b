"""Synthetic models, scenes and cameras of SURVEY.md §8(d) (numpy only; no oracle, no CUDA).

The reference loads weights and the occupancy grid only from a snapshot
(/root/reference/src/nerf_render.cu:431-473, include/nerf-cuda/nerf_network.h:424-443); there is no network
access for real checkpoints, so benches and tests use random-init weights in the reference's flat parameter
order (density MLP, rgb MLP, hash table — nerf_network.h:273-291) and analytic occupancy grids indexed like
the reference's float density grid (level*H^3 + nx*H^2 + ny*H + nz, include/nerf-cuda/render_utils.h:613-616).
"""
from __future__ import annotations

import ctypes
import math
from dataclasses import dataclass, field

import numpy as np

NRF_ACT_NONE, NRF_ACT_RELU, NRF_ACT_EXPONENTIAL, NRF_ACT_SIGMOID = 0, 1, 2, 3
MLP_PARAMS = 64 * 32 + 16 * 64 + 64 * 32 + 64 * 64 + 16 * 64  # 10240


class ModelDesc(ctypes.Structure):
    """Mirror of `nrf_model_desc` (include/nerf_b200.h)."""

    _fields_ = [
        ("n_levels", ctypes.c_uint32),
        ("n_features_per_level", ctypes.c_uint32),
        ("log2_hashmap_size", ctypes.c_uint32),
        ("base_resolution", ctypes.c_uint32),
        ("per_level_scale", ctypes.c_float),
        ("n_neurons", ctypes.c_uint32),
        ("density_hidden_layers", ctypes.c_uint32),
        ("rgb_hidden_layers", ctypes.c_uint32),
        ("sh_degree", ctypes.c_uint32),
        ("rgb_activation", ctypes.c_uint32),
        ("sigma_activation", ctypes.c_uint32),
        ("bound", ctypes.c_float),
        ("scale", ctypes.c_float),
        ("cascades", ctypes.c_uint32),
        ("grid_size", ctypes.c_uint32),
        ("mean_density", ctypes.c_float),
    ]


def make_desc(log2_hashmap_size=19, bound=1.0, cascades=1, rgb_activation=NRF_ACT_NONE, grid_size=128,
              base_resolution=16, per_level_scale=0.0, mean_density=1.0, scale=0.33) -> ModelDesc:
    """configs/nerf/base.json architecture (/root/reference/configs/nerf/base.json:23-58)."""
    return ModelDesc(16, 2, log2_hashmap_size, base_resolution, per_level_scale, 64, 1, 2, 4, rgb_activation,
                     NRF_ACT_EXPONENTIAL, bound, scale, cascades, grid_size, mean_density)


def level_table(desc: ModelDesc):
    """Python restatement of tcnn grid.h:899-931 / nerf_render.cu:147-165 in float32 (host logic, used for sizing
    synthetic parameter arrays; the C ABI's nrf_level_table is authoritative and tested against this)."""
    f32 = np.float32
    base = desc.base_resolution or (1 << (desc.log2_hashmap_size // 3))
    pls = f32(desc.per_level_scale)
    if pls <= 0 and desc.n_levels > 1:
        pls = np.exp(np.log(f32(2048.0) * f32(desc.bound) / f32(base)) / f32(desc.n_levels - 1), dtype=f32)
    log2_pls = np.log2(pls, dtype=f32)
    offsets, scales, res = [0], [], []
    for i in range(desc.n_levels):
        scale = f32(np.exp2(f32(i) * log2_pls, dtype=f32) * f32(base)) - f32(1.0)
        r = int(math.ceil(float(scale))) + 1
        n = r ** 3 if float(r) ** 3 <= float(0xFFFFFFFF // 2) else 0xFFFFFFFF // 2
        n = (n + 7) // 8 * 8
        n = min(n, 1 << desc.log2_hashmap_size)
        offsets.append(offsets[-1] + n)
        scales.append(float(scale))
        res.append(r)
    return np.array(offsets, np.uint32), np.array(scales, np.float32), np.array(res, np.uint32), float(pls)


def param_count(desc: ModelDesc) -> int:
    return MLP_PARAMS + 2 * int(level_table(desc)[0][-1])


def random_params(desc: ModelDesc, seed=1337, table_range=1.0) -> np.ndarray:
    """Flat float32 params in the reference's order W0[64x32] W1[16x64] V0[64x32] V1[64x64] V2[16x64] table.
    MLP: Xavier-uniform (tcnn gpu_matrix.h:292-307).  Table: U(-r, r); r=1e-4 is tcnn's init (grid.h:1280), the
    default r=1 is the "stress" set of SURVEY.md §8(d) so that the gather is visible after fp16 rounding."""
    rng = np.random.default_rng(seed)
    parts = []
    for n_out, n_in in ((64, 32), (16, 64), (64, 32), (64, 64), (16, 64)):
        s = math.sqrt(6.0 / (n_in + n_out))
        parts.append(rng.uniform(-s, s, size=n_out * n_in).astype(np.float32))
    n_table = 2 * int(level_table(desc)[0][-1])
    parts.append(rng.uniform(-table_range, table_range, size=n_table).astype(np.float32))
    return np.concatenate(parts)


def _cell_centres(H: int, half_extent: float):
    c = ((np.arange(H, dtype=np.float64) + 0.5) / H * 2.0 - 1.0) * half_extent
    return np.meshgrid(c, c, c, indexing="ij")  # index order nx, ny, nz


def occupancy_grid(desc: ModelDesc, scene="sphere") -> np.ndarray:
    """Float density grid [C*H^3], 1.0 where the cell centre satisfies the scene predicate."""
    H, C = desc.grid_size, desc.cascades
    out = np.zeros((C, H, H, H), np.float32)
    for c in range(C):
        mb = min(2.0 ** c, desc.bound)
        x, y, z = _cell_centres(H, mb)
        r = np.sqrt(x * x + y * y + z * z)
        occ = r < 0.6
        if scene == "unbounded":
            occ |= (r > 6.0) & (r < 7.0)
            occ |= (y > -0.1 * 1.0) & (y < 0.0) & (np.abs(x) < 12.0) & (np.abs(z) < 12.0)
            # the slab is thinner than a coarse cell: mark cells whose extent straddles it
            cell = 2.0 * mb / H
            occ |= (np.abs(y + 0.05) < 0.5 * cell + 0.05) & (np.abs(x) < 12.0) & (np.abs(z) < 12.0)
        elif scene != "sphere":
            raise ValueError(scene)
        out[c] = occ.astype(np.float32)
    return out.reshape(-1)


def aabb_of(desc: ModelDesc) -> np.ndarray:
    b = desc.bound
    return np.array([-b, -b, -b, b, b, b], np.float32)


def intrinsics(W: int, H: int, fov_deg=50.0) -> np.ndarray:
    fl = 0.5 * W / math.tan(math.radians(fov_deg) / 2.0)
    return np.array([fl, fl, W / 2.0, H / 2.0], np.float32)


def look_at_pose(radius=4.0, elevation_deg=20.0, azimuth_deg=0.0) -> np.ndarray:
    """Row-major 4x4 camera-to-world, NeRF/Blender convention (-z forward, +y up), looking at the origin."""
    el, az = math.radians(elevation_deg), math.radians(azimuth_deg)
    eye = np.array([radius * math.cos(el) * math.sin(az), radius * math.sin(el), radius * math.cos(el) * math.cos(az)])
    fwd = -eye / np.linalg.norm(eye)
    right = np.cross(fwd, np.array([0.0, 1.0, 0.0]))
    right /= np.linalg.norm(right)
    up = np.cross(right, fwd)
    P = np.eye(4)
    P[:3, 0], P[:3, 1], P[:3, 2], P[:3, 3] = right, up, -fwd, eye
    return P.astype(np.float32).reshape(-1)


@dataclass
class Scene:
    desc: ModelDesc
    params: np.ndarray
    grid: np.ndarray
    aabb: np.ndarray
    name: str = ""
    extra: dict = field(default_factory=dict)


def make_opaque(params: np.ndarray, gain=48.0) -> np.ndarray:
    """"Opaque" variant of a random-init model: the density row of the density network's output layer (W1 row 0, flat
    offset 2048..2112) is scaled so that raw densities are O(+-10) and sigma = exp(.) is either huge or ~0 — a ray
    terminates (T < 1e-4) at its first sample with a positive raw density, like a ray hitting a trained surface.
    Random-init weights alone give sigma ~ 1, where nothing ever terminates."""
    p = np.array(params, np.float32, copy=True)
    p[2048:2048 + 64] *= np.float32(gain)
    return p


def make_scene(config="C2", seed=1337, rgb_activation=NRF_ACT_NONE, table_range=1.0, opaque=False) -> Scene:
    """BASELINE.json configs: C1/C2 bound 1, one cascade, T=2^19; C3/C4 bound 16, 5 cascades; C5 T=2^24;
    'tiny' = T=2^14 for fast CPU tests.  opaque: see make_opaque."""
    if config in ("C1", "C2"):
        d = make_desc(19, 1.0, 1, rgb_activation)
        scene = "sphere"
    elif config in ("C3", "C4"):
        d = make_desc(19, 16.0, 5, rgb_activation)
        scene = "unbounded"
    elif config == "C5":
        d = make_desc(24, 1.0, 1, rgb_activation)
        scene = "sphere"
    elif config == "tiny":
        d = make_desc(14, 1.0, 1, rgb_activation)
        scene = "sphere"
    elif config == "tiny16":
        d = make_desc(14, 16.0, 5, rgb_activation)
        scene = "unbounded"
    elif config == "tiny3":   # bound is not 2^(cascades-1): the outer cascade is capped at `bound` (generic march step)
        d = make_desc(14, 3.0, 3, rgb_activation)
        scene = "unbounded"
    else:
        raise ValueError(config)
    params = random_params(d, seed, table_range)
    if opaque:
        params = make_opaque(params)
    return Scene(d, params, occupancy_grid(d, scene), aabb_of(d), config + ("-opaque" if opaque else ""))


def camera_radius(config: str) -> float:
    return 12.0 if config in ("C3", "C4", "tiny16") else (7.0 if config == "tiny3" else 4.0)
