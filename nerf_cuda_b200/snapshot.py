"""Writer for the reference's snapshot format (test / tooling side; the READER is C++, csrc/nrf_snapshot.h).

Schema (/root/reference/src/nerf_render.cu:431-473, include/nerf-cuda/nerf_network.h:424-431,
configs/nerf/base.json:23-58): msgpack map with the four network blocks of base.json and
"snapshot": {aabb[6], bound, scale, cascade, density_grid_size, mean_density, density_grid[C*H^3], params[...]}
where both big arrays are msgpack ARRAYS OF NUMBERS (not binary blobs).  The reference has no writer; this one exists
so that synthetic models can be fed through `reload_network_from_file` exactly like a trained export.
"""
from __future__ import annotations

import struct

import numpy as np

from .synthetic import NRF_ACT_EXPONENTIAL, NRF_ACT_NONE, NRF_ACT_RELU, NRF_ACT_SIGMOID, ModelDesc

_ACT = {NRF_ACT_NONE: "None", NRF_ACT_RELU: "ReLU", NRF_ACT_EXPONENTIAL: "Exponential", NRF_ACT_SIGMOID: "Sigmoid"}


def _f32_array(a: np.ndarray) -> bytes:
    """msgpack array32 of float32 elements (0xca + big-endian bits), vectorised."""
    a = np.ascontiguousarray(a, np.float32).reshape(-1)
    rec = np.empty(a.size, dtype=[("t", "u1"), ("v", ">f4")])
    rec["t"] = 0xCA
    rec["v"] = a
    return b"\xdd" + struct.pack(">I", a.size) + rec.tobytes()


def write_snapshot(path: str, desc: ModelDesc, params, grid, aabb) -> None:
    import msgpack
    pk = msgpack.Packer(use_single_float=True)
    top = {
        "encoding": {"otype": "HashGrid", "n_levels": desc.n_levels, "n_features_per_level": desc.n_features_per_level,
                     "log2_hashmap_size": desc.log2_hashmap_size, "base_resolution": desc.base_resolution},
        "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None",
                    "n_neurons": desc.n_neurons, "n_hidden_layers": desc.density_hidden_layers,
                    "sigma_activation": _ACT[desc.sigma_activation]},
        "dir_encoding": {"otype": "Composite", "nested": [{"n_dims_to_encode": 3, "otype": "SphericalHarmonics",
                                                            "degree": desc.sh_degree}, {"otype": "Identity"}]},
        "rgb_network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": _ACT[desc.rgb_activation],
                        "n_neurons": desc.n_neurons, "n_hidden_layers": desc.rgb_hidden_layers},
    }
    if desc.per_level_scale > 0:
        top["encoding"]["per_level_scale"] = float(desc.per_level_scale)
    with open(path, "wb") as f:
        f.write(pk.pack_map_header(len(top) + 1))
        for k, v in top.items():
            f.write(pk.pack(k))
            f.write(pk.pack(v))
        f.write(pk.pack("snapshot"))
        f.write(pk.pack_map_header(8))
        for k, v in (("aabb", [float(x) for x in np.asarray(aabb).reshape(6)]), ("bound", float(desc.bound)),
                     ("scale", float(desc.scale)), ("cascade", int(desc.cascades)),
                     ("density_grid_size", int(desc.grid_size)), ("mean_density", float(desc.mean_density))):
            f.write(pk.pack(k))
            f.write(pk.pack(v))
        f.write(pk.pack("density_grid"))
        f.write(_f32_array(grid))
        f.write(pk.pack("params"))
        f.write(_f32_array(params))
