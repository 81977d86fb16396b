"""nerf_cuda_b200 — B200-native NeRF inference hot path behind the reference's NerfRender interface.

Only what the path needs lives here: `csrc/` (sm_100a kernels + the C ABI), `capi.py` (ctypes binding + the Python
mirror of ngp::NerfRender), `synthetic.py` (synthetic snapshots/cameras), `snapshot.py` (snapshot writer for tests),
`build.py` (in-tree nvcc build).  Nothing in this package imports the CPU oracle.
"""
from .capi import NerfError, NerfRender, Stats, level_table, lib, param_count  # noqa: F401
from .synthetic import ModelDesc, make_scene  # noqa: F401
