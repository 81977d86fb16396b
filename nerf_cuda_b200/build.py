"""In-tree build of the CUDA hot path: nvcc -> nerf_cuda_b200/libnerf_b200.so (sm_100a only, no torch dependency)."""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libnerf_b200.so")
SOURCES = ["nrf_api.cu", "kernels_exact.cu", "kernels_network.cu", "kernels_composite.cu", "kernels_probe.cu", "kernels_post.cu"]
HEADERS = ["nrf_device.cuh", "nrf_snapshot.h", "../../include/nerf_b200.h", "../../include/nerf_b200_ext.h",
           "../../include/nerf-cuda/nerf_render.h", "../host/render_server.cpp", "../host/testbed.cpp"]
# host executables (C++ over the C ABI): the reference's render_server / testbed counterparts
HOST_EXES = {"render_server": "host/render_server.cpp", "testbed": "host/testbed.cpp"}
# -fmad=false: no contraction by the compiler; the FMAs the reference's nvcc build contracts are spelled __fmaf_rn in the
# sources, so the march / encode / SH arithmetic is bit-exact against a run of the reference's CUDA build.
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-fmad=false",
              "-Xcompiler", "-fPIC,-ffp-contract=off", "-Xptxas", "-v"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def needs_build() -> bool:
    if not os.path.exists(LIB) or not all(os.path.exists(os.path.join(HERE, e)) for e in HOST_EXES):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def declared_symbols():
    """Every function include/nerf_b200.h and include/nerf_b200_ext.h declare: the library's whole export list."""
    import re
    names = []
    for h in ("nerf_b200.h", "nerf_b200_ext.h"):
        src = open(os.path.join(HERE, "..", "include", h)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names += re.findall(r"\b(nrf_[a-z0-9_]+)\s*\(", src)
    return sorted(set(names))


def export_map() -> str:
    """Linker version script: exactly the declared C ABI is visible; kernels' host stubs, cross-TU launch helpers and
    experiment hooks stay local to the library."""
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    path = os.path.join(HERE, "build", "exports.map")
    with open(path, "w") as f:
        f.write("{\n  global:\n" + "".join(f"    {n};\n" for n in declared_symbols()) + "  local: *;\n};\n")
    return path


# Experiment builds (A/B evidence for DESIGN.md §4, never the product): name -> extra -D flags for the kernel TUs.
VARIANTS = {
    "floor_magic": ["-DNRF_EXP_FLOOR_MAGIC=1"],        # floor/F2I (XU pipe) -> FADD.RZ magic constant (FMA pipe)
    "evict": ["-DNRF_EXP_EVICT=1"],                    # table loads L2::evict_last, sample stream loads/stores evict_first
    "smem_l01": ["-DNRF_EXP_SMEM_L01=1"],              # dense levels 0-1 staged in shared memory (LDS instead of LDG)
    "floor_magic_evict": ["-DNRF_EXP_FLOOR_MAGIC=1", "-DNRF_EXP_EVICT=1"],
    "march_magic": ["-DNRF_EXP_MARCH_MAGIC=1"],
    "no_tail_sync": ["-DNRF_NO_TAIL_SYNC=1"],        # k_network: drop the group barrier at the end of a tile
    "arrive_sync": ["-DNRF_ARRIVE_SYNC=1"],          # k_network: group barriers as mbarrier arrive + wait by the issuing thread
    "tcache128": ["-DNRF_TCACHE=128"],                # 128 instead of 64 cached sample t's per ray (1 GB at 1080p)
    "tcache32": ["-DNRF_TCACHE=32"],      # march step: F2I.TRUNC x3 (XU pipe) -> FADD.RZ + mantissa mask
}


def build_variant(name: str) -> str:
    """libnerf_b200_<name>.so next to the product library (load it with NRF_LIB=<path>)."""
    out = os.path.join(HERE, f"libnerf_b200_{name}.so")
    objs = []
    os.makedirs(os.path.join(HERE, "build", name), exist_ok=True)
    for s in SOURCES:
        o = os.path.join(HERE, "build", name, s.replace(".cu", ".o"))
        r = subprocess.run([_nvcc(), *NVCC_FLAGS, *VARIANTS[name], "-c", os.path.join(CSRC, s), "-o", o], capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError(f"nvcc failed on {s} ({name})")
        objs.append(o)
        with open(os.path.join(HERE, "build", name, s.replace(".cu", ".ptxas.log")), "w") as f:
            f.write(r.stderr)
    r = subprocess.run([_nvcc(), "-shared", "-o", out, *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-lcudart",
                        "-Xlinker", f"--version-script={export_map()}"], capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("nvcc link failed")
    return out


def build(force=False, verbose=False) -> str:
    if not force and not needs_build():
        return LIB
    objs = []
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    log = []
    for s in SOURCES:
        o = os.path.join(HERE, "build", s.replace(".cu", ".o"))
        cmd = [_nvcc(), *NVCC_FLAGS, "-c", os.path.join(CSRC, s), "-o", o]
        r = subprocess.run(cmd, capture_output=True, text=True)
        log.append(r.stderr)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError(f"nvcc failed on {s}")
        objs.append(o)
    cmd = [_nvcc(), "-shared", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-lcudart",
           "-Xlinker", f"--version-script={export_map()}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("nvcc link failed")
    for exe, src in HOST_EXES.items():
        cmd = ["g++", "-std=c++17", "-O2", "-Wall", os.path.join(HERE, src), "-o", os.path.join(HERE, exe), "-L", HERE,
               "-lnerf_b200", "-Wl,-rpath,$ORIGIN"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError(f"g++ failed on {src}")
    with open(os.path.join(HERE, "build", "ptxas.log"), "w") as f:
        f.write("\n".join(log))
    if verbose:
        print("\n".join(log))
    return LIB


if __name__ == "__main__":
    if "--variant" in sys.argv:
        for v in sys.argv[sys.argv.index("--variant") + 1].split(","):
            print(build_variant(v))
    else:
        print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
