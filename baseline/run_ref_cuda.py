#!/usr/bin/env python
"""Runs the REFERENCE's own CUDA renderer (baseline/_ref/ref_cuda_driver{,_d1}, built by baseline/build_ref_cuda.sh)
on a B200 and turns its outputs into (a) golden vectors that pin the network / frame of this repo and of the oracle
against an EXECUTION of the reference (tests/golden/ref_cuda_*.npz) and (b) the reference-CUDA baseline timings
(profiles/).  Runs on the GPU box only; needs neither /root/reference nor the oracle.

    python baseline/run_ref_cuda.py golden  [--out gpurun_out/ref_cuda]      # tiny + tiny16: net / frames / march dumps
    python baseline/run_ref_cuda.py time    [--workloads C2,C3] [--reps 3]    # ms/frame as shipped and D1-clean

Synthetic models are handed to the reference exactly like a trained export: as a .msgpack snapshot
(nerf_cuda_b200/snapshot.py) through NerfRender::reload_network_from_file.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from nerf_cuda_b200 import synthetic as syn  # noqa: E402
from nerf_cuda_b200.snapshot import write_snapshot  # noqa: E402

REF = os.path.join(ROOT, "baseline", "_ref")
NET_B = 4096
FRAME = 64


def have_ref() -> bool:
    return os.path.exists(os.path.join(REF, "ref_cuda_driver")) and os.path.exists(os.path.join(REF, "ref_cuda_driver_d1"))


def run_driver(exe, args, log_path=None):
    cmd = [os.path.join(REF, exe)] + [str(a) for a in args]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if log_path:
        with open(log_path, "w") as f:
            f.write(" ".join(cmd) + "\n" + r.stdout + "\n--- stderr ---\n" + r.stderr)
    if r.returncode != 0:
        raise RuntimeError(f"{exe} failed ({r.returncode}): {r.stderr[-2000:]}")
    events = []
    for line in r.stdout.splitlines():
        line = line.strip()
        if line.startswith("{"):
            try:
                events.append(json.loads(line))
            except json.JSONDecodeError:
                pass
    return events, r.stdout


def net_inputs(seed, B=NET_B):
    """B seeded samples: position uniform in [0,1]^3, direction uniform on the sphere, normalised input d*0.5+0.5
    (nerf_render.cu:311-314).  Returned as the [6 x B] SoA matrix the reference network takes."""
    rng = np.random.default_rng(seed)
    pos = rng.uniform(0.0, 1.0, size=(B, 3)).astype(np.float32)
    d = rng.normal(size=(B, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    dirs01 = (d * 0.5 + 0.5).astype(np.float32)
    return pos, dirs01, np.ascontiguousarray(np.concatenate([pos.T, dirs01.T], axis=0), np.float32)


def scene_variants():
    # (name, scene config, rgb activation, table range) — "stress" tables U(-1,1) and tcnn's init range U(-1e-4,1e-4)
    return [("tiny", "tiny", syn.NRF_ACT_NONE, 1.0), ("tiny16", "tiny16", syn.NRF_ACT_SIGMOID, 1.0),
            ("tiny_init", "tiny", syn.NRF_ACT_NONE, 1e-4)]


def golden(out_dir):
    os.makedirs(out_dir, exist_ok=True)
    summary = {}
    for name, cfg, act, trange in scene_variants():
        sc = syn.make_scene(cfg, rgb_activation=act, table_range=trange)
        with tempfile.TemporaryDirectory() as tmp:
            snap = os.path.join(tmp, f"{name}.msgpack")
            write_snapshot(snap, sc.desc, sc.params, sc.grid, sc.aabb)
            pos, dirs01, mat = net_inputs(20221 + len(name))
            mat.tofile(os.path.join(tmp, "net_in.bin"))
            rad = syn.camera_radius(cfg)
            poses = np.stack([syn.look_at_pose(rad, 20.0, az) for az in (30.0, 200.0)]).astype(np.float32)
            poses.tofile(os.path.join(tmp, "poses.bin"))
            poses[0].tofile(os.path.join(tmp, "pose0.bin"))
            intr = syn.intrinsics(FRAME, FRAME)
            cam = [float(x) for x in intr]
            ev, stdout = run_driver("ref_cuda_driver_d1",
                                    [snap, tmp, "net", os.path.join(tmp, "net_in.bin"), NET_B,
                                     "frames", FRAME, FRAME, *cam, os.path.join(tmp, "poses.bin"), 2, 1,
                                     "march", 32, 32, *[float(x) for x in syn.intrinsics(32, 32)], os.path.join(tmp, "pose0.bin"), 1024],
                                    log_path=os.path.join(out_dir, f"{name}_driver.log"))
            f16 = lambda fn, rows: np.fromfile(os.path.join(tmp, fn), np.uint16).reshape(rows, NET_B)  # noqa: E731
            net_ev = [e for e in ev if e.get("event") == "net"][0]
            out16 = f16("net_out.f16", net_ev["out_rows"])
            enc = np.fromfile(os.path.join(tmp, "net_enc.f16"), np.uint16)
            den = np.fromfile(os.path.join(tmp, "net_density.f16"), np.uint16)
            sh = np.fromfile(os.path.join(tmp, "net_sh.f16"), np.uint16)
            # encodings prefer AoS/CM ([B][rows] in memory), the networks SoA/RM: normalise everything to [B, rows]
            enc = enc.reshape(NET_B, 32) if net_ev["enc_layout"] == "CM" else enc.reshape(32, NET_B).T
            den = den.reshape(NET_B, 16) if net_ev["enc_layout"] == "CM" else den.reshape(16, NET_B).T
            sh = sh.reshape(NET_B, 16) if net_ev["sh_layout"] == "CM" else sh.reshape(16, NET_B).T
            frames_rgb = np.stack([np.fromfile(os.path.join(tmp, f"frame_{FRAME}x{FRAME}_{k}_rgb.f32"), np.float32).reshape(-1, 3) for k in range(2)])
            frames_dep = np.stack([np.fromfile(os.path.join(tmp, f"frame_{FRAME}x{FRAME}_{k}_depth.f32"), np.float32) for k in range(2)])
            n32 = 32 * 32
            mx = np.fromfile(os.path.join(tmp, "march_xyz.f32"), np.float32).reshape(n32, 1024, 3)
            md = np.fromfile(os.path.join(tmp, "march_deltas.f32"), np.float32).reshape(n32, 1024, 2)
            counts = (md[:, :, 0] != 0).sum(axis=1).astype(np.int32)
            keep = int(max(1, min(1024, counts.max())))
            np.savez_compressed(
                os.path.join(out_dir, f"ref_cuda_{name}.npz"),
                scene=cfg, rgb_activation=act, table_range=trange, seed=1337,
                net_pos=pos, net_dirs01=dirs01, net_out_f16=np.ascontiguousarray(out16[:4].T),   # [B,4] = r,g,b,sigma
                net_enc_f16=np.ascontiguousarray(enc), net_density_f16=np.ascontiguousarray(den), net_sh_f16=np.ascontiguousarray(sh),
                frame_size=FRAME, frame_intr=intr, frame_poses=poses, frame_rgb=frames_rgb, frame_depth=frames_dep,
                march_size=32, march_pose=poses[0], march_counts=counts,
                march_xyz=mx[:, :keep].astype(np.float32), march_deltas=md[:, :keep].astype(np.float32),
                march_nears=np.fromfile(os.path.join(tmp, "march_nears.f32"), np.float32),
                march_fars=np.fromfile(os.path.join(tmp, "march_fars.f32"), np.float32),
                march_rays_d=np.ascontiguousarray(np.fromfile(os.path.join(tmp, "march_rays_d.f32"), np.float32).reshape(3, n32).T),
                level_scales=np.fromfile(os.path.join(tmp, "level_scales.f32"), np.float32),   # device exp2f (MUFU.EX2)
            )
            probes = [ln for ln in stdout.splitlines() if "Inference" in ln]
            summary[name] = {"events": ev, "test_probe_lines": probes, "max_count": int(counts.max()), "samples": int(counts.sum())}
    with open(os.path.join(out_dir, "golden_summary.json"), "w") as f:
        json.dump(summary, f, indent=1)
    print(json.dumps({k: {"max_count": v["max_count"], "samples": v["samples"]} for k, v in summary.items()}))


WORKLOADS = {"C2": ("C2", 800, 800, 8), "C3": ("C3", 1920, 1080, 8), "C1": ("C1", 200, 200, 1)}


def time_workload(wl, reps, variants=("ref_cuda_driver", "ref_cuda_driver_d1"), snap_dir=None, n_poses=None):
    """ms/frame of NerfRender::render_frame (wall clock around the call, as src/main.cu:159-164 times it), per pose,
    first rep (buffer growth / first-touch) excluded when reps > 1."""
    cfg, W, H, K = WORKLOADS[wl]
    K = n_poses or K
    sc = syn.make_scene(cfg)
    tmp_ctx = tempfile.TemporaryDirectory() if snap_dir is None else None
    tmp = snap_dir or tmp_ctx.name
    snap = os.path.join(tmp, f"{wl}.msgpack")
    if not os.path.exists(snap):
        write_snapshot(snap, sc.desc, sc.params, sc.grid, sc.aabb)
    rad = syn.camera_radius(cfg)
    poses = np.stack([syn.look_at_pose(rad, 20.0, 360.0 * k / WORKLOADS[wl][3]) for k in range(K)]).astype(np.float32)
    pp = os.path.join(tmp, f"{wl}_poses.bin")
    poses.tofile(pp)
    cam = [float(x) for x in syn.intrinsics(W, H)]
    res = {}
    for exe in variants:
        ev, _ = run_driver(exe, [snap, tmp, "frames", W, H, *cam, pp, K, reps])
        fr = [e for e in ev if e.get("event") == "frame"]
        per_pose = {}
        for e in fr:
            per_pose.setdefault(e["pose"], []).append(e["ms"])
        steady = [min(v[1:]) if len(v) > 1 else v[0] for _, v in sorted(per_pose.items())]
        res[exe] = {"ms_per_pose_best": steady, "ms_per_frame_mean": float(np.mean(steady)),
                    "rays_per_s": W * H / (float(np.mean(steady)) * 1e-3),
                    "load_s": [e for e in ev if e.get("event") == "loaded"][0]["seconds"]}
    if tmp_ctx:
        tmp_ctx.cleanup()
    return {"workload": wl, "W": W, "H": H, "poses": K, "reps": reps, **res}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("mode", choices=["golden", "time"])
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "ref_cuda"))
    ap.add_argument("--workloads", default="C2,C3")
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    if not have_ref():
        raise SystemExit("baseline/_ref/ref_cuda_driver missing: run baseline/build_ref_cuda.sh in the dev container")
    os.makedirs(a.out, exist_ok=True)
    if a.mode == "golden":
        golden(a.out)
    else:
        out = [time_workload(w, a.reps) for w in a.workloads.split(",")]
        with open(os.path.join(a.out, "ref_cuda_timing.json"), "w") as f:
            json.dump(out, f, indent=1)
        print(json.dumps(out))


if __name__ == "__main__":
    main()
