#!/usr/bin/env python
"""bench.py — headline benchmark of the NeRF inference hot path (BASELINE.json: rays/s & ms/frame at 1080p).

A step = one frame = one pass of the hot path over W*H rays of the workload:
  default workload C3: 1920x1080, aabb_scale 16 (5 cascades), T=2^19, random-init weights, synthetic occupancy.
  N>1 (torchrun, one process per GPU): every rank renders its interleaved pixel blocks of the SAME frame
  (strong scaling); the u8 framebuffer is gathered to rank 0 over NCCL/NVLink.

JSON line (rank 0): value = rays/s with everything device-resident (the only host traffic inside a frame is the
8-byte sample-count read-back the algorithm itself needs); e2e = same metric through the reference-facing call
nrf_render_frame with host pose in / host u8 frame out inside the timed region.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (scene config, W, H, n_poses cycled)
    "C1": ("C1", 200, 200, 1),
    "C2": ("C2", 800, 800, 8),
    "C3": ("C3", 1920, 1080, 8),
    "C4": ("C4", 3840, 2160, 32),
    "C5": ("C5", 800, 800, 64),
}
GATHER_BYTES_PER_SAMPLE = 512      # 16 levels x 8 corners x 2 features x 2 B  
STREAM_BYTES_PER_SAMPLE = 16 + 8   # sample record read (float4 {u,v,w, ray id}) + {r,g,b,sigma} write
MLP_FLOP_PER_SAMPLE = 20480


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tf=d["bf16_tflops"], tf_sus=d["bf16_tflops_sustained"], src="measured")
    return dict(hbm=6650.0, tf=1590.0, tf_sus=1400.0, src="fallback")


def ncu_traffic(samples_per_launch):
    """DRAM traffic of the dominant kernel per launch, from the committed ncu --set full capture of one full-size launch
    (profiles/traffic_k_network.json, written by scripts/ncu_traffic.py): bytes per sample x this run's samples per launch."""
    p = os.path.join(ROOT, "profiles", "traffic_k_network.json")
    if not os.path.exists(p):
        return {"traffic": None}
    d = json.load(open(p))
    return {"traffic": d["dram_bytes_per_sample"] * samples_per_launch,
            "traffic_source": f"{d['dram_bytes_per_sample']:.1f} B/sample ({d['source']}, {d['workload']}) x samples per launch; "
                              f"L1 wavefront pipe {d['l1_wavefronts_pct_of_peak']:.0f} % of peak, tensor pipe {d['tensor_pipe_pct']:.1f} % in that capture"}


class ClockSampler:
    """nvidia-smi clocks during the timed region (B200_PROFILING.md clocks line)."""

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, f"/tmp/nrf_clocks_{os.getpid()}.csv"

    def start(self):
        q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.index)], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.proc.wait()
        sm, mx, reasons = [], [], set()
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.path)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def poses_for(workload):
    from nerf_cuda_b200 import synthetic as syn
    cfg, W, H, K = WORKLOADS[workload]
    rad = syn.camera_radius(cfg)
    return [syn.look_at_pose(rad, 20.0, 360.0 * k / K) for k in range(K)]


# --------------------------------------------------------------------------------------------- reference / CPU arm
def cpu_arm(workload, steps, warmup, target_s=4.0):
    """Times the oracle (CPU port of the reference's path, OpenMP over rays) on a bounded sample of the workload:
    a band of full pixel rows centred on the image, sized from a probe so one step is ~target_s seconds."""
    from nerf_cuda_b200 import synthetic as syn
    from oracle import orc
    cfg, W, H, _ = WORKLOADS[workload]
    sc = syn.make_scene(cfg)
    O = orc.Oracle(sc.desc, sc.params, sc.grid, sc.aabb)
    intr, poses = syn.intrinsics(W, H), poses_for(workload)
    mid = (H // 2) * W
    t0 = time.perf_counter()
    O.render(intr, poses[0], W, H, mid, W)
    probe = time.perf_counter() - t0
    rows = int(max(1, min(H // 2, target_s / max(probe, 1e-6))))
    pix0, npix = (H // 2 - rows // 2) * W, rows * W
    times, samples = [], 0
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        _, _, cnt = O.render(intr, poses[i % len(poses)], W, H, pix0, npix)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
            samples += int(cnt.sum())
    tot = sum(times)
    return dict(value=npix * len(times) / tot, ms_per_step=1e3 * tot / len(times), cores=O.num_threads(),
                sample=f"{rows} full pixel rows ({npix} rays) centred in the {W}x{H} frame per step, "
                       f"{samples / max(len(times), 1) / npix:.1f} samples/ray",
                samples_per_s=samples / tot)


# --------------------------------------------------------------------------------------------- GPU arm
class _DevPtr:
    """Zero-copy view of a device buffer owned by the C library, for torch (NCCL gather)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def gpu_arm(args):
    import torch
    import torch.distributed as dist
    from nerf_cuda_b200 import NerfRender, synthetic as syn

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cfg, W, H, _ = WORKLOADS[args.workload]
    sc = syn.make_scene(cfg)
    intr, poses = syn.intrinsics(W, H), poses_for(args.workload)
    r = NerfRender(1, device_ids=[local])
    r.load_raw(sc.desc, sc.params, sc.grid, sc.aabb)
    if world > 1:
        r.set_shard(rank, world)
    r.set_resolution((W, H))
    n_local = r.local_rays()
    ext = torch.cuda.ExternalStream(r.stream(), device=torch.device("cuda", local))
    rgb_ptr, dep_ptr, _ = r.device_outputs()
    gatherer = local_rgb = host_frame = fence = None
    peer = world > 1 and args.gather == "peer"
    if peer:
        # the gather is fused into the composite kernel: rank 0 exports its frame buffer (CUDA IPC), every other rank
        # maps it and stores its pixels there over NVLink; the only collective left is a 4-byte completion fence
        handle = torch.zeros(64, dtype=torch.uint8, device="cuda")
        if rank == 0:
            handle.copy_(torch.frombuffer(bytearray(r.peer_frame_export()), dtype=torch.uint8))
        dist.broadcast(handle, 0)
        if rank != 0:
            r.peer_frame_attach(bytes(handle.cpu().numpy().tobytes()))
        with torch.cuda.stream(ext):
            fence = torch.zeros(1, dtype=torch.int32, device="cuda")
        dist.barrier()
    elif world > 1:
        from nerf_cuda_b200.multigpu import FrameGather, FrameSharding
        sh = FrameSharding(W, H, world)
        with torch.cuda.stream(ext):
            gatherer = FrameGather(sh, rank, torch.device("cuda", local))
        local_rgb = torch.as_tensor(_DevPtr(rgb_ptr, n_local * 3), device="cuda").view(-1, 32, 32, 3)
        host_frame = torch.empty(H, W, 3, dtype=torch.uint8).pin_memory() if rank == 0 else None

    def frame(i, to_host):
        """One step.  to_host: reference-facing semantics (host pose in, host u8 frame out on rank 0)."""
        pose = poses[i % len(poses)]
        with torch.cuda.stream(ext):
            if world == 1:
                if to_host:
                    return r.render_frame(intr, pose)
                r.render_frame_device(intr, pose)
                return None
            r.render_frame_device(intr, pose)
            if peer:
                dist.all_reduce(fence)                      # completion fence (stream-ordered): every rank's stores landed
                if to_host and rank == 0:                   # the peer frame is double-buffered: the next frame's stores go
                    r.peer_frame_to_host()                  # to the other half, and ITS fence orders this copy before the
                    ext.synchronize()                       # half is written again — one collective per frame
                return None
            full = gatherer(local_rgb)                      # NCCL gather of packed u8 tiles + tile transpose on rank 0
            if to_host and rank == 0:
                host_frame.copy_(full, non_blocking=True)
                ext.synchronize()
        return None

    def timed(to_host, steps, warmup, collect_stats=False):
        for i in range(warmup):
            frame(i, to_host)
        r.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        net_ms, samples = [], []
        e0.record(ext)
        for i in range(steps):
            frame(warmup + i, to_host)
            if collect_stats:
                st = r.stats()
                net_ms.append(st.ms_network); samples.append(int(st.n_samples))
        e1.record(ext)
        e1.synchronize()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, net_ms, samples

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for i in range(len(poses)):          # size every buffer for the largest pose before anything is timed
        frame(i, False)
    r.synchronize()                      # sync point: the buffers grow to the largest frame enqueued so far
    ovf0 = r.overflowed_frames()
    ms_dev, _, _ = timed(False, args.steps, args.warmup)
    ms_e2e, net_ms, samples = timed(True, args.steps, max(3, args.warmup // 2), collect_stats=True)
    # frames are enqueued without host syncs: one that did not fit the sample buffers would have been rendered
    # incompletely (and too fast) — none may have occurred inside the timed regions
    ovf = r.overflowed_frames() - ovf0
    if ovf:
        raise RuntimeError(f"{ovf} timed frame(s) overflowed the sample buffers: the measurement is invalid")
    batch_rps = None
    if world == 1:   # the same host-in / host-out work through the pipelined batch call (render_server --batch)
        pb = np.stack([poses[(args.warmup + i) % len(poses)] for i in range(args.steps)])
        r.render_batch(intr, pb[: min(len(pb), len(poses))])
        tb = time.perf_counter()
        r.render_batch(intr, pb)
        batch_rps = W * H * len(pb) / (time.perf_counter() - tb)
    clocks = sampler.stop() if rank == 0 else None
    # stage rooflines on the last frame's resident samples (kernel timed alone -> burst peaks)
    stage = {}
    if rank == 0 or world == 1:
        stage = {k: r.bench_stage(w, 0, 3) for k, w in (("encode_ms", 0), ("mlp_ms", 1), ("fused_ms", 2))}
    # the samples the stage bench replayed: those of the frame rendered LAST (after the batch call that is its last pose)
    smp_local = int(r.stats().n_samples) if world == 1 else (samples[-1] if samples else 0)
    if world > 1:
        t = torch.tensor([float(np.mean(samples))], device="cuda", dtype=torch.float64)
        dist.all_reduce(t)
        smp_total = float(t.item())
    else:
        smp_total = float(np.mean(samples))
    rays = W * H
    pk = peaks()
    table_mb = 2 * int(syn.level_table(sc.desc)[0][-1]) * 2 / 1e6
    out = None
    if rank == 0:
        net = float(np.mean(net_ms))
        s_k = float(np.mean(samples))  # rank 0's samples per launch
        alg_bytes = s_k * (GATHER_BYTES_PER_SAMPLE + STREAM_BYTES_PER_SAMPLE)
        achieved = alg_bytes / (net * 1e-3) / 1e9
        out = {
            "metric": "rays_per_second_1080p" if args.workload == "C3" else f"rays_per_second_{args.workload}",
            "value": rays * args.steps / (ms_dev * 1e-3), "unit": "rays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f16", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {W}x{H} frame, bound {sc.desc.bound:g}, {sc.desc.cascades} cascades, "
                                   f"log2_hashmap_size {sc.desc.log2_hashmap_size}, base.json MLPs, random-init weights, "
                                   f"synthetic occupancy, {len(poses)} orbit poses cycled",
                       "sharding": "whole frame" if world == 1 else (
                           f"32x32-pixel tiles round-robin over {world} ranks; every rank's composite kernel stores its u8 pixels straight into "
                           f"rank 0's frame over NVLink (CUDA IPC peer stores), 4-byte NCCL all-reduce as completion fence" if peer else
                           f"32x32-pixel tiles round-robin over {world} ranks, NCCL gather of packed u8 tiles to rank 0"),
                       "samples_per_frame": smp_total, "samples_per_ray": smp_total / rays,
                       "l2": f"inputs larger than L2: each frame streams {smp_total / world * 56 / 1e6:.0f} MB of sample records per GPU (20 B written by the march, 16 read + 8 written by the network, 12 read by the composite) "
                             f"and the pose changes every step; the {table_mb:.1f} MB hash table is "
                             + ("L2-resident by design" if table_mb < 100 else "far larger than the 126 MB L2 (HBM-resident gather)")},
            "e2e": {"value": rays * args.steps / (ms_e2e * 1e-3), "unit": "rays/s", "ms_per_step": ms_e2e / args.steps,
                    "batch_call_value": batch_rps, "h2d_bytes_per_step": 80, "d2h_bytes_per_step": rays * 4 + 8 if (world == 1 or peer) else rays * 3 + 8},
            "gpu_launches": 5 * args.steps,
            "overflowed_timed_frames": int(ovf),
            "clocks": clocks,
            "roofline": {"kernel": "k_network<2> (hash-grid gather + tcgen05 MLP, fused)", "bound": "hbm",
                         "achieved": achieved, "peak": pk["hbm"], "unit": "GB/s", "frac": achieved / pk["hbm"],
                         **(ncu_traffic(s_k) if table_mb < 100 else {"traffic": None}), "peak_source": pk["src"] + (" (HBM copy; the table is L2-resident, so >1 is possible)" if table_mb < 100 else " (HBM copy)"),
                         "algorithmic_bytes_per_sample": GATHER_BYTES_PER_SAMPLE + STREAM_BYTES_PER_SAMPLE,
                         "samples_per_launch": s_k, "ms_per_launch": net,
                         "mlp_tflops_in_fused": s_k * MLP_FLOP_PER_SAMPLE / (net * 1e-3) / 1e12,
                         "mlp_frac_of_sustained_bf16_peak": s_k * MLP_FLOP_PER_SAMPLE / (net * 1e-3) / 1e12 / pk["tf_sus"]},
            "stages_alone": {
                **stage,
                "encode_gather_gbs": smp_local * GATHER_BYTES_PER_SAMPLE / (stage["encode_ms"] * 1e-3) / 1e9,
                "encode_frac_of_hbm_peak": smp_local * GATHER_BYTES_PER_SAMPLE / (stage["encode_ms"] * 1e-3) / 1e9 / pk["hbm"],
                "mlp_tflops": smp_local * MLP_FLOP_PER_SAMPLE / (stage["mlp_ms"] * 1e-3) / 1e12,
                "mlp_frac_of_burst_bf16_peak": smp_local * MLP_FLOP_PER_SAMPLE / (stage["mlp_ms"] * 1e-3) / 1e12 / pk["tf"],
            },
        }
        st = r.stats()
        out["frame_breakdown_ms"] = {"rays+count+scan": st.ms_setup, "readback+emit": st.ms_march, "network": st.ms_network,
                                     "composite": st.ms_composite, "total": st.ms_total}
    if world > 1:
        dist.barrier()
    if rank == 0 and world == 1 and not args.no_cpu:
        c = cpu_arm(args.workload, steps=2, warmup=1, target_s=4.0)
        out["cpu_baseline"] = {"value": c["value"], "unit": "rays/s", "cores": c["cores"], "kind": "port", "sample": c["sample"]}
    if rank == 0:
        print(json.dumps(out), flush=True)
    # teardown order matters: every torch tensor that was used on the library's stream must be released (and the
    # caching allocator's blocks for that stream returned) BEFORE nrf_destroy destroys the stream
    gatherer = local_rgb = host_frame = frame = timed = fence = None
    if peer:
        torch.cuda.synchronize()
        dist.barrier()
        if rank != 0:
            r.peer_frame_detach()
        dist.barrier()
    import gc
    gc.collect()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    torch.cuda.empty_cache()
    r.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C3", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--gather", default="nccl", choices=["peer", "nccl"],
                    help="N>1: nccl = gather of packed u8 tiles to rank 0 (default: measured 3 %% faster, the non-root ranks are not "
                         "held in lockstep); peer = composite kernels store into rank 0's frame over NVLink, 4-byte all-reduce fence")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:
            return
        cfg, W, H, _ = WORKLOADS[args.workload]
        steps, warm = min(args.steps, 8), min(args.warmup, 2)
        c = cpu_arm(args.workload, steps=steps, warmup=warm, target_s=6.0)
        print(json.dumps({
            "impl": "reference", "metric": "rays_per_second_1080p" if args.workload == "C3" else f"rays_per_second_{args.workload}",
            "value": c["value"], "unit": "rays/s", "n_gpus": args.gpus, "steps": steps, "warmup": warm,
            "ms_per_step": c["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f16",
            "data": "synthetic", "config": {"workload": f"{args.workload}: {W}x{H} frame (CPU arm renders a bounded row band per step)"},
            "cpu_baseline": {"value": c["value"], "unit": "rays/s", "cores": c["cores"], "kind": "port", "sample": c["sample"]},
            "e2e": {"value": c["value"], "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "CPU port of the reference's path (oracle/, OpenMP over rays); the reference's own CUDA build needs "
                    "cmake + CUDA-11 APIs and is recorded as unbuildable here (DESIGN.md)"}))
        return
    gpu_arm(args)


if __name__ == "__main__":
    main()
