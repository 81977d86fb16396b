#!/usr/bin/env python
"""bench.py — headline benchmark of the NeRF inference hot path (BASELINE.json: rays/s & ms/frame at 1080p).

A step = one frame = one pass of the hot path over W*H rays of the workload:
  default workload C3: 1920x1080, aabb_scale 16 (5 cascades), T=2^19, random-init weights, synthetic occupancy.
  N>1 (torchrun, one process per GPU): every rank renders its interleaved pixel blocks of the SAME frame
  (strong scaling); every rank's composite kernel stores its u8 pixels into rank 0's frame over NVLink (CUDA IPC peer stores,
  one-sided sequence-flag fence, no collective; --gather nccl = round 1's gather of packed tiles).

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


def ncu_capture():
    """Figures of the committed ncu --set full capture of one full-size k_network launch (profiles/traffic_k_network.json,
    written by scripts/ncu_traffic.py): DRAM bytes per sample, L1 wavefront pipe and tensor pipe utilisation."""
    p = os.path.join(ROOT, "profiles", "traffic_k_network.json")
    return json.load(open(p)) if os.path.exists(p) else None


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
    O.set_num_threads(os.cpu_count() or 1)     # torchrun exports OMP_NUM_THREADS=1 to its workers: use every host core anyway
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


def ref_cuda_arm(workload, steps, warmup, exe="ref_cuda_driver_d1"):
    """Times the reference's OWN CUDA renderer (baseline/_ref/<exe>: src/nerf_render.cu + tiny-cuda-nn compiled for sm_100
    by baseline/build_ref_cuda.sh) on this box's GPU 0 through its public call NerfRender::render_frame — host pose in,
    host u8 frame out, wall clock around the calls as src/main.cu:159-164 times it.  The model reaches it the way a trained
    export does: as a .msgpack snapshot through reload_network_from_file.  Returns None when the binary or a GPU is absent."""
    import tempfile
    from nerf_cuda_b200 import synthetic as syn
    from nerf_cuda_b200.snapshot import write_snapshot
    path = os.path.join(ROOT, "baseline", "_ref", exe)
    if not os.path.exists(path):
        return None
    try:
        if subprocess.run(["nvidia-smi", "-L"], capture_output=True, text=True).stdout.count("GPU ") < 1:
            return None
    except OSError:
        return None
    cfg, W, H, _ = WORKLOADS[workload]
    sc = syn.make_scene(cfg)
    poses = np.stack(poses_for(workload)).astype(np.float32)
    with tempfile.TemporaryDirectory() as tmp:
        snap, pp = os.path.join(tmp, "model.msgpack"), os.path.join(tmp, "poses.bin")
        write_snapshot(snap, sc.desc, sc.params, sc.grid, sc.aabb)
        poses.tofile(pp)
        cam = [repr(float(x)) for x in syn.intrinsics(W, H)]
        env = dict(os.environ, CUDA_VISIBLE_DEVICES=os.environ.get("CUDA_VISIBLE_DEVICES", "0").split(",")[0])
        r = subprocess.run([path, snap, tmp, "bench", str(W), str(H), *cam, pp, str(len(poses)), str(warmup), str(steps)],
                           capture_output=True, text=True, env=env, timeout=1500)
    if r.returncode != 0:
        return {"error": f"{exe} exited {r.returncode}: {r.stderr[-300:]}"}
    ev = [json.loads(ln) for ln in r.stdout.splitlines() if ln.startswith("{")]
    b = [e for e in ev if e.get("event") == "bench"][0]
    return dict(value=W * H / (b["ms_per_step"] * 1e-3), ms_per_step=b["ms_per_step"], exe=exe,
                load_s=[e for e in ev if e.get("event") == "loaded"][0]["seconds"])


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
    sc = syn.make_scene(cfg, opaque=args.scene == "opaque")
    intr, poses = syn.intrinsics(W, H), poses_for(args.workload)
    r = NerfRender(1, device_ids=[local])
    r.load_raw(sc.desc, sc.params, sc.grid, sc.aabb)
    pose_par = world > 1 and args.shard == "poses"   # whole frames per rank (pose k -> rank k % world), no intra-frame traffic
    if world > 1 and not pose_par:
        r.set_shard(rank, world)
    r.set_resolution((W, H))
    n_local = r.local_rays()
    ext = torch.cuda.ExternalStream(r.stream(), device=torch.device("cuda", local))
    rgb_ptr, dep_ptr, _ = r.device_outputs()
    gatherer = local_rgb = host_frame = None
    peer = world > 1 and args.gather == "peer" and not pose_par
    if peer:
        # the gather is fused into the composite kernel: rank 0 exports its frame buffer (CUDA IPC), every other rank
        # maps it and stores its pixels there over NVLink; the only collective left is a 4-byte completion fence
        handle = torch.zeros(64, dtype=torch.uint8, device="cuda")
        if rank == 0:
            handle.copy_(torch.frombuffer(bytearray(r.peer_frame_export()), dtype=torch.uint8))
        dist.broadcast(handle, 0)
        if rank != 0:
            r.peer_frame_attach(bytes(handle.cpu().numpy().tobytes()))
        if rank == 0:
            r.peer_frame_copy_out(True)   # every frame is copied to a pinned host slot behind its fence, on the copy stream
        dist.barrier()
    elif world > 1 and not pose_par:
        from nerf_cuda_b200.multigpu import FrameGather, FrameSharding
        sh = FrameSharding(W, H, world)
        with torch.cuda.stream(ext):
            gatherer = FrameGather(sh, rank, torch.device("cuda", local))
        local_rgb = torch.as_tensor(_DevPtr(rgb_ptr, n_local * 3), device="cuda").view(-1, 32, 32, 3)
        host_frame = torch.empty(H, W, 3, dtype=torch.uint8).pin_memory() if rank == 0 else None

    def frame(i, to_host):
        """One step.  to_host: reference-facing semantics (host pose in, host u8 frame out on rank 0)."""
        pose = poses[(i * world + rank) % len(poses)] if pose_par else poses[i % len(poses)]
        with torch.cuda.stream(ext):
            if world == 1 or pose_par:
                if to_host:
                    return r.render_frame(intr, pose)
                r.render_frame_device(intr, pose)
                return None
            r.render_frame_device(intr, pose)
            if peer:
                # nothing else on the device side: the pixel stores ARE the gather and the fence is one-sided (each rank's
                # stream publishes its frame number after its composite; rank 0's copy stream waits for all of them).
                # Reference-facing: the frame's D2H runs behind its fence while the next frame renders; the step hands
                # over the PREVIOUS frame (two pinned slots), the last one is drained inside the timed region.
                if to_host:
                    return r.peer_frame_collect(1)
                return None
            full = gatherer(local_rgb)                      # NCCL gather of packed u8 tiles + tile transpose on rank 0
            if to_host and rank == 0:
                host_frame.copy_(full, non_blocking=True)
                ext.synchronize()
        return None

    def drain(to_host):
        """End of a timed loop, inside the timed region: everything the last frame still has in flight."""
        if peer:
            if to_host:
                r.peer_frame_collect(0)
            r.frame_join()

    def timed(to_host, steps, warmup):
        for i in range(warmup):
            frame(i, to_host)
        r.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(ext)
        for i in range(steps):
            frame(warmup + i, to_host)
        drain(to_host)
        e1.record(ext)
        e1.synchronize()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    def frame_stats(steps, warmup):
        """Untimed pass over the frames of the timed loop, one host sync per frame, for the per-frame counters and the
        per-stage event times (reading them inside the timed loop would serialise it)."""
        keys = ("ms_setup", "ms_march", "ms_network", "ms_composite", "ms_total", "n_samples", "n_samples_marched", "n_samples_used", "n_rounds")
        acc = {k: [] for k in keys}
        for i in range(steps):
            frame(warmup + i, False)
            st = r.stats()
            for k in keys:
                acc[k].append(float(getattr(st, k)))
        drain(False)
        r.synchronize()
        return acc

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for i in range(len(poses)):          # size every buffer for the largest pose before anything is timed
        frame(i, False)
    r.synchronize()                      # sync point: the buffers grow to the largest frame enqueued so far
    ovf0 = r.overflowed_frames()
    ms_dev = timed(False, args.steps, args.warmup)
    ms_e2e = timed(True, args.steps, max(3, args.warmup // 2))
    fs = frame_stats(args.steps, args.warmup)
    net_ms, samples, marched, used, rounds = fs["ms_network"], fs["n_samples"], fs["n_samples_marched"], fs["n_samples_used"], fs["n_rounds"]
    per_rank = {k: float(np.mean(v)) for k, v in fs.items()}
    if world > 1:   # every rank's counters and stage times on the JSON line (load balance of the tile sharding)
        allr = [None] * world
        dist.all_gather_object(allr, per_rank)
        per_rank = allr
    else:
        per_rank = [per_rank]
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
    rays = W * H * (world if pose_par else 1)   # per step: one frame, or one frame per rank when poses are sharded
    pk = peaks()
    table_bytes = 2 * int(syn.level_table(sc.desc)[0][-1]) * 2
    table_mb = table_bytes / 1e6
    l2_resident = table_mb < 100
    # roofline denominators measured NOW on this GPU (kernels_probe.cu): L2->SM streaming bandwidth over a window the size
    # of the table (every CTA streams the whole window, so no SM's L1 can hold it) or HBM streaming when the table does not
    # fit L2; and the random-sector gather rate over the same window (the access shape of a hashed level)
    mem = {}
    if rank == 0:
        win = max(8 << 20, (table_bytes + (1 << 20) - 1) >> 20 << 20)
        mem["stream_gbs"] = r.bench_memory(2 if l2_resident else 0, win if l2_resident else 4 << 30, 3)[0]
        mem["gather_gloads"] = r.bench_memory(1, win, 3)[1]
    out = None
    if rank == 0:
        net = float(np.mean(net_ms))
        s_k = float(np.mean(samples))  # rank 0's samples per launch
        n_launch = float(np.mean(rounds))   # k_network launches per frame (one per evaluation round); s_k, net are per frame
        alg_bytes = s_k * GATHER_BYTES_PER_SAMPLE
        achieved = alg_bytes / (net * 1e-3) / 1e9
        ncu = ncu_capture()
        out = {
            "metric": "rays_per_second_1080p" if args.workload == "C3" else f"rays_per_second_{args.workload}",
            "value": rays * args.steps / (ms_dev * 1e-3), "unit": "rays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak" if pose_par else "strong",
            "vs_baseline": None, "dtype": "f16", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {W}x{H} frame, bound {sc.desc.bound:g}, {sc.desc.cascades} cascades, "
                                   f"log2_hashmap_size {sc.desc.log2_hashmap_size}, base.json MLPs, random-init weights, "
                                   f"synthetic occupancy, {len(poses)} orbit poses cycled",
                       "sharding": "whole frame" if world == 1 else (f"pose-parallel: every step renders {world} frames, pose k on rank k % {world}, whole frames, "
                                                                     f"each rank returns its own frame to its host; no intra-frame traffic" if pose_par else
                           f"32x32-pixel tiles round-robin over {world} ranks; every rank's composite kernel stores its u8 pixels straight into "
                           f"rank 0's frame over NVLink (CUDA IPC peer stores); completion fence one-sided: per-rank st.release.sys sequence "
                           f"flags polled by a one-warp kernel on rank 0's copy stream, no collective; rank 0's D2H double-buffered behind the fence" if peer else
                           f"32x32-pixel tiles round-robin over {world} ranks, NCCL gather of packed u8 tiles to rank 0"),
                       "scene": args.scene, "samples_per_frame": smp_total, "samples_per_ray": smp_total / rays,
                       "samples_rank0": {"marched": float(np.mean(marched)), "evaluated": s_k, "composited": float(np.mean(used)),
                                         "evaluation_rounds": n_launch},
                       "l2": f"inputs larger than L2: each frame streams {smp_total / world * 56 / 1e6:.0f} MB of sample records per GPU (20 B written by the march, 16 read + 8 written by the network, 12 read by the composite) "
                             f"and the pose changes every step; the {table_mb:.1f} MB hash table is "
                             + ("L2-resident by design" if table_mb < 100 else "far larger than the 126 MB L2 (HBM-resident gather)")},
            "e2e": {"value": rays * args.steps / (ms_e2e * 1e-3), "unit": "rays/s", "ms_per_step": ms_e2e / args.steps,
                    "batch_call_value": batch_rps, "h2d_bytes_per_step": 80, "d2h_bytes_per_step": rays * 4 + 8 if (world == 1 or peer) else rays * 3 + 8},
            "per_rank": [{"rank": k, "samples_evaluated": p["n_samples"], "samples_marched": p["n_samples_marched"],
                          "stage_ms": {"rays+count+scan": p["ms_setup"], "emit": p["ms_march"], "network": p["ms_network"],
                                       "composite": p["ms_composite"], "total": p["ms_total"]}} for k, p in enumerate(per_rank)],
            "gpu_launches": 5 * int(sum(rounds)) + (int(3 * args.steps) if peer else 0),   # rank 0, timed device loop; per frame and round: count (or round sums), scan, emit, network, composite (+ 3 one-warp fence kernels per frame with the peer frame)
            "overflowed_timed_frames": int(ovf),
            "clocks": clocks,
            "roofline": {"kernel": "k_network<2> (hash-grid gather + tcgen05 MLP, fused)",
                         "bound": "l2" if l2_resident else "hbm",
                         "achieved": achieved, "peak": mem["stream_gbs"], "unit": "GB/s", "frac": achieved / mem["stream_gbs"],
                         "peak_source": (f"L2->SM streaming read bandwidth over a {table_mb:.0f} MB window, measured live on this GPU "
                                         "(nrf_bench_memory mode 2); the table is L2-resident, so HBM is not the bound" if l2_resident else
                                         "HBM streaming read bandwidth over a 4 GB window, measured live on this GPU (nrf_bench_memory mode 0)"),
                         "hbm_copy_peak": pk["hbm"], "hbm_copy_peak_source": pk["src"], "frac_of_hbm_copy_peak": achieved / pk["hbm"],
                         "algorithmic_bytes_per_sample": GATHER_BYTES_PER_SAMPLE,
                         "stream_bytes_per_sample_not_counted": STREAM_BYTES_PER_SAMPLE,
                         "samples_per_launch": s_k / n_launch, "ms_per_launch": net / n_launch, "launches_per_frame": n_launch,
                         "traffic": (ncu["dram_bytes_per_sample"] * s_k / n_launch) if ncu and l2_resident else None,
                         "traffic_source": (f"{ncu['dram_bytes_per_sample']:.1f} DRAM B/sample ({ncu['source']}, {ncu['workload']}) x samples per launch"
                                            if ncu and l2_resident else None),
                         # the pipes that actually limit the kernel (ncu --set full of one full-size launch, profiles/)
                         "l1_wavefront_frac": ncu["l1_wavefronts_pct_of_peak"] / 100 if ncu else None,
                         "tensor_pipe_pct": ncu["tensor_pipe_pct"] if ncu else None,
                         "gather_lookups_per_ns": s_k * 128 / (net * 1e-3) / 1e9,
                         "random_sector_gather_loads_per_ns_measured": mem["gather_gloads"],
                         "mlp_tflops_in_fused": s_k * MLP_FLOP_PER_SAMPLE / (net * 1e-3) / 1e12,
                         "mlp_frac_of_sustained_bf16_peak": s_k * MLP_FLOP_PER_SAMPLE / (net * 1e-3) / 1e12 / pk["tf_sus"]},
            "stages_alone": {
                **stage,
                "encode_gather_gbs": smp_local * GATHER_BYTES_PER_SAMPLE / (stage["encode_ms"] * 1e-3) / 1e9,
                "encode_frac_of_measured_stream_peak": smp_local * GATHER_BYTES_PER_SAMPLE / (stage["encode_ms"] * 1e-3) / 1e9 / mem["stream_gbs"],
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
    gatherer = local_rgb = host_frame = frame = timed = drain = None
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


def inproc_arm(args):
    """--mode inproc: ONE process drives all N GPUs through nrf_create(n_gpus) — the path ngp::NerfRender(n) and
    `render_server --gpus n` use (the reference's own multi-GPU shape: one process, NGPU devices, common.h:91).  Tiles are
    sharded round-robin over the devices, every device's composite stores its pixels into device 0's frame over NVLink,
    device 0's stream waits on the other devices' events; no collective, no second process."""
    import torch
    from nerf_cuda_b200 import NerfRender, synthetic as syn
    n = args.gpus
    cfg, W, H, _ = WORKLOADS[args.workload]
    sc = syn.make_scene(cfg, opaque=args.scene == "opaque")
    intr, poses = syn.intrinsics(W, H), poses_for(args.workload)
    torch.cuda.set_device(0)
    r = NerfRender(n)
    r.load_raw(sc.desc, sc.params, sc.grid, sc.aabb)
    r.set_resolution((W, H))
    ext = torch.cuda.ExternalStream(r.stream(), device=torch.device("cuda", 0))
    sampler = ClockSampler(0)
    sampler.start()
    for i in range(len(poses)):
        r.render_frame_device(intr, poses[i])
    r.synchronize()
    ovf0 = r.overflowed_frames()

    def timed(to_host, steps, warmup):
        for i in range(warmup):
            (r.render_frame if to_host else r.render_frame_device)(intr, poses[i % len(poses)])
        r.synchronize()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(ext)
        for i in range(steps):
            (r.render_frame if to_host else r.render_frame_device)(intr, poses[(warmup + i) % len(poses)])
        r.frame_join()                     # device 0's stream waits for the other devices' last kernels
        e1.record(ext)
        e1.synchronize()
        for d in range(n):
            torch.cuda.synchronize(d)
        return e0.elapsed_time(e1)

    ms_dev = timed(False, args.steps, args.warmup)
    ms_e2e = timed(True, args.steps, max(3, args.warmup // 2))
    ovf = r.overflowed_frames() - ovf0
    if ovf:
        raise RuntimeError(f"{ovf} timed frame(s) overflowed the sample buffers: the measurement is invalid")
    pb = np.stack([poses[(args.warmup + i) % len(poses)] for i in range(args.steps)])
    r.render_batch(intr, pb[: min(len(pb), len(poses))])
    tb = time.perf_counter()
    r.render_batch(intr, pb)
    batch_rps = W * H * len(pb) / (time.perf_counter() - tb)
    clocks = sampler.stop()
    rounds = []
    for i in range(args.steps):
        r.render_frame_device(intr, poses[(args.warmup + i) % len(poses)])
        rounds.append(int(r.stats().n_rounds))
    st = r.stats()
    rays = W * H
    out = {"metric": "rays_per_second_1080p" if args.workload == "C3" else f"rays_per_second_{args.workload}",
           "value": rays * args.steps / (ms_dev * 1e-3), "unit": "rays/s", "n_gpus": n, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f16",
           "data": "synthetic",
           "config": {"workload": f"{args.workload}: {W}x{H} frame, bound {sc.desc.bound:g}, {sc.desc.cascades} cascades, log2_hashmap_size "
                                  f"{sc.desc.log2_hashmap_size}, base.json MLPs, random-init weights, synthetic occupancy, {len(poses)} orbit poses cycled",
                      "mode": "inproc: one process, nrf_create(n_gpus) — the ngp::NerfRender(n) / render_server --gpus n path",
                      "sharding": f"32x32-pixel tiles round-robin over {n} devices of one process; every device's composite kernel stores its u8 "
                                  f"pixels into device 0's frame over NVLink (peer access), device 0 waits on the other devices' events; no collective",
                      "scene": args.scene, "samples_per_frame": float(st.n_samples), "samples_per_ray": float(st.n_samples) / rays,
                      "l2": "inputs larger than L2 (see the default mode's line); the pose changes every step"},
           "e2e": {"value": rays * args.steps / (ms_e2e * 1e-3), "unit": "rays/s", "ms_per_step": ms_e2e / args.steps,
                   "batch_call_value": batch_rps, "h2d_bytes_per_step": 80, "d2h_bytes_per_step": rays * 4 + 8 * n,
                   "note": "value: nrf_render_frame per pose (returns the host frame, one sync per frame); batch_call_value: the same poses "
                           "through nrf_render_batch (frame i's D2H on a copy stream beside frame i+1's kernels), wall clock"},
           "gpu_launches": 5 * int(sum(rounds)) * n, "overflowed_timed_frames": int(ovf), "clocks": clocks,
           "frame_breakdown_ms_device0": {"rays+count+scan": st.ms_setup, "emit": st.ms_march, "network": st.ms_network,
                                          "composite": st.ms_composite, "total": st.ms_total}}
    print(json.dumps(out), flush=True)
    r.close()


def reference_arm(args):
    """--impl reference.  The reference has no CPU implementation of this path — it IS a CUDA renderer — so the arm runs
    the reference's own CUDA code (baseline/_ref, built from /root/reference by baseline/build_ref_cuda.sh) on GPU 0 of
    this box when the binary travelled here; the CPU port (oracle/) is timed beside it as cpu_baseline, and is the arm
    itself only where the binary or a GPU is missing.  The reference's GPU count is a compile-time constant (NGPU,
    common.h:91; built with 1), so for --gpus N > 1 the line is still its single-GPU number and says so."""
    cfg, W, H, _ = WORKLOADS[args.workload]
    metric = "rays_per_second_1080p" if args.workload == "C3" else f"rays_per_second_{args.workload}"
    steps, warm = max(1, args.steps), max(0, args.warmup)
    base = {"impl": "reference", "metric": metric, "unit": "rays/s", "n_gpus": args.gpus, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f16", "data": "synthetic"}
    g = ref_cuda_arm(args.workload, steps, warm, "ref_cuda_driver_d1")
    if g and "error" not in g:
        shipped = ref_cuda_arm(args.workload, max(2, min(steps, 8)), min(warm, 2), "ref_cuda_driver")
        c = None if args.no_cpu else cpu_arm(args.workload, steps=2, warmup=1, target_s=4.0)
        return {**base, "value": g["value"], "steps": steps, "warmup": warm, "ms_per_step": g["ms_per_step"],
                "config": {"workload": f"{args.workload}: {W}x{H} frame, same model, occupancy grid and {WORKLOADS[args.workload][3]} orbit poses as "
                                       f"the default arm, handed over as a .msgpack snapshot",
                           "implementation": "the reference's own CUDA renderer (src/nerf_render.cu + tiny-cuda-nn, compiled for sm_100 by "
                                             "baseline/build_ref_cuda.sh), NerfRender::render_frame on ONE GPU (NGPU is a compile-time constant)",
                           "variant": "D1-clean: one cudaMemsetAsync of the deltas buffer before the march so no uninitialised samples are "
                                      "evaluated (SURVEY.md defect D1) — the FASTER of the two builds, i.e. the conservative baseline; "
                                      "`as_shipped` is the unmodified source"},
                "as_shipped": None if not shipped or "error" in shipped else
                              {"value": shipped["value"], "ms_per_step": shipped["ms_per_step"], "unit": "rays/s"},
                "load_seconds": g["load_s"],
                "cpu_baseline": ({"value": c["value"], "unit": "rays/s", "cores": c["cores"], "kind": "port", "sample": c["sample"]} if c else None),
                "e2e": {"value": g["value"], "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                        "note": "render_frame itself takes a host pose and returns the host u8 frame: value is already end to end"}}
    steps, warm = min(steps, 8), min(warm, 2)
    c = cpu_arm(args.workload, steps=steps, warmup=warm, target_s=6.0)
    return {**base, "value": c["value"], "steps": steps, "warmup": warm, "ms_per_step": c["ms_per_step"],
            "config": {"workload": f"{args.workload}: {W}x{H} frame (CPU arm renders a bounded row band per step)",
                       "implementation": "CPU port of the reference's path (oracle/, OpenMP over rays): "
                                         + (g["error"] if g else "baseline/_ref/ref_cuda_driver_d1 or a GPU is not present")},
            "cpu_baseline": {"value": c["value"], "unit": "rays/s", "cores": c["cores"], "kind": "port", "sample": c["sample"]},
            "e2e": {"value": c["value"], "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C3", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--scene", default="random", choices=["random", "opaque"],
                    help="random = random-init weights (BASELINE.json's config: semi-transparent fog, nothing terminates early); opaque = same "
                         "model with the density head scaled so rays saturate after a few samples (a trained scene's behaviour)")
    ap.add_argument("--mode", default="ranks", choices=["ranks", "inproc"],
                    help="N>1: ranks (default) = one process per GPU (torchrun); inproc = one process drives all N GPUs through "
                         "nrf_create(N), the path render_server --gpus N uses")
    ap.add_argument("--shard", default="tiles", choices=["tiles", "poses"],
                    help="N>1: tiles (default) = every rank renders its 32x32-pixel tiles of the SAME frame (strong scaling, latency); "
                         "poses = rank k renders whole frames of poses k, k+N, ... (throughput; no intra-frame traffic)")
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl"],
                    help="N>1: peer (default) = composite kernels store into rank 0's frame over NVLink, one-sided sequence-flag fence, "
                         "no collective; nccl = gather of packed u8 tiles to rank 0 + tile transpose (the round-1 path)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:
            return
        print(json.dumps(reference_arm(args)), flush=True)
        return
    if args.mode == "inproc" and args.gpus > 1:
        if int(os.environ.get("RANK", "0")) == 0:   # under torchrun the other ranks have nothing to do
            inproc_arm(args)
        return
    gpu_arm(args)


if __name__ == "__main__":
    main()
