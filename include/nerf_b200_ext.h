/*
 * nerf_b200_ext.h — multi-process / device-resident extensions of the C ABI (no counterpart in the reference, whose
 * multi-GPU is a compile-time `#define NGPU 2` with host threads, include/nerf-cuda/common.h:91 and
 * src/nerf_render.cu:252-253,345-358).  One process per GPU (torchrun) uses these to render its shard of the frame
 * and to expose the device buffers that the NVLink gather (NCCL) reads.
 */
#ifndef NERF_B200_EXT_H
#define NERF_B200_EXT_H
#include "nerf_b200.h"
#ifdef __cplusplus
extern "C" {
#endif

/* This context renders only the 32x32-pixel image tiles t with t % n_shards == shard (tiles numbered row-major;
 * reference: pixel = NGPU*tid + gpu, render_utils.h:37).  Call before nrf_set_resolution; the context must own exactly
 * one device.  The u8 outputs of a sharded context are PACKED: its tiles in local order, 1024 pixels each, row-major
 * inside the tile (pixels of an edge tile that fall outside the image are present and zero). */
int nrf_set_shard(nrf_ctx* ctx, int shard, int n_shards);
/* Number of ray slots this context renders per frame (after nrf_set_resolution): its tiles x 1024, including the
 * out-of-image slots of edge tiles. */
int nrf_local_rays(nrf_ctx* ctx, int* n_local);
/* Device pointers of the last frame's outputs on device 0 of the context: rgb u8 / depth u8 (whole frame in pixel
 * order for an unsharded context, packed tiles [n_local*3] / [n_local] for a sharded one) and rgb f32 [n_local*3] in
 * local ray order.  Valid until nrf_set_resolution / nrf_destroy. */
int nrf_device_outputs(nrf_ctx* ctx, void** rgb_u8, void** depth_u8, void** rgb_f32);
/* CUDA stream (cudaStream_t) of device 0, for event timing on the launching stream. */
int nrf_stream(nrf_ctx* ctx, void** stream);
/* Global pixel index of local ray q, or -1 for a slot outside the image (host arithmetic; same map the kernels use:
 * q = local_tile*1024 + patch*32 + lane, patches of 8x4 pixels, 4x8 patches per tile). */
int nrf_shard_pixel(nrf_ctx* ctx, int q, int* pixel);

/* ---- peer frame: the gather fused into the composite kernel (one process per GPU) ---------------------------------
 * The destination rank ("owner") exports ONE buffer holding two frames used alternately (each: rgb W*H*3 bytes followed
 * by depth W*H bytes, pixel order) plus a few 32-bit sequence flags, as a 64-byte CUDA IPC handle; every other rank
 * attaches it.  From then on each rank's composite kernel stores its u8 pixels straight into that buffer over NVLink (peer
 * stores): no collective moves pixels, and none orders them either — the completion fence is ONE-SIDED.  After its composite
 * of frame k every rank's stream stores k into its own flag (st.release.sys); the owner's copy stream waits until every
 * flag has reached k (a one-warp kernel polling with ld.acquire.sys), copies the slot to the host if asked to, and
 * publishes ack = k; a rank waits for ack >= k-2 before its composite overwrites the slot frame k-2 was in, so the owner
 * may still be copying frame k-1 out while everybody renders frame k.  Ranks are never held in lockstep and no rank waits
 * for a frame it does not write.  Every rank must render every frame (frames are counted, not named).  A wait gives up
 * after 20 s and the owner's next sync point fails instead of hanging the GPU.
 * The exporting context writes into its own buffer directly.  Call after nrf_set_shard + nrf_set_resolution; detach
 * before changing either.  A frame whose samples do not fit a rank's buffers is re-rendered at that rank's next host sync
 * point, which the owner cannot see: size the buffers with an untimed pass over the poses first (as bench.py does; frames
 * that overflowed are counted by nrf_overflowed_frames).
 * (Reference: host threads + 2 D2H copies per GPU + CPU re-interleave, nerf_render.cu:345-358.) */
int nrf_peer_frame_export(nrf_ctx* ctx, void* handle64);
int nrf_peer_frame_attach(nrf_ctx* ctx, const void* handle64);
int nrf_peer_frame_detach(nrf_ctx* ctx);
/* Owner: device pointer of the latest frame's slot. */
int nrf_peer_frame_device(nrf_ctx* ctx, void** frame);
/* Owner: the latest frame in context-owned pinned memory, complete (every rank's fence passed) when the call returns. */
int nrf_peer_frame_to_host(nrf_ctx* ctx, uint8_t** rgb, uint8_t** depth);
/* Owner: copy every later frame to the host behind its fence, on the copy stream, overlapping the next frame (two pinned
 * slots); nrf_peer_frame_collect(back) then returns frame (latest - back), back = 0 or 1, waiting only for that frame.  On
 * the other ranks collect waits for the rank's own kernels of that frame and returns NULL pointers. */
int nrf_peer_frame_copy_out(nrf_ctx* ctx, int on);
int nrf_peer_frame_collect(nrf_ctx* ctx, int back, uint8_t** rgb, uint8_t** depth);
/* Make device 0's stream wait for whatever the latest frame still has in flight elsewhere: the other devices' kernels (one
 * process, several devices) and, on the owner of a peer frame, that frame's fence.  An event recorded on nrf_stream()
 * afterwards covers the whole frame. */
int nrf_frame_join(nrf_ctx* ctx);

/* Asynchronous device frames (nrf_render_frame_device without a sync in between) whose samples did not fit the sample
 * buffers: such a frame is rendered incompletely; the LATEST one is re-rendered at the next sync point, earlier ones
 * are not (their pixels were already overwritten), and the buffers grow to the largest frame seen.  Cumulative count
 * since nrf_set_resolution; a caller that times or keeps asynchronous frames checks that it did not move.  The
 * reference has no equivalent: it blocks on a device->host read after every <= 8 march steps (nerf_render.cu:291). */
int nrf_overflowed_frames(nrf_ctx* ctx, uint64_t* n_frames);

/* Per-level hash-grid scale[L] and resolution[L] of the loaded model AS THE DEVICE EVALUATES THEM.  The reference derives
 * both inside kernel_grid with the device's exp2f (tcnn grid.h:186-187: exp2f(level*log2(per_level_scale))*base - 1),
 * which differs from the host's exp2f by an ulp on some levels; nrf_level_table (host arithmetic, usable without a GPU)
 * returns the host values, which is what the OFFSET table is built from (grid.h:899-931).  Either pointer may be NULL. */
int nrf_get_level_scales(nrf_ctx* ctx, float* scales, uint32_t* resolutions);

/* nrf_render_batch with the depth channel: depth_out (may be NULL) receives n * W*H bytes, frame after frame, the u8
 * depth of struct Image (common.h:76-89) that the reference computes but never sends (render_server.cu:101). */
int nrf_render_batch_rgbd(nrf_ctx* ctx, const float intr[4], const float* poses_rowmajor, int n_poses, uint8_t* rgb_out,
                          uint8_t* depth_out);

/* Termination-aware evaluation.  The reference kills a ray as soon as its transmittance falls below 1e-4
 * (render_utils.h:727, rays_t = -1) and compacts it away before its next wavefront of <= 8 steps (kernel_compact_rays,
 * render_utils.h:394-415; nerf_render.cu:278-302), so samples behind an opaque surface never reach its network.  Here a
 * frame is evaluated in ROUNDS of samples per ray — by default [0,8), [8,32), [32,128), [128,1024) — each round marching,
 * evaluating and compositing only the rays still alive, all enqueued without a host round trip.  mode 0: one round (whole
 * rays; right when almost nothing terminates, e.g. random-init weights); 1: rounds always; 2 (default): automatic — rounds
 * are switched on when the last validated frame composited < 3/4 of the samples its march found and off again above
 * 9/10.  The pixels, depths and composited counts are identical in every mode.  bounds/n_bounds (NULL/0 = keep): the
 * cumulative per-ray sample boundaries of all rounds but the last (1..3 values, ascending, < NRF_MAX_STEPS). */
int nrf_set_termination(nrf_ctx* ctx, int mode, const int* bounds, int n_bounds);

/* nrf_eval_samples_noalloc — NO COUNTERPART IN THE REFERENCE.
 * Evaluates n samples (hash-grid encode -> density + colour MLP) WITHOUT ANY ALLOCATION, for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it): the intermediate fp16 features
 * (64 bytes per sample) go to a fixed file-scope device array inside the library instead of an allocated buffer.  That array
 * (NRF_NOALLOC_MAX_SAMPLES x 64 B = 4 MB) is held by every GPU that loads the library until the process exits.
 *   positions_dev : DEVICE pointer, n x float4 {u, v, w, ignored}, positions normalised to [0,1]^3, 16-byte aligned
 *   dirs_dev      : DEVICE pointer, n x float4 {dx, dy, dz, ignored}, direction normalised to [0,1] (d*0.5+0.5), 16-byte aligned
 *   out_dev       : DEVICE pointer, n x 4 fp16 {r, g, b, sigma} (8 bytes per sample), 8-byte aligned
 * The caller owns all three.  The two kernels are enqueued on the context's stream (nrf_stream) of device 0 and the call
 * returns without synchronising; it calls no cudaMalloc / cudaFree or allocating library routine (the first call on a device
 * resolves the array's device address with cudaGetSymbolAddress).  n > NRF_NOALLOC_MAX_SAMPLES is rejected with NRF_EINVAL:
 * split the batch.  Results equal nrf_probe_samples on the same inputs.  Every call in the process (any context on that GPU)
 * shares the one array: it keeps the last call's contents (nothing needs clearing: a call overwrites the n entries it reads),
 * and calls that overlap on the GPU — other streams, other contexts on the same device — race on it; calls on one context
 * are ordered by its stream. */
#define NRF_NOALLOC_MAX_SAMPLES 65536
int nrf_eval_samples_noalloc(nrf_ctx* ctx, int n, const void* positions_dev, const void* dirs_dev, void* out_dev);

/* Memory micro-benchmarks on device 0 (roofline denominators for the hash gather).  mode 0: coalesced
 * 16-byte streaming reads over a window split across the CTAs (window >> L2 => HBM read bandwidth); mode 2: every CTA
 * streams the WHOLE window (per-SM footprint = window, so the SM's own L1 cannot serve it: window <= L2 => L2->SM bandwidth);
 * mode 1: one random 8-byte slot per lane (one 32-byte sector in its own 128-byte line: the access shape of a hashed
 * grid level).  *gbs = requested bytes per second / 1e9; *gloads (may be NULL) = loads per nanosecond. */
int nrf_bench_memory(nrf_ctx* ctx, int mode, size_t window_bytes, int iters, float* gbs, float* gloads);

#ifdef __cplusplus
}
#endif
#endif
