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

/* This context renders only pixel blocks b with b % n_shards == shard, block = `block_pixels` consecutive pixels in
 * row-major order (reference: pixel = NGPU*tid + gpu, render_utils.h:37, i.e. block_pixels = 1).  Call before
 * nrf_set_resolution.  Outputs of a sharded context are packed in local ray order. */
int nrf_set_shard(nrf_ctx* ctx, int shard, int n_shards, int block_pixels);
/* Number of rays this context renders per frame (after nrf_set_resolution). */
int nrf_local_rays(nrf_ctx* ctx, int* n_local);
/* Device pointers of the last frame's packed outputs on device 0 of the context: rgb u8 [n_local*3],
 * depth u8 [n_local], rgb f32 [n_local*3].  Valid until nrf_set_resolution / nrf_destroy. */
int nrf_device_outputs(nrf_ctx* ctx, void** rgb_u8, void** depth_u8, void** rgb_f32);
/* CUDA stream (cudaStream_t) of device 0, for event timing on the launching stream. */
int nrf_stream(nrf_ctx* ctx, void** stream);
/* Global pixel index of local ray q (host arithmetic; same map the kernels use). */
int nrf_shard_pixel(nrf_ctx* ctx, int q, int* pixel);

#ifdef __cplusplus
}
#endif
#endif
