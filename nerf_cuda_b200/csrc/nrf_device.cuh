// nrf_device.cuh — device-side parameter blocks and small helpers shared by the sm_100a kernels.
//
// HBM data layout (one replica per GPU):
//   table      : __half2[sum T_l]         hash-grid entries, level l at lv[l].offset (tcnn grid.h layout, F=2)
//   mlp_packed : uint4[1280] (20 KB)      the five weight matrices pre-packed into the tcgen05 K-major
//                                         SWIZZLE_NONE core-matrix order, copied verbatim into shared memory
//   mlp_rowmaj : __half[10240]            same weights, reference order [out][in] (SIMT check kernel only)
//   occupancy  : uint32[C*H^3/32]         bit == (density > min(0.01, mean_density)), same linear cell index as
//                                         the reference's float grid (render_utils.h:613-619)
//   rays       : float4[N] {dx,dy,dz,near} + float[N] far
//   samples    : float4[S] {u,v,w (position normalised to [0,1]), dt} + float[S] t_after + int[S] ray
//   sample_out : uint2[S]  four halves {r,g,b,sigma}
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#define NRF_L 16          // hash-grid levels (base.json)
#define NRF_TILE 128      // samples per MMA tile == UMMA M
#define NRF_MAX_STEPS_D 1024

struct LevelInfo {
  uint32_t offset;  // first entry of the level (in entries)
  uint32_t size;    // T_l entries
  uint32_t res;     // grid resolution
  float scale;      // exp2f(l*log2(b))*N_min - 1, computed on the HOST (see DESIGN.md numerics contract)
  uint32_t hashed;  // 1 if res^3 > T_l (fast_hash), 0 = dense index
  uint32_t pad[3];
};

struct NetParams {
  LevelInfo lv[NRF_L];
  const __half2* table;
  const uint4* mlp_packed;
  const __half* mlp_rowmaj;
  uint32_t rgb_act;
  uint32_t sigma_act;
};

struct MarchParams {
  float o[3];       // ray origin (camera centre in ngp space), constant per frame
  float R[9];       // rotation rows (ngp space)
  float intr[4];    // fl_x, fl_y, cx, cy
  float aabb[6];
  float bound, dt_gamma, dt_min, dt_max, min_near, pos_w;
  uint32_t C, H;
  int W;            // image width
  const uint32_t* occupancy;
};

// activation ids mirror include/nerf_b200.h
#define ACT_NONE 0
#define ACT_RELU 1
#define ACT_EXP 2
#define ACT_SIGMOID 3

// Maps local ray index q of this shard to the global pixel index.  Shards interleave blocks of `block` pixels
// round-robin (balance + locality); n_shards == 1 => identity.
struct ShardMap {
  int shard, n_shards, block, n_pixels;
};
__host__ __device__ inline int shard_local_count(const ShardMap& m) {
  if (m.n_shards == 1) return m.n_pixels;
  const int n_blocks = (m.n_pixels + m.block - 1) / m.block;
  int cnt = 0;
  // blocks b with b % n_shards == shard
  const int full = n_blocks / m.n_shards, rem = n_blocks % m.n_shards;
  const int mine = full + (m.shard < rem ? 1 : 0);
  cnt = mine * m.block;
  // the last block overall may be partial
  const int last = n_blocks - 1;
  if (last >= 0 && last % m.n_shards == m.shard) cnt -= n_blocks * m.block - m.n_pixels;
  return cnt;
}
__host__ __device__ inline int shard_pixel(const ShardMap& m, int q) {
  if (m.n_shards == 1) return q;
  const int lb = q / m.block, off = q - lb * m.block;
  return (lb * m.n_shards + m.shard) * m.block + off;
}
