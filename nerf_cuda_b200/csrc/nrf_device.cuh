// nrf_device.cuh — device-side parameter blocks and small helpers shared by the sm_100a kernels.
//
// HBM data layout (one replica per GPU):
//   table      : __half2[sum T_l]         hash-grid entries, level l at lv[l].offset (tcnn grid.h layout, F=2)
//   mlp_packed : uint4[1280] (20 KB)      the five weight matrices pre-packed into the tcgen05 K-major
//                                         SWIZZLE_NONE core-matrix order, copied verbatim into shared memory
//   mlp_rowmaj : __half[10240]            same weights, reference order [out][in] (SIMT check kernel only)
//   occupancy  : uint32[C*H^3/32]         bit == (density > min(0.01, mean_density)), same linear cell index as
//                                         the reference's float grid (render_utils.h:613-619)
//   rays       : float4[N] {dx,dy,dz,near} + float2[N] {far, t_first}; N = local rays in patch-major order (ShardMap)
//   samples    : float4[S] {u,v,w (position normalised to [0,1]), ray index} + float[S] t (SampleSrc), in
//                WARP-INTERLEAVED order: for each 32-ray patch, step k of all its live rays, then step k+1, ...
//                (stage probes: float4 {u,v,w,dt} + float[S] t_after + int[S] ray)
//   warp_base  : int[N/32] first sample of each patch
//   tcache     : float[N/32][NRF_TCACHE][32] t of each ray's first NRF_TCACHE samples (pass 1 -> pass 2)
//   ray_sh     : uint4[2N] the 16 fp16 SH coefficients of every ray's direction
//   sample_out : uint2[S]  four halves {r,g,b,sigma}
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#define NRF_L 16          // hash-grid levels (base.json)
#define NRF_TILE 128      // samples per MMA tile == UMMA M
#define NRF_MAX_STEPS_D 1024
#ifndef NRF_TCACHE
#define NRF_TCACHE 64     // sample t values cached per ray between the two march passes (kernels_exact.cu)
#endif

struct LevelInfo {
  uint32_t offset;  // first entry of the level (in entries)
  uint32_t size;    // T_l entries
  uint32_t res;     // grid resolution
  float scale;      // exp2f(l*log2(b))*N_min - 1 evaluated on the DEVICE like the reference's kernel_grid does (grid.h:186;
                    // MUFU.EX2 and glibc's exp2f differ by an ulp on some levels), once at load time (k_level_scales)
  uint32_t hashed;  // 0 dense index; 1 fast_hash with paired 8-byte slot loads; 2 fast_hash, 4-byte loads; 3 wrapped dense index & (T-1)
  uint32_t pad;
  const uint32_t* tab;  // first entry of the level (one __half2 = one uint32 per entry): address = tab + index, 1 IMAD.WIDE
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
  float rbound;     // 1/bound (IEEE, host)
  const uint32_t* occupancy;
  const float* div_lut;  // [0..H]: (float)m / (float)(H-1) (IEEE division done once on the host); [H+1..2H+1]: fl(2*that - 1)
  uint32_t fast;         // bound == 2^(C-1), H and 1/dt_gamma powers of two: the trimmed march step (kernels_exact.cu)
  float t_lo, t_hi;      // dt_min/dt_gamma, dt_max/dt_gamma: between them t*dt_gamma needs no clamp
};

// Frame path: a sample is the 16-byte record {u,v,w, ray index} k_network reads, plus the t it was taken at (4 B) for
// k_composite: dt and t_after are functions of t and the per-frame constants below and are recomputed there with
// exactly the expressions of the march step (20 B per sample written instead of 24, and the composite reads 12 B per
// sample in whole sectors instead of 28 B in quarter-used ones).
struct SampleSrc {
  const float* t0;     // null: explicit arrays instead (stage probes, density grid): samples[i].w = dt, sample_ray[i]
  float dt_gamma, dt_min, dt_max;
};

// One evaluation round of a frame: the samples with step index in [s0, s1) of every ray that is still alive.  A frame
// is one round {0, NRF_MAX_STEPS_D, first, last} unless the context runs termination-aware (nrf_api.cu round_schedule):
// then rays whose transmittance fell below 1e-4 in one round are not marched / evaluated / composited in the next.
struct RoundWin {
  int s0, s1;
  int first, last;
  float4* st;      // per-ray accumulators carried between rounds (2 float4 per ray), null for a single round
  uint8_t* dead;   // per-ray "terminated" flag written by k_composite, read by k_round_sums / k_emit / k_composite
};

// activation ids mirror include/nerf_b200.h
#define ACT_NONE 0
#define ACT_RELU 1
#define ACT_EXP 2
#define ACT_SIGMOID 3

#ifdef __CUDACC__
// dt of the occupied branch of the march step (kernels_exact.cu march_step) at a recorded t
__device__ __forceinline__ float record_dt(const SampleSrc& S, float t) {
  return fminf(S.dt_max, fmaxf(S.dt_min, t * S.dt_gamma));
}
__device__ __forceinline__ uint32_t pack_h2(__half a, __half b) {
  return (uint32_t)__half_as_ushort(a) | ((uint32_t)__half_as_ushort(b) << 16);
}
// tcnn kernel_sh degree 4 (spherical_harmonics.h:66-96); in = normalised direction input in [0,1]; 16 halves out.
// Evaluated once per RAY (k_rays_count) — the reference evaluates it once per sample although the direction is per ray.
__device__ __forceinline__ void sh4(float i0, float i1, float i2, uint32_t out[8]) {
  const float x = i0 * 2.f - 1.f, y = i1 * 2.f - 1.f, z = i2 * 2.f - 1.f;
  const float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
  float v[16];
  v[0] = 0.28209479177387814f;
  v[1] = -0.48860251190291987f * y;
  v[2] = 0.48860251190291987f * z;
  v[3] = -0.48860251190291987f * x;
  v[4] = 1.0925484305920792f * xy;
  v[5] = -1.0925484305920792f * yz;
  // the fused multiply-adds are those of the reference's kernel_sh<__half> SASS (nvcc -fmad=true): entries 6, 8, 9, 11,
  // 12, 13, 15; everything else is a plain product (the TUs that include this are compiled with -fmad=false)
  v[6] = __fmaf_rn(0.94617469575755997f, z2, -0.31539156525251999f);
  v[7] = -1.0925484305920792f * xz;
  v[8] = __fmaf_rn(0.54627421529603959f, x2, -(0.54627421529603959f * y2));
  v[9] = 0.59004358992664352f * y * __fmaf_rn(-3.0f, x2, y2);
  v[10] = 2.8906114426405538f * xy * z;
  v[11] = 0.45704579946446572f * y * __fmaf_rn(-5.0f, z2, 1.0f);
  v[12] = 0.3731763325901154f * z * __fmaf_rn(5.0f, z2, -3.0f);
  v[13] = 0.45704579946446572f * x * __fmaf_rn(-5.0f, z2, 1.0f);
  v[14] = 1.4453057213202769f * z * (x2 - y2);
  v[15] = 0.59004358992664352f * x * __fmaf_rn(3.0f, y2, -x2);
#pragma unroll
  for (int k = 0; k < 8; ++k) out[k] = pack_h2(__float2half_rn(v[2 * k]), __float2half_rn(v[2 * k + 1]));
}
#endif

// Ray order and sharding.  The image is cut into 32x32-pixel tiles; tile t belongs to shard t % n_shards (round-robin
// for balance; the reference interleaves single pixels, pixel = NGPU*tid + gpu, render_utils.h:37).
// Inside a tile rays are ordered patch-major: 32 patches of 8x4 pixels, 32 rays (one WARP) per patch, so that the 32
// lanes of every march / composite warp — and, through the step-interleaved sample list, of every gather warp — are
// spatial neighbours.  Local ray index q = local_tile*1024 + patch*32 + lane.  Tiles that stick out of
// the image contain invalid rays (pixel -1, zero samples).
#define NRF_TILE_PX 32
struct ShardMap {
  int shard, n_shards, W, H, tiles_x, tiles_y;
};
__host__ __device__ inline ShardMap make_shard_map(int shard, int n_shards, int W, int H) {
  ShardMap m;
  m.shard = shard; m.n_shards = n_shards; m.W = W; m.H = H;
  m.tiles_x = (W + NRF_TILE_PX - 1) / NRF_TILE_PX;
  m.tiles_y = (H + NRF_TILE_PX - 1) / NRF_TILE_PX;
  return m;
}
__host__ __device__ inline int shard_tile_count(const ShardMap& m) {
  const int n = m.tiles_x * m.tiles_y;
  return n / m.n_shards + (m.shard < n % m.n_shards ? 1 : 0);
}
__host__ __device__ inline int shard_local_count(const ShardMap& m) { return shard_tile_count(m) * (NRF_TILE_PX * NRF_TILE_PX); }
// local ray q -> pixel coordinates; returns false for rays outside the image
__host__ __device__ inline bool shard_pixel_xy(const ShardMap& m, int q, int& px, int& py) {
  const int lt = q >> 10, w = q & 1023, patch = w >> 5, lane = w & 31;
  const int tile = lt * m.n_shards + m.shard;
  const int tx = tile % m.tiles_x, ty = tile / m.tiles_x;
  px = tx * NRF_TILE_PX + (patch & 3) * 8 + (lane & 7);
  py = ty * NRF_TILE_PX + (patch >> 2) * 4 + (lane >> 3);
  return px < m.W && py < m.H;
}
__host__ __device__ inline int shard_pixel(const ShardMap& m, int q) {
  int px, py;
  return shard_pixel_xy(m, q, px, py) ? py * m.W + px : -1;
}
// Position of local ray q in the PACKED u8 output of a sharded context: tiles in local order, pixels row-major inside
// the 32x32 tile (what multigpu.FrameSharding.assemble un-tiles after the gather).
__host__ __device__ inline int shard_packed_index(int q) {
  const int lt = q >> 10, w = q & 1023, patch = w >> 5, lane = w & 31;
  return lt * 1024 + ((patch >> 2) * 4 + (lane >> 3)) * NRF_TILE_PX + (patch & 3) * 8 + (lane & 7);
}
