// kernels_exact.cu — ray setup, occupancy march (count + emit) and the occupancy bitfield builder.
//
// Numerics: ray directions, near/far, sample positions, step sizes and sample counts are BIT-EXACT against the
// reference's CUDA build executed on a B200 (tests/golden/ref_cuda_*.npz) and against the oracle in its "cuda" mode.
// The reference is compiled with nvcc's default -fmad=true, so its kernels contract a*b+c into one FFMA wherever the
// compiler chose to; this translation unit is compiled with -fmad=false (nothing is contracted implicitly, IEEE
// div/sqrt) and every fused multiply-add of the reference's sm_100 SASS is written out as an explicit __fmaf_rn:
// set_rays_d (norm, R*v), kernel_march_rays (o + t*d, x*rb + 1, (.)*mip_bound - x), linear_transformer (w*x + b).
//
// Replaces the reference kernels K1-K7 of SURVEY.md §2c:
//   set_rays_o/set_rays_d (include/nerf-cuda/render_utils.h:31-66), kernel_near_far_from_aabb (:338-392),
//   init_step0 (:221-239), kernel_compact_rays (:394-415), kernel_march_rays (:524-655),
//   linear_transformer x2 (include/nerf-cuda/common_device.cuh:31-36), concat_network_in_and_out (:280-306).
// Design: one pass over whole rays instead of a host-driven <=8-step wavefront with a blocking D2H per iteration
// (src/nerf_render.cu:269-338).  Pass 1 counts samples per ray, caches the t of each ray's first NRF_TCACHE samples and
// reduces per block; a one-block scan gives block bases; pass 2 replays the cached samples (marching again only beyond
// them) and writes a COMPACT sample list, step-interleaved inside each warp of 32 neighbouring rays (ballot/popc
// position arithmetic, see k_emit).
#include "nrf_device.cuh"

#ifndef NRF_EXP_MARCH_MAGIC
#define NRF_EXP_MARCH_MAGIC 0
#endif
#define RAYS_BLOCK 128
#define RAYS_MIN_BLOCKS 8
#define EMIT_CHUNK 8
#ifndef COUNT_MIN_BLOCKS
#define COUNT_MIN_BLOCKS 8   // k_rays_count: 47 registers = 10 resident blocks; asking for 10 or 12 measured 3 % / 14 % slower
#endif
#define NRF_MAX_GRID 512  // largest density_grid_size (check_arch)

__device__ __forceinline__ float clampf(float x, float lo, float hi) { return fminf(hi, fmaxf(lo, x)); }

// ---- K1-K5: ray direction + slab test ---------------------------------------------------------------------------
__device__ __forceinline__ void ray_dir(const MarchParams& P, int p, float d[3]) {
  const float i = (float)(p % P.W) + 0.5f;
  const float j = (float)(p / P.W) + 0.5f;
  const float xs = (i - P.intr[2]) / P.intr[0] * 1.0f;
  const float ys = (j - P.intr[3]) / P.intr[1] * 1.0f;
  const float zs = 1.0f;
  // Eigen fixed-size-3 reduction order a0 + (a1 + a2); reference SASS: FFMA(ys,ys,1), FFMA(xs,xs,.) and per row
  // FMUL(vy,R1), FFMA(vz,R2,.), FFMA(vx,R0,.)
  const float nrm = sqrtf(__fmaf_rn(xs, xs, __fmaf_rn(ys, ys, zs * zs)));
  const float v0 = xs / nrm, v1 = ys / nrm, v2 = zs / nrm;
#pragma unroll
  for (int r = 0; r < 3; ++r) d[r] = __fmaf_rn(P.R[r * 3 + 0], v0, __fmaf_rn(P.R[r * 3 + 2], v2, P.R[r * 3 + 1] * v1));
}

__device__ __forceinline__ void near_far(const MarchParams& P, const float d[3], float& near_o, float& far_o) {
  const float rdx = 1 / d[0], rdy = 1 / d[1], rdz = 1 / d[2];
  float near = (P.aabb[0] - P.o[0]) * rdx, far = (P.aabb[3] - P.o[0]) * rdx;
  if (near > far) { float t = near; near = far; far = t; }
  float near_y = (P.aabb[1] - P.o[1]) * rdy, far_y = (P.aabb[4] - P.o[1]) * rdy;
  if (near_y > far_y) { float t = near_y; near_y = far_y; far_y = t; }
  if (near > far_y || near_y > far) { near_o = far_o = 3.402823466e+38f; return; }
  if (near_y > near) near = near_y;
  if (far_y < far) far = far_y;
  float near_z = (P.aabb[2] - P.o[2]) * rdz, far_z = (P.aabb[5] - P.o[2]) * rdz;
  if (near_z > far_z) { float t = near_z; near_z = far_z; far_z = t; }
  if (near > far_z || near_z > far) { near_o = far_o = 3.402823466e+38f; return; }
  if (near_z > near) near = near_z;
  if (far_z < far) far = far_z;
  if (near < P.min_near) near = P.min_near;
  near_o = near;
  far_o = far;
}

// ---- K7: one iteration of the march loop (render_utils.h:596-651) ---------------------------------------------------
struct Step {
  float x, y, z, dt;
  uint32_t cell;
  int level;
};

// Two IEEE divisions of the reference are replaced by values that are bit-identical by construction:
//   1/mip_bound           mip_bound = min(2^level, bound): an exact power of two (reciprocal = exponent flip) or `bound`
//                         itself (P.rbound = 1.0f/bound computed once on the host, IEEE)
//   (n + .5 + .5*sign)/(H-1)   the numerator is the exact integer n or n+1, so the quotient comes from a table
//                         lut[m] = (float)m / (float)(H-1) filled by the host's IEEE division (H+1 entries, shared memory)
//
// FAST variant (P.fast: bound == 2^(C-1), H and 1/dt_gamma powers of two — every BASELINE config): the kernel is bound
// by instruction issue (ncu: issue slots 80 % busy), so every step is trimmed WITHOUT changing a single rounding:
//   * mip_bound is never capped, so it and its reciprocal are powers of two and products by them are exact;
//     RN(a*p2 + c) is then one FFMA instead of FMUL+FADD (cell coordinate, cell-exit plane);
//   * |x| <= mip_bound, so (x*rb+1)*H/2 >= 0 and the lower clamp of the cell coordinate is dead;
//   * the position clamp to [-bound, bound] is the identity unless max|x_i| > bound (rounding at the box faces only):
//     one compare on the max-norm the mip level needs anyway, the six FMNMX only in that rare case;
//   * the table holds fl(2*lut[m] - 1) (same two roundings, done on the host) and is addressed from a per-ray base;
//   * while dt_min <= t*dt_gamma <= dt_max the step t += clamp(t*dt_gamma) is RN(t + t*2^-k) = one FFMA.
// The sample set depends on the whole chain of evaluated points (the reference's cell-exit planes use 1/(H-1) and do
// not coincide with the cell faces, so a step may land one cell further), which rules out any hierarchical skipping:
// the chain itself is reproduced, only cheaper.
__device__ __forceinline__ float lds_f32(uint32_t saddr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(saddr));
  return v;
}

// The march loop is issue-bound and ptxas re-reads kernel parameters from the constant bank on every use (one LDC/LDCU
// issue slot each, ~10 per step).  The per-frame constants the FAST step needs are therefore staged behind the table in
// shared memory (load_lut) and read ONCE per ray with a volatile shared load, which pins them in vector registers.
#define LUT_CONST 16  // o[3], bound, dt_gamma, halfH, hm1, t_lo, t_hi, H, C-1, dt_min, dt_max, occupancy pointer (2), (pad) behind the H+1 table entries
struct RayCtx {
  float d[3], rd[3];
  uint32_t lut_a[3];  // shared-memory byte address of lut[1 - signbit(d_a)]: entry n + 1 (d >= +0) or n (sign set)
  float o[3], bound, dt_gamma, halfH, hm1, t_lo, t_hi, dt_min, dt_max;   // pinned copies of per-frame constants
  uint32_t H;
  uint32_t rbh_base;    // FAST: bits(0.5*H) + bits(1.0): minus bits(2^level) it is bits(0.5*H * 2^-level)
  int cm1;
  const uint32_t* occ;  // P.occupancy, pinned like the constants above
};
__device__ __forceinline__ void ray_ctx(const MarchParams& P, const float* lut, const float d[3], RayCtx& R) {
  const uint32_t base = (uint32_t)__cvta_generic_to_shared(lut);
  const uint32_t cb = base + 4u * (P.H + 1);
  R.o[0] = lds_f32(cb); R.o[1] = lds_f32(cb + 4); R.o[2] = lds_f32(cb + 8);
  R.bound = lds_f32(cb + 12); R.dt_gamma = lds_f32(cb + 16); R.halfH = lds_f32(cb + 20); R.hm1 = lds_f32(cb + 24);
  R.t_lo = lds_f32(cb + 28); R.t_hi = lds_f32(cb + 32);
  R.H = __float_as_uint(lds_f32(cb + 36)); R.cm1 = (int)__float_as_uint(lds_f32(cb + 40));
  R.dt_min = lds_f32(cb + 44); R.dt_max = lds_f32(cb + 48);
  R.rbh_base = __float_as_uint(R.halfH) + 0x3f800000u;
  R.occ = reinterpret_cast<const uint32_t*>(((unsigned long long)__float_as_uint(lds_f32(cb + 56)) << 32) |
                                            __float_as_uint(lds_f32(cb + 52)));
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    R.d[a] = d[a];
    R.rd[a] = 1 / d[a];
    // n + 0.5f + 0.5f*copysignf(1, d) is the integer n+1 (sign bit clear, incl. +0) or n (sign bit set, incl. -0)
    R.lut_a[a] = base + 4u * (1u - (__float_as_uint(d[a]) >> 31));
  }
}

// rare path of the FAST step (a coordinate left [-bound, bound] by rounding at a box face): out of line and by value
// (registers in, registers out), so that the common path does not issue seven predicated-off instructions per step
__device__ __noinline__ float4 clamp_to_box(float bound, float x, float y, float z) {
  x = clampf(x, -bound, bound); y = clampf(y, -bound, bound); z = clampf(z, -bound, bound);
  return make_float4(x, y, z, fmaxf(fabsf(x), fmaxf(fabsf(y), fabsf(z))));
}

template <bool FAST>
__device__ __forceinline__ bool march_step(const MarchParams& P, const RayCtx& R, float& t, Step& s) {
  // reference SASS: FFMA(d, t, o)
  float x = __fmaf_rn(t, R.d[0], R.o[0]), y = __fmaf_rn(t, R.d[1], R.o[1]), z = __fmaf_rn(t, R.d[2], R.o[2]);
  float mx = fmaxf(fabsf(x), fmaxf(fabsf(y), fabsf(z)));
  if (FAST) {
    if (mx > R.bound) { const float4 c = clamp_to_box(R.bound, x, y, z); x = c.x; y = c.y; z = c.z; mx = c.w; }
  } else {
    x = clampf(x, -R.bound, R.bound); y = clampf(y, -R.bound, R.bound); z = clampf(z, -R.bound, R.bound);
    mx = fmaxf(fabsf(x), fmaxf(fabsf(y), fabsf(z)));
  }
  // mip_from_pos: frexpf exponent of the max-norm, clamped to [0, C-1]  (exact, done on the bit pattern)
  const int e = (int)(__float_as_uint(mx) >> 23) - 126;  // mx >= 0: the sign bit is clear, no mask needed
  const int level = max(0, min(FAST ? R.cm1 : (int)P.C - 1, e));
  const float pow2 = __uint_as_float((uint32_t)(127 + level) << 23);
  const bool capped = !FAST && pow2 > P.bound;
  const float mip_bound = capped ? P.bound : pow2;
  // 2^-level from 2^level: the two bit patterns add up to 0x7f000000 — one IADD; kept opaque, or the compiler turns it
  // back into a constant load + IMAD
  uint32_t rb_bits;
  asm("sub.u32 %0, 0x7f000000, %1;" : "=r"(rb_bits) : "r"(__float_as_uint(pow2)));
  const float mip_rbound = capped ? P.rbound : __uint_as_float(rb_bits);
  const uint32_t H = FAST ? R.H : P.H;
  // reference: (int)clamp(0.5 /*double*/ * (x*rb+1) * H, 0, H-1).  0.5*H is exact in float and the double product is
  // exact, so one float multiply by (0.5f*H) rounds identically (DESIGN.md numerics contract).
  const float halfH = R.halfH, hm1 = R.hm1;
  int nx, ny, nz;
  if (FAST) {
#if NRF_EXP_MARCH_MAGIC
    // experiment: the three F2I.TRUNC (quarter-rate XU pipe) as FADD.RZ with 2^23 — the operands lie in [0, H-1], so the
    // truncated integer is the low mantissa of the sum
    nx = (int)(__float_as_uint(__fadd_rz(fminf(__fmaf_rn(x, mip_rbound, 1.0f) * halfH, hm1), 8388608.f)) & 0x7fffffu);
    ny = (int)(__float_as_uint(__fadd_rz(fminf(__fmaf_rn(y, mip_rbound, 1.0f) * halfH, hm1), 8388608.f)) & 0x7fffffu);
    nz = (int)(__float_as_uint(__fadd_rz(fminf(__fmaf_rn(z, mip_rbound, 1.0f) * halfH, hm1), 8388608.f)) & 0x7fffffu);
#else
    // RN(x*rb + 1) * (0.5*H) == RN(x*(rb*0.5*H) + 0.5*H): rb and 0.5*H are powers of two (FAST requires it), and scaling by a
    // power of two commutes with rounding (no overflow / subnormals: the values lie in [0, H]) — one FFMA instead of FFMA+FMUL
    uint32_t rbh_bits;
    asm("sub.u32 %0, %1, %2;" : "=r"(rbh_bits) : "r"(R.rbh_base), "r"(__float_as_uint(pow2)));
    const float rbh = __uint_as_float(rbh_bits);
    nx = (int)fminf(__fmaf_rn(x, rbh, halfH), hm1);
    ny = (int)fminf(__fmaf_rn(y, rbh, halfH), hm1);
    nz = (int)fminf(__fmaf_rn(z, rbh, halfH), hm1);
#endif
  } else {
    nx = (int)clampf(__fmaf_rn(x, mip_rbound, 1.0f) * halfH, 0.0f, hm1);   // reference SASS: FFMA(x, rb, 1), DMUL x2
    ny = (int)clampf(__fmaf_rn(y, mip_rbound, 1.0f) * halfH, 0.0f, hm1);
    nz = (int)clampf(__fmaf_rn(z, mip_rbound, 1.0f) * halfH, 0.0f, hm1);
  }
  const uint32_t cell = ((level * H + nx) * H + ny) * H + nz;
  uint32_t occ_u;  // ((word >> (cell & 31)) & 1) != 0, kept opaque: SHF + predicate-writing LOP3 (2 issue slots, not 3)
  asm("{ .reg .pred p; .reg .b32 m; shr.b32 m, %1, %2; and.b32 m, m, 1; setp.ne.u32 p, m, 0; selp.u32 %0, 1, 0, p; }"
      : "=r"(occ_u) : "r"(__ldg(R.occ + (cell >> 5))), "r"(cell & 31u));
  const bool occ = occ_u != 0u;
  if (occ) {
    s.x = x; s.y = y; s.z = z;
    const float dt = clampf(t * R.dt_gamma, R.dt_min, R.dt_max);
    t += dt;
    s.dt = dt;
    s.cell = cell;
    s.level = level;
    return true;
  }
  const float qx = lds_f32(R.lut_a[0] + 4u * nx), qy = lds_f32(R.lut_a[1] + 4u * ny), qz = lds_f32(R.lut_a[2] + 4u * nz);
  float tx, ty, tz;
  if (FAST) {  // table entry = fl(2q - 1); (.)*mip_bound is exact, so RN((.)*mip_bound - x) is one FFMA
    tx = __fmaf_rn(qx, mip_bound, -x) * R.rd[0];
    ty = __fmaf_rn(qy, mip_bound, -y) * R.rd[1];
    tz = __fmaf_rn(qz, mip_bound, -z) * R.rd[2];
  } else {
    tx = __fmaf_rn(qx * 2.0f - 1.0f, mip_bound, -x) * R.rd[0];   // reference SASS: FFMA(q,2,-1) [exact], FFMA(.,mip_bound,-x)
    ty = __fmaf_rn(qy * 2.0f - 1.0f, mip_bound, -y) * R.rd[1];
    tz = __fmaf_rn(qz * 2.0f - 1.0f, mip_bound, -z) * R.rd[2];
  }
  const float tt = t + fmaxf(0.0f, fminf(tx, fminf(ty, tz)));
  // tt >= t (the increment is max(0, .), and max(0, NaN) = 0), so every t the loop multiplies lies in [t, tt]
  if (FAST && t >= R.t_lo && tt <= R.t_hi) {
    do { t = __fmaf_rn(t, R.dt_gamma, t); } while (t < tt);   // dt_min <= t*dt_gamma <= dt_max on every step taken
  } else {
    do {
      const float dt = clampf(t * R.dt_gamma, R.dt_min, R.dt_max);
      t += dt;
    } while (t < tt);
  }
  return false;
}

// lut (shared): FAST -> fl(2*m/(H-1) - 1), else m/(H-1); both tables sit back to back in P.div_lut
__device__ __forceinline__ void load_lut(const MarchParams& P, float* lut) {
  const float* src = P.div_lut + (P.fast ? P.H + 1 : 0);
  for (uint32_t i = threadIdx.x; i <= P.H; i += blockDim.x) lut[i] = __ldg(src + i);
  if (threadIdx.x == 0) {
    float* c = lut + P.H + 1;
    c[0] = P.o[0]; c[1] = P.o[1]; c[2] = P.o[2]; c[3] = P.bound; c[4] = P.dt_gamma;
    c[5] = 0.5f * (float)P.H; c[6] = (float)(P.H - 1); c[7] = P.t_lo; c[8] = P.t_hi;
    c[9] = __uint_as_float(P.H); c[10] = __uint_as_float(P.C - 1u); c[11] = P.dt_min; c[12] = P.dt_max;
    const unsigned long long op = (unsigned long long)P.occupancy;
    c[13] = __uint_as_float((uint32_t)op); c[14] = __uint_as_float((uint32_t)(op >> 32));
  }
  __syncthreads();
}

// Whole-ray count; t_first = the t at which the first sample was taken (pass 2 restarts there: the march state is t)
// tc (may be null): this ray's column of the sample-t cache — the t of its first NRF_TCACHE samples, entry k at
// tc[32*k] (patch-interleaved like the sample list), so that pass 2 replays the samples instead of marching the
// empty space between them a second time.
template <bool FAST>
__device__ __forceinline__ int count_ray_t(const MarchParams& P, const RayCtx& R, float near, float far, float& t_first,
                                           float* __restrict__ tc) {
  float t = near;
  int c = 0;
  t_first = near;
  Step s;
  while (t < far) {  // the sample cap is tested where c changes, not on every empty step
    const float t0 = t;
    if (march_step<FAST>(P, R, t, s)) {
      if (c == 0) t_first = t0;
      if (tc && c < NRF_TCACHE) tc[32 * c] = t0;
      if (++c >= NRF_MAX_STEPS_D) break;
    }
  }
  return c;
}
__device__ __forceinline__ int count_ray(const MarchParams& P, const float* lut, const float d[3], float near, float far,
                                         float& t_first, float* __restrict__ tc = nullptr) {
  RayCtx R;
  ray_ctx(P, lut, d, R);
  return P.fast ? count_ray_t<true>(P, R, near, far, t_first, tc) : count_ray_t<false>(P, R, near, far, t_first, tc);
}

// block-wide sum / exclusive scan helpers (128 threads, warp shuffles)
__device__ __forceinline__ int warp_incl_scan(int v) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int n = __shfl_up_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) >= o) v += n;
  }
  return v;
}
__device__ __forceinline__ void block_total(int c, int* __restrict__ block_sums) {
  __shared__ int wsum[RAYS_BLOCK / 32];
  const int incl = warp_incl_scan(c);
  if ((threadIdx.x & 31) == 31) wsum[threadIdx.x >> 5] = incl;
  __syncthreads();
  if (threadIdx.x == 0) {
    int s = 0;
#pragma unroll
    for (int w = 0; w < RAYS_BLOCK / 32; ++w) s += wsum[w];
    block_sums[blockIdx.x] = s;
  }
}

// Pass 1: rays of this shard (patch-major order, ShardMap) + sample counts + per-block totals.
// rays: float4 {dx,dy,dz,near}; far_tf: float2 {far, t_first}; counts; block_sums[gridDim.x]
// s1: samples per ray the FIRST evaluation round takes (NRF_MAX_STEPS_D = the whole ray; termination-aware frames take
// fewer, see k_round_sums); block_sums then holds the round's totals.  frame_tot[0] += all samples found (one atomic
// per block), so the host learns how many samples the march produced whatever the round schedule was.
extern "C" __global__ void __launch_bounds__(RAYS_BLOCK, COUNT_MIN_BLOCKS)
k_rays_count(const __grid_constant__ MarchParams P, ShardMap map, int n_local, float4* __restrict__ rays,
             float2* __restrict__ far_tf, int* __restrict__ counts, int* __restrict__ block_sums,
             uint4* __restrict__ ray_sh, float* __restrict__ tcache, int s1, unsigned long long* __restrict__ frame_tot) {
  __shared__ float lut[NRF_MAX_GRID + 1 + LUT_CONST];
  load_lut(P, lut);
  const int q = blockIdx.x * RAYS_BLOCK + threadIdx.x;
  float* tc = tcache ? tcache + (size_t)(q >> 5) * (32 * NRF_TCACHE) + (q & 31) : nullptr;
  int c = 0;
  if (q < n_local) {
    const int p = shard_pixel(map, q);
    float d[3] = {0.f, 0.f, 1.f}, near = 0.f, far = 0.f, tf = 0.f;
    if (p >= 0) {
      ray_dir(P, p, d);
      near_far(P, d, near, far);
      c = count_ray(P, lut, d, near, far, tf, tc);
    }
    rays[q] = make_float4(d[0], d[1], d[2], near);
    far_tf[q] = make_float2(far, tf);
    counts[q] = c;
    if (c > 0) {  // SH of the normalised direction d*0.5+0.5 (nerf_render.cu:313-314), once per ray
      uint32_t sh[8];
      sh4(0.5f * d[0] + 0.5f, 0.5f * d[1] + 0.5f, 0.5f * d[2] + 0.5f, sh);
      ray_sh[2 * (size_t)q] = make_uint4(sh[0], sh[1], sh[2], sh[3]);
      ray_sh[2 * (size_t)q + 1] = make_uint4(sh[4], sh[5], sh[6], sh[7]);
    }
  }
  block_total(min(c, s1), block_sums);
  {
    __shared__ int marched[RAYS_BLOCK / 32];
    const int wsum_c = __reduce_add_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0) marched[threadIdx.x >> 5] = wsum_c;
    __syncthreads();
    if (threadIdx.x == 0 && frame_tot) {
      int m = 0;
#pragma unroll
      for (int w = 0; w < RAYS_BLOCK / 32; ++w) m += marched[w];
      if (m) atomicAdd(frame_tot, (unsigned long long)m);
    }
  }
}

// Later evaluation rounds of a termination-aware frame: per-block totals of the samples with step index in [s0, s1) of
// the rays that are still alive (dead[q] set by k_composite when a ray's transmittance fell below 1e-4).
extern "C" __global__ void __launch_bounds__(RAYS_BLOCK)
k_round_sums(int n, const int* __restrict__ counts, const uint8_t* __restrict__ dead, int s0, int s1,
             int* __restrict__ block_sums) {
  const int q = blockIdx.x * RAYS_BLOCK + threadIdx.x;
  int c = 0;
  if (q < n && !dead[q]) c = max(0, min(counts[q], s1) - s0);
  block_total(c, block_sums);
}

// Rays only, for pixels [pix0, pix0+n) in plain pixel order (stage probe nrf_probe_rays).
extern "C" __global__ void __launch_bounds__(RAYS_BLOCK)
k_rays_given_pixels(const __grid_constant__ MarchParams P, int pix0, int n, float4* __restrict__ rays,
                    float* __restrict__ fars) {
  const int q = blockIdx.x * RAYS_BLOCK + threadIdx.x;
  if (q >= n) return;
  float d[3], near, far;
  ray_dir(P, pix0 + q, d);
  near_far(P, d, near, far);
  rays[q] = make_float4(d[0], d[1], d[2], near);
  fars[q] = far;
}

// Count-only variant for externally supplied rays (stage probe nrf_probe_march).
extern "C" __global__ void __launch_bounds__(RAYS_BLOCK, RAYS_MIN_BLOCKS)
k_count_given(const __grid_constant__ MarchParams P, int n, const float4* __restrict__ rays,
              float2* __restrict__ far_tf, int* __restrict__ counts, int* __restrict__ block_sums) {
  __shared__ float lut[NRF_MAX_GRID + 1 + LUT_CONST];
  load_lut(P, lut);
  const int q = blockIdx.x * RAYS_BLOCK + threadIdx.x;
  int c = 0;
  if (q < n) {
    const float4 r = rays[q];
    const float d[3] = {r.x, r.y, r.z};
    float tf;
    c = count_ray(P, lut, d, r.w, far_tf[q].x, tf);
    far_tf[q].y = tf;
    counts[q] = c;
  }
  block_total(c, block_sums);
}

// One-block exclusive scan of the per-block totals (in place).  n_blocks <= 1024*64.  The running prefix is kept in 64
// bits and the stored block bases saturate at INT_MAX, so a frame with more than 2^31 samples cannot wrap to a small
// positive base: its patches fail k_emit's capacity test (capacity < 2^31) and the frame is reported as overflowed.
// total[] (TOT_* in nrf_api.cu): [0] samples of this round; [1] running maximum of [0] since the host last looked
// (buffer growth: frames are enqueued without host syncs); [2] rounds that did not fit `capacity`, cumulative;
// [4] += samples of this round (evaluated samples of the frame); [6] = 1 if a round of this frame did not fit.
extern "C" __global__ void __launch_bounds__(1024) k_scan_blocks(int* __restrict__ block_sums, int n_blocks,
                                                                 unsigned long long* __restrict__ total,
                                                                 unsigned long long capacity) {
  __shared__ int wsum[32];
  __shared__ long long carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (int base = 0; base < n_blocks; base += 1024) {
    const int i = base + threadIdx.x;
    const int v = (i < n_blocks) ? block_sums[i] : 0;
    const int incl = warp_incl_scan(v);   // <= 1024 blocks x 131072 samples = 2^27: no overflow inside one chunk
    if ((threadIdx.x & 31) == 31) wsum[threadIdx.x >> 5] = incl;
    __syncthreads();
    if (threadIdx.x < 32) {
      int w = wsum[threadIdx.x];
      int wi = warp_incl_scan(w);
      wsum[threadIdx.x] = wi - w;  // exclusive warp bases
    }
    __syncthreads();
    const long long excl = carry_s + wsum[threadIdx.x >> 5] + incl - v;
    if (i < n_blocks) block_sums[i] = (int)min(excl, (long long)0x7fffffff);
    __syncthreads();
    if (threadIdx.x == 1023) carry_s = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const unsigned long long s = (unsigned long long)carry_s;
    total[0] = s;
    if (s > total[1]) total[1] = s;
    total[4] += s;
    if (s > capacity) { total[2] += 1ull; total[6] = 1ull; }
  }
}

// Pass 2: replay the t-cache (re-march only beyond it; the probe path re-marches from t_first) and emit the compact
// sample list in WARP-INTERLEAVED order: the samples of one warp
// (32 neighbouring rays, an 8x4 pixel patch) are stored step-major — step k of every ray that has a k-th sample, in
// lane order, then step k+1 ... — so 32 consecutive samples are spatial neighbours (one thin slab of the patch), which
// is what makes the hash-grid gather cache-coherent.  Position of (lane r, step k):
//     warp_base + sum_{r'} min(c_r', k) + #{r' < r : c_r' > k}
// evaluated incrementally per lane from a per-warp table died[k] = {r' : c_r' == k+1} in shared memory, so lanes do not
// march in lockstep.  warp_base[] (first sample of every warp) is kept for the composite kernel.
// Frame path (t_rec != null): samples float4 {u,v,w, ray id} + t_rec = the sample's t (SampleSrc, nrf_device.cuh).
// Probe path: samples float4 {u,v,w,dt} (position already normalised: x*pos_w+0.5, nerf_render.cu:311-312), t_after, ray id.
// dbg (optional, probe only): per sample {x,y,z,dt,t_after,level,cell,ray} raw, RAY-major at dbg_offsets[q] + k.
// Evaluation rounds (termination-aware frames): only the samples with step index in [s0, s1) of rays that are still
// alive are emitted (dead may be null = all alive); the list, warp_base and the block bases are those of THIS round.
// A ray with samples beyond s1 leaves its march state (t after its last emitted sample) in t_resume for the next round.
extern "C" __global__ void __launch_bounds__(RAYS_BLOCK, RAYS_MIN_BLOCKS)
k_emit(const __grid_constant__ MarchParams P, int n, const float4* __restrict__ rays, const float2* __restrict__ far_tf,
       const int* __restrict__ counts, const int* __restrict__ block_bases, int* __restrict__ warp_base,
       float4* __restrict__ samples, float* __restrict__ t_after, int* __restrict__ sample_ray,
       float* __restrict__ dbg, const long long* __restrict__ dbg_offsets, long long capacity,
       const float* __restrict__ tcache, float* __restrict__ t_rec, int s0, int s1,
       const uint8_t* __restrict__ dead, float* __restrict__ t_resume) {
  __shared__ float lut[NRF_MAX_GRID + 1 + LUT_CONST];
  __shared__ int wsum[RAYS_BLOCK / 32];
  __shared__ uint32_t died[RAYS_BLOCK / 32][NRF_MAX_STEPS_D];
  load_lut(P, lut);
  const int q = blockIdx.x * RAYS_BLOCK + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c_all = (q < n) ? counts[q] : 0;
  const int c_total = (q < n && !(dead && dead[q])) ? max(0, min(c_all, s1) - s0) : 0;   // this round's samples of the ray
  const int incl = warp_incl_scan(c_total);
  if (lane == 31) wsum[warp] = incl;
  // per-warp table of the lanes that run out of samples after step k
  int cmax = c_total;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) cmax = max(cmax, __shfl_xor_sync(0xffffffffu, cmax, o));
  for (int k = lane; k < cmax; k += 32) died[warp][k] = 0u;
  __syncwarp();
  if (c_total > 0) atomicOr(&died[warp][c_total - 1], 1u << lane);
  __syncthreads();
  long long base = block_bases[blockIdx.x];
  for (int w = 0; w < warp; ++w) base += wsum[w];
  // a patch whose samples do not fit the buffers is skipped and marked (warp_base = -1: k_composite skips it too); the
  // host sees total > capacity at the end of the frame, grows the buffers and renders the frame again
  const bool overflow = base + wsum[warp] > capacity;
  if (lane == 0 && q < n) warp_base[q >> 5] = overflow ? -1 : (int)base;
  uint32_t alive = __ballot_sync(0xffffffffu, c_total > 0);
  if (c_total == 0 || overflow) return;
  const uint32_t lt = (1u << lane) - 1u;
  const float4 r = rays[q];
  const float d[3] = {r.x, r.y, r.z};
  RayCtx R;
  ray_ctx(P, lut, d, R);
  const float2 ft = far_tf[q];
  const float far = ft.x;
  float t = s0 == 0 ? ft.y : t_resume[q];
  int c = 0;
  long long off = base;
  const long long dbg0 = dbg ? dbg_offsets[q] : 0;
  Step s;
  const bool fast = P.fast;
  if (tcache && s0 < NRF_TCACHE) {
    // replay the cached samples: position, dt and t_after are the occupied branch of march_step evaluated at the
    // cached t (the clamp to the box is the identity inside it, so both march variants give these same values)
    const float* tc = tcache + (size_t)(q >> 5) * (32 * NRF_TCACHE) + lane + 32 * s0;
    const int nrep = min(c_total, NRF_TCACHE - s0);
    // chunks of EMIT_CHUNK samples: all their cached t in flight at once (one L2 / HBM latency per chunk, not per sample)
    while (c < nrep) {
      float tq[EMIT_CHUNK];
#pragma unroll
      for (int j = 0; j < EMIT_CHUNK; ++j) tq[j] = (c + j < nrep) ? __ldg(tc + 32 * (c + j)) : 0.f;
#pragma unroll
      for (int j = 0; j < EMIT_CHUNK; ++j) {
        if (c < nrep) {
          const float t0 = tq[j];
          const float x = clampf(__fmaf_rn(t0, R.d[0], R.o[0]), -R.bound, R.bound);
          const float y = clampf(__fmaf_rn(t0, R.d[1], R.o[1]), -R.bound, R.bound);
          const float z = clampf(__fmaf_rn(t0, R.d[2], R.o[2]), -R.bound, R.bound);
          const float dt = clampf(t0 * R.dt_gamma, R.dt_min, R.dt_max);
          t = t0 + dt;
          const long long k = off + __popc(alive & lt);
          if (t_rec) {   // frame path (SampleSrc): ray index in the fourth component, the sample's t beside it
            samples[k] = make_float4(__fmaf_rn(P.pos_w, x, 0.5f), __fmaf_rn(P.pos_w, y, 0.5f), __fmaf_rn(P.pos_w, z, 0.5f), __int_as_float(q));
            t_rec[k] = t0;
          } else {
            samples[k] = make_float4(__fmaf_rn(P.pos_w, x, 0.5f), __fmaf_rn(P.pos_w, y, 0.5f), __fmaf_rn(P.pos_w, z, 0.5f), dt);
            t_after[k] = t;
            sample_ray[k] = q;
          }
          off += __popc(alive);
          alive &= ~died[warp][c];
          ++c;
        }
      }
    }
  }
  while (t < far && c < c_total) {
    const float t0 = t;
    if (!(fast ? march_step<true>(P, R, t, s) : march_step<false>(P, R, t, s))) continue;
    const long long k = off + __popc(alive & lt);
    if (t_rec) {
      samples[k] = make_float4(__fmaf_rn(P.pos_w, s.x, 0.5f), __fmaf_rn(P.pos_w, s.y, 0.5f), __fmaf_rn(P.pos_w, s.z, 0.5f), __int_as_float(q));
      t_rec[k] = t0;
    } else {
      samples[k] = make_float4(__fmaf_rn(P.pos_w, s.x, 0.5f), __fmaf_rn(P.pos_w, s.y, 0.5f), __fmaf_rn(P.pos_w, s.z, 0.5f), s.dt);
      t_after[k] = t;
      sample_ray[k] = q;
    }
    if (dbg) {
      float* o = dbg + (dbg0 + c) * 8;
      o[0] = s.x; o[1] = s.y; o[2] = s.z; o[3] = s.dt; o[4] = t;
      o[5] = __int_as_float(s.level);
      o[6] = __uint_as_float(s.cell);
      o[7] = __int_as_float(q);
    }
    off += __popc(alive);
    alive &= ~died[warp][c];
    ++c;
  }
  if (t_resume && c_all > s1) t_resume[q] = t;   // march state for the next round: t after sample s1-1
}

// Load-time: float density grid -> occupancy bitfield, bit == (density > thresh)   (render_utils.h:560,619)
extern "C" __global__ void k_build_occupancy(const float* __restrict__ grid, uint32_t n_cells, float thresh,
                                             uint32_t* __restrict__ bits) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool occ = (i < n_cells) && (grid[i] > thresh);
  const uint32_t m = __ballot_sync(0xffffffffu, occ);
  if ((threadIdx.x & 31) == 0 && i < n_cells) bits[i >> 5] = m;
}

// Load-time: float32 params -> fp16 (nerf_network.h:434-436)
extern "C" __global__ void k_cast_half(const float* __restrict__ in, __half* __restrict__ out, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = __float2half_rn(in[i]);
}

// ---- density-grid regeneration (NerfRender::generate_density_grid, src/nerf_render.cu:388-429) ---------------------
// pcg32 (tcnn default_rng_t, random.h:36-65): LCG 6364136223846793005 with stream constant inc, XSH-RR output,
// O(log n) skip-ahead.  next_float() = bits (u >> 9) | 0x3f800000 minus 1.
struct Pcg32 {
  unsigned long long state, inc;
  __device__ Pcg32(unsigned long long initstate, unsigned long long initseq = 1ull) {
    state = 0ull; inc = (initseq << 1u) | 1ull;
    next_uint(); state += initstate; next_uint();
  }
  __device__ uint32_t next_uint() {
    const unsigned long long old = state;
    state = old * 0x5851f42d4c957f2dull + inc;
    const uint32_t xorshifted = (uint32_t)(((old >> 18u) ^ old) >> 27u), rot = (uint32_t)(old >> 59u);
    return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31u));
  }
  __device__ float next_float() { return __uint_as_float((next_uint() >> 9) | 0x3f800000u) - 1.0f; }
  __device__ void advance(unsigned long long delta) {
    unsigned long long cur_mult = 0x5851f42d4c957f2dull, cur_plus = inc, acc_mult = 1ull, acc_plus = 0ull;
    while (delta > 0) {
      if (delta & 1ull) { acc_mult *= cur_mult; acc_plus = acc_plus * cur_mult + cur_plus; }
      cur_plus = (cur_mult + 1ull) * cur_plus;
      cur_mult *= cur_mult;
      delta >>= 1;
    }
    state = acc_mult * state + acc_plus;
  }
};

// Sample positions of one cascade: init_xyzs (render_utils.h:88-105) -> dd_scale by (bound_c - half) (:79-86,
// nerf_render.cu:411) -> add_random with k = half (:107-117, nerf_render.cu:412; the generator is advanced to the
// element index, tcnn style, instead of handing every thread the same state) -> normalised to [0,1] like every network
// input (nerf_render.cu:311-312).  Cell order == the march's cell index ((x*H + y)*H + z, render_utils.h:613-616).
extern "C" __global__ void k_dg_positions(uint32_t H, float bound_c, float half, float pos_w, unsigned long long seed,
                                          int jitter, unsigned long long elem0, float4* __restrict__ samples) {
  const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= H * H * H) return;
  const uint32_t id[3] = {idx / (H * H), idx % (H * H) / H, idx % H};
  float u[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    float p = -1.f + 2.f / (H - 1) * id[a];
    p = (bound_c - half) * p + 0.f;
    if (jitter) {
      Pcg32 rng(seed);
      rng.advance(elem0 + 3ull * idx + a);
      p += half * (2.f * rng.next_float() - 1.f) + 0.f;
    }
    u[a] = pos_w * p + 0.5f;
  }
  samples[idx] = make_float4(u[0], u[1], u[2], 0.f);
}

// dd_scale by 0.001691 (nerf_render.cu:417) + dg_update (render_utils.h:119-128): EMA-max of the decayed grid and the
// freshly evaluated density (sigma = 4th half of the network output).  Also accumulates the sum for mean_density.
extern "C" __global__ void k_dg_update(uint32_t n, const uint2* __restrict__ sample_out, float decay, int reset,
                                       float* __restrict__ grid, double* __restrict__ sum) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  float v = 0.f;
  if (i < n) {
    const float sigma = __half2float(__ushort_as_half((unsigned short)(sample_out[i].y >> 16)));
    const float tmp = 0.001691f * sigma + 0.f;
    const float dg = reset ? 1.0f / 64 : grid[i];
    v = dg;
    if (dg >= 0) v = (dg * decay > tmp) ? dg * decay : tmp;
    grid[i] = v;
  }
  // block sum in double (order-independent enough for a threshold that is clamped to 0.01 anyway)
  double d = (double)v;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
  __shared__ double ws[32];
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = d;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0;
    for (uint32_t w = 0; w < (blockDim.x + 31) / 32; ++w) t += ws[w];
    atomicAdd(sum, t);
  }
}
