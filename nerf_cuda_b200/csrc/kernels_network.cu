// kernels_network.cu — multiresolution hash-grid gather + fused density/colour MLP on tcgen05 (sm_100a).
//
// Replaces, per sample, the reference's K10-K15:
//   tcnn::kernel_grid<__half,3,2>            dependencies/tiny-cuda-nn/include/tiny-cuda-nn/encodings/grid.h:139-268
//   kernel_mlp_fused (density 32->64->16)    dependencies/tiny-cuda-nn/src/fully_fused_mlp.cu:500-558
//   tcnn::kernel_sh<__half> (degree 4)       .../encodings/spherical_harmonics.h:46-152
//   kernel_mlp_fused (rgb 32->64->64->16)    fully_fused_mlp.cu:500-558
//   extract_density                          include/nerf-cuda/nerf_network.h:49-61,187-195
//   decompose_network_in_and_out             include/nerf-cuda/render_utils.h:308-334
// and removes every global-memory round trip between them: the encoded features go from registers straight into the
// shared-memory A operand of the first UMMA, each layer's accumulator lives in TMEM, the epilogue (ReLU -> fp16) writes
// the next layer's A operand back to shared memory, and only {r,g,b,sigma} (4 halves) per sample reach HBM.
//
// One CTA = 128 threads = one 128-sample tile at a time (UMMA M = 128); thread t owns sample row t end to end
// (gather -> A tile row t -> TMEM lane t).  All weights (20 KB, pre-packed on the host into the K-major SWIZZLE_NONE
// core-matrix order) stay resident in shared memory for the lifetime of the persistent CTA.
//
// Numerics: hash indices and fp16 interpolation follow the reference bit-exactly (this TU is built with -fmad=false);
// the MLP accumulates in fp32 (TMEM) and rounds to fp16 at each layer boundary like the oracle.
#include "nrf_device.cuh"

#define NET_THREADS 128
#define SMEM_W_BYTES 20480
#define SMEM_A_BYTES 16384
// byte offsets of the packed weight matrices inside the 20 KB block
#define W0_OFF 0       // [64 x 32]
#define W1_OFF 4096    // [16 x 64]
#define V0_OFF 6144    // [64 x 32]
#define V1_OFF 10240   // [64 x 64]
#define V2_OFF 18432   // [16 x 64]

// ------------------------------------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "DONE:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// K-major, SWIZZLE_NONE shared-memory matrix descriptor (cute/arch/mma_sm100_desc.hpp SmemDescriptor):
// start>>4 [0,14) | LBO>>4 [16,30) (stride between the two 8-element K chunks) | SBO>>4 [32,46) (stride between
// 8-row groups) | version=1 [46,48) | layout_type=0 [61,64).
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo_bytes >> 4) << 16) | ((uint64_t)(sbo_bytes >> 4) << 32) |
         (1ull << 46);
}
// instruction descriptor, kind::f16: D=f32 (bit 4), A=B=f16 (0), both K-major, N>>3 at [17,23), M>>4 at [24,29)
__host__ __device__ constexpr uint32_t umma_idesc(int M, int N) {
  return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// One layer: D[128 x N] = A[128 x K] * W[N x K]^T, A tile at a_addr (LBO 2048, SBO 128), W at b_addr (LBO N*16, SBO 128)
template <int N, int K>
__device__ __forceinline__ void issue_layer(uint32_t tmem_d, uint32_t a_addr, uint32_t b_addr) {
  constexpr uint32_t idesc = umma_idesc(NRF_TILE, N);
#pragma unroll
  for (int kk = 0; kk < K / 16; ++kk) {
    const uint64_t ad = umma_desc(a_addr + kk * 2 * (NRF_TILE * 16), NRF_TILE * 16, 128);
    const uint64_t bd = umma_desc(b_addr + kk * 2 * (N * 16), N * 16, 128);
    umma_f16(tmem_d, ad, bd, idesc, kk > 0);
  }
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ------------------------------------------------------------------------------------------------ numerics helpers
__device__ __forceinline__ uint32_t pack_h2(__half a, __half b) {
  return (uint32_t)__half_as_ushort(a) | ((uint32_t)__half_as_ushort(b) << 16);
}
// tcnn ReLU in half: x * (x > 0)  (common_device.h:71-76); result bits: x for x>0, else (+/-)0
__device__ __forceinline__ __half relu_h(float acc) {
  const __half h = __float2half_rn(acc);
  return __hgt(h, __float2half_rn(0.f)) ? h : __hmul(h, __float2half_rn(0.f));
}
__device__ __forceinline__ __half act_h(uint32_t act, __half v) {
  switch (act) {
    case ACT_RELU: return __hgt(v, __float2half_rn(0.f)) ? v : __float2half_rn(0.f);
    case ACT_EXP: return __float2half_rn(expf(__half2float(v)));
    case ACT_SIGMOID: return __float2half_rn(1.0f / (1.0f + expf(-__half2float(v))));
    default: return v;
  }
}

// One level of tcnn kernel_grid (grid.h:186-268) + grid_index/fast_hash (grid.h:81-117): trilinear blend accumulated
// in fp16 in corner order 0..7 (corner bit d selects g[d]+1 and weight w[d], else 1-w[d]; weight = (f0*f1)*f2).
// Index math: hashed levels  (x ^ y*2654435761 ^ z*805459861) & (T-1)   (T is a power of two);
//             dense levels   x + y*res + z*res^2, which is < 2T, so one conditional subtract equals the reference's % T.
// The two x-corners of a cell are adjacent entries whenever they fall in the same aligned 8-byte slot (always when x is
// even on hashed levels — the hash prime for x is 1, grid.h:88 — and for even base indices on dense levels): one
// 8-byte load then serves both corners and costs the L1 a single sector request.  Values and order are unchanged.
__device__ __forceinline__ uint32_t encode_level(const LevelInfo& lv, const __half2* __restrict__ table, float u0,
                                                 float u1, float u2, uint32_t* idx_out) {
  const float p0 = u0 * lv.scale + 0.5f, p1 = u1 * lv.scale + 0.5f, p2 = u2 * lv.scale + 0.5f;  // -fmad=false
  const float f0 = floorf(p0), f1 = floorf(p1), f2 = floorf(p2);
  const uint32_t g0 = (uint32_t)(int)f0, g1 = (uint32_t)(int)f1, g2 = (uint32_t)(int)f2;
  const float w0 = p0 - f0, w1 = p1 - f1, w2 = p2 - f2;
  const float o0 = 1.0f - w0, o1 = 1.0f - w1, o2 = 1.0f - w2;
  const float wxy[4] = {o0 * o1, w0 * o1, o0 * w1, w0 * w1};  // (f0 * f1), corner bits 0,1
  const uint32_t* tab32 = reinterpret_cast<const uint32_t*>(table + lv.offset);
  uint32_t e[8];
  if (lv.hashed) {
    const uint32_t m = lv.size - 1u;
    const uint32_t hy0 = g1 * 2654435761u, hy1 = hy0 + 2654435761u, hz0 = g2 * 805459861u, hz1 = hz0 + 805459861u;
    const uint32_t yz[4] = {hy0 ^ hz0, hy1 ^ hz0, hy0 ^ hz1, hy1 ^ hz1};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      e[2 * j] = (g0 ^ yz[j]) & m;
      e[2 * j + 1] = ((g0 + 1u) ^ yz[j]) & m;
    }
  } else {
    const uint32_t r1 = lv.res, r2 = lv.res * lv.res, T = lv.size;
    const uint32_t b = g0 + g1 * r1 + g2 * r2;
    const uint32_t off[4] = {0u, r1, r2, r1 + r2};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      uint32_t a = b + off[j], c = a + 1u;
      e[2 * j] = a >= T ? a - T : a;
      e[2 * j + 1] = c >= T ? c - T : c;
    }
  }
  uint32_t raw[8];
#pragma unroll
  for (int k = 0; k < 8; k += 2) {
    const uint32_t ea = e[k], eb = e[k + 1];
    if ((ea ^ eb) == 1u) {
      const uint2 q = __ldg(reinterpret_cast<const uint2*>(tab32 + (ea & ~1u)));
      raw[k] = (ea & 1u) ? q.y : q.x;
      raw[k + 1] = (ea & 1u) ? q.x : q.y;
    } else {
      raw[k] = __ldg(tab32 + ea);
      raw[k + 1] = __ldg(tab32 + eb);
    }
  }
  if (idx_out) {
#pragma unroll
    for (int k = 0; k < 8; ++k) idx_out[k] = e[k];
  }
  __half2 acc = __floats2half2_rn(0.f, 0.f);
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const float weight = wxy[k & 3] * ((k & 4) ? w2 : o2);
    const float2 v = __half22float2(*reinterpret_cast<const __half2*>(&raw[k]));
    acc = __hadd2(acc, __floats2half2_rn(weight * v.x, weight * v.y));  // per-component fp16 RN add, as __half +=
  }
  return *reinterpret_cast<uint32_t*>(&acc);
}

// tcnn kernel_sh degree 4 (spherical_harmonics.h:66-96); in = normalised direction input in [0,1]
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
  v[6] = 0.94617469575755997f * z2 - 0.31539156525251999f;
  v[7] = -1.0925484305920792f * xz;
  v[8] = 0.54627421529603959f * x2 - 0.54627421529603959f * y2;
  v[9] = 0.59004358992664352f * y * (-3.0f * x2 + y2);
  v[10] = 2.8906114426405538f * xy * z;
  v[11] = 0.45704579946446572f * y * (1.0f - 5.0f * z2);
  v[12] = 0.3731763325901154f * z * (5.0f * z2 - 3.0f);
  v[13] = 0.45704579946446572f * x * (1.0f - 5.0f * z2);
  v[14] = 1.4453057213202769f * z * (x2 - y2);
  v[15] = 0.59004358992664352f * x * (-x2 + 3.0f * y2);
#pragma unroll
  for (int k = 0; k < 8; ++k) out[k] = pack_h2(__float2half_rn(v[2 * k]), __float2half_rn(v[2 * k + 1]));
}

// ------------------------------------------------------------------------------------------------ epilogue helpers
// TMEM -> registers for this thread's row: NCOLS fp32 accumulators starting at column 0 of the CTA's allocation.
template <int NCOLS>
__device__ __forceinline__ void load_acc(uint32_t taddr, uint32_t* acc) {
#pragma unroll
  for (int c = 0; c < NCOLS; c += 16) tmem_ld16(taddr + c, acc + c);
  tmem_wait_ld();
}
// ReLU -> fp16 -> this row's K-major chunks of the A tile: chunk kc (8 halves) lives at kc*2048 + row*16.
template <int NCOLS>
__device__ __forceinline__ void store_relu_row(uint8_t* sA, int row, const uint32_t* acc) {
  // ReLU then fp16 RN.  tcnn computes half(x) * (half(x) > 0) (common_device.h:71-76); max(x,0) then rounding gives
  // the same value for every finite x (rounding is monotonic and keeps the sign), except that a negative input yields
  // +0 instead of -0, which is indistinguishable as an MMA operand.
#pragma unroll
  for (int kc = 0; kc < NCOLS / 8; ++kc) {
    uint32_t w[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const __half2 h = __floats2half2_rn(fmaxf(__uint_as_float(acc[kc * 8 + 2 * j]), 0.f),
                                          fmaxf(__uint_as_float(acc[kc * 8 + 2 * j + 1]), 0.f));
      w[j] = *reinterpret_cast<const uint32_t*>(&h);
    }
    *reinterpret_cast<uint4*>(sA + kc * (NRF_TILE * 16) + row * 16) = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

// MODE 0: encode only  (positions -> enc_io, 64 B per sample, sample-major)          [hash-gather roofline probe]
// MODE 1: MLP only     (enc_io -> out)                                               [tensor-pipe probe]
// MODE 2: fused        (positions -> out), nothing but 8 B per sample leaves the SM  [production path]
// dirs_mode 0: direction = rays[sample_ray[i]].xyz (raw ray direction, normalised here as d*0.5+0.5, nerf_render.cu:313)
// dirs_mode 1: direction input already normalised, rays[i].xyz (stage probes)
template <int MODE>
__global__ void __launch_bounds__(NET_THREADS)
k_network(const __grid_constant__ NetParams P, int n, const float4* __restrict__ samples,
          const int* __restrict__ sample_ray, const float4* __restrict__ rays, int dirs_mode,
          uint4* __restrict__ enc_io, uint2* __restrict__ out, uint32_t* __restrict__ idx_out) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* sW = smem;
  uint8_t* sA = smem + SMEM_W_BYTES;
  uint64_t* mbar = reinterpret_cast<uint64_t*>(smem + SMEM_W_BYTES + SMEM_A_BYTES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + SMEM_W_BYTES + SMEM_A_BYTES + 8);

  const int tid = threadIdx.x, warp = tid >> 5;
  const int n_tiles = (n + NRF_TILE - 1) / NRF_TILE;
  uint32_t tmem_base = 0, phase = 0;
  const uint32_t bar = smem_u32(mbar);

  if (MODE != 0) {
    // weights: 20 KB, already in UMMA core-matrix order -> straight copy
    const uint4* src = P.mlp_packed;
    for (int i = tid; i < SMEM_W_BYTES / 16; i += NET_THREADS) reinterpret_cast<uint4*>(sW)[i] = __ldg(src + i);
    if (tid == 0) mbar_init(bar, 1);
    if (warp == 0) {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(smem_u32(tmem_slot)) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    tmem_base = *tmem_slot;
  }
  const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16);
  const uint32_t a_addr = smem_u32(sA), w_addr = smem_u32(sW);

  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int i = tile * NRF_TILE + tid;
    const bool valid = i < n;
    // ---- stage A0: 32 encoded features of this thread's sample ---------------------------------------------------
    // (the A tile is free here: the previous tile's last MMA was waited on and its TMEM reads are fenced below)
    if (MODE == 1) {
#pragma unroll
      for (int kc = 0; kc < 4; ++kc)
        *reinterpret_cast<uint4*>(sA + kc * (NRF_TILE * 16) + tid * 16) =
            valid ? __ldg(enc_io + (size_t)i * 4 + kc) : make_uint4(0, 0, 0, 0);
    } else {
      const float4 s = valid ? __ldg(samples + i) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
      for (int kc = 0; kc < 4; ++kc) {  // rolled: keeps the gather body (4 levels) resident in the instruction cache
        uint32_t f[4];
#pragma unroll
        for (int li = 0; li < 4; ++li) {
          const int l = kc * 4 + li;
          f[li] = encode_level(P.lv[l], P.table, s.x, s.y, s.z,
                               (idx_out && valid) ? idx_out + ((size_t)i * NRF_L + l) * 8 : nullptr);
        }
        const uint4 v = make_uint4(f[0], f[1], f[2], f[3]);
        if (MODE == 0) {
          if (valid) enc_io[(size_t)i * 4 + kc] = v;
        } else {
          *reinterpret_cast<uint4*>(sA + kc * (NRF_TILE * 16) + tid * 16) = v;
        }
      }
    }
    if (MODE == 0) continue;
    // ---- direction -> SH (per sample here; the direction is per ray) ---------------------------------------------
    uint32_t sh[8];
    {
      float dx = 0.5f, dy = 0.5f, dz = 0.5f;
      if (valid) {
        if (dirs_mode == 0) {
          const float4 r = __ldg(rays + __ldg(sample_ray + i));
          dx = 0.5f * r.x + 0.5f; dy = 0.5f * r.y + 0.5f; dz = 0.5f * r.z + 0.5f;
        } else {
          const float4 r = __ldg(rays + i);
          dx = r.x; dy = r.y; dz = r.z;
        }
      }
      sh4(dx, dy, dz, sh);
    }
    fence_proxy_async();
    __syncthreads();

    uint32_t acc[64];
    // ---- density layer 0: 32 -> 64, ReLU -----------------------------------------------------------------------
    if (tid == 0) { tc_fence_after(); issue_layer<64, 32>(tmem_base, a_addr, w_addr + W0_OFF); umma_commit(bar); }
    mbar_wait(bar, phase); phase ^= 1; tc_fence_after();
    load_acc<64>(taddr, acc);
    store_relu_row<64>(sA, tid, acc);
    tc_fence_before(); fence_proxy_async(); __syncthreads();
    // ---- density layer 1: 64 -> 16, no activation; [den16 | sh16] is the rgb network input ----------------------
    if (tid == 0) { tc_fence_after(); issue_layer<16, 64>(tmem_base, a_addr, w_addr + W1_OFF); umma_commit(bar); }
    mbar_wait(bar, phase); phase ^= 1; tc_fence_after();
    load_acc<16>(taddr, acc);
    __half sigma;
    {
      uint32_t d[8];
#pragma unroll
      for (int k = 0; k < 8; ++k)
        d[k] = pack_h2(__float2half_rn(__uint_as_float(acc[2 * k])), __float2half_rn(__uint_as_float(acc[2 * k + 1])));
      sigma = act_h(P.sigma_act, __ushort_as_half((unsigned short)(d[0] & 0xffffu)));
      *reinterpret_cast<uint4*>(sA + 0 * (NRF_TILE * 16) + tid * 16) = make_uint4(d[0], d[1], d[2], d[3]);
      *reinterpret_cast<uint4*>(sA + 1 * (NRF_TILE * 16) + tid * 16) = make_uint4(d[4], d[5], d[6], d[7]);
      *reinterpret_cast<uint4*>(sA + 2 * (NRF_TILE * 16) + tid * 16) = make_uint4(sh[0], sh[1], sh[2], sh[3]);
      *reinterpret_cast<uint4*>(sA + 3 * (NRF_TILE * 16) + tid * 16) = make_uint4(sh[4], sh[5], sh[6], sh[7]);
    }
    tc_fence_before(); fence_proxy_async(); __syncthreads();
    // ---- rgb layer 0: 32 -> 64, ReLU ---------------------------------------------------------------------------
    if (tid == 0) { tc_fence_after(); issue_layer<64, 32>(tmem_base, a_addr, w_addr + V0_OFF); umma_commit(bar); }
    mbar_wait(bar, phase); phase ^= 1; tc_fence_after();
    load_acc<64>(taddr, acc);
    store_relu_row<64>(sA, tid, acc);
    tc_fence_before(); fence_proxy_async(); __syncthreads();
    // ---- rgb layer 1: 64 -> 64, ReLU ---------------------------------------------------------------------------
    if (tid == 0) { tc_fence_after(); issue_layer<64, 64>(tmem_base, a_addr, w_addr + V1_OFF); umma_commit(bar); }
    mbar_wait(bar, phase); phase ^= 1; tc_fence_after();
    load_acc<64>(taddr, acc);
    store_relu_row<64>(sA, tid, acc);
    tc_fence_before(); fence_proxy_async(); __syncthreads();
    // ---- rgb layer 2: 64 -> 16 (3 used), output activation ------------------------------------------------------
    if (tid == 0) { tc_fence_after(); issue_layer<16, 64>(tmem_base, a_addr, w_addr + V2_OFF); umma_commit(bar); }
    mbar_wait(bar, phase); phase ^= 1; tc_fence_after();
    load_acc<16>(taddr, acc);
    if (valid) {
      const __half r = act_h(P.rgb_act, __float2half_rn(__uint_as_float(acc[0])));
      const __half g = act_h(P.rgb_act, __float2half_rn(__uint_as_float(acc[1])));
      const __half b = act_h(P.rgb_act, __float2half_rn(__uint_as_float(acc[2])));
      out[i] = make_uint2(pack_h2(r, g), pack_h2(b, sigma));
    }
    // the next tile's A0 store and first MMA must not overtake this tile's TMEM reads
    tc_fence_before();
    __syncthreads();
  }

  if (MODE != 0) {
    tc_fence_before();
    __syncthreads();
    if (warp == 0)
      asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(tmem_base) : "memory");
  }
}


// ------------------------------------------------------------------------------------------------ SIMT check kernel
// Scalar fp32-accumulate evaluation of the same network from the row-major weights; used only by the stage probe
// nrf_probe_network(impl=1) to cross-check the tcgen05 path on the GPU.  Not on the frame path.
__device__ __forceinline__ void dense_simt(const __half* __restrict__ W, int n_out, int n_in, const __half* x,
                                           __half* y, bool relu) {
  for (int o = 0; o < n_out; ++o) {
    float acc = 0.f;
    for (int k = 0; k < n_in; ++k) acc = acc + __half2float(__ldg(W + o * n_in + k)) * __half2float(x[k]);
    y[o] = relu ? relu_h(acc) : __float2half_rn(acc);
  }
}

extern "C" __global__ void __launch_bounds__(128)
k_network_simt(const __grid_constant__ NetParams P, int n, const uint4* __restrict__ enc,
               const float4* __restrict__ dirs01, uint2* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  __half x[32], h[64], h2[64], rin[32], o16[16];
#pragma unroll
  for (int kc = 0; kc < 4; ++kc) {
    const uint4 v = __ldg(enc + (size_t)i * 4 + kc);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      x[kc * 8 + 2 * j] = __ushort_as_half((unsigned short)(w[j] & 0xffffu));
      x[kc * 8 + 2 * j + 1] = __ushort_as_half((unsigned short)(w[j] >> 16));
    }
  }
  const __half* W = P.mlp_rowmaj;
  dense_simt(W, 64, 32, x, h, true);
  dense_simt(W + 2048, 16, 64, h, rin, false);
  const __half sigma = act_h(P.sigma_act, rin[0]);
  uint32_t sh[8];
  const float4 d = __ldg(dirs01 + i);
  sh4(d.x, d.y, d.z, sh);
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    rin[16 + 2 * k] = __ushort_as_half((unsigned short)(sh[k] & 0xffffu));
    rin[16 + 2 * k + 1] = __ushort_as_half((unsigned short)(sh[k] >> 16));
  }
  dense_simt(W + 3072, 64, 32, rin, h, true);
  dense_simt(W + 5120, 64, 64, h, h2, true);
  dense_simt(W + 9216, 16, 64, h2, o16, false);
  out[i] = make_uint2(pack_h2(act_h(P.rgb_act, o16[0]), act_h(P.rgb_act, o16[1])),
                      pack_h2(act_h(P.rgb_act, o16[2]), sigma));
}

// ------------------------------------------------------------------------------------------------ host launchers
extern "C" size_t nrf_network_smem_bytes() { return SMEM_W_BYTES + SMEM_A_BYTES + 64; }

static void nrf_network_setup() {
  static bool attr_set = false;
  if (attr_set) return;
  const size_t smem = nrf_network_smem_bytes();
  {
    cudaFuncSetAttribute(k_network<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(k_network<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(k_network<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    // ask for the full shared-memory carveout: several 37 KB CTAs per SM must be resident to overlap one tile's
    // gather with another tile's MLP (the occupancy query reports 1 CTA/SM under the default carveout)
    cudaFuncSetAttribute(k_network<0>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaFuncSetAttribute(k_network<1>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaFuncSetAttribute(k_network<2>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    attr_set = true;
  }
}

extern "C" cudaError_t nrf_launch_network(int mode, int grid, cudaStream_t stream, const NetParams* P, int n,
                                          const float4* samples, const int* sample_ray, const float4* rays,
                                          int dirs_mode, uint4* enc_io, uint2* out, uint32_t* idx_out) {
  const size_t smem = nrf_network_smem_bytes();
  nrf_network_setup();
  if (n <= 0) return cudaSuccess;
  switch (mode) {
    case 0: k_network<0><<<grid, NET_THREADS, smem, stream>>>(*P, n, samples, sample_ray, rays, dirs_mode, enc_io, out, idx_out); break;
    case 1: k_network<1><<<grid, NET_THREADS, smem, stream>>>(*P, n, samples, sample_ray, rays, dirs_mode, enc_io, out, idx_out); break;
    default: k_network<2><<<grid, NET_THREADS, smem, stream>>>(*P, n, samples, sample_ray, rays, dirs_mode, enc_io, out, idx_out); break;
  }
  return cudaGetLastError();
}

extern "C" cudaError_t nrf_launch_network_simt(cudaStream_t stream, const NetParams* P, int n, const uint4* enc,
                                               const float4* dirs01, uint2* out) {
  if (n <= 0) return cudaSuccess;
  k_network_simt<<<(n + 127) / 128, 128, 0, stream>>>(*P, n, enc, dirs01, out);
  return cudaGetLastError();
}

extern "C" int nrf_network_occupancy(int mode) {
  nrf_network_setup();
  int nb = 0;
  const size_t smem = nrf_network_smem_bytes();
  cudaError_t e;
  if (mode == 0) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_network<0>, NET_THREADS, smem);
  else if (mode == 1) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_network<1>, NET_THREADS, smem);
  else e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_network<2>, NET_THREADS, smem);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return -(int)e;
  }
  return nb;
}
