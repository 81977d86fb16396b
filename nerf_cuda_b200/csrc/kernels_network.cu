// kernels_network.cu — multiresolution hash-grid gather + fused density/colour MLP on tcgen05 (sm_100a).
//
// Replaces, per sample, the reference's K10-K15:
//   tcnn::kernel_grid<__half,3,2>            dependencies/tiny-cuda-nn/include/tiny-cuda-nn/encodings/grid.h:139-268
//   kernel_mlp_fused (density 32->64->16)    dependencies/tiny-cuda-nn/src/fully_fused_mlp.cu:500-558
//   tcnn::kernel_sh<__half> (degree 4)       .../encodings/spherical_harmonics.h:46-152   (per RAY, see k_rays_count)
//   kernel_mlp_fused (rgb 32->64->64->16)    fully_fused_mlp.cu:500-558
//   extract_density                          include/nerf-cuda/nerf_network.h:49-61,187-195
//   decompose_network_in_and_out             include/nerf-cuda/render_utils.h:308-334
// and removes every global-memory round trip between them: only {r,g,b,sigma} (4 halves) per sample reach HBM.
//
// Shape of the kernel (k_network): ONE persistent CTA per SM with 7 tile groups of 128 threads (896 threads x 72
// registers; 8 groups x 64 registers spill and measure 3-4 % slower, 6 groups lose more to occupancy); a group
// processes one 128-sample tile at a time (UMMA M = 128) and thread t of the group owns sample row t end to end
// (gather -> A-operand row t -> accumulator lane t).  All groups share one 20 KB copy of the weights in shared memory
// (pre-packed on the host into the K-major SWIZZLE_NONE core-matrix order); nothing else lives in shared memory, so
// ~200 KB of the SM's unified array stays L1 for the table gathers.  Every activation lives in TMEM:
//   * the CTA owns all 512 TMEM columns, 64 per group: columns [0,32) = A operand (fp16 pairs, row = lane: the layout
//     tcgen05.mma reads for an A-from-TMEM operand), columns [32,64) = fp32 accumulator;
//   * a layer is tcgen05.st (thread t writes row t of A) -> tcgen05.mma A[tmem] x W[smem] -> tcgen05.ld (thread t reads
//     row t of D) -> ReLU/pack (cvt.rn.relu.f16x2.f32) -> next tcgen05.st.  64-wide layers run as two N=32 halves so
//     that A (<= 32 columns) and D (32 columns) fit the group's 64 columns;
//   * groups synchronise with their own named barrier and their own mbarrier, never with the whole CTA: while one
//     group waits for an MMA the other 28 warps keep the gather's loads in flight.
// The G groups of a CTA (7 on the frame path) walk ONE contiguous chunk of tiles interleaved (tile = chunk + G*j + g): at any time an SM works
// on neighbouring patches / steps of the warp-interleaved sample list, which is what the L1 hit rate lives on.
//
// Numerics: hash indices and fp16 interpolation are bit-exact against the reference's CUDA build executed on a B200
// (tests/golden/ref_cuda_*.npz): the per-level scale comes from the device's exp2f (k_level_scales below), the position
// is one FFMA like in the reference's SASS, everything else is uncontracted (this TU is built with -fmad=false);
// the MLP accumulates in fp32 (TMEM) and rounds to fp16 at each layer boundary like the oracle (the reference
// accumulates in fp16 inside wmma: outputs agree to ~5e-4 absolute, tests/test_ref_cuda_golden.py).
#include <cstdlib>
#include <mutex>

#include "nrf_device.cuh"

#define NET_GROUPS_DEFAULT 8   // template default (stage probes MODE 0/1 run 8 groups x 128 threads); the frame path launches 7 (g_net_groups), 6 / 8 via NRF_NET_GROUPS
#define SMEM_W_BYTES 20480
#define NET_SMEM (SMEM_W_BYTES + 256)   // weights | 8 MMA mbarriers | TMEM slot | (+128) 8 "operands ready" mbarriers
#define NET_SMEM_L01 (66 * 1024)   // experiment NRF_EXP_SMEM_L01: room for the two coarsest (dense) levels behind the weights
#ifdef NRF_EXP_SMEM_L01
#if NRF_EXP_SMEM_L01
#define NET_SMEM_LAUNCH (NET_SMEM + NET_SMEM_L01)
#endif
#endif
#ifndef NET_SMEM_LAUNCH
#define NET_SMEM_LAUNCH NET_SMEM
#endif
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
__device__ __forceinline__ void bar_sync_group(int id) { asm volatile("bar.sync %0, 128;" ::"r"(id) : "memory"); }
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
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
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
// A operand from TENSOR MEMORY, B from shared memory
__device__ __forceinline__ void umma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// D[128 x N] = A[128 x K] (TMEM, K/2 columns from tmem_a) * W[n0 .. n0+N)[K]^T.  W is stored for NFULL rows in the K-major
// core-matrix order (LBO = NFULL*16 between 8-element K chunks, SBO = 128 between 8-row groups).
// w16 = (shared-memory address of the weight block) >> 4, computed once per kernel; byte_off = compile-time offset of the
// matrix (+ row offset) inside the block.  The descriptor's low word is then ONE add per MMA — start address, in 16-byte
// units, plus a constant that carries the leading-dimension field — and its high word a constant (the 20 KB block sits
// far below 2^18, so the 14-bit address field never carries into the LBO field).
template <int N, int K, int NFULL>
__device__ __forceinline__ void issue_layer_ts(uint32_t tmem_d, uint32_t tmem_a, uint32_t w16, int byte_off) {
  constexpr uint32_t idesc = umma_idesc(NRF_TILE, N);
  constexpr uint32_t hi = (uint32_t)(128 >> 4) | (1u << 14);   // SBO = 128 B at [32,46), version 1 at [46,48)
#pragma unroll
  for (int kk = 0; kk < K / 16; ++kk) {
    const uint32_t lo = w16 + (uint32_t)((byte_off + kk * 2 * (NFULL * 16)) >> 4) + ((uint32_t)((NFULL * 16) >> 4) << 16);
    umma_f16_ts(tmem_d, tmem_a + kk * 8, ((uint64_t)hi << 32) | lo, idesc, kk > 0);
  }
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// ------------------------------------------------------------------------------------------------ numerics helpers
// ReLU + fp16 RN of two accumulators in one instruction.  tcnn computes half(x) * (half(x) > 0) (common_device.h:71-76);
// cvt.rn.relu gives the same value for every finite x except that a negative input yields +0 instead of -0, which is
// indistinguishable as an MMA operand.  Result {lo, hi} = {a, b}.
__device__ __forceinline__ uint32_t relu_pack(uint32_t a_bits, uint32_t b_bits) {
  uint32_t r;
  asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(__uint_as_float(b_bits)), "f"(__uint_as_float(a_bits)));
  return r;
}
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
// two fp32 products with one FMUL2: {v.x*w, v.y*w}  (IEEE RN each, identical to two FMULs)
__device__ __forceinline__ float2 mul2(float2 v, float w) {
  unsigned long long r;
  const unsigned long long x = ((unsigned long long)__float_as_uint(v.y) << 32) | __float_as_uint(v.x);
  const unsigned long long y = ((unsigned long long)__float_as_uint(w) << 32) | __float_as_uint(w);
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(x), "l"(y));
  return make_float2(__uint_as_float((uint32_t)r), __uint_as_float((uint32_t)(r >> 32)));
}

// One level of tcnn kernel_grid (grid.h:186-268) + grid_index/fast_hash (grid.h:81-117): trilinear blend accumulated
// in fp16 in corner order 0..7 (corner bit d selects g[d]+1 and weight w[d], else 1-w[d]; weight = (f0*f1)*f2).
// Index math: hashed levels  (x ^ y*2654435761 ^ z*805459861) & (T-1)   (T is a power of two);
//             dense levels   x + y*res + z*res^2, which is < 2T, so min(a, a-T) (unsigned) equals the reference's % T;
//             mode 3         the same dense index in wrapping 32-bit arithmetic & (T-1), where the reference's stride overflowed.
// Straight-line code, no divergence.  On hashed levels the two x-corners of a cell share an aligned 8-byte slot exactly
// when x is even (the hash prime for x is 1, grid.h:88): one LDG.64 fetches the slot of the x corner and a LDG.32
// predicated on "x odd" fetches the x+1 corner when it lives elsewhere; the predicate is the same for all four pairs.
// Split in two halves so that a caller can put the loads of two levels in flight before blending either:
//   level_load  : positions -> 8 raw entries (loads issued) + the three fractional weights
//   level_blend : fp16-accumulated trilinear blend of the 8 entries
struct LevelTaps {
  uint32_t raw[8];
  float w0, w1, w2;
};

// ---- experiment hooks (nerf_cuda_b200/build.py VARIANTS; all off in the product build; results in DESIGN.md §4) ----
#ifndef NRF_EXP_FLOOR_MAGIC
#define NRF_EXP_FLOOR_MAGIC 0
#endif
#ifndef NRF_EXP_EVICT
#define NRF_EXP_EVICT 0
#endif
#ifndef NRF_EXP_SMEM_L01
#define NRF_EXP_SMEM_L01 0
#endif
#ifndef NRF_NO_TAIL_SYNC      // 1: no group barrier at the end of a tile (the next tile's first one already orders the TMEM reads)
#define NRF_NO_TAIL_SYNC 0
#endif
#ifndef NRF_ARRIVE_SYNC       // 1: "operands ready" as mbarrier arrive (one lane per warp) + wait by the issuing thread only
#define NRF_ARRIVE_SYNC 0
#endif
#if NRF_EXP_EVICT
// table entries: keep in L2 (evict_last); sample records / outputs: streamed once (evict_first, no L1 allocation)
__device__ __forceinline__ uint64_t policy_evict_last() { uint64_t p; asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p)); return p; }
__device__ __forceinline__ uint64_t policy_evict_first() { uint64_t p; asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p)); return p; }
__device__ __forceinline__ uint32_t tab_ld32(const uint32_t* a) {
  uint32_t v;
  asm("ld.global.nc.L2::cache_hint.b32 %0, [%1], %2;" : "=r"(v) : "l"(a), "l"(policy_evict_last()));
  return v;
}
__device__ __forceinline__ uint2 tab_ld64(const uint2* a) {
  uint2 v;
  asm("ld.global.nc.L2::cache_hint.v2.b32 {%0,%1}, [%2], %3;" : "=r"(v.x), "=r"(v.y) : "l"(a), "l"(policy_evict_last()));
  return v;
}
__device__ __forceinline__ float4 sample_ld(const float4* a) {
  float4 f;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;"
               : "=f"(f.x), "=f"(f.y), "=f"(f.z), "=f"(f.w) : "l"(a), "l"(policy_evict_first()));
  return f;
}
__device__ __forceinline__ void out_st(uint2* a, uint2 v) {
  asm volatile("st.global.L2::cache_hint.v2.b32 [%0], {%1,%2}, %3;" ::"l"(a), "r"(v.x), "r"(v.y), "l"(policy_evict_first()) : "memory");
}
#else
__device__ __forceinline__ uint32_t tab_ld32(const uint32_t* a) { return __ldg(a); }
__device__ __forceinline__ uint2 tab_ld64(const uint2* a) { return __ldg(a); }
__device__ __forceinline__ float4 sample_ld(const float4* a) { return __ldg(a); }
__device__ __forceinline__ void out_st(uint2* a, uint2 v) { *a = v; }
#endif

__device__ __forceinline__ uint32_t lds32(uint32_t saddr) {
  uint32_t v;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(saddr));
  return v;
}
// stab != 0 (experiment NRF_EXP_SMEM_L01): this DENSE level's table sits in shared memory at that address
// LM >= 0: the level's addressing mode is known at compile time (kernel specialised for the model's level layout, see
// NRF_LMODES_T19): no jump table per level and half the code; LM < 0: read from the level descriptor.
template <bool IDX, int LM = -1>
__device__ __forceinline__ void level_load(const LevelInfo& lv, float u0, float u1, float u2, LevelTaps& T, uint32_t* idx_out,
                                           uint32_t stab = 0) {
  const uint32_t mode = LM >= 0 ? (uint32_t)LM : lv.hashed;
  // pos_fract (common_device.h:424-431): input*scale + 0.5 is ONE FFMA in the reference's kernel_grid SASS
  const float p0 = __fmaf_rn(u0, lv.scale, 0.5f), p1 = __fmaf_rn(u1, lv.scale, 0.5f), p2 = __fmaf_rn(u2, lv.scale, 0.5f);
#if NRF_EXP_FLOOR_MAGIC
  // 0 <= p < 2^23: p + 2^23 rounded toward zero IS floor(p) + 2^23 (ulp 1 there) — FADD.RZ on the FMA pipe instead of
  // FRND + F2I on the quarter-rate XU pipe; the integer is the low mantissa
  const float m0 = __fadd_rz(p0, 8388608.f), m1 = __fadd_rz(p1, 8388608.f), m2 = __fadd_rz(p2, 8388608.f);
  const float f0 = m0 - 8388608.f, f1 = m1 - 8388608.f, f2 = m2 - 8388608.f;
  const uint32_t g0 = __float_as_uint(m0) & 0x7fffffu, g1 = __float_as_uint(m1) & 0x7fffffu, g2 = __float_as_uint(m2) & 0x7fffffu;
#else
  const float f0 = floorf(p0), f1 = floorf(p1), f2 = floorf(p2);
  const uint32_t g0 = (uint32_t)(int)f0, g1 = (uint32_t)(int)f1, g2 = (uint32_t)(int)f2;
#endif
  T.w0 = p0 - f0; T.w1 = p1 - f1; T.w2 = p2 - f2;
  const uint32_t* __restrict__ tl = lv.tab;
  uint32_t e[8];
  if (mode == 1) {
    const uint32_t m = lv.size - 1u;
    // (a ^ b ^ c) & m == (a & m) ^ (b & m) ^ (c & m): six masks + eight 3-input XORs instead of eight XORs + eight masks
    const uint32_t hy0 = g1 * 2654435761u, hy1 = hy0 + 2654435761u, hz0 = g2 * 805459861u, hz1 = hz0 + 805459861u;
    const uint32_t my[2] = {hy0 & m, hy1 & m}, mz[2] = {hz0 & m, hz1 & m};
    const uint32_t x0 = g0 & m, x1 = (g0 + 1u) & m;
    const bool odd = g0 & 1u;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      e[2 * j] = x0 ^ my[j & 1] ^ mz[j >> 1];
      e[2 * j + 1] = x1 ^ my[j & 1] ^ mz[j >> 1];
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint2 q = tab_ld64(reinterpret_cast<const uint2*>(tl) + (e[2 * j] >> 1));
      uint32_t far = 0;
      if (odd) far = tab_ld32(tl + e[2 * j + 1]);
      const bool up = e[2 * j] & 1u;
      T.raw[2 * j] = up ? q.y : q.x;
      T.raw[2 * j + 1] = odd ? far : (up ? q.x : q.y);
    }
  } else if (mode == 2) {
    // hashed level, plain 4-byte loads (no slot pairing): fewer instructions, more L1 wavefronts — chosen per level on
    // the host for the levels whose lanes share cache lines anyway (NRF_PAIR_FROM)
    const uint32_t m = lv.size - 1u;
    const uint32_t hy0 = g1 * 2654435761u, hy1 = hy0 + 2654435761u, hz0 = g2 * 805459861u, hz1 = hz0 + 805459861u;
    const uint32_t my[2] = {hy0 & m, hy1 & m}, mz[2] = {hz0 & m, hz1 & m};
    const uint32_t x0 = g0 & m, x1 = (g0 + 1u) & m;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      e[2 * j] = x0 ^ my[j & 1] ^ mz[j >> 1];
      e[2 * j + 1] = x1 ^ my[j & 1] ^ mz[j >> 1];
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) T.raw[k] = tab_ld32(tl + e[k]);
  } else if (mode == 3) {
    // res^3 wrapped below T in the reference's 32-bit stride (grid.h:100-117): dense index, 32-bit wrap, mod T (a power of 2)
    const uint32_t r1 = lv.res, r2 = lv.res * lv.res, m = lv.size - 1u;
    const uint32_t b = g0 + g1 * r1 + g2 * r2;
    const uint32_t off[4] = {0u, r1, r2, r1 + r2};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      e[2 * j] = (b + off[j]) & m;
      e[2 * j + 1] = (b + off[j] + 1u) & m;
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) T.raw[k] = tab_ld32(tl + e[k]);
  } else {
    const uint32_t r1 = lv.res, r2 = lv.res * lv.res, Tl = lv.size;
    const uint32_t b = g0 + g1 * r1 + g2 * r2;
    const uint32_t off[4] = {0u, r1, r2, r1 + r2};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t a = b + off[j], c = a + 1u;
      e[2 * j] = min(a, a - Tl);
      e[2 * j + 1] = min(c, c - Tl);
    }
    if (NRF_EXP_SMEM_L01 && stab) {
#pragma unroll
      for (int k = 0; k < 8; ++k) T.raw[k] = lds32(stab + e[k] * 4u);
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k) T.raw[k] = tab_ld32(tl + e[k]);
    }
  }
  if (IDX && idx_out) {
#pragma unroll
    for (int k = 0; k < 8; ++k) idx_out[k] = e[k];
  }
}

__device__ __forceinline__ uint32_t level_blend(const LevelTaps& T) {
  const float o0 = 1.0f - T.w0, o1 = 1.0f - T.w1, o2 = 1.0f - T.w2;
  const float2 wlo = mul2(make_float2(o0, T.w0), o1), whi = mul2(make_float2(o0, T.w0), T.w1);  // (f0*f1), corner bits 0,1
  const float2 w01 = mul2(wlo, o2), w23 = mul2(whi, o2), w45 = mul2(wlo, T.w2), w67 = mul2(whi, T.w2);
  const float wk[8] = {w01.x, w01.y, w23.x, w23.y, w45.x, w45.y, w67.x, w67.y};
  __half2 acc = __floats2half2_rn(0.f, 0.f);
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const float2 v = mul2(__half22float2(*reinterpret_cast<const __half2*>(&T.raw[k])), wk[k]);
    acc = __hadd2(acc, __floats2half2_rn(v.x, v.y));  // per-component fp16 RN add, as __half +=
  }
  return *reinterpret_cast<uint32_t*>(&acc);
}

// 32 fp32 accumulator columns of this thread's row -> ReLU -> 16 packed fp16 pairs
__device__ __forceinline__ void ld_relu_pack32(uint32_t taddr, uint32_t* packed) {
  uint32_t a[32];
  tmem_ld16(taddr, a);
  tmem_ld16(taddr + 16, a + 16);
  tmem_wait_ld();
#pragma unroll
  for (int j = 0; j < 16; ++j) packed[j] = relu_pack(a[2 * j], a[2 * j + 1]);
}

// MODE 0: encode only  (positions -> enc_io, 64 B per sample, sample-major)          [hash-gather roofline probe]
// MODE 1: MLP only     (enc_io -> out)                                               [tensor-pipe probe]
// MODE 2: fused        (positions -> out), nothing but 8 B per sample leaves the SM  [production path]
// dirs_mode 0: SH coefficients of the sample's ray from ray_sh[sample_ray[i]] (computed once per ray by k_rays_count)
// dirs_mode 1: direction input already normalised, rays[i].xyz; SH evaluated here (stage probes)
//
// MODE 2 is software-pipelined inside every group: while the tensor core works on a layer of tile j, the group's
// threads gather 3-4 levels of tile j+G (the group's next tile, G = groups per CTA).  The gather slices sit in the five MMA waits where no
// activation registers are live (after L0-lo, L1, V0-lo, V1-lo, V2 were issued), so the MLP's round-trip latency is
// hidden behind the group's own gather instead of idling 4 of the SM's 32 warps.
// level boundaries of the gather slices hidden in the five MMA waits of a tile: 0-4 | 4-SL1 | SL1-SL2 | SL2-SL3 | SL3-16
#ifndef SL1
#define SL1 7
#define SL2 10
#define SL3 13
#endif
#define NRF_LMODES_RUNTIME 0xFFFFFFFFu
// 2 bits per level: the layouts of the log2_hashmap_size = 19 models with the base.json grid.  aabb_scale 1 (C2): levels 0-4
// dense, 5-11 hashed with 4-byte loads, 12-15 hashed with slot-paired loads (NRF_PAIR_FROM = 12); aabb_scale 16 (C3, C4: finer
// per_level_scale): levels 0-3 dense, 4-11 hashed, 12-15 slot-paired
#define NRF_LMODES_T19 0x55AAA800u
#define NRF_LMODES_T19_B16 0x55AAAA00u
// log2_hashmap_size = 24, aabb_scale 1 (C5): levels 0-8 dense, 9-11 hashed, 12-14 slot-paired, 15 the wrapped dense index
#define NRF_LMODES_T24 0xD5A80000u
template <int L, int L1, bool IDX, uint32_t LMODES>
__device__ __forceinline__ void gather_pairs(const NetParams& P, const float4& s, uint32_t* f, uint32_t* idx_out, uint32_t stab0,
                                             uint32_t stab1) {
  // two levels' loads in flight before either blend: measured 0.7 % faster than level by level (ptxas already
  // overlaps most of it; 7 groups x 72 registers then spill 4 bytes)
  if constexpr (L < L1) {
    constexpr int M0 = LMODES == NRF_LMODES_RUNTIME ? -1 : (int)((LMODES >> (2 * L)) & 3u);
    constexpr int M1 = LMODES == NRF_LMODES_RUNTIME ? -1 : (int)((LMODES >> (2 * ((L + 1) & 15))) & 3u);
    LevelTaps A, B;
    level_load<IDX, M0>(P.lv[L], s.x, s.y, s.z, A, IDX && idx_out ? idx_out + L * 8 : nullptr, L == 0 ? stab0 : (L == 1 ? stab1 : 0u));
    if constexpr (L + 1 < L1)
      level_load<IDX, M1>(P.lv[L + 1], s.x, s.y, s.z, B, IDX && idx_out ? idx_out + (L + 1) * 8 : nullptr, L + 1 == 1 ? stab1 : 0u);
    f[L] = level_blend(A);
    if constexpr (L + 1 < L1) f[L + 1] = level_blend(B);
    gather_pairs<L + 2, L1, IDX, LMODES>(P, s, f, idx_out, stab0, stab1);
  }
}
template <int L0, int L1, bool IDX, uint32_t LMODES = NRF_LMODES_RUNTIME>
__device__ __forceinline__ void gather_slice(const NetParams& P, const float4& s, uint32_t* f, uint32_t* idx_out,
                                             uint32_t stab0 = 0, uint32_t stab1 = 0) {
  gather_pairs<L0, L1, IDX, LMODES>(P, s, f, idx_out, stab0, stab1);
}

// Occupancy experiment (not on the frame path): encode-only with THREADS per CTA and MINB CTAs per SM.
template <int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB)
k_encode_exp(const __grid_constant__ NetParams P, int n, const float4* __restrict__ samples, uint4* __restrict__ enc_io) {
  const int n_tiles = (n + NRF_TILE - 1) / NRF_TILE;
  const int per_cta = (n_tiles + gridDim.x - 1) / gridDim.x;
  const int t0 = blockIdx.x * per_cta, t1 = min(n_tiles, t0 + per_cta);
  for (int tile = t0 + (threadIdx.x >> 7); tile < t1; tile += THREADS / 128) {
    const int i = tile * NRF_TILE + (threadIdx.x & 127);
    if (i >= n) continue;
    const float4 s = __ldg(samples + i);
    uint32_t f[16];
    gather_slice<0, 8, false>(P, s, f, nullptr);
    gather_slice<8, 16, false>(P, s, f, nullptr);
#pragma unroll
    for (int kc = 0; kc < 4; ++kc)
      enc_io[(size_t)i * 4 + kc] = make_uint4(f[kc * 4], f[kc * 4 + 1], f[kc * 4 + 2], f[kc * 4 + 3]);
  }
}
extern "C" cudaError_t nrf_launch_encode_exp(int variant, int sms, cudaStream_t stream, const NetParams* P, int n,
                                             const float4* samples, uint4* enc_io) {
  switch (variant) {
    case 0: k_encode_exp<1024, 1><<<sms, 1024, 0, stream>>>(*P, n, samples, enc_io); break;      // 32 warps, 64 regs
    case 1: k_encode_exp<768, 2><<<sms * 2, 768, 0, stream>>>(*P, n, samples, enc_io); break;    // 48 warps, 40 regs
    case 2: k_encode_exp<1024, 2><<<sms * 2, 1024, 0, stream>>>(*P, n, samples, enc_io); break;  // 64 warps, 32 regs
    case 3: k_encode_exp<512, 1><<<sms, 512, 0, stream>>>(*P, n, samples, enc_io); break;        // 16 warps
    default: k_encode_exp<768, 1><<<sms, 768, 0, stream>>>(*P, n, samples, enc_io); break;       // 24 warps
  }
  return cudaGetLastError();
}

template <int MODE, int NET_GROUPS = NET_GROUPS_DEFAULT, uint32_t LMODES = NRF_LMODES_RUNTIME>
__global__ void __launch_bounds__(NET_GROUPS * 128, 1)
k_network(const __grid_constant__ NetParams P, int n, const float4* __restrict__ samples,
          const int* __restrict__ sample_ray, const float4* __restrict__ rays, const uint4* __restrict__ ray_sh,
          int dirs_mode, uint4* __restrict__ enc_io, uint2* __restrict__ out, uint32_t* __restrict__ idx_out,
          const unsigned long long* __restrict__ n_dev, const __grid_constant__ SampleSrc S) {
  // frame path: the sample count is still in device memory when this kernel is enqueued (no host round trip in the
  // middle of a frame); `n` is then the capacity of the sample buffers
  // (a total above the capacity means the list has holes — k_emit skipped the patches that did not fit — and the host
  // will render the frame again with larger buffers: nothing to do here)
  if (n_dev) {
    const unsigned long long total = __ldg(n_dev);
    n = total > (unsigned long long)n ? 0 : (int)total;
  }
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* sW = smem;
  uint64_t* mbars = reinterpret_cast<uint64_t*>(smem + SMEM_W_BYTES);           // one per group
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + SMEM_W_BYTES + 64);

  // warp and group index read back from lane 0: the compiler then knows them (and the TMEM addresses derived from them) to
  // be warp-uniform and keeps them in uniform registers — the tcgen05 instructions take uniform operands, and without this
  // every MMA issue moved its addresses there with an ELECT / R2UR.BROADCAST retry loop
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), g = warp >> 2, gt = tid & 127;
  const int n_tiles = (n + NRF_TILE - 1) / NRF_TILE;
  // contiguous chunk of tiles per CTA, interleaved over its NET_GROUPS groups
  const int per_cta = (n_tiles + gridDim.x - 1) / gridDim.x;
  const int t0 = blockIdx.x * per_cta, t1 = min(n_tiles, t0 + per_cta);

  if (MODE == 0) {
    for (int tile = t0 + g; tile < t1; tile += NET_GROUPS) {
      const int i = tile * NRF_TILE + gt;
      if (i >= n) continue;
      const float4 s = __ldg(samples + i);
      uint32_t f[16];
      uint32_t* io = idx_out ? idx_out + (size_t)i * NRF_L * 8 : nullptr;
      gather_slice<0, 8, true>(P, s, f, io);
      gather_slice<8, 16, true>(P, s, f, io);
#pragma unroll
      for (int kc = 0; kc < 4; ++kc)
        enc_io[(size_t)i * 4 + kc] = make_uint4(f[kc * 4], f[kc * 4 + 1], f[kc * 4 + 2], f[kc * 4 + 3]);
    }
    return;
  }

  uint32_t phase = 0;
  const uint32_t bar = smem_u32(mbars + g);
#if NRF_ARRIVE_SYNC
  uint32_t rphase = 0;
  const uint32_t rbar = smem_u32(smem + SMEM_W_BYTES + 128) + 8u * (uint32_t)g;
#endif
  {
    // weights: 20 KB, already in UMMA core-matrix order -> straight copy
    const uint4* src = P.mlp_packed;
    for (int i = tid; i < SMEM_W_BYTES / 16; i += NET_GROUPS * 128) reinterpret_cast<uint4*>(sW)[i] = __ldg(src + i);
#if NRF_EXP_SMEM_L01
    if (MODE == 2 && P.lv[0].hashed == 0 && P.lv[1].hashed == 0 && (P.lv[0].size + P.lv[1].size) * 4u <= NET_SMEM_L01) {
      uint32_t* st = reinterpret_cast<uint32_t*>(smem + NET_SMEM);
      const uint32_t n01 = P.lv[0].size + P.lv[1].size;   // level 1 follows level 0 in the table as well
      for (uint32_t i = tid; i < n01; i += NET_GROUPS * 128) st[i] = __ldg(P.lv[0].tab + i);
    }
#endif
    if (gt == 0) mbar_init(bar, 1);
#if NRF_ARRIVE_SYNC
    if (gt == 0) mbar_init(rbar, 4);   // one arrival per warp of the group
#endif
    if (warp == 0) {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmem_slot)) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
  }
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
  const uint32_t t_a = tmem_base + (uint32_t)(g * 64);         // A operand columns (lane field 0: the MMA's view)
  const uint32_t t_d = t_a + 32;                                // accumulator columns
  const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;  // this warp's 32 TMEM lanes
  const uint32_t w16 = (smem_u32(sW) & 0x3FFFFu) >> 4;   // the weight block's address field of every UMMA descriptor
  const int bar_id = 1 + g;

#if NRF_ARRIVE_SYNC
  // a warp whose TMEM stores / loads are done announces it and goes on (to its gather slice or to the MMA wait); only the
  // issuing thread waits for all four warps.  Between two uses every warp passes an MMA wait, so the phases cannot overlap.
#define GROUP_SYNC() do { tc_fence_before(); __syncwarp(); if ((tid & 31) == 0) mbar_arrive(rbar); } while (0)
#define ISSUE(N_, K_, NFULL_, OFF_, N0_) \
  do { if (gt == 0) { mbar_wait(rbar, rphase); tc_fence_after(); issue_layer_ts<N_, K_, NFULL_>(t_d, t_a, w16, (OFF_) + (N0_) * 16); umma_commit(bar); } rphase ^= 1; } while (0)
#else
#define GROUP_SYNC() do { tc_fence_before(); bar_sync_group(bar_id); } while (0)
#define ISSUE(N_, K_, NFULL_, OFF_, N0_) \
  do { if (gt == 0) { tc_fence_after(); issue_layer_ts<N_, K_, NFULL_>(t_d, t_a, w16, (OFF_) + (N0_) * 16); umma_commit(bar); } } while (0)
#endif
#define MMA_WAIT() do { mbar_wait(bar, phase); phase ^= 1; tc_fence_after(); } while (0)

#if NRF_EXP_SMEM_L01
  const bool staged = MODE == 2 && P.lv[0].hashed == 0 && P.lv[1].hashed == 0 && (P.lv[0].size + P.lv[1].size) * 4u <= NET_SMEM_L01;
  const uint32_t stab0 = staged ? smem_u32(smem + NET_SMEM) : 0u, stab1 = staged ? stab0 + P.lv[0].size * 4u : 0u;
#else
  const uint32_t stab0 = 0, stab1 = 0;
#endif
  // features of the group's first tile (pipeline prologue)
  uint32_t fn[16];
  float4 sn = make_float4(0.f, 0.f, 0.f, 0.f);
  int tile = t0 + g;
  if (MODE == 2 && tile < t1) {
    const int i = tile * NRF_TILE + gt;
    if (i < n) sn = sample_ld(samples + i);
    gather_slice<0, 8, false, LMODES>(P, sn, fn, nullptr, stab0, stab1);
    gather_slice<8, 16, false, LMODES>(P, sn, fn, nullptr);
  }

  for (; tile < t1; tile += NET_GROUPS) {
    const int i = tile * NRF_TILE + gt;
    const bool valid = i < n;
    // ---- this tile's features -> A operand; next tile's sample position ---------------------------------------------
    if (MODE == 1) {
#pragma unroll
      for (int kc = 0; kc < 4; ++kc) {
        const uint4 v = valid ? __ldg(enc_io + (size_t)i * 4 + kc) : make_uint4(0, 0, 0, 0);
        fn[kc * 4 + 0] = v.x; fn[kc * 4 + 1] = v.y; fn[kc * 4 + 2] = v.z; fn[kc * 4 + 3] = v.w;
      }
    }
    tmem_st16(t_a + lane_off, fn);
    const bool has_next = MODE == 2 && tile + NET_GROUPS < t1;   // uniform over the group
    if (has_next) {
      const int in = i + NET_GROUPS * NRF_TILE;
      sn = in < n ? sample_ld(samples + in) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    tmem_wait_st();
    // ---- density layer 0: 32 -> 64 (two N=32 halves), ReLU ---------------------------------------------------------
    GROUP_SYNC();
    ISSUE(32, 32, 64, W0_OFF, 0);
    if (has_next) gather_slice<0, 4, false, LMODES>(P, sn, fn, nullptr, stab0, stab1);
    uint32_t h[32];
    MMA_WAIT();
    ld_relu_pack32(t_d + lane_off, h);
    GROUP_SYNC();
    ISSUE(32, 32, 64, W0_OFF, 32);
    MMA_WAIT();
    ld_relu_pack32(t_d + lane_off, h + 16);
    // ---- density layer 1: 64 -> 16, no activation ------------------------------------------------------------------
    tmem_st16(t_a + lane_off, h);
    tmem_st16(t_a + 16 + lane_off, h + 16);
    tmem_wait_st();
    GROUP_SYNC();
    ISSUE(16, 64, 16, W1_OFF, 0);
    if (has_next) gather_slice<4, SL1, false, LMODES>(P, sn, fn, nullptr);
    // SH coefficients of the direction (per ray)
    uint32_t sh[8];
    if (dirs_mode == 0) {
      uint4 a = make_uint4(0, 0, 0, 0), b = a;
      if (valid) {
        // frame path: the ray index rides in the sample record's fourth component (SampleSrc)
        const int ray = S.t0 ? __float_as_int(__ldg(&samples[i].w)) : __ldg(sample_ray + i);
        a = __ldg(ray_sh + 2 * (size_t)ray);
        b = __ldg(ray_sh + 2 * (size_t)ray + 1);
      }
      sh[0] = a.x; sh[1] = a.y; sh[2] = a.z; sh[3] = a.w; sh[4] = b.x; sh[5] = b.y; sh[6] = b.z; sh[7] = b.w;
    } else {
      float dx = 0.5f, dy = 0.5f, dz = 0.5f;
      if (valid) { const float4 r = __ldg(rays + i); dx = r.x; dy = r.y; dz = r.z; }
      sh4(dx, dy, dz, sh);
    }
    MMA_WAIT();
    __half sigma;
    {
      uint32_t a[16];
      tmem_ld16(t_d + lane_off, a);
      tmem_wait_ld();
#pragma unroll
      for (int k = 0; k < 8; ++k)
        h[k] = pack_h2(__float2half_rn(__uint_as_float(a[2 * k])), __float2half_rn(__uint_as_float(a[2 * k + 1])));
      sigma = act_h(P.sigma_act, __ushort_as_half((unsigned short)(h[0] & 0xffffu)));
#pragma unroll
      for (int k = 0; k < 8; ++k) h[8 + k] = sh[k];
    }
    // ---- rgb layer 0: [den16 | sh16] 32 -> 64, ReLU ----------------------------------------------------------------
    tmem_st16(t_a + lane_off, h);
    tmem_wait_st();
    GROUP_SYNC();
    ISSUE(32, 32, 64, V0_OFF, 0);
    if (has_next) gather_slice<SL1, SL2, false, LMODES>(P, sn, fn, nullptr);
    MMA_WAIT();
    ld_relu_pack32(t_d + lane_off, h);
    GROUP_SYNC();
    ISSUE(32, 32, 64, V0_OFF, 32);
    MMA_WAIT();
    ld_relu_pack32(t_d + lane_off, h + 16);
    // ---- rgb layer 1: 64 -> 64, ReLU -------------------------------------------------------------------------------
    tmem_st16(t_a + lane_off, h);
    tmem_st16(t_a + 16 + lane_off, h + 16);
    tmem_wait_st();
    GROUP_SYNC();
    ISSUE(32, 64, 64, V1_OFF, 0);
    if (has_next) gather_slice<SL2, SL3, false, LMODES>(P, sn, fn, nullptr);
    MMA_WAIT();
    ld_relu_pack32(t_d + lane_off, h);
    GROUP_SYNC();
    ISSUE(32, 64, 64, V1_OFF, 32);
    MMA_WAIT();
    ld_relu_pack32(t_d + lane_off, h + 16);
    // ---- rgb layer 2: 64 -> 16 (3 used), output activation ---------------------------------------------------------
    tmem_st16(t_a + lane_off, h);
    tmem_st16(t_a + 16 + lane_off, h + 16);
    tmem_wait_st();
    GROUP_SYNC();
    ISSUE(16, 64, 16, V2_OFF, 0);
    if (has_next) gather_slice<SL3, 16, false, LMODES>(P, sn, fn, nullptr);
    MMA_WAIT();
    {
      uint32_t a[16];
      tmem_ld16(t_d + lane_off, a);
      tmem_wait_ld();
      if (valid) {
        const __half r = act_h(P.rgb_act, __float2half_rn(__uint_as_float(a[0])));
        const __half gg = act_h(P.rgb_act, __float2half_rn(__uint_as_float(a[1])));
        const __half b = act_h(P.rgb_act, __float2half_rn(__uint_as_float(a[2])));
        out_st(out + i, make_uint2(pack_h2(r, gg), pack_h2(b, sigma)));
      }
    }
#if !NRF_NO_TAIL_SYNC && !NRF_ARRIVE_SYNC
    // the next tile's first tcgen05.st / MMA of this group must not overtake these TMEM reads
    GROUP_SYNC();
#endif
    // (without it: a warp's next tcgen05.st touches only its own 32 lanes of the A columns, which no MMA reads any more — the
    // tile's last MMA was waited for above — and the next MMA is issued behind the next tile's first group barrier, which every
    // warp reaches only after its tcgen05.wait::ld)
  }
#undef GROUP_SYNC
#undef ISSUE
#undef MMA_WAIT

  tc_fence_before();
  __syncthreads();
  if (warp == 0)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
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

// Load-time: per-level scale and resolution exactly as the reference's kernel_grid derives them ON THE DEVICE
// (grid.h:186-187): scale = exp2f(level * log2_per_level_scale) * base_resolution - 1 (FMUL, MUFU.EX2, FFMA in its SASS),
// resolution = (uint32_t)ceil(scale) + 1.  The host's exp2f differs from MUFU.EX2 by an ulp on some levels.
extern "C" __global__ void k_level_scales(int n_levels, float log2_per_level_scale, uint32_t base_resolution,
                                          float* __restrict__ scales, uint32_t* __restrict__ resolutions) {
  const int level = threadIdx.x;
  if (level >= n_levels) return;
  const float scale = __fmaf_rn(exp2f((float)level * log2_per_level_scale), (float)base_resolution, -1.0f);
  scales[level] = scale;
  resolutions[level] = (uint32_t)ceilf(scale) + 1u;
}

// Stage probe (nrf_probe_sh): the 16 SH coefficients of n normalised directions, as the frame path's k_rays_count stores them.
extern "C" __global__ void k_probe_sh(int n, const float4* __restrict__ dirs01, uint4* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 d = dirs01[i];
  uint32_t sh[8];
  sh4(d.x, d.y, d.z, sh);
  out[2 * (size_t)i] = make_uint4(sh[0], sh[1], sh[2], sh[3]);
  out[2 * (size_t)i + 1] = make_uint4(sh[4], sh[5], sh[6], sh[7]);
}

// ------------------------------------------------------------------------------------------------ host launchers
template <int MODE, int G, uint32_t LM = NRF_LMODES_RUNTIME>
static void net_attr() {
  // 20.6 KB of shared memory per SM; leave the rest of the unified array to L1 (carveout in percent of 228 KB)
  cudaFuncSetAttribute(k_network<MODE, G, LM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)NET_SMEM_LAUNCH);
  static const int carveout = getenv("NRF_NET_CARVEOUT") ? atoi(getenv("NRF_NET_CARVEOUT")) : (NRF_EXP_SMEM_L01 ? 40 : 16);
  cudaFuncSetAttribute(k_network<MODE, G, LM>, cudaFuncAttributePreferredSharedMemoryCarveout, carveout);
}
static bool g_specialise = true;   // NRF_NO_SPECIALISE=1: always the generic kernel (A/B, tests of the generic path at full size)
static int g_net_groups = 7;   // measured on B200 (C3): 7 groups = 896 threads x 72 registers, no spills: 3-4 % faster than 8 x 64 (spills), 6 is slower again
// Function attributes (shared-memory size, carveout preference) belong to the CURRENT device: set once per device, and safely
// when several host threads enqueue the first frames of a multi-device context at the same time.
static void nrf_network_setup() {
  static std::mutex m;
  static bool attr_set[64] = {};
  static bool env_read = false;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
  std::lock_guard<std::mutex> l(m);
  if (!env_read) {
    g_specialise = getenv("NRF_NO_SPECIALISE") == nullptr;
    if (const char* e = getenv("NRF_NET_GROUPS")) {   // experiment: fewer groups = more registers per thread
      const int g = atoi(e);
      if (g == 6 || g == 7 || g == 8) g_net_groups = g;
    }
    env_read = true;
  }
  if (attr_set[dev]) return;
  net_attr<0, 8>(); net_attr<1, 8>(); net_attr<2, 8>(); net_attr<2, 7>(); net_attr<2, 6>(); net_attr<2, 7, NRF_LMODES_T19>(); net_attr<2, 7, NRF_LMODES_T19_B16>(); net_attr<2, 7, NRF_LMODES_T24>();
  attr_set[dev] = true;
}

// grid = number of persistent CTAs (<= SM count); each walks a contiguous chunk of ceil(tiles/grid) tiles.
extern "C" cudaError_t nrf_launch_network(int mode, int grid, cudaStream_t stream, const NetParams* P, int n,
                                          const float4* samples, const int* sample_ray, const float4* rays,
                                          const uint4* ray_sh, int dirs_mode, uint4* enc_io, uint2* out,
                                          uint32_t* idx_out, const unsigned long long* n_dev, const SampleSrc* src) {
  nrf_network_setup();
  SampleSrc S{};
  if (src) S = *src;
  if (n <= 0) return cudaSuccess;
#define NET_ARGS *P, n, samples, sample_ray, rays, ray_sh, dirs_mode, enc_io, out, idx_out, n_dev, S
  switch (mode) {
    case 0: k_network<0, 8><<<grid, 1024, NET_SMEM_LAUNCH, stream>>>(NET_ARGS); break;
    case 1: k_network<1, 8><<<grid, 1024, NET_SMEM_LAUNCH, stream>>>(NET_ARGS); break;
    default:
      if (g_net_groups == 7) {
        // the level layout of this model as 2 bits per level; the layout of the T = 2^19 models has its own instantiation
        // (measured: network C2 5.11 -> 4.88 ms, C3 8.97 -> 8.61, C4 23.8 -> 22.9; half the code: 4 992 instead of 9 384 SASS instructions)
        uint32_t lm = 0;
        for (int l = 0; l < NRF_L; ++l) lm |= (P->lv[l].hashed & 3u) << (2 * l);
        if (g_specialise && lm == NRF_LMODES_T19) k_network<2, 7, NRF_LMODES_T19><<<grid, 896, NET_SMEM_LAUNCH, stream>>>(NET_ARGS);
        else if (g_specialise && lm == NRF_LMODES_T19_B16) k_network<2, 7, NRF_LMODES_T19_B16><<<grid, 896, NET_SMEM_LAUNCH, stream>>>(NET_ARGS);
        else if (g_specialise && lm == NRF_LMODES_T24) k_network<2, 7, NRF_LMODES_T24><<<grid, 896, NET_SMEM_LAUNCH, stream>>>(NET_ARGS);
        else k_network<2, 7><<<grid, 896, NET_SMEM_LAUNCH, stream>>>(NET_ARGS);
      }
      else if (g_net_groups == 6) k_network<2, 6><<<grid, 768, NET_SMEM_LAUNCH, stream>>>(NET_ARGS);
      else k_network<2, 8><<<grid, 1024, NET_SMEM_LAUNCH, stream>>>(NET_ARGS);
      break;
  }
#undef NET_ARGS
  return cudaGetLastError();
}

extern "C" cudaError_t nrf_launch_network_simt(cudaStream_t stream, const NetParams* P, int n, const uint4* enc,
                                               const float4* dirs01, uint2* out) {
  if (n <= 0) return cudaSuccess;
  k_network_simt<<<(n + 127) / 128, 128, 0, stream>>>(*P, n, enc, dirs01, out);
  return cudaGetLastError();
}
