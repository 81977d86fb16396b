// kernels_composite.cu — front-to-back alpha compositing with early termination, finalisation and u8 packing.
//
// Replaces kernel_composite_rays (include/nerf-cuda/render_utils.h:658-751), get_image_and_depth (:241-265) and the
// host-side u8 cast loop (src/nerf_render.cu:352-358).
//
// One THREAD per ray, one WARP per 8x4-pixel patch, the warp walking the patch's step-interleaved sample list in
// lockstep: at step k the live lanes' samples are contiguous (k_emit's layout), so every load is one coalesced
// segment, and the position of a lane's sample is a ballot + popc — the warp-cooperative part.  Each lane accumulates
// its own ray sequentially in registers with exactly the reference's operations and order (T = 1 - sum(w), add the
// sample, THEN stop if T < 1e-4, render_utils.h:717-727), so colour/depth differ from the oracle only through
// __expf (ex2.approx) versus libm expf.
#include "nrf_device.cuh"

#define COMP_CHUNK 8

// Evaluation rounds (termination-aware frames, RoundWin R): this launch composites the samples with step index in
// [RW.s0, RW.s1) — the list k_emit wrote for the same window — and carries each ray's accumulators between rounds in
// RW.st (two float4 per ray: {wsum, depth, t, last_t}, {r, g, b, samples used}).  A ray whose transmittance fell below
// 1e-4 is marked in RW.dead: the next round neither marches, evaluates nor composites it — the reference drops such a
// ray from its next wavefront the same way (render_utils.h:727 rays_t = -1, kernel_compact_rays :394-415).  The
// operations on a ray and their order do not depend on the round schedule, so every schedule gives identical pixels.
// RW.first: start from fresh accumulators; RW.last: finalise and write the pixels.  A whole-ray frame is one round with both.
extern "C" __global__ void __launch_bounds__(128)
k_composite(int n_rays, ShardMap map, const int* __restrict__ warp_base, const int* __restrict__ counts,
            const float4* __restrict__ rays, const float2* __restrict__ far_tf, const __grid_constant__ SampleSrc S,
            const uint2* __restrict__ sample_out, float* __restrict__ rgb_f32,
            float* __restrict__ depth_f32, int* __restrict__ used_counts, uint8_t* __restrict__ rgb_u8,
            uint8_t* __restrict__ depth_u8, int u8_pixel_order, const __grid_constant__ RoundWin RW,
            unsigned long long* __restrict__ used_total) {
  const int lane = threadIdx.x & 31;
  const int q = blockIdx.x * blockDim.x + threadIdx.x;   // n_rays is a multiple of 32: whole warps only
  if (q >= n_rays) return;
  const int wb = warp_base[q >> 5];
  const bool was_dead = !RW.first && RW.dead[q];
  // wb < 0: the patch did not fit the sample buffers (frame will be redone)
  const int cnt = (wb < 0 || was_dead) ? 0 : max(0, min(counts[q], RW.s1) - RW.s0);
  const float4 ray = rays[q];
  const float near = ray.w, far = far_tf[q].x;
  int cmax = cnt;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) cmax = max(cmax, __shfl_xor_sync(0xffffffffu, cmax, o));
  const uint32_t lt = (1u << lane) - 1u;
  long long off = wb;
  float wsum = 0.f, dep = 0.f, cr = 0.f, cg = 0.f, cb = 0.f;
  float tc = near, last_t = near;
  int used = 0;
  bool running = !was_dead;
  if (!RW.first) {
    const float4 a = RW.st[2 * (size_t)q], b = RW.st[2 * (size_t)q + 1];
    wsum = a.x; dep = a.y; tc = a.z; last_t = a.w;
    cr = b.x; cg = b.y; cb = b.z; used = __float_as_int(b.w);
  }
  const int used0 = used;
  // The steps of a ray are sequential, but WHERE its samples live is not data-dependent (ballot/popc of the counts), so
  // the loop runs in chunks of COMP_CHUNK steps: all indices first, then all loads in flight at once, then the
  // sequential accumulation.  Without this the kernel's critical path is one DRAM latency per step of its longest ray
  // (1024 steps = 0.28 ms), which is what one frame shard took on 8 GPUs.
  for (int k0 = 0; k0 < cmax; k0 += COMP_CHUNK) {
    long long idx[COMP_CHUNK];
    bool has[COMP_CHUNK];
#pragma unroll
    for (int j = 0; j < COMP_CHUNK; ++j) {
      has[j] = k0 + j < cnt;
      const uint32_t alive = __ballot_sync(0xffffffffu, has[j]);
      idx[j] = off + __popc(alive & lt);
      off += __popc(alive);
    }
    uint2 so[COMP_CHUNK];
    float dts[COMP_CHUNK], tas[COMP_CHUNK];
#pragma unroll
    for (int j = 0; j < COMP_CHUNK; ++j) {
      if (has[j] && running) {
        so[j] = __ldg(sample_out + idx[j]);
        tas[j] = __ldg(S.t0 + idx[j]);   // the t the sample was taken at
      }
    }
#pragma unroll
    for (int j = 0; j < COMP_CHUNK; ++j) {  // dt and t_after as the march step computed them (record_dt)
      dts[j] = record_dt(S, tas[j]);
      tas[j] += dts[j];
    }
#pragma unroll
    for (int j = 0; j < COMP_CHUNK; ++j) {
      if (has[j] && running) {
        const __half2 rg = *reinterpret_cast<const __half2*>(&so[j].x);
        const __half2 bs = *reinterpret_cast<const __half2*>(&so[j].y);
        const float alpha = 1.0f - __expf(-__high2float(bs) * dts[j]);
        const float T = 1.0f - wsum;
        const float w = alpha * T;
        wsum += w;
        tc += tas[j] - last_t;   // deltas(1) = t - last_t (render_utils.h:629-631), accumulated like the reference
        last_t = tas[j];
        dep += w * tc;
        cr += w * __low2float(rg);
        cg += w * __high2float(rg);
        cb += w * __low2float(bs);
        ++used;
        if ((double)T < 1e-4) running = false;   // the reference compares the float against a double literal
      }
    }
    if (!__any_sync(0xffffffffu, running && k0 + COMP_CHUNK < cnt)) break;
  }
  if (used_total) {   // samples composited by this launch (one atomic per warp that used any)
    const int wu = __reduce_add_sync(0xffffffffu, used - used0);
    if (lane == 0 && wu) atomicAdd(used_total, (unsigned long long)wu);
  }
  if (!RW.last) {
    RW.st[2 * (size_t)q] = make_float4(wsum, dep, tc, last_t);
    RW.st[2 * (size_t)q + 1] = make_float4(cr, cg, cb, __int_as_float(used));
    RW.dead[q] = running ? 0 : 1;
    return;
  }
  const float bg = 1.0f;  // m_bg_color, nerf_render.h:74
  const float R = cr + (1.0f - wsum) * bg, G = cg + (1.0f - wsum) * bg, B = cb + (1.0f - wsum) * bg;
  // defect D3: missed rays (near == far == FLT_MAX) give 0/0 in the reference; defined as 0 here and in the oracle
  const float D = (far == near) ? 0.0f : fmaxf(dep - near, 0.0f) / (far - near);
  rgb_f32[q * 3 + 0] = R; rgb_f32[q * 3 + 1] = G; rgb_f32[q * 3 + 2] = B;
  depth_f32[q] = D;
  used_counts[q] = used;
  // (unsigned char)(255.0 * v): double multiply + truncation (nerf_render.cu:355-358); clamped where that is UB (D4)
  int px, py;
  const bool inside = shard_pixel_xy(map, q, px, py);
  const float v4[4] = {R, G, B, D};
  uint32_t u[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const double x = 255.0 * (double)v4[k];
    u[k] = !(x >= 0.0) ? 0u : (x >= 255.0 ? 255u : (uint32_t)(int)x);
  }
  // Word-packed stores.  A patch row is 8 consecutive pixels = 24 rgb bytes + 8 depth bytes, 4-byte aligned whenever the
  // row stride is (always in the packed tile layout; in pixel order when W % 4 == 0): lanes 0-5 of the row store the six
  // rgb words, lanes 6-7 the two depth words, the bytes collected with warp shuffles.  32 bytes per row in 8 stores
  // instead of 32 single-byte stores — these stores cross NVLink when the frame lives on another GPU.
  const int col = lane & 7, row0 = lane & 24;
  const uint32_t rgb24 = inside ? (u[0] | (u[1] << 8) | (u[2] << 16)) : 0u;
  const uint32_t dep8 = inside ? u[3] : 0u;
  const int c6 = min(col, 5), p0 = (4 * c6) / 3;                      // first pixel contributing to rgb word c6
  const uint32_t lo = __shfl_sync(0xffffffffu, rgb24, row0 + p0), hi = __shfl_sync(0xffffffffu, rgb24, row0 + min(p0 + 1, 7));
  const uint32_t rgb_word = (uint32_t)((((unsigned long long)hi << 24) | lo) >> (8 * (4 * c6 - 3 * p0)));
  const int d4 = (col & 1) * 4;
  uint32_t dep_word = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) dep_word |= __shfl_sync(0xffffffffu, dep8, row0 + d4 + k) << (8 * k);
  // the row's first pixel decides: whole row inside the image (or packed layout, where outside slots exist and hold 0)
  const int px0 = __shfl_sync(0xffffffffu, px, row0), py0 = __shfl_sync(0xffffffffu, py, row0);
  const bool row_inside = px0 + 8 <= map.W && py0 < map.H;
  const bool vec = u8_pixel_order ? (row_inside && (map.W & 3) == 0) : true;
  if (vec) {
    const long long base = u8_pixel_order ? (long long)py0 * map.W + px0 : (long long)shard_packed_index(q - col);
    if (col < 6) reinterpret_cast<uint32_t*>(rgb_u8 + base * 3)[col] = rgb_word;
    else reinterpret_cast<uint32_t*>(depth_u8 + base)[col - 6] = dep_word;
  } else if (inside) {
    const long long dst = (long long)py * map.W + px;
    rgb_u8[dst * 3 + 0] = (uint8_t)u[0]; rgb_u8[dst * 3 + 1] = (uint8_t)u[1]; rgb_u8[dst * 3 + 2] = (uint8_t)u[2];
    depth_u8[dst] = (uint8_t)u[3];
  }
}
