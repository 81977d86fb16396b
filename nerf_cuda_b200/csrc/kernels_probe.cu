// kernels_probe.cu — memory micro-benchmarks that give the hash-gather roofline its denominators (SURVEY.md §8d:
// "L2 bandwidth measured on the box").  Not on the frame path; called through nrf_bench_memory (nerf_b200_ext.h).
//
//   mode 0  streaming read   : every thread reads coalesced 16-byte words (L1 bypassed) over a window; a window that fits the 126 MB
//                              L2 measures L2->SM bandwidth, a window far larger than L2 measures HBM read bandwidth.
//   mode 2  streaming read, every CTA over the WHOLE window (rotated start): the per-SM footprint is the window itself, so
//                              a 24 MB window cannot be served by the SM's own L1 (mode 0 splits the window over the
//                              CTAs — 24 MB / 148 = 162 KB per SM, which fits L1, and ld.cg did not keep it out: the
//                              round-1 log shows 28 TB/s for 24 MB against 12.9 TB/s for 64 MB).  This is the L2->SM figure.
//   mode 1  sector gather    : every lane reads ONE 8-byte slot at a pseudo-random position of the window, i.e. one
//                              32-byte sector in its own 128-byte line per lane — the access shape of a hashed level of
//                              the grid (kernels_network.cu level_load: one LDG.64 per (y,z) corner pair).  It is bound by
//                              L1 wavefronts (one 128-byte line per cycle per SM), not by bytes.
#include <cuda_runtime.h>
#include <stdint.h>

#define PROBE_THREADS 1024

__global__ void __launch_bounds__(PROBE_THREADS, 1)
k_probe_stream(const uint4* __restrict__ buf, size_t n_vec, int passes, uint32_t* __restrict__ sink) {
  uint32_t acc = 0;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (int p = 0; p < passes; ++p) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
#pragma unroll 4
    for (; i < n_vec; i += stride) {
      const uint4 v = __ldcg(buf + i);  // ld.global.cg: L2 only, so a window that fits L1 still measures L2
      acc ^= v.x ^ v.y ^ v.z ^ v.w;
    }
  }
  if (acc == 0x9e3779b9u) sink[0] = acc;  // never true for the zero-filled buffer; keeps the loads alive
}

__global__ void __launch_bounds__(PROBE_THREADS, 1)
k_probe_stream_whole(const uint4* __restrict__ buf, size_t n_vec, int passes, uint32_t* __restrict__ sink) {
  uint32_t acc = 0;
  const size_t start = (n_vec / gridDim.x) * blockIdx.x;   // rotated start: the CTAs do not hit the same L2 lines in lockstep
  for (int p = 0; p < passes; ++p) {
#pragma unroll 4
    for (size_t k = threadIdx.x; k < n_vec; k += PROBE_THREADS) {
      size_t i = start + k;
      if (i >= n_vec) i -= n_vec;
      const uint4 v = __ldcg(buf + i);
      acc ^= v.x ^ v.y ^ v.z ^ v.w;
    }
  }
  if (acc == 0x9e3779b9u) sink[0] = acc;
}

__device__ __forceinline__ uint32_t mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}

__global__ void __launch_bounds__(PROBE_THREADS, 1)
k_probe_gather(const uint2* __restrict__ buf, uint32_t slot_mask, int loads_per_thread, uint32_t* __restrict__ sink) {
  uint32_t acc = 0;
  uint32_t s = mix32((blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u);
  for (int k = 0; k < loads_per_thread; k += 8) {
    uint2 v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {  // 8 independent loads in flight per thread, like one grid level
      s = s * 1664525u + 1013904223u;
      v[j] = __ldg(buf + (mix32(s) & slot_mask));
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) acc ^= v[j].x ^ v[j].y;
  }
  if (acc == 0x9e3779b9u) sink[0] = acc;
}

extern "C" cudaError_t nrf_launch_probe(int mode, int sms, cudaStream_t stream, const void* buf, size_t bytes, int work,
                                        uint32_t* sink) {
  if (mode == 2) {
    k_probe_stream_whole<<<sms, PROBE_THREADS, 0, stream>>>(reinterpret_cast<const uint4*>(buf), bytes / 16, work, sink);
  } else if (mode == 0) {
    k_probe_stream<<<sms, PROBE_THREADS, 0, stream>>>(reinterpret_cast<const uint4*>(buf), bytes / 16, work, sink);
  } else {
    uint32_t slots = 1;
    while ((size_t)slots * 2 * 8 <= bytes) slots *= 2;  // largest power of two of 8-byte slots inside the window
    k_probe_gather<<<sms, PROBE_THREADS, 0, stream>>>(reinterpret_cast<const uint2*>(buf), slots - 1, work, sink);
  }
  return cudaGetLastError();
}
