// kernels_post.cu — frame post-processing on linear device buffers: progressive accumulation and tonemapping.
//
// Per-pixel math of the reference's render buffer (src/render_buffer.cu): accumulate_kernel :224-258, tonemap :260-338,
// tonemap_kernel :529-555, with linear_to_srgb / srgb_to_linear of include/nerf-cuda/common_device.cuh:38-60.  The
// reference writes the result into a CUDA surface bound to a GL texture (or a DLSS/Vulkan image); that plumbing needs a
// graphics stack and is left out — here input, accumulator and output are plain float4 arrays in HBM.
// One thread per pixel, one 16-byte load/store per buffer: a pure HBM stream (48 B/pixel for accumulate, 32 B/pixel
// for tonemap).  Built with -fmad=false like the other exact kernels so the arithmetic is the expression tree as written.
#include <cuda_runtime.h>
#include <stdint.h>

#define CS_LINEAR 0
#define CS_SRGB 1
#define CS_VISPOSNEG 2
#define TM_IDENTITY 0
#define TM_ACES 1
#define TM_HABLE 2
#define TM_REINHARD 3

__device__ __forceinline__ float srgb_to_linear(float s) { return s <= 0.04045f ? s / 12.92f : powf((s + 0.055f) / 1.055f, 2.4f); }
__device__ __forceinline__ float linear_to_srgb(float l) { return l < 0.0031308f ? 12.92f * l : 1.055f * powf(l, 0.41666f) - 0.055f; }

// accumulate_kernel, render_buffer.cu:224-258: running mean over sample_count previous frames
extern "C" __global__ void k_accumulate(int n, const float4* __restrict__ frame, float4* __restrict__ accum, float sample_count,
                                        int color_space) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float4 c = frame[i], t = accum[i];
  if (color_space == CS_VISPOSNEG) {
    const float val = c.x - c.y;
    float tv = t.x - t.y;
    tv = (tv * sample_count + val) / (sample_count + 1);
    t.x = fmaxf(tv, 0.0f);
    t.y = fmaxf(-tv, 0.0f);
  } else {
    if (color_space == CS_SRGB) { c.x = linear_to_srgb(c.x); c.y = linear_to_srgb(c.y); c.z = linear_to_srgb(c.z); }
    t.x = (t.x * sample_count + c.x) / (sample_count + 1);
    t.y = (t.y * sample_count + c.y) / (sample_count + 1);
    t.z = (t.z * sample_count + c.z) / (sample_count + 1);
  }
  t.w = (t.w * sample_count + c.w) / (sample_count + 1);
  accum[i] = t;
}

// tonemap(x, curve), render_buffer.cu:260-318
__device__ __forceinline__ void tonemap_curve(float& r, float& g, float& b, int curve) {
  if (curve == TM_IDENTITY) return;
  r = fmaxf(r, 0.f); g = fmaxf(g, 0.f); b = fmaxf(b, 0.f);
  float k0, k1, k2, k3, k4, k5;
  if (curve == TM_ACES) {
    k0 = 0.6f * 0.6f * 2.51f; k1 = 0.6f * 0.03f; k2 = 0.0f; k3 = 0.6f * 0.6f * 2.43f; k4 = 0.6f * 0.59f; k5 = 0.14f;
  } else if (curve == TM_HABLE) {
    const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
    k0 = A * F - A * E; k1 = C * B * F - B * E; k2 = 0.0f; k3 = A * F; k4 = B * F; k5 = D * F * F;
    const float W = 11.2f;
    const float nom = k0 * (W * W) + k1 * W + k2, denom = k3 * (W * W) + k4 * W + k5;
    const float white_scale = denom / nom;
    k0 = 4.0f * k0 * white_scale; k1 = 2.0f * k1 * white_scale; k2 = k2 * white_scale; k3 = 4.0f * k3; k4 = 2.0f * k4;
  } else {  // Reinhard
    const float Y = 0.2126f * r + (0.7152f * g + 0.0722f * b);  // Eigen fixed-size-3 reduction order: a0 + (a1 + a2)
    const float s = 1.f / (Y + 1.0f);
    r *= s; g *= s; b *= s;
    return;
  }
  float* ch[3] = {&r, &g, &b};
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const float x = *ch[k], sq = x * x;
    const float nom = sq * k0 + k1 * x + k2, denom = k3 * sq + k4 * x + k5;
    *ch[k] = nom / denom;
  }
}

// tonemap_kernel, render_buffer.cu:529-555 (+ the 5-argument tonemap :320-338): background blend, exposure, curve,
// output colour space, optional clamp.  background is given in sRGB (:539-543).
extern "C" __global__ void k_tonemap(int n, const float4* __restrict__ accum, float exposure, float4 background, int color_space,
                                     int output_color_space, int curve, int clamp_output, float4* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (color_space != CS_SRGB) {
    background.x = srgb_to_linear(background.x); background.y = srgb_to_linear(background.y); background.z = srgb_to_linear(background.z);
  }
  float4 c = accum[i];
  const float weight = (1 - c.w) * background.w;
  c.x += background.x * weight; c.y += background.y * weight; c.z += background.z * weight;
  c.w += weight;
  if (color_space == CS_SRGB) { c.x = srgb_to_linear(c.x); c.y = srgb_to_linear(c.y); c.z = srgb_to_linear(c.z); }
  const float e = powf(2.0f, exposure);
  c.x *= e; c.y *= e; c.z *= e;
  tonemap_curve(c.x, c.y, c.z, curve);
  if (output_color_space == CS_SRGB) { c.x = linear_to_srgb(c.x); c.y = linear_to_srgb(c.y); c.z = linear_to_srgb(c.z); }
  if (clamp_output) {
    c.x = fminf(fmaxf(c.x, 0.0f), 1.0f); c.y = fminf(fmaxf(c.y, 0.0f), 1.0f);
    c.z = fminf(fmaxf(c.z, 0.0f), 1.0f); c.w = fminf(fmaxf(c.w, 0.0f), 1.0f);
  }
  out[i] = c;
}
