// nerf_oracle.cpp — CPU restatement of the reference's per-frame hot path.  TEST INFRASTRUCTURE ONLY.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this
// library.  The product (nerf_cuda_b200/csrc) never links, imports or calls it; it has no CPU path.
//
// Parity status: PINNED BY EXECUTING REFERENCE CODE.  The reference ships no golden vectors;
// every stage here except the MLP is checked bit-exactly against the reference's own kernels compiled for the
// host (oracle/_ref, built by oracle/build_ref.sh from /root/reference in the dev container) and against the
// fixtures those produced (tests/golden/*.npz).  The MLP cannot run on the host (wmma fragments); it follows
// the layout and wiring cited below, accumulates in fp32 and rounds to fp16 at every layer boundary.
//
// Build: g++ -O2 -ffp-contract=off -fopenmp -shared -fPIC  (no IMPLICIT contraction: every fused multiply-add below
// is an explicit fmaf; see DESIGN.md "numerics contract").
//
// Two numerics modes (orc_set_numerics):
//   host (default): expression trees as the reference WRITES them, no contraction — what its kernels give when compiled
//                   for the host; pinned bit-exactly by oracle/_ref and tests/golden/ref_tiny*.npz.
//   cuda          : as nvcc COMPILES them for the GPU (default -fmad=true): a*b+c contracted into one fma exactly where
//                   the reference's sm_100 SASS has an FFMA (read off `cuobjdump -sass baseline/_ref/ref_cuda_driver`:
//                   set_rays_d, kernel_march_rays, kernel_grid, kernel_sh, linear_transformer), and the per-level grid
//                   scale taken from the DEVICE's exp2f (MUFU.EX2; supplied by the caller, it cannot be computed here).
//                   Pinned bit-exactly by an execution of the reference's CUDA build on a B200
//                   (tests/golden/ref_cuda_*.npz, baseline/run_ref_cuda.py).  This is what the product follows.
//
// Every function cites the reference file:line it restates (paths relative to /root/reference,
// tcnn/ = dependencies/tiny-cuda-nn/).
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/nerf_b200.h"

namespace {

// ---- fp16 helpers (IEEE binary16, round-to-nearest-even; same as __half on host, cuda_fp16.hpp) ----------
typedef _Float16 half_t;
inline uint16_t h2bits(half_t h) { uint16_t b; std::memcpy(&b, &h, 2); return b; }
inline half_t bits2h(uint16_t b) { half_t h; std::memcpy(&h, &b, 2); return h; }
inline half_t f2h(float f) { return (half_t)f; }
inline float h2f(half_t h) { return (float)h; }
// __half + __half: exact sum rounded once to fp16 (float add of two halves then RN to half is innocuous
// double rounding because 24 >= 2*11+2).
inline half_t hadd(half_t a, half_t b) { return f2h(h2f(a) + h2f(b)); }

inline float clampf(float x, float lo, float hi) { return fminf(hi, fmaxf(lo, x)); }  // render_utils.h:141-144

// numerics mode (see header): g_fma = contract like nvcc; g_scales = per-level grid scales as the device computes them
bool g_fma = false;
bool g_have_scales = false;
float g_scales[NRF_MAX_LEVELS];
// a*b + c: one rounding where nvcc contracts, two as written otherwise
inline float mad(float a, float b, float c) { return g_fma ? fmaf(a, b, c) : a * b + c; }

struct Levels {
  uint32_t offsets[NRF_MAX_LEVELS + 1];
  float scales[NRF_MAX_LEVELS];
  uint32_t res[NRF_MAX_LEVELS];
};

inline uint32_t powi_u32(uint32_t base, uint32_t e) {
  uint32_t r = 1;
  for (uint32_t i = 0; i < e; ++i) r *= base;
  return r;
}

// nerf_render.cu:147-165 (defaults) + tcnn/include/tiny-cuda-nn/encodings/grid.h:899-931 (offset table ctor)
// + grid.h:186-187 (per-level scale / resolution inside kernel_grid).
int build_levels(const nrf_model_desc& d, Levels& lv, float* pls_out) {
  if (d.n_levels == 0 || d.n_levels > NRF_MAX_LEVELS) return NRF_EINVAL;
  uint32_t base = d.base_resolution;
  if (!base) base = 1u << (d.log2_hashmap_size / 3);
  float pls = d.per_level_scale;
  if (pls <= 0.0f && d.n_levels > 1) {
    pls = std::exp(std::log(2048.0f * (float)d.bound / (float)base) / (d.n_levels - 1));
  }
  if (pls_out) *pls_out = pls;
  const float log2_pls = std::log2(pls);
  uint32_t offset = 0;
  for (uint32_t i = 0; i < d.n_levels; ++i) {
    const float scale = exp2f(i * log2_pls) * base - 1.0f;
    const uint32_t resolution = (uint32_t)(ceilf(scale)) + 1;
    const uint32_t max_params = std::numeric_limits<uint32_t>::max() / 2;
    uint32_t params_in_level =
        std::pow((float)resolution, 3) > (float)max_params ? max_params : powi_u32(resolution, 3);
    params_in_level = ((params_in_level + 7u) / 8u) * 8u;
    params_in_level = std::min(params_in_level, (1u << d.log2_hashmap_size));
    lv.offsets[i] = offset;   // grid.h:899-931: the offset table is built on the HOST, with the host's exp2f
    lv.scales[i] = scale;
    lv.res[i] = resolution;
    if (g_have_scales) {      // grid.h:186-187: scale and resolution are re-derived inside the kernel, on the DEVICE
      lv.scales[i] = g_scales[i];
      lv.res[i] = (uint32_t)(ceilf(g_scales[i])) + 1;
    }
    offset += params_in_level;
  }
  lv.offsets[d.n_levels] = offset;
  return NRF_OK;
}

int check_arch(const nrf_model_desc& d) {
  if (d.n_levels != 16 || d.n_features_per_level != 2 || d.n_neurons != 64 || d.density_hidden_layers != 1 ||
      d.rgb_hidden_layers != 2 || d.sh_degree != 4)
    return NRF_EINVAL;
  return NRF_OK;
}

// ---- ray generation: render_utils.h:31-77, nerf_render.cu:369-386 ----------------------------------------
struct Frame {
  float R[3][3];
  float o[3];
};

// render_utils.h:68-77 nerf_matrix_to_ngp (offset = 0): rows (1,2,0), columns 1,2 negated, translation*scale.
void pose_to_ngp(const float* P /*row-major 4x4*/, float scale, Frame& f) {
  const int rows[3] = {1, 2, 0};
  for (int r = 0; r < 3; ++r) {
    const float* src = P + rows[r] * 4;
    f.R[r][0] = src[0];
    f.R[r][1] = -src[1];
    f.R[r][2] = -src[2];
    f.o[r] = src[3] * scale + 0.0f;
  }
}

// set_rays_d, render_utils.h:31-52 (NGPU=1, gpuid=0).  Eigen's fixed-size-3 reductions associate as
// a0 + (a1 + a2) (redux_novec_unroller splits [0,1) | [1,3)); pinned against the host-compiled reference.
inline void ray_dir(const Frame& f, const float intr[4], int W, int p, float d[3]) {
  float i = (p % W) + 0.5;
  float j = (p / W) + 0.5;
  float zs = 1;
  float xs = (i - intr[2]) / intr[0] * zs;
  float ys = (j - intr[3]) / intr[1] * zs;
  // cuda: FFMA(ys,ys,1) then FFMA(xs,xs,.); row r: FMUL(vy,R[r][1]), FFMA(vz,R[r][2],.), FFMA(vx,R[r][0],.)
  float nrm = sqrtf(mad(xs, xs, mad(ys, ys, zs * zs)));
  float v[3] = {xs / nrm, ys / nrm, zs / nrm};
  for (int r = 0; r < 3; ++r) d[r] = mad(f.R[r][0], v[0], mad(f.R[r][2], v[2], f.R[r][1] * v[1]));
}

// kernel_near_far_from_aabb, render_utils.h:338-392
inline void near_far(const float o[3], const float d[3], const float* aabb, float min_near, float& near_out,
                     float& far_out) {
  const float rdx = 1 / d[0], rdy = 1 / d[1], rdz = 1 / d[2];
  float near = (aabb[0] - o[0]) * rdx, far = (aabb[3] - o[0]) * rdx;
  if (near > far) std::swap(near, far);
  float near_y = (aabb[1] - o[1]) * rdy, far_y = (aabb[4] - o[1]) * rdy;
  if (near_y > far_y) std::swap(near_y, far_y);
  if (near > far_y || near_y > far) { near_out = far_out = FLT_MAX; return; }
  if (near_y > near) near = near_y;
  if (far_y < far) far = far_y;
  float near_z = (aabb[2] - o[2]) * rdz, far_z = (aabb[5] - o[2]) * rdz;
  if (near_z > far_z) std::swap(near_z, far_z);
  if (near > far_z || near_z > far) { near_out = far_out = FLT_MAX; return; }
  if (near_z > near) near = near_z;
  if (far_z < far) far = far_z;
  if (near < min_near) near = min_near;
  near_out = near;
  far_out = far;
}

// ---- march: kernel_march_rays, render_utils.h:524-655 ----------------------------------------------------
struct MarchConst {
  float bound, dt_gamma, dt_min, dt_max, thresh;
  uint32_t C, H;
};

MarchConst march_const(const nrf_model_desc& d) {
  MarchConst m;
  m.bound = d.bound;
  m.dt_gamma = 1.0 / 128;                          // nerf_render.h:66 (float m_dt_gamma = 1.0/128)
  m.dt_min = 2 * 1.7320508075688772f / 1024;       // MIN_STEPSIZE(), render_utils.h:179-183
  m.dt_max = 2 * d.bound / d.grid_size;            // render_utils.h:581
  m.thresh = fminf(0.01f, d.mean_density);         // render_utils.h:560
  m.C = d.cascades;
  m.H = d.grid_size;
  return m;
}

struct Sample {
  float x, y, z, dt, t_after;
  int level;
  uint32_t cell;
};

// One loop iteration of render_utils.h:596-651.  Returns true when a sample was emitted.
inline bool march_step(const MarchConst& m, const float* grid, const float o[3], const float d[3],
                       const float rd[3], float& t, Sample& s) {
  const float x = clampf(mad(t, d[0], o[0]), -m.bound, m.bound);   // cuda: FFMA(d, t, o)
  const float y = clampf(mad(t, d[1], o[1]), -m.bound, m.bound);
  const float z = clampf(mad(t, d[2], o[2]), -m.bound, m.bound);
  // mip_from_pos, render_utils.h:148-155
  const float mx = fmaxf(fabsf(x), fmaxf(fabsf(y), fabsf(z)));
  int exponent;
  frexpf(mx, &exponent);
  const int level = fminf((float)m.C - 1, fmaxf(0, exponent));
  const float mip_bound = fminf(exp2f((float)level), m.bound);
  const float mip_rbound = 1 / mip_bound;
  const uint32_t H = m.H;
  // the literal 0.5 is a double: (x*rb+1) float, then double multiply by 0.5 and H, rounded to float by clamp()
  const int nx = clampf(0.5 * mad(x, mip_rbound, 1) * H, 0.0f, (float)(H - 1));   // cuda: FFMA(x, rb, 1), then DMUL x2
  const int ny = clampf(0.5 * mad(y, mip_rbound, 1) * H, 0.0f, (float)(H - 1));
  const int nz = clampf(0.5 * mad(z, mip_rbound, 1) * H, 0.0f, (float)(H - 1));
  const uint32_t cell = level * H * H * H + nx * H * H + ny * H + nz;
  if (grid[cell] > m.thresh) {
    s.x = x; s.y = y; s.z = z;
    const float dt = clampf(t * m.dt_gamma, m.dt_min, m.dt_max);
    t += dt;
    s.dt = dt;
    s.t_after = t;
    s.level = level;
    s.cell = cell;
    return true;
  }
  // cuda: FADD(n,.5), FFMA(sign,.5,.) [exact], IEEE div, FFMA(q,2,-1) [exact], FFMA(.,mip_bound,-x), FMUL(.,rd)
  const float tx = mad(mad((nx + 0.5f + 0.5f * copysignf(1.0, d[0])) / (H - 1), 2, -1), mip_bound, -x) * rd[0];
  const float ty = mad(mad((ny + 0.5f + 0.5f * copysignf(1.0, d[1])) / (H - 1), 2, -1), mip_bound, -y) * rd[1];
  const float tz = mad(mad((nz + 0.5f + 0.5f * copysignf(1.0, d[2])) / (H - 1), 2, -1), mip_bound, -z) * rd[2];
  const float tt = t + fmaxf(0.0f, fminf(tx, fminf(ty, tz)));
  do {
    const float dt = clampf(t * m.dt_gamma, m.dt_min, m.dt_max);
    t += dt;
  } while (t < tt);
  return false;
}

// ---- hash grid: tcnn grid.h:81-117 (fast_hash, grid_index), 139-268 (kernel_grid), common_device.h:424-431 --
inline uint32_t grid_entry(uint32_t T_l, uint32_t res, const uint32_t c[3]) {
  uint32_t stride = 1, index = 0;
  for (uint32_t dim = 0; dim < 3 && stride <= T_l; ++dim) {
    index += c[dim] * stride;
    stride *= res;
  }
  if (T_l < stride) index = (c[0] * 1u) ^ (c[1] * 2654435761u) ^ (c[2] * 805459861u);
  return index % T_l;
}

// One sample: u[3] in [0,1] -> 32 halves (row l*2+f).  idx_out (optional) L*8 entry indices (level-local).
void encode_sample(const Levels& lv, uint32_t L, const uint16_t* table, const float u[3], uint16_t* out,
                   uint32_t* idx_out) {
  for (uint32_t l = 0; l < L; ++l) {
    const uint32_t T_l = lv.offsets[l + 1] - lv.offsets[l];
    const uint16_t* tab = table + (size_t)lv.offsets[l] * 2;
    const float scale = lv.scales[l];
    const uint32_t res = lv.res[l];
    float w[3];
    uint32_t g[3];
    for (int dim = 0; dim < 3; ++dim) {  // pos_fract, identity interpolation
      float p = mad(u[dim], scale, 0.5f);   // cuda: FFMA(input, scale, .5)
      int tmp = floorf(p);
      g[dim] = (uint32_t)tmp;
      w[dim] = p - (float)tmp;
    }
    half_t acc0 = f2h(0.0f), acc1 = f2h(0.0f);
    for (uint32_t k = 0; k < 8; ++k) {
      float weight = 1;
      uint32_t c[3];
      for (int dim = 0; dim < 3; ++dim) {
        if ((k & (1u << dim)) == 0) { weight *= 1 - w[dim]; c[dim] = g[dim]; }
        else { weight *= w[dim]; c[dim] = g[dim] + 1; }
      }
      const uint32_t e = grid_entry(T_l, res, c);
      if (idx_out) idx_out[l * 8 + k] = e;
      acc0 = hadd(acc0, f2h(weight * h2f(bits2h(tab[e * 2 + 0]))));
      acc1 = hadd(acc1, f2h(weight * h2f(bits2h(tab[e * 2 + 1]))));
    }
    out[l * 2 + 0] = h2bits(acc0);
    out[l * 2 + 1] = h2bits(acc1);
  }
}

// ---- SH degree 4: tcnn spherical_harmonics.h:46-96.  in[3] is the normalised direction input in [0,1]. -----
void sh4(const float in[3], uint16_t out[16]) {
  float x = in[0] * 2.f - 1.f, y = in[1] * 2.f - 1.f, z = in[2] * 2.f - 1.f;
  float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
  float v[16];
  if (g_fma) {   // the contractions of kernel_sh<__half>'s SASS: entries 6, 8, 9, 11, 12, 13, 15
    v[0] = 0.28209479177387814f;
    v[1] = -0.48860251190291987f * y;
    v[2] = 0.48860251190291987f * z;
    v[3] = -0.48860251190291987f * x;
    v[4] = 1.0925484305920792f * xy;
    v[5] = -1.0925484305920792f * yz;
    v[6] = fmaf(0.94617469575755997f, z2, -0.31539156525251999f);
    v[7] = -1.0925484305920792f * xz;
    v[8] = fmaf(0.54627421529603959f, x2, -(0.54627421529603959f * y2));
    v[9] = 0.59004358992664352f * y * fmaf(-3.0f, x2, y2);
    v[10] = 2.8906114426405538f * xy * z;
    v[11] = 0.45704579946446572f * y * fmaf(-5.0f, z2, 1.0f);
    v[12] = 0.3731763325901154f * z * fmaf(5.0f, z2, -3.0f);
    v[13] = 0.45704579946446572f * x * fmaf(-5.0f, z2, 1.0f);
    v[14] = 1.4453057213202769f * z * (x2 - y2);
    v[15] = 0.59004358992664352f * x * fmaf(3.0f, y2, -x2);
    for (int k = 0; k < 16; ++k) out[k] = h2bits(f2h(v[k]));
    return;
  }
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
  for (int k = 0; k < 16; ++k) out[k] = h2bits(f2h(v[k]));
}

// ---- MLP: tcnn/src/fully_fused_mlp.cu:316-558 (inference), weights row-major [out][in] (:662-686,874-883),
//      wiring nerf_network.h:148-196, activation common_device.h:68-114 ------------------------------------
struct Mlp {
  // weights converted to float once and transposed to [in][out] so the per-output fp32 accumulation (k ascending,
  // identical order to a row-major dot product) vectorises across outputs
  std::vector<float> Wt[5];
};
inline Mlp mlp_view(const uint16_t* w) {
  static const int offs[5] = {0, 2048, 3072, 5120, 9216}, n_out[5] = {64, 16, 64, 64, 16}, n_in[5] = {32, 64, 32, 64, 64};
  Mlp m;
  for (int l = 0; l < 5; ++l) {
    m.Wt[l].resize((size_t)n_out[l] * n_in[l]);
    for (int o = 0; o < n_out[l]; ++o)
      for (int k = 0; k < n_in[l]; ++k) m.Wt[l][(size_t)k * n_out[l] + o] = h2f(bits2h(w[offs[l] + o * n_in[l] + k]));
  }
  return m;
}

inline void dense(const std::vector<float>& Wt, int n_out, int n_in, const half_t* x, half_t* y, bool relu) {
  float acc[64];
  for (int o = 0; o < n_out; ++o) acc[o] = 0.f;
  for (int k = 0; k < n_in; ++k) {
    const float xk = h2f(x[k]);
    const float* row = Wt.data() + (size_t)k * n_out;
    for (int o = 0; o < n_out; ++o) acc[o] += row[o] * xk;
  }
  for (int o = 0; o < n_out; ++o) {
    half_t h = f2h(acc[o]);
    if (relu) h = (h2f(h) > 0.0f) ? h : f2h(0.0f) * h;  // x * (x > 0): keeps -0 for negatives like the reference
    y[o] = h;
  }
}

inline float logistic(float x) { return 1.0f / (1.0f + expf(-x)); }  // tcnn common_device.h logistic

inline half_t act_out(uint32_t act, half_t v) {  // warp_activation / wrap_a_activation
  switch (act) {
    case NRF_ACT_RELU: return (h2f(v) > 0.0f) ? v : f2h(0.0f);
    case NRF_ACT_EXPONENTIAL: return f2h(expf(h2f(v)));
    case NRF_ACT_SIGMOID: return f2h(logistic(h2f(v)));
    default: return v;
  }
}

// enc: 32 halves, sh: 16 halves -> out {r,g,b,sigma} as float(half)
void network_sample(const nrf_model_desc& d, const Mlp& m, const uint16_t* enc, const uint16_t* sh, float out[4]) {
  half_t x[32], h[64], den[16], rin[32], h1[64], h2[64], rgb[16];
  for (int k = 0; k < 32; ++k) x[k] = bits2h(enc[k]);
  dense(m.Wt[0], 64, 32, x, h, true);
  dense(m.Wt[1], 16, 64, h, den, false);
  for (int k = 0; k < 16; ++k) { rin[k] = den[k]; rin[16 + k] = bits2h(sh[k]); }
  dense(m.Wt[2], 64, 32, rin, h1, true);
  dense(m.Wt[3], 64, 64, h1, h2, true);
  dense(m.Wt[4], 16, 64, h2, rgb, false);
  for (int k = 0; k < 3; ++k) out[k] = h2f(act_out(d.rgb_activation, rgb[k]));
  out[3] = h2f(act_out(d.sigma_activation, den[0]));  // extract_density, nerf_network.h:49-61,187-195
}

struct Model {
  nrf_model_desc d;
  Levels lv;
  std::vector<uint16_t> mlp;    // 10240 halves
  std::vector<uint16_t> table;  // 2*sum T_l halves
};

}  // namespace

extern "C" {

int orc_level_table(const nrf_model_desc* desc, uint32_t* offsets, float* scales, uint32_t* resolutions,
                    float* per_level_scale) {
  Levels lv;
  int rc = build_levels(*desc, lv, per_level_scale);
  if (rc) return rc;
  for (uint32_t i = 0; i <= desc->n_levels; ++i) offsets[i] = lv.offsets[i];
  for (uint32_t i = 0; i < desc->n_levels; ++i) { scales[i] = lv.scales[i]; resolutions[i] = lv.res[i]; }
  return 0;
}

// float32 -> fp16 bit patterns (the load-time cast, nerf_network.h:434-436)
void orc_cast_f16(const float* in, uint16_t* out, size_t n) {
  for (size_t i = 0; i < n; ++i) out[i] = h2bits(f2h(in[i]));
}
void orc_f16_to_f32(const uint16_t* in, float* out, size_t n) {
  for (size_t i = 0; i < n; ++i) out[i] = h2f(bits2h(in[i]));
}

// rays for pixels [pix0, pix0+n): rays_d n*3, nears n, fars n
int orc_rays(const nrf_model_desc* desc, const float* aabb, const float intr[4], const float* pose, int W, int H,
             int pix0, int n, float rays_o[3], float* rays_d, float* nears, float* fars) {
  (void)H;
  Frame f;
  pose_to_ngp(pose, desc->scale, f);
  for (int k = 0; k < 3; ++k) rays_o[k] = f.o[k];
#pragma omp parallel for schedule(static)
  for (int q = 0; q < n; ++q) {
    float d[3];
    ray_dir(f, intr, W, pix0 + q, d);
    rays_d[q * 3 + 0] = d[0]; rays_d[q * 3 + 1] = d[1]; rays_d[q * 3 + 2] = d[2];
    near_far(f.o, d, aabb, 0.2f /* m_min_near, nerf_render.h:76 */, nears[q], fars[q]);
  }
  return 0;
}

// whole-ray march.  counts[n]; samples (optional) 8 floats per sample, ray-major:
// {x,y,z,dt,t_after,level(int bits),cell(uint bits),ray(int bits)}.  Returns total samples, or -1 if max_samples
// is too small.
long long orc_march(const nrf_model_desc* desc, const float* grid, int n, const float rays_o[3],
                    const float* rays_d, const float* nears, const float* fars, int32_t* counts, float* samples,
                    size_t max_samples) {
  const MarchConst m = march_const(*desc);
  std::vector<long long> offs((size_t)n + 1, 0);
#pragma omp parallel for schedule(dynamic, 64)
  for (int r = 0; r < n; ++r) {
    const float* d = rays_d + (size_t)r * 3;
    const float rd[3] = {1 / d[0], 1 / d[1], 1 / d[2]};
    float t = nears[r];
    int c = 0;
    Sample s;
    while (t < fars[r] && c < NRF_MAX_STEPS)
      if (march_step(m, grid, rays_o, d, rd, t, s)) ++c;
    counts[r] = c;
  }
  for (int r = 0; r < n; ++r) offs[r + 1] = offs[r] + counts[r];
  if (!samples) return offs[n];
  if ((size_t)offs[n] > max_samples) return -1;
#pragma omp parallel for schedule(dynamic, 64)
  for (int r = 0; r < n; ++r) {
    const float* d = rays_d + (size_t)r * 3;
    const float rd[3] = {1 / d[0], 1 / d[1], 1 / d[2]};
    float t = nears[r];
    int c = 0;
    Sample s;
    float* out = samples + (size_t)offs[r] * 8;
    while (t < fars[r] && c < NRF_MAX_STEPS) {
      if (march_step(m, grid, rays_o, d, rd, t, s)) {
        float* q = out + (size_t)c * 8;
        q[0] = s.x; q[1] = s.y; q[2] = s.z; q[3] = s.dt; q[4] = s.t_after;
        std::memcpy(q + 5, &s.level, 4);
        std::memcpy(q + 6, &s.cell, 4);
        std::memcpy(q + 7, &r, 4);
        ++c;
      }
    }
  }
  return offs[n];
}

// positions n*3 in [0,1] -> enc n*32 halves (sample-major), indices n*L*8 (optional)
int orc_encode(const nrf_model_desc* desc, const uint16_t* table, int n, const float* positions, uint16_t* enc,
               uint32_t* indices) {
  Levels lv;
  int rc = build_levels(*desc, lv, nullptr);
  if (rc) return rc;
  const uint32_t L = desc->n_levels;
#pragma omp parallel for schedule(static)
  for (int i = 0; i < n; ++i)
    encode_sample(lv, L, table, positions + (size_t)i * 3, enc + (size_t)i * 2 * L,
                  indices ? indices + (size_t)i * L * 8 : nullptr);
  return 0;
}

// dirs n*3 normalised direction input in [0,1] -> sh n*16 halves
void orc_sh(int n, const float* dirs, uint16_t* out) {
  for (int i = 0; i < n; ++i) sh4(dirs + (size_t)i * 3, out + (size_t)i * 16);
}

// enc n*32 halves, dirs n*3 in [0,1] -> out n*4 {r,g,b,sigma}
int orc_network(const nrf_model_desc* desc, const uint16_t* mlp_weights, int n, const uint16_t* enc,
                const float* dirs, float* out) {
  if (check_arch(*desc)) return NRF_EINVAL;
  const Mlp m = mlp_view(mlp_weights);
#pragma omp parallel for schedule(static)
  for (int i = 0; i < n; ++i) {
    uint16_t sh[16];
    sh4(dirs + (size_t)i * 3, sh);
    network_sample(*desc, m, enc + (size_t)i * 32, sh, out + (size_t)i * 4);
  }
  return 0;
}

// positions, dirs (both normalised to [0,1]) -> {r,g,b,sigma}
int orc_samples(const nrf_model_desc* desc, const uint16_t* mlp_weights, const uint16_t* table, int n,
                const float* positions, const float* dirs, float* out) {
  if (check_arch(*desc)) return NRF_EINVAL;
  Levels lv;
  build_levels(*desc, lv, nullptr);
  const Mlp m = mlp_view(mlp_weights);
#pragma omp parallel for schedule(static)
  for (int i = 0; i < n; ++i) {
    uint16_t enc[32], sh[16];
    encode_sample(lv, 16, table, positions + (size_t)i * 3, enc, nullptr);
    sh4(dirs + (size_t)i * 3, sh);
    network_sample(*desc, m, enc, sh, out + (size_t)i * 4);
  }
  return 0;
}

// composite one ray's samples: kernel_composite_rays render_utils.h:691-749 + get_image_and_depth :241-265.
// sig_rgb n*4 {r,g,b,sigma}, dts n, t_afters n (march t after each sample).  Returns composited sample count.
int orc_composite_ray(int n, const float* sig_rgb, const float* dts, const float* t_afters, float near,
                      float far, float out_rgb[3], float* out_depth, float* out_wsum) {
  float wsum = 0, dep = 0, r = 0, g = 0, b = 0;
  float t = near, last_t = near;
  int used = 0;
  for (int s = 0; s < n; ++s) {
    const float alpha = 1.0f - expf(-sig_rgb[s * 4 + 3] * dts[s]);
    const float T = 1 - wsum;
    const float w = alpha * T;
    wsum += w;
    t += t_afters[s] - last_t;  // deltas(1) = t - last_t  (render_utils.h:629-631)
    last_t = t_afters[s];
    dep += w * t;
    r += w * sig_rgb[s * 4 + 0];
    g += w * sig_rgb[s * 4 + 1];
    b += w * sig_rgb[s * 4 + 2];
    ++used;
    if (T < 1e-4) break;
  }
  const float bg = 1;  // m_bg_color, nerf_render.h:74
  out_rgb[0] = r + (1 - wsum) * bg;
  out_rgb[1] = g + (1 - wsum) * bg;
  out_rgb[2] = b + (1 - wsum) * bg;
  // defect D3: missed rays have near=far=FLT_MAX -> 0/0; defined as 0 here and in the product.
  *out_depth = (far == near) ? 0.0f : fmaxf(dep - near, 0.0f) / (far - near);
  if (out_wsum) *out_wsum = wsum;
  return used;
}

// Whole frame, per-ray sequential restatement of nerf_render.cu:238-367 under D1-clean / D8 (cap = 1024 samples).
// rgb W*H*3, depth W*H, counts W*H (composited samples), all optional except rgb.  Returns total composited samples.
long long orc_render_frame(const nrf_model_desc* desc, const uint16_t* mlp_weights, const uint16_t* table,
                           const float* grid, const float* aabb, const float intr[4], const float* pose, int W,
                           int H, int pix0, int npix, float* rgb, float* depth, int32_t* counts) {
  if (check_arch(*desc)) return NRF_EINVAL;
  Levels lv;
  build_levels(*desc, lv, nullptr);
  const Mlp mlp = mlp_view(mlp_weights);
  const MarchConst m = march_const(*desc);
  Frame f;
  pose_to_ngp(pose, desc->scale, f);
  (void)H;
  const float pos_w = 1.0 / (2 * desc->bound);  // nerf_render.cu:311-312: 1.0/(2*m_bound) -> float weight
  long long total = 0;
#pragma omp parallel for schedule(dynamic, 32) reduction(+ : total)
  for (int q = 0; q < npix; ++q) {
    float d[3], near, far;
    ray_dir(f, intr, W, pix0 + q, d);
    near_far(f.o, d, aabb, 0.2f, near, far);
    const float rd[3] = {1 / d[0], 1 / d[1], 1 / d[2]};
    const float dn[3] = {0.5f * d[0] + 0.5f, 0.5f * d[1] + 0.5f, 0.5f * d[2] + 0.5f};  // nerf_render.cu:313-314
    uint16_t sh[16];
    sh4(dn, sh);
    float wsum = 0, dep = 0, cr = 0, cg = 0, cb = 0;
    float t = near, tc = near, last_t = near;
    int c = 0;
    Sample s;
    while (t < far && c < NRF_MAX_STEPS) {
      if (!march_step(m, grid, f.o, d, rd, t, s)) continue;
      ++c;
      // linear_transformer (common_device.cuh:31-36): weight*input + bias — cuda: one FFMA
      const float u[3] = {mad(pos_w, s.x, 0.5f), mad(pos_w, s.y, 0.5f), mad(pos_w, s.z, 0.5f)};
      uint16_t enc[32];
      float o4[4];
      encode_sample(lv, 16, table, u, enc, nullptr);
      network_sample(*desc, mlp, enc, sh, o4);
      const float alpha = 1.0f - expf(-o4[3] * s.dt);
      const float T = 1 - wsum;
      const float w = alpha * T;
      wsum += w;
      tc += s.t_after - last_t;
      last_t = s.t_after;
      dep += w * tc;
      cr += w * o4[0]; cg += w * o4[1]; cb += w * o4[2];
      if (T < 1e-4) break;
    }
    total += c;
    rgb[q * 3 + 0] = cr + (1 - wsum) * 1;
    rgb[q * 3 + 1] = cg + (1 - wsum) * 1;
    rgb[q * 3 + 2] = cb + (1 - wsum) * 1;
    if (depth) depth[q] = (far == near) ? 0.0f : fmaxf(dep - near, 0.0f) / (far - near);
    if (counts) counts[q] = c;
  }
  return total;
}

// u8 conversion, nerf_render.cu:355-358: (unsigned char)(255.0*v); clamped where the reference is UB (defect D4).
void orc_to_u8(const float* in, uint8_t* out, size_t n) {
  for (size_t i = 0; i < n; ++i) {
    double v = 255.0 * in[i];
    out[i] = (v != v || v < 0.0) ? 0 : (v >= 255.0 ? 255 : (uint8_t)v);
  }
}

// ---- density-grid regeneration: NerfRender::generate_density_grid, nerf_render.cu:388-429, with the kernels it
// launches: init_xyzs render_utils.h:88-105, dd_scale :79-86, add_random :107-117, dg_update :119-128.  The reference
// routine is dead code (no caller; the density evaluation is commented out at nerf_render.cu:415), so this restates
// what the code intends, with the decisions listed in DESIGN.md: positions normalised to [0,1] (nerf_render.cu:311-312)
// before the network, density = sigma after the configured activation, scale 0.001691 (:417), pcg32 jitter advanced to
// the element index 3*(cascade*H^3 + cell) + axis (tcnn random.h:36-65 generator; pcg32 skip-ahead), reset = the 1/64
// fill of :395.  grid [cascades*H^3] is updated in place; returns the new mean density.
struct Pcg32 {
  uint64_t state, inc;
  explicit Pcg32(uint64_t initstate, uint64_t initseq = 1) {
    state = 0; inc = (initseq << 1u) | 1u;
    next_uint(); state += initstate; next_uint();
  }
  uint32_t next_uint() {
    const uint64_t old = state;
    state = old * 0x5851f42d4c957f2dULL + inc;
    const uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u), rot = (uint32_t)(old >> 59u);
    return (xs >> rot) | (xs << ((~rot + 1u) & 31));
  }
  float next_float() {
    const uint32_t b = (next_uint() >> 9) | 0x3f800000u;
    float f; std::memcpy(&f, &b, 4);
    return f - 1.0f;
  }
  void advance(uint64_t delta) {
    uint64_t cur_mult = 0x5851f42d4c957f2dULL, cur_plus = inc, acc_mult = 1, acc_plus = 0;
    while (delta > 0) {
      if (delta & 1) { acc_mult *= cur_mult; acc_plus = acc_plus * cur_mult + cur_plus; }
      cur_plus = (cur_mult + 1) * cur_plus;
      cur_mult *= cur_mult;
      delta >>= 1;
    }
    state = acc_mult * state + acc_plus;
  }
};

float orc_density_grid(const nrf_model_desc* desc, const uint16_t* mlp_weights, const uint16_t* table, float decay,
                       int reset, uint64_t seed, int jitter, float* grid) {
  if (check_arch(*desc)) return -1.f;
  Levels lv;
  build_levels(*desc, lv, nullptr);
  const Mlp m = mlp_view(mlp_weights);
  const uint32_t H = desc->grid_size, H3 = H * H * H;
  const float pos_w = (float)(1.0 / (2 * desc->bound));
  double sum = 0;
  for (uint32_t cas = 0; cas < desc->cascades; ++cas) {
    const float bound_c = (float)(1u << cas) < desc->bound ? (float)(1u << cas) : desc->bound;
    const float half = bound_c / H;
#pragma omp parallel for schedule(static) reduction(+ : sum)
    for (long long idx = 0; idx < (long long)H3; ++idx) {
      const uint32_t id[3] = {(uint32_t)idx / (H * H), (uint32_t)idx % (H * H) / H, (uint32_t)idx % H};
      float u[3];
      for (int a = 0; a < 3; ++a) {
        float p = -1.f + 2.f / (H - 1) * id[a];
        p = (bound_c - half) * p + 0.f;
        if (jitter) {
          Pcg32 rng(seed);
          rng.advance(3ull * cas * H3 + 3ull * (uint64_t)idx + a);
          p += half * (2.f * rng.next_float() - 1.f) + 0.f;
        }
        u[a] = pos_w * p + 0.5f;
      }
      uint16_t enc[32], sh[16] = {0};
      float out[4];
      encode_sample(lv, 16, table, u, enc, nullptr);
      network_sample(*desc, m, enc, sh, out);
      const float tmp = 0.001691f * out[3] + 0.f;
      float* g = grid + (size_t)cas * H3 + idx;
      const float dg = reset ? 1.0f / 64 : *g;
      float v = dg;
      if (dg >= 0) v = (dg * decay > tmp) ? dg * decay : tmp;
      *g = v;
      sum += (double)v;
    }
  }
  return (float)(sum / ((double)H3 * desc->cascades));
}

// ---- post-process: src/render_buffer.cu accumulate_kernel :224-258, tonemap :260-338, tonemap_kernel :529-555;
// linear_to_srgb / srgb_to_linear include/nerf-cuda/common_device.cuh:38-60.  RGBA float buffers, n pixels.
static inline float srgb_to_linear(float s) { return s <= 0.04045f ? s / 12.92f : std::pow((s + 0.055f) / 1.055f, 2.4f); }
static inline float linear_to_srgb(float l) { return l < 0.0031308f ? 12.92f * l : 1.055f * std::pow(l, 0.41666f) - 0.055f; }

void orc_accumulate(int n, const float* frame, float* accum, float sample_count, int color_space) {
  for (int i = 0; i < n; ++i) {
    float c[4], *t = accum + (size_t)i * 4;
    std::memcpy(c, frame + (size_t)i * 4, 16);
    if (color_space == 2) {  // VisPosNeg
      const float val = c[0] - c[1];
      float tv = t[0] - t[1];
      tv = (tv * sample_count + val) / (sample_count + 1);
      t[0] = fmaxf(tv, 0.0f);
      t[1] = fmaxf(-tv, 0.0f);
    } else {
      if (color_space == 1) for (int k = 0; k < 3; ++k) c[k] = linear_to_srgb(c[k]);
      for (int k = 0; k < 3; ++k) t[k] = (t[k] * sample_count + c[k]) / (sample_count + 1);
    }
    t[3] = (t[3] * sample_count + c[3]) / (sample_count + 1);
  }
}

static void tonemap_curve(float x[3], int curve) {
  if (curve == 0) return;
  for (int k = 0; k < 3; ++k) x[k] = fmaxf(x[k], 0.f);
  float k0, k1, k2, k3, k4, k5;
  if (curve == 1) {
    k0 = 0.6f * 0.6f * 2.51f; k1 = 0.6f * 0.03f; k2 = 0.0f; k3 = 0.6f * 0.6f * 2.43f; k4 = 0.6f * 0.59f; k5 = 0.14f;
  } else if (curve == 2) {
    const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
    k0 = A * F - A * E; k1 = C * B * F - B * E; k2 = 0.0f; k3 = A * F; k4 = B * F; k5 = D * F * F;
    const float W = 11.2f;
    const float nom = k0 * (W * W) + k1 * W + k2, denom = k3 * (W * W) + k4 * W + k5;
    const float white_scale = denom / nom;
    k0 = 4.0f * k0 * white_scale; k1 = 2.0f * k1 * white_scale; k2 = k2 * white_scale; k3 = 4.0f * k3; k4 = 2.0f * k4;
  } else {
    const float Y = 0.2126f * x[0] + (0.7152f * x[1] + 0.0722f * x[2]);
    const float s = 1.f / (Y + 1.0f);
    for (int k = 0; k < 3; ++k) x[k] *= s;
    return;
  }
  for (int k = 0; k < 3; ++k) {
    const float v = x[k], sq = v * v;
    const float nom = sq * k0 + k1 * v + k2, denom = k3 * sq + k4 * v + k5;
    x[k] = nom / denom;
  }
}

void orc_tonemap(int n, const float* accum, float exposure, const float background_in[4], int color_space,
                 int output_color_space, int curve, int clamp_output, float* out) {
  float bg[4];
  std::memcpy(bg, background_in, 16);
  if (color_space != 1) for (int k = 0; k < 3; ++k) bg[k] = srgb_to_linear(bg[k]);
  for (int i = 0; i < n; ++i) {
    float c[4];
    std::memcpy(c, accum + (size_t)i * 4, 16);
    const float weight = (1 - c[3]) * bg[3];
    for (int k = 0; k < 3; ++k) c[k] += bg[k] * weight;
    c[3] += weight;
    if (color_space == 1) for (int k = 0; k < 3; ++k) c[k] = srgb_to_linear(c[k]);
    const float e = std::pow(2.0f, exposure);
    for (int k = 0; k < 3; ++k) c[k] *= e;
    tonemap_curve(c, curve);
    if (output_color_space == 1) for (int k = 0; k < 3; ++k) c[k] = linear_to_srgb(c[k]);
    if (clamp_output) for (int k = 0; k < 4; ++k) c[k] = fminf(fmaxf(c[k], 0.0f), 1.0f);
    std::memcpy(out + (size_t)i * 4, c, 16);
  }
}

// Numerics mode of every later call (see header): fma != 0 -> "cuda", else "host"; scales (may be NULL) = the n per-level
// grid scales as the DEVICE's exp2f gives them (from the product, or from tests/golden/ref_cuda_*.npz).
int orc_set_numerics(int fma, const float* scales, int n) {
  g_fma = fma != 0;
  g_have_scales = scales != nullptr && n > 0 && n <= NRF_MAX_LEVELS;
  if (g_have_scales) for (int i = 0; i < n; ++i) g_scales[i] = scales[i];
  return 0;
}

int orc_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

// Explicit thread count for the timed CPU legs: torch.distributed.run exports OMP_NUM_THREADS=1 to every rank, which
// would silently time the oracle on one core (bench.py sets this to the cores the process may run on).
int orc_set_num_threads(int n) {
#ifdef _OPENMP
  if (n >= 1) omp_set_num_threads(n);
  return omp_get_max_threads();
#else
  (void)n;
  return 1;
#endif
}

}  // extern "C"

// pcg32 stream for known-answer tests (tests/test_oracle_golden.py): n outputs after `skip` advance steps.
extern "C" void orc_pcg32(uint64_t initstate, uint64_t initseq, uint64_t skip, int n, uint32_t* out) {
  Pcg32 r(initstate, initseq);
  if (skip) r.advance(skip);
  for (int i = 0; i < n; ++i) out[i] = r.next_uint();
}
