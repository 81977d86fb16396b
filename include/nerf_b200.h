/*
 * nerf_b200.h — C ABI of the B200-native NeRF inference hot path.
 *
 * This is the drop-in boundary.  The reference has no FFI layer: its boundary is
 * the C++ class ngp::NerfRender (/root/reference/include/nerf-cuda/nerf_render.h:29-50) plus the
 * render_server TCP protocol (/root/reference/src/render_server.cu:93-101).  The C++ facade in
 * include/nerf-cuda/nerf_render.h of this repo keeps the reference's method names and forwards to the
 * entry points below; every entry point cites the reference code it replaces.
 *
 * Rules of the ABI: plain C types only, no Eigen / json / STL / torch types.  Every function returns
 * 0 on success or a negative NRF_E* code; the message is available from nrf_last_error() (thread local).
 * A context is not thread-safe (the reference's NerfRender is not either, nerf_render.h:80).
 * There is NO CPU fallback: every compute entry point fails with NRF_ENODEV when no CUDA device is usable.
 */
#ifndef NERF_B200_H
#define NERF_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NRF_OK 0
#define NRF_EINVAL (-1)   /* bad argument / unsupported architecture                         */
#define NRF_ENODEV (-2)   /* no CUDA device (there is no CPU path)                            */
#define NRF_ECUDA (-3)    /* CUDA runtime error, see nrf_last_error()                         */
#define NRF_EIO (-4)      /* snapshot file missing / unreadable (reference: std::runtime_error) */
#define NRF_EFORMAT (-5)  /* snapshot content invalid (no "snapshot", size mismatches)        */
#define NRF_ESTATE (-6)   /* call order violated (e.g. render before load / set_resolution)   */

/* activation ids (tcnn::Activation subset used by base.json, tcnn common.h) */
#define NRF_ACT_NONE 0
#define NRF_ACT_RELU 1
#define NRF_ACT_EXPONENTIAL 2
#define NRF_ACT_SIGMOID 3

#define NRF_MAX_LEVELS 16
#define NRF_MAX_STEPS 1024 /* m_max_infer_steps, nerf_render.h:78; per-ray sample cap */

/*
 * Model + scene description: what reset_network() (nerf_render.cu:111-184) derives from the config json
 * and what load_snapshot() (nerf_render.cu:431-473) reads from "snapshot".
 * The architecture is the reference's base.json family: HashGrid(L levels, F=2) -> density MLP
 * 32->64->16, SH degree 4, rgb MLP 32->64->64->16(3 used).  Other shapes are rejected with NRF_EINVAL.
 */
typedef struct nrf_model_desc {
  uint32_t n_levels;              /* "encoding.n_levels" (16)                                    */
  uint32_t n_features_per_level;  /* "encoding.n_features_per_level" (2)                         */
  uint32_t log2_hashmap_size;     /* "encoding.log2_hashmap_size"                                */
  uint32_t base_resolution;       /* "encoding.base_resolution"; 0 => 1<<(log2T/3) (nerf_render.cu:147-152) */
  float per_level_scale;          /* <=0 => exp(log(2048*bound/base)/(L-1))   (nerf_render.cu:154-165)     */
  uint32_t n_neurons;             /* 64                                                          */
  uint32_t density_hidden_layers; /* "network.n_hidden_layers" (1)                               */
  uint32_t rgb_hidden_layers;     /* "rgb_network.n_hidden_layers" (2)                           */
  uint32_t sh_degree;             /* 4                                                           */
  uint32_t rgb_activation;        /* "rgb_network.output_activation": NRF_ACT_NONE | NRF_ACT_SIGMOID */
  uint32_t sigma_activation;      /* "network.sigma_activation", default Exponential (nerf_network.h:125)  */
  float bound;                    /* snapshot.bound            (m_bound)                         */
  float scale;                    /* snapshot.scale            (m_scale, 0.33)                   */
  uint32_t cascades;              /* snapshot.cascade          (m_dg_cascade)                    */
  uint32_t grid_size;             /* snapshot.density_grid_size (m_dg_h, 128)                    */
  float mean_density;             /* snapshot.mean_density                                       */
} nrf_model_desc;

/* Per-frame statistics (SURVEY.md §5 metrics row). Times are CUDA-event milliseconds on GPU 0's stream. */
typedef struct nrf_stats {
  uint64_t n_rays;        /* rays of the last frame (W*H)                                   */
  uint64_t n_samples;     /* samples evaluated by the network (all GPUs)                    */
  uint64_t n_hit_rays;    /* rays with at least one sample                                  */
  float ms_total;         /* whole frame, device time                                       */
  float ms_setup;         /* ray generation + aabb                                          */
  float ms_march;         /* occupancy march (count + emit)                                 */
  float ms_network;       /* hash-grid encode + fused MLP                                   */
  float ms_composite;     /* composite + finalise                                           */
  uint32_t n_launches;    /* kernels launched for the last frame                            */
  uint32_t n_gpus;
  uint64_t n_samples_marched; /* samples the occupancy march found on all rays (schedule-independent)          */
  uint64_t n_samples_used;    /* samples composited before the rays terminated (T < 1e-4, render_utils.h:727)  */
  uint32_t n_rounds;          /* evaluation rounds of the frame: 1 = whole rays, > 1 = termination-aware        */
  uint32_t reserved;
} nrf_stats;

typedef struct nrf_ctx nrf_ctx;

const char* nrf_last_error(void);
const char* nrf_version(void);

/* Level geometry of the hash grid, host-side (tcnn grid.h:892-940 ctor + grid.h:186-187 per-level scale).
 * offsets[L+1] in entries, scales[L], resolutions[L].  Pure host arithmetic, usable without a GPU. */
int nrf_level_table(const nrf_model_desc* desc, uint32_t* offsets, float* scales, uint32_t* resolutions);
/* Number of float params the flat "params" array must hold: 10240 MLP + 2*sum(T_l) (nerf_network.h:273-291). */
int nrf_param_count(const nrf_model_desc* desc, uint64_t* n_params);

/* NerfRender::NerfRender() (nerf_render.cu:46-57). n_gpus is runtime (reference: compile-time NGPU, common.h:91).
 * device_ids may be NULL (=> 0..n_gpus-1). */
int nrf_create(int n_gpus, const int* device_ids, nrf_ctx** out);
/* NerfRender::~NerfRender() */
int nrf_destroy(nrf_ctx* ctx);

/* NerfRender::reload_network_from_file (nerf_render.cu:93-109): .msgpack snapshot in the reference schema. */
int nrf_load_snapshot(nrf_ctx* ctx, const char* path);
/* Same, from memory: params in tcnn order (density MLP, rgb MLP, hash table), float32, cast to fp16 on load
 * (nerf_network.h:424-443); density_grid float [cascades*H^3] (nerf_render.cu:446-462); aabb[6]. */
int nrf_load_raw(nrf_ctx* ctx, const nrf_model_desc* desc, const float* params, size_t n_params,
                 const float* density_grid, size_t n_grid, const float aabb[6]);
int nrf_get_model_desc(nrf_ctx* ctx, nrf_model_desc* out);
/* Host-only: parse a snapshot file (load_snapshot, nerf_render.cu:431-473 + load_network_config :66-91) without
 * touching a device; reports the model description, array lengths, aabb and (for integrity checks) the double-precision
 * sums of the two arrays.  Any output pointer may be NULL.  Same error codes / messages as nrf_load_snapshot. */
int nrf_snapshot_inspect(const char* path, nrf_model_desc* desc, uint64_t* n_params, uint64_t* n_grid, float aabb[6],
                         double* params_sum, double* grid_sum);

/* NerfRender::set_resolution (nerf_render.cu:186-236); idempotent here (reference defect D7). */
int nrf_set_resolution(nrf_ctx* ctx, int width, int height);

/* NerfRender::render_frame (nerf_render.cu:238-367).  intr = {fl_x, fl_y, cx, cy} (struct Camera, common.h:68-74);
 * pose = row-major 4x4 camera-to-world in NeRF convention (render_server.cu:93-99).
 * *rgb (W*H*3 u8) and *depth (W*H u8) point into context-owned pinned host memory, valid until the next
 * render call / destroy (same ownership as struct Image, common.h:76-89). */
int nrf_render_frame(nrf_ctx* ctx, const float intr[4], const float pose_rowmajor[16], uint8_t** rgb,
                     uint8_t** depth);
/* Batch of poses (render_server batch extension).  rgb_out: caller-owned host buffer n*W*H*3 bytes. */
int nrf_render_batch(nrf_ctx* ctx, const float intr[4], const float* poses_rowmajor, int n_poses,
                     uint8_t* rgb_out);
/* Device-resident variant for measurement: renders without host copies; result stays in HBM. */
int nrf_render_frame_device(nrf_ctx* ctx, const float intr[4], const float pose_rowmajor[16]);
/* Float image of the last frame, before u8 quantisation (parity is judged on floats, reference defects D3/D4):
 * rgb_f32 W*H*3, depth_f32 W*H, samples_per_ray W*H (composited samples). Any pointer may be NULL. */
int nrf_get_float_frame(nrf_ctx* ctx, float* rgb_f32, float* depth_f32, int32_t* samples_per_ray);
int nrf_get_stats(nrf_ctx* ctx, nrf_stats* out);
int nrf_synchronize(nrf_ctx* ctx);

/* NerfRender::generate_density_grid (nerf_render.cu:388-429; render_utils.h:79-128 init_xyzs / dd_scale / add_random /
 * dg_update): re-evaluates the density network at the centre of every occupancy cell of every cascade (jittered by
 * +-half a cell with pcg32(jitter_seed) when jitter != 0), scales by 0.001691, EMA-max-updates the resident float grid
 * (grid = max(grid*decay, density) where grid >= 0; reset != 0 starts from the reference's 1/64 fill), then refreshes
 * mean_density and the occupancy bitfield the march reads.  The reference routine is dead code; see DESIGN.md for the
 * decisions taken where it is broken.  *mean_density_out (may be NULL) = new mean of the grid. */
int nrf_generate_density_grid(nrf_ctx* ctx, float decay, int reset, unsigned long long jitter_seed, int jitter,
                              float* mean_density_out);
/* Resident float density grid [cascades*H^3] and occupancy bitfield [cascades*H^3/32] (either pointer may be NULL). */
int nrf_get_density_grid(nrf_ctx* ctx, float* grid_out, uint32_t* bits_out);

/* Frame post-processing on linear RGBA float buffers (CudaRenderBuffer::accumulate / ::tonemap, src/render_buffer.cu:
 * accumulate_kernel :224-258, tonemap :260-338, tonemap_kernel :529-555; the GL-surface / DLSS output plumbing is not
 * built — no graphics stack).  All buffers are host pointers to n_pixels*4 floats.  If frame_rgba != NULL the frame is
 * folded into accum_rgba as the (sample_count+1)-th sample (running mean; colour spaces as EColorSpace, common.h:93-97:
 * 0 Linear, 1 SRGB, 2 VisPosNeg).  If out_rgba != NULL the accumulator is blended over background_rgba (given in sRGB),
 * exposed by 2^exposure, tonemapped (ETonemapCurve, common.h:100-105: 0 Identity, 1 ACES, 2 Hable, 3 Reinhard),
 * converted to output_color_space and optionally clamped to [0,1]. */
int nrf_postprocess(nrf_ctx* ctx, int n_pixels, const float* frame_rgba, float* accum_rgba, float sample_count,
                    int color_space, float exposure, const float background_rgba[4], int output_color_space,
                    int tonemap_curve, int clamp_output, float* out_rgba);

/* ---- stage probes (test / bench entry points; host pointers in, host pointers out) -------------------- */

/* K1-K5 (render_utils.h:31-66,338-392): rays for pixels [pix0, pix0+n). rays_d n*3 (AoS), near/far n. */
int nrf_probe_rays(nrf_ctx* ctx, const float intr[4], const float pose_rowmajor[16], int width, int height,
                   int pix0, int n, float rays_o[3], float* rays_d, float* nears, float* fars);
/* K7 (render_utils.h:524-655) over whole rays: counts[n] = samples per ray (<= NRF_MAX_STEPS);
 * if samples != NULL it receives, ray-major compacted, 8 floats per sample:
 * {x, y, z, dt, t_after, level, cell_index, ray} (x,y,z un-normalised, ints stored as float bit patterns via memcpy). */
int nrf_probe_march(nrf_ctx* ctx, int n, const float rays_o[3], const float* rays_d, const float* nears,
                    const float* fars, int32_t* counts, float* samples, size_t max_samples);
/* K10 (tcnn grid.h:139-268): positions n*3 in [0,1]; enc_out n*32 fp16 bit patterns (sample-major);
 * indices_out (may be NULL) n*L*8 uint32 entry indices (corner order of grid.h:239-251). */
int nrf_probe_encode(nrf_ctx* ctx, int n, const float* positions, uint16_t* enc_out, uint32_t* indices_out);
/* tcnn kernel_sh<__half>, degree 4 (spherical_harmonics.h:46-152): dirs n*3 in [0,1] (normalised direction input);
 * sh_out n*16 fp16 bit patterns.  The frame path evaluates this once per ray. */
int nrf_probe_sh(nrf_ctx* ctx, int n, const float* dirs, uint16_t* sh_out);
/* K11-K14 (fully_fused_mlp.cu, spherical_harmonics.h, nerf_network.h:148-196): enc n*32 fp16, dirs n*3 in [0,1]
 * (normalised direction input); out n*4 float {r,g,b,sigma}.  impl: 0 = tcgen05 kernel, 1 = SIMT check kernel. */
int nrf_probe_network(nrf_ctx* ctx, int n, const uint16_t* enc, const float* dirs, float* out, int impl);
/* Whole per-sample network: positions n*3 [0,1], dirs n*3 [0,1] -> n*4 float {r,g,b,sigma}. */
int nrf_probe_samples(nrf_ctx* ctx, int n, const float* positions, const float* dirs, float* out);

/* ---- device-resident benchmarking entry points (inputs already in HBM; no host copies) ---------------- */
/* Runs the encode(+MLP) stage on n pseudo-random resident samples `iters` times; returns mean device ms.
 * what: 0 = encode only, 1 = MLP only, 2 = fused encode+MLP. */
int nrf_bench_stage(nrf_ctx* ctx, int what, int n, int iters, float* ms_out);

#ifdef __cplusplus
}
#endif
#endif /* NERF_B200_H */
