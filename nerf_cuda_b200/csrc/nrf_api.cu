// nrf_api.cu — the C ABI (include/nerf_b200.h, include/nerf_b200_ext.h): context, model upload, per-frame
// orchestration and the stage probes.  Host-side counterpart of ngp::NerfRender
// (/root/reference/include/nerf-cuda/nerf_render.h:29-133, /root/reference/src/nerf_render.cu).
//
// Differences from the reference's host loop (src/nerf_render.cu:238-367), all deliberate:
//   * 5 kernel launches and ONE 8-byte D2H read per frame instead of ~13 launches + a blocking D2H per wavefront
//     iteration (up to 1024 iterations);
//   * the network runs on exactly the live samples, not on the padded N_pad buffer every iteration;
//   * n_gpus is a runtime value; each device renders interleaved pixel blocks and copies its u8 pixels straight
//     into the pinned host frame (strided 2-D copy) — no CPU de-interleave loop;
//   * there is no CPU path: without a CUDA device every compute entry point returns NRF_ENODEV.
#include <cuda_runtime.h>
#include <sys/stat.h>

#include <algorithm>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <thread>
#include <cmath>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <memory>
#include <string>
#include <vector>

#include "../../include/nerf_b200.h"
#include "../../include/nerf_b200_ext.h"
#include "nrf_device.cuh"
#include "nrf_snapshot.h"

// kernels (kernels_exact.cu, kernels_network.cu, kernels_composite.cu)
extern "C" {
__global__ void k_rays_count(const MarchParams P, ShardMap map, int n_local, float4* rays, float2* far_tf, int* counts,
                             int* block_sums, uint4* ray_sh, float* tcache, int s1, unsigned long long* frame_tot);
__global__ void k_round_sums(int n, const int* counts, const uint8_t* dead, int s0, int s1, int* block_sums);
__global__ void k_rays_given_pixels(const MarchParams P, int pix0, int n, float4* rays, float* fars);
__global__ void k_count_given(const MarchParams P, int n, const float4* rays, float2* far_tf, int* counts,
                              int* block_sums);
__global__ void k_scan_blocks(int* block_sums, int n_blocks, unsigned long long* total, unsigned long long capacity);
__global__ void k_emit(const MarchParams P, int n, const float4* rays, const float2* far_tf, const int* counts,
                       const int* block_bases, int* warp_base, float4* samples, float* t_after, int* sample_ray,
                       float* dbg, const long long* dbg_offsets, long long capacity, const float* tcache,
                       float* t_rec, int s0, int s1, const uint8_t* dead, float* t_resume);
__global__ void k_build_occupancy(const float* grid, uint32_t n_cells, float thresh, uint32_t* bits);
__global__ void k_cast_half(const float* in, __half* out, size_t n);
__global__ void k_probe_sh(int n, const float4* dirs01, uint4* out);
__global__ void k_level_scales(int n_levels, float log2_per_level_scale, uint32_t base_resolution, float* scales,
                               uint32_t* resolutions);
__global__ void k_accumulate(int n, const float4* frame, float4* accum, float sample_count, int color_space);
__global__ void k_tonemap(int n, const float4* accum, float exposure, float4 background, int color_space,
                          int output_color_space, int curve, int clamp_output, float4* out);
__global__ void k_dg_positions(uint32_t H, float bound_c, float half, float pos_w, unsigned long long seed, int jitter,
                               unsigned long long elem0, float4* samples);
__global__ void k_dg_update(uint32_t n, const uint2* sample_out, float decay, int reset, float* grid, double* sum);
__global__ void k_composite(int n_rays, ShardMap map, const int* warp_base, const int* counts, const float4* rays,
                            const float2* far_tf, const SampleSrc S, const uint2* sample_out,
                            float* rgb_f32, float* depth_f32, int* used_counts, uint8_t* rgb_u8, uint8_t* depth_u8,
                            int u8_pixel_order, const RoundWin R, unsigned long long* used_total);
size_t nrf_network_smem_bytes();
cudaError_t nrf_launch_network(int mode, int grid, cudaStream_t stream, const NetParams* P, int n,
                               const float4* samples, const int* sample_ray, const float4* rays, const uint4* ray_sh,
                               int dirs_mode, uint4* enc_io, uint2* out, uint32_t* idx_out,
                               const unsigned long long* n_dev, const SampleSrc* src);
cudaError_t nrf_launch_encode_exp(int variant, int sms, cudaStream_t stream, const NetParams* P, int n,
                                  const float4* samples, uint4* enc_io);
cudaError_t nrf_launch_network_simt(cudaStream_t stream, const NetParams* P, int n, const uint4* enc,
                                    const float4* dirs01, uint2* out);
cudaError_t nrf_launch_probe(int mode, int sms, cudaStream_t stream, const void* buf, size_t bytes, int work,
                             uint32_t* sink);
}

// ---------------------------------------------------------------------------------------------------- errors
static thread_local char g_err[512] = "";
static int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
#define CU(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e__ = (call);                                                                      \
    if (e__ != cudaSuccess)                                                                        \
      return fail(NRF_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
  } while (0)

extern "C" const char* nrf_last_error(void) { return g_err; }
extern "C" const char* nrf_version(void) { return "nerf_b200 0.1 (sm_100a)"; }

// ---------------------------------------------------------------------------------------------------- level table
// nerf_render.cu:147-165 (defaults) + tcnn grid.h:899-931 (offset table) + grid.h:186-187 (scale / resolution).
// Computed on the HOST once; the kernels read scale/res from the table instead of re-deriving them with the
// device's exp2f (DESIGN.md numerics contract).
struct HostLevels {
  uint32_t offsets[NRF_MAX_LEVELS + 1];
  float scales[NRF_MAX_LEVELS];
  uint32_t res[NRF_MAX_LEVELS];
  float per_level_scale;
};

static int build_levels(const nrf_model_desc& d, HostLevels& lv) {
  if (d.n_levels == 0 || d.n_levels > NRF_MAX_LEVELS) return fail(NRF_EINVAL, "n_levels=%u out of range", d.n_levels);
  if (d.log2_hashmap_size == 0 || d.log2_hashmap_size > 28) return fail(NRF_EINVAL, "log2_hashmap_size out of range");
  uint32_t base = d.base_resolution ? d.base_resolution : 1u << (d.log2_hashmap_size / 3);
  float pls = d.per_level_scale;
  if (pls <= 0.0f && d.n_levels > 1) pls = std::exp(std::log(2048.0f * (float)d.bound / (float)base) / (d.n_levels - 1));
  lv.per_level_scale = pls;
  const float log2_pls = std::log2(pls);
  uint32_t offset = 0;
  for (uint32_t i = 0; i < d.n_levels; ++i) {
    const float scale = exp2f(i * log2_pls) * base - 1.0f;
    const uint32_t resolution = (uint32_t)(ceilf(scale)) + 1;
    const uint32_t max_params = std::numeric_limits<uint32_t>::max() / 2;
    uint32_t n = std::pow((float)resolution, 3) > (float)max_params ? max_params : resolution * resolution * resolution;
    n = (n + 7u) / 8u * 8u;
    n = std::min(n, 1u << d.log2_hashmap_size);
    lv.offsets[i] = offset;
    lv.scales[i] = scale;
    lv.res[i] = resolution;
    offset += n;
  }
  lv.offsets[d.n_levels] = offset;
  return NRF_OK;
}

static int check_arch(const nrf_model_desc& d) {
  if (d.n_levels != 16 || d.n_features_per_level != 2 || d.n_neurons != 64 || d.density_hidden_layers != 1 ||
      d.rgb_hidden_layers != 2 || d.sh_degree != 4)
    return fail(NRF_EINVAL,
                "unsupported architecture (need L=16 F=2 width=64 density_hidden=1 rgb_hidden=2 sh_degree=4, got L=%u "
                "F=%u width=%u dh=%u rh=%u sh=%u)",
                d.n_levels, d.n_features_per_level, d.n_neurons, d.density_hidden_layers, d.rgb_hidden_layers,
                d.sh_degree);
  if (d.rgb_activation > NRF_ACT_SIGMOID || d.sigma_activation > NRF_ACT_SIGMOID)
    return fail(NRF_EINVAL, "unsupported activation id");
  if (!(d.bound > 0) || d.cascades == 0 || d.cascades > 8 || d.grid_size == 0 || d.grid_size > 512 ||
      d.grid_size < 2 || (d.grid_size * d.grid_size * d.grid_size) % 32 != 0)
    return fail(NRF_EINVAL, "bad scene description (bound/cascades/grid_size)");
  return NRF_OK;
}

extern "C" int nrf_level_table(const nrf_model_desc* desc, uint32_t* offsets, float* scales, uint32_t* resolutions) {
  if (!desc) return fail(NRF_EINVAL, "null desc");
  HostLevels lv;
  int rc = build_levels(*desc, lv);
  if (rc) return rc;
  for (uint32_t i = 0; i <= desc->n_levels; ++i)
    if (offsets) offsets[i] = lv.offsets[i];
  for (uint32_t i = 0; i < desc->n_levels; ++i) {
    if (scales) scales[i] = lv.scales[i];
    if (resolutions) resolutions[i] = lv.res[i];
  }
  return NRF_OK;
}

extern "C" int nrf_param_count(const nrf_model_desc* desc, uint64_t* n_params) {
  if (!desc || !n_params) return fail(NRF_EINVAL, "null argument");
  HostLevels lv;
  int rc = build_levels(*desc, lv);
  if (rc) return rc;
  *n_params = 10240ull + 2ull * lv.offsets[desc->n_levels];
  return NRF_OK;
}

// fp32 -> fp16 bits on the host (RN), used only to pre-pack the 20 KB of MLP weights
static uint16_t f2h_bits(float f) {
  __half h = __float2half_rn(f);
  uint16_t b;
  std::memcpy(&b, &h, 2);
  return b;
}

// Pack W[n_out][n_in] (row-major, reference order fully_fused_mlp.cu:662-686) into the tcgen05 K-major SWIZZLE_NONE
// core-matrix order used by kernels_network.cu: byte offset (k/8)*(N*16) + n*16 + (k%8)*2.
static void pack_weight(const float* W, int n_out, int n_in, uint16_t* dst) {
  for (int n = 0; n < n_out; ++n)
    for (int k = 0; k < n_in; ++k) dst[(k / 8) * (n_out * 8) + n * 8 + (k % 8)] = f2h_bits(W[n * n_in + k]);
}

// ---------------------------------------------------------------------------------------------------- context
#define NRF_MAX_ROUNDS 4
#define NRF_BATCH_STAGES 3   // host staging slots of nrf_render_batch; < 4 = the ring of per-frame counters (Device::total_h)
// device counters of a frame (Device::total; written by k_scan_blocks / k_rays_count / k_composite)
enum { TOT_ROUND = 0, TOT_MAX = 1, TOT_OVF_EVENTS = 2, TOT_MARCHED = 3, TOT_EVAL = 4, TOT_USED = 5, TOT_FRAME_OVF = 6, TOT_PEER_TIMEOUT = 7, TOT_N = 8 };

template <typename T>
struct DevBuf {
  T* p = nullptr;
  size_t n = 0;
  cudaError_t ensure(size_t want) {
    if (want <= n) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr;
    n = 0;
    cudaError_t e = cudaMalloc(&p, want * sizeof(T));
    if (e == cudaSuccess) n = want;
    return e;
  }
  void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
};

struct Device {
  int dev = 0, n_sms = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[6] = {};
  cudaEvent_t done = nullptr;  // composite finished (cross-device wait before the frame copy)
  ShardMap map{0, 1, 0, 0, 0, 0};
  int n_local = 0;
  // model replica
  DevBuf<__half> params_h;     // all params as fp16 (mlp row-major 10240 | table)
  DevBuf<uint4> mlp_packed;    // 1280 x 16 B
  DevBuf<uint32_t> occupancy;
  DevBuf<float> density_grid;  // float grid [C*H^3], kept for generate_density_grid's EMA update
  DevBuf<float> div_lut;
  NetParams net{};
  MarchParams march{};
  // per-frame buffers
  DevBuf<float4> rays, samples;
  DevBuf<uint4> ray_sh;
  DevBuf<float2> far_tf;
  DevBuf<float> t0rec;          // the t of every sample of the frame (SampleSrc)
  DevBuf<float> t_after, rgb_f32, depth_f32, tcache;
  DevBuf<int> counts, warp_base, block_sums, sample_ray, used;
  DevBuf<uint2> sample_out;
  DevBuf<uint8_t> rgb_u8, depth_u8;
  DevBuf<unsigned long long> total;        // TOT_N counters (below)
  unsigned long long* total_h = nullptr;  // pinned: ring of 4 frames x TOT_N
  // termination-aware frames (round schedule): per-ray accumulators, terminated flags and march state between rounds
  DevBuf<float4> ray_state;
  DevBuf<uint8_t> dead;
  DevBuf<float> t_resume;
  cudaEvent_t rev[NRF_MAX_ROUNDS][4] = {};   // rounds >= 1: start, after emit, after network, after composite
  int rounds_timed = 1;                     // rounds of the frame the events above belong to
  unsigned long long last_marched = 0, last_used = 0;
  // probe scratch
  DevBuf<uint4> enc;
  DevBuf<uint32_t> idx;
  DevBuf<float> dbg;
  unsigned long long last_samples = 0;
  unsigned frame_no = 0;          // frames enqueued; the counters of frame f land in total_h[TOT_N * (f & 3)]
  unsigned long long overflowed = 0;  // frames that did not fit the sample buffers (as of the last sync point)
};

// One enqueuing thread per device beyond the first (one process, several devices): the ~20 runtime calls a frame takes per
// device would otherwise be issued by ONE host thread for all devices, and at 8 devices that thread, not the GPUs, sets the
// frame rate (the reference spawns NGPU std::threads per frame for the same reason, nerf_render.cu:252-253).
struct DevWorker {
  std::thread th;
  std::mutex m;
  std::condition_variable cv;
  bool has_job = false, stop = false, done = true;
  float intr[4] = {}, pose[16] = {};
  int rc = 0;
  char err[512] = "";
};

struct nrf_ctx {
  std::vector<Device> devs;
  std::vector<std::unique_ptr<DevWorker>> workers;   // [g] for device g >= 1, created with the first multi-device frame
  // Evaluation rounds of a frame (cumulative step boundaries; the last is always NRF_MAX_STEPS_D).  One round = whole
  // rays, what the headline benchmark runs.  Termination-aware: {8, 32, 128, 1024} — rays that turned opaque in a round are
  // neither marched, evaluated nor composited in the next (the reference drops them from its next <= 8-step wavefront,
  // render_utils.h:727 + :394-415).  term_mode 0 = off, 1 = on, 2 = auto: switched on when the last validated frame
  // composited less than 3/4 of the samples its march produced, off again above 9/10.  Pixels do not depend on it.
  int term_mode = 2;
  bool chunked = false;
  int round_bounds[NRF_MAX_ROUNDS] = {8, 32, 128, NRF_MAX_STEPS_D};
  int n_round_bounds = 4;
  nrf_model_desc desc{};
  HostLevels levels{};
  float aabb[6] = {};
  bool loaded = false;
  int W = 0, H = 0;
  int shard = 0, n_shards = 1;    // process-level sharding (one process per GPU); requires a single-device context
  // sample-t cache between the two march passes (NRF_TCACHE floats per ray; NRF_NO_TCACHE=1 re-marches instead)
  bool use_tcache = std::getenv("NRF_NO_TCACHE") == nullptr;
  // peer frame of the one-process-per-GPU mode: ONE W*H*4-byte buffer (rgb W*H*3 | depth W*H) owned by the destination
  // rank and mapped into every other rank through CUDA IPC; composites store their pixels straight into it over NVLink
  uint8_t* peer_frame = nullptr;  // device pointer valid in THIS process (own allocation or IPC mapping)
  bool peer_owned = false, peer_mapped = false;
  int peer_slot = 0;              // the buffer holds TWO frames used alternately (see nrf_peer_frame_export)
  unsigned peer_seq = 0;          // frames rendered into the peer frame so far (slot = seq & 1); equal on every rank
  uint8_t* peer_host = nullptr;   // pinned 2 x W*H*4 (one per slot), destination rank only
  // one-sided completion fence (no collective): 32-bit sequence numbers behind the two frames.  Rank r's stream stores
  // `seq` into stored[r] (release, system scope) after its composite of frame `seq`; the owner's copy stream waits for
  // every stored[r] >= seq, copies the frame out if asked to, then publishes ack = seq; a rank waits for ack >= seq - 2
  // before its composite overwrites the slot frame seq - 2 was in.
  cudaStream_t peer_copy_stream = nullptr;   // owner only: fence wait + D2H run here, beside the next frame's kernels
  cudaEvent_t peer_copied[2] = {nullptr, nullptr};   // owner: slot consumed (recorded on peer_copy_stream)
  bool peer_copied_valid[2] = {false, false};
  cudaEvent_t peer_done[2] = {nullptr, nullptr};     // every rank: this rank's own kernels of the frame in that slot finished
  bool peer_copy_out = false;     // owner: copy every frame to peer_host[slot] behind the fence
  uint8_t* rgb_host = nullptr;    // pinned, W*H*3 (full frame when n_shards == 1, else this shard's packed tiles)
  uint8_t* depth_host = nullptr;  // pinned, W*H
  uint8_t* batch_stage[NRF_BATCH_STAGES] = {};   // pinned staging of nrf_render_batch's pipeline
  cudaEvent_t batch_ev[NRF_BATCH_STAGES] = {};
  cudaStream_t batch_copy_stream = nullptr;   // device 0: the frame's D2H runs here, beside the next frame's kernels
  // set while nrf_render_batch pipelines frames on several devices: the event after which device 0's frame may be
  // overwritten (its copy to the staging buffer has finished); the other devices' pixel-writing composite waits on it
  cudaEvent_t pixel_wait = nullptr;
  size_t batch_bytes = 0;
  nrf_stats stats{};
  // a frame that was enqueued without a host sync (its sample total is still unchecked against the buffer capacity)
  bool pending = false, pending_to_host = false;
  float pending_intr[4] = {}, pending_pose[16] = {};
};

static int device_count() {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

struct nrf_ctx;
static int peer_detach(nrf_ctx* c);
static void stop_workers(nrf_ctx* c);
static int frame_resolve(nrf_ctx* c);  // host sync point of an asynchronously enqueued frame (defined with the frame code)

extern "C" int nrf_create(int n_gpus, const int* device_ids, nrf_ctx** out) {
  if (!out || n_gpus < 1) return fail(NRF_EINVAL, "nrf_create: bad arguments");
  const int avail = device_count();
  if (avail < 1) return fail(NRF_ENODEV, "no CUDA device available (this library has no CPU path)");
  if (n_gpus > avail && !device_ids) return fail(NRF_ENODEV, "requested %d GPUs, %d available", n_gpus, avail);
  nrf_ctx* c = new nrf_ctx();
  c->devs.resize(n_gpus);
  for (int g = 0; g < n_gpus; ++g) {
    Device& D = c->devs[g];
    D.dev = device_ids ? device_ids[g] : g;
    cudaError_t e = cudaSetDevice(D.dev);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&D.stream, cudaStreamNonBlocking);
    for (int k = 0; k < 6 && e == cudaSuccess; ++k) e = cudaEventCreate(&D.ev[k]);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&D.done, cudaEventDisableTiming);
    for (int r = 1; r < NRF_MAX_ROUNDS && e == cudaSuccess; ++r)
      for (int k = 0; k < 4 && e == cudaSuccess; ++k) e = cudaEventCreate(&D.rev[r][k]);
    if (e == cudaSuccess) e = cudaMallocHost(&D.total_h, 4 * TOT_N * sizeof(unsigned long long));
    if (e == cudaSuccess) std::memset(D.total_h, 0, 4 * TOT_N * sizeof(unsigned long long));
    cudaDeviceProp prop;
    if (e == cudaSuccess) e = cudaGetDeviceProperties(&prop, D.dev);
    if (e != cudaSuccess) {
      delete c;
      return fail(NRF_ECUDA, "device %d init failed: %s", D.dev, cudaGetErrorString(e));
    }
    if (prop.major != 10) {
      delete c;
      return fail(NRF_ENODEV, "device %d is sm_%d%d; this library contains sm_100a code only", D.dev, prop.major, prop.minor);
    }
    D.n_sms = prop.multiProcessorCount;
  }
  *out = c;
  return NRF_OK;
}

static void free_device(Device& D) {
  cudaSetDevice(D.dev);
  D.params_h.release(); D.mlp_packed.release(); D.occupancy.release(); D.density_grid.release(); D.div_lut.release();
  D.rays.release(); D.ray_sh.release(); D.tcache.release(); D.t0rec.release(); D.samples.release(); D.far_tf.release(); D.t_after.release(); D.rgb_f32.release();
  D.depth_f32.release(); D.counts.release(); D.warp_base.release(); D.block_sums.release();
  D.sample_ray.release(); D.used.release(); D.sample_out.release(); D.rgb_u8.release(); D.depth_u8.release();
  D.total.release(); D.enc.release(); D.idx.release(); D.dbg.release();
  D.ray_state.release(); D.dead.release(); D.t_resume.release();
  for (int r = 1; r < NRF_MAX_ROUNDS; ++r) for (int k = 0; k < 4; ++k) if (D.rev[r][k]) cudaEventDestroy(D.rev[r][k]);
  if (D.total_h) cudaFreeHost(D.total_h);
  for (int k = 0; k < 6; ++k) if (D.ev[k]) cudaEventDestroy(D.ev[k]);
  if (D.done) cudaEventDestroy(D.done);
  if (D.stream) cudaStreamDestroy(D.stream);
}

extern "C" int nrf_destroy(nrf_ctx* c) {
  if (!c) return NRF_OK;
  c->pending = false;
  stop_workers(c);
  for (Device& D : c->devs) { cudaSetDevice(D.dev); if (D.stream) cudaStreamSynchronize(D.stream); }
  if (c->peer_frame) peer_detach(c);
  if (c->peer_host) cudaFreeHost(c->peer_host);
  cudaSetDevice(c->devs[0].dev);
  if (c->batch_copy_stream) { cudaStreamSynchronize(c->batch_copy_stream); cudaStreamDestroy(c->batch_copy_stream); }
  for (int k = 0; k < NRF_BATCH_STAGES; ++k) {
    if (c->batch_stage[k]) cudaFreeHost(c->batch_stage[k]);
    if (c->batch_ev[k]) cudaEventDestroy(c->batch_ev[k]);
  }
  for (Device& D : c->devs) free_device(D);
  if (c->rgb_host) cudaFreeHost(c->rgb_host);
  if (c->depth_host) cudaFreeHost(c->depth_host);
  delete c;
  return NRF_OK;
}

// Upload a model replica to one device: params fp32 -> fp16 (nerf_network.h:434-436), MLP pre-packed for UMMA,
// density grid -> occupancy bitfield with thresh = min(0.01, mean_density) (render_utils.h:560,619).
static int upload_model(nrf_ctx* c, Device& D, const float* params, size_t n_params, const float* grid, size_t n_grid) {
  CU(cudaSetDevice(D.dev));
  const nrf_model_desc& d = c->desc;
  DevBuf<float> tmp;
  CU(tmp.ensure(n_params));
  CU(D.params_h.ensure(n_params));
  CU(cudaMemcpyAsync(tmp.p, params, n_params * sizeof(float), cudaMemcpyHostToDevice, D.stream));
  k_cast_half<<<(unsigned)((n_params + 255) / 256), 256, 0, D.stream>>>(tmp.p, D.params_h.p, n_params);
  CU(cudaGetLastError());
  CU(cudaStreamSynchronize(D.stream));
  // occupancy
  CU(D.occupancy.ensure(n_grid / 32));
  CU(D.density_grid.ensure(n_grid));
  CU(cudaMemcpyAsync(D.density_grid.p, grid, n_grid * sizeof(float), cudaMemcpyHostToDevice, D.stream));
  k_build_occupancy<<<(unsigned)((n_grid + 255) / 256), 256, 0, D.stream>>>(D.density_grid.p, (uint32_t)n_grid,
                                                                          fminf(0.01f, d.mean_density), D.occupancy.p);
  CU(cudaGetLastError());
  // packed MLP weights
  std::vector<uint16_t> packed(10240);
  pack_weight(params + 0, 64, 32, packed.data() + 0);
  pack_weight(params + 2048, 16, 64, packed.data() + 2048);
  pack_weight(params + 3072, 64, 32, packed.data() + 3072);
  pack_weight(params + 5120, 64, 64, packed.data() + 5120);
  pack_weight(params + 9216, 16, 64, packed.data() + 9216);
  // division table of the march: lut[m] = (float)m / (float)(H-1), IEEE division on the host (kernels_exact.cu)
  // second half: fl(2*lut[m] - 1), the cell-exit plane in [-1,1] (render_utils.h:636-638), same two roundings
  std::vector<float> lut(2 * (d.grid_size + 1));
  for (uint32_t m = 0; m <= d.grid_size; ++m) {
    const float q = (float)m / (float)(d.grid_size - 1);
    lut[m] = q;
    lut[d.grid_size + 1 + m] = q * 2.0f - 1.0f;
  }
  CU(D.div_lut.ensure(lut.size()));
  CU(cudaMemcpyAsync(D.div_lut.p, lut.data(), lut.size() * sizeof(float), cudaMemcpyHostToDevice, D.stream));
  CU(D.mlp_packed.ensure(1280));
  CU(cudaMemcpyAsync(D.mlp_packed.p, packed.data(), 20480, cudaMemcpyHostToDevice, D.stream));
  CU(cudaStreamSynchronize(D.stream));
  tmp.release();
  // per-level scale / resolution as the reference's kernel derives them on the device (grid.h:186-187); the offset
  // table stays the host's (grid.h:899-931 runs in the GridEncoding constructor)
  {
    DevBuf<float> sc;
    DevBuf<uint32_t> rs;
    CU(sc.ensure(NRF_MAX_LEVELS)); CU(rs.ensure(NRF_MAX_LEVELS));
    k_level_scales<<<1, 32, 0, D.stream>>>((int)d.n_levels, std::log2(c->levels.per_level_scale), c->desc.base_resolution, sc.p, rs.p);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(c->levels.scales, sc.p, d.n_levels * sizeof(float), cudaMemcpyDeviceToHost, D.stream));
    CU(cudaMemcpyAsync(c->levels.res, rs.p, d.n_levels * sizeof(uint32_t), cudaMemcpyDeviceToHost, D.stream));
    CU(cudaStreamSynchronize(D.stream));
    sc.release(); rs.release();
  }
  // parameter blocks
  NetParams& N = D.net;
  for (uint32_t l = 0; l < d.n_levels; ++l) {
    LevelInfo& L = N.lv[l];
    L.offset = c->levels.offsets[l];
    L.size = c->levels.offsets[l + 1] - c->levels.offsets[l];
    L.res = c->levels.res[l];
    L.scale = c->levels.scales[l];
    // grid_index (grid.h:100-117) decides in 32-bit arithmetic: stride *= res for every dimension while stride <= T_l,
    // then "hash iff T_l < stride".  The early stop keeps most fine levels from overflowing, but when res^2 <= T_l < res^3
    // the third multiplication can still wrap below T_l (res 2048 at T = 2^22..2^24: stride = 2^33 -> 0; res 2581 at
    // bound 16): the reference then uses the WRAPPED dense index (x + y*res + z*res^2) mod T_l instead of the hash.
    // Followed as is (mode 3) — the reference's exported tables were trained with exactly that addressing.
    uint32_t stride = 1, dims = 0;
    for (; dims < 3 && stride <= L.size; ++dims) stride *= L.res;
    const bool ref_hashed = L.size < stride;
    const double dense = (double)L.res * L.res * L.res;
    L.hashed = ref_hashed ? 1u : (dense > (double)L.size ? 3u : 0u);
    if (L.hashed == 3u && dims != 3) return fail(NRF_EINVAL, "level %u (res %u, size %u): unsupported index overflow", l, L.res, L.size);
    if (L.hashed && (L.size & (L.size - 1)) != 0) return fail(NRF_EINVAL, "hashed level %u size %u not a power of two", l, L.size);
    // hashed levels below pair_from use plain 4-byte loads (2), from pair_from on the paired 8-byte slot loads (1)
    // (measured on B200: C3 network 5.68 / 5.62 / 5.69 / 5.78 ms at pair_from 0 / 8 / 12->5.57 / 16; C2 prefers 16 by 3 %:
    // pairing trades ~5 instructions per corner pair for fewer L1 wavefronts, which pays only on the finest levels)
    static const int pair_from = std::getenv("NRF_PAIR_FROM") ? std::atoi(std::getenv("NRF_PAIR_FROM")) : 12;
    if (L.hashed == 1u && (int)l < pair_from) L.hashed = 2u;
  }
  N.mlp_rowmaj = D.params_h.p;
  N.table = reinterpret_cast<const __half2*>(D.params_h.p + 10240);
  // experiment knob (A/B in DESIGN.md §4, off by default): NRF_L2_PERSIST=<MB> sets aside that much L2 for persisting lines
  // and marks the hash table as the stream's persisting access window, everything else as streaming
  if (const char* e = std::getenv("NRF_L2_PERSIST")) {
    const size_t want = (size_t)std::atoi(e) << 20;
    if (want > 0) {
      cudaDeviceProp prop;
      CU(cudaGetDeviceProperties(&prop, D.dev));
      const size_t table_bytes = (size_t)c->levels.offsets[d.n_levels] * 4;
      CU(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, std::min(want, (size_t)prop.persistingL2CacheMaxSize)));
      cudaStreamAttrValue attr{};
      attr.accessPolicyWindow.base_ptr = const_cast<__half*>(D.params_h.p + 10240);
      attr.accessPolicyWindow.num_bytes = std::min(table_bytes, (size_t)prop.accessPolicyMaxWindowSize);
      attr.accessPolicyWindow.hitRatio = std::min(1.0f, (float)std::min(want, (size_t)prop.persistingL2CacheMaxSize) / (float)table_bytes);
      attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
      attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
      CU(cudaStreamSetAttribute(D.stream, cudaStreamAttributeAccessPolicyWindow, &attr));
    }
  }
  for (uint32_t l = 0; l < d.n_levels; ++l) N.lv[l].tab = reinterpret_cast<const uint32_t*>(N.table) + N.lv[l].offset;
  N.mlp_packed = D.mlp_packed.p;
  N.rgb_act = d.rgb_activation;
  N.sigma_act = d.sigma_activation;
  MarchParams& M = D.march;
  std::memcpy(M.aabb, c->aabb, 24);
  M.bound = d.bound;
  M.dt_gamma = 1.0 / 128;                         // nerf_render.h:66
  M.dt_min = 2 * 1.7320508075688772f / 1024;      // render_utils.h:179-183
  M.dt_max = 2 * d.bound / d.grid_size;           // render_utils.h:581
  M.min_near = 0.2f;                              // nerf_render.h:76
  M.pos_w = 1.0 / (2 * d.bound);                  // nerf_render.cu:311-312
  M.C = d.cascades;
  M.H = d.grid_size;
  M.rbound = 1.0f / d.bound;
  M.occupancy = D.occupancy.p;
  M.div_lut = D.div_lut.p;
  {
    auto pow2f = [](float v) { int e; return v > 0 && std::frexp(v, &e) == 0.5f; };
    const bool h_pow2 = (d.grid_size & (d.grid_size - 1)) == 0;
    M.fast = (pow2f(M.dt_gamma) && h_pow2 && d.cascades <= 24 && d.bound == (float)(1u << (d.cascades - 1)) &&
              !std::getenv("NRF_MARCH_GENERIC")) ? 1u : 0u;
    M.t_lo = M.dt_min / M.dt_gamma;   // exact: dt_gamma is a power of two whenever `fast` is set
    M.t_hi = M.dt_max / M.dt_gamma;
  }
  return NRF_OK;
}

extern "C" int nrf_load_raw(nrf_ctx* c, const nrf_model_desc* desc, const float* params, size_t n_params,
                            const float* density_grid, size_t n_grid, const float aabb[6]) {
  if (!c || !desc || !params || !density_grid || !aabb) return fail(NRF_EINVAL, "nrf_load_raw: null argument");
  { int rc_ = frame_resolve(c); if (rc_) return rc_; }
  int rc = check_arch(*desc);
  if (rc) return rc;
  HostLevels lv;
  rc = build_levels(*desc, lv);
  if (rc) return rc;
  const size_t want = 10240 + 2 * (size_t)lv.offsets[desc->n_levels];
  if (n_params != want)  // nerf_network.h:425-427 param-count check
    return fail(NRF_EFORMAT, "Can't set params because CPU buffer has the wrong size: got %zu, expected %zu", n_params, want);
  const size_t cells = (size_t)desc->grid_size * desc->grid_size * desc->grid_size * desc->cascades;
  if (n_grid != cells) return fail(NRF_EFORMAT, "Incompatible number of grid cascades.");  // nerf_render.cu:467-469
  c->desc = *desc;
  c->desc.per_level_scale = lv.per_level_scale;
  if (!c->desc.base_resolution) c->desc.base_resolution = 1u << (desc->log2_hashmap_size / 3);
  c->levels = lv;
  std::memcpy(c->aabb, aabb, 24);
  for (Device& D : c->devs) {
    rc = upload_model(c, D, params, n_params, density_grid, n_grid);
    if (rc) return rc;
  }
  c->loaded = true;
  return NRF_OK;
}

// Reads and parses a snapshot file on the host (no device needed); uses / refreshes the binary side-cache.
static int read_snapshot(const char* path, nrf::Snapshot& snap) {
  const std::string p(path);
  // reload_network_from_file accepts only .msgpack (nerf_render.cu:98,105-107)
  if (!nrf::has_ext(p, "msgpack")) return fail(NRF_EIO, "Network config %s must be a .msgpack snapshot.", path);
  try {
    struct stat st;
    if (::stat(path, &st) != 0) return fail(NRF_EIO, "Network config %s does not exist.", path);  // nerf_render.cu:73-76
    const bool use_cache = nrf::cache_enabled((size_t)st.st_size);
    if (use_cache && nrf::cache_load(p, snap)) return NRF_OK;
    const std::vector<uint8_t> buf = nrf::read_file(p);
    nrf::parse_snapshot(buf.data(), buf.size(), p, snap);
    if (use_cache) nrf::cache_store(p, snap);
  } catch (const std::exception& e) {
    const bool missing = std::strstr(e.what(), "does not exist") != nullptr;
    return fail(missing ? NRF_EIO : NRF_EFORMAT, "%s", e.what());
  }
  return NRF_OK;
}

extern "C" int nrf_load_snapshot(nrf_ctx* c, const char* path) {
  if (!c || !path) return fail(NRF_EINVAL, "nrf_load_snapshot: null argument");
  nrf::Snapshot snap;
  int rc = read_snapshot(path, snap);
  if (rc) return rc;
  return nrf_load_raw(c, &snap.desc, snap.params.data(), snap.params.size(), snap.density_grid.data(),
                      snap.density_grid.size(), snap.aabb);
}

extern "C" int nrf_snapshot_inspect(const char* path, nrf_model_desc* desc, uint64_t* n_params, uint64_t* n_grid,
                                    float aabb[6], double* params_sum, double* grid_sum) {
  if (!path) return fail(NRF_EINVAL, "nrf_snapshot_inspect: null path");
  const std::string p(path);
  if (nrf::has_ext(p, "json")) {
    // load_network_config's .json branch (nerf_render.cu:79-82): comments allowed, "parent" chain merged (RFC 7386).
    // A .json config carries no weights: the array outputs are zero.
    try {
      const nrf::Val cfg = nrf::load_json_config(p);
      nrf_model_desc d;
      nrf::describe_model(cfg, d);
      if (desc) *desc = d;
    } catch (const std::exception& e) {
      const bool missing = std::strstr(e.what(), "does not exist") != nullptr;
      return fail(missing ? NRF_EIO : NRF_EFORMAT, "%s", e.what());
    }
    if (n_params) *n_params = 0;
    if (n_grid) *n_grid = 0;
    if (aabb) std::memset(aabb, 0, 24);
    if (params_sum) *params_sum = 0;
    if (grid_sum) *grid_sum = 0;
    return NRF_OK;
  }
  nrf::Snapshot snap;
  int rc = read_snapshot(path, snap);
  if (rc) return rc;
  if (desc) *desc = snap.desc;
  if (n_params) *n_params = snap.params.size();
  if (n_grid) *n_grid = snap.density_grid.size();
  if (aabb) std::memcpy(aabb, snap.aabb, 24);
  if (params_sum) { double s = 0; for (float v : snap.params) s += v; *params_sum = s; }
  if (grid_sum) { double s = 0; for (float v : snap.density_grid) s += v; *grid_sum = s; }
  return NRF_OK;
}

// Per-level grid scale and resolution of the loaded model as evaluated on the device (see upload_model).
extern "C" int nrf_get_level_scales(nrf_ctx* c, float* scales, uint32_t* resolutions) {
  if (!c) return fail(NRF_EINVAL, "null ctx");
  if (!c->loaded) return fail(NRF_ESTATE, "no model loaded");
  for (uint32_t l = 0; l < c->desc.n_levels; ++l) {
    if (scales) scales[l] = c->levels.scales[l];
    if (resolutions) resolutions[l] = c->levels.res[l];
  }
  return NRF_OK;
}

extern "C" int nrf_get_model_desc(nrf_ctx* c, nrf_model_desc* out) {
  if (!c || !out) return fail(NRF_EINVAL, "null argument");
  if (!c->loaded) return fail(NRF_ESTATE, "no model loaded");
  *out = c->desc;
  return NRF_OK;
}

// ---------------------------------------------------------------------------------------------------- resolution
extern "C" int nrf_set_shard(nrf_ctx* c, int shard, int n_shards) {
  if (!c || n_shards < 1 || shard < 0 || shard >= n_shards) return fail(NRF_EINVAL, "nrf_set_shard: bad arguments");
  { int rc_ = frame_resolve(c); if (rc_) return rc_; }
  if (n_shards > 1 && c->devs.size() != 1)
    return fail(NRF_EINVAL, "process-level sharding needs a single-device context (one process per GPU)");
  if (c->peer_frame) return fail(NRF_ESTATE, "nrf_set_shard: detach the peer frame first");
  c->shard = shard;
  c->n_shards = n_shards;
  c->W = c->H = 0;  // force re-allocation on the next set_resolution
  return NRF_OK;
}

// NerfRender::set_resolution (nerf_render.cu:186-236).  Idempotent (reference defect D7).
extern "C" int nrf_set_resolution(nrf_ctx* c, int W, int H) {
  if (!c || W < 1 || H < 1) return fail(NRF_EINVAL, "nrf_set_resolution: bad arguments");
  if (!c->loaded) return fail(NRF_ESTATE, "set_resolution must follow reload_network_from_file");  // nerf_render.cu:226
  if ((long long)W * H > (1ll << 29)) return fail(NRF_EINVAL, "resolution too large");
  { int rc = frame_resolve(c); if (rc) return rc; }
  if (c->W == W && c->H == H) return NRF_OK;
  if (c->peer_frame) return fail(NRF_ESTATE, "nrf_set_resolution: detach the peer frame before changing the resolution");
  if (c->peer_host) { cudaFreeHost(c->peer_host); c->peer_host = nullptr; }
  const int n_dev = (int)c->devs.size();
  const int total_shards = c->n_shards * n_dev;
  // Several devices in one process write their pixels straight into device 0's frame over NVLink (peer stores)
  for (int g = 1; g < n_dev; ++g) {
    int can = 0;
    CU(cudaDeviceCanAccessPeer(&can, c->devs[g].dev, c->devs[0].dev));
    if (!can) return fail(NRF_ENODEV, "device %d cannot access device %d (peer access is required for n_gpus > 1)",
                          c->devs[g].dev, c->devs[0].dev);
    CU(cudaSetDevice(c->devs[g].dev));
    cudaError_t e = cudaDeviceEnablePeerAccess(c->devs[0].dev, 0);
    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CU(e);
    cudaGetLastError();
  }
  for (int g = 0; g < n_dev; ++g) {
    Device& D = c->devs[g];
    CU(cudaSetDevice(D.dev));
    D.map = make_shard_map(c->shard * n_dev + g, total_shards, W, H);
    D.n_local = shard_local_count(D.map);
    const size_t N = (size_t)std::max(D.n_local, 32);
    const size_t nb = (N + 127) / 128;
    CU(D.rays.ensure(N)); CU(D.ray_sh.ensure(2 * N)); CU(D.far_tf.ensure(N)); CU(D.counts.ensure(N)); CU(D.warp_base.ensure((N + 31) / 32));
    CU(D.block_sums.ensure(nb)); CU(D.used.ensure(N)); CU(D.rgb_f32.ensure(N * 3)); CU(D.depth_f32.ensure(N));
    CU(D.total.ensure(TOT_N));
    CU(cudaMemsetAsync(D.total.p, 0, TOT_N * sizeof(unsigned long long), D.stream));
    CU(D.ray_state.ensure(2 * N)); CU(D.dead.ensure(N)); CU(D.t_resume.ensure(N));
    if (c->use_tcache) CU(D.tcache.ensure(((N + 31) / 32) * 32 * NRF_TCACHE));
    // u8 outputs: whole-frame contexts keep ONE frame in pixel order on device 0; a sharded context keeps its packed tiles
    if (g == 0) {
      const size_t px = c->n_shards == 1 ? (size_t)W * H : N;
      CU(D.rgb_u8.ensure(px * 3)); CU(D.depth_u8.ensure(px));
      CU(cudaMemsetAsync(D.rgb_u8.p, 0, px * 3, D.stream)); CU(cudaMemsetAsync(D.depth_u8.p, 0, px, D.stream));
    }
    D.march.W = W;
  }
  if (c->rgb_host) cudaFreeHost(c->rgb_host);
  if (c->depth_host) cudaFreeHost(c->depth_host);
  c->rgb_host = c->depth_host = nullptr;
  const size_t host_px = c->n_shards == 1 ? (size_t)W * H : (size_t)std::max(c->devs[0].n_local, 1);
  CU(cudaMallocHost(&c->rgb_host, host_px * 3));
  CU(cudaMallocHost(&c->depth_host, host_px));
  c->W = W;
  c->H = H;
  return NRF_OK;
}

extern "C" int nrf_local_rays(nrf_ctx* c, int* n_local) {
  if (!c || !n_local) return fail(NRF_EINVAL, "null argument");
  if (!c->W) return fail(NRF_ESTATE, "set_resolution not called");
  int n = 0;
  for (Device& D : c->devs) n += D.n_local;
  *n_local = n;
  return NRF_OK;
}

extern "C" int nrf_shard_pixel(nrf_ctx* c, int q, int* pixel) {
  if (!c || !pixel || !c->W) return fail(NRF_EINVAL, "bad argument / state");
  if (q < 0 || q >= c->devs[0].n_local) return fail(NRF_EINVAL, "local ray index out of range");
  *pixel = shard_pixel(c->devs[0].map, q);
  return NRF_OK;
}

extern "C" int nrf_device_outputs(nrf_ctx* c, void** rgb_u8, void** depth_u8, void** rgb_f32) {
  if (!c || !c->W) return fail(NRF_ESTATE, "set_resolution not called");
  if (rgb_u8) *rgb_u8 = c->devs[0].rgb_u8.p;
  if (depth_u8) *depth_u8 = c->devs[0].depth_u8.p;
  if (rgb_f32) *rgb_f32 = c->devs[0].rgb_f32.p;
  return NRF_OK;
}

extern "C" int nrf_stream(nrf_ctx* c, void** stream) {
  if (!c || !stream) return fail(NRF_EINVAL, "null argument");
  *stream = (void*)c->devs[0].stream;
  return NRF_OK;
}

extern "C" int nrf_synchronize(nrf_ctx* c) {
  if (!c) return fail(NRF_EINVAL, "null ctx");
  { int rc = frame_resolve(c); if (rc) return rc; }
  for (Device& D : c->devs) {
    CU(cudaSetDevice(D.dev));
    CU(cudaStreamSynchronize(D.stream));
  }
  if (c->peer_copy_stream) CU(cudaStreamSynchronize(c->peer_copy_stream));
  return NRF_OK;
}

// ---------------------------------------------------------------------------------------------------- frame
// nerf_matrix_to_ngp (render_utils.h:68-77): rows (1,2,0), columns 1,2 negated, translation * scale (+ offset 0).
static void set_pose(nrf_ctx* c, MarchParams& M, const float intr[4], const float* P) {
  const int rows[3] = {1, 2, 0};
  for (int r = 0; r < 3; ++r) {
    const float* src = P + rows[r] * 4;
    M.R[r * 3 + 0] = src[0];
    M.R[r * 3 + 1] = -src[1];
    M.R[r * 3 + 2] = -src[2];
    M.o[r] = src[3] * c->desc.scale + 0.0f;
  }
  std::memcpy(M.intr, intr, 16);
}

// k_network is persistent: one 896-thread CTA per SM (7 tile groups of 128 threads), fewer CTAs when there are fewer than 8 tiles per SM.
static int network_grid(nrf_ctx*, const Device& D, int n_samples) {
  const int tiles = (n_samples + NRF_TILE - 1) / NRF_TILE;
  return std::max(1, std::min(D.n_sms, (tiles + 7) / 8));
}

static inline uint8_t* peer_slot_ptr(nrf_ctx* c) { return c->peer_frame + (size_t)c->peer_slot * c->W * c->H * 4; }

// ---- one-sided fence of the peer frame (see nrf_ctx::peer_seq) ----
#define NRF_PEER_MAX_RANKS 64
#define NRF_PEER_FLAG_STRIDE 8   // 32-bit words between two ranks' flags (one 32-byte sector each)
static inline unsigned* peer_flags(nrf_ctx* c) {
  const size_t frames = ((size_t)c->W * c->H * 4 * 2 + 255) & ~(size_t)255;
  return reinterpret_cast<unsigned*>(c->peer_frame + frames);
}
static inline size_t peer_alloc_bytes(nrf_ctx* c) {
  return (((size_t)c->W * c->H * 4 * 2 + 255) & ~(size_t)255) + (NRF_PEER_MAX_RANKS + 1) * NRF_PEER_FLAG_STRIDE * 4;
}
static inline unsigned* peer_stored_flag(nrf_ctx* c, int rank) { return peer_flags(c) + (size_t)rank * NRF_PEER_FLAG_STRIDE; }
static inline unsigned* peer_ack_flag(nrf_ctx* c) { return peer_flags(c) + (size_t)NRF_PEER_MAX_RANKS * NRF_PEER_FLAG_STRIDE; }

// everything this stream wrote before (the composite's pixel stores, local or over NVLink) is visible system-wide before
// the flag is
__global__ void k_peer_signal(unsigned* flag, unsigned seq) {
  __threadfence_system();
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(seq) : "memory");
}
// lane i polls flags[(first + i) * stride] until it has reached `seq` (wrap-safe); gives up after timeout_ns and raises the
// device counter TOT_PEER_TIMEOUT instead of hanging the GPU when a rank died
__global__ void k_peer_wait(const unsigned* flags, int first, int n, unsigned seq, unsigned long long timeout_ns,
                            unsigned long long* timed_out) {
  const int i = threadIdx.x;
  if (i >= n) return;
  const unsigned* f = flags + (size_t)(first + i) * NRF_PEER_FLAG_STRIDE;
  unsigned long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (;;) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
    if ((int)(v - seq) >= 0) break;
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    if (t - t0 > timeout_ns) { atomicAdd(timed_out, 1ull); break; }
    __nanosleep(200);
  }
}
static const unsigned long long kPeerTimeoutNs = 20ull * 1000 * 1000 * 1000;

// One frame on one device, enqueued WITHOUT a host round trip: rays + count + scan leave the sample total in device
// memory, k_emit / k_network / k_composite read it there; the sample buffers keep the capacity of earlier frames.
// If the total exceeds the capacity the overflowing patches are skipped on the device (k_emit marks them) and
// frame_check() — at the next host sync — grows the buffers and asks for the frame to be rendered again.
// samples one frame can hold: the three per-sample buffers normally grow together, but the density-grid pass sizes
// two of them on its own
static size_t frame_capacity(const Device& D) { return std::min(D.samples.n, std::min(D.t0rec.n, D.sample_out.n)); }

static SampleSrc sample_src(const Device& D) {
  SampleSrc s{};
  s.t0 = D.t0rec.p;
  s.dt_gamma = D.march.dt_gamma; s.dt_min = D.march.dt_min; s.dt_max = D.march.dt_max;
  return s;
}

static int frame_enqueue(nrf_ctx* c, Device& D, const float intr[4], const float* pose) {
  CU(cudaSetDevice(D.dev));
  set_pose(c, D.march, intr, pose);
  const int nb = (D.n_local + 127) / 128;
  if (D.t0rec.n == 0) {  // first frame at this resolution: start from 32 samples per ray
    const size_t want = std::max<size_t>((size_t)D.n_local * 32, 4096);
    CU(D.samples.ensure(want)); CU(D.t0rec.ensure(want)); CU(D.sample_out.ensure(want));
  }
  const size_t cap = std::min<size_t>(frame_capacity(D), (1ull << 31) - 256);
  const SampleSrc src = sample_src(D);
  // round schedule of this frame: whole rays, or the termination-aware windows (nrf_ctx::round_bounds)
  int bounds[NRF_MAX_ROUNDS] = {NRF_MAX_STEPS_D};
  int n_rounds = 1;
  if (c->chunked) {
    n_rounds = c->n_round_bounds;
    for (int r = 0; r < n_rounds; ++r) bounds[r] = c->round_bounds[r];
  }
  D.rounds_timed = n_rounds;
  Device& D0 = c->devs[0];
  uint8_t* rgb_dst = c->peer_frame ? peer_slot_ptr(c) : D0.rgb_u8.p;
  uint8_t* dep_dst = c->peer_frame ? peer_slot_ptr(c) + (size_t)c->W * c->H * 3 : D0.depth_u8.p;
  const int u8_pixel_order = (c->n_shards == 1 || c->peer_frame) ? 1 : 0;
  CU(cudaEventRecord(D.ev[0], D.stream));
  // this frame's counters: marched / evaluated / used / overflow flag
  CU(cudaMemsetAsync(D.total.p + TOT_MARCHED, 0, 4 * sizeof(unsigned long long), D.stream));
  for (int r = 0; r < n_rounds; ++r) {
    const int s0 = r == 0 ? 0 : bounds[r - 1], s1 = bounds[r];
    RoundWin R{s0, s1, r == 0, r == n_rounds - 1, n_rounds > 1 ? D.ray_state.p : nullptr, n_rounds > 1 ? D.dead.p : nullptr};
    if (r > 0) CU(cudaEventRecord(D.rev[r][0], D.stream));
    if (D.n_local > 0) {
      if (r == 0) {
        // experiment knob: unused dynamic shared memory caps the resident blocks per SM (registers allow 10)
        static const int count_dyn_smem = std::getenv("NRF_COUNT_DYN_SMEM") ? std::atoi(std::getenv("NRF_COUNT_DYN_SMEM")) : 0;
        k_rays_count<<<nb, 128, count_dyn_smem, D.stream>>>(D.march, D.map, D.n_local, D.rays.p, D.far_tf.p, D.counts.p,
                                                            D.block_sums.p, D.ray_sh.p, D.tcache.p, s1, D.total.p + TOT_MARCHED);
      } else {
        k_round_sums<<<nb, 128, 0, D.stream>>>(D.n_local, D.counts.p, D.dead.p, s0, s1, D.block_sums.p);
      }
      CU(cudaGetLastError());
    }
    k_scan_blocks<<<1, 1024, 0, D.stream>>>(D.block_sums.p, D.n_local > 0 ? nb : 0, D.total.p, (unsigned long long)cap);
    CU(cudaGetLastError());
    if (r == 0) CU(cudaEventRecord(D.ev[1], D.stream));
    if (D.n_local > 0) {
      k_emit<<<nb, 128, 0, D.stream>>>(D.march, D.n_local, D.rays.p, D.far_tf.p, D.counts.p, D.block_sums.p,
                                      D.warp_base.p, D.samples.p, nullptr, nullptr, nullptr, nullptr,
                                      (long long)cap, D.tcache.p, D.t0rec.p, s0, s1, r == 0 ? nullptr : D.dead.p,
                                      n_rounds > 1 ? D.t_resume.p : nullptr);
      CU(cudaGetLastError());
    }
    CU(cudaEventRecord(r == 0 ? D.ev[2] : D.rev[r][1], D.stream));
    CU(nrf_launch_network(2, network_grid(c, D, (int)cap), D.stream, &D.net, (int)cap, D.samples.p, nullptr, D.rays.p,
                          D.ray_sh.p, 0, nullptr, D.sample_out.p, nullptr, D.total.p + TOT_ROUND, &src));
    CU(cudaEventRecord(r == 0 ? D.ev[3] : D.rev[r][2], D.stream));
    if (D.n_local > 0) {
      // stores into device 0's frame must not overtake the copy-out of the previous frame, which runs on the copy
      // stream beside this frame's kernels (nrf_render_batch)
      if (R.last && c->pixel_wait) CU(cudaStreamWaitEvent(D.stream, c->pixel_wait, 0));
      if (R.last && c->peer_frame && c->peer_seq > 2) {   // the slot still holds frame seq - 2: wait until it was consumed
        if (c->peer_owned) { if (c->peer_copied_valid[c->peer_slot]) CU(cudaStreamWaitEvent(D.stream, c->peer_copied[c->peer_slot], 0)); }
        else k_peer_wait<<<1, 32, 0, D.stream>>>(peer_ack_flag(c), 0, 1, c->peer_seq - 2, kPeerTimeoutNs, D.total.p + TOT_PEER_TIMEOUT);
      }
      k_composite<<<nb, 128, 0, D.stream>>>(D.n_local, D.map, D.warp_base.p, D.counts.p, D.rays.p, D.far_tf.p, src,
                                            D.sample_out.p, D.rgb_f32.p, D.depth_f32.p, D.used.p, rgb_dst, dep_dst,
                                            u8_pixel_order, R, D.total.p + TOT_USED);
      CU(cudaGetLastError());
    }
    CU(cudaEventRecord(r == 0 ? D.ev[4] : D.rev[r][3], D.stream));
  }
  if (c->peer_frame && !c->peer_owned) {   // this rank's pixels of frame peer_seq are in the owner's buffer
    k_peer_signal<<<1, 1, 0, D.stream>>>(peer_stored_flag(c, c->shard), c->peer_seq);
    CU(cudaGetLastError());
  }
  ++D.frame_no;
  CU(cudaMemcpyAsync(D.total_h + TOT_N * (D.frame_no & 3u), D.total.p, TOT_N * sizeof(unsigned long long), cudaMemcpyDeviceToHost, D.stream));
  CU(cudaEventRecord(D.done, D.stream));
  if (c->peer_frame) CU(cudaEventRecord(c->peer_done[c->peer_slot], D.stream));
  if (c->peer_frame && c->peer_owned) {
    // owner: behind its own composite, on the copy stream (so the next frame's kernels are not held up): wait for every
    // other rank's flag, copy the frame out if asked to, release the slot
    cudaStream_t cs = c->peer_copy_stream;
    CU(cudaStreamWaitEvent(cs, D.done, 0));
    k_peer_signal<<<1, 1, 0, cs>>>(peer_stored_flag(c, c->shard), c->peer_seq);   // the owner's own pixels
    if (c->n_shards > 1) {
      for (int first = 0; first < c->n_shards; first += 32)
        k_peer_wait<<<1, 32, 0, cs>>>(peer_flags(c), first, std::min(32, c->n_shards - first), c->peer_seq, kPeerTimeoutNs,
                                      D.total.p + TOT_PEER_TIMEOUT);
      CU(cudaGetLastError());
    }
    if (c->peer_copy_out)
      CU(cudaMemcpyAsync(c->peer_host + (size_t)c->peer_slot * c->W * c->H * 4, peer_slot_ptr(c), (size_t)c->W * c->H * 4,
                         cudaMemcpyDeviceToHost, cs));
    k_peer_signal<<<1, 1, 0, cs>>>(peer_ack_flag(c), c->peer_seq);
    CU(cudaGetLastError());
    CU(cudaEventRecord(c->peer_copied[c->peer_slot], cs));
    c->peer_copied_valid[c->peer_slot] = true;
  }
  return NRF_OK;
}

// After the device's stream has been synchronised: validates the LATEST frame's sample total (redo = it overflowed the
// buffers and must be rendered again) and grows the buffers to the largest total of ANY frame enqueued since the last
// check (a running maximum kept on the device: frames may be enqueued back to back without a host sync; only the
// latest one's result is still observable, earlier frames that overflowed are counted — nrf_overflowed_frames).
static int frame_check(nrf_ctx*, Device& D, bool& redo) {
  CU(cudaSetDevice(D.dev));
  const unsigned long long* slot = D.total_h + TOT_N * (D.frame_no & 3u);
  const unsigned long long mx = slot[TOT_MAX];   // largest round of any frame since the last check
  D.overflowed = slot[TOT_OVF_EVENTS];
  CU(cudaMemsetAsync(D.total.p + TOT_MAX, 0, sizeof(unsigned long long), D.stream));   // restart the running maximum
  if (mx > (1ull << 31) - 256) return fail(NRF_EINVAL, "frame produces %llu samples in one round (> 2^31)", mx);
  D.last_samples = slot[TOT_EVAL];
  D.last_marched = slot[TOT_MARCHED];
  D.last_used = slot[TOT_USED];
  const size_t cap = frame_capacity(D);
  if (mx > cap) {
    const size_t want = (size_t)mx + (size_t)mx / 4;
    CU(D.samples.ensure(want)); CU(D.t0rec.ensure(want)); CU(D.sample_out.ensure(want));
  }
  if (slot[TOT_FRAME_OVF]) redo = true;   // a round of the LATEST frame did not fit
  if (slot[TOT_PEER_TIMEOUT]) return fail(NRF_ECUDA, "peer frame fence timed out (%llu waits): a rank of the job did not deliver its pixels",
                                          slot[TOT_PEER_TIMEOUT]);
  return NRF_OK;
}

// Frame copy to the pinned host buffers (device 0, after every device's composite).
static int frame_to_host(nrf_ctx* c) {
  Device& D0 = c->devs[0];
  CU(cudaSetDevice(D0.dev));
  for (size_t g = 1; g < c->devs.size(); ++g) CU(cudaStreamWaitEvent(D0.stream, c->devs[g].done, 0));
  const size_t px = c->n_shards == 1 ? (size_t)c->W * c->H : (size_t)D0.n_local;
  if (px > 0) {
    CU(cudaMemcpyAsync(c->rgb_host, D0.rgb_u8.p, px * 3, cudaMemcpyDeviceToHost, D0.stream));
    CU(cudaMemcpyAsync(c->depth_host, D0.depth_u8.p, px, cudaMemcpyDeviceToHost, D0.stream));
  }
  CU(cudaEventRecord(D0.ev[5], D0.stream));
  return NRF_OK;
}

static void stop_workers(nrf_ctx* c) {
  for (auto& w : c->workers) {
    if (!w) continue;
    { std::lock_guard<std::mutex> l(w->m); w->stop = true; }
    w->cv.notify_all();
    if (w->th.joinable()) w->th.join();
  }
  c->workers.clear();
}

static void start_workers(nrf_ctx* c) {
  c->workers.resize(c->devs.size());
  for (size_t g = 1; g < c->devs.size(); ++g) {
    c->workers[g].reset(new DevWorker());
    DevWorker* w = c->workers[g].get();
    Device* D = &c->devs[g];
    w->th = std::thread([c, w, D]() {
      for (;;) {
        {
          std::unique_lock<std::mutex> l(w->m);
          w->cv.wait(l, [&] { return w->stop || w->has_job; });
          if (w->stop) return;
          w->has_job = false;
        }
        const int rc = frame_enqueue(c, *D, w->intr, w->pose);
        {
          std::lock_guard<std::mutex> l(w->m);
          w->rc = rc;
          if (rc) std::snprintf(w->err, sizeof(w->err), "%s", g_err);
          w->done = true;
        }
        w->cv.notify_all();
      }
    });
  }
}

static int frame_enqueue_all(nrf_ctx* c, const float intr[4], const float* pose, bool to_host) {
  const bool threaded = std::getenv("NRF_NO_ENQUEUE_THREADS") == nullptr;
  const char* from_env = std::getenv("NRF_ENQUEUE_THREADS_FROM");   // tests: exercise the threads on a 2-GPU box
  const size_t threads_from = from_env ? (size_t)std::max(2, std::atoi(from_env)) : 4;
  // measured (profiles/r2_ab_enqueue_threads.json): 8 devices, synchronous per-frame call 967 -> 1 040 M rays/s, batch call
  // 1 103 -> 1 125 M; 2 devices 334 -> 331 / 342 -> 334 (the hand-off costs more than two devices' launches): from 4 devices up
  if (c->devs.size() >= threads_from && threaded) {
    if (c->workers.size() != c->devs.size()) start_workers(c);
    for (size_t g = 1; g < c->devs.size(); ++g) {
      DevWorker* w = c->workers[g].get();
      {
        std::lock_guard<std::mutex> l(w->m);
        std::memcpy(w->intr, intr, 16);
        std::memcpy(w->pose, pose, 64);
        w->done = false;
        w->has_job = true;
      }
      w->cv.notify_all();
    }
    int rc = frame_enqueue(c, c->devs[0], intr, pose);
    for (size_t g = 1; g < c->devs.size(); ++g) {   // every device's `done` event is recorded before anyone waits on it
      DevWorker* w = c->workers[g].get();
      std::unique_lock<std::mutex> l(w->m);
      w->cv.wait(l, [&] { return w->done; });
      if (w->rc && !rc) { rc = w->rc; std::snprintf(g_err, sizeof(g_err), "%s", w->err); }
    }
    if (rc) return rc;
  } else {
    for (Device& D : c->devs) { int rc = frame_enqueue(c, D, intr, pose); if (rc) return rc; }
  }
  if (to_host) return frame_to_host(c);
  CU(cudaSetDevice(c->devs[0].dev));
  CU(cudaEventRecord(c->devs[0].ev[5], c->devs[0].stream));
  return NRF_OK;
}

// Host sync point: wait for the enqueued frame, validate its sample total, re-render if the buffers were too small
// (at most twice: the exact total is known after the first attempt).
static int frame_resolve(nrf_ctx* c) {
  if (!c->pending) return NRF_OK;
  for (int attempt = 0; attempt < 3; ++attempt) {
    bool redo = false;
    for (Device& D : c->devs) {
      CU(cudaSetDevice(D.dev));
      CU(cudaStreamSynchronize(D.stream));
      int rc = frame_check(c, D, redo);
      if (rc) { c->pending = false; return rc; }
    }
    // termination-aware rounds, decided from the frame just validated (both counts are schedule-independent)
    if (c->term_mode == 2 && !redo) {   // (a frame that overflowed composited only part of its samples: no basis for the ratio)
      unsigned long long marched = 0, used = 0;
      for (Device& D : c->devs) { marched += D.last_marched; used += D.last_used; }
      if (marched > 0) {
        if (4 * used < 3 * marched) c->chunked = true;
        else if (10 * used > 9 * marched) c->chunked = false;
      }
    }
    if (!redo) { c->pending = false; return NRF_OK; }
    int rc = frame_enqueue_all(c, c->pending_intr, c->pending_pose, c->pending_to_host);
    if (rc) { c->pending = false; return rc; }
  }
  c->pending = false;
  return fail(NRF_ECUDA, "frame did not fit its sample buffers after regrowing them");
}

static int render_common(nrf_ctx* c, const float intr[4], const float* pose, bool to_host, bool sync) {
  if (!c || !intr || !pose) return fail(NRF_EINVAL, "render: null argument");
  if (!c->loaded) return fail(NRF_ESTATE, "render_frame before reload_network_from_file");
  if (!c->W) return fail(NRF_ESTATE, "render_frame before set_resolution");
  // with a peer frame attached the pixels go to the owner's buffer, not to this context's frame: the host copy of the
  // synchronous call would return stale bytes
  if (to_host && c->peer_frame)
    return fail(NRF_ESTATE, "a peer frame is attached: render with nrf_render_frame_device and read the frame with nrf_peer_frame_collect");
  // No host sync here: an earlier asynchronous frame that is still unvalidated is superseded (its buffers are reused
  // in stream order; its total stays in the ring as a growth hint for the next sync point).
  int rc;
  c->peer_slot = (int)(++c->peer_seq & 1u);   // every rank renders every frame, so all ranks alternate in step
  rc = frame_enqueue_all(c, intr, pose, to_host);
  if (rc) return rc;
  c->pending = true;
  c->pending_to_host = to_host;
  std::memcpy(c->pending_intr, intr, 16);
  std::memcpy(c->pending_pose, pose, 64);
  return sync ? frame_resolve(c) : NRF_OK;
}

extern "C" int nrf_render_frame(nrf_ctx* c, const float intr[4], const float pose[16], uint8_t** rgb, uint8_t** depth) {
  int rc = render_common(c, intr, pose, true, true);
  if (rc) return rc;
  if (rgb) *rgb = c->rgb_host;
  if (depth) *depth = c->depth_host;
  return NRF_OK;
}

extern "C" int nrf_render_frame_device(nrf_ctx* c, const float intr[4], const float pose[16]) {
  return render_common(c, intr, pose, false, false);
}

// Batch of poses (render_server --batch): a two-deep pipeline.  Frame i is enqueued (no host sync), its u8 pixels are
// copied to one of two pinned staging buffers on the same stream, and while the GPU renders frame i+1 the host copies
// frame i-1 out of the other staging buffer into the caller's memory.  A frame whose sample total exceeded the buffers
// (possible only while the buffers are still growing) is rendered again synchronously at the end.
extern "C" int nrf_render_batch_rgbd(nrf_ctx* c, const float intr[4], const float* poses, int n_poses, uint8_t* rgb_out,
                                     uint8_t* depth_out) {
  if (!c || !intr || !rgb_out || n_poses < 0 || (n_poses > 0 && !poses)) return fail(NRF_EINVAL, "nrf_render_batch: bad arguments");
  int n_local = 0;
  int rc = nrf_local_rays(c, &n_local);
  if (rc) return rc;
  if (!c->loaded) return fail(NRF_ESTATE, "render_batch before reload_network_from_file");
  if (c->peer_frame) return fail(NRF_ESTATE, "nrf_render_batch: detach the peer frame first (its pixels live in the owner's buffer)");
  const size_t px = c->n_shards == 1 ? (size_t)c->W * c->H : (size_t)n_local;
  const size_t frame_bytes = px * 3, stage_bytes = px * (depth_out ? 4 : 3);
  if (n_poses == 0) return NRF_OK;
  // the first frame synchronously: it sizes the sample buffers
  rc = render_common(c, intr, poses, true, true);
  if (rc) return rc;
  std::memcpy(rgb_out, c->rgb_host, frame_bytes);
  if (depth_out) std::memcpy(depth_out, c->depth_host, px);
  if (n_poses == 1) return NRF_OK;
  Device& D0 = c->devs[0];
  CU(cudaSetDevice(D0.dev));
  if (c->batch_bytes < stage_bytes) {
    for (int k = 0; k < NRF_BATCH_STAGES; ++k) {
      if (c->batch_stage[k]) cudaFreeHost(c->batch_stage[k]);
      c->batch_stage[k] = nullptr;
      CU(cudaMallocHost(&c->batch_stage[k], stage_bytes));
      if (!c->batch_ev[k]) CU(cudaEventCreateWithFlags(&c->batch_ev[k], cudaEventDisableTiming));
    }
    c->batch_bytes = stage_bytes;
  }
  if (!c->batch_copy_stream) CU(cudaStreamCreateWithFlags(&c->batch_copy_stream, cudaStreamNonBlocking));
  // Collector thread: waits for a frame's copy event and moves the frame from its pinned staging slot into the caller's
  // (pageable) memory, while this thread goes on enqueuing frames.  8 MB of host memcpy per 1080p frame is invisible next to
  // a 12 ms frame on one GPU, but it is half of a 1.7 ms frame on eight.
  struct Job { int j, slot; unsigned fno[8]; };
  struct Collector {
    std::mutex m;
    std::condition_variable cv_job, cv_done;
    std::deque<Job> jobs;
    bool closing = false;
    int collected = 0;            // jobs finished, in order
    cudaError_t err = cudaSuccess;
    std::vector<int> redo;
    std::thread th;
    ~Collector() {
      { std::lock_guard<std::mutex> l(m); closing = true; }
      cv_job.notify_all();
      if (th.joinable()) th.join();
    }
  } col;
  const size_t n_dev = std::min<size_t>(c->devs.size(), 8);
  col.th = std::thread([&]() {
    for (;;) {
      Job job;
      {
        std::unique_lock<std::mutex> l(col.m);
        col.cv_job.wait(l, [&] { return col.closing || !col.jobs.empty(); });
        if (col.jobs.empty()) return;
        job = col.jobs.front();
        col.jobs.pop_front();
      }
      const cudaError_t e = cudaEventSynchronize(c->batch_ev[job.slot]);
      bool over = false;
      for (size_t g = 0; g < n_dev; ++g) over |= c->devs[g].total_h[TOT_N * (job.fno[g] & 3u) + TOT_FRAME_OVF] != 0;
      if (e == cudaSuccess && !over) {
        std::memcpy(rgb_out + (size_t)job.j * frame_bytes, c->batch_stage[job.slot], frame_bytes);
        if (depth_out) std::memcpy(depth_out + (size_t)job.j * px, c->batch_stage[job.slot] + frame_bytes, px);
      }
      {
        std::lock_guard<std::mutex> l(col.m);
        if (e != cudaSuccess && col.err == cudaSuccess) col.err = e;
        if (over) col.redo.push_back(job.j);
        ++col.collected;
      }
      col.cv_done.notify_all();
    }
  });
  auto wait_collected = [&](int n) {   // frames 1..n are in the caller's memory (or marked for a repeat)
    std::unique_lock<std::mutex> l(col.m);
    col.cv_done.wait(l, [&] { return col.collected >= n; });
  };
  for (int i = 1; i < n_poses; ++i) {
    const int slot = i % NRF_BATCH_STAGES;
    if (i > NRF_BATCH_STAGES) wait_collected(i - NRF_BATCH_STAGES);   // frees this staging slot
    c->peer_slot = (int)(++c->peer_seq & 1u);
    // frame i-1's copy out of device 0's frame (its batch_ev) must finish before any device's composite stores frame i's
    // pixels there; frame 0 was rendered and copied synchronously
    c->pixel_wait = i >= 2 ? c->batch_ev[(i - 1) % NRF_BATCH_STAGES] : nullptr;
    rc = frame_enqueue_all(c, intr, poses + (size_t)i * 16, false);
    c->pixel_wait = nullptr;
    if (rc) return rc;
    Job job{i, slot, {}};
    for (size_t g = 0; g < n_dev; ++g) job.fno[g] = c->devs[g].frame_no;
    CU(cudaSetDevice(D0.dev));
    for (size_t g = 0; g < c->devs.size(); ++g) CU(cudaStreamWaitEvent(c->batch_copy_stream, c->devs[g].done, 0));
    CU(cudaMemcpyAsync(c->batch_stage[slot], D0.rgb_u8.p, frame_bytes, cudaMemcpyDeviceToHost, c->batch_copy_stream));
    if (depth_out) CU(cudaMemcpyAsync(c->batch_stage[slot] + frame_bytes, D0.depth_u8.p, px, cudaMemcpyDeviceToHost, c->batch_copy_stream));
    CU(cudaEventRecord(c->batch_ev[slot], c->batch_copy_stream));
    c->pending = true;
    c->pending_to_host = false;
    std::memcpy(c->pending_intr, intr, 16);
    std::memcpy(c->pending_pose, poses + (size_t)i * 16, 64);
    { std::lock_guard<std::mutex> l(col.m); col.jobs.push_back(job); }
    col.cv_job.notify_one();
  }
  wait_collected(n_poses - 1);
  if (col.err != cudaSuccess) return fail(NRF_ECUDA, "nrf_render_batch: frame copy failed: %s", cudaGetErrorString(col.err));
  const std::vector<int> redo = col.redo;
  rc = frame_resolve(c);   // validates the last frame, grows the buffers to the largest total seen
  if (rc) return rc;
  for (int j : redo) {
    rc = render_common(c, intr, poses + (size_t)j * 16, true, true);
    if (rc) return rc;
    std::memcpy(rgb_out + (size_t)j * frame_bytes, c->rgb_host, frame_bytes);
    if (depth_out) std::memcpy(depth_out + (size_t)j * px, c->depth_host, px);
  }
  return NRF_OK;
}

extern "C" int nrf_render_batch(nrf_ctx* c, const float intr[4], const float* poses, int n_poses, uint8_t* rgb_out) {
  return nrf_render_batch_rgbd(c, intr, poses, n_poses, rgb_out, nullptr);
}

extern "C" int nrf_get_float_frame(nrf_ctx* c, float* rgb_f32, float* depth_f32, int32_t* samples_per_ray) {
  if (!c || !c->W) return fail(NRF_ESTATE, "no frame rendered");
  { int rc = frame_resolve(c); if (rc) return rc; }
  // gathers in GLOBAL pixel order when the context covers the whole frame, else in local order
  const bool whole = c->n_shards == 1;
  size_t off = 0;
  for (Device& D : c->devs) {
    CU(cudaSetDevice(D.dev));
    CU(cudaStreamSynchronize(D.stream));
    const int n = D.n_local;
    std::vector<float> r((size_t)n * 3), d(n);
    std::vector<int> u(n);
    if (n) {
      CU(cudaMemcpy(r.data(), D.rgb_f32.p, (size_t)n * 12, cudaMemcpyDeviceToHost));
      CU(cudaMemcpy(d.data(), D.depth_f32.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
      CU(cudaMemcpy(u.data(), D.used.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
    }
    for (int q = 0; q < n; ++q) {
      // whole-frame context: global pixel order; sharded context: the packed tile order of the u8 outputs
      const long long pp = whole ? (long long)shard_pixel(D.map, q) : (long long)off + shard_packed_index(q);
      if (pp < 0) continue;  // ray of an edge tile outside the image
      const size_t p = (size_t)pp;
      if (rgb_f32) std::memcpy(rgb_f32 + p * 3, &r[(size_t)q * 3], 12);
      if (depth_f32) depth_f32[p] = d[q];
      if (samples_per_ray) samples_per_ray[p] = u[q];
    }
    off += n;
  }
  return NRF_OK;
}

extern "C" int nrf_get_stats(nrf_ctx* c, nrf_stats* out) {
  if (!c || !out) return fail(NRF_EINVAL, "null argument");
  if (!c->W) return fail(NRF_ESTATE, "no frame rendered");
  { int rc = frame_resolve(c); if (rc) return rc; }
  nrf_stats s{};
  s.n_rays = (uint64_t)c->W * c->H;
  s.n_gpus = (uint32_t)c->devs.size();
  for (Device& D : c->devs) {
    CU(cudaSetDevice(D.dev));
    CU(cudaStreamSynchronize(D.stream));
    s.n_samples += D.last_samples;
  }
  Device& D0 = c->devs[0];
  CU(cudaSetDevice(D0.dev));
  float ms;
  CU(cudaEventElapsedTime(&ms, D0.ev[0], D0.ev[1])); s.ms_setup = ms;   // rays + count + scan (march pass 1)
  CU(cudaEventElapsedTime(&ms, D0.ev[1], D0.ev[2])); s.ms_march = ms;   // emit (march pass 2)
  CU(cudaEventElapsedTime(&ms, D0.ev[2], D0.ev[3])); s.ms_network = ms;
  CU(cudaEventElapsedTime(&ms, D0.ev[3], D0.ev[4])); s.ms_composite = ms;
  for (int r = 1; r < D0.rounds_timed; ++r) {   // later rounds of a termination-aware frame: sums + scan + emit | network | composite
    CU(cudaEventElapsedTime(&ms, D0.rev[r][0], D0.rev[r][1])); s.ms_march += ms;
    CU(cudaEventElapsedTime(&ms, D0.rev[r][1], D0.rev[r][2])); s.ms_network += ms;
    CU(cudaEventElapsedTime(&ms, D0.rev[r][2], D0.rev[r][3])); s.ms_composite += ms;
  }
  CU(cudaEventElapsedTime(&ms, D0.ev[0], D0.ev[5])); s.ms_total = ms;
  s.n_rounds = (uint32_t)D0.rounds_timed;
  s.n_launches = (5 * s.n_rounds) * s.n_gpus;
  for (Device& D : c->devs) { s.n_samples_marched += D.last_marched; s.n_samples_used += D.last_used; }
  *out = s;
  return NRF_OK;
}

// Termination-aware evaluation rounds (nerf_b200_ext.h).  mode 0 = whole rays, 1 = rounds always, 2 = automatic.
// bounds (may be NULL = keep): cumulative samples-per-ray boundaries of all rounds but the last, ascending, < 1024.
extern "C" int nrf_set_termination(nrf_ctx* c, int mode, const int* bounds, int n_bounds) {
  if (!c || mode < 0 || mode > 2) return fail(NRF_EINVAL, "nrf_set_termination: bad arguments");
  if (bounds) {
    if (n_bounds < 1 || n_bounds > NRF_MAX_ROUNDS - 1) return fail(NRF_EINVAL, "nrf_set_termination: 1..%d round boundaries", NRF_MAX_ROUNDS - 1);
    for (int i = 0; i < n_bounds; ++i)
      if (bounds[i] < 1 || bounds[i] >= NRF_MAX_STEPS_D || (i > 0 && bounds[i] <= bounds[i - 1]))
        return fail(NRF_EINVAL, "nrf_set_termination: boundaries must be ascending in [1, %d)", NRF_MAX_STEPS_D);
  }
  { int rc = frame_resolve(c); if (rc) return rc; }
  if (bounds) {
    for (int i = 0; i < n_bounds; ++i) c->round_bounds[i] = bounds[i];
    c->round_bounds[n_bounds] = NRF_MAX_STEPS_D;
    c->n_round_bounds = n_bounds + 1;
  }
  c->term_mode = mode;
  c->chunked = mode == 1;
  return NRF_OK;
}

extern "C" int nrf_overflowed_frames(nrf_ctx* c, uint64_t* out) {
  if (!c || !out) return fail(NRF_EINVAL, "null argument");
  { int rc = frame_resolve(c); if (rc) return rc; }
  uint64_t n = 0;
  for (Device& D : c->devs) n += D.overflowed;
  *out = n;
  return NRF_OK;
}

// ---------------------------------------------------------------------------------------------------- stage probes
#define NEED_MODEL(c)                                                        \
  do {                                                                       \
    if (!(c)) return fail(NRF_EINVAL, "null ctx");                           \
    if (!(c)->loaded) return fail(NRF_ESTATE, "no model loaded");            \
  } while (0)

extern "C" int nrf_probe_rays(nrf_ctx* c, const float intr[4], const float pose[16], int W, int H, int pix0, int n,
                              float rays_o[3], float* rays_d, float* nears, float* fars) {
  NEED_MODEL(c);
  if (n < 0 || pix0 < 0 || (long long)pix0 + n > (long long)W * H) return fail(NRF_EINVAL, "pixel range out of frame");
  Device& D = c->devs[0];
  CU(cudaSetDevice(D.dev));
  MarchParams M = D.march;
  set_pose(c, M, intr, pose);
  M.W = W;
  for (int k = 0; k < 3; ++k) rays_o[k] = M.o[k];
  if (n == 0) return NRF_OK;
  DevBuf<float4> r; DevBuf<float> f;
  CU(r.ensure(n)); CU(f.ensure(n));
  k_rays_given_pixels<<<(n + 127) / 128, 128, 0, D.stream>>>(M, pix0, n, r.p, f.p);
  CU(cudaGetLastError());
  std::vector<float4> hr(n);
  CU(cudaMemcpyAsync(hr.data(), r.p, (size_t)n * 16, cudaMemcpyDeviceToHost, D.stream));
  CU(cudaMemcpyAsync(fars, f.p, (size_t)n * 4, cudaMemcpyDeviceToHost, D.stream));
  CU(cudaStreamSynchronize(D.stream));
  for (int i = 0; i < n; ++i) {
    rays_d[i * 3 + 0] = hr[i].x; rays_d[i * 3 + 1] = hr[i].y; rays_d[i * 3 + 2] = hr[i].z;
    nears[i] = hr[i].w;
  }
  r.release(); f.release();
  return NRF_OK;
}

extern "C" int nrf_probe_march(nrf_ctx* c, int n, const float rays_o[3], const float* rays_d, const float* nears,
                               const float* fars, int32_t* counts, float* samples, size_t max_samples) {
  NEED_MODEL(c);
  if (n <= 0) return n == 0 ? NRF_OK : fail(NRF_EINVAL, "n < 0");
  Device& D = c->devs[0];
  CU(cudaSetDevice(D.dev));
  MarchParams M = D.march;
  for (int k = 0; k < 3; ++k) M.o[k] = rays_o[k];
  std::vector<float4> hr(n);
  std::vector<float2> hf(n);
  for (int i = 0; i < n; ++i) {
    hr[i] = make_float4(rays_d[i * 3], rays_d[i * 3 + 1], rays_d[i * 3 + 2], nears[i]);
    hf[i] = make_float2(fars[i], nears[i]);
  }
  DevBuf<float4> r, smp; DevBuf<float2> f; DevBuf<float> ta, dbg; DevBuf<int> cnt, bs, wb, sray;
  DevBuf<unsigned long long> tot; DevBuf<long long> doff;
  const int nb = (n + 127) / 128;
  CU(r.ensure(n)); CU(f.ensure(n)); CU(cnt.ensure(n)); CU(bs.ensure(nb)); CU(wb.ensure((n + 31) / 32)); CU(tot.ensure(TOT_N));
  CU(cudaMemsetAsync(tot.p, 0, TOT_N * sizeof(unsigned long long), D.stream));
  CU(cudaMemcpyAsync(r.p, hr.data(), (size_t)n * 16, cudaMemcpyHostToDevice, D.stream));
  CU(cudaMemcpyAsync(f.p, hf.data(), (size_t)n * 8, cudaMemcpyHostToDevice, D.stream));
  k_count_given<<<nb, 128, 0, D.stream>>>(M, n, r.p, f.p, cnt.p, bs.p);
  CU(cudaGetLastError());
  k_scan_blocks<<<1, 1024, 0, D.stream>>>(bs.p, nb, tot.p, ~0ull);
  CU(cudaGetLastError());
  unsigned long long S = 0;
  CU(cudaMemcpyAsync(&S, tot.p, 8, cudaMemcpyDeviceToHost, D.stream));
  CU(cudaMemcpyAsync(counts, cnt.p, (size_t)n * 4, cudaMemcpyDeviceToHost, D.stream));
  CU(cudaStreamSynchronize(D.stream));
  int rc = NRF_OK;
  if (samples) {
    if (S > max_samples) rc = fail(NRF_EINVAL, "sample buffer too small: need %llu", S);
    else if (S > 0) {
      // the probe returns the samples RAY-major (like the oracle); the frame path's interleaved list is also written
      std::vector<long long> ho(n);
      long long acc = 0;
      for (int i = 0; i < n; ++i) { ho[i] = acc; acc += counts[i]; }
      CU(doff.ensure(n));
      CU(cudaMemcpyAsync(doff.p, ho.data(), (size_t)n * 8, cudaMemcpyHostToDevice, D.stream));
      CU(smp.ensure(S)); CU(ta.ensure(S)); CU(sray.ensure(S)); CU(dbg.ensure(S * 8));
      k_emit<<<nb, 128, 0, D.stream>>>(M, n, r.p, f.p, cnt.p, bs.p, wb.p, smp.p, ta.p, sray.p, dbg.p, doff.p, (long long)S, nullptr, nullptr,
                                      0, NRF_MAX_STEPS_D, nullptr, nullptr);
      CU(cudaGetLastError());
      CU(cudaMemcpyAsync(samples, dbg.p, (size_t)S * 32, cudaMemcpyDeviceToHost, D.stream));
      CU(cudaStreamSynchronize(D.stream));
    }
  }
  r.release(); smp.release(); f.release(); ta.release(); dbg.release(); cnt.release(); bs.release(); wb.release();
  sray.release(); tot.release(); doff.release();
  return rc;
}

static int upload_vec3_as_float4(Device& D, int n, const float* v3, DevBuf<float4>& out, float w = 0.f) {
  std::vector<float4> h(n);
  for (int i = 0; i < n; ++i) h[i] = make_float4(v3[i * 3], v3[i * 3 + 1], v3[i * 3 + 2], w);
  CU(out.ensure(n));
  CU(cudaMemcpyAsync(out.p, h.data(), (size_t)n * 16, cudaMemcpyHostToDevice, D.stream));
  CU(cudaStreamSynchronize(D.stream));
  return NRF_OK;
}

static void unpack_out(const std::vector<uint2>& h, float* out) {
  for (size_t i = 0; i < h.size(); ++i) {
    const uint32_t w[2] = {h[i].x, h[i].y};
    for (int k = 0; k < 4; ++k) {
      const uint16_t b = (uint16_t)(w[k >> 1] >> ((k & 1) * 16));
      __half hh;
      std::memcpy(&hh, &b, 2);
      out[i * 4 + k] = __half2float(hh);
    }
  }
}

extern "C" int nrf_probe_encode(nrf_ctx* c, int n, const float* positions, uint16_t* enc_out, uint32_t* indices_out) {
  NEED_MODEL(c);
  if (n <= 0) return n == 0 ? NRF_OK : fail(NRF_EINVAL, "n < 0");
  Device& D = c->devs[0];
  CU(cudaSetDevice(D.dev));
  DevBuf<float4> pos;
  int rc = upload_vec3_as_float4(D, n, positions, pos);
  if (rc) return rc;
  CU(D.enc.ensure((size_t)n * 4));
  if (indices_out) CU(D.idx.ensure((size_t)n * NRF_L * 8));
  CU(nrf_launch_network(0, network_grid(c, D, n), D.stream, &D.net, n, pos.p, nullptr, nullptr, nullptr, 1, D.enc.p,
                        nullptr, indices_out ? D.idx.p : nullptr, nullptr, nullptr));
  CU(cudaMemcpyAsync(enc_out, D.enc.p, (size_t)n * 64, cudaMemcpyDeviceToHost, D.stream));
  if (indices_out) CU(cudaMemcpyAsync(indices_out, D.idx.p, (size_t)n * NRF_L * 8 * 4, cudaMemcpyDeviceToHost, D.stream));
  CU(cudaStreamSynchronize(D.stream));
  pos.release();
  return NRF_OK;
}

extern "C" int nrf_probe_sh(nrf_ctx* c, int n, const float* dirs, uint16_t* sh_out) {
  if (!c || n < 0 || (n > 0 && (!dirs || !sh_out))) return fail(NRF_EINVAL, "nrf_probe_sh: bad arguments");
  if (n == 0) return NRF_OK;
  Device& D = c->devs[0];
  CU(cudaSetDevice(D.dev));
  DevBuf<float4> d4; DevBuf<uint4> o;
  int rc = upload_vec3_as_float4(D, n, dirs, d4);
  if (rc) return rc;
  CU(o.ensure(2 * (size_t)n));
  k_probe_sh<<<(n + 127) / 128, 128, 0, D.stream>>>(n, d4.p, o.p);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(sh_out, o.p, (size_t)n * 32, cudaMemcpyDeviceToHost, D.stream));
  CU(cudaStreamSynchronize(D.stream));
  d4.release(); o.release();
  return NRF_OK;
}

extern "C" int nrf_probe_network(nrf_ctx* c, int n, const uint16_t* enc, const float* dirs, float* out, int impl) {
  NEED_MODEL(c);
  if (n <= 0) return n == 0 ? NRF_OK : fail(NRF_EINVAL, "n < 0");
  Device& D = c->devs[0];
  CU(cudaSetDevice(D.dev));
  DevBuf<float4> d4; DevBuf<uint2> o;
  int rc = upload_vec3_as_float4(D, n, dirs, d4);
  if (rc) return rc;
  CU(D.enc.ensure((size_t)n * 4)); CU(o.ensure(n));
  CU(cudaMemcpyAsync(D.enc.p, enc, (size_t)n * 64, cudaMemcpyHostToDevice, D.stream));
  if (impl == 1) CU(nrf_launch_network_simt(D.stream, &D.net, n, D.enc.p, d4.p, o.p));
  else CU(nrf_launch_network(1, network_grid(c, D, n), D.stream, &D.net, n, nullptr, nullptr, d4.p, nullptr, 1, D.enc.p, o.p, nullptr, nullptr, nullptr));
  std::vector<uint2> h(n);
  CU(cudaMemcpyAsync(h.data(), o.p, (size_t)n * 8, cudaMemcpyDeviceToHost, D.stream));
  CU(cudaStreamSynchronize(D.stream));
  unpack_out(h, out);
  d4.release(); o.release();
  return NRF_OK;
}

extern "C" int nrf_probe_samples(nrf_ctx* c, int n, const float* positions, const float* dirs, float* out) {
  NEED_MODEL(c);
  if (n <= 0) return n == 0 ? NRF_OK : fail(NRF_EINVAL, "n < 0");
  Device& D = c->devs[0];
  CU(cudaSetDevice(D.dev));
  DevBuf<float4> p4, d4; DevBuf<uint2> o;
  int rc = upload_vec3_as_float4(D, n, positions, p4);
  if (!rc) rc = upload_vec3_as_float4(D, n, dirs, d4);
  if (rc) return rc;
  CU(o.ensure(n));
  CU(nrf_launch_network(2, network_grid(c, D, n), D.stream, &D.net, n, p4.p, nullptr, d4.p, nullptr, 1, nullptr, o.p, nullptr, nullptr, nullptr));
  std::vector<uint2> h(n);
  CU(cudaMemcpyAsync(h.data(), o.p, (size_t)n * 8, cudaMemcpyDeviceToHost, D.stream));
  CU(cudaStreamSynchronize(D.stream));
  unpack_out(h, out);
  p4.release(); d4.release(); o.release();
  return NRF_OK;
}

// ---------------------------------------------------------------------------------------------------- no-allocation entry point
// nrf_eval_samples_noalloc (include/nerf_b200_ext.h): the two-kernel form of the sample evaluation — k_network<0> writes the
// fp16 hash-grid features of every sample (64 B) to scratch, k_network<1> runs the MLP from that scratch — with the scratch in
// a file-scope device array instead of a cudaMalloc'ed buffer, for callers that must not allocate (CUDA graph capture, tight
// serving loops).  Every GPU that loads the library holds the array (4 MB) until exit.  In host code the array's NAME is a
// host-side stand-in, not a device address: its device address is taken once per device with cudaGetSymbolAddress (with that
// device current) and that pointer is what the kernels get.
static __device__ __align__(256) uint4 g_noalloc_enc[(size_t)NRF_NOALLOC_MAX_SAMPLES * 4];
static uint4* g_noalloc_ptr[64] = {};   // device address of g_noalloc_enc per CUDA device ordinal
static std::mutex g_noalloc_mutex;

static int noalloc_scratch(int dev, uint4** out) {   // call with device `dev` current
  if (dev < 0 || dev >= 64) return fail(NRF_EINVAL, "nrf_eval_samples_noalloc: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> l(g_noalloc_mutex);
  if (!g_noalloc_ptr[dev]) {
    void* p = nullptr;
    CU(cudaGetSymbolAddress(&p, g_noalloc_enc));
    if (!p) return fail(NRF_ECUDA, "cudaGetSymbolAddress returned a null address for the no-allocation scratch");
    g_noalloc_ptr[dev] = static_cast<uint4*>(p);
  }
  *out = g_noalloc_ptr[dev];
  return NRF_OK;
}

extern "C" int nrf_eval_samples_noalloc(nrf_ctx* c, int n, const void* positions_dev, const void* dirs_dev, void* out_dev) {
  NEED_MODEL(c);
  if (n < 0) return fail(NRF_EINVAL, "nrf_eval_samples_noalloc: n < 0");
  if (n > NRF_NOALLOC_MAX_SAMPLES)
    return fail(NRF_EINVAL, "nrf_eval_samples_noalloc: %d samples exceed the fixed scratch of %d (NRF_NOALLOC_MAX_SAMPLES); split the call",
                n, NRF_NOALLOC_MAX_SAMPLES);
  if (n == 0) return NRF_OK;
  if (!positions_dev || !dirs_dev || !out_dev) return fail(NRF_EINVAL, "nrf_eval_samples_noalloc: null pointer");
  if (((uintptr_t)positions_dev & 15) || ((uintptr_t)dirs_dev & 15) || ((uintptr_t)out_dev & 7))
    return fail(NRF_EINVAL, "nrf_eval_samples_noalloc: positions / dirs must be 16-byte aligned, out 8-byte aligned");
  Device& D = c->devs[0];
  CU(cudaSetDevice(D.dev));
  uint4* scratch = nullptr;
  { int rc = noalloc_scratch(D.dev, &scratch); if (rc) return rc; }
  const int grid = network_grid(c, D, n);
  // stage 1 writes entries [0, n) of the scratch in full and stage 2 reads only those: nothing has to start at zero, and what
  // an earlier call left beyond n is never read
  CU(nrf_launch_network(0, grid, D.stream, &D.net, n, static_cast<const float4*>(positions_dev), nullptr, nullptr, nullptr, 1,
                        scratch, nullptr, nullptr, nullptr, nullptr));
  CU(nrf_launch_network(1, grid, D.stream, &D.net, n, nullptr, nullptr, static_cast<const float4*>(dirs_dev), nullptr, 1,
                        scratch, static_cast<uint2*>(out_dev), nullptr, nullptr, nullptr));
  return NRF_OK;
}

// ---------------------------------------------------------------------------------------------------- stage bench
// Device-resident stage timing on the samples of the LAST rendered frame (so the access pattern is the real one):
// what 0 = hash-grid gather only, 1 = MLP only (reads the encoded features), 2 = fused.  CUDA events on the launching
// stream; between iterations nothing else touches L2 than the stage itself (the sample stream is >> L2 at full size).
extern "C" int nrf_bench_stage(nrf_ctx* c, int what, int n, int iters, float* ms_out) {
  NEED_MODEL(c);
  { int rc_ = frame_resolve(c); if (rc_) return rc_; }
  if (!ms_out || iters < 1 || what < 0 || what > 2) return fail(NRF_EINVAL, "nrf_bench_stage: bad arguments");
  Device& D = c->devs[0];
  CU(cudaSetDevice(D.dev));
  const int S = (int)std::min<unsigned long long>(D.last_samples, n > 0 ? (unsigned long long)n : ~0ull);
  if (S <= 0) return fail(NRF_ESTATE, "render a frame first (the stage bench replays its samples)");
  CU(D.enc.ensure((size_t)S * 4));
  const int grid = network_grid(c, D, S);
  const SampleSrc src = sample_src(D);
  if (what == 1)  // produce the features once
    CU(nrf_launch_network(0, grid, D.stream, &D.net, S, D.samples.p, nullptr, D.rays.p, D.ray_sh.p, 0, D.enc.p, nullptr, nullptr, nullptr, &src));
  cudaEvent_t e0, e1;
  CU(cudaEventCreate(&e0)); CU(cudaEventCreate(&e1));
  CU(nrf_launch_network(what, grid, D.stream, &D.net, S, D.samples.p, nullptr, D.rays.p, D.ray_sh.p, 0, D.enc.p, D.sample_out.p, nullptr, nullptr, &src));
  CU(cudaEventRecord(e0, D.stream));
  for (int i = 0; i < iters; ++i)
    CU(nrf_launch_network(what, grid, D.stream, &D.net, S, D.samples.p, nullptr, D.rays.p, D.ray_sh.p, 0, D.enc.p, D.sample_out.p, nullptr, nullptr, &src));
  CU(cudaEventRecord(e1, D.stream));
  CU(cudaEventSynchronize(e1));
  float ms = 0;
  CU(cudaEventElapsedTime(&ms, e0, e1));
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  *ms_out = ms / iters;
  return NRF_OK;
}

// Experiment hook (not part of the public header): time the gather stage on caller-supplied positions.
extern "C" int nrf_bench_encode_positions(nrf_ctx* c, int n, const float* positions, int iters, float* ms_out) {
  NEED_MODEL(c);
  Device& D = c->devs[0];
  CU(cudaSetDevice(D.dev));
  DevBuf<float4> pos;
  int rc = upload_vec3_as_float4(D, n, positions, pos);
  if (rc) return rc;
  CU(D.enc.ensure((size_t)n * 4));
  const int grid = network_grid(c, D, n);
  cudaEvent_t e0, e1;
  CU(cudaEventCreate(&e0)); CU(cudaEventCreate(&e1));
  CU(nrf_launch_network(0, grid, D.stream, &D.net, n, pos.p, nullptr, nullptr, nullptr, 1, D.enc.p, nullptr, nullptr, nullptr, nullptr));
  CU(cudaEventRecord(e0, D.stream));
  for (int i = 0; i < iters; ++i)
    CU(nrf_launch_network(0, grid, D.stream, &D.net, n, pos.p, nullptr, nullptr, nullptr, 1, D.enc.p, nullptr, nullptr, nullptr, nullptr));
  CU(cudaEventRecord(e1, D.stream));
  CU(cudaEventSynchronize(e1));
  float ms = 0;
  CU(cudaEventElapsedTime(&ms, e0, e1));
  *ms_out = ms / iters;
  pos.release();
  return NRF_OK;
}

// Experiment hook (not in the public header): encode-only occupancy variants on the last frame's samples.
extern "C" int nrf_exp_encode(nrf_ctx* c, int variant, int iters, float* ms_out) {
  NEED_MODEL(c);
  { int rc_ = frame_resolve(c); if (rc_) return rc_; }
  Device& D = c->devs[0];
  CU(cudaSetDevice(D.dev));
  const int S = (int)D.last_samples;
  if (S <= 0) return fail(NRF_ESTATE, "render a frame first");
  CU(D.enc.ensure((size_t)S * 4));
  cudaEvent_t e0, e1;
  CU(cudaEventCreate(&e0)); CU(cudaEventCreate(&e1));
  CU(nrf_launch_encode_exp(variant, D.n_sms, D.stream, &D.net, S, D.samples.p, D.enc.p));
  CU(cudaEventRecord(e0, D.stream));
  for (int i = 0; i < iters; ++i) CU(nrf_launch_encode_exp(variant, D.n_sms, D.stream, &D.net, S, D.samples.p, D.enc.p));
  CU(cudaEventRecord(e1, D.stream));
  CU(cudaEventSynchronize(e1));
  float ms = 0;
  CU(cudaEventElapsedTime(&ms, e0, e1));
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  *ms_out = ms / iters;
  return NRF_OK;
}

// ---------------------------------------------------------------------------------------------------- memory probes
// Roofline denominators measured on the box (kernels_probe.cu).  mode 0: streaming 16-byte reads over `window_bytes`
// split over the CTAs (>> L2 => HBM); mode 2: every CTA streams the whole window (<= L2 => L2->SM bandwidth); mode 1: one random 8-byte slot (= one 32-byte sector, own 128-byte line) per
// lane.  *gbs = bytes REQUESTED by the threads per second / 1e9 (mode 1: 8 B per load); *gloads = loads per ns.
extern "C" int nrf_bench_memory(nrf_ctx* c, int mode, size_t window_bytes, int iters, float* gbs, float* gloads) {
  if (!c || !gbs || mode < 0 || mode > 2 || window_bytes < (1u << 20) || iters < 1)
    return fail(NRF_EINVAL, "nrf_bench_memory: bad arguments");
  Device& D = c->devs[0];
  CU(cudaSetDevice(D.dev));
  void* buf = nullptr;
  uint32_t* sink = nullptr;
  CU(cudaMalloc(&buf, window_bytes));
  CU(cudaMalloc(&sink, 4));
  CU(cudaMemsetAsync(buf, 0, window_bytes, D.stream));
  // work per launch: mode 0 = passes over the window (>= ~4 GB read in total), mode 1 = loads per thread
  const int work = mode == 0 ? (int)std::max<size_t>(1, (4ull << 30) / window_bytes)
                   : mode == 2 ? (int)std::max<size_t>(1, (64ull << 20) / window_bytes) : 4096;   // mode 2: every CTA reads the window `work` times
  cudaEvent_t e0, e1;
  CU(cudaEventCreate(&e0)); CU(cudaEventCreate(&e1));
  CU(nrf_launch_probe(mode, D.n_sms, D.stream, buf, window_bytes, work, sink));  // warm-up (fills L2)
  CU(cudaEventRecord(e0, D.stream));
  for (int i = 0; i < iters; ++i) CU(nrf_launch_probe(mode, D.n_sms, D.stream, buf, window_bytes, work, sink));
  CU(cudaEventRecord(e1, D.stream));
  CU(cudaEventSynchronize(e1));
  float ms = 0;
  CU(cudaEventElapsedTime(&ms, e0, e1));
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaFree(buf); cudaFree(sink);
  const double s = ms * 1e-3 / iters;
  const double loads = mode == 0 ? (double)(window_bytes / 16) * work
                       : mode == 2 ? (double)(window_bytes / 16) * work * D.n_sms : (double)D.n_sms * 1024.0 * work;
  const double bytes = loads * (mode == 1 ? 8.0 : 16.0);
  *gbs = (float)(bytes / s / 1e9);
  if (gloads) *gloads = (float)(loads / s / 1e9);
  return NRF_OK;
}

// ---------------------------------------------------------------------------------------------------- density grid
// NerfRender::generate_density_grid (nerf_render.cu:388-429; dead code in the reference: no caller, the density call
// is commented out at :415 and only the last cascade's values would reach dg_update).  Built here as the code intends:
// per cascade, evaluate the density network at the (optionally jittered) centre of every cell, scale by 0.001691,
// EMA-max into the float grid (dg_update), then refresh mean_density and the occupancy bitfield the march reads.
// Decisions: positions are normalised to [0,1] like every other network input; density = sigma after the
// configured activation; jitter uses pcg32(seed) advanced to the element index (3*(cascade*H^3 + cell) + axis);
// reset != 0 starts from the reference's 1/64 fill (:395), else from the resident grid.
extern "C" int nrf_generate_density_grid(nrf_ctx* c, float decay, int reset, unsigned long long jitter_seed, int jitter,
                                         float* mean_density_out) {
  NEED_MODEL(c);
  { int rc_ = frame_resolve(c); if (rc_) return rc_; }
  const nrf_model_desc& d = c->desc;
  const uint32_t H = d.grid_size, H3 = H * H * H;
  float mean = 0.f;
  for (size_t g = 0; g < c->devs.size(); ++g) {
    Device& D = c->devs[g];
    CU(cudaSetDevice(D.dev));
    CU(D.samples.ensure(H3)); CU(D.sample_out.ensure(H3)); CU(D.sample_ray.ensure(H3)); CU(D.t_after.ensure(H3));
    CU(D.ray_sh.ensure(2));
    DevBuf<double> sum;
    CU(sum.ensure(1));
    CU(cudaMemsetAsync(sum.p, 0, sizeof(double), D.stream));
    CU(cudaMemsetAsync(D.sample_ray.p, 0, (size_t)H3 * sizeof(int), D.stream));   // every sample "belongs" to ray 0
    CU(cudaMemsetAsync(D.ray_sh.p, 0, 2 * sizeof(uint4), D.stream));              // direction features: irrelevant to sigma
    for (uint32_t cas = 0; cas < d.cascades; ++cas) {
      const float bound_c = (float)(1u << cas) < d.bound ? (float)(1u << cas) : d.bound;  // nerf_render.cu:408
      const float half = bound_c / H;                                                      // :409
      k_dg_positions<<<(H3 + 255) / 256, 256, 0, D.stream>>>(H, bound_c, half, D.march.pos_w, jitter_seed, jitter,
                                                           3ull * cas * H3, D.samples.p);
      CU(cudaGetLastError());
      CU(nrf_launch_network(2, network_grid(c, D, (int)H3), D.stream, &D.net, (int)H3, D.samples.p, D.sample_ray.p,
                            D.rays.p, D.ray_sh.p, 0, nullptr, D.sample_out.p, nullptr, nullptr, nullptr));
      k_dg_update<<<(H3 + 255) / 256, 256, 0, D.stream>>>(H3, D.sample_out.p, decay, reset,
                                                        D.density_grid.p + (size_t)cas * H3, sum.p);
      CU(cudaGetLastError());
    }
    double hsum = 0;
    CU(cudaMemcpyAsync(&hsum, sum.p, sizeof(double), cudaMemcpyDeviceToHost, D.stream));
    CU(cudaStreamSynchronize(D.stream));
    mean = (float)(hsum / ((double)H3 * d.cascades));
    k_build_occupancy<<<(unsigned)(((size_t)H3 * d.cascades + 255) / 256), 256, 0, D.stream>>>(
        D.density_grid.p, (uint32_t)((size_t)H3 * d.cascades), fminf(0.01f, mean), D.occupancy.p);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(D.stream));
    sum.release();
    D.last_samples = 0;  // the per-frame sample buffers were used as scratch
  }
  c->desc.mean_density = mean;
  if (mean_density_out) *mean_density_out = mean;
  return NRF_OK;
}

// Resident float density grid [cascades*H^3] and occupancy bitfield [cascades*H^3/32] of device 0 (either may be NULL).
extern "C" int nrf_get_density_grid(nrf_ctx* c, float* grid_out, uint32_t* bits_out) {
  NEED_MODEL(c);
  Device& D = c->devs[0];
  CU(cudaSetDevice(D.dev));
  const size_t n = (size_t)c->desc.grid_size * c->desc.grid_size * c->desc.grid_size * c->desc.cascades;
  if (grid_out) CU(cudaMemcpyAsync(grid_out, D.density_grid.p, n * sizeof(float), cudaMemcpyDeviceToHost, D.stream));
  if (bits_out) CU(cudaMemcpyAsync(bits_out, D.occupancy.p, n / 32 * sizeof(uint32_t), cudaMemcpyDeviceToHost, D.stream));
  CU(cudaStreamSynchronize(D.stream));
  return NRF_OK;
}

// ---------------------------------------------------------------------------------------------------- post-process
// CudaRenderBuffer::accumulate + ::tonemap (src/render_buffer.cu:598-640 -> accumulate_kernel :224, tonemap_kernel :529)
// on linear RGBA float buffers.  Host pointers in / out (n_pixels * 4 floats each); accum is read and updated.
// frame may be NULL (tonemap only); out may be NULL (accumulate only).
extern "C" int nrf_postprocess(nrf_ctx* c, int n_pixels, const float* frame_rgba, float* accum_rgba, float sample_count,
                               int color_space, float exposure, const float background_rgba[4], int output_color_space,
                               int tonemap_curve, int clamp_output, float* out_rgba) {
  if (!c || n_pixels < 0 || !accum_rgba) return fail(NRF_EINVAL, "nrf_postprocess: bad arguments");
  if (color_space < 0 || color_space > 2 || output_color_space < 0 || output_color_space > 2 || tonemap_curve < 0 || tonemap_curve > 3)
    return fail(NRF_EINVAL, "nrf_postprocess: bad colour space / tonemap curve");
  if (out_rgba && !background_rgba) return fail(NRF_EINVAL, "nrf_postprocess: background required for tonemapping");
  if (n_pixels == 0) return NRF_OK;
  Device& D = c->devs[0];
  CU(cudaSetDevice(D.dev));
  DevBuf<float4> f, a, o;
  const size_t bytes = (size_t)n_pixels * 16;
  CU(a.ensure(n_pixels));
  CU(cudaMemcpyAsync(a.p, accum_rgba, bytes, cudaMemcpyHostToDevice, D.stream));
  const int nb = (n_pixels + 255) / 256;
  if (frame_rgba) {
    CU(f.ensure(n_pixels));
    CU(cudaMemcpyAsync(f.p, frame_rgba, bytes, cudaMemcpyHostToDevice, D.stream));
    k_accumulate<<<nb, 256, 0, D.stream>>>(n_pixels, f.p, a.p, sample_count, color_space);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(accum_rgba, a.p, bytes, cudaMemcpyDeviceToHost, D.stream));
  }
  if (out_rgba) {
    CU(o.ensure(n_pixels));
    const float4 bg = make_float4(background_rgba[0], background_rgba[1], background_rgba[2], background_rgba[3]);
    k_tonemap<<<nb, 256, 0, D.stream>>>(n_pixels, a.p, exposure, bg, color_space, output_color_space, tonemap_curve,
                                       clamp_output, o.p);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(out_rgba, o.p, bytes, cudaMemcpyDeviceToHost, D.stream));
  }
  CU(cudaStreamSynchronize(D.stream));
  f.release(); a.release(); o.release();
  return NRF_OK;
}

// ---------------------------------------------------------------------------------------------------- peer frame (IPC)
// One process per GPU: instead of gathering packed tiles with a collective after the frame, every rank's composite
// kernel stores its u8 pixels straight into ONE frame buffer that lives on the destination rank's GPU (NVLink peer
// stores through a CUDA-IPC mapping) — the compute step and the gather are the same kernel, exactly like the
// one-process n-GPU mode.  The caller orders "all ranks finished" -> "destination reads" with one stream-ordered
// collective of 4 bytes (bench.py) or a barrier.
static int peer_detach(nrf_ctx* c) {
  if (!c->peer_frame) return NRF_OK;
  CU(cudaSetDevice(c->devs[0].dev));
  CU(cudaStreamSynchronize(c->devs[0].stream));
  if (c->peer_copy_stream) {
    CU(cudaStreamSynchronize(c->peer_copy_stream));
    CU(cudaStreamDestroy(c->peer_copy_stream));
    c->peer_copy_stream = nullptr;
  }
  for (int k = 0; k < 2; ++k) {
    if (c->peer_copied[k]) CU(cudaEventDestroy(c->peer_copied[k]));
    if (c->peer_done[k]) CU(cudaEventDestroy(c->peer_done[k]));
    c->peer_copied[k] = c->peer_done[k] = nullptr;
    c->peer_copied_valid[k] = false;
  }
  if (c->peer_mapped) CU(cudaIpcCloseMemHandle(c->peer_frame));
  else if (c->peer_owned) CU(cudaFree(c->peer_frame));
  c->peer_frame = nullptr;
  c->peer_owned = c->peer_mapped = false;
  c->peer_copy_out = false;
  c->peer_seq = 0;
  c->peer_slot = 0;
  return NRF_OK;
}

extern "C" int nrf_peer_frame_export(nrf_ctx* c, void* handle64) {
  if (!c || !handle64) return fail(NRF_EINVAL, "nrf_peer_frame_export: null argument");
  { int rc_ = frame_resolve(c); if (rc_) return rc_; }
  if (!c->W) return fail(NRF_ESTATE, "set_resolution not called");
  if (c->devs.size() != 1) return fail(NRF_EINVAL, "peer frame needs a single-device context (one process per GPU)");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  int rc = peer_detach(c);
  if (rc) return rc;
  Device& D = c->devs[0];
  CU(cudaSetDevice(D.dev));
  // TWO frames, written alternately: the destination can still be copying frame i out while the other ranks store
  // frame i+1 — the completion fence of frame i+1 then orders "copy of i finished" before anyone stores frame i+2
  if (c->n_shards > NRF_PEER_MAX_RANKS) return fail(NRF_EINVAL, "peer frame supports up to %d ranks", NRF_PEER_MAX_RANKS);
  const size_t bytes = peer_alloc_bytes(c);   // two frames + the fence flags
  c->peer_slot = 0;
  c->peer_seq = 0;
  CU(cudaMalloc(&c->peer_frame, bytes));
  c->peer_owned = true;
  CU(cudaMemsetAsync(c->peer_frame, 0, bytes, D.stream));
  CU(cudaStreamSynchronize(D.stream));
  if (!c->peer_host) CU(cudaMallocHost(&c->peer_host, (size_t)c->W * c->H * 4 * 2));
  CU(cudaStreamCreateWithFlags(&c->peer_copy_stream, cudaStreamNonBlocking));
  for (int k = 0; k < 2; ++k) {
    CU(cudaEventCreateWithFlags(&c->peer_copied[k], cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&c->peer_done[k], cudaEventDisableTiming));
  }
  cudaIpcMemHandle_t h;
  CU(cudaIpcGetMemHandle(&h, c->peer_frame));
  std::memcpy(handle64, &h, 64);
  return NRF_OK;
}

extern "C" int nrf_peer_frame_attach(nrf_ctx* c, const void* handle64) {
  if (!c || !handle64) return fail(NRF_EINVAL, "nrf_peer_frame_attach: null argument");
  { int rc_ = frame_resolve(c); if (rc_) return rc_; }
  if (!c->W) return fail(NRF_ESTATE, "set_resolution not called");
  if (c->devs.size() != 1) return fail(NRF_EINVAL, "peer frame needs a single-device context (one process per GPU)");
  int rc = peer_detach(c);
  if (rc) return rc;
  CU(cudaSetDevice(c->devs[0].dev));
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle64, 64);
  void* p = nullptr;
  CU(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  c->peer_frame = (uint8_t*)p;
  c->peer_mapped = true;
  c->peer_slot = 0;
  c->peer_seq = 0;
  for (int k = 0; k < 2; ++k) CU(cudaEventCreateWithFlags(&c->peer_done[k], cudaEventDisableTiming));
  return NRF_OK;
}

extern "C" int nrf_peer_frame_detach(nrf_ctx* c) {
  if (!c) return fail(NRF_EINVAL, "null ctx");
  { int rc_ = frame_resolve(c); if (rc_) return rc_; }
  return peer_detach(c);
}

// Destination rank: device pointer of the frame (rgb W*H*3 | depth W*H), and an async copy of it to pinned host memory
// on the context's stream (call after the completion fence; synchronise the stream before reading *rgb / *depth).
extern "C" int nrf_peer_frame_device(nrf_ctx* c, void** frame) {
  if (!c || !frame) return fail(NRF_EINVAL, "null argument");
  if (!c->peer_frame) return fail(NRF_ESTATE, "no peer frame");
  *frame = peer_slot_ptr(c);
  return NRF_OK;
}
// Owner: the LATEST frame on the host.  Waits for the one-sided fence of that frame (every rank's pixels have landed),
// copies the slot out unless nrf_peer_frame_copy_out already did so behind the fence, and returns when the bytes are there.
extern "C" int nrf_peer_frame_to_host(nrf_ctx* c, uint8_t** rgb, uint8_t** depth) {
  return nrf_peer_frame_collect(c, 0, rgb, depth);
}

// Owner: switch the automatic copy-out on or off: every later frame is copied to its own pinned host slot on the copy
// stream as soon as the fence of that frame has passed, while the next frame is already rendering.
extern "C" int nrf_peer_frame_copy_out(nrf_ctx* c, int on) {
  if (!c) return fail(NRF_EINVAL, "null ctx");
  if (!c->peer_frame || !c->peer_owned) return fail(NRF_ESTATE, "not the owner of a peer frame");
  c->peer_copy_out = on != 0;
  return NRF_OK;
}

// Frame `back` frames before the latest (0 or 1; the buffer holds two).  Owner: waits until that frame's fence has passed
// and it is on the host, returns pointers into its pinned slot (valid until two more frames have been rendered).  Other
// ranks: waits until this rank's own kernels of that frame have finished (rgb/depth are set to NULL) — what keeps a rank's
// host loop from running more than a frame ahead of its GPU.
extern "C" int nrf_peer_frame_collect(nrf_ctx* c, int back, uint8_t** rgb, uint8_t** depth) {
  if (!c || back < 0 || back > 1) return fail(NRF_EINVAL, "nrf_peer_frame_collect: bad arguments");
  if (!c->peer_frame) return fail(NRF_ESTATE, "no peer frame");
  Device& D = c->devs[0];
  CU(cudaSetDevice(D.dev));
  if (rgb) *rgb = nullptr;
  if (depth) *depth = nullptr;
  if (c->peer_seq <= (unsigned)back) return fail(NRF_ESTATE, "nrf_peer_frame_collect: no such frame yet");
  const int slot = (int)((c->peer_seq - (unsigned)back) & 1u);
  if (!c->peer_owned) { CU(cudaEventSynchronize(c->peer_done[slot])); return NRF_OK; }
  uint8_t* host = c->peer_host + (size_t)slot * c->W * c->H * 4;
  // the latest frame must be a valid one first: a frame of this rank that did not fit its sample buffers is rendered again
  // here (same slot, same sequence number; its fence and copy are enqueued again behind it)
  if (back == 0) { int rc_ = frame_resolve(c); if (rc_) return rc_; }
  if (!c->peer_copy_out) {   // not copied behind the fence: do it now, after the slot's fence, on the copy stream
    if (back != 0) return fail(NRF_ESTATE, "nrf_peer_frame_collect: an earlier frame needs nrf_peer_frame_copy_out");
    CU(cudaMemcpyAsync(host, c->peer_frame + (size_t)slot * c->W * c->H * 4, (size_t)c->W * c->H * 4, cudaMemcpyDeviceToHost,
                       c->peer_copy_stream));
    CU(cudaStreamSynchronize(c->peer_copy_stream));
  } else {
    CU(cudaEventSynchronize(c->peer_copied[slot]));
  }
  if (rgb) *rgb = host;
  if (depth) *depth = host + (size_t)c->W * c->H * 3;
  return NRF_OK;
}

// Makes device 0's stream wait for everything the latest frame still has in flight elsewhere — the other devices' kernels
// (one process, several devices) and the owner's fence of the peer frame (one process per GPU; a no-op on the other
// ranks) — so that an event recorded on that stream afterwards covers the whole frame: what bench.py brackets.
extern "C" int nrf_frame_join(nrf_ctx* c) {
  if (!c) return fail(NRF_EINVAL, "null ctx");
  Device& D = c->devs[0];
  CU(cudaSetDevice(D.dev));
  for (size_t g = 1; g < c->devs.size(); ++g) CU(cudaStreamWaitEvent(D.stream, c->devs[g].done, 0));
  if (c->peer_frame && c->peer_owned && c->peer_seq) CU(cudaStreamWaitEvent(D.stream, c->peer_copied[c->peer_seq & 1u], 0));
  return NRF_OK;
}
