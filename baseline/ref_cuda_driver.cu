// ref_cuda_driver.cu — drives the REFERENCE's own CUDA renderer on a B200.  Baseline / parity tooling only.
//
// The reference's src/nerf_render.cu is compiled INTO this translation unit (NRF_RENDER_SRC = the file where it lies
// under /root/reference, or the sed-generated D1-clean variant, see build_ref_cuda.sh), so the driver can call
// ngp::NerfRender exactly as src/render_server.cu:44-101 / src/main.cu:147-161 do and additionally reach the private
// network object for stage dumps.  Nothing here is linked into libnerf_b200.so.
//
//   ref_cuda_driver <snapshot.msgpack> <out_dir> [frames W H fl_x fl_y cx cy poses.bin n_poses reps]
//                                                [net in.bin B] [march W H fl_x fl_y cx cy pose.bin max_step]
//                                                [bench W H fl_x fl_y cx cy poses.bin n_poses warmup steps]
//
//   frames : render_frame (nerf_render.cu:238-367) per pose; float image/depth (image_h/deep_h) and the u8 frame are
//            written for the first rep, wall-clock ms of every rep printed as JSON lines.
//   net    : NerfNetwork<__half>::inference_mixed_precision_impl (nerf_network.h:148-196) on B samples
//            (in.bin = [6 x B] f32 SoA: rows 0-2 position in [0,1], rows 3-5 direction*0.5+0.5); dumps the full
//            [16 x B] fp16 output plus the two intermediates (hash encoding [32 x B], density net output [16 x B]).
//   march  : set_rays / near_far / init_step0 / kernel_march_rays over WHOLE rays (n_step = max_step, buffers zeroed)
//            -> per-ray sample counts, xyz and deltas: the reference GPU's (FMA-contracted) traversal.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <memory>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

// Second CUDA-12 incompatibility of the reference (besides cuda_graph.h): NerfNetwork calls tcnn::parallel_for_gpu
// unqualified from namespace ngp (nerf_network.h:434,467,...), which only older nvcc front ends resolved through the
// later `using namespace tcnn`.  Made visible here, without touching the reference file.
#include <tiny-cuda-nn/common.h>
#include <tiny-cuda-nn/common_device.h>
namespace ngp { using tcnn::parallel_for_gpu; }

#define private public
#define protected public
#include NRF_RENDER_SRC
#undef private
#undef protected

using namespace ngp;

// The per-level grid scale exactly as tcnn's kernel_grid evaluates it on the device (grid.h:159,186: `level` is a
// uint32_t, scale = exp2f(level * log2_per_level_scale) * base_resolution - 1.0f), compiled with the same nvcc defaults.
// The reference never exposes this value; the oracle needs it because MUFU.EX2 cannot be reproduced on the host.
__global__ void ref_level_scales(uint32_t n_levels, uint32_t base_resolution, float log2_per_level_scale, float* out) {
  const uint32_t level = threadIdx.x;
  if (level >= n_levels) return;
  const float scale = exp2f(level * log2_per_level_scale) * base_resolution - 1.0f;
  out[level] = scale;
}

static void write_file(const std::string& path, const void* p, size_t bytes) {
  FILE* f = fopen(path.c_str(), "wb");
  if (!f) { fprintf(stderr, "cannot write %s\n", path.c_str()); exit(2); }
  fwrite(p, 1, bytes, f);
  fclose(f);
}
static std::vector<float> read_floats(const std::string& path, size_t n) {
  std::vector<float> v(n);
  FILE* f = fopen(path.c_str(), "rb");
  if (!f || fread(v.data(), 4, n, f) != n) { fprintf(stderr, "cannot read %zu floats from %s\n", n, path.c_str()); exit(2); }
  fclose(f);
  return v;
}
static Eigen::Matrix<float, 4, 4> pose_from(const float* p) {  // render_server.cu:95-99: row-major 16 floats
  Eigen::Matrix<float, 4, 4> m;
  for (int r = 0; r < 4; ++r)
    for (int c = 0; c < 4; ++c) m(r, c) = p[r * 4 + c];
  return m;
}

int main(int argc, char** argv) {
  if (argc < 4) { fprintf(stderr, "usage: see header\n"); return 2; }
  const std::string snapshot = argv[1], out = argv[2];
  NerfRender* render = new NerfRender();
  auto t0 = std::chrono::steady_clock::now();
  render->reload_network_from_file(snapshot);
  double load_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  printf("{\"event\": \"loaded\", \"seconds\": %.3f, \"n_params\": %zu}\n", load_s, render->m_nerf_network[0]->n_params());
  {
    // encoding parameters as reset_network derives them (nerf_render.cu:125-170)
    const json& enc = render->m_network_config["encoding"];
    const uint32_t n_levels = enc.value("n_levels", 16u), log2_hashmap_size = enc.value("log2_hashmap_size", 15);
    uint32_t base = enc.value("base_resolution", 0);
    if (!base) base = 1u << (log2_hashmap_size / 3);
    float pls = enc.value("per_level_scale", 0.0f);
    if (pls <= 0.0f && n_levels > 1) pls = std::exp(std::log(2048.0f * (float)render->m_bound / (float)base) / (n_levels - 1));
    tcnn::GPUMemory<float> d_scales(n_levels);
    ref_level_scales<<<1, 32>>>(n_levels, base, std::log2(pls), d_scales.data());
    std::vector<float> h(n_levels);
    d_scales.copy_to_host(h);
    write_file(out + "/level_scales.f32", h.data(), h.size() * 4);
    printf("{\"event\": \"level_scales\", \"n_levels\": %u, \"base_resolution\": %u, \"per_level_scale\": %.9g}\n", n_levels, base, pls);
  }
  int a = 3;
  while (a < argc) {
    const std::string mode = argv[a++];
    if (mode == "frames") {
      const int W = atoi(argv[a]), H = atoi(argv[a + 1]);
      Camera cam{(float)atof(argv[a + 2]), (float)atof(argv[a + 3]), (float)atof(argv[a + 4]), (float)atof(argv[a + 5])};
      const std::string poses_path = argv[a + 6];
      const int n_poses = atoi(argv[a + 7]), reps = atoi(argv[a + 8]);
      a += 9;
      auto poses = read_floats(poses_path, (size_t)16 * n_poses);
      render->set_resolution(Eigen::Vector2i(W, H));
      const size_t N = (size_t)W * H;
      for (int k = 0; k < n_poses; ++k) {
        for (int r = 0; r < reps; ++r) {
          cudaDeviceSynchronize();
          auto s = std::chrono::steady_clock::now();
          Image img = render->render_frame(cam, pose_from(&poses[(size_t)16 * k]));
          double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - s).count();
          printf("{\"event\": \"frame\", \"W\": %d, \"H\": %d, \"pose\": %d, \"rep\": %d, \"ms\": %.4f}\n", W, H, k, r, ms);
          if (r == 0) {
            char nm[64];
            snprintf(nm, sizeof nm, "/frame_%dx%d_%d", W, H, k);
            write_file(out + nm + "_rgb.f32", render->image_h, N * 3 * 4);
            write_file(out + nm + "_depth.f32", render->deep_h, N * 4);
            write_file(out + nm + "_rgb.u8", img.rgb, N * 3);
          }
        }
        fflush(stdout);
      }
    } else if (mode == "bench") {
      // bench.py contract: W untimed warm-up frames, then K timed frames cycling the poses, wall clock around the calls
      const int W = atoi(argv[a]), H = atoi(argv[a + 1]);
      Camera cam{(float)atof(argv[a + 2]), (float)atof(argv[a + 3]), (float)atof(argv[a + 4]), (float)atof(argv[a + 5])};
      const std::string poses_path = argv[a + 6];
      const int n_poses = atoi(argv[a + 7]), warm = atoi(argv[a + 8]), steps = atoi(argv[a + 9]);
      a += 10;
      auto poses = read_floats(poses_path, (size_t)16 * n_poses);
      render->set_resolution(Eigen::Vector2i(W, H));
      for (int i = 0; i < warm; ++i) render->render_frame(cam, pose_from(&poses[(size_t)16 * (i % n_poses)]));
      cudaDeviceSynchronize();
      auto s0 = std::chrono::steady_clock::now();
      for (int i = 0; i < steps; ++i) render->render_frame(cam, pose_from(&poses[(size_t)16 * ((warm + i) % n_poses)]));
      cudaDeviceSynchronize();
      double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - s0).count();
      printf("{\"event\": \"bench\", \"W\": %d, \"H\": %d, \"warmup\": %d, \"steps\": %d, \"ms_total\": %.4f, \"ms_per_step\": %.4f}\n",
             W, H, warm, steps, ms, ms / steps);
    } else if (mode == "net") {
      const std::string in_path = argv[a];
      const uint32_t B = (uint32_t)atoi(argv[a + 1]);
      a += 2;
      if (B % 128) { fprintf(stderr, "net: B must be a multiple of 128\n"); return 2; }
      auto in = read_floats(in_path, (size_t)6 * B);
      auto& net = *render->m_nerf_network[0];
      cudaStream_t st = render->m_inference_stream[0];
      tcnn::GPUMatrixDynamic<float> input(6, B, tcnn::RM);
      tcnn::GPUMatrixDynamic<precision_t> output(net.padded_output_width(), B, tcnn::RM);
      cudaMemcpy(input.data(), in.data(), in.size() * 4, cudaMemcpyHostToDevice);
      net.inference_mixed_precision_impl(st, input, output);
      cudaStreamSynchronize(st);
      std::vector<uint16_t> h((size_t)net.padded_output_width() * B);
      cudaMemcpy(h.data(), output.data(), h.size() * 2, cudaMemcpyDeviceToHost);
      write_file(out + "/net_out.f16", h.data(), h.size() * 2);
      // intermediates with the same calls inference_mixed_precision_impl makes (nerf_network.h:155-175)
      tcnn::GPUMatrixDynamic<precision_t> enc(net.m_pos_encoding->padded_output_width(), B, st,
                                               net.m_pos_encoding->preferred_output_layout());
      tcnn::GPUMatrixDynamic<precision_t> den(net.m_density_network->padded_output_width(), B, st,
                                               net.m_pos_encoding->preferred_output_layout());
      net.m_pos_encoding->inference_mixed_precision(st, input.slice_rows(0, net.m_pos_encoding->input_width()), enc, true);
      net.m_density_network->inference_mixed_precision(st, enc, den, true);
      tcnn::GPUMatrixDynamic<precision_t> sh(net.m_dir_encoding->padded_output_width(), B, st,
                                              net.m_dir_encoding->preferred_output_layout());
      net.m_dir_encoding->inference_mixed_precision(st, input.slice_rows(net.m_dir_offset, net.m_dir_encoding->input_width()), sh, true);
      cudaStreamSynchronize(st);
      std::vector<uint16_t> he((size_t)enc.m() * B), hd((size_t)den.m() * B), hs((size_t)sh.m() * B);
      cudaMemcpy(he.data(), enc.data(), he.size() * 2, cudaMemcpyDeviceToHost);
      cudaMemcpy(hd.data(), den.data(), hd.size() * 2, cudaMemcpyDeviceToHost);
      cudaMemcpy(hs.data(), sh.data(), hs.size() * 2, cudaMemcpyDeviceToHost);
      write_file(out + "/net_enc.f16", he.data(), he.size() * 2);
      write_file(out + "/net_density.f16", hd.data(), hd.size() * 2);
      write_file(out + "/net_sh.f16", hs.data(), hs.size() * 2);
      printf("{\"event\": \"net\", \"B\": %u, \"out_rows\": %u, \"enc_rows\": %u, \"enc_layout\": \"%s\", \"density_rows\": %u, "
             "\"sh_rows\": %u, \"sh_layout\": \"%s\"}\n",
             B, net.padded_output_width(), enc.m(), enc.layout() == tcnn::CM ? "CM" : "RM", den.m(), sh.m(),
             sh.layout() == tcnn::CM ? "CM" : "RM");
    } else if (mode == "march") {
      const int W = atoi(argv[a]), H = atoi(argv[a + 1]);
      Camera cam{(float)atof(argv[a + 2]), (float)atof(argv[a + 3]), (float)atof(argv[a + 4]), (float)atof(argv[a + 5])};
      auto pose = read_floats(argv[a + 6], 16);
      const int max_step = atoi(argv[a + 7]);
      a += 8;
      render->set_resolution(Eigen::Vector2i(W, H));
      const int N = W * H;
      const size_t g = render->rays_o.size() - 1;  // defect D7: set_resolution appends; use the buffers just made
      cudaStream_t st = render->m_inference_stream[0];
      // generate_rays always uses index [0] of the buffer vectors (D7); call the kernels directly on index g instead
      Eigen::Matrix<float, 4, 4> new_pose = nerf_matrix_to_ngp(pose_from(pose.data()), render->m_scale);
      const int T = render->m_num_thread, grid_size = (N + T) / T;
      set_rays_o<<<grid_size, T, 0, st>>>(render->rays_o[g].view(), new_pose.block<3, 1>(0, 3), N);
      set_rays_d<<<grid_size, T, 0, st>>>(render->rays_d[g].view(), cam, new_pose.block<3, 3>(0, 0), W, N, 0);
      kernel_near_far_from_aabb<<<tcnn::div_round_up(N, T), T, 0, st>>>(render->rays_o[g].view(), render->rays_d[g].view(),
                                                                       render->m_aabb[0].data(), N, render->m_min_near,
                                                                       render->nears[g].view(), render->fars[g].view());
      cudaMemsetAsync(render->rays_alive[g].data(), 0, sizeof(int) * 2 * N, st);
      cudaMemsetAsync(render->rays_t[g].data(), 0, sizeof(float) * 2 * N, st);
      init_step0<<<tcnn::div_round_up(N, T), T, 0, st>>>(render->rays_alive[g].view(), render->rays_t[g].view(), N,
                                                          render->nears[g].view());
      const size_t B = (size_t)N * max_step;
      tcnn::GPUMatrixDynamic<float> xyzs(3, B, tcnn::CM), dirs(3, B, tcnn::CM), deltas(2, B, tcnn::CM);
      cudaMemsetAsync(xyzs.data(), 0, 3 * B * 4, st);
      cudaMemsetAsync(deltas.data(), 0, 2 * B * 4, st);
      kernel_march_rays<<<tcnn::div_round_up(N, T), T, 0, st>>>(
          N, max_step, render->rays_alive[g].view(), render->rays_t[g].view(), render->rays_o[g].view(),
          render->rays_d[g].view(), render->m_bound, render->m_dt_gamma, render->m_dg_cascade, render->m_dg_h,
          render->m_density_grid[0].data(), render->m_mean_density, render->nears[g].view(), render->fars[g].view(),
          xyzs.view(), dirs.view(), deltas.view(), 0, 0);
      cudaStreamSynchronize(st);
      std::vector<float> hx(3 * B), hd(2 * B), hn(N), hf(N), hrd(3 * (size_t)N);
      cudaMemcpy(hx.data(), xyzs.data(), hx.size() * 4, cudaMemcpyDeviceToHost);
      cudaMemcpy(hd.data(), deltas.data(), hd.size() * 4, cudaMemcpyDeviceToHost);
      cudaMemcpy(hn.data(), render->nears[g].data(), N * 4, cudaMemcpyDeviceToHost);
      cudaMemcpy(hf.data(), render->fars[g].data(), N * 4, cudaMemcpyDeviceToHost);
      cudaMemcpy(hrd.data(), render->rays_d[g].data(), hrd.size() * 4, cudaMemcpyDeviceToHost);
      write_file(out + "/march_xyz.f32", hx.data(), hx.size() * 4);
      write_file(out + "/march_deltas.f32", hd.data(), hd.size() * 4);
      write_file(out + "/march_nears.f32", hn.data(), hn.size() * 4);
      write_file(out + "/march_fars.f32", hf.data(), hf.size() * 4);
      write_file(out + "/march_rays_d.f32", hrd.data(), hrd.size() * 4);
      printf("{\"event\": \"march\", \"N\": %d, \"max_step\": %d, \"cuda\": \"%s\"}\n", N, max_step,
             cudaGetErrorString(cudaGetLastError()));
    } else {
      fprintf(stderr, "unknown mode %s\n", mode.c_str());
      return 2;
    }
  }
  printf("{\"event\": \"done\", \"cuda\": \"%s\"}\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
