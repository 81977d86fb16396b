// testbed — headless counterpart of the reference's src/main.cu:131-207: load a snapshot, render one frame through
// ngp::NerfRender, print the wall time per frame and write the image as binary PPM/PGM (the reference writes PNGs with
// stb and then runs a GLFW/Vulkan/DLSS pass, which needs a graphics stack: unavailable on this box).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <string>

#include "../../include/nerf-cuda/nerf_render.h"

int main(int argc, char** argv) {
  std::string snapshot = argc > 1 ? argv[1] : "./freality.msgpack";
  const int W = argc > 2 ? std::atoi(argv[2]) : 500, H = argc > 3 ? std::atoi(argv[3]) : 500;  // main.cu:153
  const std::string out = argc > 4 ? argv[4] : "image";
  try {
    ngp::NerfRender render;
    render.reload_network_from_file(snapshot);
    ngp::Camera cam = {1375.52f * W / 500 / 2.75f, 1374.49f * H / 500 / 2.75f, W * 0.5f, H * 0.5f};
    // main.cu:150-152 pose (camera-to-world, NeRF convention)
    const float pos[16] = {0.8926439112348871f, 0.08799600283226543f, 0.4420900262071262f, 3.168359405609479f,
                           0.4464189982715247f, -0.03675452191179031f, -0.8940689141475064f, -5.4794898611466945f,
                           -0.062425682580756266f, 0.995442519072023f, -0.07209178487538156f, -0.9791660699008925f,
                           0.0f, 0.0f, 0.0f, 1.0f};
    render.set_resolution(W, H);
    render.render_frame(cam, pos);  // warm-up (allocations)
    const auto t0 = std::chrono::steady_clock::now();
    ngp::Image img = render.render_frame(cam, pos);
    const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    const nrf_stats st = render.stats();
    std::printf("Process time : %f s / frame  (%llu rays, %llu samples, device %.3f ms)\n", s,
                (unsigned long long)st.n_rays, (unsigned long long)st.n_samples, st.ms_total);
    FILE* f = std::fopen((out + ".ppm").c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + out + ".ppm");
    std::fprintf(f, "P6\n%d %d\n255\n", img.W, img.H);
    std::fwrite(img.rgb, 1, (size_t)3 * img.W * img.H, f);
    std::fclose(f);
    f = std::fopen((out + "_depth.pgm").c_str(), "wb");
    std::fprintf(f, "P5\n%d %d\n255\n", img.W, img.H);
    std::fwrite(img.depth, 1, (size_t)img.W * img.H, f);
    std::fclose(f);
  } catch (const std::exception& e) {
    std::fprintf(stderr, "%s\n", e.what());
    return 1;
  }
  return 0;
}
