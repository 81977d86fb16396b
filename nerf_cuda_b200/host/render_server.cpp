// render_server — TCP pose -> RGB frame server with the reference's wire protocol, on POSIX sockets.
//
// Reference: src/render_server.cu:41-107 (sockpp acceptor, one client at a time).  Protocol kept bit-for-bit:
//   request  = 64 bytes: 16 little-endian float32, ROW-MAJOR 4x4 camera-to-world pose (NeRF convention)   (:93-99)
//   response = 3*W*H bytes u8 RGB, row-major, no header                                                     (:101)
//   defaults = port 12345, 1080x1080, intrinsics {840,840,339,590}, snapshot ./freality.msgpack            (:45-64)
// Added (all optional, defaults reproduce the reference): --snapshot, --width/--height, --fl/--cx/--cy, --gpus, and
// --batch N: a request carries N poses (N*64 bytes) and the response is N frames back to back, rendered through the
//            pipelined nrf_render_batch (frame i+1 renders while frame i leaves the GPU);
// --depth  : every frame is followed by its W*H depth bytes (struct Image::depth, common.h:76-89 — the reference computes
//            it and never sends it, render_server.cu:101).
// Unlike the reference, short reads are handled (a pose may arrive in several TCP segments) and a client disconnect
// or a render error does not terminate the server.
#include <arpa/inet.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <sys/socket.h>
#include <unistd.h>

#include <csignal>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/nerf-cuda/nerf_render.h"

static bool read_full(int fd, void* buf, size_t n) {
  char* p = static_cast<char*>(buf);
  while (n) {
    ssize_t r = ::read(fd, p, n);
    if (r <= 0) return false;
    p += r;
    n -= (size_t)r;
  }
  return true;
}
static bool write_full(int fd, const void* buf, size_t n) {
  const char* p = static_cast<const char*>(buf);
  while (n) {
    ssize_t r = ::write(fd, p, n);
    if (r <= 0) return false;
    p += r;
    n -= (size_t)r;
  }
  return true;
}

int main(int argc, char** argv) {
  std::string snapshot = "./freality.msgpack";
  int port = 12345, W = 1080, H = 1080, gpus = 1, batch = 1, max_requests = -1;
  bool depth = false;
  ngp::Camera cam = {840, 840, 339, 590};
  for (int i = 1; i < argc; ++i) {
    std::string a = argv[i];
    auto val = [&](const char* name) -> const char* {
      if (a == name && i + 1 < argc) return argv[++i];
      return nullptr;
    };
    if (const char* v = val("--snapshot")) snapshot = v;
    else if (const char* v = val("--port")) port = std::atoi(v);
    else if (const char* v = val("--width")) W = std::atoi(v);
    else if (const char* v = val("--height")) H = std::atoi(v);
    else if (const char* v = val("--fl")) cam.fl_x = cam.fl_y = (float)std::atof(v);
    else if (const char* v = val("--cx")) cam.cx = (float)std::atof(v);
    else if (const char* v = val("--cy")) cam.cy = (float)std::atof(v);
    else if (const char* v = val("--gpus")) gpus = std::atoi(v);
    else if (const char* v = val("--batch")) batch = std::atoi(v);
    else if (a == "--depth") depth = true;
    else if (const char* v = val("--max-requests")) max_requests = std::atoi(v);  // for tests
    else if (i == 1 && a[0] != '-') port = std::atoi(argv[1]);                     // reference: argv[1] = port
    else { std::fprintf(stderr, "unknown argument %s\n", a.c_str()); return 2; }
  }
  if (batch < 1 || W < 1 || H < 1 || gpus < 1) { std::fprintf(stderr, "--batch, --width, --height and --gpus must be >= 1\n"); return 2; }
  std::signal(SIGPIPE, SIG_IGN);
  try {
    ngp::NerfRender render(gpus);
    render.reload_network_from_file(snapshot);
    render.set_resolution(W, H);

    int srv = ::socket(AF_INET, SOCK_STREAM, 0);
    int one = 1;
    ::setsockopt(srv, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
    sockaddr_in addr{};
    addr.sin_family = AF_INET;
    addr.sin_addr.s_addr = htonl(INADDR_ANY);
    addr.sin_port = htons((uint16_t)port);
    if (::bind(srv, reinterpret_cast<sockaddr*>(&addr), sizeof(addr)) != 0 || ::listen(srv, 4) != 0) {
      std::perror("Error creating the acceptor");
      return 1;
    }
    std::printf("Awaiting connections on port %d...\n", port);
    std::fflush(stdout);
    std::vector<float> poses((size_t)batch * 16);
    const size_t px = (size_t)W * H;
    std::vector<unsigned char> frames(batch > 1 ? (size_t)batch * px * 3 : 0), depths(batch > 1 && depth ? (size_t)batch * px : 0);
    int served = 0;
    while (max_requests < 0 || served < max_requests) {
      sockaddr_in peer{};
      socklen_t plen = sizeof(peer);
      int sock = ::accept(srv, reinterpret_cast<sockaddr*>(&peer), &plen);
      if (sock < 0) { std::perror("Error accepting incoming connection"); continue; }
      ::setsockopt(sock, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
      char ip[64];
      inet_ntop(AF_INET, &peer.sin_addr, ip, sizeof(ip));
      std::printf("Received a connection request from %s:%d\n", ip, ntohs(peer.sin_port));
      while ((max_requests < 0 || served < max_requests) && read_full(sock, poses.data(), poses.size() * 4)) {
        bool ok = true;
        try {
          if (batch == 1) {
            ngp::Image img = render.render_frame(cam, poses.data());
            ok = write_full(sock, img.rgb, px * 3) && (!depth || write_full(sock, img.depth, px));
          } else {
            render.render_batch(cam, poses.data(), batch, frames.data(), depth ? depths.data() : nullptr);
            for (int b = 0; b < batch && ok; ++b)
              ok = write_full(sock, frames.data() + (size_t)b * px * 3, px * 3) &&
                   (!depth || write_full(sock, depths.data() + (size_t)b * px, px));
          }
        } catch (const std::exception& e) {
          std::fprintf(stderr, "render_frame failed: %s\n", e.what());
          ok = false;
        }
        ++served;
        if (!ok) break;
      }
      ::close(sock);
      std::printf("Connection closed from %s\n", ip);
      std::fflush(stdout);
    }
    ::close(srv);
  } catch (const std::exception& e) {
    std::fprintf(stderr, "%s\n", e.what());
    return 1;
  }
  return 0;
}
