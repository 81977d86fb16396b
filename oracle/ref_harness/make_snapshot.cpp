// make_snapshot.cpp — writes snapshot fixtures with the REFERENCE's own serialiser: the vendored nlohmann json
// (/root/reference/dependencies/tiny-cuda-nn/dependencies/json/json.hpp), i.e. the library whose from_msgpack decodes
// snapshots in NerfRender::load_network_config (src/nerf_render.cu:66-91).  TEST TOOLING ONLY (dev container).
//
// The product's reader (nerf_cuda_b200/csrc/nrf_snapshot.h) is hand-written; what a hand-written reader gets wrong are
// exactly the encodings this library chooses on its own: float64 (0xcb) for doubles that are not float-exact, float32
// for those that are, positive / negative fixint, uint8/16/32 and int8/16 for whole numbers, str8 for long keys, map16 /
// array32 headers, alphabetical key order.  The DOM below is what the reference expects (nerf_render.cu:438-453,
// nerf_network.h:424-431, configs/nerf/base.json); numbers are chosen so that every one of those encodings occurs.
//
//   make_snapshot <out.msgpack> <child.json> <parent.json>
// out.msgpack : the snapshot; child.json + parent.json : the same network config split over a "parent" chain with
//               comments (load_network_config's .json path: json::parse(f, nullptr, true, true) + merge_patch, :30-44).
// The arrays follow closed formulas (below) that tests/test_snapshot_fixtures.py restates to know the expected values.
#include <cstdint>
#include <cstdio>
#include <fstream>
#include <json/json.hpp>
#include <vector>

using json = nlohmann::json;

int main(int argc, char** argv) {
  if (argc < 4) { std::fprintf(stderr, "usage: make_snapshot out.msgpack child.json parent.json\n"); return 2; }
  const int L = 16, log2T = 8, H = 32, C = 2;
  // level sizes of tcnn grid.h:899-931 for base 16, per_level_scale 1.5: every level is capped at 2^8 entries
  const size_t n_table = (size_t)L * (1u << log2T) * 2, n_params = 10240 + n_table, n_grid = (size_t)C * H * H * H;
  json cfg;
  cfg["encoding"] = {{"otype", "HashGrid"}, {"n_levels", L}, {"n_features_per_level", 2}, {"log2_hashmap_size", log2T},
                     {"base_resolution", 16}, {"per_level_scale", 1.5}};
  cfg["network"] = {{"otype", "FullyFusedMLP"}, {"activation", "ReLU"}, {"output_activation", "None"}, {"n_neurons", 64},
                    {"n_hidden_layers", 1}};
  cfg["dir_encoding"] = {{"otype", "Composite"},
                         {"nested", json::array({json{{"n_dims_to_encode", 3}, {"otype", "SphericalHarmonics"}, {"degree", 4}},
                                                 json{{"otype", "Identity"}, {"n_bins", 4}, {"degree", 4}}})}};
  cfg["rgb_network"] = {{"otype", "FullyFusedMLP"}, {"activation", "ReLU"}, {"output_activation", "Sigmoid"},
                        {"n_neurons", 64}, {"n_hidden_layers", 2}};
  cfg["loss"] = {{"otype", "Huber"}};   // keys the renderer ignores
  cfg["a_key_that_is_longer_than_thirty_one_bytes_to_force_str8"] = json::array({1, 2.5, "x", nullptr, true});
  json snap;
  snap["aabb"] = json::array({-2, -2, -2, 2, 2, 2});   // negative fixint / positive fixint
  snap["bound"] = 2;                                    // integer where the reader wants a float
  snap["scale"] = 0.33;                                 // not float-exact -> float64
  snap["cascade"] = C;
  snap["density_grid_size"] = H;
  snap["mean_density"] = 0.0078125;                     // float-exact -> float32
  std::vector<json> grid(n_grid), params(n_params);
  for (size_t i = 0; i < n_grid; ++i) {
    // mixture: ints 0 / 1 / 300 / 70000 (fixint, uint16, uint32), -3 (negative fixint), -200 (int16), 0.5 (float32), 0.1*k (float64)
    switch (i % 8) {
      case 0: grid[i] = 0; break;
      case 1: grid[i] = 1; break;
      case 2: grid[i] = 0.5; break;
      case 3: grid[i] = 0.1 * (double)(i % 13); break;
      case 4: grid[i] = (i % 64 == 4) ? 300 : 0; break;
      case 5: grid[i] = (i % 128 == 5) ? 70000 : 1; break;
      case 6: grid[i] = (i % 256 == 6) ? -3 : 0; break;
      default: if (i % 512 == 7) grid[i] = -200; else grid[i] = 0.25; break;
    }
  }
  for (size_t i = 0; i < n_params; ++i) {
    const double v = ((double)((i * 2654435761ull) % 2001) - 1000.0) / 4096.0;   // in [-0.245, 0.245], float-exact (k/4096)
    params[i] = (i % 5 == 0) ? v * 1.1 : v;                                      // every fifth one is not float-exact -> float64
  }
  snap["density_grid"] = grid;
  snap["params"] = params;
  cfg["snapshot"] = snap;
  {
    std::vector<uint8_t> bytes = json::to_msgpack(cfg);
    std::ofstream f(argv[1], std::ios::binary);
    f.write(reinterpret_cast<const char*>(bytes.data()), (std::streamsize)bytes.size());
  }
  // .json with a parent chain: the child overrides two keys and adds one; comments are legal (ignore_comments = true)
  {
    std::ofstream p(argv[3]);
    json parent = cfg;
    parent.erase("snapshot");
    parent["encoding"]["log2_hashmap_size"] = 19;       // overridden by the child
    parent["rgb_network"]["output_activation"] = "None";  // overridden by the child
    p << "// parent network config (comments allowed: json::parse(f, nullptr, true, true))\n" << parent.dump(1, '\t') << "\n";
    std::ofstream c(argv[2]);
    const std::string parent_name = std::string(argv[3]).substr(std::string(argv[3]).find_last_of('/') + 1);
    c << "{\n\t/* child: merge_patch over the parent (nerf_render.cu:30-44) */\n\t\"parent\": \"" << parent_name
      << "\",\n\t\"encoding\": {\"log2_hashmap_size\": " << log2T << "}, // RFC 7386: objects merge key by key\n"
      << "\t\"rgb_network\": {\"output_activation\": \"Sigmoid\"},\n\t\"loss\": null\n}\n";
  }
  return 0;
}
