// nrf_snapshot.h — reader for the reference's snapshot file (msgpack, the project's torch-ngp export schema).
//
// Schema (src/nerf_render.cu:431-473, include/nerf-cuda/nerf_network.h:424-431, configs/nerf/base.json):
//   { "encoding": {otype,n_levels,n_features_per_level,log2_hashmap_size,base_resolution[,per_level_scale]},
//     "network": {otype,activation,output_activation,n_neurons,n_hidden_layers[,sigma_activation]},
//     "dir_encoding": {...}, "rgb_network": {otype,activation,output_activation,n_neurons,n_hidden_layers},
//     "snapshot": {aabb:[6], bound, scale, cascade, density_grid_size, mean_density,
//                  density_grid:[C*H^3 numbers], params:[numbers]} }
// All arrays are msgpack arrays of numbers (float32/float64/ints), NOT binary blobs.  The reference decodes with
// nlohmann::json::from_msgpack (nerf_render.cu:85-86) into a DOM; here the two big arrays are streamed straight into
// std::vector<float> (12-266 M numbers) and everything else goes through a tiny generic skipper.
#pragma once
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/nerf_b200.h"

namespace nrf {

struct Snapshot {
  nrf_model_desc desc;
  std::vector<float> params, density_grid;
  float aabb[6];
  bool has_snapshot = false;
};

class MsgpackReader {
 public:
  MsgpackReader(const uint8_t* p, size_t n) : p_(p), end_(p + n) {}

  bool done() const { return p_ >= end_; }
  uint8_t peek() const { need(1); return *p_; }

  // returns element count if the next object is a map, else throws
  uint32_t read_map() {
    uint8_t b = take();
    if ((b & 0xf0) == 0x80) return b & 0x0f;
    if (b == 0xde) return be16();
    if (b == 0xdf) return be32();
    throw std::runtime_error("msgpack: expected map");
  }
  uint32_t read_array() {
    uint8_t b = take();
    if ((b & 0xf0) == 0x90) return b & 0x0f;
    if (b == 0xdc) return be16();
    if (b == 0xdd) return be32();
    throw std::runtime_error("msgpack: expected array");
  }
  std::string read_str() {
    uint8_t b = take();
    uint32_t n;
    if ((b & 0xe0) == 0xa0) n = b & 0x1f;
    else if (b == 0xd9) n = take();
    else if (b == 0xda) n = be16();
    else if (b == 0xdb) n = be32();
    else throw std::runtime_error("msgpack: expected string");
    need(n);
    std::string s(reinterpret_cast<const char*>(p_), n);
    p_ += n;
    return s;
  }
  bool is_number() const {
    uint8_t b = peek();
    return b <= 0x7f || b >= 0xe0 || (b >= 0xca && b <= 0xd3) || b == 0xc2 || b == 0xc3;
  }
  bool is_str() const {
    uint8_t b = peek();
    return (b & 0xe0) == 0xa0 || b == 0xd9 || b == 0xda || b == 0xdb;
  }
  double read_number() {
    uint8_t b = take();
    if (b <= 0x7f) return b;
    if (b >= 0xe0) return (int8_t)b;
    switch (b) {
      case 0xc2: return 0;
      case 0xc3: return 1;
      case 0xca: { uint32_t u = be32(); float f; std::memcpy(&f, &u, 4); return f; }
      case 0xcb: { uint64_t u = be64(); double d; std::memcpy(&d, &u, 8); return d; }
      case 0xcc: return take();
      case 0xcd: return be16();
      case 0xce: return be32();
      case 0xcf: return (double)be64();
      case 0xd0: return (int8_t)take();
      case 0xd1: return (int16_t)be16();
      case 0xd2: return (int32_t)be32();
      case 0xd3: return (double)(int64_t)be64();
    }
    throw std::runtime_error("msgpack: expected number");
  }
  void read_float_array(std::vector<float>& out) {
    const uint32_t n = read_array();
    out.resize(n);
    for (uint32_t i = 0; i < n; ++i) {
      // fast path: float32
      if (p_ + 5 <= end_ && *p_ == 0xca) {
        uint32_t u = (uint32_t)p_[1] << 24 | (uint32_t)p_[2] << 16 | (uint32_t)p_[3] << 8 | p_[4];
        std::memcpy(&out[i], &u, 4);
        p_ += 5;
      } else {
        out[i] = (float)read_number();  // nlohmann: number -> float conversion, same as .at(i) into float
      }
    }
  }
  void skip() {
    uint8_t b = take();
    if (b <= 0x7f || b >= 0xe0 || b == 0xc0 || b == 0xc2 || b == 0xc3) return;
    if ((b & 0xe0) == 0xa0) { adv(b & 0x1f); return; }
    if ((b & 0xf0) == 0x90) { for (uint32_t i = 0, n = b & 0x0f; i < n; ++i) skip(); return; }
    if ((b & 0xf0) == 0x80) { for (uint32_t i = 0, n = (b & 0x0f) * 2; i < n; ++i) skip(); return; }
    switch (b) {
      case 0xc4: case 0xd9: adv(take()); return;
      case 0xc5: case 0xda: adv(be16()); return;
      case 0xc6: case 0xdb: adv(be32()); return;
      case 0xca: case 0xce: case 0xd2: adv(4); return;
      case 0xcb: case 0xcf: case 0xd3: adv(8); return;
      case 0xcc: case 0xd0: adv(1); return;
      case 0xcd: case 0xd1: adv(2); return;
      case 0xd4: adv(2); return;
      case 0xd5: adv(3); return;
      case 0xd6: adv(5); return;
      case 0xd7: adv(9); return;
      case 0xd8: adv(17); return;
      case 0xc7: { uint32_t n = take(); adv(n + 1); return; }
      case 0xc8: { uint32_t n = be16(); adv(n + 1); return; }
      case 0xc9: { uint32_t n = be32(); adv((size_t)n + 1); return; }
      case 0xdc: { uint32_t n = be16(); for (uint32_t i = 0; i < n; ++i) skip(); return; }
      case 0xdd: { uint32_t n = be32(); for (uint32_t i = 0; i < n; ++i) skip(); return; }
      case 0xde: { uint32_t n = be16(); for (uint32_t i = 0; i < 2 * n; ++i) skip(); return; }
      case 0xdf: { uint32_t n = be32(); for (uint64_t i = 0; i < 2ull * n; ++i) skip(); return; }
    }
    throw std::runtime_error("msgpack: unsupported type byte");
  }

 private:
  void need(size_t n) const { if ((size_t)(end_ - p_) < n) throw std::runtime_error("msgpack: truncated input"); }
  uint8_t take() { need(1); return *p_++; }
  void adv(size_t n) { need(n); p_ += n; }
  uint32_t be16() { need(2); uint32_t v = (uint32_t)p_[0] << 8 | p_[1]; p_ += 2; return v; }
  uint32_t be32() { need(4); uint32_t v = (uint32_t)p_[0] << 24 | (uint32_t)p_[1] << 16 | (uint32_t)p_[2] << 8 | p_[3]; p_ += 4; return v; }
  uint64_t be64() { uint64_t hi = be32(); uint64_t lo = be32(); return hi << 32 | lo; }
  const uint8_t* p_;
  const uint8_t* end_;
};

inline bool ieq(const std::string& a, const char* b) {
  size_t n = std::strlen(b);
  if (a.size() != n) return false;
  for (size_t i = 0; i < n; ++i)
    if (std::tolower((unsigned char)a[i]) != std::tolower((unsigned char)b[i])) return false;
  return true;
}
inline uint32_t activation_id(const std::string& s) {
  if (ieq(s, "None")) return NRF_ACT_NONE;
  if (ieq(s, "ReLU")) return NRF_ACT_RELU;
  if (ieq(s, "Exponential")) return NRF_ACT_EXPONENTIAL;
  if (ieq(s, "Sigmoid")) return NRF_ACT_SIGMOID;
  throw std::runtime_error("unsupported activation '" + s + "'");
}

// Parses a whole snapshot buffer.  Throws std::runtime_error with the reference's messages where it has one.
inline void parse_snapshot(const uint8_t* data, size_t size, const std::string& path, Snapshot& out) {
  nrf_model_desc& d = out.desc;
  std::memset(&d, 0, sizeof(d));
  // defaults: reset_network (nerf_render.cu:111-165), NerfNetwork ctor (nerf_network.h:95-135), nerf_render.h:55-78
  d.n_levels = 16; d.n_features_per_level = 2; d.log2_hashmap_size = 15; d.base_resolution = 0; d.per_level_scale = 0;
  d.n_neurons = 128; d.density_hidden_layers = 5; d.rgb_hidden_layers = 5;  // tcnn network.cu defaults
  d.sh_degree = 4; d.rgb_activation = NRF_ACT_NONE; d.sigma_activation = NRF_ACT_EXPONENTIAL;
  d.bound = 1; d.scale = 0.33f; d.cascades = 1; d.grid_size = 128; d.mean_density = 1.e-4f;
  uint32_t rgb_neurons = 128;
  MsgpackReader r(data, size);
  const uint32_t n_top = r.read_map();
  for (uint32_t i = 0; i < n_top; ++i) {
    const std::string key = r.read_str();
    if (key == "encoding" || key == "network" || key == "rgb_network" || key == "dir_encoding") {
      const uint32_t n = r.read_map();
      for (uint32_t j = 0; j < n; ++j) {
        const std::string k = r.read_str();
        if (key == "dir_encoding") {
          if (k == "nested") {  // [{SphericalHarmonics, degree}, {Identity}]  (composite.h:139-216)
            const uint32_t na = r.read_array();
            for (uint32_t a = 0; a < na; ++a) {
              const uint32_t nm = r.read_map();
              std::string otype; double degree = -1;
              for (uint32_t m = 0; m < nm; ++m) {
                const std::string kk = r.read_str();
                if (kk == "otype") otype = r.read_str();
                else if (kk == "degree" && r.is_number()) degree = r.read_number();
                else r.skip();
              }
              if (ieq(otype, "SphericalHarmonics") && degree >= 0) d.sh_degree = (uint32_t)degree;
            }
          } else if (k == "degree" && r.is_number()) d.sh_degree = (uint32_t)r.read_number();
          else r.skip();
          continue;
        }
        if (!r.is_number() && !r.is_str()) { r.skip(); continue; }
        if (r.is_str()) {
          const std::string v = r.read_str();
          if (k == "output_activation" && key == "rgb_network") d.rgb_activation = activation_id(v);
          if (k == "sigma_activation" && key == "network") d.sigma_activation = activation_id(v);
          if (k == "activation" && !ieq(v, "ReLU")) throw std::runtime_error("unsupported hidden activation '" + v + "'");
          if (k == "otype" && key == "encoding" && !ieq(v, "HashGrid"))
            throw std::runtime_error("unsupported encoding otype '" + v + "'");
          continue;
        }
        const double v = r.read_number();
        if (key == "encoding") {
          if (k == "n_levels") d.n_levels = (uint32_t)v;
          else if (k == "n_features_per_level") d.n_features_per_level = (uint32_t)v;
          else if (k == "log2_hashmap_size") d.log2_hashmap_size = (uint32_t)v;
          else if (k == "base_resolution") d.base_resolution = (uint32_t)v;
          else if (k == "per_level_scale") d.per_level_scale = (float)v;
        } else if (key == "network") {
          if (k == "n_neurons") d.n_neurons = (uint32_t)v;
          else if (k == "n_hidden_layers") d.density_hidden_layers = (uint32_t)v;
        } else {
          if (k == "n_neurons") rgb_neurons = (uint32_t)v;
          else if (k == "n_hidden_layers") d.rgb_hidden_layers = (uint32_t)v;
        }
      }
    } else if (key == "snapshot") {
      out.has_snapshot = true;
      const uint32_t n = r.read_map();
      for (uint32_t j = 0; j < n; ++j) {
        const std::string k = r.read_str();
        if (k == "aabb") {
          std::vector<float> a;
          r.read_float_array(a);
          if (a.size() != 6) throw std::runtime_error("snapshot.aabb must have 6 entries");
          std::memcpy(out.aabb, a.data(), 24);
        } else if (k == "density_grid") r.read_float_array(out.density_grid);
        else if (k == "params") r.read_float_array(out.params);
        else if (k == "bound") d.bound = (float)r.read_number();
        else if (k == "scale") d.scale = (float)r.read_number();
        else if (k == "cascade") d.cascades = (uint32_t)r.read_number();
        else if (k == "density_grid_size") d.grid_size = (uint32_t)r.read_number();
        else if (k == "mean_density") d.mean_density = (float)r.read_number();
        else r.skip();
      }
    } else {
      r.skip();
    }
  }
  if (!out.has_snapshot) throw std::runtime_error("File " + path + " does not contain a snapshot.");  // nerf_render.cu:435
  if (rgb_neurons != d.n_neurons) throw std::runtime_error("density and rgb networks must have the same width");
  if (out.density_grid.size() != (size_t)d.grid_size * d.grid_size * d.grid_size * d.cascades)
    throw std::runtime_error("Incompatible number of grid cascades.");  // nerf_render.cu:468
}

}  // namespace nrf
