// nrf_snapshot.h — reader for the reference's network config / snapshot files.
//
// Reference: NerfRender::load_network_config (src/nerf_render.cu:66-91: ".json" -> json::parse with comments + the
// "parent" chain of merge_parent_network_config :30-44; ".msgpack" -> json::from_msgpack), load_snapshot (:431-473),
// reset_network (:111-184) and NerfNetwork::deserialize (include/nerf-cuda/nerf_network.h:424-443).  Schema:
//   { "encoding": {otype,n_levels|n_features,n_features_per_level,log2_hashmap_size,base_resolution[,per_level_scale]},
//     "network": {otype,activation,output_activation,n_neurons,n_hidden_layers[,sigma_activation]},
//     "dir_encoding": {...}, "rgb_network": {otype,activation,output_activation,n_neurons,n_hidden_layers},
//     "snapshot": {aabb:[6], bound, scale, cascade, density_grid_size, mean_density,
//                  density_grid:[C*H^3 numbers], params:[numbers]} }
// All arrays are msgpack arrays of NUMBERS (nlohmann writes float32 / float64 / any integer width as it sees fit), not
// binary blobs.  The reference decodes everything into a nlohmann DOM; here the config goes into a small DOM of its own
// (Val) — the same one the JSON text parser fills, so the parent merge (RFC 7386 merge_patch) and the model description
// are taken from one place — while the two big arrays (12-266 M numbers) are streamed straight into std::vector<float>.
// Parsed snapshots can be kept in a binary side-file next to the source (SnapshotCache below): a 266 M-number msgpack
// array costs seconds to decode, the cache is one fread.
#pragma once
#include <cctype>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include <sys/stat.h>

#include "../../include/nerf_b200.h"

namespace nrf {

struct Snapshot {
  nrf_model_desc desc;
  std::vector<float> params, density_grid;
  float aabb[6] = {0, 0, 0, 0, 0, 0};
  bool has_snapshot = false;
  bool from_cache = false;
};

// ---------------------------------------------------------------------------------------------------- config DOM
struct Val {
  enum Type { Null, Bool, Num, Str, Arr, Obj } type = Null;
  double num = 0;
  bool b = false;
  std::string str;
  std::vector<Val> arr;
  std::vector<std::pair<std::string, Val>> obj;   // insertion order; keys unique

  const Val* find(const std::string& k) const {
    if (type != Obj) return nullptr;
    for (const auto& kv : obj) if (kv.first == k) return &kv.second;
    return nullptr;
  }
  Val* find(const std::string& k) { return const_cast<Val*>(static_cast<const Val*>(this)->find(k)); }
  Val& at_or_insert(const std::string& k) {
    if (Val* v = find(k)) return *v;
    obj.emplace_back(k, Val{});
    return obj.back().second;
  }
  void erase(const std::string& k) {
    for (size_t i = 0; i < obj.size(); ++i) if (obj[i].first == k) { obj.erase(obj.begin() + (long)i); return; }
  }
  // nlohmann basic_json::merge_patch (RFC 7386), what merge_parent_network_config applies (nerf_render.cu:42)
  void merge_patch(const Val& patch) {
    if (patch.type != Obj) { *this = patch; return; }
    if (type != Obj) { *this = Val{}; type = Obj; }
    for (const auto& kv : patch.obj) {
      if (kv.second.type == Null) erase(kv.first);
      else at_or_insert(kv.first).merge_patch(kv.second);
    }
  }
};

// JSON text, with // and /* */ comments ignored like json::parse(f, nullptr, true, /*ignore_comments=*/true)
class JsonText {
 public:
  JsonText(const char* p, size_t n) : p_(p), end_(p + n) {}
  Val parse() {
    Val v = value();
    ws();
    if (p_ != end_) err("trailing characters");
    return v;
  }

 private:
  [[noreturn]] void err(const char* what) const { throw std::runtime_error(std::string("json: ") + what); }
  void ws() {
    for (;;) {
      while (p_ < end_ && (*p_ == ' ' || *p_ == '\t' || *p_ == '\n' || *p_ == '\r')) ++p_;
      if (p_ + 1 < end_ && p_[0] == '/' && p_[1] == '/') { while (p_ < end_ && *p_ != '\n') ++p_; continue; }
      if (p_ + 1 < end_ && p_[0] == '/' && p_[1] == '*') {
        p_ += 2;
        while (p_ + 1 < end_ && !(p_[0] == '*' && p_[1] == '/')) ++p_;
        if (p_ + 1 >= end_) err("unterminated comment");
        p_ += 2;
        continue;
      }
      return;
    }
  }
  bool lit(const char* s) {
    const size_t n = std::strlen(s);
    if ((size_t)(end_ - p_) >= n && std::memcmp(p_, s, n) == 0) { p_ += n; return true; }
    return false;
  }
  std::string string() {
    if (p_ >= end_ || *p_ != '"') err("expected string");
    ++p_;
    std::string s;
    while (p_ < end_ && *p_ != '"') {
      if (*p_ == '\\') {
        if (++p_ >= end_) err("bad escape");
        switch (*p_) {
          case 'n': s += '\n'; break; case 't': s += '\t'; break; case 'r': s += '\r'; break;
          case 'b': s += '\b'; break; case 'f': s += '\f'; break;
          case 'u': {   // BMP code point -> UTF-8 (config keys / file names are ASCII in practice)
            if (end_ - p_ < 5) err("bad \\u escape");
            unsigned cp = 0;
            for (int i = 1; i <= 4; ++i) {
              const char c = p_[i];
              cp = cp * 16 + (unsigned)(c >= '0' && c <= '9' ? c - '0' : (std::tolower((unsigned char)c) - 'a' + 10) & 15);
            }
            p_ += 4;
            if (cp < 0x80) s += (char)cp;
            else if (cp < 0x800) { s += (char)(0xc0 | (cp >> 6)); s += (char)(0x80 | (cp & 0x3f)); }
            else { s += (char)(0xe0 | (cp >> 12)); s += (char)(0x80 | ((cp >> 6) & 0x3f)); s += (char)(0x80 | (cp & 0x3f)); }
            break;
          }
          default: s += *p_;
        }
        ++p_;
      } else {
        s += *p_++;
      }
    }
    if (p_ >= end_) err("unterminated string");
    ++p_;
    return s;
  }
  Val value() {
    ws();
    if (p_ >= end_) err("unexpected end");
    Val v;
    const char c = *p_;
    if (c == '{') {
      ++p_;
      v.type = Val::Obj;
      ws();
      if (p_ < end_ && *p_ == '}') { ++p_; return v; }
      for (;;) {
        ws();
        std::string k = string();
        ws();
        if (p_ >= end_ || *p_ != ':') err("expected ':'");
        ++p_;
        v.at_or_insert(k) = value();   // duplicate keys: the last one wins, like nlohmann
        ws();
        if (p_ < end_ && *p_ == ',') { ++p_; continue; }
        if (p_ < end_ && *p_ == '}') { ++p_; return v; }
        err("expected ',' or '}'");
      }
    }
    if (c == '[') {
      ++p_;
      v.type = Val::Arr;
      ws();
      if (p_ < end_ && *p_ == ']') { ++p_; return v; }
      for (;;) {
        v.arr.push_back(value());
        ws();
        if (p_ < end_ && *p_ == ',') { ++p_; continue; }
        if (p_ < end_ && *p_ == ']') { ++p_; return v; }
        err("expected ',' or ']'");
      }
    }
    if (c == '"') { v.type = Val::Str; v.str = string(); return v; }
    if (lit("true")) { v.type = Val::Bool; v.b = true; return v; }
    if (lit("false")) { v.type = Val::Bool; v.b = false; return v; }
    if (lit("null")) return v;
    // number
    const char* s = p_;
    while (p_ < end_ && (std::isdigit((unsigned char)*p_) || *p_ == '-' || *p_ == '+' || *p_ == '.' || *p_ == 'e' || *p_ == 'E')) ++p_;
    if (p_ == s) err("unexpected character");
    v.type = Val::Num;
    v.num = std::strtod(std::string(s, p_).c_str(), nullptr);
    return v;
  }
  const char* p_;
  const char* end_;
};

// ---------------------------------------------------------------------------------------------------- msgpack
class MsgpackReader {
 public:
  MsgpackReader(const uint8_t* p, size_t n) : p_(p), end_(p + n) {}

  uint8_t peek() const { need(1); return *p_; }
  bool next_is_array() const { const uint8_t b = peek(); return (b & 0xf0) == 0x90 || b == 0xdc || b == 0xdd; }
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
  // one number in any of the encodings nlohmann's to_msgpack picks; bool counts like nlohmann's implicit conversion
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
  // streams an array of numbers into floats: values_full[i] = data["params"].at(i) (nerf_network.h:430-432) is a
  // number -> float conversion, i.e. (float) of the stored double / integer
  void read_float_array(std::vector<float>& out) {
    const uint32_t n = read_array();
    out.resize(n);
    for (uint32_t i = 0; i < n; ++i) {
      if (p_ + 5 <= end_ && *p_ == 0xca) {   // fast path: float32
        uint32_t u = (uint32_t)p_[1] << 24 | (uint32_t)p_[2] << 16 | (uint32_t)p_[3] << 8 | p_[4];
        std::memcpy(&out[i], &u, 4);
        p_ += 5;
      } else {
        out[i] = (float)read_number();
      }
    }
  }
  // generic value -> DOM.  big_keys: arrays found under these keys of the CURRENT map are streamed into *big[i] and
  // left in the DOM as an empty array
  Val read_value(int depth = 0, const char* const* big_keys = nullptr, std::vector<float>* const* big = nullptr, int n_big = 0) {
    if (depth > 64) throw std::runtime_error("msgpack: nesting too deep");
    Val v;
    const uint8_t b = peek();
    if ((b & 0xf0) == 0x80 || b == 0xde || b == 0xdf) {
      const uint32_t n = read_map();
      v.type = Val::Obj;
      for (uint32_t i = 0; i < n; ++i) {
        const std::string k = read_str();
        int hit = -1;
        for (int j = 0; j < n_big; ++j) if (k == big_keys[j]) hit = j;
        Val& dst = v.at_or_insert(k);
        if (hit >= 0 && next_is_array()) {
          read_float_array(*big[hit]);
          dst = Val{};
          dst.type = Val::Arr;
        } else {
          dst = read_value(depth + 1);
        }
      }
      return v;
    }
    if (next_is_array()) {
      const uint32_t n = read_array();
      v.type = Val::Arr;
      v.arr.reserve(std::min<uint32_t>(n, 1u << 16));
      for (uint32_t i = 0; i < n; ++i) v.arr.push_back(read_value(depth + 1));
      return v;
    }
    if ((b & 0xe0) == 0xa0 || b == 0xd9 || b == 0xda || b == 0xdb) { v.type = Val::Str; v.str = read_str(); return v; }
    if (b == 0xc0) { take(); return v; }
    if (b == 0xc2 || b == 0xc3) { take(); v.type = Val::Bool; v.b = b == 0xc3; return v; }
    if (b == 0xc4 || b == 0xc5 || b == 0xc6) {   // bin: kept as an opaque string
      take();
      const uint32_t n = b == 0xc4 ? take() : (b == 0xc5 ? be16() : be32());
      need(n);
      v.type = Val::Str;
      v.str.assign(reinterpret_cast<const char*>(p_), n);
      p_ += n;
      return v;
    }
    v.type = Val::Num;
    v.num = read_number();
    return v;
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

 private:
  void need(size_t n) const { if ((size_t)(end_ - p_) < n) throw std::runtime_error("msgpack: truncated input"); }
  uint8_t take() { need(1); return *p_++; }
  uint32_t be16() { need(2); uint32_t v = (uint32_t)p_[0] << 8 | p_[1]; p_ += 2; return v; }
  uint32_t be32() { need(4); uint32_t v = (uint32_t)p_[0] << 24 | (uint32_t)p_[1] << 16 | (uint32_t)p_[2] << 8 | p_[3]; p_ += 4; return v; }
  uint64_t be64() { uint64_t hi = be32(); uint64_t lo = be32(); return hi << 32 | lo; }
  const uint8_t* p_;
  const uint8_t* end_;
};

// ---------------------------------------------------------------------------------------------------- model description
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
inline double num_or(const Val* o, const char* k, double dflt) {
  const Val* v = o ? o->find(k) : nullptr;
  if (!v) return dflt;
  if (v->type == Val::Num) return v->num;
  if (v->type == Val::Bool) return v->b ? 1 : 0;
  throw std::runtime_error(std::string("config key '") + k + "' must be a number");
}
inline std::string str_or(const Val* o, const char* k, const char* dflt) {
  const Val* v = o ? o->find(k) : nullptr;
  if (!v) return dflt;
  if (v->type != Val::Str) throw std::runtime_error(std::string("config key '") + k + "' must be a string");
  return v->str;
}

// What reset_network (nerf_render.cu:111-184), the NerfNetwork ctor (nerf_network.h:95-135) and load_snapshot
// (nerf_render.cu:431-473) take from the config.  Settings this renderer does not implement are REJECTED, not ignored.
inline void describe_model(const Val& cfg, nrf_model_desc& d) {
  std::memset(&d, 0, sizeof(d));
  const Val *enc = cfg.find("encoding"), *net = cfg.find("network"), *rgb = cfg.find("rgb_network"), *dir = cfg.find("dir_encoding");
  // encoding: nerf_render.cu:125-170
  const std::string eo = str_or(enc, "otype", "HashGrid");
  if (!ieq(eo, "HashGrid")) throw std::runtime_error("unsupported encoding otype '" + eo + "'");
  d.n_features_per_level = (uint32_t)num_or(enc, "n_features_per_level", 2);
  const double n_features = num_or(enc, "n_features", 0);
  if (n_features > 0) {   // nerf_render.cu:131-137: n_levels = n_features / n_features_per_level
    if (d.n_features_per_level == 0) throw std::runtime_error("encoding.n_features_per_level must be > 0");
    d.n_levels = (uint32_t)n_features / d.n_features_per_level;
  } else {
    d.n_levels = (uint32_t)num_or(enc, "n_levels", 16);
  }
  d.log2_hashmap_size = (uint32_t)num_or(enc, "log2_hashmap_size", 15);
  d.base_resolution = (uint32_t)num_or(enc, "base_resolution", 0);
  d.per_level_scale = (float)num_or(enc, "per_level_scale", 0.0);
  const std::string interp = str_or(enc, "interpolation", "Linear");
  if (!ieq(interp, "Linear")) throw std::runtime_error("unsupported encoding interpolation '" + interp + "'");
  // density network: tcnn network.cu defaults 128 neurons / 5 hidden layers
  d.n_neurons = (uint32_t)num_or(net, "n_neurons", 128);
  d.density_hidden_layers = (uint32_t)num_or(net, "n_hidden_layers", 5);
  d.sigma_activation = activation_id(str_or(net, "sigma_activation", "Exponential"));   // nerf_network.h:125
  for (const Val* n : {net, rgb}) {
    const std::string a = str_or(n, "activation", "ReLU");
    if (!ieq(a, "ReLU")) throw std::runtime_error("unsupported hidden activation '" + a + "'");
  }
  // the density network's output activation applies to all 16 of its outputs, which feed the rgb network
  const std::string doa = str_or(net, "output_activation", "None");
  if (!ieq(doa, "None")) throw std::runtime_error("unsupported density-network output_activation '" + doa + "' (only None)");
  if (num_or(net, "n_output_dims", 16) != 16) throw std::runtime_error("unsupported density-network n_output_dims (only 16)");
  d.rgb_hidden_layers = (uint32_t)num_or(rgb, "n_hidden_layers", 5);
  d.rgb_activation = activation_id(str_or(rgb, "output_activation", "None"));
  if ((uint32_t)num_or(rgb, "n_neurons", 128) != d.n_neurons) throw std::runtime_error("density and rgb networks must have the same width");
  // direction encoding: Composite[SphericalHarmonics(degree), Identity] (base.json:36-50) or plain SphericalHarmonics
  d.sh_degree = 4;
  if (dir) {
    const Val* nested = dir->find("nested");
    if (nested && nested->type == Val::Arr) {
      for (const Val& e : nested->arr)
        if (ieq(str_or(&e, "otype", ""), "SphericalHarmonics")) d.sh_degree = (uint32_t)num_or(&e, "degree", 4);
    } else if (dir->find("degree")) {
      d.sh_degree = (uint32_t)num_or(dir, "degree", 4);
    }
  }
  // scene: nerf_render.h:55-78 defaults, overwritten by the snapshot block
  const Val* snap = cfg.find("snapshot");
  d.bound = (float)num_or(snap, "bound", 1);
  d.scale = (float)num_or(snap, "scale", 0.33);
  d.cascades = (uint32_t)num_or(snap, "cascade", 1);
  d.grid_size = (uint32_t)num_or(snap, "density_grid_size", 128);
  d.mean_density = (float)num_or(snap, "mean_density", 1.e-4);
}

inline std::vector<uint8_t> read_file(const std::string& path) {
  FILE* f = std::fopen(path.c_str(), "rb");
  // load_network_config, nerf_render.cu:73-76
  if (!f) throw std::runtime_error("Network config \"" + path + "\" does not exist.");
  std::fseek(f, 0, SEEK_END);
  const long sz = std::ftell(f);
  std::fseek(f, 0, SEEK_SET);
  std::vector<uint8_t> buf((size_t)std::max(0L, sz));
  const size_t got = sz > 0 ? std::fread(buf.data(), 1, (size_t)sz, f) : 0;
  std::fclose(f);
  if ((long)got != sz) throw std::runtime_error("short read on " + path);
  return buf;
}
inline bool has_ext(const std::string& p, const char* ext) {   // equals_case_insensitive(path.extension(), ext)
  const size_t dot = p.find_last_of('.');
  const size_t sep = p.find_last_of('/');
  if (dot == std::string::npos || (sep != std::string::npos && dot < sep)) return false;
  return ieq(p.substr(dot + 1), ext);
}

// merge_parent_network_config (nerf_render.cu:30-44): parent path relative to the child's directory, recursively
inline Val load_json_config(const std::string& path, int depth = 0) {
  if (depth > 16) throw std::runtime_error("network config parent chain too deep at " + path);
  const std::vector<uint8_t> buf = read_file(path);
  Val child = JsonText(reinterpret_cast<const char*>(buf.data()), buf.size()).parse();
  const Val* parent_name = child.find("parent");
  if (!parent_name) return child;
  if (parent_name->type != Val::Str) throw std::runtime_error("network config \"parent\" must be a string in " + path);
  const size_t sep = path.find_last_of('/');
  const std::string dir = sep == std::string::npos ? std::string() : path.substr(0, sep + 1);
  Val parent = load_json_config(dir + parent_name->str, depth + 1);
  parent.merge_patch(child);
  return parent;
}

// Parses a .msgpack snapshot buffer.  Throws std::runtime_error with the reference's messages where it has one.
inline void parse_snapshot(const uint8_t* data, size_t size, const std::string& path, Snapshot& out) {
  // top-level map: every block goes into the DOM; the snapshot's two big arrays are streamed
  MsgpackReader r(data, size);
  Val cfg;
  cfg.type = Val::Obj;
  static const char* const big_keys[] = {"params", "density_grid"};
  std::vector<float>* const big[] = {&out.params, &out.density_grid};
  const uint32_t n_top = r.read_map();
  for (uint32_t i = 0; i < n_top; ++i) {
    const std::string key = r.read_str();
    Val& dst = cfg.at_or_insert(key);
    if (key == "snapshot") dst = r.read_value(1, big_keys, big, 2);
    else dst = r.read_value(1);
  }
  const Val* snap = cfg.find("snapshot");
  if (!snap || snap->type != Val::Obj) throw std::runtime_error("File " + path + " does not contain a snapshot.");  // nerf_render.cu:434-436
  out.has_snapshot = true;
  describe_model(cfg, out.desc);
  const Val* aabb = snap->find("aabb");
  if (!aabb || aabb->type != Val::Arr || aabb->arr.size() != 6) throw std::runtime_error("snapshot.aabb must be an array of 6 numbers");
  for (int k = 0; k < 6; ++k) {
    if (aabb->arr[(size_t)k].type != Val::Num) throw std::runtime_error("snapshot.aabb must be an array of 6 numbers");
    out.aabb[k] = (float)aabb->arr[(size_t)k].num;
  }
  if (!snap->find("density_grid")) throw std::runtime_error("snapshot.density_grid missing");
  if (!snap->find("params")) throw std::runtime_error("snapshot.params missing");
  const nrf_model_desc& d = out.desc;
  if (out.density_grid.size() != (size_t)d.grid_size * d.grid_size * d.grid_size * d.cascades)
    throw std::runtime_error("Incompatible number of grid cascades.");  // nerf_render.cu:467-469
}

// ---------------------------------------------------------------------------------------------------- binary side-cache
// "<file>.nrfcache": the parsed snapshot as raw floats.  A cache is used only if it was made from exactly this source:
// the header records the source's size, modification time and an FNV-1a hash over its first and last 64 KB; anything
// else (truncated cache, other version, touched or rewritten source) is rejected and the source is parsed again.
struct CacheHeader {
  char magic[8];          // "NRFCACH2"
  uint64_t src_size, src_mtime_ns, src_hash;
  nrf_model_desc desc;
  float aabb[6];
  uint64_t n_params, n_grid;
};
inline uint64_t fnv1a(const uint8_t* p, size_t n, uint64_t h = 1469598103934665603ull) {
  for (size_t i = 0; i < n; ++i) { h ^= p[i]; h *= 1099511628211ull; }
  return h;
}
inline bool source_identity(const std::string& path, uint64_t& size, uint64_t& mtime_ns, uint64_t& hash) {
  struct stat st;
  if (::stat(path.c_str(), &st) != 0) return false;
  size = (uint64_t)st.st_size;
  mtime_ns = (uint64_t)st.st_mtim.tv_sec * 1000000000ull + (uint64_t)st.st_mtim.tv_nsec;
  FILE* f = std::fopen(path.c_str(), "rb");
  if (!f) return false;
  std::vector<uint8_t> buf(65536);
  uint64_t h = fnv1a(reinterpret_cast<const uint8_t*>(&size), 8);
  size_t got = std::fread(buf.data(), 1, buf.size(), f);
  h = fnv1a(buf.data(), got, h);
  if (size > buf.size()) {
    std::fseek(f, -(long)buf.size(), SEEK_END);
    got = std::fread(buf.data(), 1, buf.size(), f);
    h = fnv1a(buf.data(), got, h);
  }
  std::fclose(f);
  hash = h;
  return true;
}
inline bool cache_load(const std::string& path, Snapshot& out) {
  uint64_t size, mtime, hash;
  if (!source_identity(path, size, mtime, hash)) return false;
  FILE* f = std::fopen((path + ".nrfcache").c_str(), "rb");
  if (!f) return false;
  CacheHeader h;
  bool ok = std::fread(&h, sizeof(h), 1, f) == 1 && std::memcmp(h.magic, "NRFCACH2", 8) == 0 && h.src_size == size &&
            h.src_mtime_ns == mtime && h.src_hash == hash && h.n_params < (1ull << 34) && h.n_grid < (1ull << 34);
  if (ok) {
    out.params.resize((size_t)h.n_params);
    out.density_grid.resize((size_t)h.n_grid);
    ok = std::fread(out.params.data(), 4, out.params.size(), f) == out.params.size() &&
         std::fread(out.density_grid.data(), 4, out.density_grid.size(), f) == out.density_grid.size();
    uint8_t extra;
    ok = ok && std::fread(&extra, 1, 1, f) == 0;   // nothing may follow
  }
  std::fclose(f);
  if (!ok) { out.params.clear(); out.density_grid.clear(); return false; }
  out.desc = h.desc;
  std::memcpy(out.aabb, h.aabb, 24);
  out.has_snapshot = true;
  out.from_cache = true;
  return true;
}
inline void cache_store(const std::string& path, const Snapshot& s) {   // best effort: failures are silent
  CacheHeader h;
  std::memset(&h, 0, sizeof(h));
  std::memcpy(h.magic, "NRFCACH2", 8);
  if (!source_identity(path, h.src_size, h.src_mtime_ns, h.src_hash)) return;
  h.desc = s.desc;
  std::memcpy(h.aabb, s.aabb, 24);
  h.n_params = s.params.size();
  h.n_grid = s.density_grid.size();
  const std::string tmp = path + ".nrfcache.tmp";
  FILE* f = std::fopen(tmp.c_str(), "wb");
  if (!f) return;
  const bool ok = std::fwrite(&h, sizeof(h), 1, f) == 1 && std::fwrite(s.params.data(), 4, s.params.size(), f) == s.params.size() &&
                  std::fwrite(s.density_grid.data(), 4, s.density_grid.size(), f) == s.density_grid.size();
  std::fclose(f);
  if (ok) std::rename(tmp.c_str(), (path + ".nrfcache").c_str());
  else std::remove(tmp.c_str());
}
// NRF_SNAPSHOT_CACHE: "0" never, "1" always, unset: for sources of at least 8 MB (where decoding takes longer than a read)
inline bool cache_enabled(size_t source_bytes) {
  const char* e = std::getenv("NRF_SNAPSHOT_CACHE");
  if (e && e[0] == '0') return false;
  if (e && e[0] == '1') return true;
  return source_bytes >= (8u << 20);
}

}  // namespace nrf
