// vpt_archive.h — ".vptscene" named-array container (header-only, no deps).
//
// A scene travels between the exporter (reference loaders, built only where
// /root/reference exists), the Python fixture generators, the product host
// library and the oracle tools as a flat list of named, typed, n-d arrays.
// Layout (little endian):
//   char[8]  "VPTARCH1"
//   u32      number of arrays
//   per array: u32 name_len, name bytes, u32 dtype, u32 ndim, u64 dims[ndim],
//              u64 nbytes, raw data
// dtype: 0 = float32, 1 = int32, 2 = uint8, 3 = uint64.
// The Python mirror is volumetric-path-tracer_b200/archive.py.
#pragma once

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace vpt {

enum archive_dtype : uint32_t { dt_f32 = 0, dt_i32 = 1, dt_u8 = 2, dt_u64 = 3 };

inline size_t archive_dtype_size(uint32_t dt) {
  switch (dt) {
    case dt_f32: return 4;
    case dt_i32: return 4;
    case dt_u8: return 1;
    case dt_u64: return 8;
    default: return 0;
  }
}

struct archive_array {
  std::string           name;
  uint32_t              dtype = dt_f32;
  std::vector<uint64_t> dims;
  std::vector<uint8_t>  data;

  size_t count() const {
    size_t n = 1;
    for (auto d : dims) n *= (size_t)d;
    return dims.empty() ? 0 : n;
  }
  const float*    f32() const { return (const float*)data.data(); }
  const int32_t*  i32() const { return (const int32_t*)data.data(); }
  const uint8_t*  u8() const { return data.data(); }
  const uint64_t* u64() const { return (const uint64_t*)data.data(); }
};

struct archive {
  std::vector<archive_array> arrays;

  const archive_array* find(const std::string& name) const {
    for (auto& a : arrays)
      if (a.name == name) return &a;
    return nullptr;
  }
  bool has(const std::string& name) const { return find(name) != nullptr; }

  void add(const std::string& name, uint32_t dtype,
      const std::vector<uint64_t>& dims, const void* ptr) {
    auto& a = arrays.emplace_back();
    a.name  = name;
    a.dtype = dtype;
    a.dims  = dims;
    auto n  = a.count() * archive_dtype_size(dtype);
    a.data.resize(n);
    if (n) memcpy(a.data.data(), ptr, n);
  }
  void add_f32(const std::string& name, const std::vector<uint64_t>& dims,
      const float* p) {
    add(name, dt_f32, dims, p);
  }
  void add_i32(const std::string& name, const std::vector<uint64_t>& dims,
      const int32_t* p) {
    add(name, dt_i32, dims, p);
  }

  bool save(const std::string& path, std::string& error) const {
    auto fs = fopen(path.c_str(), "wb");
    if (!fs) {
      error = "cannot create " + path;
      return false;
    }
    auto ok = true;
    auto wr = [&](const void* p, size_t n) {
      if (n && fwrite(p, 1, n, fs) != n) ok = false;
    };
    wr("VPTARCH1", 8);
    auto n = (uint32_t)arrays.size();
    wr(&n, 4);
    for (auto& a : arrays) {
      auto len = (uint32_t)a.name.size();
      wr(&len, 4);
      wr(a.name.data(), len);
      wr(&a.dtype, 4);
      auto nd = (uint32_t)a.dims.size();
      wr(&nd, 4);
      wr(a.dims.data(), 8 * nd);
      auto nb = (uint64_t)a.data.size();
      wr(&nb, 8);
      wr(a.data.data(), a.data.size());
    }
    fclose(fs);
    if (!ok) error = "write error on " + path;
    return ok;
  }

  bool load(const std::string& path, std::string& error) {
    arrays.clear();
    auto fs = fopen(path.c_str(), "rb");
    if (!fs) {
      error = "cannot open " + path;
      return false;
    }
    auto ok = true;
    auto rd = [&](void* p, size_t n) {
      if (n && fread(p, 1, n, fs) != n) ok = false;
    };
    char magic[8];
    rd(magic, 8);
    if (!ok || memcmp(magic, "VPTARCH1", 8) != 0) {
      fclose(fs);
      error = "not a vptscene archive: " + path;
      return false;
    }
    uint32_t n = 0;
    rd(&n, 4);
    for (uint32_t i = 0; ok && i < n; i++) {
      auto&    a   = arrays.emplace_back();
      uint32_t len = 0, nd = 0;
      rd(&len, 4);
      if (!ok || len > 4096) { ok = false; break; }
      a.name.resize(len);
      rd(a.name.data(), len);
      rd(&a.dtype, 4);
      rd(&nd, 4);
      if (!ok || nd > 8) { ok = false; break; }
      a.dims.resize(nd);
      rd(a.dims.data(), 8 * nd);
      uint64_t nb = 0;
      rd(&nb, 8);
      if (!ok || nb != a.count() * archive_dtype_size(a.dtype)) { ok = false; break; }
      a.data.resize(nb);
      rd(a.data.data(), nb);
    }
    fclose(fs);
    if (!ok) error = "corrupt vptscene archive: " + path;
    return ok;
  }
};

}  // namespace vpt
