// TENF ("tensorfile") reader for BEAR data files, with a minimal JSON parser.
// Format: bear/tensorfile.py:21-58 in the reference; error behaviour mirrors
// visr_bear/src/tensorfile.cpp:187-215 (what() strings start with "tensorfile format error: ").
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace tenf {

struct FormatError : std::runtime_error {
  explicit FormatError(const std::string &what) : std::runtime_error("tensorfile format error: " + what) {}
};

struct Json {
  enum Kind { Null, Bool, Number, String, Array, Object } kind = Null;
  bool b = false;
  double num = 0.0;
  bool is_int = false;
  int64_t inum = 0;
  std::string str;
  std::vector<Json> arr;
  std::vector<std::pair<std::string, Json>> obj;

  const Json *find(const std::string &key) const;
  const Json &at(const std::string &key) const;  // throws std::runtime_error if missing
  bool has(const std::string &key) const { return find(key) != nullptr; }
  bool is_tensor() const;
};

Json parse_json(const char *begin, size_t len);

// An array materialised as float32 in C order (the reference accesses tensors through strides;
// Panner additionally asserts C order for the filter tables, panner.cpp:252,260,410).
struct FloatTensor {
  std::vector<size_t> shape;
  std::vector<float> data;
  size_t size() const { return data.size(); }
};

class File {
 public:
  explicit File(const std::string &path);
  const Json &metadata() const { return meta_; }
  FloatTensor unpack_f32(const Json &node) const;

 private:
  std::vector<unsigned char> bytes_;
  Json meta_;
};

}  // namespace tenf
