#include "tenf.hpp"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace tenf {

// ---------------------------------------------------------------- JSON
namespace {
struct Parser {
  const char *p, *end, *begin;
  [[noreturn]] void fail(const char *what)
  {
    throw FormatError("could not parse JSON metadata at " + std::to_string(p - begin) + ": " + what);
  }
  void ws()
  {
    while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) p++;
  }
  bool lit(const char *s)
  {
    size_t n = std::strlen(s);
    if ((size_t)(end - p) >= n && std::memcmp(p, s, n) == 0) {
      p += n;
      return true;
    }
    return false;
  }
  std::string string()
  {
    if (p >= end || *p != '"') fail("expected string");
    p++;
    std::string out;
    while (p < end && *p != '"') {
      if (*p == '\\') {
        if (++p >= end) fail("bad escape");
        switch (*p) {
          case 'n': out += '\n'; break;
          case 't': out += '\t'; break;
          case 'r': out += '\r'; break;
          case 'b': out += '\b'; break;
          case 'f': out += '\f'; break;
          case 'u': {
            if (end - p < 5) fail("bad unicode escape");
            unsigned cp = (unsigned)std::strtoul(std::string(p + 1, p + 5).c_str(), nullptr, 16);
            p += 4;
            if (cp < 0x80) out += (char)cp;
            else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
            else { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
            break;
          }
          default: out += *p;
        }
        p++;
      } else {
        out += *p++;
      }
    }
    if (p >= end) fail("unterminated string");
    p++;
    return out;
  }
  Json value()
  {
    ws();
    if (p >= end) fail("unexpected end");
    Json j;
    if (*p == '{') {
      j.kind = Json::Object;
      p++;
      ws();
      if (p < end && *p == '}') { p++; return j; }
      for (;;) {
        ws();
        std::string key = string();
        ws();
        if (p >= end || *p != ':') fail("expected ':'");
        p++;
        j.obj.emplace_back(std::move(key), value());
        ws();
        if (p < end && *p == ',') { p++; continue; }
        if (p < end && *p == '}') { p++; break; }
        fail("expected ',' or '}'");
      }
    } else if (*p == '[') {
      j.kind = Json::Array;
      p++;
      ws();
      if (p < end && *p == ']') { p++; return j; }
      for (;;) {
        j.arr.push_back(value());
        ws();
        if (p < end && *p == ',') { p++; continue; }
        if (p < end && *p == ']') { p++; break; }
        fail("expected ',' or ']'");
      }
    } else if (*p == '"') {
      j.kind = Json::String;
      j.str = string();
    } else if (lit("true")) {
      j.kind = Json::Bool; j.b = true;
    } else if (lit("false")) {
      j.kind = Json::Bool; j.b = false;
    } else if (lit("null")) {
      j.kind = Json::Null;
    } else if (lit("NaN")) {
      j.kind = Json::Number; j.num = std::nan("");
    } else {
      const char *s = p;
      bool integral = true;
      if (p < end && (*p == '-' || *p == '+')) p++;
      while (p < end && ((*p >= '0' && *p <= '9') || *p == '.' || *p == 'e' || *p == 'E' || *p == '-' || *p == '+')) {
        if (*p == '.' || *p == 'e' || *p == 'E') integral = false;
        p++;
      }
      if (p == s) fail("invalid value");
      std::string tok(s, p);
      j.kind = Json::Number;
      j.num = std::strtod(tok.c_str(), nullptr);
      j.is_int = integral;
      if (integral) j.inum = std::strtoll(tok.c_str(), nullptr, 10);
    }
    return j;
  }
};
}  // namespace

Json parse_json(const char *begin, size_t len)
{
  Parser ps{begin, begin + len, begin};
  Json j = ps.value();
  ps.ws();
  if (ps.p != ps.end) ps.fail("trailing characters");
  return j;
}

const Json *Json::find(const std::string &key) const
{
  if (kind != Object) return nullptr;
  for (auto &kv : obj)
    if (kv.first == key) return &kv.second;
  return nullptr;
}
const Json &Json::at(const std::string &key) const
{
  const Json *j = find(key);
  if (!j) throw std::runtime_error("BEAR data file error: missing key '" + key + "'");
  return *j;
}
bool Json::is_tensor() const
{
  const Json *t = find("_tenf_type");
  return t && t->kind == String;
}

// ---------------------------------------------------------------- container
static uint64_t read_le(const unsigned char *d, int n)
{
  uint64_t acc = 0;
  for (int i = 0; i < n; i++) acc |= (uint64_t)d[i] << (8 * i);
  return acc;
}

File::File(const std::string &path)
{
  FILE *f = std::fopen(path.c_str(), "rb");
  if (!f) throw std::runtime_error("could not open data file: " + path);
  std::fseek(f, 0, SEEK_END);
  long len = std::ftell(f);
  std::fseek(f, 0, SEEK_SET);
  bytes_.resize(len > 0 ? (size_t)len : 0);
  size_t got = bytes_.empty() ? 0 : std::fread(bytes_.data(), 1, bytes_.size(), f);
  std::fclose(f);
  if (got != bytes_.size()) throw std::runtime_error("could not read data file: " + path);

  const size_t header = 24;
  if (bytes_.size() < header) throw FormatError("file not long enough");
  if (std::memcmp(bytes_.data(), "TENF", 4) != 0) throw FormatError("magic number not found");
  if (read_le(bytes_.data() + 4, 4) != 0) throw FormatError("unknown version number");
  uint64_t dlen = read_le(bytes_.data() + 8, 8), jlen = read_le(bytes_.data() + 16, 8);
  // (compared by subtraction: the sum of two untrusted 64-bit lengths can wrap)
  const uint64_t room = bytes_.size() - header;
  if (dlen > room || jlen > room - dlen) throw FormatError("file not long enough");
  if (jlen == 0) throw FormatError("no JSON metadata found");
  meta_ = parse_json(reinterpret_cast<const char *>(bytes_.data()) + header + dlen, jlen);
}

template <class T>
static double load_as_double(const unsigned char *p, bool swap)
{
  unsigned char tmp[sizeof(T)];
  if (swap)
    for (size_t i = 0; i < sizeof(T); i++) tmp[i] = p[sizeof(T) - 1 - i];
  else
    std::memcpy(tmp, p, sizeof(T));
  T v;
  std::memcpy(&v, tmp, sizeof(T));
  return (double)v;
}

FloatTensor File::unpack_f32(const Json &v) const
{
  if (!v.is_tensor() || v.at("_tenf_type").str != "array") throw FormatError("unknown type");
  // shape, strides and base come from the file: non-negative integers that fit the file, or a format
  // error (a negative or huge double converted to size_t would be undefined behaviour)
  auto index_field = [&](const Json &j, const char *what) {
    const double d = j.num;
    if (j.kind != Json::Number || !(d >= 0.0) || d != std::floor(d) || d > 9007199254740992.0)
      throw FormatError(std::string(what) + " must be a non-negative integer");
    return (size_t)d;
  };
  std::vector<size_t> shape, strides;
  for (auto &s : v.at("shape").arr) shape.push_back(index_field(s, "shape"));
  for (auto &s : v.at("strides").arr) strides.push_back(index_field(s, "strides"));
  if (shape.size() != strides.size()) throw FormatError("mismatched strides and shape");
  const std::string dtype = v.at("dtype").str;
  if (dtype.size() < 3) throw FormatError("unknown byte order");
  const char order = dtype[0];
  if (order != '<' && order != '>' && order != '|') throw FormatError("unknown byte order");
  const std::string type = dtype.substr(1);
  const size_t base = index_field(v.at("base"), "base");

  size_t item = 0;
  double (*load)(const unsigned char *, bool) = nullptr;
  if (type == "f4") { item = 4; load = load_as_double<float>; }
  else if (type == "f8") { item = 8; load = load_as_double<double>; }
  else if (type == "u8") { item = 8; load = load_as_double<uint64_t>; }
  else if (type == "u4") { item = 4; load = load_as_double<uint32_t>; }
  else if (type == "u2") { item = 2; load = load_as_double<uint16_t>; }
  else if (type == "u1") { item = 1; load = load_as_double<uint8_t>; }
  else if (type == "i8") { item = 8; load = load_as_double<int64_t>; }
  else if (type == "i4") { item = 4; load = load_as_double<int32_t>; }
  else if (type == "i2") { item = 2; load = load_as_double<int16_t>; }
  else if (type == "i1") { item = 1; load = load_as_double<int8_t>; }
  else throw FormatError("unknown dtype " + dtype);
  for (size_t s : strides)
    if (s % item) throw FormatError("stride is not a multiple of the type size");

  const bool swap = order == '>';  // hosts running CUDA are little endian
  // element count and byte span with overflow checks: neither can exceed the file (every element of a
  // non-degenerate array occupies bytes of it; zero strides are allowed, so the count is bounded separately)
  const size_t fsize = bytes_.size();
  size_t total = 1, span = item;
  for (size_t i = 0; i < shape.size(); i++) {
    if (shape[i] && total > (size_t)1 << 40) throw FormatError("array too large");
    total *= shape[i];
    if (shape[i] > 1) {
      if (strides[i] && shape[i] - 1 > (fsize - std::min(fsize, span)) / strides[i]) throw FormatError("file not long enough");
      span += strides[i] * (shape[i] - 1);
    }
  }
  if (total > (size_t)1 << 40) throw FormatError("array too large");
  if (total && (base > fsize || span > fsize - base)) throw FormatError("file not long enough");

  FloatTensor out;
  out.shape = shape;
  out.data.resize(total);
  // fast path: little-endian float32 in C order
  bool c_order = type == "f4" && !swap;
  for (size_t i = shape.size(), expect = item; i-- > 0 && c_order; expect *= shape[i])
    if (shape[i] > 1 && strides[i] != expect) c_order = false;
  if (c_order) {
    if (total) std::memcpy(out.data.data(), bytes_.data() + base, total * sizeof(float));
    return out;
  }
  std::vector<size_t> idx(shape.size(), 0);
  for (size_t n = 0; n < total; n++) {
    size_t off = base;
    for (size_t i = 0; i < shape.size(); i++) off += idx[i] * strides[i];
    out.data[n] = (float)load(bytes_.data() + off, swap);
    for (int i = (int)shape.size() - 1; i >= 0; i--) {
      if (++idx[i] < shape[i]) break;
      idx[i] = 0;
    }
  }
  return out;
}

}  // namespace tenf
