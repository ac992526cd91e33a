// go_canon.cpp — the canonicaliser: Go value (transport form) -> canonical bytes.
//
// The drift check compares values with reflect.DeepEqual (controller.go:539,600,725,795).  The
// GPU compares bytes.  This encoder makes the two agree: it is injective and canonical, so two
// values are DeepEqual exactly when their encodings are byte-identical.
//   node      = tag, content
//   lengths   = u32 little endian;  ints = width byte + i64 little endian
//   nil map / slice / []byte / pointer / interface have their own tags (nil != empty in Go)
//   map       = entries sorted by the canonical bytes of the key (Go map order is random)
//   float     = IEEE bits with -0 folded into +0; NaN is refused (NaN is not DeepEqual to itself)
// It streams over the input once; no value tree is built.
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <cmath>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace ncdh {

enum : uint8_t {
  T_BOOL = 1, T_INT, T_STRING, T_BYTES, T_NILBYTES, T_SLICE, T_NILSLICE, T_MAP, T_NILMAP, T_PTR, T_NILPTR,
  T_STRUCT, T_FLOAT, T_IFACE, T_NILIFACE
};

namespace {
struct Cursor {
  const uint8_t* p;
  const uint8_t* end;
  void need(uint64_t n) const {
    if ((uint64_t)(end - p) < n) throw std::runtime_error("value is truncated");
  }
  uint8_t byte() {
    need(1);
    return *p++;
  }
  uint32_t len32() {
    need(4);
    uint32_t v;
    memcpy(&v, p, 4);
    p += 4;
    return v;
  }
  void copy(std::string& out, uint64_t n) {
    need(n);
    out.append(reinterpret_cast<const char*>(p), n);
    p += n;
  }
};

void encode(Cursor& in, std::string& out, int depth) {
  if (depth > 200) throw std::runtime_error("value nested too deeply");
  const uint8_t tag = in.byte();
  out.push_back((char)tag);
  switch (tag) {
    case T_BOOL: out.push_back(in.byte() ? 1 : 0); return;
    case T_INT: in.copy(out, 9); return;
    case T_STRING:
    case T_BYTES: {
      const uint32_t n = in.len32();
      out.append(reinterpret_cast<const char*>(&n), 4);
      in.copy(out, n);
      return;
    }
    case T_NILBYTES: case T_NILSLICE: case T_NILMAP: case T_NILPTR: case T_NILIFACE: return;
    case T_SLICE:
    case T_STRUCT: {
      const uint32_t n = in.len32();
      out.append(reinterpret_cast<const char*>(&n), 4);
      for (uint32_t i = 0; i < n; ++i) encode(in, out, depth + 1);
      return;
    }
    case T_PTR: encode(in, out, depth + 1); return;
    case T_IFACE: {
      const uint32_t n = in.len32();
      out.append(reinterpret_cast<const char*>(&n), 4);
      in.copy(out, n);
      encode(in, out, depth + 1);
      return;
    }
    case T_FLOAT: {
      in.need(8);
      double f;
      memcpy(&f, in.p, 8);
      in.p += 8;
      if (std::isnan(f)) throw std::runtime_error("NaN cannot be canonicalised: it is not DeepEqual to itself");
      if (f == 0) f = 0.0;
      out.append(reinterpret_cast<const char*>(&f), 8);
      return;
    }
    case T_MAP: {
      const uint32_t n = in.len32();
      std::vector<std::pair<std::string, std::string>> entries(n);
      for (auto& e : entries) {
        encode(in, e.first, depth + 1);
        encode(in, e.second, depth + 1);
      }
      std::sort(entries.begin(), entries.end(),
                [](const auto& a, const auto& b) { return a.first < b.first; });
      for (size_t i = 1; i < entries.size(); ++i)
        if (entries[i].first == entries[i - 1].first) throw std::runtime_error("map holds a key twice");
      out.append(reinterpret_cast<const char*>(&n), 4);
      for (const auto& e : entries) {
        out += e.first;
        out += e.second;
      }
      return;
    }
    default: throw std::runtime_error("unknown tag in value");
  }
}
}  // namespace

// Throws std::runtime_error on malformed input.
std::string canonicalise(const uint8_t* value, uint64_t len) {
  Cursor in{value, value + len};
  std::string out;
  out.reserve(len + 16);
  encode(in, out, 0);
  if (in.p != in.end) throw std::runtime_error("bytes left over after the value");
  return out;
}

}  // namespace ncdh
