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
#include <type_traits>
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
  const uint8_t* take(uint64_t n) {
    need(n);
    const uint8_t* at = p;
    p += n;
    return at;
  }
};

// Where encoded bytes go: into a string, or nowhere (a pass that only validates and measures).
struct StrSink {
  std::string& s;
  void put(const void* b, uint64_t n) { s.append(reinterpret_cast<const char*>(b), n); }
  void byte(uint8_t b) { s.push_back((char)b); }
};
struct BufSink {  // a caller-sized buffer: the canonical form is never longer than the transport form
  uint8_t* p;
  uint8_t* end;
  void put(const void* b, uint64_t n) {
    if ((uint64_t)(end - p) < n) throw std::runtime_error("canonical form does not fit the buffer");
    memcpy(p, b, n);
    p += n;
  }
  void byte(uint8_t b) {
    if (p == end) throw std::runtime_error("canonical form does not fit the buffer");
    *p++ = b;
  }
};
struct NullSink {
  void put(const void*, uint64_t) {}
  void byte(uint8_t) {}
};

template <class Sink>
void encode(Cursor& in, Sink& out, int depth) {
  if (depth > 200) throw std::runtime_error("value nested too deeply");
  const uint8_t tag = in.byte();
  out.byte(tag);
  switch (tag) {
    case T_BOOL: out.byte(in.byte() ? 1 : 0); return;
    case T_INT: out.put(in.take(9), 9); return;
    case T_STRING:
    case T_BYTES: {
      const uint32_t n = in.len32();
      out.put(&n, 4);
      out.put(in.take(n), n);
      return;
    }
    case T_NILBYTES: case T_NILSLICE: case T_NILMAP: case T_NILPTR: case T_NILIFACE: return;
    case T_SLICE:
    case T_STRUCT: {
      const uint32_t n = in.len32();
      out.put(&n, 4);
      for (uint32_t i = 0; i < n; ++i) encode(in, out, depth + 1);
      return;
    }
    case T_PTR: encode(in, out, depth + 1); return;
    case T_IFACE: {
      const uint32_t n = in.len32();
      out.put(&n, 4);
      out.put(in.take(n), n);
      encode(in, out, depth + 1);
      return;
    }
    case T_FLOAT: {
      double f;
      memcpy(&f, in.take(8), 8);
      if (std::isnan(f)) throw std::runtime_error("NaN cannot be canonicalised: it is not DeepEqual to itself");
      if (f == 0) f = 0.0;
      out.put(&f, 8);
      return;
    }
    case T_MAP: {
      // Keys are encoded apart (they order the entries); a value is only walked here — checked and
      // measured — and encoded straight from the input once its place in the output is known, so
      // a large value (a 1 MiB Secret) is copied once, not once per container it sits in.
      struct Entry {
        std::string key;
        const uint8_t* vbegin;
        const uint8_t* vend;
      };
      const uint32_t n = in.len32();
      if ((uint64_t)n > (uint64_t)(in.end - in.p)) throw std::runtime_error("value is truncated");  // >= 1 byte per entry
      std::vector<Entry> entries(n);
      NullSink nowhere;
      for (Entry& e : entries) {
        StrSink ks{e.key};
        encode(in, ks, depth + 1);
        e.vbegin = in.p;
        encode(in, nowhere, depth + 1);
        e.vend = in.p;
      }
      std::sort(entries.begin(), entries.end(), [](const Entry& a, const Entry& b) { return a.key < b.key; });
      for (size_t i = 1; i < entries.size(); ++i)
        if (entries[i].key == entries[i - 1].key) throw std::runtime_error("map holds a key twice");
      if constexpr (std::is_same<Sink, NullSink>::value) return;  // checked above; nothing to write, no second walk
      out.put(&n, 4);
      for (const Entry& e : entries) {
        out.put(e.key.data(), e.key.size());
        Cursor v{e.vbegin, e.vend};
        encode(v, out, depth + 1);
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
  StrSink sink{out};
  encode(in, sink, 0);
  if (in.p != in.end) throw std::runtime_error("bytes left over after the value");
  return out;
}

// The same into a buffer of `cap` bytes (cap = len always suffices: every node's canonical form is
// exactly as long as its transport form, map entries are only reordered).  Returns the bytes written.
uint64_t canonicalise_into(const uint8_t* value, uint64_t len, uint8_t* out, uint64_t cap) {
  Cursor in{value, value + len};
  BufSink sink{out, out + cap};
  encode(in, sink, 0);
  if (in.p != in.end) throw std::runtime_error("bytes left over after the value");
  return (uint64_t)(sink.p - out);
}

}  // namespace ncdh
