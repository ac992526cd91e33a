// oracle/go_values.cpp — Go values, reflect.DeepEqual and the canonical encoding, on the CPU.
// TEST INFRASTRUCTURE (see ncd_oracle.cpp for who may load it).
//
// The compare arithmetic of the hot path is reflect.DeepEqual (controller.go:539,600,725,795).
// It lives in the Go standard library (toolchain go 1.24.4, go.mod:3), which is not under
// /root/reference, and no Go toolchain exists in this image, so the rules are restated here
// from the documented behaviour of reflect.DeepEqual:
//   Map     nil-ness must match; lengths must match; every key of a is present in b with a
//           deeply equal value
//   Slice   nil-ness must match (a nil slice is NOT equal to an empty one); lengths match;
//           elements pairwise deeply equal ([]byte: bytewise)
//   String  byte equality          Bool, Int  ==          Float  == (NaN != NaN, +0 == -0)
//   Pointer both nil, or both non-nil with deeply equal pointees
//   Struct  every field, in order
//   Interface  both nil, or same dynamic type and deeply equal values
// Values of different kinds are never equal (DeepEqual requires identical types).
// What pins these rules to the reference: tests/golden/reference_cases.json (the handler tests
// of controller_test.go, which exercise empty-vs-empty maps, non-empty-vs-empty maps, *bool
// dereference and slice length) — checked in tests/test_oracle_golden.py.  Rules no reference
// test exercises (nil vs empty, key order, same-length byte changes) are marked "extrapolated"
// there.
//
// Values arrive in the tag-length-value transport written by tests/govalues.py.
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <cmath>
#include <memory>
#include <stdexcept>
#include <chrono>
#include <string>
#include <thread>
#include <vector>

namespace {

enum Tag : uint8_t {
  T_BOOL = 1, T_INT, T_STRING, T_BYTES, T_NILBYTES, T_SLICE, T_NILSLICE, T_MAP, T_NILMAP, T_PTR, T_NILPTR,
  T_STRUCT, T_FLOAT, T_IFACE, T_NILIFACE
};

struct Node {
  uint8_t tag = 0;
  uint8_t width = 0;          // T_INT
  int64_t i = 0;              // T_BOOL, T_INT
  double f = 0;               // T_FLOAT
  std::string s;              // T_STRING, T_BYTES, T_IFACE (type name)
  std::vector<Node> kids;     // T_SLICE, T_STRUCT, T_PTR (1), T_IFACE (1), T_MAP (k0 v0 k1 v1 ...)
};

struct Reader {
  const uint8_t* p;
  const uint8_t* end;
  uint8_t u8() {
    if (p >= end) throw std::runtime_error("truncated value");
    return *p++;
  }
  uint32_t u32() {
    if (end - p < 4) throw std::runtime_error("truncated value");
    uint32_t v;
    memcpy(&v, p, 4);
    p += 4;
    return v;
  }
  std::string bytes(uint32_t n) {
    if ((uint64_t)(end - p) < n) throw std::runtime_error("truncated value");
    std::string s(reinterpret_cast<const char*>(p), n);
    p += n;
    return s;
  }
};

Node parse(Reader& r, int depth = 0) {
  if (depth > 200) throw std::runtime_error("value nested too deeply");
  Node n;
  n.tag = r.u8();
  switch (n.tag) {
    case T_BOOL: n.i = r.u8() != 0; break;
    case T_INT: {
      n.width = r.u8();
      std::string b = r.bytes(8);
      memcpy(&n.i, b.data(), 8);
      break;
    }
    case T_STRING: case T_BYTES: n.s = r.bytes(r.u32()); break;
    case T_NILBYTES: case T_NILSLICE: case T_NILMAP: case T_NILPTR: case T_NILIFACE: break;
    case T_SLICE: case T_STRUCT: {
      uint32_t c = r.u32();
      for (uint32_t k = 0; k < c; ++k) n.kids.push_back(parse(r, depth + 1));
      break;
    }
    case T_MAP: {
      uint32_t c = r.u32();
      for (uint32_t k = 0; k < 2 * c; ++k) n.kids.push_back(parse(r, depth + 1));
      break;
    }
    case T_PTR: n.kids.push_back(parse(r, depth + 1)); break;
    case T_FLOAT: {
      std::string b = r.bytes(8);
      memcpy(&n.f, b.data(), 8);
      break;
    }
    case T_IFACE: n.s = r.bytes(r.u32()); n.kids.push_back(parse(r, depth + 1)); break;
    default: throw std::runtime_error("unknown value tag");
  }
  return n;
}

bool deep_equal(const Node& a, const Node& b) {
  if (a.tag != b.tag) return false;  // covers nil vs non-nil of every nilable kind
  switch (a.tag) {
    case T_BOOL: return a.i == b.i;
    case T_INT: return a.width == b.width && a.i == b.i;
    case T_STRING: case T_BYTES: return a.s == b.s;
    case T_FLOAT: return a.f == b.f;
    case T_NILBYTES: case T_NILSLICE: case T_NILMAP: case T_NILPTR: case T_NILIFACE: return true;
    case T_SLICE: case T_STRUCT:
      if (a.kids.size() != b.kids.size()) return false;
      for (size_t k = 0; k < a.kids.size(); ++k)
        if (!deep_equal(a.kids[k], b.kids[k])) return false;
      return true;
    case T_PTR: return deep_equal(a.kids[0], b.kids[0]);
    case T_IFACE: return a.s == b.s && deep_equal(a.kids[0], b.kids[0]);
    case T_MAP: {
      if (a.kids.size() != b.kids.size()) return false;
      for (size_t k = 0; k < a.kids.size(); k += 2) {
        bool found = false;
        for (size_t m = 0; m < b.kids.size(); m += 2)
          if (deep_equal(a.kids[k], b.kids[m])) {  // map lookup by key ==
            if (!deep_equal(a.kids[k + 1], b.kids[m + 1])) return false;
            found = true;
            break;
          }
        if (!found) return false;
      }
      return true;
    }
  }
  return false;
}

void put_u32(std::string& o, uint32_t v) { o.append(reinterpret_cast<const char*>(&v), 4); }

// Canonical encoding: tag, then content; lengths as u32 little endian; map entries sorted by the
// canonical bytes of their keys.  Prefix-free per node, hence injective on value trees up to map
// order: canon(a) == canon(b)  <=>  DeepEqual(a, b)  (floats: -0 is written as +0, NaN is refused
// because NaN is not DeepEqual to itself — no float is reachable from the compared types).
void canon(const Node& n, std::string& o) {
  o.push_back((char)n.tag);
  switch (n.tag) {
    case T_BOOL: o.push_back((char)n.i); break;
    case T_INT: o.push_back((char)n.width); o.append(reinterpret_cast<const char*>(&n.i), 8); break;
    case T_STRING: case T_BYTES: put_u32(o, (uint32_t)n.s.size()); o += n.s; break;
    case T_NILBYTES: case T_NILSLICE: case T_NILMAP: case T_NILPTR: case T_NILIFACE: break;
    case T_SLICE: case T_STRUCT:
      put_u32(o, (uint32_t)n.kids.size());
      for (const Node& k : n.kids) canon(k, o);
      break;
    case T_PTR: canon(n.kids[0], o); break;
    case T_IFACE: put_u32(o, (uint32_t)n.s.size()); o += n.s; canon(n.kids[0], o); break;
    case T_FLOAT: {
      if (std::isnan(n.f)) throw std::runtime_error("NaN has no canonical encoding (NaN != NaN)");
      double f = n.f == 0 ? 0.0 : n.f;
      o.append(reinterpret_cast<const char*>(&f), 8);
      break;
    }
    case T_MAP: {
      std::vector<std::pair<std::string, std::string>> ents;
      for (size_t k = 0; k < n.kids.size(); k += 2) {
        std::string ck, cv;
        canon(n.kids[k], ck);
        canon(n.kids[k + 1], cv);
        ents.emplace_back(std::move(ck), std::move(cv));
      }
      std::sort(ents.begin(), ents.end());
      for (size_t k = 1; k < ents.size(); ++k)
        if (ents[k].first == ents[k - 1].first) throw std::runtime_error("duplicate map key");
      put_u32(o, (uint32_t)ents.size());
      for (auto& e : ents) {
        o += e.first;
        o += e.second;
      }
      break;
    }
  }
}

thread_local std::string g_err;

Node parse_all(const uint8_t* p, uint64_t n) {
  Reader r{p, p + n};
  Node v = parse(r);
  if (r.p != r.end) throw std::runtime_error("trailing bytes after value");
  return v;
}

}  // namespace

extern "C" {

const char* oracle_values_error(void) { return g_err.c_str(); }

// 1 equal, 0 different, -1 malformed input.
int oracle_deep_equal(const uint8_t* a, uint64_t alen, const uint8_t* b, uint64_t blen) {
  try {
    return deep_equal(parse_all(a, alen), parse_all(b, blen)) ? 1 : 0;
  } catch (const std::exception& e) {
    g_err = e.what();
    return -1;
  }
}

// Timing helper for the CPU baseline: the walk reflect.DeepEqual does over values that are already
// in memory (controller.go:539,600,725,795).  npairs separate copies of both values are parsed
// first (untimed; separate allocations, so a sweep over them streams from memory like a sweep over
// an informer cache does), then `nthreads` threads compare pair i of their share, `reps` sweeps.
// Returns the wall-clock seconds of the sweeps only, or -1 on malformed input.
double oracle_deep_equal_bench(const uint8_t* a, uint64_t alen, const uint8_t* b, uint64_t blen, uint64_t npairs,
                               uint64_t reps, int nthreads, int* equal_out) {
  try {
    if (nthreads < 1) nthreads = 1;
    std::vector<Node> as, bs;
    as.reserve(npairs);
    bs.reserve(npairs);
    for (uint64_t i = 0; i < npairs; ++i) {
      as.push_back(parse_all(a, alen));
      bs.push_back(parse_all(b, blen));
    }
    std::vector<uint64_t> hits((size_t)nthreads, 0);
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; ++t)
      th.emplace_back([&, t] {
        const uint64_t lo = npairs * (uint64_t)t / (uint64_t)nthreads, hi = npairs * (uint64_t)(t + 1) / (uint64_t)nthreads;
        uint64_t h = 0;
        for (uint64_t r = 0; r < reps; ++r)
          for (uint64_t i = lo; i < hi; ++i) {
            h += deep_equal(as[i], bs[i]) ? 1 : 0;
            asm volatile("" ::: "memory");  // every comparison really walks the values
          }
        hits[(size_t)t] = h;
      });
    for (auto& x : th) x.join();
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    uint64_t total = 0;
    for (uint64_t h : hits) total += h;
    if (equal_out) *equal_out = total == npairs * reps ? 1 : 0;
    return sec;
  } catch (const std::exception& e) {
    g_err = e.what();
    return -1.0;
  }
}

// Writes the canonical encoding to out (capacity cap); returns its length, or -1 on error, or
// the needed length when cap is too small (nothing written).
int64_t oracle_canon(const uint8_t* v, uint64_t vlen, uint8_t* out, uint64_t cap) {
  try {
    std::string o;
    canon(parse_all(v, vlen), o);
    if (o.size() <= cap && out) memcpy(out, o.data(), o.size());
    return (int64_t)o.size();
  } catch (const std::exception& e) {
    g_err = e.what();
    return -1;
  }
}

}  // extern "C"

// ------------------------------------------------------------------ fleet: the handlers over parsed values
// The CPU arm of the steady-state bench and a second, independent statement of the decisions:
// every cluster's informer cache as a map of PARSED objects (what a lister holds), informer events
// applied to it, and a sweep that walks every template and workgroup of the controller cache over
// every shard the way templateSyncHandler / workgroupSyncHandler do (controller.go:790-831 with
// :504-626 and :484-502; :720-745) — lister lookups by key and reflect.DeepEqual on the value
// trees, no canonical bytes, no arenas.  Events arrive in the same ncdh_event records the GPU arm's
// host layer takes (include/ncdhost.h is included for that struct alone).
#include <atomic>
#include <unordered_map>

#include "../include/ncdhost.h"
#include "../include/ncdrift.h"

namespace {

struct FleetObj {
  uint32_t kind = 0;
  std::string key, uid;
  Node value;
  std::vector<std::pair<std::string, std::string>> owners;  // (Kind, UID)
  std::vector<std::string> deps;                            // "ns/name", secrets first
  uint32_t nsecrets = 0;
};
using FleetStore = std::unordered_map<std::string, FleetObj>;  // (kind digit + "ns/name") -> object

inline std::string fkey(uint32_t kind, const std::string& key) { return std::string(1, (char)('0' + kind)) + key; }

inline uint64_t fmix(uint64_t x) {
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdull;
  x ^= x >> 33;
  x *= 0xc4ceb9fe1a85ec53ull;
  return x ^ (x >> 33);
}
inline uint64_t key_fnv(const std::string& skey) {
  uint64_t h = 0xcbf29ce484222325ull;
  for (unsigned char c : skey) h = (h ^ c) * 0x100000001b3ull;
  return h;
}
// The checksum both arms compute over their diff lists: order independent, keyed by NAME (not by
// arena index), so two implementations with different layouts can be compared.
inline uint64_t entry_digest(const std::string& skey, uint32_t dep, uint32_t shard, uint32_t code) {
  return fmix(key_fnv(skey) ^ ((uint64_t)dep * 0x9E3779B185EBCA87ull) ^ ((uint64_t)shard * 0xC2B2AE3D27D4EB4Full) ^
              ((uint64_t)code * 0x165667B19E3779F9ull));
}

// isMissingOwnership, controller.go:484-502
uint32_t fleet_ownership(const FleetObj& dep, const FleetObj* shard_tmpl) {
  if (dep.owners.empty()) return NCD_ERR_ROGUE;  // :494-499
  if (shard_tmpl)
    for (const auto& o : dep.owners)  // :489-492
      if (o.first == "NexusAlgorithmTemplate" && o.second == shard_tmpl->uid) return 0;
  return NCD_UPDATE_OWNER;  // :501
}

}  // namespace

struct oracle_fleet {
  FleetStore controller;
  std::vector<FleetStore> shards;
};

extern "C" {

oracle_fleet* oracle_fleet_create(int nshards) {
  oracle_fleet* f = new oracle_fleet();
  f->shards.resize(nshards < 0 ? 0 : nshards);
  return f;
}
void oracle_fleet_destroy(oracle_fleet* f) { delete f; }

// Applies informer events in order: values are parsed (in parallel, the analogue of an informer
// decoding an object), then stored.  Returns 0, or -1 on a malformed value (nothing applied).
int oracle_fleet_apply(oracle_fleet* f, const ncdh_event* ev, uint64_t n, int nthreads) {
  if (nthreads < 1) nthreads = 1;
  std::vector<FleetObj> objs(n);
  std::atomic<int> bad{0};
  auto work = [&](uint64_t lo, uint64_t hi) {
    for (uint64_t i = lo; i < hi; ++i) {
      const ncdh_event& e = ev[i];
      FleetObj& o = objs[i];
      o.kind = e.kind;
      o.key = std::string(e.ns) + "/" + e.name;
      if (e.op != NCDH_EV_PUT) continue;
      try {
        o.value = parse_all(e.value, e.value_len);
      } catch (const std::exception& ex) {
        g_err = ex.what();
        bad.store(1);
        return;
      }
      o.uid = e.uid;
      for (int k = 0; k < e.nowners; ++k) o.owners.emplace_back(e.owner_kinds[k], e.owner_uids[k]);
      for (int k = 0; k < e.nsecrets; ++k) o.deps.push_back(std::string(e.ns) + "/" + e.secret_names[k]);
      for (int k = 0; k < e.nconfigmaps; ++k) o.deps.push_back(std::string(e.ns) + "/" + e.configmap_names[k]);
      o.nsecrets = (uint32_t)e.nsecrets;
    }
  };
  std::vector<std::thread> th;
  for (int t = 0; t < nthreads; ++t) th.emplace_back(work, n * t / nthreads, n * (t + 1) / nthreads);
  for (auto& x : th) x.join();
  if (bad.load()) return -1;
  for (uint64_t i = 0; i < n; ++i) {
    const int c = ev[i].cluster;
    if (c < -1 || c >= (int)f->shards.size()) return -1;
    FleetStore& st = c < 0 ? f->controller : f->shards[c];
    const std::string k = fkey(objs[i].kind, objs[i].key);
    if (ev[i].op == NCDH_EV_PUT) st[k] = std::move(objs[i]);
    else st.erase(k);
  }
  return 0;
}

// One resync sweep: every template (with its dependents) and workgroup of the controller cache on
// every shard.  Returns the seconds the sweep took; decisions = (row, shard) pairs decided, ndiff =
// those != SKIP, dirty = templates / workgroups with at least one, digest = checksum of the diff list.
double oracle_fleet_sweep(oracle_fleet* f, int nthreads, uint64_t* decisions, uint64_t* ndiff, uint64_t* dirty, uint64_t* digest) {
  if (nthreads < 1) nthreads = 1;
  std::vector<const FleetObj*> heads;
  for (const auto& kv : f->controller)
    if (kv.second.kind == NCD_KIND_TEMPLATE || kv.second.kind == NCD_KIND_WORKGROUP) heads.push_back(&kv.second);
  const size_t S = f->shards.size();
  std::atomic<uint64_t> t_dec{0}, t_diff{0}, t_dirty{0}, t_dig{0};
  const auto t0 = std::chrono::steady_clock::now();
  auto work = [&](size_t lo, size_t hi) {
    uint64_t dec = 0, nd = 0, dirt = 0, dig = 0;
    for (size_t h = lo; h < hi; ++h) {
      const FleetObj& t = *heads[h];
      const std::string tk = fkey(t.kind, t.key);
      bool any = false;
      for (size_t s = 0; s < S; ++s) {  // :790 / :720
        const FleetStore& sh = f->shards[s];
        auto sti = sh.find(tk);  // :792 / :722
        const FleetObj* st = sti == sh.end() ? nullptr : &sti->second;
        uint32_t code = !st ? (uint32_t)NCD_CREATE : (deep_equal(st->value, t.value) ? (uint32_t)NCD_SKIP : (uint32_t)NCD_UPDATE_DATA);  // :795 / :725
        ++dec;
        if (code) {
          ++nd;
          any = true;
          dig += entry_digest(tk, 0, (uint32_t)s, code);
        }
        if (t.kind != NCD_KIND_TEMPLATE) continue;
        for (size_t j = 0; j < t.deps.size(); ++j) {  // :505 / :567
          const uint32_t dk = j < t.nsecrets ? NCD_KIND_SECRET : NCD_KIND_CONFIGMAP;
          const std::string k = fkey(dk, t.deps[j]);
          auto cdi = f->controller.find(k);  // :507 / :569
          if (cdi == f->controller.end()) {
            code = NCD_ERR_MISSING_REF;  // :509-514 / :571-576
          } else {
            auto sdi = sh.find(k);  // :516 / :578
            if (sdi == sh.end()) {
              code = NCD_CREATE;  // :518-520 / :580-582
            } else {
              const uint32_t own = fleet_ownership(sdi->second, st);  // :529 / :591
              if (own == NCD_ERR_ROGUE) code = NCD_ERR_ROGUE;
              else code = (deep_equal(cdi->second.value, sdi->second.value) ? 0u : (uint32_t)NCD_UPDATE_DATA) | own;  // :539 / :600
            }
          }
          ++dec;
          if (code) {
            ++nd;
            any = true;
            dig += entry_digest(tk, (uint32_t)j + 1, (uint32_t)s, code);
          }
        }
      }
      if (any) ++dirt;
    }
    t_dec += dec;
    t_diff += nd;
    t_dirty += dirt;
    t_dig += dig;
  };
  std::vector<std::thread> th;
  for (int t = 0; t < nthreads; ++t) th.emplace_back(work, heads.size() * t / nthreads, heads.size() * (t + 1) / nthreads);
  for (auto& x : th) x.join();
  const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  if (decisions) *decisions = t_dec.load();
  if (ndiff) *ndiff = t_diff.load();
  if (dirty) *dirty = t_dirty.load();
  if (digest) *digest = t_dig.load();
  return sec;
}

}  // extern "C"
