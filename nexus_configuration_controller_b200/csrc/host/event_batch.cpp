// event_batch.cpp — bench / test utility: informer events synthesised from arena-shaped arrays.
//
// The drop-in's input is informer events carrying live Go objects.  There is no Go here, so the
// bench and the tests need another source of events at scale: this turns objects laid out like an
// arena (the synthetic generator's output, downloaded to host memory) into ncdh_event records
// whose values are Go values in transport form wrapping the raw record bytes:
//   Secret      map[string][]byte{"data": <record>}      (corev1.Secret.Data)
//   ConfigMap   map[string]string{"data": <record>}      (corev1.ConfigMap.Data)
//   Template /  struct{ Payload []byte; Tag string }     (a stand-in for .Spec)
//   Workgroup
// Names, UIDs, ownerReferences and dependent names are copied from the arrays.  Not part of the
// drop-in surface (like ncd_synth_fill in ncdrift.h).
#include <stdint.h>
#include <string.h>

#include <memory>
#include <string>
#include <vector>

#include "../../../include/ncdhost.h"
#include "../../../include/ncdrift.h"
#include "cluster_image.hpp"

struct ncdh_event_batch {
  std::vector<ncdh_event> events;
  std::vector<uint8_t> wire;           // every value, back to back
  std::vector<char> strings;           // NUL-terminated names, uids, owner kinds
  std::vector<const char*> ptr_lists;  // owner_kinds / owner_uids / secret_names / configmap_names arrays
};

namespace {

inline uint64_t mix(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
inline uint64_t key_id(uint32_t kind, const uint8_t* key, uint32_t len) {  // the same for every cluster's copy of an object
  uint64_t h = 0xcbf29ce484222325ull ^ kind;
  for (uint32_t i = 0; i < len; ++i) h = (h ^ key[i]) * 0x100000001b3ull;
  return h;
}
inline void put_u32(uint8_t*& p, uint32_t v) {
  memcpy(p, &v, 4);
  p += 4;
}
constexpr uint8_t T_STRING = 3, T_BYTES = 4, T_MAP = 8, T_STRUCT = 12;

uint64_t wire_size(uint32_t kind, uint32_t len) {
  if (kind == NCD_KIND_SECRET || kind == NCD_KIND_CONFIGMAP) return 1 + 4 + (1 + 4 + 4) + (1 + 4) + len;
  return 1 + 4 + (1 + 4) + len + (1 + 4 + 4);
}
void write_wire(uint8_t* p, uint32_t kind, const uint8_t* rec, uint32_t len, uint64_t salt, uint64_t id) {
  uint8_t* data;
  if (kind == NCD_KIND_SECRET || kind == NCD_KIND_CONFIGMAP) {
    *p++ = T_MAP;
    put_u32(p, 1);
    *p++ = T_STRING;
    put_u32(p, 4);
    memcpy(p, "data", 4);
    p += 4;
    *p++ = kind == NCD_KIND_SECRET ? T_BYTES : T_STRING;
    put_u32(p, len);
    data = p;
    memcpy(p, rec, len);
  } else {
    *p++ = T_STRUCT;
    put_u32(p, 2);
    *p++ = T_BYTES;
    put_u32(p, len);
    data = p;
    memcpy(p, rec, len);
    p += len;
    *p++ = T_STRING;
    put_u32(p, 4);
    memcpy(p, "spec", 4);
  }
  if (salt && len) {  // the change this event carries: one byte, same length
    const uint64_t h = mix(salt * 0x100000001B3ull ^ id);
    data[h % len] ^= (uint8_t)(1 + (h >> 32) % 255);
  }
}

}  // namespace

extern "C" {

int ncdh_batch_from_arrays(const ncd_arena_view* meta, const ncd_arena_view* values, int cluster, const uint32_t* obj_idx,
                           uint64_t n, uint64_t salt, int threads, ncdh_event_batch** out) {
  if (!meta || !out) return -1;
  if (!values) values = meta;
  if (!obj_idx) n = meta->nobj;
  std::unique_ptr<ncdh_event_batch> b(new ncdh_event_batch);
  b->events.resize(n);
  // pass 1 (serial, cheap): sizes and offsets
  std::vector<uint64_t> woff(n + 1, 0), soff(n + 1, 0), poff(n + 1, 0);
  for (uint64_t k = 0; k < n; ++k) {
    const uint32_t i = obj_idx ? obj_idx[k] : (uint32_t)k;
    if (i >= meta->nobj || i >= values->nobj) return -1;
    const uint32_t nown = meta->own_start[i + 1] - meta->own_start[i], ndep = meta->dep_start[i + 1] - meta->dep_start[i];
    uint64_t s = (uint64_t)meta->key_len[i] + 2 + meta->uid_len[i] + 1;  // "ns\0name\0", uid
    for (uint32_t o = meta->own_start[i]; o < meta->own_start[i + 1]; ++o) s += meta->own_uid_len[o] + 1;
    for (uint32_t d = meta->dep_start[i]; d < meta->dep_start[i + 1]; ++d) s += meta->dep_key_len[d] + 1;
    woff[k + 1] = woff[k] + wire_size(meta->kind[i], values->rec_len[i]);
    soff[k + 1] = soff[k] + s;
    poff[k + 1] = poff[k] + 2 * nown + ndep;
  }
  static const char kTemplate[] = "NexusAlgorithmTemplate", kOther[] = "Other";
  b->wire.resize(woff[n] + 16);
  b->strings.resize(soff[n] + 16);
  b->ptr_lists.resize(poff[n] + 4);
  // pass 2 (parallel): the copies
  ncdh::parallel_for(n, threads <= 0 ? 8 : threads, 256, [&](size_t k) {
    const uint32_t i = obj_idx ? obj_idx[k] : (uint32_t)k;
    ncdh_event& e = b->events[k];
    memset(&e, 0, sizeof(e));
    char* s = b->strings.data() + soff[k];
    const char** pl = b->ptr_lists.data() + poff[k];
    auto str = [&](const uint8_t* p, uint32_t len) {
      char* at = s;
      memcpy(s, p, len);
      s[len] = 0;
      s += len + 1;
      return at;
    };
    const uint8_t* key = meta->aux + meta->key_off[i];
    const uint32_t klen = meta->key_len[i];
    uint32_t slash = 0;
    while (slash < klen && key[slash] != '/') ++slash;
    e.cluster = cluster;
    e.kind = meta->kind[i];
    e.op = NCDH_EV_PUT;
    e.ns = str(key, slash);
    e.name = str(key + (slash < klen ? slash + 1 : klen), slash < klen ? klen - slash - 1 : 0);
    e.uid = str(meta->aux + meta->uid_off[i], meta->uid_len[i]);
    const uint32_t o0 = meta->own_start[i], o1 = meta->own_start[i + 1];
    e.nowners = (int32_t)(o1 - o0);
    e.owner_kinds = pl;
    e.owner_uids = pl + (o1 - o0);
    for (uint32_t o = o0; o < o1; ++o) {
      pl[o - o0] = meta->own_is_template[o] ? kTemplate : kOther;
      pl[(o1 - o0) + (o - o0)] = str(meta->aux + meta->own_uid_off[o], meta->own_uid_len[o]);
    }
    pl += 2 * (o1 - o0);
    const uint32_t d0 = meta->dep_start[i], d1 = meta->dep_start[i + 1];
    uint32_t nsec = 0;
    for (uint32_t d = d0; d < d1; ++d) {  // secrets first, then configmaps (arena order)
      const uint8_t* dk = meta->aux + meta->dep_key_off[d];
      const uint32_t dl = meta->dep_key_len[d];
      uint32_t ds = 0;
      while (ds < dl && dk[ds] != '/') ++ds;
      pl[d - d0] = str(dk + (ds < dl ? ds + 1 : dl), ds < dl ? dl - ds - 1 : 0);
      if (meta->dep_kind[d] == NCD_KIND_SECRET) ++nsec;
    }
    e.secret_names = pl;
    e.nsecrets = (int32_t)nsec;
    e.configmap_names = pl + nsec;
    e.nconfigmaps = (int32_t)(d1 - d0 - nsec);
    uint8_t* w = b->wire.data() + woff[k];
    write_wire(w, e.kind, values->payload + values->rec_off[i], values->rec_len[i], salt, key_id(e.kind, key, klen));
    e.value = w;
    e.value_len = woff[k + 1] - woff[k];
  });
  b->wire.resize(woff[n]);
  *out = b.release();
  return 0;
}

const ncdh_event* ncdh_batch_events(const ncdh_event_batch* b, uint64_t* n) {
  if (n) *n = b ? b->events.size() : 0;
  return b ? b->events.data() : nullptr;
}
uint64_t ncdh_batch_value_bytes(const ncdh_event_batch* b) { return b ? b->wire.size() : 0; }
void ncdh_batch_free(ncdh_event_batch* b) { delete b; }

}  // extern "C"
