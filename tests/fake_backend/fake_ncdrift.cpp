// fake_ncdrift.cpp — TEST DOUBLE of the part of libncdrift.so the host layer calls, for machines
// without a GPU.  It exists so that the host layer's own logic (packing, incremental uploads,
// snapshots, the freshness rule, locking) can be tested on CPU and under ThreadSanitizer; it is
// never built into, linked with or loaded by the product.  Arenas keep a "device" copy apart from
// the pinned host arrays, filled only by commit / recommit / sync exactly as the real library
// moves bytes — a host layer that forgets an upload scans stale data here too.  Decisions come
// from the CPU oracle (oracle/ncd_oracle.cpp, compiled into the same test library).
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <thread>
#include <vector>

#include "../../include/ncdrift.h"

extern "C" uint64_t oracle_rows(const ncd_arena_view* ctrl);
extern "C" uint64_t oracle_scan(const ncd_arena_view* ctrl, const ncd_arena_view* const* shards, int nshards,
                                uint8_t* codes_out, ncd_entry* entries_out, int nthreads, uint64_t* bytes_out);
extern "C" uint64_t oracle_unadopted(const ncd_arena_view* c, uint32_t* out, uint64_t cap);

static thread_local char g_err[512] = "";
static int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

struct ncd_ctx {
  std::vector<ncd_entry> list;
  ncd_scan_stats stats;
  int scan_sleep_us;
};
struct ncd_arena {
  ncd_ctx* ctx;
  uint64_t payload_cap, nobj_cap, aux_cap, nowner_cap, ndep_cap;
  ncd_arena_view h, d;
  bool committed;
};

namespace {
struct Field {
  size_t offset, elem;
  int cap_kind;  // 0 payload, 1 nobj, 2 nobj+1, 3 nowner, 4 ndep, 5 aux
};
#define F(member, cap) {offsetof(ncd_arena_view, member), sizeof(*((ncd_arena_view*)0)->member), cap}
const Field kFields[] = {F(payload, 0),     F(rec_off, 1),     F(rec_len, 1),   F(kind, 1),     F(key_off, 1),     F(key_len, 1),
                         F(uid_off, 1),     F(uid_len, 1),     F(own_start, 2), F(own_is_template, 3), F(own_uid_off, 3),
                         F(own_uid_len, 3), F(dep_start, 2),   F(dep_kind, 4),  F(dep_key_off, 4), F(dep_key_len, 4), F(aux, 5)};
#undef F
void** fptr(ncd_arena_view* v, const Field& f) { return (void**)((char*)v + f.offset); }
uint64_t fcount(const Field& f, uint64_t payload, uint64_t nobj, uint64_t aux, uint64_t nowner, uint64_t ndep) {
  switch (f.cap_kind) {
    case 0: return (payload + 15) / 16 * 16;
    case 1: return nobj;
    case 2: return nobj + 1;
    case 3: return nowner;
    case 4: return ndep;
    default: return aux;
  }
}
int upload(ncd_arena* a, uint64_t nobj, uint64_t payload_bytes, uint64_t aux_bytes, uint64_t nowner, uint64_t ndep,
           uint64_t payload_from) {
  if (nobj > a->nobj_cap || payload_bytes > a->payload_cap || aux_bytes > a->aux_cap || nowner > a->nowner_cap || ndep > a->ndep_cap)
    return fail(NCD_E_CAPACITY, "fake: contents exceed the arena's capacity");
  for (const Field& f : kFields) {
    uint64_t bytes = fcount(f, payload_bytes, nobj, aux_bytes, nowner, ndep) * f.elem, from = 0;
    if (f.cap_kind == 0) {
      from = payload_from;
      bytes = payload_bytes - payload_from;
    }
    if (bytes) memcpy((char*)*fptr(&a->d, f) + from, (const char*)*fptr(&a->h, f) + from, bytes);
  }
  a->d.nobj = nobj;
  a->d.payload_bytes = payload_bytes;
  a->d.aux_bytes = aux_bytes;
  a->d.nowner = nowner;
  a->d.ndep = ndep;
  for (uint64_t i = 0; i < nobj; ++i)
    if ((a->d.rec_off[i] & 15) || a->d.rec_off[i] + a->d.rec_len[i] > payload_bytes)
      return fail(NCD_E_INVALID, "fake: record %llu outside the payload or not 16-byte aligned", (unsigned long long)i);
  a->committed = true;
  return NCD_OK;
}
}  // namespace

extern "C" {

const char* ncd_last_error(void) { return g_err; }

int ncd_init(int device, ncd_ctx** out) {
  if (!out || device != 0) return fail(NCD_E_INVALID, "fake: only device 0");
  *out = new ncd_ctx();
  const char* s = getenv("NCD_FAKE_SCAN_SLEEP_US");
  (*out)->scan_sleep_us = s ? atoi(s) : 0;
  return NCD_OK;
}
void ncd_destroy(ncd_ctx* ctx) { delete ctx; }

int ncd_arena_create(ncd_ctx* ctx, uint64_t payload_cap, uint64_t nobj_cap, uint64_t aux_cap, uint64_t nowner_cap,
                     uint64_t ndep_cap, uint32_t flags, ncd_arena** out) {
  if (!ctx || !out || flags != NCD_ARENA_HOST_PINNED) return fail(NCD_E_INVALID, "fake: bad argument");
  ncd_arena* a = new ncd_arena();
  memset(a, 0, sizeof(*a));
  a->ctx = ctx;
  a->payload_cap = (payload_cap + 15) / 16 * 16;
  a->nobj_cap = nobj_cap;
  a->aux_cap = aux_cap;
  a->nowner_cap = nowner_cap;
  a->ndep_cap = ndep_cap;
  for (const Field& f : kFields) {
    uint64_t bytes = fcount(f, payload_cap, nobj_cap, aux_cap, nowner_cap, ndep_cap) * f.elem;
    if (!bytes) bytes = 16;
    *fptr(&a->h, f) = malloc(bytes);
    *fptr(&a->d, f) = malloc(bytes);
    memset(*fptr(&a->h, f), 0xA5, bytes);  // what the host layer does not write is garbage, as pinned memory is
    memset(*fptr(&a->d, f), 0x5A, bytes);
  }
  *out = a;
  return NCD_OK;
}
void ncd_arena_destroy(ncd_arena* a) {
  if (!a) return;
  for (const Field& f : kFields) {
    free(*fptr(&a->h, f));
    free(*fptr(&a->d, f));
  }
  delete a;
}
int ncd_arena_host_view(ncd_arena* a, ncd_arena_view* out) {
  if (!a || !out) return fail(NCD_E_INVALID, "fake: null");
  *out = a->h;
  out->nobj = a->nobj_cap;
  out->payload_bytes = a->payload_cap;
  out->aux_bytes = a->aux_cap;
  out->nowner = a->nowner_cap;
  out->ndep = a->ndep_cap;
  return NCD_OK;
}
int ncd_arena_commit(ncd_arena* a, uint64_t nobj, uint64_t payload_bytes, uint64_t aux_bytes, uint64_t nowner, uint64_t ndep) {
  if (!a) return fail(NCD_E_INVALID, "fake: null");
  return upload(a, nobj, payload_bytes, aux_bytes, nowner, ndep, 0);
}
int ncd_arenas_commit(int count, ncd_arena* const* arenas, const uint64_t* nobj, const uint64_t* payload_bytes,
                      const uint64_t* aux_bytes, const uint64_t* nowner, const uint64_t* ndep) {
  for (int k = 0; k < count; ++k) {
    const int rc = ncd_arena_commit(arenas[k], nobj[k], payload_bytes[k], aux_bytes[k], nowner[k], ndep[k]);
    if (rc != NCD_OK) return rc;
  }
  return NCD_OK;
}
int ncd_arena_recommit(ncd_arena* a, uint64_t nobj, uint64_t payload_bytes, uint64_t aux_bytes, uint64_t nowner, uint64_t ndep,
                       uint64_t payload_dirty_from) {
  if (!a || !a->committed) return fail(NCD_E_STATE, "fake: recommit before commit");
  if (payload_dirty_from > payload_bytes || (payload_dirty_from & 15)) return fail(NCD_E_INVALID, "fake: bad payload_dirty_from");
  return upload(a, nobj, payload_bytes, aux_bytes, nowner, ndep, payload_dirty_from);
}
// Test hook: the next `n` value syncs fail the way a CUDA error would, nothing copied.
static std::atomic<int> g_fail_syncs{0};
void ncd_fake_fail_next_syncs(int n) { g_fail_syncs.store(n); }

int ncd_arenas_sync_records(int count, ncd_arena* const* arenas, const uint64_t* n, const uint32_t* const* obj_idx,
                            const uint64_t* new_payload_bytes) {
  if (g_fail_syncs.load() > 0) {
    g_fail_syncs.fetch_sub(1);
    return fail(NCD_E_CUDA, "fake: injected sync failure");
  }
  for (int k = 0; k < count; ++k) {
    ncd_arena* a = arenas[k];
    if (!a || !a->committed) return fail(NCD_E_STATE, "fake: sync before commit");
    if (new_payload_bytes[k] > a->payload_cap || (new_payload_bytes[k] & 15)) return fail(NCD_E_CAPACITY, "fake: payload beyond capacity");
    for (uint64_t q = 0; q < n[k]; ++q) {
      const uint32_t i = obj_idx[k][q];
      if (i >= a->d.nobj) return fail(NCD_E_INVALID, "fake: record index out of range");
      const uint64_t off = a->h.rec_off[i], len = a->h.rec_len[i], padded = (len + 15) / 16 * 16;
      if ((off & 15) || off + padded > new_payload_bytes[k]) return fail(NCD_E_INVALID, "fake: record refused");
      a->d.rec_off[i] = off;
      a->d.rec_len[i] = (uint32_t)len;
      memcpy(a->d.payload + off, a->h.payload + off, padded);
    }
    a->d.payload_bytes = new_payload_bytes[k];
  }
  return NCD_OK;
}

int ncd_arena_compact(ncd_arena* a, uint64_t* new_payload_bytes) {
  if (!a || !a->committed) return fail(NCD_E_STATE, "fake: compact of an uncommitted arena");
  const uint64_t n = a->d.nobj;
  std::vector<uint32_t> order(n);
  for (uint64_t i = 0; i < n; ++i) order[i] = (uint32_t)i;
  std::sort(order.begin(), order.end(), [&](uint32_t x, uint32_t y) {
    return a->h.rec_off[x] != a->h.rec_off[y] ? a->h.rec_off[x] < a->h.rec_off[y] : a->h.rec_len[x] < a->h.rec_len[y];
  });
  uint64_t cur = 0;
  std::vector<uint8_t> tmp(a->payload_cap + 16);
  std::vector<uint64_t> noff(n);
  for (uint32_t i : order) {  // the "device" copy is gathered from ITS offsets, as the real library does, the slots
                              // sized by the HOST's lengths (a record rewritten in place may not be synced yet)
    const uint64_t padded = ((uint64_t)a->h.rec_len[i] + 15) / 16 * 16;
    if (a->d.rec_off[i] != a->h.rec_off[i]) return fail(NCD_E_STATE, "fake: compact of an arena with a record moved and not synced");
    memcpy(tmp.data() + cur, a->d.payload + a->d.rec_off[i], padded);
    noff[i] = cur;
    cur += padded;
  }
  memcpy(a->d.payload, tmp.data(), cur);
  for (uint32_t i : order) {
    const uint64_t padded = ((uint64_t)a->h.rec_len[i] + 15) / 16 * 16;
    memmove(a->h.payload + noff[i], a->h.payload + a->h.rec_off[i], padded);
  }
  for (uint64_t i = 0; i < n; ++i) a->h.rec_off[i] = a->d.rec_off[i] = noff[i];
  a->d.payload_bytes = cur;
  if (new_payload_bytes) *new_payload_bytes = cur;
  return NCD_OK;
}

// Test hook: the next `n` scans fail the way a CUDA error would (after the uploads went through).
static int g_fail_scans = 0;
void ncd_fake_fail_next_scans(int n) { g_fail_scans = n; }

int ncd_scan_host(ncd_ctx* ctx, const ncd_arena* ctrl, const ncd_arena* const* shards, int nshards, uint32_t mode,
                  const ncd_entry** entries, uint64_t* n) {
  (void)mode;
  if (g_fail_scans > 0) {
    --g_fail_scans;
    return fail(NCD_E_CUDA, "fake: injected scan failure");
  }
  if (!ctx || !ctrl || !entries || !n || !ctrl->committed) return fail(NCD_E_STATE, "fake: scan of an uncommitted arena");
  std::vector<const ncd_arena_view*> sv;
  for (int s = 0; s < nshards; ++s) {
    if (!shards[s] || !shards[s]->committed) return fail(NCD_E_STATE, "fake: shard %d not committed", s);
    sv.push_back(&shards[s]->d);
  }
  const uint64_t rows = oracle_rows(&ctrl->d);
  std::vector<uint8_t> codes(rows * (uint64_t)(nshards ? nshards : 1) + 1);
  ctx->list.resize(rows * (uint64_t)(nshards ? nshards : 1) + 1);
  const uint64_t found = oracle_scan(&ctrl->d, sv.data(), nshards, codes.data(), ctx->list.data(), 1, nullptr);
  if (ctx->scan_sleep_us) std::this_thread::sleep_for(std::chrono::microseconds(ctx->scan_sleep_us));
  memset(&ctx->stats, 0, sizeof(ctx->stats));
  ctx->stats.rows = rows;
  ctx->stats.pairs = rows * (uint64_t)nshards;
  ctx->stats.ndiff = found;
  *entries = ctx->list.data();
  *n = found;
  return NCD_OK;
}
int ncd_scan_last_stats(ncd_ctx* ctx, ncd_scan_stats* out) {
  *out = ctx->stats;
  return NCD_OK;
}
int ncd_find_unadopted(ncd_ctx* ctx, const ncd_arena* ctrl, ncd_adoption* out, uint64_t cap, uint64_t* n) {
  if (!ctx || !ctrl || !n || !ctrl->committed) return fail(NCD_E_STATE, "fake: unadopted of an uncommitted arena");
  *n = oracle_unadopted(&ctrl->d, (uint32_t*)out, cap);
  return NCD_OK;
}

}  // extern "C"
