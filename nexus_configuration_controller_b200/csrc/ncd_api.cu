// ncd_api.cu — context and arena entry points of the C ABI (include/ncdrift.h).
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "ncd_common.cuh"

// ------------------------------------------------------------------ errors
static thread_local char g_err[512] = "";

void ncd_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* ncd_last_error(void) { return g_err; }
extern "C" const char* ncd_version(void) { return "ncdrift 0.1 (sm_100a)"; }

extern "C" uint64_t ncd_keyhash(uint32_t kind, const void* key, uint32_t key_len) {
  return ncd_keyhash_impl(kind, (const uint8_t*)key, key_len);
}

// ------------------------------------------------------------------ context
extern "C" int ncd_init(int device, ncd_ctx** out) {
  if (!out) {
    ncd_set_error("ncd_init: null out");
    return NCD_E_INVALID;
  }
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    ncd_set_error("ncd_init: no CUDA device (%s); this library has no CPU path",
                  e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    return NCD_E_CUDA;
  }
  if (device < 0 || device >= ndev) {
    ncd_set_error("ncd_init: device %d out of range (0..%d)", device, ndev - 1);
    return NCD_E_INVALID;
  }
  NCD_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  NCD_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) {
    ncd_set_error("ncd_init: device %d is sm_%d%d; the kernels are built for sm_100a only", device,
                  prop.major, prop.minor);
    return NCD_E_CUDA;
  }
  ncd_ctx* ctx = (ncd_ctx*)calloc(1, sizeof(ncd_ctx));
  if (!ctx) return NCD_E_NOMEM;
  ctx->device = device;
  ctx->sm_count = prop.multiProcessorCount;
  ctx->compare_variant = -1;  // -1: chosen per scan from the mean tile fill
  ctx->compare_ctas_per_sm = 8;
  ctx->fuse_small = 1;
  if (const char* v = getenv("NCD_FUSE_SMALL")) ctx->fuse_small = atoi(v) != 0;  // experiment knob: 0 = k_join + k_compare_small
  ctx->line_cache = 1;
  if (const char* v = getenv("NCD_LINE_CACHE")) ctx->line_cache = atoi(v) != 0;  // experiment knob
  ctx->ss_lpo = 0;
  if (const char* v = getenv("NCD_SS_LPO")) {  // experiment knob: lanes per object in k_scan_small
    const int l = atoi(v);
    if (l == 4 || l == 8 || l == 16 || l == 32) ctx->ss_lpo = l;
  }
  if (const char* v = getenv("NCD_COMPARE_VARIANT")) {
    // experiment knob: only the kernels that exist may be named (ncd_scan_stats.compare_variant)
    const int k = atoi(v);
    if (k != -1 && k != 0 && k != 1 && k != 2 && k != 4 && k != 8 && k != 16) {
      ncd_set_error("ncd_init: NCD_COMPARE_VARIANT=%s is not one of -1, 0, 1, 2, 4, 8, 16", v);
      free(ctx);
      return NCD_E_INVALID;
    }
    ctx->compare_variant = k;
  }
  if (const char* v = getenv("NCD_COMPARE_CTAS")) ctx->compare_ctas_per_sm = atoi(v) > 0 ? atoi(v) : 8;
  NCD_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
  for (int i = 0; i < 8; ++i) NCD_CUDA(cudaEventCreate(&ctx->ev[i]));
  for (int r = 0; r < NCD_EV_RING; ++r)
    for (int i = 0; i < 5; ++i) NCD_CUDA(cudaEventCreate(&ctx->evr[r][i]));
  NCD_CUDA(cudaMalloc((void**)&ctx->d_counters, 8 * sizeof(uint64_t)));
  NCD_CUDA(cudaMallocHost((void**)&ctx->h_counters, NCD_EV_RING * 8 * sizeof(uint64_t)));
  // keep stream-ordered allocations cached: the per-scan diff buffers come from this pool
  cudaMemPool_t pool;
  if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
    uint64_t thr = UINT64_MAX;
    cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
  }
  *out = ctx;
  return NCD_OK;
}

extern "C" void ncd_destroy(ncd_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  if (ctx->pending) ncd_diff_destroy(ctx->pending);
  if (ctx->scratch) cudaFree(ctx->scratch);
  if (ctx->d_shards) cudaFree(ctx->d_shards);
  free(ctx->h_shards_last);
  if (ctx->d_jobs) cudaFree(ctx->d_jobs);
  if (ctx->d_counters) cudaFree(ctx->d_counters);
  if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
  if (ctx->list_dev) cudaFree(ctx->list_dev);
  if (ctx->list_host) cudaFreeHost(ctx->list_host);
  for (int i = 0; i < 4; ++i)
    if (ctx->hpool_ptr[i]) cudaFreeHost(ctx->hpool_ptr[i]);
  if (ctx->h_counters) cudaFreeHost(ctx->h_counters);
  for (int i = 0; i < 8; ++i) cudaEventDestroy(ctx->ev[i]);
  for (int r = 0; r < NCD_EV_RING; ++r)
    for (int i = 0; i < 5; ++i) cudaEventDestroy(ctx->evr[r][i]);
  cudaStreamDestroy(ctx->stream);
  free(ctx);
}

extern "C" void* ncd_ctx_stream(ncd_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
extern "C" int ncd_ctx_device(ncd_ctx* ctx) { return ctx ? ctx->device : -1; }
extern "C" int ncd_ctx_sync(ncd_ctx* ctx) {
  if (!ctx) {
    ncd_set_error("ncd_ctx_sync: null ctx");
    return NCD_E_INVALID;
  }
  NCD_CUDA(cudaStreamSynchronize(ctx->stream));
  return NCD_OK;
}

int ncd_ctx_scratch(ncd_ctx* ctx, size_t bytes, void** out) {
  if (bytes > ctx->scratch_bytes) {
    NCD_CUDA(cudaStreamSynchronize(ctx->stream));
    if (ctx->scratch) cudaFree(ctx->scratch);
    ctx->scratch = nullptr;
    ctx->scratch_bytes = 0;
    size_t want = bytes + bytes / 8;
    cudaError_t e = cudaMalloc(&ctx->scratch, want);
    if (e != cudaSuccess) {
      ncd_set_error("scan scratch of %zu bytes: %s", want, cudaGetErrorString(e));
      return NCD_E_NOMEM;
    }
    ctx->scratch_bytes = want;
  }
  *out = ctx->scratch;
  return NCD_OK;
}

// ------------------------------------------------------------------ arenas
namespace {
struct Field {
  size_t offset;      // offset of the pointer inside ncd_arena_view
  size_t elem;        // element size
  int cap_kind;       // 0 payload, 1 nobj, 2 nobj+1, 3 nowner, 4 ndep, 5 aux
};
#define F(member, cap) {offsetof(ncd_arena_view, member), sizeof(*((ncd_arena_view*)0)->member), cap}
const Field kFields[] = {
    F(payload, 0),     F(rec_off, 1),     F(rec_len, 1),     F(kind, 1),        F(key_off, 1),
    F(key_len, 1),     F(uid_off, 1),     F(uid_len, 1),     F(own_start, 2),   F(own_is_template, 3),
    F(own_uid_off, 3), F(own_uid_len, 3), F(dep_start, 2),   F(dep_kind, 4),    F(dep_key_off, 4),
    F(dep_key_len, 4), F(aux, 5)};
#undef F
constexpr int kNumFields = sizeof(kFields) / sizeof(kFields[0]);

inline void** field_ptr(ncd_arena_view* v, const Field& f) { return (void**)((char*)v + f.offset); }
inline uint64_t field_count(const Field& f, uint64_t payload, uint64_t nobj, uint64_t aux,
                            uint64_t nowner, uint64_t ndep) {
  switch (f.cap_kind) {
    case 0: return ncd_align_up(payload, 16);
    case 1: return nobj;
    case 2: return nobj + 1;
    case 3: return nowner;
    case 4: return ndep;
    default: return aux;
  }
}
}  // namespace

extern "C" int ncd_arena_create(ncd_ctx* ctx, uint64_t payload_cap, uint64_t nobj_cap, uint64_t aux_cap,
                                uint64_t nowner_cap, uint64_t ndep_cap, uint32_t flags, ncd_arena** out) {
  if (!ctx || !out) {
    ncd_set_error("ncd_arena_create: null argument");
    return NCD_E_INVALID;
  }
  *out = nullptr;
  if (nobj_cap >= 0xFFFFFFFFull || aux_cap >= 0xFFFFFFFFull || nowner_cap >= 0xFFFFFFFFull ||
      ndep_cap >= 0xFFFFFFFFull) {
    ncd_set_error("ncd_arena_create: object / aux / owner / dependent capacities are 32-bit");
    return NCD_E_CAPACITY;
  }
  if (flags > NCD_ARENA_DEVICE_ONLY) {
    ncd_set_error("ncd_arena_create: unknown flags %u", flags);
    return NCD_E_INVALID;
  }
  NCD_CUDA(cudaSetDevice(ctx->device));
  ncd_arena* a = (ncd_arena*)calloc(1, sizeof(ncd_arena));
  if (!a) return NCD_E_NOMEM;
  a->ctx = ctx;
  a->flags = flags;
  a->payload_cap = ncd_align_up(payload_cap, 16);
  a->nobj_cap = nobj_cap;
  a->aux_cap = aux_cap;
  a->nowner_cap = nowner_cap;
  a->ndep_cap = ndep_cap;
  for (int i = 0; i < kNumFields; ++i) {
    const Field& f = kFields[i];
    uint64_t bytes = field_count(f, payload_cap, nobj_cap, aux_cap, nowner_cap, ndep_cap) * f.elem;
    if (bytes == 0) bytes = 16;
    cudaError_t e = cudaMalloc(field_ptr(&a->d, f), bytes);
    if (e == cudaSuccess && flags == NCD_ARENA_HOST_PINNED) e = cudaMallocHost(field_ptr(&a->h, f), bytes);
    if (e != cudaSuccess) {
      ncd_set_error("ncd_arena_create: allocating %llu bytes: %s", (unsigned long long)bytes,
                    cudaGetErrorString(e));
      ncd_arena_destroy(a);
      return NCD_E_NOMEM;
    }
  }
  *out = a;
  return NCD_OK;
}

extern "C" void ncd_arena_destroy(ncd_arena* a) {
  if (!a) return;
  cudaSetDevice(a->ctx->device);
  cudaStreamSynchronize(a->ctx->stream);
  for (int i = 0; i < kNumFields; ++i) {
    void* d = *field_ptr(&a->d, kFields[i]);
    void* h = *field_ptr(&a->h, kFields[i]);
    if (d) cudaFree(d);
    if (h) cudaFreeHost(h);
  }
  void* idx[] = {a->meta,    a->table,     a->row_start,  a->row_obj, a->row_dep, a->row_cidx,
                 a->tile_obj, a->tile_chunk, a->tile_start, a->h64,     a->sha,     a->xxh,     a->jc_idx,  a->uid_id,  a->own_packed, a->jc_line, a->jc_rev,
                 a->ctile_obj, a->ctile_chunk, a->ctile_start};
  for (void* p : idx)
    if (p) cudaFree(p);
  free(a);
}

extern "C" int ncd_arena_host_view(ncd_arena* a, ncd_arena_view* out) {
  if (!a || !out) {
    ncd_set_error("ncd_arena_host_view: null argument");
    return NCD_E_INVALID;
  }
  if (a->flags != NCD_ARENA_HOST_PINNED) {
    ncd_set_error("ncd_arena_host_view: arena is device-only");
    return NCD_E_STATE;
  }
  *out = a->h;
  out->nobj = a->nobj_cap;
  out->payload_bytes = a->payload_cap;
  out->aux_bytes = a->aux_cap;
  out->nowner = a->nowner_cap;
  out->ndep = a->ndep_cap;
  return NCD_OK;
}

static int check_counts(ncd_arena* a, uint64_t nobj, uint64_t payload_bytes, uint64_t aux_bytes,
                        uint64_t nowner, uint64_t ndep) {
  if (nobj > a->nobj_cap || payload_bytes > a->payload_cap || aux_bytes > a->aux_cap ||
      nowner > a->nowner_cap || ndep > a->ndep_cap) {
    ncd_set_error("ncd_arena_commit: contents (%llu objects, %llu payload bytes, %llu aux bytes, "
                  "%llu owners, %llu dependents) exceed the arena's capacity",
                  (unsigned long long)nobj, (unsigned long long)payload_bytes,
                  (unsigned long long)aux_bytes, (unsigned long long)nowner, (unsigned long long)ndep);
    return NCD_E_CAPACITY;
  }
  return NCD_OK;
}

// Shared by ncd_arena_commit and the synthetic generator (which fills the device arrays itself).
int ncd_arena_commit_device(ncd_arena* a, uint64_t nobj, uint64_t payload_bytes, uint64_t aux_bytes,
                            uint64_t nowner, uint64_t ndep) {
  NCD_TRY(check_counts(a, nobj, payload_bytes, aux_bytes, nowner, ndep));
  a->committed = false;
  a->nobj = nobj;
  a->payload_bytes = payload_bytes;
  a->aux_bytes = aux_bytes;
  a->nowner = nowner;
  a->ndep = ndep;
  NCD_TRY(ncd_build_index(a));
  a->committed = true;
  return NCD_OK;
}

extern "C" int ncd_arena_commit(ncd_arena* a, uint64_t nobj, uint64_t payload_bytes, uint64_t aux_bytes,
                                uint64_t nowner, uint64_t ndep) {
  if (!a) {
    ncd_set_error("ncd_arena_commit: null arena");
    return NCD_E_INVALID;
  }
  if (a->flags != NCD_ARENA_HOST_PINNED) {
    ncd_set_error("ncd_arena_commit: device-only arenas are committed by ncd_synth_fill");
    return NCD_E_STATE;
  }
  NCD_TRY(check_counts(a, nobj, payload_bytes, aux_bytes, nowner, ndep));
  NCD_CUDA(cudaSetDevice(a->ctx->device));
  cudaStream_t st = a->ctx->stream;
  for (int i = 0; i < kNumFields; ++i) {
    const Field& f = kFields[i];
    uint64_t bytes = field_count(f, payload_bytes, nobj, aux_bytes, nowner, ndep) * f.elem;
    if (f.cap_kind == 0) bytes = payload_bytes;  // the padding of the last record is part of it
    if (bytes)
      NCD_CUDA(cudaMemcpyAsync(*field_ptr(&a->d, f), *field_ptr(&a->h, f), bytes, cudaMemcpyHostToDevice, st));
  }
  return ncd_arena_commit_device(a, nobj, payload_bytes, aux_bytes, nowner, ndep);
}

// Several arenas of one context committed in one call (a controller without resident state uploads
// the controller cluster and every shard before each sweep; the host layer does it on its first
// scan).  Every upload is enqueued up front on a copy stream, so the PCIe link never idles while
// the index of an earlier arena is being built on the main stream (ncd_arena_commit alone
// alternates copy -> index build -> copy ...).
extern "C" int ncd_arenas_commit(int count, ncd_arena* const* arenas, const uint64_t* nobj, const uint64_t* payload_bytes,
                                 const uint64_t* aux_bytes, const uint64_t* nowner, const uint64_t* ndep) {
  if (count < 0 || (count > 0 && (!arenas || !nobj || !payload_bytes || !aux_bytes || !nowner || !ndep))) {
    ncd_set_error("ncd_arenas_commit: null argument");
    return NCD_E_INVALID;
  }
  if (count == 0) return NCD_OK;
  ncd_ctx* ctx = nullptr;
  for (int k = 0; k < count; ++k) {
    ncd_arena* a = arenas[k];
    if (!a || a->flags != NCD_ARENA_HOST_PINNED) {
      ncd_set_error("ncd_arenas_commit: arena %d has no pinned host arrays", k);
      return NCD_E_STATE;
    }
    if (ctx && a->ctx != ctx) {
      ncd_set_error("ncd_arenas_commit: arenas of different contexts");
      return NCD_E_INVALID;
    }
    for (int q = 0; q < k; ++q)
      if (arenas[q] == a) {
        ncd_set_error("ncd_arenas_commit: arena %d is listed twice", k);
        return NCD_E_INVALID;
      }
    ctx = a->ctx;
    NCD_TRY(check_counts(a, nobj[k], payload_bytes[k], aux_bytes[k], nowner[k], ndep[k]));
  }
  NCD_CUDA(cudaSetDevice(ctx->device));
  if (!ctx->copy_stream) NCD_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
  NCD_CUDA(cudaStreamSynchronize(ctx->stream));  // nothing on the main stream still reads what is about to be overwritten
  std::vector<cudaEvent_t> done((size_t)count);
  for (int k = 0; k < count; ++k) {
    ncd_arena* a = arenas[k];
    a->committed = false;
    for (int i = 0; i < kNumFields; ++i) {
      const Field& f = kFields[i];
      uint64_t bytes = field_count(f, payload_bytes[k], nobj[k], aux_bytes[k], nowner[k], ndep[k]) * f.elem;
      if (f.cap_kind == 0) bytes = payload_bytes[k];
      if (bytes)
        NCD_CUDA(cudaMemcpyAsync(*field_ptr(&a->d, f), *field_ptr(&a->h, f), bytes, cudaMemcpyHostToDevice, ctx->copy_stream));
    }
    NCD_CUDA(cudaEventCreateWithFlags(&done[k], cudaEventDisableTiming));
    NCD_CUDA(cudaEventRecord(done[k], ctx->copy_stream));
  }
  int rc = NCD_OK;
  for (int k = 0; k < count; ++k) {
    if (rc == NCD_OK) {
      cudaStreamWaitEvent(ctx->stream, done[k], 0);
      rc = ncd_arena_commit_device(arenas[k], nobj[k], payload_bytes[k], aux_bytes[k], nowner[k], ndep[k]);
      if (rc != NCD_OK) cudaStreamSynchronize(ctx->copy_stream);  // the later uploads finish before anybody reuses the host arrays
    }
    cudaEventDestroy(done[k]);
  }
  return rc;
}

extern "C" int ncd_arena_recommit(ncd_arena* a, uint64_t nobj, uint64_t payload_bytes, uint64_t aux_bytes,
                                  uint64_t nowner, uint64_t ndep, uint64_t payload_dirty_from) {
  if (!a) {
    ncd_set_error("ncd_arena_recommit: null arena");
    return NCD_E_INVALID;
  }
  if (a->flags != NCD_ARENA_HOST_PINNED) {
    ncd_set_error("ncd_arena_recommit: the arena has no pinned host arrays");
    return NCD_E_STATE;
  }
  if (!a->committed) {
    ncd_set_error("ncd_arena_recommit: nothing is resident yet, use ncd_arena_commit");
    return NCD_E_STATE;
  }
  if (payload_dirty_from > payload_bytes || (payload_dirty_from & 15)) {
    ncd_set_error("ncd_arena_recommit: payload_dirty_from must be a multiple of 16 within the payload");
    return NCD_E_INVALID;
  }
  NCD_TRY(check_counts(a, nobj, payload_bytes, aux_bytes, nowner, ndep));
  NCD_CUDA(cudaSetDevice(a->ctx->device));
  cudaStream_t st = a->ctx->stream;
  // Digests the arena holds are carried over to the records that survive (ncd_digests_carry): keep
  // the old offsets and lengths, the only thing that identifies a record once objects are renumbered.
  const bool had_h64 = a->h64_valid, had_sha = a->sha_valid, had_xxh = a->xxh_valid;
  const uint64_t old_n = a->nobj;
  uint64_t* old_off = nullptr;
  uint32_t* old_len = nullptr;
  if ((had_h64 || had_sha || had_xxh) && old_n) {
    NCD_CUDA(cudaMallocAsync((void**)&old_off, old_n * 8, st));
    NCD_CUDA(cudaMallocAsync((void**)&old_len, old_n * 4, st));
    NCD_CUDA(cudaMemcpyAsync(old_off, a->d.rec_off, old_n * 8, cudaMemcpyDeviceToDevice, st));
    NCD_CUDA(cudaMemcpyAsync(old_len, a->d.rec_len, old_n * 4, cudaMemcpyDeviceToDevice, st));
  }
  for (int i = 0; i < kNumFields; ++i) {
    const Field& f = kFields[i];
    uint64_t bytes = field_count(f, payload_bytes, nobj, aux_bytes, nowner, ndep) * f.elem;
    uint64_t from = 0;
    if (f.cap_kind == 0) {  // payload: only what was appended
      from = payload_dirty_from;
      bytes = payload_bytes - payload_dirty_from;
    }
    if (bytes)
      NCD_CUDA(cudaMemcpyAsync((char*)*field_ptr(&a->d, f) + from, (const char*)*field_ptr(&a->h, f) + from, bytes,
                               cudaMemcpyHostToDevice, st));
  }
  int rc = ncd_arena_commit_device(a, nobj, payload_bytes, aux_bytes, nowner, ndep);
  if (rc == NCD_OK && old_off)
    rc = ncd_digests_carry(a, old_off, old_len, old_n, payload_dirty_from, had_h64, had_sha, had_xxh, &a->last_rehashed);
  if (old_off) {
    cudaFreeAsync(old_off, st);
    cudaFreeAsync(old_len, st);
  }
  return rc;
}

extern "C" int ncd_arena_counts(ncd_arena* a, ncd_arena_view* c) {
  if (!a || !c) {
    ncd_set_error("ncd_arena_counts: null argument");
    return NCD_E_INVALID;
  }
  memset(c, 0, sizeof(*c));
  c->nobj = a->nobj;
  c->payload_bytes = a->payload_bytes;
  c->aux_bytes = a->aux_bytes;
  c->nowner = a->nowner;
  c->ndep = a->ndep;
  return NCD_OK;
}

extern "C" int ncd_arena_download(ncd_arena* a, ncd_arena_view* dst) {
  if (!a || !dst) {
    ncd_set_error("ncd_arena_download: null argument");
    return NCD_E_INVALID;
  }
  if (!a->committed) {
    ncd_set_error("ncd_arena_download: arena not committed");
    return NCD_E_STATE;
  }
  NCD_CUDA(cudaSetDevice(a->ctx->device));
  cudaStream_t st = a->ctx->stream;
  for (int i = 0; i < kNumFields; ++i) {
    const Field& f = kFields[i];
    void* to = *field_ptr(dst, f);
    if (!to) continue;
    uint64_t bytes = field_count(f, a->payload_bytes, a->nobj, a->aux_bytes, a->nowner, a->ndep) * f.elem;
    if (f.cap_kind == 0) bytes = a->payload_bytes;
    if (bytes) NCD_CUDA(cudaMemcpyAsync(to, *field_ptr(&a->d, f), bytes, cudaMemcpyDeviceToHost, st));
  }
  NCD_CUDA(cudaStreamSynchronize(st));
  dst->nobj = a->nobj;
  dst->payload_bytes = a->payload_bytes;
  dst->aux_bytes = a->aux_bytes;
  dst->nowner = a->nowner;
  dst->ndep = a->ndep;
  return NCD_OK;
}

extern "C" int ncd_arena_rows(ncd_arena* a, uint64_t* rows) {
  if (!a || !rows) {
    ncd_set_error("ncd_arena_rows: null argument");
    return NCD_E_INVALID;
  }
  if (!a->committed) {
    ncd_set_error("ncd_arena_rows: arena not committed");
    return NCD_E_STATE;
  }
  *rows = a->nrows;
  return NCD_OK;
}

// ------------------------------------------------------------------ snapshots
// One arena as a file, for offline diffing: what a cluster's caches held at some moment, compared
// later (or elsewhere) against another snapshot or a live arena.  Layout, little-endian:
//   header   "NCDARENA", u32 version (1), u32 field count (17), u64 nobj, payload_bytes,
//            aux_bytes, nowner, ndep
//   fields   in ncd_arena_view order; each: u64 byte count, the bytes, zero padding to 16
//   trailer  u64 checksum of every field byte (word-wise multiply-xor), "NCDAEND\0"
// Streams through a bounded pinned buffer, so a 10 GB arena needs no 10 GB of host memory.
// No reference counterpart (the reference keeps no record of what its caches held).
namespace {
constexpr size_t kSnapChunk = 32u << 20;
struct SnapHeader {
  char magic[8];
  uint32_t version, nfields;
  uint64_t nobj, payload_bytes, aux_bytes, nowner, ndep;
};
inline uint64_t snap_mix(uint64_t h, const void* p, size_t n) {
  const unsigned char* b = (const unsigned char*)p;
  size_t i = 0;
  for (; i + 8 <= n; i += 8) {
    uint64_t w;
    memcpy(&w, b + i, 8);
    h = (h ^ w) * 0x9E3779B185EBCA87ull;
    h ^= h >> 29;
  }
  uint64_t w = 0;
  if (i < n) {
    memcpy(&w, b + i, n - i);
    h = (h ^ w ^ ((uint64_t)(n - i) << 56)) * 0x9E3779B185EBCA87ull;
    h ^= h >> 29;
  }
  return h;
}
struct SnapFile {
  FILE* f = nullptr;
  void* stage = nullptr;
  ~SnapFile() {
    if (f) fclose(f);
    if (stage) cudaFreeHost(stage);
  }
};
}  // namespace

extern "C" int ncd_arena_save(ncd_arena* a, const char* path) {
  if (!a || !path) {
    ncd_set_error("ncd_arena_save: null argument");
    return NCD_E_INVALID;
  }
  if (!a->committed) {
    ncd_set_error("ncd_arena_save: arena not committed");
    return NCD_E_STATE;
  }
  NCD_CUDA(cudaSetDevice(a->ctx->device));
  cudaStream_t st = a->ctx->stream;
  SnapFile sf;
  sf.f = fopen(path, "wb");
  if (!sf.f) {
    ncd_set_error("ncd_arena_save: cannot open %s for writing", path);
    return NCD_E_INVALID;
  }
  NCD_CUDA(cudaMallocHost(&sf.stage, kSnapChunk));
  SnapHeader h;
  memset(&h, 0, sizeof(h));
  memcpy(h.magic, "NCDARENA", 8);
  h.version = 1;
  h.nfields = kNumFields;
  h.nobj = a->nobj;
  h.payload_bytes = a->payload_bytes;
  h.aux_bytes = a->aux_bytes;
  h.nowner = a->nowner;
  h.ndep = a->ndep;
  bool ok = fwrite(&h, sizeof(h), 1, sf.f) == 1;
  uint64_t sum = 0x6e6364615f736e61ull;
  const char zeros[16] = {0};
  for (int i = 0; i < kNumFields && ok; ++i) {
    const Field& f = kFields[i];
    uint64_t bytes = field_count(f, a->payload_bytes, a->nobj, a->aux_bytes, a->nowner, a->ndep) * f.elem;
    if (f.cap_kind == 0) bytes = a->payload_bytes;
    ok = fwrite(&bytes, 8, 1, sf.f) == 1;
    const char* src = (const char*)*field_ptr(&a->d, f);
    for (uint64_t off = 0; off < bytes && ok; off += kSnapChunk) {
      const size_t n = (size_t)(bytes - off < kSnapChunk ? bytes - off : kSnapChunk);
      NCD_CUDA(cudaMemcpyAsync(sf.stage, src + off, n, cudaMemcpyDeviceToHost, st));
      NCD_CUDA(cudaStreamSynchronize(st));
      sum = snap_mix(sum, sf.stage, n);
      ok = fwrite(sf.stage, 1, n, sf.f) == n;
    }
    const size_t pad = (size_t)((16 - bytes % 16) % 16);
    if (ok && pad) ok = fwrite(zeros, 1, pad, sf.f) == pad;
  }
  if (ok) ok = fwrite(&sum, 8, 1, sf.f) == 1 && fwrite("NCDAEND\0", 8, 1, sf.f) == 1;
  if (ok) ok = fflush(sf.f) == 0;
  if (!ok) {
    ncd_set_error("ncd_arena_save: write to %s failed", path);
    return NCD_E_INVALID;
  }
  return NCD_OK;
}

extern "C" int ncd_arena_load(ncd_ctx* ctx, const char* path, uint32_t flags, uint64_t payload_slack, ncd_arena** out) {
  if (!ctx || !path || !out) {
    ncd_set_error("ncd_arena_load: null argument");
    return NCD_E_INVALID;
  }
  *out = nullptr;
  NCD_CUDA(cudaSetDevice(ctx->device));
  SnapFile sf;
  sf.f = fopen(path, "rb");
  if (!sf.f) {
    ncd_set_error("ncd_arena_load: cannot open %s", path);
    return NCD_E_INVALID;
  }
  SnapHeader h;
  if (fread(&h, sizeof(h), 1, sf.f) != 1 || memcmp(h.magic, "NCDARENA", 8) != 0 || h.version != 1 ||
      h.nfields != (uint32_t)kNumFields) {
    ncd_set_error("ncd_arena_load: %s is not an arena snapshot of version 1", path);
    return NCD_E_INVALID;
  }
  ncd_arena* a = nullptr;
  NCD_TRY(ncd_arena_create(ctx, h.payload_bytes + payload_slack, h.nobj, h.aux_bytes, h.nowner, h.ndep, flags, &a));
  struct Guard {
    ncd_arena* a;
    ~Guard() {
      if (a) ncd_arena_destroy(a);
    }
  } guard{a};
  NCD_CUDA(cudaMallocHost(&sf.stage, kSnapChunk));
  cudaStream_t st = ctx->stream;
  uint64_t sum = 0x6e6364615f736e61ull;
  for (int i = 0; i < kNumFields; ++i) {
    const Field& f = kFields[i];
    uint64_t want = field_count(f, h.payload_bytes, h.nobj, h.aux_bytes, h.nowner, h.ndep) * f.elem;
    if (f.cap_kind == 0) want = h.payload_bytes;
    uint64_t bytes = 0;
    if (fread(&bytes, 8, 1, sf.f) != 1 || bytes != want) {
      ncd_set_error("ncd_arena_load: %s: field %d holds %llu bytes, the header implies %llu", path, i,
                    (unsigned long long)bytes, (unsigned long long)want);
      return NCD_E_INVALID;
    }
    char* dev = (char*)*field_ptr(&a->d, f);
    char* host = flags == NCD_ARENA_HOST_PINNED ? (char*)*field_ptr(&a->h, f) : nullptr;
    for (uint64_t off = 0; off < bytes; off += kSnapChunk) {
      const size_t n = (size_t)(bytes - off < kSnapChunk ? bytes - off : kSnapChunk);
      void* into = host ? (void*)(host + off) : sf.stage;
      if (fread(into, 1, n, sf.f) != n) {
        ncd_set_error("ncd_arena_load: %s is truncated", path);
        return NCD_E_INVALID;
      }
      sum = snap_mix(sum, into, n);
      NCD_CUDA(cudaMemcpyAsync(dev + off, into, n, cudaMemcpyHostToDevice, st));
      if (!host) NCD_CUDA(cudaStreamSynchronize(st));  // the staging buffer is reused
    }
    const size_t pad = (size_t)((16 - bytes % 16) % 16);
    char skip[16];
    if (pad && fread(skip, 1, pad, sf.f) != pad) {
      ncd_set_error("ncd_arena_load: %s is truncated", path);
      return NCD_E_INVALID;
    }
  }
  uint64_t stored = 0;
  char endm[8];
  if (fread(&stored, 8, 1, sf.f) != 1 || fread(endm, 8, 1, sf.f) != 1 || memcmp(endm, "NCDAEND\0", 8) != 0) {
    ncd_set_error("ncd_arena_load: %s has no trailer (truncated?)", path);
    return NCD_E_INVALID;
  }
  if (stored != sum) {
    ncd_set_error("ncd_arena_load: %s is corrupt (checksum %016llx, expected %016llx)", path, (unsigned long long)sum,
                  (unsigned long long)stored);
    return NCD_E_INVALID;
  }
  NCD_CUDA(cudaStreamSynchronize(st));
  // the arrays are on the device already: validate them and build the index there
  NCD_TRY(ncd_arena_commit_device(a, h.nobj, h.payload_bytes, h.aux_bytes, h.nowner, h.ndep));
  guard.a = nullptr;
  *out = a;
  return NCD_OK;
}
