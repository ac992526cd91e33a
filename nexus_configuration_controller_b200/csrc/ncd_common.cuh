// ncd_common.cuh — internal types shared by the translation units of libncdrift.so.
// Nothing here is part of the C ABI (include/ncdrift.h is).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/ncdrift.h"

// Phase ranges for profilers (nsys, ncu --nvtx): hash / join / compare / classify / exchange / sync.
// Compiled out unless the library is built with -DNCD_NVTX (NCD_NVCC_EXTRA=-DNCD_NVTX python -m
// nexus_configuration_controller_b200.build); nvtx3 is header only.
#ifdef NCD_NVTX
#include <nvtx3/nvToolsExt.h>
struct NcdRange {
  explicit NcdRange(const char* name) { nvtxRangePushA(name); }
  ~NcdRange() { nvtxRangePop(); }
};
#define NCD_RANGE(var, name) NcdRange var(name)
#define NCD_RANGE_PUSH(name) nvtxRangePushA(name)
#define NCD_RANGE_POP() nvtxRangePop()
#else
#define NCD_RANGE(var, name) \
  do {                       \
  } while (0)
#define NCD_RANGE_PUSH(name) \
  do {                       \
  } while (0)
#define NCD_RANGE_POP() \
  do {                  \
  } while (0)
#endif

#define NCD_NOTFOUND 0xFFFFFFFFu
#define NCD_NO_OFF 0xFFFFFFFFFFFFFFFFull

// Data verdict of one (controller object, shard) pair, the input of the decision codes.
enum : uint8_t { DV_EQUAL = 0, DV_DIFF = 1, DV_NOTFOUND = 2 };

// Tile of the compare / hash kernels: a controller record is cut into tiles of this many
// bytes; one warp owns one tile (32 lanes x 16 B x 8 vectors).
constexpr uint32_t NCD_TILE_BYTES = 4096;
constexpr uint32_t NCD_TILE_VECS = NCD_TILE_BYTES / 16;

// Compare tiles: the same 4 KB cut, except that a short remainder (< 2 KB) is merged into the
// tile before it, so no warp is sent off for a few hundred bytes: a record of 8704 bytes is
// tiles of 4096 + 4608, not 4096 + 4096 + 512.  Tiles of a multi-tile record are 2049..6144 bytes.
__host__ __device__ __forceinline__ uint32_t ncd_ctiles_of(uint32_t len) {
  return len <= 6144u ? (len ? 1u : 0u) : (uint32_t)(((uint64_t)len + 2047u) >> 12);
}

// Everything a probe needs about one object, in one 64-byte line: the join reads the table
// slot, then this, and nothing else unless the key is longer than NCD_KEY_INLINE bytes.
// The first 32-byte sector holds everything a scan needs once the join is cached (record
// offset / length and the ownership tag); the second sector is the inline key for probes.
constexpr uint32_t NCD_KEY_INLINE = 32;
// Ownership tag of an object, interned at commit: what isMissingOwnership (controller.go:484-502)
// needs, in one word.  Template: the id of its own metadata.uid.  Any other object: its single
// ownerReference packed as (Kind == NexusAlgorithmTemplate) << 31 | uid id, NCD_TAG_ROGUE when it
// has none, NCD_TAG_MANY when it has several (the decision kernel then walks the list).
#define NCD_TAG_ROGUE 0xFFFFFFFFu
#define NCD_TAG_MANY 0xFFFFFFFEu
struct __align__(16) ObjMeta {
  uint64_t keyhash;
  uint64_t rec_off;
  uint32_t rec_len;
  uint32_t key_off;   // in aux (whole key)
  uint16_t key_len;
  uint8_t kind;
  uint8_t pad;
  uint32_t tag;
  uint8_t key[NCD_KEY_INLINE];  // first bytes of the key, zero padded
};
static_assert(sizeof(ObjMeta) == 64, "ObjMeta must be one 64-byte line");
#define NCD_TABLE_EMPTY 0xFFFFFFFFFFFFFFFFull

// Device-visible description of a committed arena (all pointers are device pointers).
struct DevArena {
  const uint8_t* payload;
  const uint64_t* rec_off;
  const uint32_t* rec_len;
  const uint8_t* kind;
  const uint32_t* key_off;
  const uint32_t* key_len;
  const uint32_t* uid_off;
  const uint32_t* uid_len;
  const uint32_t* own_start;
  const uint8_t* own_is_template;
  const uint32_t* own_uid_off;
  const uint32_t* own_uid_len;
  const uint32_t* dep_start;
  const uint8_t* dep_kind;
  const uint32_t* dep_key_off;
  const uint32_t* dep_key_len;
  const uint8_t* aux;
  // index built by commit
  const ObjMeta* meta;      // [nobj]
  const uint64_t* table;    // open addressing; slot = (keyhash >> 32) << 32 | index, all ones = empty
  uint32_t table_mask;
  uint32_t nobj;
  // UID interning built by commit: equal ids <=> byte-equal UID strings within this arena
  const uint32_t* uid_id;      // [nobj] id of the object's own metadata.uid
  const uint32_t* own_packed;  // [nowner] (is_template << 31) | id of the owner's UID
  // join cache (shard side): the controller-object -> shard-object map of the previous scan,
  // valid while neither arena's keys changed; exactly one of the two is set by ncd_scan
  const uint32_t* join_cache;  // [controller nobj] read instead of probing
  uint32_t* join_cache_out;    // [controller nobj] filled by this scan's probes
  // and what the join reads from the found object's line — {rec_off lo, rec_off hi, rec_len, tag} —
  // cached the same way, so that a steady-state scan streams 20 bytes per pair instead of touching
  // one random 32-byte sector of the shard's object lines per pair.  Kept current by
  // k_pull_records through the reverse map (shard object -> controller object).
  const uint4* join_line;    // [controller nobj] read instead of the object line (only with join_cache)
  uint4* join_line_out;      // [controller nobj] filled by this scan
  uint32_t* join_rev_out;    // [shard nobj] filled by this scan
  // digests, filled by the hash kernels
  const uint64_t* h64;      // [nobj]
  const uint32_t* sha;      // [nobj][8]
  const uint64_t* xxh;      // [nobj] XXH64
};

constexpr int NCD_EV_RING = 64;

struct ncd_ctx {
  int device;
  int sm_count;
  int compare_variant;      // -1 = pick per scan, 0 = per-shard loop, 1/2/4 = grouped with CH vectors per lane
  int compare_ctas_per_sm;  // grid = sm_count * this
  int fuse_small;           // 1: small records in direct mode take k_scan_small (join + compare in one kernel)
  int ss_lpo;               // lanes per object in k_scan_small (0: from the mean record length)
  int line_cache;           // 1: the join caches the found objects' record offset / length / tag per pair
  cudaStream_t stream;
  cudaEvent_t ev[8];                        // ev[5], ev[6]: ncd_hash_arena
  cudaEvent_t evr[NCD_EV_RING][5];          // per-scan phase events, ring indexed by scan_seq
  ncd_scan_stats stats_ring[NCD_EV_RING];   // host-known part of each scan's stats
  uint64_t scan_seq;                        // scans enqueued so far
  // scratch of the scan, grown on demand and kept
  void* scratch;
  size_t scratch_bytes;
  DevArena* d_shards;  // device array of shard descriptors
  DevArena* h_shards_last;  // what it holds (host copy), to skip the upload when nothing changed
  uint32_t h_shards_last_n;
  void* d_jobs;        // device array of hash jobs (ncd_hash.cu)
  int d_jobs_cap;
  int d_shards_cap;
  uint64_t* d_counters;  // [8] device counters (diff count, collisions, ...)
  uint64_t* h_counters;  // pinned mirror
  // pending scan
  bool scan_pending;
  struct ncd_diff* pending;
  ncd_scan_stats stats;
  // pinned host buffers of destroyed diff lists, reused by the next ones (pinning memory is slow
  // and synchronises the device: a steady stream of scans must not do it per scan)
  void* hpool_ptr[4];
  size_t hpool_cap[4];
  // ncd_scan_host: the diff list buffer (device, header + entries) and its pinned host mirror,
  // kept between scans and grown only when a scan finds more than they hold
  cudaStream_t copy_stream;  // ncd_arenas_commit: uploads run here while the index of an earlier arena is built
  void* list_dev;
  void* list_host;
  uint64_t list_cap;  // entries
};

struct ncd_arena {
  ncd_ctx* ctx;
  uint32_t flags;
  // capacities
  uint64_t payload_cap, nobj_cap, aux_cap, nowner_cap, ndep_cap;
  // committed counts
  uint64_t nobj, payload_bytes, aux_bytes, nowner, ndep;
  uint64_t len_sum;  // sum of rec_len: the arena's share of the algorithmic bytes
  uint32_t len_max;  // longest record
  uint64_t last_rehashed;  // records the last re-commit had to hash again (the others kept their digests)
  bool committed;
  ncd_arena_view h;  // pinned host arrays (HOST_PINNED only)
  ncd_arena_view d;  // device arrays
  // index (device)
  ObjMeta* meta;
  uint64_t* table;
  uint32_t table_size;
  uint32_t dup_keys;
  // controller-side lists (device), built by commit
  uint64_t nrows;
  uint32_t* row_start;  // [nobj+1] first row of object i
  uint32_t* row_obj;    // [nrows] owning template / workgroup
  uint32_t* row_dep;    // [nrows] NCD_NOTFOUND = the object itself, else dep entry index
  uint32_t* row_cidx;   // [nrows] controller index of the compared object (NCD_NOTFOUND = missing ref)
  uint64_t ntiles;
  uint32_t* tile_obj;   // [ntiles]
  uint32_t* tile_chunk; // [ntiles]
  uint32_t* tile_start; // [nobj+1]
  uint64_t nctiles;     // compare tiles (short remainders merged, ncd_ctiles_of)
  uint32_t* ctile_obj;
  uint32_t* ctile_chunk;
  uint32_t* ctile_start;
  uint64_t tile_cap, ctile_cap;  // allocated entries of the tile lists
  // digests (device), lazily allocated
  uint64_t* h64;
  uint32_t* sha;
  bool h64_valid, sha_valid, xxh_valid;
  uint64_t* xxh;
  // index generation (bumped by every commit / recommit) and the join cache kept for one controller arena
  uint64_t generation;
  uint32_t* uid_id;      // [nobj]
  uint32_t* own_packed;  // [nowner]
  uint32_t* jc_idx;
  const ncd_arena* jc_ctrl;
  uint64_t jc_ctrl_gen, jc_gen, jc_n;
  uint4* jc_line;     // [jc_n] see DevArena::join_line
  uint32_t* jc_rev;   // [jc_rev_n] controller object of each object of this arena, NCD_NOTFOUND if none
  uint64_t jc_rev_n, jc_line_n;
  bool jc_line_ok;    // jc_line / jc_rev describe the records as they are now (for the pairs jc_idx describes)
};

struct ncd_diff {
  ncd_ctx* ctx;
  uint64_t rows, nshards, n;
  ncd_entry* d_entries;  // device, capacity rows*nshards lazily bounded by n
  ncd_entry* h_entries;  // pinned, filled on first view
  size_t h_cap;          // bytes of h_entries
  uint8_t* d_codes;      // device [rows*nshards]
  bool h_valid;
  uint32_t hdr[4];       // staging for ncd_diff_pack
};

// ------------------------------------------------------------------ errors
void ncd_set_error(const char* fmt, ...);
#define NCD_CUDA(call)                                                              \
  do {                                                                              \
    cudaError_t _e = (call);                                                        \
    if (_e != cudaSuccess) {                                                        \
      ncd_set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e), __FILE__, \
                    __LINE__);                                                      \
      return NCD_E_CUDA;                                                            \
    }                                                                               \
  } while (0)
#define NCD_TRY(call)          \
  do {                         \
    int _s = (call);           \
    if (_s != NCD_OK) return _s; \
  } while (0)

// ------------------------------------------------------------------ shared host/device helpers
__host__ __device__ __forceinline__ uint64_t ncd_fmix64(uint64_t x) {
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdull;
  x ^= x >> 33;
  x *= 0xc4ceb9fe1a85ec53ull;
  x ^= x >> 33;
  return x;
}
__host__ __device__ __forceinline__ uint64_t ncd_splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
// FNV-1a over the kind byte and the key, then a finaliser.
__host__ __device__ __forceinline__ uint64_t ncd_keyhash_impl(uint32_t kind, const uint8_t* key,
                                                             uint32_t len) {
  uint64_t h = 0xcbf29ce484222325ull;
  h = (h ^ (uint64_t)(kind & 0xff)) * 0x100000001b3ull;
  for (uint32_t i = 0; i < len; ++i) h = (h ^ key[i]) * 0x100000001b3ull;
  return ncd_fmix64(h);
}

static inline uint64_t ncd_align_up(uint64_t x, uint64_t a) { return (x + a - 1) / a * a; }

// ------------------------------------------------------------------ internal entry points
// ncd_index.cu
int ncd_build_index(ncd_arena* a);
int ncd_rebuild_tiles(ncd_arena* a);
int ncd_exclusive_scan_u32(ncd_ctx* ctx, const uint32_t* in, uint32_t* out, uint64_t n,
                           uint32_t* total_out_dev);
int ncd_exclusive_scan_u64(ncd_ctx* ctx, const uint32_t* in, uint64_t* out, uint64_t n,
                           uint64_t* total_out_dev);
void ncd_fill_dev_arena(const ncd_arena* a, DevArena* out);
int ncd_ctx_scratch(ncd_ctx* ctx, size_t bytes, void** out);
// ncd_compare_tma.cu
int ncd_launch_compare_tma(ncd_ctx* ctx, const DevArena& dctrl, uint32_t S, const uint32_t* tile_obj,
                           const uint32_t* tile_chunk, uint64_t ntiles, const uint64_t* sh_off, uint8_t* dv,
                           uint32_t* counters, uint32_t count_collisions);
// ncd_hash.cu
int ncd_hash_launch(ncd_ctx* ctx, ncd_arena* a, uint32_t mode, uint32_t* launches);
int ncd_hash_records_many(ncd_ctx* ctx, int count, ncd_arena* const* arenas, const uint32_t* const* d_idx,
                          const uint64_t* n);  // ncd_hash.cu: named records only
int ncd_hash_launch_many(ncd_ctx* ctx, ncd_arena* const* arenas, int count, uint32_t mode, uint32_t* launches);
int ncd_digests_carry(ncd_arena* a, const uint64_t* old_off, const uint32_t* old_len, uint64_t old_n, uint64_t dirty_from,
                      bool had_h64, bool had_sha, bool had_xxh, uint64_t* rehashed);

// ------------------------------------------------------------------ device helpers
#ifdef __CUDACC__
__device__ __forceinline__ bool ncd_bytes_equal(const uint8_t* a, const uint8_t* b, uint32_t n) {
  for (uint32_t i = 0; i < n; ++i)
    if (a[i] != b[i]) return false;
  return true;
}

// lister.Get on the GPU: exact ("kind", "namespace/name") lookup in an arena's table.
// Linear probing.  A slot carries the upper half of the key hash, so almost every foreign slot
// is rejected without touching the object's line; a candidate matches when hash, kind, key
// length and every key byte match (inline part first, the rest from aux).
__device__ __forceinline__ uint32_t ncd_probe(const DevArena& t, uint32_t kind, const uint8_t* key,
                                              uint32_t key_len, uint64_t hash) {
  if (t.nobj == 0) return NCD_NOTFOUND;
  uint32_t slot = (uint32_t)hash & t.table_mask;
  const uint32_t hi = (uint32_t)(hash >> 32);
  for (;;) {
    const uint64_t e = t.table[slot];
    if (e == NCD_TABLE_EMPTY) return NCD_NOTFOUND;
    if ((uint32_t)(e >> 32) == hi) {
      const uint32_t idx = (uint32_t)e;
      const ObjMeta& m = t.meta[idx];
      if (m.keyhash == hash && m.kind == kind && m.key_len == key_len) {
        const uint32_t inl = key_len < NCD_KEY_INLINE ? key_len : NCD_KEY_INLINE;
        if (ncd_bytes_equal(m.key, key, inl) &&
            (key_len <= NCD_KEY_INLINE ||
             ncd_bytes_equal(t.aux + m.key_off + NCD_KEY_INLINE, key + NCD_KEY_INLINE, key_len - NCD_KEY_INLINE)))
          return idx;
      }
    }
    slot = (slot + 1) & t.table_mask;
  }
}

// Same lookup when the key is itself an object's line (the join: controller object -> shard).
__device__ __forceinline__ uint32_t ncd_probe_meta(const DevArena& t, const ObjMeta& q, const uint8_t* q_aux) {
  if (t.nobj == 0) return NCD_NOTFOUND;
  uint32_t slot = (uint32_t)q.keyhash & t.table_mask;
  const uint32_t hi = (uint32_t)(q.keyhash >> 32);
  const uint4* qw = reinterpret_cast<const uint4*>(&q);
  for (;;) {
    const uint64_t e = t.table[slot];
    if (e == NCD_TABLE_EMPTY) return NCD_NOTFOUND;
    if ((uint32_t)(e >> 32) == hi) {
      const uint32_t idx = (uint32_t)e;
      const uint4* mw = reinterpret_cast<const uint4*>(&t.meta[idx]);
      // words 0-1: keyhash; word 6: key_len | kind | pad; word 7: tag (not part of the key);
      // words 8-15: inline key (zero padded)
      const uint4 m0 = mw[0], m1 = mw[1], m2 = mw[2], m3 = mw[3];
      const uint4 q0 = qw[0], q1 = qw[1], q2 = qw[2], q3 = qw[3];
      const uint32_t diff = (m0.x ^ q0.x) | (m0.y ^ q0.y) | ((m1.z ^ q1.z) & 0x00FFFFFFu) |
                            (m2.x ^ q2.x) | (m2.y ^ q2.y) | (m2.z ^ q2.z) | (m2.w ^ q2.w) |
                            (m3.x ^ q3.x) | (m3.y ^ q3.y) | (m3.z ^ q3.z) | (m3.w ^ q3.w);
      if (diff == 0) {
        const uint32_t klen = q.key_len;
        if (klen <= NCD_KEY_INLINE ||
            ncd_bytes_equal(t.aux + m1.y + NCD_KEY_INLINE, q_aux + q.key_off + NCD_KEY_INLINE, klen - NCD_KEY_INLINE))
          return idx;
      }
    }
    slot = (slot + 1) & t.table_mask;
  }
}

__device__ __forceinline__ uint4 ncd_ldg_stream(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
#endif
