// ncd_patch.cu — incremental arenas: keep every cluster's arena
// resident in HBM and move only the records an informer event rewrote.
//
// The reference re-walks every object on every 30 s resync (main.go:70-71; UpdateFunc without a
// ResourceVersion filter, controller.go:288-290, 301-303), which costs it nothing because its
// caches already sit in host RAM.  A GPU path that re-uploaded everything would be PCIe bound
// (about 50 GB/s) and slower than the host cores; with resident arenas only the changed records
// cross PCIe.  The packer rewrites a changed object's record in the pinned host arrays (in place
// when it fits, else appended behind the payload end) and names the object here; one kernel
// pulls exactly those records over PCIe from the pinned (device-mapped) host memory.
#include <algorithm>
#include <vector>

#include "ncd_common.cuh"

// One CTA per rewritten record.  h_* are the pinned host arrays (unified addressing makes them
// readable from the device); the record lands at the same offset in the device payload.
__global__ void __launch_bounds__(128)
k_pull_records(uint32_t* __restrict__ idx, const uint8_t* h_payload, const uint64_t* h_rec_off,
               const uint32_t* h_rec_len, uint8_t* __restrict__ d_payload, uint64_t* __restrict__ d_rec_off,
               uint32_t* __restrict__ d_rec_len, ObjMeta* __restrict__ meta, uint32_t nobj,
               uint64_t payload_limit, uint32_t* __restrict__ bad /*[0] refused, [1] lengths changed*/,
               const uint32_t* __restrict__ jc_rev, uint4* __restrict__ jc_line /* the arena's pair cache, or null */) {
  const uint32_t i = idx[blockIdx.x];
  if (i >= nobj) {
    if (threadIdx.x == 0) {
      atomicAdd(bad, 1u);
      idx[blockIdx.x] = 0;  // the list is read again by the re-hash: keep it inside the arena
    }
    return;
  }
  const uint64_t off = h_rec_off[i];
  const uint32_t len = h_rec_len[i];
  const uint64_t padded = ((uint64_t)len + 15) & ~15ull;
  if ((off & 15) || off + padded > payload_limit) {
    if (threadIdx.x == 0) atomicAdd(bad, 1u);
    return;
  }
  if (threadIdx.x == 0) {
    if (d_rec_len[i] != len) atomicAdd(bad + 1, 1u);  // tile lists depend on lengths only
    d_rec_off[i] = off;
    d_rec_len[i] = len;
    meta[i].rec_off = off;
    meta[i].rec_len = len;
    if (jc_rev) {  // the join's cached copy of this object's line follows (DevArena::join_line)
      const uint32_t c = jc_rev[i];
      if (c != NCD_NOTFOUND) jc_line[c] = make_uint4((uint32_t)off, (uint32_t)(off >> 32), len, meta[i].tag);
    }
  }
  const uint4* src = reinterpret_cast<const uint4*>(h_payload + off);
  uint4* dst = reinterpret_cast<uint4*>(d_payload + off);
  for (uint32_t v = threadIdx.x; v < (uint32_t)(padded >> 4); v += blockDim.x) dst[v] = src[v];
}

// Checks one arena's arguments for a sync (no CUDA calls).
static int sync_check(const ncd_arena* a, uint64_t n, const uint32_t* obj_idx, uint64_t new_payload_bytes) {
  if (!a || (!obj_idx && n)) {
    ncd_set_error("ncd_arena_sync_records: null argument");
    return NCD_E_INVALID;
  }
  if (a->flags != NCD_ARENA_HOST_PINNED) {
    ncd_set_error("ncd_arena_sync_records: the arena has no pinned host arrays to pull from");
    return NCD_E_STATE;
  }
  if (!a->committed) {
    ncd_set_error("ncd_arena_sync_records: arena not committed");
    return NCD_E_STATE;
  }
  if (new_payload_bytes > a->payload_cap || (new_payload_bytes & 15)) {
    ncd_set_error("ncd_arena_sync_records: payload of %llu bytes exceeds the arena's capacity %llu "
                  "(or is not a multiple of 16); re-create the arena and commit",
                  (unsigned long long)new_payload_bytes, (unsigned long long)a->payload_cap);
    return NCD_E_CAPACITY;
  }
  if (n >= 0xFFFFFFFFull) {
    ncd_set_error("ncd_arena_sync_records: more than 2^32 records named");
    return NCD_E_CAPACITY;
  }
  return NCD_OK;
}

// Several arenas of one context in one pass: every pull (and the re-hash of the pulled records,
// where an arena holds digests) is enqueued first, then ONE host synchronisation reads all the
// flags back.  A sweep over S shard clusters + the controller pays one round trip, not S + 1.
int ncd_arenas_sync_records(int count, ncd_arena* const* arenas, const uint64_t* n, const uint32_t* const* obj_idx,
                            const uint64_t* new_payload_bytes) {
  if (count < 0 || (count > 0 && (!arenas || !n || !obj_idx || !new_payload_bytes))) {
    ncd_set_error("ncd_arenas_sync_records: null argument");
    return NCD_E_INVALID;
  }
  uint64_t total = 0;
  ncd_ctx* ctx = nullptr;
  for (int k = 0; k < count; ++k) {
    NCD_TRY(sync_check(arenas[k], n[k], obj_idx[k], new_payload_bytes[k]));
    for (int q = 0; q < k; ++q)
      if (arenas[q] == arenas[k]) {
        ncd_set_error("ncd_arenas_sync_records: arena %d is listed twice", k);
        return NCD_E_INVALID;
      }
    if (ctx && arenas[k]->ctx != ctx) {
      ncd_set_error("ncd_arenas_sync_records: arenas of different contexts");
      return NCD_E_INVALID;
    }
    ctx = arenas[k]->ctx;
    total += n[k];
  }
  if (total == 0) {
    for (int k = 0; k < count; ++k) arenas[k]->payload_bytes = new_payload_bytes[k];
    return NCD_OK;
  }
  NCD_CUDA(cudaSetDevice(ctx->device));
  NCD_RANGE(sync_range, "ncd:sync_records");
  cudaStream_t st = ctx->stream;
  // device: [indices of arena 0 | arena 1 | ...][2 flag words per arena]
  uint32_t* d_idx = nullptr;
  NCD_CUDA(cudaMallocAsync((void**)&d_idx, (total + 2 * (uint64_t)count) * 4, st));
  uint32_t* d_flags = d_idx + total;
  NCD_CUDA(cudaMemsetAsync(d_flags, 0, 8 * (size_t)count, st));
  uint64_t at = 0;
  std::vector<const uint32_t*> d_lists((size_t)count, nullptr);
  for (int k = 0; k < count; ++k) {
    ncd_arena* a = arenas[k];
    if (!n[k]) continue;
    // pageable source: staged by the runtime before the call returns
    NCD_CUDA(cudaMemcpyAsync(d_idx + at, obj_idx[k], n[k] * 4, cudaMemcpyHostToDevice, st));
    k_pull_records<<<(unsigned)n[k], 128, 0, st>>>(d_idx + at, a->h.payload, a->h.rec_off, a->h.rec_len, a->d.payload,
                                                  a->d.rec_off, a->d.rec_len, a->meta, (uint32_t)a->nobj,
                                                  new_payload_bytes[k], d_flags + 2 * k,
                                                  a->jc_line_ok ? a->jc_rev : nullptr, a->jc_line_ok ? a->jc_line : nullptr);
    d_lists[k] = d_idx + at;
    at += n[k];
  }
  // Digests the arenas hold stay valid: only the rewritten records are hashed again, so a
  // hash-mode scan after an informer event costs the changed bytes, not the fleet's.  (A refused
  // index was replaced by the kernel above with a valid one; that arena is uncommitted below anyway.)
  const int hrc = ncd_hash_records_many(ctx, count, arenas, d_lists.data(), n);
  uint32_t* flags = (uint32_t*)calloc((size_t)count * 2, sizeof(uint32_t));
  if (!flags) {
    cudaFreeAsync(d_idx, st);
    ncd_set_error("ncd_arenas_sync_records: out of host memory");
    return NCD_E_NOMEM;
  }
  cudaError_t e = cudaMemcpyAsync(flags, d_flags, 8 * (size_t)count, cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaFreeAsync(d_idx, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) {
    free(flags);
    ncd_set_error("ncd_arenas_sync_records: %s", cudaGetErrorString(e));
    return NCD_E_CUDA;
  }
  int rc = hrc;
  for (int k = 0; k < count; ++k) {
    ncd_arena* a = arenas[k];
    a->payload_bytes = new_payload_bytes[k];
    const uint32_t bad = flags[2 * k], len_changed = flags[2 * k + 1];
    if (bad || hrc != NCD_OK) a->h64_valid = a->sha_valid = a->xxh_valid = false;
    // same-length rewrites (the common event: a value changed in place) leave the tile lists and
    // the byte count as they are
    if (len_changed || bad) {
      const int t = ncd_rebuild_tiles(a);
      if (t != NCD_OK && rc == NCD_OK) rc = t;
    }
    if (bad) {
      a->committed = false;
      if (rc == NCD_OK || rc == hrc) {
        ncd_set_error("ncd_arena_sync_records: %u of %llu records of arena %d were refused (index out of range, "
                      "record not 16-byte aligned or outside the payload); commit the arena again",
                      bad, (unsigned long long)n[k], k);
        rc = NCD_E_INVALID;
      }
    }
  }
  free(flags);
  return rc;
}

int ncd_arena_sync_records(ncd_arena* a, uint64_t n, const uint32_t* obj_idx, uint64_t new_payload_bytes) {
  if (!a) {
    ncd_set_error("ncd_arena_sync_records: null argument");
    return NCD_E_INVALID;
  }
  return ncd_arenas_sync_records(1, &a, &n, &obj_idx, &new_payload_bytes);
}

// ------------------------------------------------------------------ reading records back
// The reverse of the pull: a few named records of a committed arena gathered on the device into one
// staging buffer and copied to the host.  For checking a resident arena's contents (the full-size
// parity tests sample records of 50-110 GB arenas this way) and for offline tools; no reference
// counterpart.
__global__ void __launch_bounds__(128)
k_gather_records(const uint32_t* __restrict__ idx, const uint64_t* __restrict__ out_off, const uint8_t* __restrict__ payload,
                 const uint64_t* __restrict__ rec_off, const uint32_t* __restrict__ rec_len, uint8_t* __restrict__ out) {
  const uint32_t i = idx[blockIdx.x];
  const uint4* src = reinterpret_cast<const uint4*>(payload + rec_off[i]);
  uint4* dst = reinterpret_cast<uint4*>(out + out_off[blockIdx.x]);
  const uint32_t nvec = (rec_len[i] + 15) >> 4;
  for (uint32_t v = threadIdx.x; v < nvec; v += blockDim.x) dst[v] = src[v];
}

int ncd_arena_fetch_records(ncd_arena* a, uint64_t n, const uint32_t* obj_idx, uint8_t* out, uint64_t out_cap, uint64_t* out_off) {
  if (!a || (n && (!obj_idx || !out_off)) || (!out && out_cap)) {
    ncd_set_error("ncd_arena_fetch_records: null argument");
    return NCD_E_INVALID;
  }
  if (!a->committed) {
    ncd_set_error("ncd_arena_fetch_records: arena not committed");
    return NCD_E_STATE;
  }
  if (n == 0) return NCD_OK;
  if (n >= 0x7FFFFFFFull) {
    ncd_set_error("ncd_arena_fetch_records: too many records in one call");
    return NCD_E_CAPACITY;
  }
  for (uint64_t k = 0; k < n; ++k)
    if (obj_idx[k] >= a->nobj) {
      ncd_set_error("ncd_arena_fetch_records: object %u of %llu does not exist", obj_idx[k], (unsigned long long)a->nobj);
      return NCD_E_INVALID;
    }
  ncd_ctx* ctx = a->ctx;
  NCD_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  // lengths first (they size the staging buffer), then the gather
  uint32_t* d_idx = nullptr;
  NCD_CUDA(cudaMallocAsync((void**)&d_idx, n * 4, st));
  NCD_CUDA(cudaMemcpyAsync(d_idx, obj_idx, n * 4, cudaMemcpyHostToDevice, st));
  std::vector<uint32_t> lens(a->nobj);
  NCD_CUDA(cudaMemcpyAsync(lens.data(), a->d.rec_len, a->nobj * 4, cudaMemcpyDeviceToHost, st));
  NCD_CUDA(cudaStreamSynchronize(st));
  uint64_t total = 0;
  for (uint64_t k = 0; k < n; ++k) {
    out_off[k] = total;
    total += ((uint64_t)lens[obj_idx[k]] + 15) & ~15ull;
  }
  out_off[n] = total;
  if (total > out_cap) {
    cudaFreeAsync(d_idx, st);
    ncd_set_error("ncd_arena_fetch_records: %llu bytes needed, buffer holds %llu", (unsigned long long)total, (unsigned long long)out_cap);
    return NCD_E_CAPACITY;
  }
  uint64_t* d_off = nullptr;
  uint8_t* d_out = nullptr;
  NCD_CUDA(cudaMallocAsync((void**)&d_off, (n + 1) * 8, st));
  NCD_CUDA(cudaMallocAsync((void**)&d_out, total ? total : 16, st));
  NCD_CUDA(cudaMemcpyAsync(d_off, out_off, (n + 1) * 8, cudaMemcpyHostToDevice, st));
  k_gather_records<<<(unsigned)n, 128, 0, st>>>(d_idx, d_off, a->d.payload, a->d.rec_off, a->d.rec_len, d_out);
  if (total) NCD_CUDA(cudaMemcpyAsync(out, d_out, total, cudaMemcpyDeviceToHost, st));
  NCD_CUDA(cudaFreeAsync(d_idx, st));
  NCD_CUDA(cudaFreeAsync(d_off, st));
  NCD_CUDA(cudaFreeAsync(d_out, st));
  NCD_CUDA(cudaStreamSynchronize(st));
  NCD_CUDA(cudaGetLastError());
  return NCD_OK;
}

// ------------------------------------------------------------------ compaction
// Dead payload — records of deleted objects, old places of records that grew and were appended —
// used to be reclaimed only by a full ncd_arena_commit, i.e. by sending the whole cluster over
// PCIe again.  Here the live records are moved to the front of the payload where they are: on the
// device by a gather through a scratch buffer, in the pinned host arrays by memmove, in ascending
// order of their offsets so that nothing is overwritten before it is moved.  Offsets change on
// both sides identically; the index, the join caches and the digests do not depend on them.
__global__ void __launch_bounds__(128)
k_compact_gather(const uint8_t* __restrict__ payload, const uint64_t* __restrict__ old_off, const uint64_t* __restrict__ new_off,
                 const uint32_t* __restrict__ rec_len, uint8_t* __restrict__ tmp) {
  const uint32_t i = blockIdx.x;
  const uint4* src = reinterpret_cast<const uint4*>(payload + old_off[i]);
  uint4* dst = reinterpret_cast<uint4*>(tmp + new_off[i]);
  const uint32_t nvec = (rec_len[i] + 15) >> 4;
  for (uint32_t v = threadIdx.x; v < nvec; v += blockDim.x) dst[v] = src[v];
}
__global__ void k_compact_offsets(uint64_t* __restrict__ rec_off, ObjMeta* __restrict__ meta, const uint64_t* __restrict__ new_off, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  rec_off[i] = new_off[i];
  meta[i].rec_off = new_off[i];
}

int ncd_arena_compact(ncd_arena* a, uint64_t* new_payload_bytes) {
  if (!a) {
    ncd_set_error("ncd_arena_compact: null arena");
    return NCD_E_INVALID;
  }
  if (a->flags != NCD_ARENA_HOST_PINNED || !a->committed) {
    ncd_set_error("ncd_arena_compact: needs a committed arena with pinned host arrays");
    return NCD_E_STATE;
  }
  const uint64_t n = a->nobj;
  ncd_ctx* ctx = a->ctx;
  NCD_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  NCD_CUDA(cudaStreamSynchronize(st));
  // new offsets: the live records in ascending order of their current offsets, packed
  std::vector<uint32_t> order(n);
  for (uint64_t i = 0; i < n; ++i) order[i] = (uint32_t)i;
  const uint64_t* off = a->h.rec_off;
  const uint32_t* len = a->h.rec_len;
  std::sort(order.begin(), order.end(), [&](uint32_t x, uint32_t y) { return off[x] != off[y] ? off[x] < off[y] : len[x] < len[y]; });
  std::vector<uint64_t> new_off(n);
  uint64_t cur = 0, prev_end = 0;
  for (uint64_t k = 0; k < n; ++k) {
    const uint32_t i = order[k];
    const uint64_t padded = ((uint64_t)len[i] + 15) & ~15ull;
    if (off[i] < prev_end && padded) {
      ncd_set_error("ncd_arena_compact: records %u overlaps the one before it", i);
      return NCD_E_INVALID;
    }
    new_off[i] = cur;
    cur += padded;
    if (padded) prev_end = off[i] + padded;
  }
  if (cur == a->payload_bytes) {  // nothing to reclaim
    if (new_payload_bytes) *new_payload_bytes = cur;
    return NCD_OK;
  }
  // device first (it still needs the old offsets, which the device arrays hold).  Lengths come from
  // the HOST arrays: a record rewritten in place on the host and not synced yet may be longer or
  // shorter there than the device still believes, and the slots are laid out for the host's lengths.
  uint64_t* d_new = nullptr;
  uint8_t* tmp = nullptr;
  NCD_CUDA(cudaMallocAsync((void**)&d_new, (n ? n : 1) * 12, st));
  uint32_t* d_len = reinterpret_cast<uint32_t*>(d_new + (n ? n : 1));
  cudaError_t e = cudaMallocAsync((void**)&tmp, cur ? cur : 16, st);
  if (e != cudaSuccess) {
    cudaFreeAsync(d_new, st);
    ncd_set_error("ncd_arena_compact: %llu bytes of device scratch: %s", (unsigned long long)cur, cudaGetErrorString(e));
    return NCD_E_NOMEM;
  }
  if (n) {
    NCD_CUDA(cudaMemcpyAsync(d_new, new_off.data(), n * 8, cudaMemcpyHostToDevice, st));
    NCD_CUDA(cudaMemcpyAsync(d_len, len, n * 4, cudaMemcpyHostToDevice, st));
    k_compact_gather<<<(unsigned)n, 128, 0, st>>>(a->d.payload, a->d.rec_off, d_new, d_len, tmp);
    if (cur) NCD_CUDA(cudaMemcpyAsync(a->d.payload, tmp, cur, cudaMemcpyDeviceToDevice, st));
    k_compact_offsets<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(a->d.rec_off, a->meta, d_new, (uint32_t)n);
  }
  NCD_CUDA(cudaFreeAsync(tmp, st));
  NCD_CUDA(cudaFreeAsync(d_new, st));
  // host: the same moves, front to back, in place (new offset <= old offset for every record)
  for (uint64_t k = 0; k < n; ++k) {
    const uint32_t i = order[k];
    const uint64_t padded = ((uint64_t)len[i] + 15) & ~15ull;
    if (padded && new_off[i] != off[i]) memmove(a->h.payload + new_off[i], a->h.payload + off[i], padded);
  }
  for (uint64_t i = 0; i < n; ++i) a->h.rec_off[i] = new_off[i];
  NCD_CUDA(cudaStreamSynchronize(st));
  NCD_CUDA(cudaGetLastError());
  a->payload_bytes = cur;
  a->jc_line_ok = false;  // the offsets the join cached per pair are gone: the next scan reads the object lines again
  if (new_payload_bytes) *new_payload_bytes = cur;
  return NCD_OK;
}
