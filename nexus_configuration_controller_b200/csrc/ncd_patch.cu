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
#include "ncd_common.cuh"

// One CTA per rewritten record.  h_* are the pinned host arrays (unified addressing makes them
// readable from the device); the record lands at the same offset in the device payload.
__global__ void __launch_bounds__(128)
k_pull_records(const uint32_t* __restrict__ idx, const uint8_t* h_payload, const uint64_t* h_rec_off,
               const uint32_t* h_rec_len, uint8_t* __restrict__ d_payload, uint64_t* __restrict__ d_rec_off,
               uint32_t* __restrict__ d_rec_len, ObjMeta* __restrict__ meta, uint32_t nobj,
               uint64_t payload_limit, uint32_t* __restrict__ bad /*[0] refused, [1] lengths changed*/) {
  const uint32_t i = idx[blockIdx.x];
  if (i >= nobj) {
    if (threadIdx.x == 0) atomicAdd(bad, 1u);
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
  }
  const uint4* src = reinterpret_cast<const uint4*>(h_payload + off);
  uint4* dst = reinterpret_cast<uint4*>(d_payload + off);
  for (uint32_t v = threadIdx.x; v < (uint32_t)(padded >> 4); v += blockDim.x) dst[v] = src[v];
}

int ncd_arena_sync_records(ncd_arena* a, uint64_t n, const uint32_t* obj_idx, uint64_t new_payload_bytes) {
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
  if (n == 0) {
    a->payload_bytes = new_payload_bytes;
    return NCD_OK;
  }
  ncd_ctx* ctx = a->ctx;
  NCD_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  uint32_t* d_idx = nullptr;
  NCD_CUDA(cudaMallocAsync((void**)&d_idx, (n + 2) * 4, st));
  NCD_CUDA(cudaMemcpyAsync(d_idx, obj_idx, n * 4, cudaMemcpyHostToDevice, st));
  NCD_CUDA(cudaMemsetAsync(d_idx + n, 0, 8, st));
  k_pull_records<<<(unsigned)n, 128, 0, st>>>(d_idx, a->h.payload, a->h.rec_off, a->h.rec_len, a->d.payload,
                                              a->d.rec_off, a->d.rec_len, a->meta, (uint32_t)a->nobj,
                                              new_payload_bytes, d_idx + n);
  uint32_t flags[2] = {0, 0};
  NCD_CUDA(cudaMemcpyAsync(flags, d_idx + n, 8, cudaMemcpyDeviceToHost, st));
  NCD_CUDA(cudaStreamSynchronize(st));
  a->payload_bytes = new_payload_bytes;
  const uint32_t bad = flags[0];
  // Digests the arena holds stay valid: only the rewritten records are hashed again, so a
  // hash-mode scan after an informer event costs the changed bytes, not the fleet's.
  int hrc = NCD_OK;
  if (bad) a->h64_valid = a->sha_valid = a->xxh_valid = false;
  else hrc = ncd_hash_records(a, d_idx, n);
  NCD_CUDA(cudaFreeAsync(d_idx, st));
  NCD_CUDA(cudaStreamSynchronize(st));
  NCD_CUDA(cudaGetLastError());
  if (hrc != NCD_OK) {
    a->h64_valid = a->sha_valid = a->xxh_valid = false;
    return hrc;
  }
  // same-length rewrites (the common event: a value changed in place) leave the tile lists and
  // the byte count as they are
  if (flags[1] || bad) NCD_TRY(ncd_rebuild_tiles(a));
  if (bad) {
    ncd_set_error("ncd_arena_sync_records: %u of %llu records were refused (index out of range, record "
                  "not 16-byte aligned or outside the payload); commit the arena again", bad,
                  (unsigned long long)n);
    a->committed = false;
    return NCD_E_INVALID;
  }
  return NCD_OK;
}
