// ncd_exchange.cu — the exchange step of the multi-GPU drift check as peer-memory stores.
//
// After a partitioned scan every rank must hold every rank's diff list (north_star: "one all-gather
// of diff lists").  The lists are small (kilobytes), so the cost of a library collective is its
// launch and rendezvous, not its bytes.  Here the compaction kernel writes the rank's list into
// the rank's own slot of an exchange buffer, and ONE more kernel (world-1 blocks) stores that slot
// into the same slot of every peer's buffer over NVLink, raises a flag there, and waits for the
// peer's flag in the local buffer.  No host round trip, no extra stream, nothing allocated per step.
//
//   buffer of one rank:  [3 generations][world slots][16-byte header + cap entries]
//                        flags  u32 [3][world]   (sequence number of the list the slot holds)
//                        error  u32              (1 = a wait timed out)
//
// Step t pushes this rank's list into generation t % 3 of every peer and then waits for the peers'
// lists of step t - 1, not of step t: the ranks run up to one step apart, so the slowest rank of a
// step (the key-hash partition balances bytes only statistically) no longer stalls everybody in
// every step.  The list of step t is complete once ncd_xchg_wait (or ncd_xchg_fetch, which calls
// it) has seen the peers' flags of step t.  A peer writes generation g again only at its step
// t + 3, which it cannot start before this rank has pushed step t + 2: a gathered list stays intact
// until this rank has enqueued two more exchanges.
//
// No reference counterpart: the reference is one process on CPU (controller.go:790-831 walks the
// shards in a loop); this replaces the NCCL all-gather of the first multi-GPU version.
#include <cstring>

#include "ncd_common.cuh"

#define NCD_XCHG_MAX_WORLD 64

struct ncd_xchg {
  ncd_ctx* ctx;
  int world, rank;
  uint64_t cap;         // entries per slot
  uint64_t slot_bytes;  // 16 + cap * 16
  uint64_t flags_off, err_off, buf_bytes;
  uint8_t* buf;                        // this rank's buffer (cudaMalloc)
  uint8_t* peers[NCD_XCHG_MAX_WORLD];  // base of every rank's buffer as seen from this device
  bool opened[NCD_XCHG_MAX_WORLD];     // mapped with cudaIpcOpenMemHandle (to be closed)
  uint8_t** d_peers;                   // device copy of peers[]
  bool connected;
  uint32_t seq;  // steps exchanged so far; step t uses generation t % 3 and flag value t
  uint64_t timeout_ns;
};

namespace {

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint64_t now_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Block b serves one peer: push this rank's list into the peer's buffer, flag it, then wait until
// the peer's list for the same step has arrived here.
__device__ __forceinline__ void wait_flag(uint8_t* mine, uint64_t flags_off, uint64_t err_off, uint32_t slot, uint32_t seq,
                                          uint64_t timeout_ns) {
  const uint32_t* flag = reinterpret_cast<const uint32_t*>(mine + flags_off) + slot;
  const uint64_t t0 = now_ns();
  while ((int32_t)(ld_acquire_sys(flag) - seq) < 0) {
    if (now_ns() - t0 > timeout_ns) {  // a peer that never comes must not hang the GPU
      atomicExch(reinterpret_cast<uint32_t*>(mine + err_off), 1u);
      break;
    }
    __nanosleep(200);
  }
}

// Block b waits until peer b's list of step `seq` (generation gen) has arrived here.
__global__ void k_xchg_wait(uint8_t* const* __restrict__ peers, int world, int rank, uint64_t flags_off, uint64_t err_off,
                            uint32_t gen, uint32_t seq, uint64_t timeout_ns) {
  const int p = (int)blockIdx.x < rank ? (int)blockIdx.x : (int)blockIdx.x + 1;
  if (threadIdx.x == 0) wait_flag(peers[rank], flags_off, err_off, gen * world + p, seq, timeout_ns);
}

__global__ void __launch_bounds__(256)
k_exchange(uint8_t* const* __restrict__ peers, int world, int rank, uint64_t slot_bytes, uint64_t cap,
           uint64_t flags_off, uint64_t err_off, uint32_t gen, uint32_t seq, uint32_t wait_gen, uint32_t wait_seq,
           uint64_t timeout_ns) {
  const int p = (int)blockIdx.x < rank ? (int)blockIdx.x : (int)blockIdx.x + 1;
  uint8_t* mine = peers[rank];
  uint8_t* theirs = peers[p];
  const uint64_t slot = ((uint64_t)gen * world + rank) * slot_bytes;
  const uint4* src = reinterpret_cast<const uint4*>(mine + slot);
  uint4* dst = reinterpret_cast<uint4*>(theirs + slot);
  const uint32_t found = reinterpret_cast<const uint32_t*>(src)[0];
  const uint64_t nvec = 1 + (found < cap ? (uint64_t)found : cap);  // header + entries actually written
  for (uint64_t i = threadIdx.x; i < nvec; i += blockDim.x) dst[i] = src[i];
  __threadfence_system();  // this thread's stores are ordered before the flag below
  __syncthreads();
  if (threadIdx.x == 0) {
    st_release_sys(reinterpret_cast<uint32_t*>(theirs + flags_off) + (gen * world + rank), seq);
    // the peer's list of the PREVIOUS step: by now it has usually been here for a whole step
    wait_flag(mine, flags_off, err_off, wait_gen * world + p, wait_seq, timeout_ns);
  }
}

int set_peers(ncd_xchg* x) {
  NCD_CUDA(cudaMemcpyAsync(x->d_peers, x->peers, sizeof(uint8_t*) * x->world, cudaMemcpyHostToDevice, x->ctx->stream));
  NCD_CUDA(cudaStreamSynchronize(x->ctx->stream));
  x->connected = true;
  return NCD_OK;
}

}  // namespace

extern "C" int ncd_xchg_create(ncd_ctx* ctx, int world, int rank, uint64_t cap_entries, ncd_xchg** out) {
  if (!ctx || !out || world < 1 || world > NCD_XCHG_MAX_WORLD || rank < 0 || rank >= world || cap_entries == 0 ||
      cap_entries > (1ull << 28)) {
    ncd_set_error("ncd_xchg_create: bad argument (world 1..%d, 0 <= rank < world, 0 < cap <= 2^28)", NCD_XCHG_MAX_WORLD);
    return NCD_E_INVALID;
  }
  NCD_CUDA(cudaSetDevice(ctx->device));
  ncd_xchg* x = (ncd_xchg*)calloc(1, sizeof(ncd_xchg));
  if (!x) {
    ncd_set_error("ncd_xchg_create: out of host memory");
    return NCD_E_NOMEM;
  }
  x->ctx = ctx;
  x->world = world;
  x->rank = rank;
  x->cap = cap_entries;
  x->slot_bytes = 16 + cap_entries * sizeof(ncd_entry);
  x->flags_off = (3 * (uint64_t)world * x->slot_bytes + 127) & ~127ull;
  x->err_off = x->flags_off + 3 * (uint64_t)world * sizeof(uint32_t);
  x->buf_bytes = (x->err_off + sizeof(uint32_t) + 127) & ~127ull;
  x->timeout_ns = 10ull * 1000 * 1000 * 1000;
  if (cudaMalloc((void**)&x->buf, x->buf_bytes) != cudaSuccess || cudaMalloc((void**)&x->d_peers, sizeof(uint8_t*) * world) != cudaSuccess) {
    ncd_set_error("ncd_xchg_create: cudaMalloc of %llu bytes failed: %s", (unsigned long long)x->buf_bytes,
                  cudaGetErrorString(cudaGetLastError()));
    if (x->buf) cudaFree(x->buf);
    free(x);
    return NCD_E_NOMEM;
  }
  NCD_CUDA(cudaMemsetAsync(x->buf, 0, x->buf_bytes, ctx->stream));
  NCD_CUDA(cudaStreamSynchronize(ctx->stream));
  x->peers[rank] = x->buf;
  if (world == 1) NCD_TRY(set_peers(x));
  *out = x;
  return NCD_OK;
}

extern "C" int ncd_xchg_buffer(ncd_xchg* x, void** dev_ptr, uint64_t* bytes) {
  if (!x) {
    ncd_set_error("ncd_xchg_buffer: null argument");
    return NCD_E_INVALID;
  }
  if (dev_ptr) *dev_ptr = x->buf;
  if (bytes) *bytes = x->buf_bytes;
  return NCD_OK;
}

extern "C" int ncd_xchg_handle(ncd_xchg* x, void* handle64) {
  if (!x || !handle64) {
    ncd_set_error("ncd_xchg_handle: null argument");
    return NCD_E_INVALID;
  }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "the ABI promises a 64-byte handle");
  NCD_CUDA(cudaSetDevice(x->ctx->device));
  cudaIpcMemHandle_t h;
  NCD_CUDA(cudaIpcGetMemHandle(&h, x->buf));
  memcpy(handle64, &h, 64);
  return NCD_OK;
}

extern "C" int ncd_xchg_connect(ncd_xchg* x, void* const* peer_bufs) {
  if (!x || !peer_bufs) {
    ncd_set_error("ncd_xchg_connect: null argument");
    return NCD_E_INVALID;
  }
  if (x->connected && x->world > 1) {
    ncd_set_error("ncd_xchg_connect: already connected");
    return NCD_E_STATE;
  }
  for (int p = 0; p < x->world; ++p) {
    if (p == x->rank) continue;
    if (!peer_bufs[p]) {
      ncd_set_error("ncd_xchg_connect: no buffer for rank %d", p);
      return NCD_E_INVALID;
    }
    x->peers[p] = (uint8_t*)peer_bufs[p];
  }
  NCD_CUDA(cudaSetDevice(x->ctx->device));
  return set_peers(x);
}

extern "C" int ncd_xchg_open(ncd_xchg* x, const void* handles) {
  if (!x || !handles) {
    ncd_set_error("ncd_xchg_open: null argument");
    return NCD_E_INVALID;
  }
  if (x->connected && x->world > 1) {
    ncd_set_error("ncd_xchg_open: already connected");
    return NCD_E_STATE;
  }
  NCD_CUDA(cudaSetDevice(x->ctx->device));
  for (int p = 0; p < x->world; ++p) {
    if (p == x->rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, (const char*)handles + 64 * (size_t)p, 64);
    void* ptr = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      ncd_set_error("ncd_xchg_open: cannot map the buffer of rank %d: %s", p, cudaGetErrorString(e));
      cudaGetLastError();
      for (int q = 0; q < p; ++q)
        if (x->opened[q]) {
          cudaIpcCloseMemHandle(x->peers[q]);
          x->opened[q] = false;
        }
      return NCD_E_CUDA;
    }
    x->peers[p] = (uint8_t*)ptr;
    x->opened[p] = true;
  }
  return set_peers(x);
}

extern "C" int ncd_scan_exchange(ncd_ctx* ctx, const ncd_arena* ctrl, const ncd_arena* const* shards, int nshards,
                                 uint32_t mode, ncd_xchg* x) {
  if (!x || x->ctx != ctx) {
    ncd_set_error("ncd_scan_exchange: exchange does not belong to this context");
    return NCD_E_INVALID;
  }
  if (!x->connected) {
    ncd_set_error("ncd_scan_exchange: exchange not connected (ncd_xchg_open / ncd_xchg_connect first)");
    return NCD_E_STATE;
  }
  const uint32_t seq = x->seq + 1, gen = seq % 3u;
  uint8_t* slot = x->buf + ((uint64_t)gen * x->world + x->rank) * x->slot_bytes;
  NCD_TRY(ncd_scan_into(ctx, ctrl, shards, nshards, mode, slot, x->cap));
  NCD_RANGE(xchg_range, "ncd:exchange");
  if (x->world > 1) {
    k_exchange<<<x->world - 1, 256, 0, ctx->stream>>>(x->d_peers, x->world, x->rank, x->slot_bytes, x->cap, x->flags_off,
                                                      x->err_off, gen, seq, (seq - 1) % 3u, seq - 1, x->timeout_ns);
    NCD_CUDA(cudaGetLastError());
  }
  x->seq = seq;
  return NCD_OK;
}

extern "C" int ncd_xchg_result(ncd_xchg* x, const void** dev_ptr, uint64_t* bytes) {
  if (!x || x->seq == 0) {
    ncd_set_error("ncd_xchg_result: no exchange has been enqueued");
    return x ? NCD_E_STATE : NCD_E_INVALID;
  }
  if (dev_ptr) *dev_ptr = x->buf + (uint64_t)(x->seq % 3u) * x->world * x->slot_bytes;
  if (bytes) *bytes = (uint64_t)x->world * x->slot_bytes;
  return NCD_OK;
}

extern "C" int ncd_xchg_wait(ncd_xchg* x) {
  if (!x || x->seq == 0) {
    ncd_set_error("ncd_xchg_wait: no exchange has been enqueued");
    return x ? NCD_E_STATE : NCD_E_INVALID;
  }
  if (x->world > 1) {
    NCD_CUDA(cudaSetDevice(x->ctx->device));
    k_xchg_wait<<<x->world - 1, 32, 0, x->ctx->stream>>>(x->d_peers, x->world, x->rank, x->flags_off, x->err_off, x->seq % 3u, x->seq,
                                                         x->timeout_ns);
    NCD_CUDA(cudaGetLastError());
  }
  return NCD_OK;
}

extern "C" int ncd_xchg_fetch(ncd_xchg* x, void* host_out, uint64_t bytes) {
  const void* src = nullptr;
  uint64_t have = 0;
  NCD_TRY(ncd_xchg_result(x, &src, &have));
  NCD_TRY(ncd_xchg_wait(x));
  if (!host_out || bytes < have) {
    ncd_set_error("ncd_xchg_fetch: output buffer too small (%llu bytes, need %llu)", (unsigned long long)bytes,
                  (unsigned long long)have);
    return NCD_E_CAPACITY;
  }
  NCD_CUDA(cudaSetDevice(x->ctx->device));
  cudaStream_t st = x->ctx->stream;
  uint32_t err = 0;
  // headers first (16 bytes per slot), then only the entries each slot really holds: the slots are
  // sized for the worst case, the lists are usually a small part of that
  NCD_CUDA(cudaMemcpy2DAsync(host_out, x->slot_bytes, src, x->slot_bytes, 16, (size_t)x->world, cudaMemcpyDeviceToHost, st));
  NCD_CUDA(cudaMemcpyAsync(&err, x->buf + x->err_off, sizeof(err), cudaMemcpyDeviceToHost, st));
  NCD_CUDA(cudaStreamSynchronize(st));
  for (int r = 0; r < x->world; ++r) {
    uint32_t found = 0;
    memcpy(&found, (const char*)host_out + (size_t)r * x->slot_bytes, 4);
    const uint64_t nent = found < x->cap ? found : x->cap;
    if (nent)
      NCD_CUDA(cudaMemcpyAsync((char*)host_out + (size_t)r * x->slot_bytes + 16, (const char*)src + (size_t)r * x->slot_bytes + 16,
                               nent * sizeof(ncd_entry), cudaMemcpyDeviceToHost, st));
  }
  NCD_CUDA(cudaStreamSynchronize(st));
  if (err) {
    // reported once: the word is cleared so that one slow peer does not fail every later step
    cudaMemsetAsync(x->buf + x->err_off, 0, sizeof(uint32_t), st);
    cudaStreamSynchronize(st);
    ncd_set_error("ncd_xchg_fetch: a peer's diff list did not arrive within %.1f s", x->timeout_ns * 1e-9);
    return NCD_E_STATE;
  }
  return NCD_OK;
}

extern "C" int ncd_xchg_set_timeout(ncd_xchg* x, double seconds) {
  if (!x || !(seconds > 0)) {
    ncd_set_error("ncd_xchg_set_timeout: bad argument");
    return NCD_E_INVALID;
  }
  x->timeout_ns = (uint64_t)(seconds * 1e9);
  return NCD_OK;
}

extern "C" void ncd_xchg_destroy(ncd_xchg* x) {
  if (!x) return;
  cudaSetDevice(x->ctx->device);
  cudaStreamSynchronize(x->ctx->stream);
  for (int p = 0; p < x->world; ++p)
    if (x->opened[p]) cudaIpcCloseMemHandle(x->peers[p]);
  if (x->d_peers) cudaFree(x->d_peers);
  if (x->buf) cudaFree(x->buf);
  free(x);
}
