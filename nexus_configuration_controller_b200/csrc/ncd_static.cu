// ncd_static.cu — ncd_hash64_static: the NCD-H64 record hash for callers that must not
// allocate.  It has no counterpart in the reference (which never hashes) and no counterpart
// in the rest of this library either: every other entry point works on arenas and contexts
// that own cudaMalloc'ed memory.
//
// Who needs it.  Some hosts of this library run with the GPU's memory already spoken for:
// a controller that shares its B200 with other tenants under a fixed reservation, a process
// that hands every device byte to its own pool at start-up (a PyTorch caching allocator, a
// Go-side slab), or an event handler that may not risk cudaMalloc's device-wide
// synchronisation and its out-of-memory failure in the middle of a reconcile.  Such a caller
// brings its own device buffers for the records and for the digests; what it cannot bring is
// the kernel's working memory (the tile index and the per-record accumulators), because the
// size of that depends on the records.  Here that working memory is a fixed array inside the
// library image, so the call itself allocates nothing: no cudaMalloc, cudaMallocAsync,
// cudaFree, no pinned host memory, no library routine that allocates.
//
// Price: the array is 768.25 KiB of every GPU that loads libncdrift.so, held until the process
// exits, and one call can hash at most NCD_STATIC_MAX_RECORDS (65 536) records.  All calls on
// a device share the array and it keeps the previous call's contents; each call re-initialises
// the parts it reads, but two calls that overlap on the GPU (different streams, or different
// threads on one stream without ordering) race on it and give wrong digests.  Serialise them.
//
// The digests are bit-identical to ncd_hash_arena(NCD_MODE_HASH64) on the same records.
#include <mutex>

#include "ncd_common.cuh"

namespace {

constexpr uint32_t MAX_RECORDS = NCD_STATIC_MAX_RECORDS;
constexpr size_t TILE_START_BYTES = ((size_t)(MAX_RECORDS + 1) * 4 + 255) / 256 * 256;
constexpr size_t ACC_BYTES = (size_t)MAX_RECORDS * 8;
constexpr size_t SCRATCH_BYTES = TILE_START_BYTES + ACC_BYTES;

// The scratch.  In HOST code the name g_scratch is only a handle for the symbol, not an
// address a kernel can use: passing it to a kernel would fault (or, where the GPU can reach
// host memory, quietly work on the wrong bytes).  The device address comes from
// cudaGetSymbolAddress, once per device, below.
static __device__ __align__(256) unsigned char g_scratch[SCRATCH_BYTES];

constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
void* g_scratch_dev[MAX_DEVICES] = {nullptr};
int g_sm_count[MAX_DEVICES] = {0};

__device__ __forceinline__ uint64_t rotl64s(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }

// One block: tiles per record, their exclusive scan (tile_start[n] = total), accumulators
// cleared.  Everything the later kernels read from the scratch is written here first, so
// nothing left over from the previous call is ever used.
__global__ void __launch_bounds__(1024)
k_static_prepare(const uint32_t* __restrict__ rec_len, uint32_t n, uint32_t* __restrict__ tile_start,
                 unsigned long long* __restrict__ acc) {
  __shared__ uint32_t warp_tot[32];
  const uint32_t per = (n + 1023) / 1024;  // consecutive records per thread
  const uint32_t b = threadIdx.x * per, e = min(n, b + per);
  uint32_t sum = 0;
  for (uint32_t i = b; i < e; ++i) sum += (rec_len[i] + NCD_TILE_BYTES - 1) / NCD_TILE_BYTES;
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t incl = sum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += o;
  }
  if (lane == 31) warp_tot[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    const uint32_t w = warp_tot[lane];
    uint32_t wi = w;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t o = __shfl_up_sync(0xffffffffu, wi, d);
      if (lane >= d) wi += o;
    }
    warp_tot[lane] = wi - w;
    if (lane == 31) tile_start[n] = wi;
  }
  __syncthreads();
  uint32_t run = warp_tot[warp] + incl - sum;
  for (uint32_t i = b; i < e; ++i) {
    tile_start[i] = run;
    acc[i] = 0ull;
    run += (rec_len[i] + NCD_TILE_BYTES - 1) / NCD_TILE_BYTES;
  }
}

// Same arithmetic as k_h64_tiles (ncd_hash.cu); the record a tile belongs to is found by
// binary search in tile_start instead of a precomputed tile list.
__global__ void __launch_bounds__(256)
k_static_h64_tiles(const uint8_t* __restrict__ payload, const uint64_t* __restrict__ rec_off,
                   const uint32_t* __restrict__ rec_len, uint32_t n, const uint32_t* __restrict__ tile_start,
                   unsigned long long* __restrict__ acc) {
  __shared__ uint4 keys[NCD_TILE_VECS];
  for (uint32_t v = threadIdx.x; v < NCD_TILE_VECS; v += blockDim.x) {
    const uint64_t k0 = ncd_splitmix64(0x6e63642d68363400ull + 2 * v);
    const uint64_t k1 = ncd_splitmix64(0x6e63642d68363400ull + 2 * v + 1);
    keys[v] = make_uint4((uint32_t)k0, (uint32_t)(k0 >> 32), (uint32_t)k1, (uint32_t)(k1 >> 32));
  }
  __syncthreads();
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t warp0 = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
  const uint32_t ntiles = tile_start[n];
  constexpr int U = NCD_TILE_VECS / 32;
  for (uint32_t t = warp0; t < ntiles; t += nwarps) {
    uint32_t lo = 0, hi = n;  // largest c with tile_start[c] <= t
    while (hi - lo > 1) {
      const uint32_t mid = (lo + hi) >> 1;
      if (tile_start[mid] <= t) lo = mid;
      else hi = mid;
    }
    const uint32_t c = lo, chunk = t - tile_start[c];
    const uint32_t len = rec_len[c];
    const uint64_t toff = (uint64_t)chunk * NCD_TILE_BYTES;
    const uint32_t nbytes = min(NCD_TILE_BYTES, len - (uint32_t)toff);
    const uint32_t nvec = (nbytes + 15) >> 4;
    const uint32_t tail = nbytes & 15;
    const uint4* p = reinterpret_cast<const uint4*>(payload + rec_off[c] + toff);
    uint4 d[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint32_t v = lane + 32 * u;
      d[u] = v < nvec ? ncd_ldg_stream(p + v) : make_uint4(0, 0, 0, 0);
    }
    uint64_t A = 0, B = 0;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint32_t v = lane + 32 * u;
      if (v < nvec) {
        uint4 x = d[u];
        if (tail && v == nvec - 1) {
          uint32_t w[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int nb = (int)tail - 4 * i;
            if (nb <= 0) w[i] = 0;
            else if (nb < 4) w[i] &= (1u << (8 * nb)) - 1u;
          }
          x = make_uint4(w[0], w[1], w[2], w[3]);
        }
        const uint4 k = keys[v];
        const uint64_t d0 = (uint64_t)x.x | ((uint64_t)x.y << 32);
        const uint64_t d1 = (uint64_t)x.z | ((uint64_t)x.w << 32);
        A += (uint64_t)(x.x ^ k.x) * (uint64_t)(x.y ^ k.y) + d1;
        B += (uint64_t)(x.z ^ k.z) * (uint64_t)(x.w ^ k.w) + d0;
      }
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
      A += __shfl_xor_sync(0xffffffffu, A, s);
      B += __shfl_xor_sync(0xffffffffu, B, s);
    }
    if (lane == 0) {
      const uint64_t T = ncd_fmix64(ncd_fmix64(A + ncd_splitmix64(2ull * chunk)) + rotl64s(B, 31) +
                                    ncd_splitmix64(2ull * chunk + 1));
      atomicAdd(&acc[c], (unsigned long long)T);
    }
  }
}

__global__ void k_static_h64_final(const uint32_t* __restrict__ rec_len, uint32_t n,
                                   const unsigned long long* __restrict__ acc, uint64_t* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  out[i] = ncd_fmix64((uint64_t)acc[i] + (uint64_t)rec_len[i] * 0x9E3779B185EBCA87ull);
}

// Device address of g_scratch on the CURRENT device, looked up on first use.
int scratch_for_current_device(void** out, int* sm_count) {
  int dev = -1;
  NCD_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= MAX_DEVICES) {
    ncd_set_error("ncd_hash64_static: device ordinal %d not supported (0..%d)", dev, MAX_DEVICES - 1);
    return NCD_E_INVALID;
  }
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_scratch_dev[dev]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_scratch);
    if (e != cudaSuccess || !p) {
      ncd_set_error("ncd_hash64_static: cudaGetSymbolAddress(scratch) on device %d failed: %s", dev,
                    cudaGetErrorString(e));
      return NCD_E_CUDA;
    }
    int sms = 0;
    NCD_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    g_sm_count[dev] = sms > 0 ? sms : 1;
    g_scratch_dev[dev] = p;
  }
  *out = g_scratch_dev[dev];
  *sm_count = g_sm_count[dev];
  return NCD_OK;
}

}  // namespace

extern "C" int ncd_hash64_static(const void* dev_payload, const uint64_t* dev_rec_off,
                                 const uint32_t* dev_rec_len, uint64_t nrec, uint64_t* dev_out,
                                 void* stream) {
  if (nrec > MAX_RECORDS) {
    ncd_set_error("ncd_hash64_static: %llu records in one call, the fixed scratch holds at most %u; "
                  "split the call or use ncd_hash_arena",
                  (unsigned long long)nrec, MAX_RECORDS);
    return NCD_E_CAPACITY;
  }
  if (nrec == 0) return NCD_OK;
  if (!dev_payload || !dev_rec_off || !dev_rec_len || !dev_out) {
    ncd_set_error("ncd_hash64_static: null device pointer");
    return NCD_E_INVALID;
  }
  void* scratch = nullptr;
  int sms = 1;
  NCD_TRY(scratch_for_current_device(&scratch, &sms));
  uint32_t* tile_start = reinterpret_cast<uint32_t*>(scratch);
  unsigned long long* acc = reinterpret_cast<unsigned long long*>((char*)scratch + TILE_START_BYTES);
  cudaStream_t st = (cudaStream_t)stream;
  const uint32_t n = (uint32_t)nrec;
  k_static_prepare<<<1, 1024, 0, st>>>(dev_rec_len, n, tile_start, acc);
  k_static_h64_tiles<<<sms * 8, 256, 0, st>>>((const uint8_t*)dev_payload, dev_rec_off, dev_rec_len, n,
                                              tile_start, acc);
  k_static_h64_final<<<(n + 255) / 256, 256, 0, st>>>(dev_rec_len, n, acc, dev_out);
  NCD_CUDA(cudaGetLastError());
  return NCD_OK;
}
