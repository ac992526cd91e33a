// ncd_index.cu — what ncd_arena_commit builds on the GPU for a freshly uploaded arena:
//   one 64-byte line per object (key hash, record offset/length, inline key) + the
//   open-addressing table that answers lister.Get
//     (reference: client-go indexer behind controller.go:507,516,569,578,702,722,766,792),
//   the row list (one row per template, per dependent reference, per workgroup;
//     reference loop structure controller.go:790-831 with :505,:567 for the dependents),
//   the tile list the compare and hash kernels iterate over.
// Also the device exclusive scan used here and by the synthetic generator.
#include <atomic>

#include "ncd_common.cuh"

// ------------------------------------------------------------------ exclusive scan
// Three launches: per-block scan (2048 items / block) + block totals, one block scans the
// totals, per-block add.  Input u32, output u32 or u64.
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

template <typename T>
__device__ __forceinline__ T block_exclusive_scan(T v, T* total, T* smem /*[SCAN_THREADS/32 + 1]*/) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  T incl = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    T o = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += o;
  }
  if (lane == 31) smem[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    T w = lane < (SCAN_THREADS / 32) ? smem[lane] : T(0);
    T wi = w;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      T o = __shfl_up_sync(0xffffffffu, wi, d);
      if (lane >= d) wi += o;
    }
    if (lane < (SCAN_THREADS / 32)) smem[lane] = wi - w;
    if (lane == 31) smem[SCAN_THREADS / 32] = wi;
  }
  __syncthreads();
  T res = smem[warp] + incl - v;
  *total = smem[SCAN_THREADS / 32];
  __syncthreads();
  return res;
}

template <typename T>
__global__ void __launch_bounds__(SCAN_THREADS)
k_scan_blocks(const uint32_t* __restrict__ in, T* __restrict__ out, uint64_t n, T* __restrict__ block_sums) {
  __shared__ T smem[SCAN_THREADS / 32 + 1];
  const uint64_t base = (uint64_t)blockIdx.x * SCAN_TILE + (uint64_t)threadIdx.x * SCAN_ITEMS;
  T v[SCAN_ITEMS];
  T sum = 0;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i) {
    v[i] = (base + i < n) ? (T)in[base + i] : T(0);
    sum += v[i];
  }
  T total;
  T excl = block_exclusive_scan<T>(sum, &total, smem);
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i) {
    if (base + i < n) out[base + i] = excl;
    excl += v[i];
  }
  if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}

template <typename T>
__global__ void __launch_bounds__(SCAN_THREADS)
k_scan_sums(T* __restrict__ block_sums, uint64_t nblocks, T* __restrict__ total_out) {
  __shared__ T smem[SCAN_THREADS / 32 + 1];
  T carry = 0;
  for (uint64_t base = 0; base < nblocks; base += SCAN_THREADS) {
    uint64_t i = base + threadIdx.x;
    T v = i < nblocks ? block_sums[i] : T(0);
    T total;
    T excl = block_exclusive_scan<T>(v, &total, smem);
    if (i < nblocks) block_sums[i] = carry + excl;
    carry += total;
  }
  if (threadIdx.x == 0 && total_out) *total_out = carry;
}

template <typename T>
__global__ void __launch_bounds__(SCAN_THREADS)
k_scan_add(T* __restrict__ out, uint64_t n, const T* __restrict__ block_sums) {
  const uint64_t base = (uint64_t)blockIdx.x * SCAN_TILE + (uint64_t)threadIdx.x * SCAN_ITEMS;
  const T add = block_sums[blockIdx.x];
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i)
    if (base + i < n) out[base + i] += add;
}

// out has n + 1 entries: out[n] = total.
template <typename T>
static int exclusive_scan(ncd_ctx* ctx, const uint32_t* in, T* out, uint64_t n, T* total_out_dev) {
  if (n == 0) {
    NCD_CUDA(cudaMemsetAsync(out, 0, sizeof(T), ctx->stream));
    if (total_out_dev) NCD_CUDA(cudaMemsetAsync(total_out_dev, 0, sizeof(T), ctx->stream));
    return NCD_OK;
  }
  const uint64_t nblocks = (n + SCAN_TILE - 1) / SCAN_TILE;
  T* sums = nullptr;
  NCD_CUDA(cudaMallocAsync((void**)&sums, nblocks * sizeof(T), ctx->stream));
  k_scan_blocks<T><<<(unsigned)nblocks, SCAN_THREADS, 0, ctx->stream>>>(in, out, n, sums);
  k_scan_sums<T><<<1, SCAN_THREADS, 0, ctx->stream>>>(sums, nblocks, out + n);
  k_scan_add<T><<<(unsigned)nblocks, SCAN_THREADS, 0, ctx->stream>>>(out, n, sums);
  if (total_out_dev)
    NCD_CUDA(cudaMemcpyAsync(total_out_dev, out + n, sizeof(T), cudaMemcpyDeviceToDevice, ctx->stream));
  NCD_CUDA(cudaFreeAsync(sums, ctx->stream));
  NCD_CUDA(cudaGetLastError());
  return NCD_OK;
}

int ncd_exclusive_scan_u32(ncd_ctx* ctx, const uint32_t* in, uint32_t* out, uint64_t n,
                           uint32_t* total_out_dev) {
  return exclusive_scan<uint32_t>(ctx, in, out, n, total_out_dev);
}
int ncd_exclusive_scan_u64(ncd_ctx* ctx, const uint32_t* in, uint64_t* out, uint64_t n,
                           uint64_t* total_out_dev) {
  return exclusive_scan<uint64_t>(ctx, in, out, n, total_out_dev);
}

// ------------------------------------------------------------------ object lines + table
__global__ void k_meta_build(DevArena a, ObjMeta* __restrict__ meta) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.nobj) return;
  const uint8_t* key = a.aux + a.key_off[i];
  const uint32_t klen = a.key_len[i];
  ObjMeta m;
  m.keyhash = ncd_keyhash_impl(a.kind[i], key, klen);
  m.rec_off = a.rec_off[i];
  m.rec_len = a.rec_len[i];
  m.key_off = a.key_off[i];
  m.key_len = (uint16_t)klen;
  m.kind = a.kind[i];
  m.pad = 0;
  if (m.kind == NCD_KIND_TEMPLATE) {
    m.tag = a.uid_id[i];
  } else {
    const uint32_t o0 = a.own_start[i], o1 = a.own_start[i + 1];
    m.tag = o0 == o1 ? NCD_TAG_ROGUE : (o1 - o0 == 1 ? a.own_packed[o0] : NCD_TAG_MANY);
  }
#pragma unroll
  for (uint32_t b = 0; b < NCD_KEY_INLINE; ++b) m.key[b] = b < klen ? key[b] : 0;
  const uint4* src = reinterpret_cast<const uint4*>(&m);
  uint4* dst = reinterpret_cast<uint4*>(&meta[i]);
#pragma unroll
  for (int w = 0; w < 4; ++w) dst[w] = src[w];
}

// Inserts every object; on a duplicate ("kind", key) the lowest index wins, as a store keyed
// by namespace/name would have kept one object only.
__global__ void k_table_build(DevArena a, unsigned long long* __restrict__ table, uint32_t* __restrict__ dup_count) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.nobj) return;
  const ObjMeta& me = a.meta[i];
  const uint64_t h = me.keyhash;
  const unsigned long long mine = ((h >> 32) << 32) | i;
  uint32_t slot = (uint32_t)h & a.table_mask;
  for (;;) {
    const unsigned long long prev = atomicCAS(&table[slot], NCD_TABLE_EMPTY, mine);
    if (prev == NCD_TABLE_EMPTY) return;
    if ((prev >> 32) == (h >> 32)) {
      const ObjMeta& o = a.meta[(uint32_t)prev];
      if (o.keyhash == h && o.kind == me.kind && o.key_len == me.key_len &&
          ncd_bytes_equal(a.aux + o.key_off, a.aux + me.key_off, me.key_len)) {
        atomicMin(&table[slot], mine);  // same upper half: orders by index
        atomicAdd(dup_count, 1u);
        return;
      }
    }
    slot = (slot + 1) & a.table_mask;
  }
}

// ------------------------------------------------------------------ UID interning
// isMissingOwnership (controller.go:484-502) compares ownerReference UIDs with the UID of the
// shard's copy of the owning template, for every (dependent, shard) of every scan.  The strings
// never change between commits, so they are interned once here: string j (j < nobj: object j's
// uid, else owner entry j - nobj) gets the smallest j' whose bytes are identical.  The scan then
// compares two integers, exactly: equal ids <=> equal bytes.
__device__ __forceinline__ void uid_string(const DevArena& a, uint32_t j, const uint8_t** p, uint32_t* len) {
  if (j < a.nobj) {
    *p = a.aux + a.uid_off[j];
    *len = a.uid_len[j];
  } else {
    *p = a.aux + a.own_uid_off[j - a.nobj];
    *len = a.own_uid_len[j - a.nobj];
  }
}
__device__ __forceinline__ uint64_t uid_hash(const uint8_t* p, uint32_t len) {
  uint64_t h = 0xcbf29ce484222325ull;
  for (uint32_t i = 0; i < len; ++i) h = (h ^ p[i]) * 0x100000001b3ull;
  return ncd_fmix64(h ^ len);
}

__global__ void k_uid_insert(DevArena a, uint32_t nstr, uint32_t* __restrict__ table, uint32_t mask) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nstr) return;
  const uint8_t* p;
  uint32_t len;
  uid_string(a, j, &p, &len);
  uint32_t slot = (uint32_t)uid_hash(p, len) & mask;
  for (;;) {
    const uint32_t prev = atomicCAS(&table[slot], NCD_NOTFOUND, j);
    if (prev == NCD_NOTFOUND) return;
    const uint8_t* q;
    uint32_t qlen;
    uid_string(a, prev, &q, &qlen);
    if (qlen == len && ncd_bytes_equal(p, q, len)) {
      atomicMin(&table[slot], j);  // any member of the group may sit here: they are byte-equal
      return;
    }
    slot = (slot + 1) & mask;
  }
}

__global__ void k_uid_resolve(DevArena a, uint32_t nstr, const uint32_t* __restrict__ table, uint32_t mask,
                              uint32_t* __restrict__ uid_id, uint32_t* __restrict__ own_packed) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nstr) return;
  const uint8_t* p;
  uint32_t len;
  uid_string(a, j, &p, &len);
  uint32_t slot = (uint32_t)uid_hash(p, len) & mask;
  uint32_t id;
  for (;;) {
    id = table[slot];  // never empty before this string's own group is met
    const uint8_t* q;
    uint32_t qlen;
    uid_string(a, id, &q, &qlen);
    if (qlen == len && ncd_bytes_equal(p, q, len)) break;
    slot = (slot + 1) & mask;
  }
  if (j < a.nobj) uid_id[j] = id;
  else own_packed[j - a.nobj] = id | (a.own_is_template[j - a.nobj] ? 0x80000000u : 0u);
}

// ------------------------------------------------------------------ rows
__global__ void k_row_counts(DevArena a, uint32_t* __restrict__ counts) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.nobj) return;
  uint32_t k = a.kind[i], c = 0;
  if (k == NCD_KIND_TEMPLATE) c = 1 + (a.dep_start[i + 1] - a.dep_start[i]);
  else if (k == NCD_KIND_WORKGROUP) c = 1;
  counts[i] = c;
}

// One thread per object writes its rows.  A dependent reference is resolved against the
// controller arena's own table here (c.secretLister.Get / c.configMapLister.Get,
// controller.go:507,569): NCD_NOTFOUND marks a reference to an object the controller
// cluster does not hold.
__global__ void k_rows_fill(DevArena a, const uint32_t* __restrict__ row_start,
                            uint32_t* __restrict__ row_obj, uint32_t* __restrict__ row_dep,
                            uint32_t* __restrict__ row_cidx) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.nobj) return;
  uint32_t k = a.kind[i];
  if (k != NCD_KIND_TEMPLATE && k != NCD_KIND_WORKGROUP) return;
  uint32_t r = row_start[i];
  row_obj[r] = i;
  row_dep[r] = NCD_NOTFOUND;
  row_cidx[r] = i;
  if (k != NCD_KIND_TEMPLATE) return;
  for (uint32_t d = a.dep_start[i]; d < a.dep_start[i + 1]; ++d) {
    ++r;
    const uint8_t* key = a.aux + a.dep_key_off[d];
    const uint32_t klen = a.dep_key_len[d], dk = a.dep_kind[d];
    row_obj[r] = i;
    row_dep[r] = d;
    row_cidx[r] = ncd_probe(a, dk, key, klen, ncd_keyhash_impl(dk, key, klen));
  }
}

// ------------------------------------------------------------------ tiles
__global__ void k_tile_counts(DevArena a, uint32_t* __restrict__ counts,
                              unsigned long long* __restrict__ len_sum, uint32_t* __restrict__ len_max) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long len = 0;
  if (i < a.nobj) {
    len = a.rec_len[i];
    counts[i] = (uint32_t)((len + NCD_TILE_BYTES - 1) / NCD_TILE_BYTES);
  }
  const uint32_t longest = __reduce_max_sync(0xffffffffu, (uint32_t)len);
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) len += __shfl_xor_sync(0xffffffffu, len, d);
  if ((threadIdx.x & 31) == 0 && len) {
    atomicAdd(len_sum, len);
    atomicMax(len_max, longest);
  }
}

__global__ void k_ctile_counts(DevArena a, uint32_t* __restrict__ counts) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < a.nobj) counts[i] = ncd_ctiles_of(a.rec_len[i]);
}

// One warp per object; lanes write the object's tiles strided.
__global__ void k_tiles_fill(uint32_t nobj, const uint32_t* __restrict__ tile_start,
                             uint32_t* __restrict__ tile_obj, uint32_t* __restrict__ tile_chunk) {
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= nobj) return;
  const uint32_t t0 = tile_start[warp], t1 = tile_start[warp + 1];
  for (uint32_t t = t0 + lane; t < t1; t += 32) {
    tile_obj[t] = warp;
    tile_chunk[t] = t - t0;
  }
}

// ------------------------------------------------------------------ validation
// Records must start on 16-byte boundaries and lie inside payload; keys / uids inside aux.
__global__ void k_validate(DevArena a, uint64_t payload_bytes, uint64_t aux_bytes, uint64_t nowner,
                           uint64_t ndep, uint32_t* __restrict__ bad) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.nobj) return;
  bool ok = (a.rec_off[i] & 15) == 0 && a.rec_off[i] + a.rec_len[i] <= payload_bytes;
  ok = ok && a.kind[i] <= NCD_KIND_CONFIGMAP;
  ok = ok && (uint64_t)a.key_off[i] + a.key_len[i] <= aux_bytes && a.key_len[i] <= 0xFFFFu;
  ok = ok && (uint64_t)a.uid_off[i] + a.uid_len[i] <= aux_bytes;
  ok = ok && a.own_start[i] <= a.own_start[i + 1] && a.own_start[i + 1] <= nowner;
  ok = ok && a.dep_start[i] <= a.dep_start[i + 1] && a.dep_start[i + 1] <= ndep;
  if (ok) {
    for (uint32_t o = a.own_start[i]; o < a.own_start[i + 1]; ++o)
      ok = ok && (uint64_t)a.own_uid_off[o] + a.own_uid_len[o] <= aux_bytes;
    for (uint32_t d = a.dep_start[i]; d < a.dep_start[i + 1]; ++d)
      ok = ok && (uint64_t)a.dep_key_off[d] + a.dep_key_len[d] <= aux_bytes &&
           (a.dep_kind[d] == NCD_KIND_SECRET || a.dep_kind[d] == NCD_KIND_CONFIGMAP);
  }
  if (!ok) atomicAdd(bad, 1u);
}

// ------------------------------------------------------------------ host side
void ncd_fill_dev_arena(const ncd_arena* a, DevArena* o) {
  const ncd_arena_view& d = a->d;
  o->payload = d.payload;
  o->rec_off = d.rec_off;
  o->rec_len = d.rec_len;
  o->kind = d.kind;
  o->key_off = d.key_off;
  o->key_len = d.key_len;
  o->uid_off = d.uid_off;
  o->uid_len = d.uid_len;
  o->own_start = d.own_start;
  o->own_is_template = d.own_is_template;
  o->own_uid_off = d.own_uid_off;
  o->own_uid_len = d.own_uid_len;
  o->dep_start = d.dep_start;
  o->dep_kind = d.dep_kind;
  o->dep_key_off = d.dep_key_off;
  o->dep_key_len = d.dep_key_len;
  o->aux = d.aux;
  o->meta = a->meta;
  o->table = a->table;
  o->table_mask = a->table_size ? a->table_size - 1 : 0;
  o->nobj = (uint32_t)a->nobj;
  o->uid_id = a->uid_id;
  o->own_packed = a->own_packed;
  o->join_cache = nullptr;
  o->join_cache_out = nullptr;
  o->join_line = nullptr;
  o->join_line_out = nullptr;
  o->join_rev_out = nullptr;
  o->h64 = a->h64;
  o->sha = a->sha;
  o->xxh = a->xxh;
}

// Index arrays are (re)allocated in stream order from the device's pool (kept cached, ncd_init): a
// commit of an arena that was committed before neither returns memory to the driver nor stalls the
// device the way cudaFree / cudaMalloc do — other streams (the uploads of the next arena,
// ncd_arenas_commit) keep running.
static thread_local cudaStream_t g_realloc_stream = nullptr;  // the committing context's stream (contexts of different GPUs commit from different threads)
template <typename T>
static int dev_realloc(T** p, uint64_t n) {
  if (*p) {
    cudaFreeAsync(*p, g_realloc_stream);
    *p = nullptr;
  }
  if (n == 0) n = 1;
  cudaError_t e = cudaMallocAsync((void**)p, n * sizeof(T), g_realloc_stream);
  if (e != cudaSuccess) {
    ncd_set_error("cudaMallocAsync of %llu bytes failed: %s", (unsigned long long)(n * sizeof(T)),
                  cudaGetErrorString(e));
    return NCD_E_NOMEM;
  }
  return NCD_OK;
}

// Tile lists (hash tiles and compare tiles) and the sum of record lengths: everything that
// depends on rec_len only.  Run by commit and again after records were replaced in place.
int ncd_rebuild_tiles(ncd_arena* a) {
  ncd_ctx* ctx = a->ctx;
  cudaStream_t st = ctx->stream;
  g_realloc_stream = st;
  const uint32_t n = (uint32_t)a->nobj;
  const unsigned blocks = (n + 255) / 256;
  DevArena da;
  ncd_fill_dev_arena(a, &da);
  uint32_t* d_small = nullptr;  // [0] tiles, [1] compare tiles, [2..3] u64 sum of rec_len, [4] longest record
  uint32_t* counts = nullptr;
  NCD_CUDA(cudaMallocAsync((void**)&d_small, 6 * sizeof(uint32_t), st));
  NCD_CUDA(cudaMemsetAsync(d_small, 0, 6 * sizeof(uint32_t), st));
  NCD_CUDA(cudaMallocAsync((void**)&counts, ((size_t)n + 1) * 4, st));
  if (!a->tile_start) NCD_TRY(dev_realloc(&a->tile_start, (uint64_t)n + 1));
  if (!a->ctile_start) NCD_TRY(dev_realloc(&a->ctile_start, (uint64_t)n + 1));
  if (n) k_tile_counts<<<blocks, 256, 0, st>>>(da, counts, (unsigned long long*)(d_small + 2), d_small + 4);
  NCD_TRY(ncd_exclusive_scan_u32(ctx, counts, a->tile_start, n, d_small + 0));
  if (n) k_ctile_counts<<<blocks, 256, 0, st>>>(da, counts);
  NCD_TRY(ncd_exclusive_scan_u32(ctx, counts, a->ctile_start, n, d_small + 1));
  uint32_t h_small[6];
  NCD_CUDA(cudaMemcpyAsync(h_small, d_small, sizeof(h_small), cudaMemcpyDeviceToHost, st));
  NCD_CUDA(cudaStreamSynchronize(st));
  memcpy(&a->len_sum, &h_small[2], 8);
  a->len_max = h_small[4];
  if (h_small[0] > a->tile_cap || !a->tile_obj) {
    a->tile_cap = h_small[0] + h_small[0] / 16;
    NCD_TRY(dev_realloc(&a->tile_obj, a->tile_cap));
    NCD_TRY(dev_realloc(&a->tile_chunk, a->tile_cap));
  }
  if (h_small[1] > a->ctile_cap || !a->ctile_obj) {
    a->ctile_cap = h_small[1] + h_small[1] / 16;
    NCD_TRY(dev_realloc(&a->ctile_obj, a->ctile_cap));
    NCD_TRY(dev_realloc(&a->ctile_chunk, a->ctile_cap));
  }
  a->ntiles = h_small[0];
  a->nctiles = h_small[1];
  if (n) {
    const unsigned wblocks = (unsigned)(((uint64_t)n * 32 + 255) / 256);
    k_tiles_fill<<<wblocks, 256, 0, st>>>(n, a->tile_start, a->tile_obj, a->tile_chunk);
    k_tiles_fill<<<wblocks, 256, 0, st>>>(n, a->ctile_start, a->ctile_obj, a->ctile_chunk);
  }
  NCD_CUDA(cudaFreeAsync(counts, st));
  NCD_CUDA(cudaFreeAsync(d_small, st));
  NCD_CUDA(cudaGetLastError());
  return NCD_OK;
}

static std::atomic<uint64_t> g_generation{0};  // contexts of different GPUs may commit from different threads

int ncd_build_index(ncd_arena* a) {
  ncd_ctx* ctx = a->ctx;
  a->generation = g_generation.fetch_add(1) + 1;  // keys may have changed: join caches built on the old index are stale
  a->jc_line_ok = false;
  cudaStream_t st = ctx->stream;
  g_realloc_stream = st;
  const uint32_t n = (uint32_t)a->nobj;
  const unsigned blocks = (n + 255) / 256;

  a->h64_valid = a->sha_valid = a->xxh_valid = false;
  NCD_TRY(dev_realloc(&a->meta, n));
  uint32_t tsize = 16;
  while ((uint64_t)tsize < 2ull * n) tsize <<= 1;
  a->table_size = tsize;
  NCD_TRY(dev_realloc(&a->table, tsize));
  NCD_CUDA(cudaMemsetAsync(a->table, 0xFF, (size_t)tsize * 8, st));

  // [0] bad records, [1] duplicate keys, [2] rows
  uint32_t* d_small = nullptr;
  NCD_CUDA(cudaMallocAsync((void**)&d_small, 8 * sizeof(uint32_t), st));
  NCD_CUDA(cudaMemsetAsync(d_small, 0, 8 * sizeof(uint32_t), st));

  DevArena da;
  ncd_fill_dev_arena(a, &da);
  uint32_t* counts = nullptr;
  NCD_CUDA(cudaMallocAsync((void**)&counts, ((size_t)n + 1) * 4, st));
  NCD_TRY(dev_realloc(&a->row_start, (uint64_t)n + 1));
  NCD_TRY(dev_realloc(&a->tile_start, (uint64_t)n + 1));
  NCD_TRY(dev_realloc(&a->ctile_start, (uint64_t)n + 1));
  a->tile_cap = a->ctile_cap = 0;

  if (n) {
    k_validate<<<blocks, 256, 0, st>>>(da, a->payload_bytes, a->aux_bytes, a->nowner, a->ndep, d_small);
    uint32_t bad = 0;
    NCD_CUDA(cudaMemcpyAsync(&bad, d_small, 4, cudaMemcpyDeviceToHost, st));
    NCD_CUDA(cudaStreamSynchronize(st));
    if (bad) {
      cudaFreeAsync(counts, st);
      cudaFreeAsync(d_small, st);
      ncd_set_error("arena commit: %u objects have records/keys outside the arena, a record not "
                    "16-byte aligned, or an unknown kind", bad);
      return NCD_E_INVALID;
    }
    {
      const uint64_t nstr = (uint64_t)n + a->nowner;
      if (nstr >= 0x7FFFFFFFull) {
        ncd_set_error("arena commit: objects + ownerReferences exceed 2^31");
        return NCD_E_CAPACITY;
      }
      NCD_TRY(dev_realloc(&a->uid_id, n));
      NCD_TRY(dev_realloc(&a->own_packed, a->nowner));
      uint32_t usize = 16;
      while ((uint64_t)usize < 2 * nstr) usize <<= 1;
      uint32_t* utable = nullptr;
      NCD_CUDA(cudaMallocAsync((void**)&utable, (size_t)usize * 4, st));
      NCD_CUDA(cudaMemsetAsync(utable, 0xFF, (size_t)usize * 4, st));
      const unsigned ublocks = (unsigned)((nstr + 255) / 256);
      k_uid_insert<<<ublocks, 256, 0, st>>>(da, (uint32_t)nstr, utable, usize - 1);
      k_uid_resolve<<<ublocks, 256, 0, st>>>(da, (uint32_t)nstr, utable, usize - 1, a->uid_id, a->own_packed);
      NCD_CUDA(cudaFreeAsync(utable, st));
    }
    ncd_fill_dev_arena(a, &da);  // uid_id / own_packed were (re)allocated above: the object lines carry tags made from them
    k_meta_build<<<blocks, 256, 0, st>>>(da, a->meta);
    k_table_build<<<blocks, 256, 0, st>>>(da, (unsigned long long*)a->table, d_small + 1);
    k_row_counts<<<blocks, 256, 0, st>>>(da, counts);
  }
  NCD_TRY(ncd_exclusive_scan_u32(ctx, counts, a->row_start, n, d_small + 2));
  uint32_t h_small[8];
  NCD_CUDA(cudaMemcpyAsync(h_small, d_small, sizeof(h_small), cudaMemcpyDeviceToHost, st));
  NCD_CUDA(cudaStreamSynchronize(st));
  a->dup_keys = h_small[1];
  a->nrows = h_small[2];

  NCD_TRY(dev_realloc(&a->row_obj, a->nrows));
  NCD_TRY(dev_realloc(&a->row_dep, a->nrows));
  NCD_TRY(dev_realloc(&a->row_cidx, a->nrows));
  if (n) k_rows_fill<<<blocks, 256, 0, st>>>(da, a->row_start, a->row_obj, a->row_dep, a->row_cidx);
  NCD_TRY(ncd_rebuild_tiles(a));
  NCD_CUDA(cudaFreeAsync(counts, st));
  NCD_CUDA(cudaFreeAsync(d_small, st));
  NCD_CUDA(cudaGetLastError());
  NCD_CUDA(cudaStreamSynchronize(st));
  return NCD_OK;
}
