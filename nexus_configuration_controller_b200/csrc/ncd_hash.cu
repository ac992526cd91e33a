// ncd_hash.cu — the hash path BASELINE.json's north_star asks for.  The reference never
// hashes (controller.go:539,600,725,795 call reflect.DeepEqual), so these digests have no
// reference counterpart: they are specified in DESIGN.md ("NCD-H64") and restated on the
// CPU in oracle/ncd_oracle.cpp, bit for bit.
//
//   k_h64_tiles    striped 64-bit hash, one warp per 4 KB tile, HBM bound (all arenas of a scan in one launch)
//   k_h64_final    one thread per object folds the tile sums and the length
//   k_sha256       FIPS 180-4 SHA-256, persistent lanes pulling records, integer-pipe bound
#include <vector>

#include "ncd_common.cuh"
#include "ncd_tma.cuh"

// ------------------------------------------------------------------ NCD-H64
// Per 16-byte vector v of a tile (words d0, d1, little endian) with keys k0[v], k1[v]:
//   A += lo32(d0 ^ k0) * hi32(d0 ^ k0) + d1
//   B += lo32(d1 ^ k1) * hi32(d1 ^ k1) + d0
// (the accumulate step of XXH3, whose sums commute, so lanes and tiles can run in any
// order).  Tile t then contributes T = fmix64(fmix64(A + sm(2t)) + rotl(B, 31) + sm(2t+1))
// and the digest is fmix64(sum of T + len * 0x9E3779B185EBCA87).
__device__ __forceinline__ uint64_t rotl64(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }

// One launch hashes every arena of a scan: a job per arena, tiles (records) numbered across
// jobs, so there is one grid, one tail and no launch gaps instead of one of each per cluster copy.
struct HashJob {
  const uint8_t* payload;
  const uint64_t* rec_off;
  const uint32_t* rec_len;
  const uint32_t* tile_obj;
  const uint32_t* tile_chunk;
  unsigned long long* acc;  // NCD-H64: per-record accumulator, becomes the digest
  uint32_t* sha;            // SHA-256 digests
  uint64_t* xxh;            // XXH64 digests
  uint64_t tile_base;       // global index of this job's first tile
  uint64_t ntiles;
  uint64_t rec_base;        // global index of this job's first record
  uint64_t nobj;
  const uint32_t* rec_list; // non-null: hash only these records (entry i of the job is record rec_list[i])
};
constexpr int MAX_HASH_JOBS = 1025;

// largest j with base[j] <= x (base is ascending, base[0] == 0)
__device__ __forceinline__ int find_job(const uint64_t* base, int njobs, uint64_t x) {
  int lo = 0, hi = njobs;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (base[mid] <= x) lo = mid;
    else hi = mid;
  }
  return lo;
}

// Contribution T of one tile (chunk-th 4 KiB of a record, nbytes of it real) to the record's digest;
// the whole warp calls it, every lane returns T.  keys: (k0, k1) per vector position, in shared memory.
// mix: splitmix64(i) for i < NCD_H64_MIXTAB in shared memory (the per-tile constants of the first 256 tiles of a record,
// which a record of up to 1 MiB never leaves).
constexpr uint32_t NCD_H64_MIXTAB = 512;
__device__ __forceinline__ uint64_t h64_mix(const uint64_t* mix, uint64_t i) {
  return i < NCD_H64_MIXTAB ? mix[i] : ncd_splitmix64(i);
}
__device__ __forceinline__ uint64_t h64_tile(const uint8_t* tile, uint32_t nbytes, uint32_t chunk, const uint4* keys,
                                             const uint64_t* mix, uint32_t lane) {
  constexpr int U = NCD_TILE_VECS / 32;
  const uint32_t nvec = (nbytes + 15) >> 4;
  const uint32_t tail = nbytes & 15;
  const uint4* p = reinterpret_cast<const uint4*>(tile);
  uint4 d[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const uint32_t v = lane + 32 * u;
    d[u] = v < nvec ? ncd_ldg_stream(p + v) : make_uint4(0, 0, 0, 0);
  }
  // bytes past rec_len in the record's last vector do not belong to the value: clear them once,
  // outside the accumulate loop.  `tail` and the vector's place are the same for the whole warp, so
  // a tile without a ragged end (two of three) skips this as a branch, not as predicated-off instructions.
  if (tail) {
    const uint32_t last = nvec - 1, ut = last >> 5, ll = last & 31;
    uint32_t m[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int nb = (int)tail - 4 * i;
      m[i] = nb <= 0 ? 0u : (nb < 4 ? (1u << (8 * nb)) - 1u : 0xffffffffu);
    }
    if (lane == ll) {
#pragma unroll
      for (int u = 0; u < U; ++u)
        if ((uint32_t)u == ut) d[u] = make_uint4(d[u].x & m[0], d[u].y & m[1], d[u].z & m[2], d[u].w & m[3]);
    }
  }
  uint64_t A = 0, B = 0;
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const uint32_t v = lane + 32 * u;
    if (v < nvec) {
      const uint4 x = d[u];
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
  return ncd_fmix64(ncd_fmix64(A + h64_mix(mix, 2ull * chunk)) + rotl64(B, 31) + h64_mix(mix, 2ull * chunk + 1));
}
__device__ __forceinline__ void h64_fill_mix(uint64_t* mix) {
  for (uint32_t i = threadIdx.x; i < NCD_H64_MIXTAB; i += blockDim.x) mix[i] = ncd_splitmix64(i);
}

__device__ __forceinline__ void h64_fill_keys(uint4* keys) {
  for (uint32_t v = threadIdx.x; v < NCD_TILE_VECS; v += blockDim.x) {
    const uint64_t k0 = ncd_splitmix64(0x6e63642d68363400ull + 2 * v);
    const uint64_t k1 = ncd_splitmix64(0x6e63642d68363400ull + 2 * v + 1);
    keys[v] = make_uint4((uint32_t)k0, (uint32_t)(k0 >> 32), (uint32_t)k1, (uint32_t)(k1 >> 32));
  }
}

__global__ void __launch_bounds__(256)
k_h64_tiles(const HashJob* __restrict__ jobs, int njobs, uint64_t ntiles) {
  __shared__ uint4 keys[NCD_TILE_VECS];  // (k0, k1) per vector position
  __shared__ uint64_t mix[NCD_H64_MIXTAB];
  extern __shared__ uint64_t s_dyn[];    // [njobs] first global tile of each job, then the jobs themselves
  uint64_t* s_base = s_dyn;
  HashJob* s_jobs = reinterpret_cast<HashJob*>(s_dyn + njobs);
  for (int j = threadIdx.x; j < njobs; j += blockDim.x) {
    s_base[j] = jobs[j].tile_base;
    s_jobs[j] = jobs[j];
  }
  h64_fill_keys(keys);
  h64_fill_mix(mix);
  __syncthreads();
  const uint32_t lane = threadIdx.x & 31;
  const uint64_t warp0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  int cur = 0;  // the job of the warp's last tile: its next tile, nwarps further on, is almost always in the same arena
  for (uint64_t g = warp0; g < ntiles; g += nwarps) {
    if (g < s_base[cur] || (cur + 1 < njobs && g >= s_base[cur + 1])) cur = find_job(s_base, njobs, g);
    const HashJob& a = s_jobs[cur];
    const uint64_t t = g - a.tile_base;
    unsigned long long* acc = a.acc;
    const uint32_t c = a.tile_obj[t], chunk = a.tile_chunk[t];
    const uint32_t len = a.rec_len[c];
    const uint64_t toff = (uint64_t)chunk * NCD_TILE_BYTES;
    const uint32_t nbytes = min(NCD_TILE_BYTES, len - (uint32_t)toff);
    const uint64_t T = h64_tile(a.payload + a.rec_off[c] + toff, nbytes, chunk, keys, mix, lane);
    if (lane == 0) atomicAdd(&acc[c], (unsigned long long)T);
  }
}

__global__ void k_h64_final(const HashJob* __restrict__ jobs, int njobs, uint64_t nrecs) {
  extern __shared__ uint64_t s_base[];  // [njobs] first global record of each job
  for (int j = threadIdx.x; j < njobs; j += blockDim.x) s_base[j] = jobs[j].rec_base;
  __syncthreads();
  const uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= nrecs) return;
  const HashJob& a = jobs[find_job(s_base, njobs, g)];
  const uint64_t i = g - a.rec_base;
  a.acc[i] = ncd_fmix64((uint64_t)a.acc[i] + (uint64_t)a.rec_len[i] * 0x9E3779B185EBCA87ull);
}

// NCD-H64 of a few named records (after a value sync): block per record, its tiles dealt to the
// warps, no atomics and no tile list — the digest equals the one k_h64_tiles + k_h64_final produce.
__global__ void __launch_bounds__(256)
k_h64_records(const uint32_t* __restrict__ idx, const uint8_t* __restrict__ payload, const uint64_t* __restrict__ rec_off,
              const uint32_t* __restrict__ rec_len, unsigned long long* __restrict__ out) {
  __shared__ uint4 keys[NCD_TILE_VECS];
  __shared__ uint64_t mix[NCD_H64_MIXTAB];
  __shared__ uint64_t part[8];
  h64_fill_keys(keys);
  h64_fill_mix(mix);
  __syncthreads();
  const uint32_t c = idx[blockIdx.x];
  const uint32_t len = rec_len[c];
  const uint8_t* rec = payload + rec_off[c];
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t ntiles = (len + NCD_TILE_BYTES - 1) / NCD_TILE_BYTES;
  uint64_t sum = 0;
  for (uint32_t chunk = warp; chunk < ntiles; chunk += 8) {
    const uint32_t toff = chunk * NCD_TILE_BYTES;
    sum += h64_tile(rec + toff, min(NCD_TILE_BYTES, len - toff), chunk, keys, mix, lane);
  }
  if (lane == 0) part[warp] = sum;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint64_t t = 0;
    for (int w = 0; w < 8; ++w) t += part[w];
    out[c] = ncd_fmix64(t + (uint64_t)len * 0x9E3779B185EBCA87ull);
  }
}

// ------------------------------------------------------------------ SHA-256 (FIPS 180-4)
__constant__ uint32_t K256[64] = {
    0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5,
    0xd807aa98, 0x12835b01, 0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174,
    0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da,
    0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967,
    0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85,
    0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070,
    0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3,
    0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2};

__device__ __forceinline__ uint32_t rotr32(uint32_t x, int r) { return __funnelshift_r(x, x, r); }
__device__ __forceinline__ uint32_t bswap32(uint32_t x) { return __byte_perm(x, 0, 0x0123); }

__device__ __forceinline__ void sha256_block(uint32_t st[8], uint32_t w[16]) {
  uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
#pragma unroll
  for (int i = 0; i < 64; ++i) {
    if (i >= 16) {
      const uint32_t w15 = w[(i + 1) & 15], w2 = w[(i + 14) & 15];
      const uint32_t s0 = rotr32(w15, 7) ^ rotr32(w15, 18) ^ (w15 >> 3);
      const uint32_t s1 = rotr32(w2, 17) ^ rotr32(w2, 19) ^ (w2 >> 10);
      w[i & 15] = w[i & 15] + s0 + w[(i + 9) & 15] + s1;
    }
    const uint32_t S1 = rotr32(e, 6) ^ rotr32(e, 11) ^ rotr32(e, 25);
    const uint32_t ch = (e & f) ^ (~e & g);
    const uint32_t t1 = h + S1 + ch + K256[i] + w[i & 15];
    const uint32_t S0 = rotr32(a, 2) ^ rotr32(a, 13) ^ rotr32(a, 22);
    const uint32_t mj = (a & b) ^ (a & c) ^ (b & c);
    const uint32_t t2 = S0 + mj;
    h = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
  }
  st[0] += a; st[1] += b; st[2] += c; st[3] += d; st[4] += e; st[5] += f; st[6] += g; st[7] += h;
}

// Persistent lanes: every lane owns one record at a time and takes the next record index from
// a global counter when it finishes, so lanes of a warp stay busy although record lengths differ
// by more than 10x (a fixed record-per-thread mapping idles most lanes while the longest record
// of the warp finishes).  Every loop iteration is one 64-byte block for every lane: a message
// block, the padded last block, or the extra length-only block (FIPS 180-4 section 5.1.1).
// Records start on 16-byte boundaries and are zero padded to one, so blocks are read as 16-byte
// vectors; bytes past rec_len are masked before use.
__global__ void __launch_bounds__(128)
k_sha256(const HashJob* __restrict__ jobs, int njobs, uint32_t nrecs, unsigned int* __restrict__ next) {
  extern __shared__ uint64_t s_base[];  // [njobs] first global record of each job
  for (int j = threadIdx.x; j < njobs; j += blockDim.x) s_base[j] = jobs[j].rec_base;
  __syncthreads();
  uint32_t st[8];
  uint32_t w[16];
  const uint4* p = nullptr;
  uint32_t* out = nullptr;
  uint32_t rec = 0, len = 0, pos = 0;
  bool have = false, extra = false, drained = false;
  for (;;) {
    if (!have && !drained) {
      const uint32_t g = atomicAdd(next, 1u);
      drained = g >= nrecs;
      if (!drained) {
        const HashJob& a = jobs[find_job(s_base, njobs, g)];
        rec = (uint32_t)(g - a.rec_base);
        if (a.rec_list) rec = a.rec_list[rec];
        out = a.sha;
        len = a.rec_len[rec];
        p = reinterpret_cast<const uint4*>(a.payload + a.rec_off[rec]);
        pos = 0;
        extra = false;
        have = true;
        st[0] = 0x6a09e667; st[1] = 0xbb67ae85; st[2] = 0x3c6ef372; st[3] = 0xa54ff53a;
        st[4] = 0x510e527f; st[5] = 0x9b05688c; st[6] = 0x1f83d9ab; st[7] = 0x5be0cd19;
      }
    }
    if (!__any_sync(0xffffffffu, have)) break;  // every lane reaches this point in every iteration
    if (!have) continue;  // this lane ran out of records; the others finish theirs
    bool last = false;
    const uint32_t rem = len - pos;  // unused when extra
    const uint32_t navail = extra ? 0u : min(4u, (rem + 15) >> 4);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      uint4 v = make_uint4(0, 0, 0, 0);
      if ((uint32_t)q < navail) v = ncd_ldg_stream(p + (pos >> 4) + q);
      w[4 * q + 0] = v.x; w[4 * q + 1] = v.y; w[4 * q + 2] = v.z; w[4 * q + 3] = v.w;
    }
    if (extra) {
      last = true;
    } else if (rem < 64) {
      // little-endian word k holds message bytes 4k..4k+3: keep the first rem bytes, then 0x80
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        const int nb = (int)rem - 4 * k;
        uint32_t x = w[k];
        if (nb <= 0) x = 0;
        else if (nb < 4) x &= (1u << (8 * nb)) - 1u;
        if (nb >= 0 && nb < 4) x |= 0x80u << (8 * nb);
        w[k] = x;
      }
      if (rem < 56) last = true;
      else extra = true;
    }
#pragma unroll
    for (int k = 0; k < 16; ++k) w[k] = bswap32(w[k]);
    if (last) {
      const uint64_t bits = (uint64_t)len * 8;
      w[14] = (uint32_t)(bits >> 32);
      w[15] = (uint32_t)bits;
    }
    sha256_block(st, w);
    pos += 64;
    if (last) {
      // digest bytes in standard order: big-endian words
      uint4* o = reinterpret_cast<uint4*>(out + (size_t)rec * 8);
      o[0] = make_uint4(bswap32(st[0]), bswap32(st[1]), bswap32(st[2]), bswap32(st[3]));
      o[1] = make_uint4(bswap32(st[4]), bswap32(st[5]), bswap32(st[6]), bswap32(st[7]));
      have = false;
    }
  }
}

// ------------------------------------------------------------------ XXH64 (xxHash specification, seed 0)
// The standard function, for callers that want digests other tools can reproduce.  It is serial
// per record (four accumulators over 32-byte stripes), so the parallelism is across records: the
// same persistent-lane scheme as k_sha256, one stripe per lane per iteration.
// 64-bit arithmetic spelled in 32-bit halves, the way the integer pipes execute it: a rotate is two
// funnel shifts, a product by a constant one wide multiply-add and two multiply-adds.
__device__ __forceinline__ uint64_t pack64(uint32_t lo, uint32_t hi) { return (uint64_t)lo | ((uint64_t)hi << 32); }
template <int R>
__device__ __forceinline__ uint64_t xxh_rotl(uint64_t x) {
  static_assert(R > 0 && R < 32, "rotations of XXH64 used on the stripe path are below 32");
  const uint32_t lo = (uint32_t)x, hi = (uint32_t)(x >> 32);
  return pack64(__funnelshift_l(hi, lo, R), __funnelshift_l(lo, hi, R));
}
// acc + x * C  (mod 2^64)
__device__ __forceinline__ uint64_t xxh_mad(uint32_t xlo, uint32_t xhi, uint64_t C, uint64_t acc) {
  const uint32_t clo = (uint32_t)C, chi = (uint32_t)(C >> 32);
  const uint64_t p = (uint64_t)xlo * clo + acc;
  return pack64((uint32_t)p, (uint32_t)(p >> 32) + xlo * chi + xhi * clo);
}
__device__ __forceinline__ uint64_t xxh_round2(uint64_t acc, uint32_t inlo, uint32_t inhi) {
  acc = xxh_rotl<31>(xxh_mad(inlo, inhi, 14029467366897019727ull, acc));
  return xxh_mad((uint32_t)acc, (uint32_t)(acc >> 32), 11400714785074694791ull, 0);
}
__device__ __forceinline__ uint64_t xxh_round(uint64_t acc, uint64_t in) {
  return xxh_round2(acc, (uint32_t)in, (uint32_t)(in >> 32));
}
__device__ __forceinline__ uint64_t xxh_merge(uint64_t h, uint64_t v) {
  return (h ^ xxh_round(0, v)) * 11400714785074694791ull + 9650029242287828579ull;
}

// Each lane owns one record, but a lane reading its own record 32 bytes at a time turns the pass
// into random-sector traffic (measured 0.8 TB/s), and 128-byte windows gathered with LDG.128 and
// shuffles reach 3.0 TB/s.  So every lane moves the next XXH_WIN contiguous bytes of its record
// into its own shared-memory row with one bulk copy (cp.async.bulk, completion counted in bytes on
// the warp's mbarrier), and then runs the stripes out of that row.  Rows are padded by one vector
// so that LDS.128 at a row stride is conflict-free.
// What held this kernel at 0.81 of the HBM peak was not the staging but DIVERGENCE: with 32 records
// of different lengths per warp, almost every round had some lane in its record's last window (a
// second, generic stripe loop ran beside the unrolled one) and some lane finishing (a hundred-odd
// instructions for one active lane): 3.8 executed instructions per byte for 1.25 of hashing.  Now
// the stripe loop is one predicated stream for all lanes and finishing is batched (see `flush`):
// 0.89-0.90.  Staging variants measured on top of that (profiles/r02_xxh64_variants.txt): windows
// of 512 bytes copied by the warp together, 16 bytes per lane (cp.async: every request one contiguous
// window) 0.90; one bulk copy per lane 0.89; two row sets of 256 bytes with the next window in flight
// while one is hashed 0.87 / 0.79 — larger windows beat overlap at the same shared memory.  The FETCH
// side runs ahead of the HASH side by NBUF - 1 windows; what a window holds (record length,
// position, where the digest goes) travels with it in registers.
#ifndef NCD_XXH_WIN
#define NCD_XXH_WIN 512
#endif
#ifndef NCD_XXH_NBUF
#define NCD_XXH_NBUF 1  // row sets per warp: 1 = stage, wait, hash; 2 = the next window in flight while one is hashed
#endif
#ifndef NCD_XXH_STAGE
#define NCD_XXH_STAGE 1  // 0: one bulk copy per lane (its own record's window); 1: the warp copies window after window
                         //    together, 16 bytes per lane (cp.async), so every request is one contiguous window
#endif
constexpr int XXH_WIN = NCD_XXH_WIN;       // bytes of one record staged per window (multiple of 32)
constexpr int XXH_ROW = XXH_WIN / 16 + 1;  // uint4 per shared-memory row
constexpr int XXH_NBUF = NCD_XXH_NBUF;
static_assert(XXH_NBUF == 1 || XXH_NBUF == 2, "one or two row sets");
static_assert(NCD_XXH_STAGE == 0 || XXH_WIN == 256 || XXH_WIN == 512, "cooperative staging covers 1 or 2 windows per instruction");
constexpr int XXH_Q = 16;                  // records a warp takes from the global counter at a time
constexpr int XXH_WARPS = 4;

struct XxhSmem {
  uint4 win[XXH_WARPS][XXH_NBUF][32][XXH_ROW];
  const unsigned char* qbase[XXH_WARPS][XXH_Q];
  uint64_t* qout[XXH_WARPS][XXH_Q];
  uint32_t qlen[XXH_WARPS][XXH_Q];
  uint4 tail[XXH_WARPS][32][2];  // the last < 32 bytes of a record parked for finishing, per lane
  uint64_t bar[XXH_WARPS][2];    // bulk-copy staging only
};

struct XxhWindow {  // what one staged window holds, per lane
  uint64_t* out;
  uint32_t len, pos;  // the record's length; the window starts at byte `pos` of it
  bool valid;
};

__global__ void __launch_bounds__(128)
k_xxh64(const HashJob* __restrict__ jobs, int njobs, uint32_t nrecs, unsigned int* __restrict__ next) {
  constexpr uint64_t P1 = 11400714785074694791ull, P2 = 14029467366897019727ull, P3 = 1609587929392839161ull,
                     P4 = 9650029242287828579ull, P5 = 2870177450012600261ull;
  extern __shared__ __align__(128) uint8_t xxh_raw[];
  XxhSmem* sm = reinterpret_cast<XxhSmem*>(xxh_raw);
  uint64_t* s_base = reinterpret_cast<uint64_t*>(sm + 1);  // [njobs] first global record of each job
  for (int j = threadIdx.x; j < njobs; j += blockDim.x) s_base[j] = jobs[j].rec_base;
  if (threadIdx.x < 2 * XXH_WARPS) mbar_init(&sm->bar[threadIdx.x >> 1][threadIdx.x & 1], 1);
  if (threadIdx.x == 0) asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const uint64_t pol = l2_evict_first_policy();
  // Records are handed out to the warp XXH_Q at a time (one global atomic and one round of index
  // loads per XXH_Q records, not per record); inside the warp, lanes whose fetch side ran out take
  // the next entries in ballot order.  head / qn / drained are warp-uniform.
  const unsigned char** qb = sm->qbase[wid];
  uint32_t* ql = sm->qlen[wid];
  uint64_t** qo = sm->qout[wid];
  uint32_t head = 0, qn = 0;
  bool drained = false;
  // fetch side: the record whose windows are being staged
  const unsigned char* f_base = nullptr;
  uint64_t* f_out = nullptr;
  uint32_t f_len = 0, f_pos = 0;
  bool f_have = false;
  // hash side
  uint64_t v1 = 0, v2 = 0, v3 = 0, v4 = 0;
  uint32_t phase[2] = {0, 0};

  // Stages the next window of every lane into row set b.  Returns the bytes the warp asked for
  // (warp-uniform; 0: nothing was armed, nobody waits).
  auto fetch = [&](int b, XxhWindow& w) -> uint32_t {
    for (int pass = 0; pass < 2; ++pass) {
      const unsigned m = __ballot_sync(0xffffffffu, !f_have);
      if (!m) break;
      const uint32_t rank = __popc(m & ((1u << lane) - 1u)), avail = qn - head;
      if (!f_have && rank < avail) {
        f_base = qb[head + rank];
        f_len = ql[head + rank];
        f_out = qo[head + rank];
        f_pos = 0;
        f_have = true;
      }
      head += min((uint32_t)__popc(m), avail);
      if ((uint32_t)__popc(m) <= avail || drained || pass) break;
      // queue empty and lanes still idle: next batch
      __syncwarp();
      uint32_t g0 = 0;
      if (lane == 0) g0 = atomicAdd(next, (unsigned)XXH_Q);
      g0 = __shfl_sync(0xffffffffu, g0, 0);
      qn = g0 < nrecs ? min((uint32_t)XXH_Q, nrecs - g0) : 0u;
      drained = qn < (uint32_t)XXH_Q;
      head = 0;
      for (uint32_t i = lane; i < qn; i += 32) {
        const HashJob& a = jobs[find_job(s_base, njobs, g0 + i)];
        uint32_t rec = (uint32_t)(g0 + i - a.rec_base);
        if (a.rec_list) rec = a.rec_list[rec];
        qb[i] = a.payload + a.rec_off[rec];
        ql[i] = a.rec_len[rec];
        qo[i] = a.xxh + rec;
      }
      __syncwarp();
    }
    // 16-byte granular (the arena pads records to 16)
    const uint32_t want = f_have ? min((uint32_t)XXH_WIN, (f_len - f_pos + 15u) & ~15u) : 0u;
    const uint32_t total = __reduce_add_sync(0xffffffffu, want);
    w.valid = f_have;
    w.out = f_out;
    w.len = f_len;
    w.pos = f_pos;
#if NCD_XXH_STAGE == 0
    if (total) {
      uint64_t* bar = &sm->bar[wid][b];
      if (lane == 0) mbar_arrive_expect_tx(bar, total);
      __syncwarp();
      if (want) bulk_g2s(sm->win[wid][b][lane], f_base + f_pos, want, bar, pol);
    }
#else
    {
      // lane l copies chunk (l % C) of the window of record (l / C) of each group of 32 / C records
      constexpr int C = XXH_WIN / 16, PER = 32 / C;  // chunks per window, windows per instruction
      const unsigned long long src = (unsigned long long)(f_base + f_pos);
      const uint32_t chunk = lane % C;
#pragma unroll
      for (int g = 0; g < 32 / PER; ++g) {
        const int r = g * PER + lane / C;
        const unsigned long long a = __shfl_sync(0xffffffffu, src, r);
        const uint32_t wr = __shfl_sync(0xffffffffu, want, r);
        if (chunk * 16u < wr) {
          const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&sm->win[wid][b][r][chunk]);
          asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;" ::"r"(dst), "l"(a + chunk * 16ull), "l"(pol)
                       : "memory");
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    }
#endif
    if (f_have) {
      if (f_len - f_pos <= (uint32_t)XXH_WIN) f_have = false;  // that was the record's last window (its tail included)
      else f_pos += XXH_WIN;
    }
    return total;
  };

  // Hashes the window staged in row set b.
  // A record's last steps (merging the accumulators, its tail of < 32 bytes, the avalanche) are a
  // hundred-odd instructions that only the finishing lane needs; run when each lane gets there, the
  // whole warp pays for them 32 times per generation of records.  So a finishing lane only parks
  // what they need (accumulators and length in registers, the tail in shared memory) and goes on
  // with its next record; the parked records are finished TOGETHER, when a lane that still has
  // one parked finishes another (and at the end).
  uint64_t s_v1 = 0, s_v2 = 0, s_v3 = 0, s_v4 = 0;
  uint64_t* s_out = nullptr;
  uint32_t s_len = 0;
  bool s_full = false;
  auto flush = [&]() {
    if (s_full) {
      uint64_t h;
      if (s_len >= 32) {
        h = rotl64(s_v1, 1) + rotl64(s_v2, 7) + rotl64(s_v3, 12) + rotl64(s_v4, 18);
        h = xxh_merge(h, s_v1);
        h = xxh_merge(h, s_v2);
        h = xxh_merge(h, s_v3);
        h = xxh_merge(h, s_v4);
      } else {
        h = P5;
      }
      h += (uint64_t)s_len;
      const uint32_t rem = s_len & 31u;
      const uint32_t* t = reinterpret_cast<const uint32_t*>(sm->tail[wid][lane]);  // the tail, as words
      uint32_t o = 0;  // byte offset inside the tail; bytes past rem (padding) are never used
      for (; o + 8 <= rem; o += 8) {
        h ^= xxh_round(0, (uint64_t)t[o >> 2] | ((uint64_t)t[(o >> 2) + 1] << 32));
        h = rotl64(h, 27) * P1 + P4;
      }
      if (o + 4 <= rem) {
        h ^= (uint64_t)t[o >> 2] * P1;
        h = rotl64(h, 23) * P2 + P3;
        o += 4;
      }
      for (; o < rem; ++o) {
        h ^= (uint64_t)((t[o >> 2] >> (8 * (o & 3))) & 0xffu) * P5;
        h = rotl64(h, 11) * P1;
      }
      h ^= h >> 33;
      h *= P2;
      h ^= h >> 29;
      h *= P3;
      h ^= h >> 32;
      *s_out = h;
      s_full = false;
    }
  };

  // Hashes the window staged in row set b.  One instruction stream for every lane: the stripes a
  // lane's window does not have (a record's last window, a lane without a record) are computed on
  // whatever the row holds and discarded.
  auto hash = [&](int b, const XxhWindow& w, uint32_t armed) {
#if NCD_XXH_STAGE == 0
    if (armed) {
      mbar_wait(&sm->bar[wid][b], phase[b]);
      phase[b] ^= 1u;
    }
#else
    // this row set's group is the oldest but (NBUF - 1): every lane's part of it has landed, then the warp's
    if (XXH_NBUF == 2) asm volatile("cp.async.wait_group 1;" ::: "memory");
    else asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    (void)armed;
#endif
    const uint4* row = sm->win[wid][b][lane];
    const uint32_t len = w.len;
    if (w.valid && w.pos == 0) {
      v1 = P1 + P2;
      v2 = P2;
      v3 = 0;
      v4 = 0ull - P1;
    }
    const uint32_t nfull = w.valid ? min((len - w.pos) >> 5, (uint32_t)(XXH_WIN / 32)) : 0u;  // whole stripes in this window
#pragma unroll
    for (int k = 0; k < XXH_WIN / 32; ++k) {
      const uint4 x = row[2 * k], y = row[2 * k + 1];
      const bool on = (uint32_t)k < nfull;
      const uint64_t n1 = xxh_round2(v1, x.x, x.y), n2 = xxh_round2(v2, x.z, x.w);
      const uint64_t n3 = xxh_round2(v3, y.x, y.y), n4 = xxh_round2(v4, y.z, y.w);
      v1 = on ? n1 : v1;
      v2 = on ? n2 : v2;
      v3 = on ? n3 : v3;
      v4 = on ? n4 : v4;
    }
    const uint32_t w_used = 32 * nfull;  // bytes of the window consumed
    const uint32_t rem = len - (w.pos + w_used);
    const bool fin = w.valid && rem < 32 && w_used + rem <= XXH_WIN;  // the tail is inside this window: the record ends here
    if (__any_sync(0xffffffffu, fin && s_full)) flush();
    if (fin) {
      const uint32_t ti = min(w_used >> 4, (uint32_t)(XXH_WIN / 16 - 1));  // rem == 0 when the window was used up
      sm->tail[wid][lane][0] = row[ti];
      sm->tail[wid][lane][1] = row[ti + 1];
      s_v1 = v1;
      s_v2 = v2;
      s_v3 = v3;
      s_v4 = v4;
      s_len = len;
      s_out = w.out;
      s_full = true;
    }
  };

  XxhWindow w0, w1;
  if (XXH_NBUF == 1) {
    for (;;) {
      const uint32_t armed = fetch(0, w0);
      if (!__any_sync(0xffffffffu, w0.valid)) break;
      hash(0, w0, armed);
      __syncwarp();  // every lane is done reading the rows before they are staged again
    }
  } else {
    uint32_t armed0 = fetch(0, w0), armed1;
    for (;;) {
      armed1 = fetch(XXH_NBUF - 1, w1);
      if (!__any_sync(0xffffffffu, w0.valid)) break;  // the fetch side ran dry one window ago: w1 is empty too
      hash(0, w0, armed0);
      __syncwarp();  // every lane is done reading row set 0 before it is staged again
      armed0 = fetch(0, w0);
      if (!__any_sync(0xffffffffu, w1.valid)) break;
      hash(XXH_NBUF - 1, w1, armed1);
      __syncwarp();
    }
  }
#if NCD_XXH_STAGE == 1
  asm volatile("cp.async.wait_group 0;" ::: "memory");
#endif
  flush();
}

// ------------------------------------------------------------------ host side
// Digests of every arena of the list that does not have valid ones yet, in ONE pass of kernels.
int ncd_hash_launch_many(ncd_ctx* ctx, ncd_arena* const* arenas, int count, uint32_t mode, uint32_t* launches) {
  cudaStream_t st = ctx->stream;
  if (mode != NCD_MODE_HASH64 && mode != NCD_MODE_SHA256 && mode != NCD_MODE_XXH64) {
    ncd_set_error("hash: unknown mode %u", mode);
    return NCD_E_INVALID;
  }
  HashJob jobs[MAX_HASH_JOBS];
  int njobs = 0;
  uint64_t tiles = 0, recs = 0;
  for (int k = 0; k < count; ++k) {
    ncd_arena* a = arenas[k];
    bool seen = false;
    for (int q = 0; q < k; ++q) seen = seen || arenas[q] == a;  // the same arena may be listed twice
    if (seen || (mode == NCD_MODE_HASH64 ? a->h64_valid : mode == NCD_MODE_SHA256 ? a->sha_valid : a->xxh_valid)) continue;
    if (njobs == MAX_HASH_JOBS) {
      ncd_set_error("hash: more than %d arenas in one scan", MAX_HASH_JOBS);
      return NCD_E_CAPACITY;
    }
    const uint64_t n = a->nobj;
    // digest buffers are allocated once and kept across commits: sized for the arena's object
    // CAPACITY, because a later commit / recommit may hold more objects than the first hashed one
    const size_t dcap = (size_t)(a->nobj_cap > n ? a->nobj_cap : (n ? n : 1));
    if (mode == NCD_MODE_HASH64) {
      if (!a->h64) NCD_CUDA(cudaMalloc((void**)&a->h64, dcap * 8));
      NCD_CUDA(cudaMemsetAsync(a->h64, 0, (size_t)(n ? n : 1) * 8, st));
      a->h64_valid = true;
    } else if (mode == NCD_MODE_SHA256) {
      if (!a->sha) NCD_CUDA(cudaMalloc((void**)&a->sha, dcap * 32));
      a->sha_valid = true;
    } else {
      if (!a->xxh) NCD_CUDA(cudaMalloc((void**)&a->xxh, dcap * 8));
      a->xxh_valid = true;
    }
    HashJob& j = jobs[njobs++];
    j.payload = a->d.payload;
    j.rec_off = a->d.rec_off;
    j.rec_len = a->d.rec_len;
    j.tile_obj = a->tile_obj;
    j.tile_chunk = a->tile_chunk;
    j.acc = (unsigned long long*)a->h64;
    j.sha = a->sha;
    j.xxh = a->xxh;
    j.tile_base = 0;  // NCD-H64 tiles are launched per arena (below), so tile indices are local
    j.ntiles = a->ntiles;
    j.rec_base = recs;
    j.nobj = n;
    j.rec_list = nullptr;
    tiles += a->ntiles;
    recs += n;
  }
  if (njobs == 0 || recs == 0) return NCD_OK;
  if (recs >= 0xFFFFFFFFull) {
    ncd_set_error("hash: more than 2^32 records in one pass");
    return NCD_E_CAPACITY;
  }
  if (njobs > ctx->d_jobs_cap) {
    if (ctx->d_jobs) cudaFree(ctx->d_jobs);
    ctx->d_jobs = nullptr;
    const int cap = njobs < 32 ? 32 : njobs;
    NCD_CUDA(cudaMalloc(&ctx->d_jobs, sizeof(HashJob) * cap));
    ctx->d_jobs_cap = cap;
  }
  // pageable source: staged by the runtime before the call returns
  NCD_CUDA(cudaMemcpyAsync(ctx->d_jobs, jobs, sizeof(HashJob) * njobs, cudaMemcpyHostToDevice, st));
  const HashJob* dj = (const HashJob*)ctx->d_jobs;
  const size_t smem = (size_t)njobs * sizeof(uint64_t);
  if (mode == NCD_MODE_HASH64) {
    // one launch per arena: measured faster for this kernel than one launch over all arenas
    // (5.6 vs 5.1 TB/s on cfg 4; the per-tile job lookup sits on the warp's dependent-load chain)
    for (int k = 0; k < njobs; ++k) {
      if (!jobs[k].ntiles) continue;
      uint64_t want = (jobs[k].ntiles + 7) / 8, cap = (uint64_t)ctx->sm_count * 8;
      k_h64_tiles<<<(unsigned)(want < cap ? want : cap), 256, sizeof(uint64_t) + sizeof(HashJob), st>>>(
          dj + k, 1, jobs[k].ntiles);
      ++*launches;
    }
    k_h64_final<<<(unsigned)((recs + 255) / 256), 256, smem, st>>>(dj, njobs, recs);
    ++*launches;
  } else {
    unsigned int* next = reinterpret_cast<unsigned int*>(ctx->d_counters) + 8;  // words 8.. are free
    NCD_CUDA(cudaMemsetAsync(next, 0, 4, st));
    if (mode == NCD_MODE_SHA256) {
      const uint64_t want = (recs + 127) / 128, cap = (uint64_t)ctx->sm_count * 16;
      k_sha256<<<(unsigned)(want < cap ? want : cap), 128, smem, st>>>(dj, njobs, (uint32_t)recs, next);
    } else {
      const size_t xs = sizeof(XxhSmem) + smem;
      NCD_CUDA(cudaFuncSetAttribute(k_xxh64, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)xs));
      const uint64_t per_sm = (227u * 1024u) / (xs + 1024) ? (227u * 1024u) / (xs + 1024) : 1;  // 1 KB per CTA is the system's
      const uint64_t want = (recs + 32 * XXH_WARPS - 1) / (32 * XXH_WARPS), cap = (uint64_t)ctx->sm_count * per_sm;
      k_xxh64<<<(unsigned)(want < cap ? want : cap), 32 * XXH_WARPS, xs, st>>>(dj, njobs, (uint32_t)recs, next);
    }
    ++*launches;
  }
  NCD_CUDA(cudaGetLastError());
  return NCD_OK;
}

// Digests of the records named by d_idx[k] (device arrays, n[k] entries) of each arena, for every
// digest kind that arena currently holds valid; the other records' digests are untouched.  Called
// by the value sync so that cached digests stay valid across informer events.  SHA-256 and XXH64
// are serial inside a record, so all arenas go into ONE launch per kind: the pass then lasts as
// long as its longest record, not the sum over arenas.
int ncd_hash_records_many(ncd_ctx* ctx, int count, ncd_arena* const* arenas, const uint32_t* const* d_idx, const uint64_t* n) {
  cudaStream_t st = ctx->stream;
  std::vector<HashJob> sha_jobs, xxh_jobs;
  uint64_t sha_recs = 0, xxh_recs = 0;
  for (int k = 0; k < count; ++k) {
    ncd_arena* a = arenas[k];
    if (n[k] == 0 || a->nobj == 0) continue;
    if (a->h64_valid)
      k_h64_records<<<(unsigned)n[k], 256, 0, st>>>(d_idx[k], a->d.payload, a->d.rec_off, a->d.rec_len, (unsigned long long*)a->h64);
    HashJob j;
    memset(&j, 0, sizeof(j));
    j.payload = a->d.payload;
    j.rec_off = a->d.rec_off;
    j.rec_len = a->d.rec_len;
    j.sha = a->sha;
    j.xxh = a->xxh;
    j.nobj = n[k];
    j.rec_list = d_idx[k];
    if (a->sha_valid) {
      j.rec_base = sha_recs;
      sha_jobs.push_back(j);
      sha_recs += n[k];
    }
    if (a->xxh_valid) {
      j.rec_base = xxh_recs;
      xxh_jobs.push_back(j);
      xxh_recs += n[k];
    }
  }
  const int njobs = (int)(sha_jobs.size() + xxh_jobs.size());
  if (njobs) {
    if (njobs > MAX_HASH_JOBS || sha_recs >= 0xFFFFFFFFull || xxh_recs >= 0xFFFFFFFFull) {
      ncd_set_error("hash: too many arenas or records in one sync");
      return NCD_E_CAPACITY;
    }
    if (njobs > ctx->d_jobs_cap) {
      if (ctx->d_jobs) cudaFree(ctx->d_jobs);
      ctx->d_jobs = nullptr;
      ctx->d_jobs_cap = 0;
      const int cap = njobs < 32 ? 32 : njobs;
      NCD_CUDA(cudaMalloc(&ctx->d_jobs, sizeof(HashJob) * cap));
      ctx->d_jobs_cap = cap;
    }
    HashJob* dj = (HashJob*)ctx->d_jobs;
    unsigned int* next = reinterpret_cast<unsigned int*>(ctx->d_counters) + 8;  // two counters: SHA, XXH
    NCD_CUDA(cudaMemsetAsync(next, 0, 8, st));
    // pageable sources: staged by the runtime before each call returns
    if (!sha_jobs.empty()) {
      const int m = (int)sha_jobs.size();
      NCD_CUDA(cudaMemcpyAsync(dj, sha_jobs.data(), sizeof(HashJob) * m, cudaMemcpyHostToDevice, st));
      const uint64_t want = (sha_recs + 127) / 128, cap = (uint64_t)ctx->sm_count * 16;
      k_sha256<<<(unsigned)(want < cap ? want : cap), 128, sizeof(uint64_t) * m, st>>>(dj, m, (uint32_t)sha_recs, next);
    }
    if (!xxh_jobs.empty()) {
      const int m = (int)xxh_jobs.size();
      HashJob* djx = dj + sha_jobs.size();
      NCD_CUDA(cudaMemcpyAsync(djx, xxh_jobs.data(), sizeof(HashJob) * m, cudaMemcpyHostToDevice, st));
      const size_t xs = sizeof(XxhSmem) + sizeof(uint64_t) * m;
      NCD_CUDA(cudaFuncSetAttribute(k_xxh64, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)xs));
      const uint64_t per_sm = (227u * 1024u) / (xs + 1024) ? (227u * 1024u) / (xs + 1024) : 1;  // 1 KB per CTA is the system's
      const uint64_t want = (xxh_recs + 32 * XXH_WARPS - 1) / (32 * XXH_WARPS), cap = (uint64_t)ctx->sm_count * per_sm;
      k_xxh64<<<(unsigned)(want < cap ? want : cap), 32 * XXH_WARPS, xs, st>>>(djx, m, (uint32_t)xxh_recs, next + 1);
    }
  }
  NCD_CUDA(cudaGetLastError());
  return NCD_OK;
}

int ncd_hash_launch(ncd_ctx* ctx, ncd_arena* a, uint32_t mode, uint32_t* launches) {
  return ncd_hash_launch_many(ctx, &a, 1, mode, launches);
}

extern "C" int ncd_hash_arena(ncd_ctx* ctx, ncd_arena* a, uint32_t mode, void* out, uint64_t out_bytes,
                              float* kernel_ms) {
  if (!ctx || !a) {
    ncd_set_error("ncd_hash_arena: null argument");
    return NCD_E_INVALID;
  }
  if (!a->committed || a->ctx != ctx) {
    ncd_set_error("ncd_hash_arena: arena not committed on this context");
    return NCD_E_STATE;
  }
  const uint32_t hm = mode & 0xff;
  if (hm != NCD_MODE_HASH64 && hm != NCD_MODE_SHA256 && hm != NCD_MODE_XXH64) {
    ncd_set_error("ncd_hash_arena: mode must be NCD_MODE_HASH64, NCD_MODE_SHA256 or NCD_MODE_XXH64");
    return NCD_E_INVALID;
  }
  const uint64_t need = a->nobj * (hm == NCD_MODE_SHA256 ? 32 : 8);
  if (out && out_bytes < need) {
    ncd_set_error("ncd_hash_arena: out holds %llu bytes, %llu needed", (unsigned long long)out_bytes,
                  (unsigned long long)need);
    return NCD_E_CAPACITY;
  }
  NCD_CUDA(cudaSetDevice(ctx->device));
  // recompute unless asked not to: this entry point is also the timing harness of the hash kernels
  if (!(mode & NCD_MODE_CACHED)) {
    if (hm == NCD_MODE_HASH64) a->h64_valid = false;
    else if (hm == NCD_MODE_SHA256) a->sha_valid = false;
    else a->xxh_valid = false;
  }
  uint32_t launches = 0;
  NCD_CUDA(cudaEventRecord(ctx->ev[5], ctx->stream));
  NCD_TRY(ncd_hash_launch(ctx, a, hm, &launches));
  NCD_CUDA(cudaEventRecord(ctx->ev[6], ctx->stream));
  if (out && need)
    NCD_CUDA(cudaMemcpyAsync(out, hm == NCD_MODE_HASH64 ? (void*)a->h64 : hm == NCD_MODE_SHA256 ? (void*)a->sha : (void*)a->xxh, need,
                             cudaMemcpyDeviceToHost, ctx->stream));
  NCD_CUDA(cudaStreamSynchronize(ctx->stream));
  if (kernel_ms) cudaEventElapsedTime(kernel_ms, ctx->ev[5], ctx->ev[6]);
  return NCD_OK;
}

// ------------------------------------------------------------------ digests across a re-commit
// ncd_arena_recommit keeps surviving records where they are on the device (same rec_off, same
// bytes below payload_dirty_from) but may renumber the objects.  The digests the arena held are
// carried over by RECORD, not by index: a table of the old (rec_off, rec_len) -> old index, one
// probe per new object, and only the records that are new (appended behind payload_dirty_from, or
// never hashed) go through the hash kernels again.  A hash-mode sweep after an informer AddFunc /
// DeleteFunc then costs the added records, not the cluster.
namespace {

__device__ __forceinline__ uint32_t carry_slot(uint64_t off, uint32_t mask) {
  return (uint32_t)ncd_fmix64(off >> 4) & mask;
}

__global__ void k_carry_insert(const uint64_t* __restrict__ old_off, uint32_t old_n, uint32_t* __restrict__ table, uint32_t mask) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= old_n) return;
  uint32_t slot = carry_slot(old_off[j], mask);
  while (atomicCAS(&table[slot], 0xFFFFFFFFu, j) != 0xFFFFFFFFu) slot = (slot + 1) & mask;
}

__global__ void k_carry_lookup(const uint64_t* __restrict__ new_off, const uint32_t* __restrict__ new_len, uint32_t new_n,
                               const uint64_t* __restrict__ old_off, const uint32_t* __restrict__ old_len,
                               const uint32_t* __restrict__ table, uint32_t mask, uint64_t dirty_from,
                               const uint64_t* __restrict__ h64_old, uint64_t* __restrict__ h64_new,
                               const uint32_t* __restrict__ sha_old, uint32_t* __restrict__ sha_new,
                               const uint64_t* __restrict__ xxh_old, uint64_t* __restrict__ xxh_new,
                               uint32_t* __restrict__ rehash, uint32_t* __restrict__ nrehash) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= new_n) return;
  const uint64_t off = new_off[i];
  const uint32_t len = new_len[i];
  uint32_t found = 0xFFFFFFFFu;
  if (off < dirty_from) {  // appended records are new bytes whatever their offset says
    for (uint32_t slot = carry_slot(off, mask);; slot = (slot + 1) & mask) {
      const uint32_t j = table[slot];
      if (j == 0xFFFFFFFFu) break;
      if (old_off[j] == off && old_len[j] == len) {
        found = j;
        break;
      }
    }
  }
  if (found == 0xFFFFFFFFu) {
    rehash[atomicAdd(nrehash, 1u)] = i;
    return;
  }
  if (h64_new) h64_new[i] = h64_old[found];
  if (xxh_new) xxh_new[i] = xxh_old[found];
  if (sha_new) {
    const uint4* s = reinterpret_cast<const uint4*>(sha_old + (size_t)found * 8);
    uint4* d = reinterpret_cast<uint4*>(sha_new + (size_t)i * 8);
    d[0] = s[0];
    d[1] = s[1];
  }
}

}  // namespace

// old_off / old_len: device copies of the arena's rec_off / rec_len as they were before the
// re-commit (old_n entries); had_*: which digests were valid then.  On return the arena holds valid
// digests of those kinds for its NEW contents.  *rehashed = records that went through the kernels.
int ncd_digests_carry(ncd_arena* a, const uint64_t* old_off, const uint32_t* old_len, uint64_t old_n, uint64_t dirty_from,
                      bool had_h64, bool had_sha, bool had_xxh, uint64_t* rehashed) {
  ncd_ctx* ctx = a->ctx;
  cudaStream_t st = ctx->stream;
  const uint64_t n = a->nobj;
  if (rehashed) *rehashed = 0;
  if (!(had_h64 || had_sha || had_xxh) || n == 0) return NCD_OK;
  const size_t dcap = (size_t)(a->nobj_cap > n ? a->nobj_cap : n);
  uint64_t *h64_new = nullptr, *xxh_new = nullptr;
  uint32_t* sha_new = nullptr;
  if (had_h64) NCD_CUDA(cudaMalloc((void**)&h64_new, dcap * 8));
  if (had_sha) NCD_CUDA(cudaMalloc((void**)&sha_new, dcap * 32));
  if (had_xxh) NCD_CUDA(cudaMalloc((void**)&xxh_new, dcap * 8));
  uint32_t tsize = 16;
  while ((uint64_t)tsize < 2 * old_n) tsize <<= 1;
  uint32_t *table = nullptr, *list = nullptr, *count = nullptr;
  NCD_CUDA(cudaMallocAsync((void**)&table, (size_t)tsize * 4, st));
  NCD_CUDA(cudaMallocAsync((void**)&list, (n + 1) * 4, st));
  NCD_CUDA(cudaMallocAsync((void**)&count, 4, st));
  NCD_CUDA(cudaMemsetAsync(table, 0xFF, (size_t)tsize * 4, st));
  NCD_CUDA(cudaMemsetAsync(count, 0, 4, st));
  if (old_n) k_carry_insert<<<(unsigned)((old_n + 255) / 256), 256, 0, st>>>(old_off, (uint32_t)old_n, table, tsize - 1);
  k_carry_lookup<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(a->d.rec_off, a->d.rec_len, (uint32_t)n, old_off, old_len, table, tsize - 1,
                                                             dirty_from, a->h64, h64_new, a->sha, sha_new, a->xxh, xxh_new, list, count);
  uint32_t nre = 0;
  NCD_CUDA(cudaMemcpyAsync(&nre, count, 4, cudaMemcpyDeviceToHost, st));
  NCD_CUDA(cudaStreamSynchronize(st));
  // the carried digests become the arena's; the old buffers go
  if (had_h64) {
    cudaFree(a->h64);
    a->h64 = h64_new;
  }
  if (had_sha) {
    cudaFree(a->sha);
    a->sha = sha_new;
  }
  if (had_xxh) {
    cudaFree(a->xxh);
    a->xxh = xxh_new;
  }
  a->h64_valid = had_h64;
  a->sha_valid = had_sha;
  a->xxh_valid = had_xxh;
  int rc = NCD_OK;
  if (nre) {
    const uint32_t* lists[1] = {list};
    const uint64_t counts[1] = {nre};
    ncd_arena* as[1] = {a};
    rc = ncd_hash_records_many(ctx, 1, as, lists, counts);
  }
  NCD_CUDA(cudaFreeAsync(table, st));
  NCD_CUDA(cudaFreeAsync(list, st));
  NCD_CUDA(cudaFreeAsync(count, st));
  NCD_CUDA(cudaStreamSynchronize(st));
  if (rc != NCD_OK) a->h64_valid = a->sha_valid = a->xxh_valid = false;
  if (rehashed) *rehashed = nre;
  return rc;
}
