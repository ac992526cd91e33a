// ncd_scan.cu — the drift check itself.
//
//   k_join      cache join: for every controller object and every shard, the index of the
//               object with the same ("kind", "namespace/name") in that shard's arena
//               (reference: shard.*Lister.Get, controller.go:516,578,722,792), NotFound
//               when absent; lengths (or digests, in the hash modes) are compared here.
//   k_compare   byte equality of the canonical records, one warp per 4 KB tile of the
//               controller record against the same tile of every shard copy
//               (reference: reflect.DeepEqual at controller.go:539,600,725,795 — equality of
//               the injective canonical encodings is DeepEqual of the values).
//               (three data-movement variants: k_compare, k_compare_grouped here,
//               k_compare_tma in ncd_compare_tma.cu)
//   k_classify  decision code per (row, shard) from the verdicts + ownerReferences
//               (reference: controller.go:484-502, 529-560, 591-622, 725-738, 795-808).
//   k_emit      ordered stream compaction of the codes != SKIP into the diff list.
#include <vector>

#include "ncd_common.cuh"

// ------------------------------------------------------------------ k_join
__global__ void __launch_bounds__(256)
k_join(DevArena ctrl, const DevArena* __restrict__ shards, uint32_t S, uint32_t mode,
       uint32_t* __restrict__ sh_idx, uint64_t* __restrict__ sh_off, uint8_t* __restrict__ dv, uint32_t* __restrict__ tag) {
  const uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= (uint64_t)ctrl.nobj * S) return;
  const uint32_t c = (uint32_t)(p / S), s = (uint32_t)(p % S);
  const DevArena& sh = shards[s];
  const ObjMeta& q = ctrl.meta[c];
  // lister.Get: probe the shard's table, or reuse the previous scan's answer while neither
  // arena's keys have changed (value rewrites do not touch keys)
  uint32_t idx;
  if (sh.join_cache) {
    idx = sh.join_cache[c];
  } else {
    idx = ncd_probe_meta(sh, q, ctrl.aux);
    if (sh.join_cache_out) sh.join_cache_out[c] = idx;
  }
  uint8_t v;
  uint64_t off = NCD_NO_OFF;
  uint32_t tg = NCD_TAG_ROGUE;
  // the found object's record offset, length and tag: from the pair's cached copy, or from its line
  uint64_t s_off = 0;
  uint32_t s_len = 0;
  if (idx != NCD_NOTFOUND) {
    if (sh.join_line) {
      const uint4 w = sh.join_line[c];
      s_off = (uint64_t)w.x | ((uint64_t)w.y << 32);
      s_len = w.z;
      tg = w.w;
    } else {
      const uint4* line = reinterpret_cast<const uint4*>(sh.meta + idx);
      const uint4 w0 = line[0], w1 = line[1];  // keyhash | rec_off ; rec_len key_off key_len/kind tag
      s_off = (uint64_t)w0.z | ((uint64_t)w0.w << 32);
      s_len = w1.x;
      tg = w1.w;
      if (sh.join_line_out) {
        sh.join_line_out[c] = make_uint4(w0.z, w0.w, s_len, tg);
        sh.join_rev_out[idx] = c;
      }
    }
  }
  if (idx == NCD_NOTFOUND) {
    v = DV_NOTFOUND;
  } else if (s_len != q.rec_len) {
    v = DV_DIFF;  // the encoding is injective: different lengths cannot be DeepEqual
  } else {
    const uint32_t hm = mode & 0xff;
    bool maybe_equal = true;
    if (hm == NCD_MODE_HASH64) {
      maybe_equal = sh.h64[idx] == ctrl.h64[c];
    } else if (hm == NCD_MODE_XXH64) {
      maybe_equal = sh.xxh[idx] == ctrl.xxh[c];
    } else if (hm == NCD_MODE_SHA256) {
      const uint4* x = reinterpret_cast<const uint4*>(sh.sha + (size_t)idx * 8);
      const uint4* y = reinterpret_cast<const uint4*>(ctrl.sha + (size_t)c * 8);
      uint4 x0 = x[0], x1 = x[1], y0 = y[0], y1 = y[1];
      maybe_equal = ((x0.x ^ y0.x) | (x0.y ^ y0.y) | (x0.z ^ y0.z) | (x0.w ^ y0.w) | (x1.x ^ y1.x) |
                     (x1.y ^ y1.y) | (x1.z ^ y1.z) | (x1.w ^ y1.w)) == 0;
    }
    v = maybe_equal ? DV_EQUAL : DV_DIFF;
    if (maybe_equal && (hm == NCD_MODE_DIRECT || (mode & NCD_MODE_CONFIRM))) off = s_off;
  }
  sh_idx[p] = idx;
  sh_off[p] = off;
  dv[p] = v;
  tag[p] = tg;
}

// ------------------------------------------------------------------ k_compare
// One warp per tile.  The controller tile is loaded once into registers (8 x 16 B per lane)
// and compared with the same tile of each shard copy that is present with an equal length,
// so every canonical byte of every copy crosses HBM once.
__device__ __forceinline__ uint4 mask_tail(uint4 v, uint32_t valid /*1..15*/) {
  uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int nb = (int)valid - 4 * i;  // valid bytes in word i
    if (nb <= 0) w[i] = 0;
    else if (nb < 4) w[i] &= (1u << (8 * nb)) - 1u;
  }
  return make_uint4(w[0], w[1], w[2], w[3]);
}

constexpr int CMP_U = NCD_TILE_VECS / 32;  // vectors per lane per tile

__global__ void __launch_bounds__(256)
k_compare(DevArena ctrl, const DevArena* __restrict__ shards, uint32_t S,
          const uint32_t* __restrict__ tile_obj, const uint32_t* __restrict__ tile_chunk,
          uint64_t ntiles, const uint64_t* __restrict__ sh_off, uint8_t* __restrict__ dv,
          uint32_t* __restrict__ counters, uint32_t count_collisions) {
  const uint32_t lane = threadIdx.x & 31;
  const uint64_t warp0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  for (uint64_t t = warp0; t < ntiles; t += nwarps) {
    const uint32_t c = tile_obj[t], chunk = tile_chunk[t];
    const uint32_t len = ctrl.rec_len[c];
    const uint64_t toff = (uint64_t)chunk * NCD_TILE_BYTES;
    const uint32_t nbytes = min(NCD_TILE_BYTES, len - (uint32_t)toff);
    const uint32_t nvec = (nbytes + 15) >> 4;
    const uint32_t tail = nbytes & 15;
    const uint4* cp = reinterpret_cast<const uint4*>(ctrl.payload + ctrl.rec_off[c] + toff);
    uint4 a[CMP_U];
#pragma unroll
    for (int u = 0; u < CMP_U; ++u) {
      const uint32_t v = lane + 32 * u;
      a[u] = v < nvec ? ncd_ldg_stream(cp + v) : make_uint4(0, 0, 0, 0);
      if (tail && v == nvec - 1) a[u] = mask_tail(a[u], tail);
    }
    for (uint32_t s0 = 0; s0 < S; s0 += 32) {
      const uint64_t my_off = (s0 + lane < S) ? sh_off[(uint64_t)c * S + s0 + lane] : NCD_NO_OFF;
      uint32_t active = __ballot_sync(0xffffffffu, my_off != NCD_NO_OFF);
      while (active) {
        const int j = __ffs(active) - 1;
        active &= active - 1;
        const uint64_t so = __shfl_sync(0xffffffffu, my_off, j);
        const uint32_t s = s0 + j;
        const uint4* sp = reinterpret_cast<const uint4*>(shards[s].payload + so + toff);
        uint4 b[CMP_U];
#pragma unroll
        for (int u = 0; u < CMP_U; ++u) {
          const uint32_t v = lane + 32 * u;
          b[u] = v < nvec ? ncd_ldg_stream(sp + v) : make_uint4(0, 0, 0, 0);
        }
        uint32_t x = 0;
#pragma unroll
        for (int u = 0; u < CMP_U; ++u) {
          const uint32_t v = lane + 32 * u;
          uint4 bb = b[u];
          if (tail && v == nvec - 1) bb = mask_tail(bb, tail);
          x |= (a[u].x ^ bb.x) | (a[u].y ^ bb.y) | (a[u].z ^ bb.z) | (a[u].w ^ bb.w);
        }
        if (__any_sync(0xffffffffu, x != 0) && lane == 0) {
          // DV_EQUAL (0) -> DV_DIFF (1) on the byte of this pair; the first tile to find a
          // difference sees the old value and, in CONFIRM mode, counts a hash collision
          const uint64_t p = (uint64_t)c * S + s;
          uint32_t* word = reinterpret_cast<uint32_t*>(dv) + (p >> 2);
          const uint32_t sh8 = 8u * (uint32_t)(p & 3);
          const uint32_t old = atomicOr(word, (uint32_t)DV_DIFF << sh8);
          if (count_collisions && ((old >> sh8) & 0xffu) == DV_EQUAL) atomicAdd(&counters[1], 1u);
        }
      }
    }
  }
}

// ------------------------------------------------------------------ k_compare_grouped
// Same work item (one warp per 4 KB controller tile) but the tile is walked in chunks of CH
// vectors per lane and, for every chunk, the loads of the controller copy and of up to G shard
// copies are all issued before any of them is used: (1 + G) * CH independent 16-byte loads in
// flight per lane and one memory latency per chunk, instead of one latency per (tile, shard).
// That is what small records need: their tiles are ragged, so a warp has few bytes per
// latency unless it covers several copies at once.
constexpr int CMP_G = 8;

template <int CH>
__global__ void __launch_bounds__(256)
k_compare_grouped(DevArena ctrl, const DevArena* __restrict__ shards, uint32_t S,
                  const uint32_t* __restrict__ tile_obj, const uint32_t* __restrict__ tile_chunk,
                  uint64_t ntiles, const uint64_t* __restrict__ sh_off, uint8_t* __restrict__ dv,
                  uint32_t* __restrict__ counters, uint32_t count_collisions) {
  extern __shared__ const uint8_t* s_payload[];  // payload base of every shard
  for (uint32_t s = threadIdx.x; s < S; s += blockDim.x) s_payload[s] = shards[s].payload;
  __syncthreads();
  const uint32_t lane = threadIdx.x & 31;
  const uint64_t warp0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  for (uint64_t t = warp0; t < ntiles; t += nwarps) {
    const uint32_t c = tile_obj[t], chunk = tile_chunk[t];
    const uint32_t len = ctrl.rec_len[c];
    const uint64_t toff = (uint64_t)chunk * NCD_TILE_BYTES;
    // the last compare tile of a record runs to its end (2049..6144 bytes, ncd_ctiles_of)
    const uint32_t nbytes = chunk + 1 == ncd_ctiles_of(len) ? len - (uint32_t)toff : NCD_TILE_BYTES;
    const uint32_t nvec = (nbytes + 15) >> 4;
    const uint32_t tail = nbytes & 15;
    const uint4* cp = reinterpret_cast<const uint4*>(ctrl.payload + ctrl.rec_off[c] + toff);
    for (uint32_t s0 = 0; s0 < S; s0 += CMP_G) {
      const uint64_t my_off = (lane < CMP_G && s0 + lane < S) ? sh_off[(uint64_t)c * S + s0 + lane] : NCD_NO_OFF;
      const uint4* sp[CMP_G];
      uint32_t x[CMP_G];
#pragma unroll
      for (int k = 0; k < CMP_G; ++k) {
        const uint64_t off = __shfl_sync(0xffffffffu, my_off, k);
        sp[k] = off == NCD_NO_OFF ? nullptr
                                  : reinterpret_cast<const uint4*>(s_payload[s0 + k] + off + toff);
        x[k] = 0;
      }
      if (__all_sync(0xffffffffu, my_off == NCD_NO_OFF)) continue;  // nothing to read in this group
      for (uint32_t v0 = 0; v0 < nvec; v0 += 32 * CH) {
        uint4 a[CH], b[CMP_G][CH];
#pragma unroll
        for (int u = 0; u < CH; ++u) {
          const uint32_t v = v0 + lane + 32 * u;
          a[u] = v < nvec ? ncd_ldg_stream(cp + v) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int k = 0; k < CMP_G; ++k)
#pragma unroll
          for (int u = 0; u < CH; ++u) {
            const uint32_t v = v0 + lane + 32 * u;
            b[k][u] = (sp[k] != nullptr && v < nvec) ? ncd_ldg_stream(sp[k] + v) : a[u];
          }
#pragma unroll
        for (int u = 0; u < CH; ++u) {
          const uint32_t v = v0 + lane + 32 * u;
          const bool last = tail && v == nvec - 1;
          if (last) a[u] = mask_tail(a[u], tail);
#pragma unroll
          for (int k = 0; k < CMP_G; ++k) {
            uint4 bb = b[k][u];
            if (last) bb = mask_tail(bb, tail);
            x[k] |= (a[u].x ^ bb.x) | (a[u].y ^ bb.y) | (a[u].z ^ bb.z) | (a[u].w ^ bb.w);
          }
        }
      }
#pragma unroll
      for (int k = 0; k < CMP_G; ++k) {
        if (__any_sync(0xffffffffu, x[k] != 0) && lane == 0) {
          const uint64_t p = (uint64_t)c * S + s0 + k;
          uint32_t* word = reinterpret_cast<uint32_t*>(dv) + (p >> 2);
          const uint32_t sh8 = 8u * (uint32_t)(p & 3);
          const uint32_t old = atomicOr(word, (uint32_t)DV_DIFF << sh8);
          if (count_collisions && ((old >> sh8) & 0xffu) == DV_EQUAL) atomicAdd(&counters[1], 1u);
        }
      }
    }
  }
}

// ------------------------------------------------------------------ k_compare_small
// Records of a few hundred bytes (template and workgroup specs, small ConfigMaps): a whole warp
// per tile leaves most lanes without a vector and the scan becomes bound per record, not per
// byte (1M records x 8 shards in 1.5 ms whatever their length).  Here eight lanes share a tile,
// so a warp walks four tiles at once: four times the tiles in flight for the same registers.
// Correct for any tile size (a lane loops over vectors l, l + 8, ...); chosen only when tiles are small.
__global__ void __launch_bounds__(256)
k_compare_small(DevArena ctrl, const DevArena* __restrict__ shards, uint32_t S,
                const uint32_t* __restrict__ tile_obj, const uint32_t* __restrict__ tile_chunk,
                uint64_t ntiles, const uint64_t* __restrict__ sh_off, uint8_t* __restrict__ dv,
                uint32_t* __restrict__ counters, uint32_t count_collisions) {
  extern __shared__ const uint8_t* s_payload[];  // payload base of every shard
  for (uint32_t s = threadIdx.x; s < S; s += blockDim.x) s_payload[s] = shards[s].payload;
  __syncthreads();
  const uint32_t lane = threadIdx.x & 31, sub = lane >> 3, l8 = lane & 7;
  const uint64_t warp0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  for (uint64_t t0 = 4 * warp0; t0 < ntiles; t0 += 4 * nwarps) {
    const uint64_t t = t0 + sub;
    const bool have = t < ntiles;
    uint32_t c = 0, nvec = 0, tail = 0;
    uint64_t toff = 0;
    const uint4* cp = nullptr;
    if (have) {
      c = tile_obj[t];
      const uint32_t chunk = tile_chunk[t];
      const uint32_t len = ctrl.rec_len[c];
      toff = (uint64_t)chunk * NCD_TILE_BYTES;
      const uint32_t nbytes = chunk + 1 == ncd_ctiles_of(len) ? len - (uint32_t)toff : NCD_TILE_BYTES;
      nvec = (nbytes + 15) >> 4;
      tail = nbytes & 15;
      cp = reinterpret_cast<const uint4*>(ctrl.payload + ctrl.rec_off[c] + toff);
    }
    // the longest of the four tiles sets the trip count for the warp (shuffles below are warp-wide)
    uint32_t maxvec = nvec;
#pragma unroll
    for (int d = 8; d < 32; d <<= 1) maxvec = max(maxvec, __shfl_xor_sync(0xffffffffu, maxvec, d));
    for (uint32_t s0 = 0; s0 < S; s0 += CMP_G) {
      const uint64_t my_off = (have && s0 + l8 < S) ? sh_off[(uint64_t)c * S + s0 + l8] : NCD_NO_OFF;
      if (__all_sync(0xffffffffu, my_off == NCD_NO_OFF)) continue;  // nothing to read for any of the four tiles
      const uint4* sp[CMP_G];
      uint32_t x[CMP_G];
#pragma unroll
      for (int k = 0; k < CMP_G; ++k) {
        const uint64_t off = __shfl_sync(0xffffffffu, my_off, k, 8);  // lane k of this tile's eight lanes
        sp[k] = off == NCD_NO_OFF ? nullptr : reinterpret_cast<const uint4*>(s_payload[min(s0 + k, S - 1)] + off + toff);
        x[k] = 0;
      }
      for (uint32_t v = l8; v < maxvec; v += 8) {
        if (v >= nvec) continue;
        uint4 a = ncd_ldg_stream(cp + v);
        uint4 b[CMP_G];
#pragma unroll
        for (int k = 0; k < CMP_G; ++k) b[k] = sp[k] != nullptr ? ncd_ldg_stream(sp[k] + v) : a;
        const bool last = tail && v == nvec - 1;
        if (last) a = mask_tail(a, tail);
#pragma unroll
        for (int k = 0; k < CMP_G; ++k) {
          uint4 bb = b[k];
          if (last && sp[k] != nullptr) bb = mask_tail(bb, tail);
          if (sp[k] != nullptr) x[k] |= (a.x ^ bb.x) | (a.y ^ bb.y) | (a.z ^ bb.z) | (a.w ^ bb.w);
        }
      }
#pragma unroll
      for (int k = 0; k < CMP_G; ++k) {
        const uint32_t differs = (__ballot_sync(0xffffffffu, x[k] != 0) >> (8 * sub)) & 0xffu;  // this tile's lanes
        if (differs && l8 == (uint32_t)k) {
          const uint64_t p = (uint64_t)c * S + s0 + k;
          uint32_t* word = reinterpret_cast<uint32_t*>(dv) + (p >> 2);
          const uint32_t sh8 = 8u * (uint32_t)(p & 3);
          const uint32_t old = atomicOr(word, (uint32_t)DV_DIFF << sh8);
          if (count_collisions && ((old >> sh8) & 0xffu) == DV_EQUAL) atomicAdd(&counters[1], 1u);
        }
      }
    }
  }
}

// ------------------------------------------------------------------ k_scan_small
// Small records (a few hundred bytes: template and workgroup specs, small ConfigMaps), direct mode:
// the join and the byte compare in ONE kernel.  With k_join + k_compare_small every 64-256 bytes
// of payload sat behind two dependent load levels (tile list -> record offset/length -> shard
// offsets written by the join), and the join itself had already touched the object's line.  Here
// eight lanes own one controller object.  Per group of up to eight shards, lane l looks the object
// up in shard l (lister.Get: the cached answer of the previous scan, or a probe of the shard's
// table) and reads the FIRST 32-byte sector of the shard object's line — record offset, length and
// ownership tag; then the eight lanes walk the record together, vector v of the controller copy
// against vector v of every shard copy, all loads of a step issued before any is used.  Verdict,
// shard index and tag go out for k_classify; no tile list, no offsets round trip.
// Correct for any record length (the lanes loop over the whole record); chosen when records are small.
// LPO lanes own one object — as many as the record has 16-byte vectors (4, 8, 16 or 32), so that
// every copy of a typical record is ONE load per lane and an object costs two memory latencies
// (object lines, then payload) however long it is; lane l answers the join for shards l, l + LPO, ...
// of the group (the lanes beyond G have no shard to look up and only compare).
// When the join is cached and the shards fit one group, the cached indices of the NEXT object are
// requested while the current one is compared, which takes one level out of the dependent chain
// (index -> object line -> payload) every object otherwise waits through.
#ifndef NCD_SS_CTAS
#define NCD_SS_CTAS 3  // resident CTAs per SM asked of the eight-shard variant (80 registers; 4 would force 64 and spills)
#endif
template <int G, int LPO>
__global__ void __launch_bounds__(256, G == 8 ? NCD_SS_CTAS : 4)
k_scan_small(DevArena ctrl, const DevArena* __restrict__ shards, uint32_t S, uint32_t* __restrict__ sh_idx,
             uint8_t* __restrict__ dv, uint32_t* __restrict__ tag) {
  constexpr int SLOTS = G > LPO ? G / LPO : 1;  // shards a lane joins per group of G
  struct ShardRef {
    const uint8_t* payload;
    const ObjMeta* meta;
    const uint32_t* join_cache;
    uint32_t* join_cache_out;
    const uint4* join_line;
    uint4* join_line_out;
    uint32_t* join_rev_out;
  };
  extern __shared__ __align__(16) uint8_t s_raw[];
  ShardRef* s_sh = reinterpret_cast<ShardRef*>(s_raw);
  for (uint32_t s = threadIdx.x; s < S; s += blockDim.x)
    s_sh[s] = ShardRef{shards[s].payload,   shards[s].meta,          shards[s].join_cache, shards[s].join_cache_out,
                       shards[s].join_line, shards[s].join_line_out, shards[s].join_rev_out};
  __syncthreads();
  const uint32_t lane = threadIdx.x & 31, l = lane & (LPO - 1), base = lane & ~(uint32_t)(LPO - 1);
  const uint64_t group0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) / LPO;
  const uint64_t ngroups = ((uint64_t)gridDim.x * blockDim.x) / LPO;
  // look-ahead only when EVERY shard's answer is cached (a shard re-committed since the last scan has
  // no valid cache while the others still do)
  bool prefetch = S <= (uint32_t)G && S > 0;
  for (uint32_t s = 0; s < S && prefetch; ++s) prefetch = s_sh[s].join_cache != nullptr;
  // ... and of the cached copies of the found objects' lines, when every shard has those too: the
  // steady state, in which an object then waits for ONE memory latency, its payload's
  bool prefetch_line = prefetch;
  for (uint32_t s = 0; s < S && prefetch_line; ++s) prefetch_line = s_sh[s].join_line != nullptr;
  uint32_t idx_next[SLOTS];
  uint4 line_next[SLOTS];
#pragma unroll
  for (int j = 0; j < SLOTS; ++j) {
    const uint32_t s = l + j * LPO;
    const bool mine_here = s < G && s < S && group0 < ctrl.nobj;
    idx_next[j] = (prefetch && mine_here) ? s_sh[s].join_cache[group0] : NCD_NOTFOUND;
    line_next[j] = (prefetch_line && mine_here) ? s_sh[s].join_line[group0] : make_uint4(0, 0, 0, 0);
  }
  // The loop is uniform over the warp (its groups hold consecutive objects; the warp goes on while its first
  // group has one): shuffles and reductions below can name the whole warp, which spares the collective
  // prologue and epilogue a partial, run-time member mask costs on every one of them.  A group past the last
  // object (the last warp's last round) walks a record of no bytes and stores nothing.
  for (uint64_t cw = group0 - lane / LPO; cw < ctrl.nobj; cw += ngroups) {
    const uint64_t c = cw + lane / LPO;
    const bool act = c < ctrl.nobj;
    const ObjMeta& q = ctrl.meta[act ? c : cw];
    const uint32_t len = act ? q.rec_len : 0u;
    const uint32_t nvec = (len + 15) >> 4, tail = len & 15;
    const uint4* cp = reinterpret_cast<const uint4*>(ctrl.payload + q.rec_off);
    uint32_t idx_cur[SLOTS];
    uint4 line_cur[SLOTS];
#pragma unroll
    for (int j = 0; j < SLOTS; ++j) {
      idx_cur[j] = idx_next[j];
      line_cur[j] = line_next[j];
    }
    if (prefetch && c + ngroups < ctrl.nobj) {
#pragma unroll
      for (int j = 0; j < SLOTS; ++j) {
        const uint32_t s = l + j * LPO;
        if (s < G && s < S) {
          idx_next[j] = s_sh[s].join_cache[c + ngroups];
          if (prefetch_line) line_next[j] = s_sh[s].join_line[c + ngroups];
        }
      }
    }
    for (uint32_t s0 = 0; s0 < S; s0 += G) {
      // ---- join: this lane answers for shards s0 + l, s0 + l + LPO, ...
      uint32_t idx[SLOTS], tg[SLOTS];
      uint8_t v[SLOTS];
      const uint4* mine[SLOTS];  // this lane's shard copies, where bytes have to be compared
#pragma unroll
      for (int j = 0; j < SLOTS; ++j) {
        const uint32_t s = s0 + l + j * LPO;
        idx[j] = NCD_NOTFOUND;
        tg[j] = NCD_TAG_ROGUE;
        v[j] = DV_NOTFOUND;
        mine[j] = nullptr;
        if (act && l + j * LPO < G && s < S) {
          const ShardRef sr = s_sh[s];
          if (prefetch) {
            idx[j] = idx_cur[j];
          } else if (sr.join_cache) {
            idx[j] = sr.join_cache[c];
          } else {
            idx[j] = ncd_probe_meta(shards[s], q, ctrl.aux);
            if (sr.join_cache_out) sr.join_cache_out[c] = idx[j];
          }
          if (idx[j] != NCD_NOTFOUND) {
            uint4 w;  // rec_off lo, hi, rec_len, tag: the pair's cached copy, or from the object's line
            if (prefetch_line) {
              w = line_cur[j];
            } else if (sr.join_line) {
              w = sr.join_line[c];
            } else {
              const uint4* line = reinterpret_cast<const uint4*>(sr.meta + idx[j]);
              const uint4 w0 = line[0], w1 = line[1];  // keyhash | rec_off ; rec_len key_off key_len/kind tag
              w = make_uint4(w0.z, w0.w, w1.x, w1.w);
              if (sr.join_line_out) {
                sr.join_line_out[c] = w;
                sr.join_rev_out[idx[j]] = (uint32_t)c;
              }
            }
            tg[j] = w.w;
            if (w.z != len) {
              v[j] = DV_DIFF;  // the encoding is injective: different lengths cannot be DeepEqual
            } else {
              v[j] = DV_EQUAL;
              mine[j] = reinterpret_cast<const uint4*>(sr.payload + ((uint64_t)w.x | ((uint64_t)w.y << 32)));
            }
          }
        }
      }
      // ---- compare: the lanes walk the record against the G shard copies together
      const uint4* sp[G];
      uint32_t x[G];
#pragma unroll
      for (int k = 0; k < G; ++k) {  // shard s0 + k was joined by lane k % LPO in its slot k / LPO
        // a copy with nothing to compare (absent, or another length) reads the controller's own bytes instead:
        // no difference can come of it, and the loop below has no per-copy condition
        const uint4* src = mine[k / LPO] != nullptr ? mine[k / LPO] : cp;
        sp[k] = reinterpret_cast<const uint4*>(__shfl_sync(0xffffffffu, (unsigned long long)src, base + (k % LPO)));
        x[k] = 0;
      }
      for (uint32_t vv = l; vv < nvec; vv += LPO) {
        uint4 a = ncd_ldg_stream(cp + vv);
        uint4 b[G];
#pragma unroll
        for (int k = 0; k < G; ++k) b[k] = ncd_ldg_stream(sp[k] + vv);
        if (tail && vv == nvec - 1) {
          a = mask_tail(a, tail);
#pragma unroll
          for (int k = 0; k < G; ++k) b[k] = mask_tail(b[k], tail);
        }
#pragma unroll
        for (int k = 0; k < G; ++k) x[k] |= (a.x ^ b[k].x) | (a.y ^ b[k].y) | (a.z ^ b[k].z) | (a.w ^ b[k].w);
      }
      uint32_t xm = 0;  // bit k: this lane saw copy k differ; one OR over the object's lanes instead of a vote per copy
#pragma unroll
      for (int k = 0; k < G; ++k) xm |= x[k] != 0 ? 1u << k : 0u;
#pragma unroll
      for (int d = 1; d < LPO; d <<= 1) xm |= __shfl_xor_sync(0xffffffffu, xm, d);  // stays inside the object's LPO lanes
#pragma unroll
      for (int j = 0; j < SLOTS; ++j)
        if (xm & (1u << ((l + j * LPO) & 31))) v[j] = DV_DIFF;  // only a copy with bytes to compare can differ; lanes past G set nothing that is stored
#pragma unroll
      for (int j = 0; j < SLOTS; ++j) {
        const uint32_t s = s0 + l + j * LPO;
        if (act && l + j * LPO < G && s < S) {
          const uint64_t p = c * S + s;
          sh_idx[p] = idx[j];
          dv[p] = v[j];
          tag[p] = tg[j];
        }
      }
    }
  }
}

// ------------------------------------------------------------------ k_classify
// One thread per ROW, walking the row's shards: the row's controller index, dependent entry and
// owning template are read once, the verdicts and tags of its S pairs are contiguous, and the grid
// is rows / 256 blocks instead of rows x shards / 256 (the compaction's block scan stays in one
// block for up to a million rows).
// The decision from values already in registers: v / mine = verdict and tag of the compared object's
// pair, tv / ttag = of the owning template's pair (dependent rows only), p = the pair's index.
__device__ __forceinline__ uint32_t decide_vals(const DevArena* __restrict__ shards, uint32_t s, uint32_t dep, uint32_t cidx,
                                                uint8_t v, uint32_t mine, uint8_t tv, uint32_t ttag,
                                                const uint32_t* __restrict__ sh_idx, uint64_t p) {
  if (dep == NCD_NOTFOUND) return v == DV_NOTFOUND ? NCD_CREATE : (v == DV_DIFF ? NCD_UPDATE_DATA : NCD_SKIP);
  if (cidx == NCD_NOTFOUND) return NCD_ERR_MISSING_REF;
  if (v == DV_NOTFOUND) return NCD_CREATE;
  if (mine == NCD_TAG_ROGUE) return NCD_ERR_ROGUE;
  bool missing = true;
  if (tv != DV_NOTFOUND) {
    const uint32_t want = ttag | 0x80000000u;
    if (mine != NCD_TAG_MANY) {
      missing = mine != want;
    } else {
      const DevArena& sh = shards[s];
      const uint32_t j = sh_idx[p];
      for (uint32_t o = sh.own_start[j]; o < sh.own_start[j + 1] && missing; ++o)
        if (sh.own_packed[o] == want) missing = false;
    }
  }
  return (v == DV_DIFF ? NCD_UPDATE_DATA : 0u) | (missing ? NCD_UPDATE_OWNER : 0u);
}

__device__ __forceinline__ uint32_t decide_pair(const DevArena* __restrict__ shards, uint32_t S, uint32_t s, uint32_t obj,
                                                uint32_t dep, uint32_t cidx, const uint32_t* __restrict__ sh_idx,
                                                const uint8_t* __restrict__ dv, const uint32_t* __restrict__ tag) {
  if (dep == NCD_NOTFOUND) {
    // template (:792-808) or workgroup (:722-738) row
    const uint8_t v = dv[(uint64_t)cidx * S + s];
    return v == DV_NOTFOUND ? NCD_CREATE : (v == DV_DIFF ? NCD_UPDATE_DATA : NCD_SKIP);
  }
  if (cidx == NCD_NOTFOUND) return NCD_ERR_MISSING_REF;  // :509-514, :571-576
  const uint64_t p = (uint64_t)cidx * S + s;
  const uint8_t v = dv[p];
  if (v == DV_NOTFOUND) return NCD_CREATE;  // :518-520, :580-582
  // The join left each shard object's ownership tag beside its verdict (ObjMeta.tag): the single
  // ownerReference of a dependent, the interned UID of a template.  The common cases are decided
  // from those two words; only a dependent with several ownerReferences walks its list.
  const uint32_t mine = tag[p];
  if (mine == NCD_TAG_ROGUE) return NCD_ERR_ROGUE;  // :494-499, checked before the data compare (:529 before :539)
  // isMissingOwnership(:484-502) against the shard's copy of the owning template.  A template that
  // is about to be created on this shard has a UID no existing ownerReference can carry, so the
  // reference is missing by definition.
  const uint64_t tp = (uint64_t)obj * S + s;
  bool missing = true;
  if (dv[tp] != DV_NOTFOUND) {
    // UIDs were interned at commit: equal ids <=> byte-equal strings (ncd_index.cu)
    const uint32_t want = tag[tp] | 0x80000000u;  // Kind == NexusAlgorithmTemplate && same UID
    if (mine != NCD_TAG_MANY) {
      missing = mine != want;
    } else {
      const DevArena& sh = shards[s];
      const uint32_t j = sh_idx[p];
      for (uint32_t o = sh.own_start[j]; o < sh.own_start[j + 1] && missing; ++o)
        if (sh.own_packed[o] == want) missing = false;
    }
  }
  return (v == DV_DIFF ? NCD_UPDATE_DATA : 0u) | (missing ? NCD_UPDATE_OWNER : 0u);
}

__global__ void __launch_bounds__(256)
k_classify(DevArena ctrl, const DevArena* __restrict__ shards, uint32_t S, uint64_t nrows,
           const uint32_t* __restrict__ row_obj, const uint32_t* __restrict__ row_dep,
           const uint32_t* __restrict__ row_cidx, const uint32_t* __restrict__ sh_idx,
           const uint8_t* __restrict__ dv, const uint32_t* __restrict__ tag, uint8_t* __restrict__ codes,
           uint32_t* __restrict__ blk_cnt, uint32_t* __restrict__ blk_off, uint32_t* __restrict__ counters,
           uint32_t fuse_scan) {
  const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t mine = 0;
  if (r < nrows) {
    const uint32_t obj = row_obj[r], dep = row_dep[r], cidx = row_cidx[r];
    if (S == 8) {
      // eight shards: a row's verdicts are 8 bytes and its tags 32, both contiguous and aligned (the
      // scratch arrays start on 256-byte boundaries) — six vector loads and one vector store per row
      // instead of 32 loads and 8 byte stores
      const bool is_dep = dep != NCD_NOTFOUND, has_c = cidx != NCD_NOTFOUND;
      uint2 vw = make_uint2(0, 0), tvw = make_uint2(0, 0);
      uint4 t0 = make_uint4(0, 0, 0, 0), t1 = t0, u0 = t0, u1 = t0;
      if (has_c) {
        vw = *reinterpret_cast<const uint2*>(dv + (uint64_t)cidx * 8);
        if (is_dep) {
          const uint4* tp4 = reinterpret_cast<const uint4*>(tag + (uint64_t)cidx * 8);
          t0 = tp4[0];
          t1 = tp4[1];
        }
      }
      if (is_dep) {
        tvw = *reinterpret_cast<const uint2*>(dv + (uint64_t)obj * 8);
        const uint4* up4 = reinterpret_cast<const uint4*>(tag + (uint64_t)obj * 8);
        u0 = up4[0];
        u1 = up4[1];
      }
      const uint32_t tg[8] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w};
      const uint32_t ug[8] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w};
      uint32_t lo = 0, hi = 0;
#pragma unroll
      for (int s = 0; s < 8; ++s) {
        const uint8_t v = (uint8_t)((s < 4 ? vw.x : vw.y) >> (8 * (s & 3)));
        const uint8_t tv = (uint8_t)((s < 4 ? tvw.x : tvw.y) >> (8 * (s & 3)));
        const uint32_t code = decide_vals(shards, (uint32_t)s, dep, cidx, v, tg[s], tv, ug[s], sh_idx, (uint64_t)cidx * 8 + s);
        if (s < 4) lo |= code << (8 * s);
        else hi |= code << (8 * (s - 4));
        mine += code != NCD_SKIP;
      }
      *reinterpret_cast<uint2*>(codes + r * 8) = make_uint2(lo, hi);
    } else {
      for (uint32_t s = 0; s < S; ++s) {
        const uint32_t code = decide_pair(shards, S, s, obj, dep, cidx, sh_idx, dv, tag);
        codes[r * S + s] = (uint8_t)code;
        mine += code != NCD_SKIP;
      }
    }
  }
  // entries this block will emit
  __shared__ uint32_t wsum[8];
  uint32_t w = mine;
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) w += __shfl_xor_sync(0xffffffffu, w, d);
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = w;
  __syncthreads();
  // Small grids (launch-latency bound scans): the block that finishes last turns the per-block
  // counts into offsets (exclusive scan) and the total into counters[0], so the ordered compaction
  // needs no separate scan launches.  Large grids keep the three-launch scan.
  __shared__ bool is_last;
  __shared__ uint32_t part[256];
  if (threadIdx.x == 0) {
    uint32_t cnt = 0;
    for (int k = 0; k < 8; ++k) cnt += wsum[k];
    blk_cnt[blockIdx.x] = cnt;
    if (!fuse_scan) return;
    __threadfence();
    is_last = atomicAdd(&counters[2], 1u) == gridDim.x - 1;
  }
  if (!fuse_scan) return;
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  // counts written by other blocks: read past L1 (__ldcg), independent loads so they pipeline
  const uint32_t nblk = gridDim.x, per = (nblk + 255) / 256;
  const uint32_t b0 = threadIdx.x * per, b1 = min(nblk, b0 + per);
  uint32_t sum = 0;
#pragma unroll 8
  for (uint32_t b = b0; b < b1; ++b) sum += __ldcg(blk_cnt + b);
  part[threadIdx.x] = sum;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t run = 0;
    for (int t = 0; t < 256; ++t) {
      const uint32_t v = part[t];
      part[t] = run;
      run += v;
    }
    counters[0] = run;
  }
  __syncthreads();
  uint32_t run = part[threadIdx.x];
#pragma unroll 8
  for (uint32_t b = b0; b < b1; ++b) {
    blk_off[b] = run;
    run += __ldcg(blk_cnt + b);
  }
}

// ------------------------------------------------------------------ k_emit
// Same block shape as k_classify (a thread per row); blk_off is the exclusive scan of blk_cnt, so
// the list comes out in (row, shard) order — the order the reference walks (template, then its
// dependents, shard by shard).
__global__ void __launch_bounds__(256)
k_emit(DevArena ctrl, uint32_t S, uint64_t nrows, const uint32_t* __restrict__ row_obj,
       const uint32_t* __restrict__ row_dep, const uint8_t* __restrict__ codes,
       const uint32_t* __restrict__ blk_off, ncd_entry* __restrict__ out, uint64_t cap,
       const uint32_t* __restrict__ counters, uint32_t* __restrict__ hdr) {
  // header of a packed diff list: {entries found, shards, rows (low 32 bits), 1 if the buffer was too small}
  if (hdr && blockIdx.x == 0 && threadIdx.x == 0) {
    const uint32_t n = counters[0];
    hdr[0] = n;
    hdr[1] = S;
    hdr[2] = (uint32_t)nrows;
    hdr[3] = n > cap ? 1u : 0u;
  }
  __shared__ uint32_t warp_cnt[8];
  const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t mine = 0;
  uint2 cw = make_uint2(0, 0);  // eight shards: the row's codes in one load
  if (r < nrows) {
    if (S == 8) {
      cw = *reinterpret_cast<const uint2*>(codes + r * 8);
#pragma unroll
      for (int s = 0; s < 8; ++s) mine += (((s < 4 ? cw.x : cw.y) >> (8 * (s & 3))) & 0xffu) != 0;
    } else {
      for (uint32_t s = 0; s < S; ++s) mine += codes[r * S + s] != 0;
    }
  }
  // exclusive scan of the per-row counts inside the block
  uint32_t incl = mine;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= (uint32_t)d) incl += o;
  }
  if (lane == 31) warp_cnt[warp] = incl;
  __syncthreads();
  if (mine == 0) return;
  uint32_t pos = blk_off[blockIdx.x] + incl - mine;
  for (uint32_t w = 0; w < warp; ++w) pos += warp_cnt[w];
  const uint32_t obj = row_obj[r], dep = row_dep[r];
  ncd_entry e;
  e.obj = obj;
  e.dep = dep == NCD_NOTFOUND ? 0u : dep - ctrl.dep_start[obj] + 1u;
  for (uint32_t s = 0; s < S; ++s) {
    const uint32_t code = S == 8 ? (((s < 4 ? cw.x : cw.y) >> (8 * (s & 3))) & 0xffu) : codes[r * S + s];
    if (code == 0) continue;
    e.shard = s;
    e.code = code;
    if (pos < cap) out[pos] = e;  // a caller-sized buffer may be too small: the header says so
    ++pos;
  }
}


// ------------------------------------------------------------------ host side
static int check_scan_args(ncd_ctx* ctx, const ncd_arena* ctrl, const ncd_arena* const* shards,
                           int nshards, uint32_t mode) {
  if (!ctx || !ctrl || (!shards && nshards > 0) || nshards < 0) {
    ncd_set_error("ncd_scan: null argument");
    return NCD_E_INVALID;
  }
  const uint32_t hm = mode & 0xff;
  if (hm > NCD_MODE_XXH64 || (mode & ~(0xffu | NCD_MODE_CONFIRM | NCD_MODE_REHASH))) {
    ncd_set_error("ncd_scan: unknown mode 0x%x", mode);
    return NCD_E_INVALID;
  }
  if (!ctrl->committed || ctrl->ctx != ctx) {
    ncd_set_error("ncd_scan: controller arena not committed on this context");
    return NCD_E_STATE;
  }
  for (int s = 0; s < nshards; ++s)
    if (!shards[s] || !shards[s]->committed || shards[s]->ctx != ctx) {
      ncd_set_error("ncd_scan: shard arena %d not committed on this context", s);
      return NCD_E_STATE;
    }
  if (ctx->scan_pending) {
    ncd_set_error("ncd_scan_launch: a scan is already pending, call ncd_scan_finish first");
    return NCD_E_STATE;
  }
  if (ctrl->nrows * (uint64_t)(nshards ? nshards : 1) >= 0xFFFFFFFFull ||
      ctrl->nobj * (uint64_t)(nshards ? nshards : 1) >= 0xFFFFFFFFull) {
    ncd_set_error("ncd_scan: rows x shards exceeds 2^32");
    return NCD_E_CAPACITY;
  }
  return NCD_OK;
}

// Enqueues one whole scan on the ctx stream.  Codes go to a buffer of rows x shards bytes: the
// diff object's, or (d == nullptr) the tail of the ctx scratch; entries go to `entries_out`
// (capacity `entry_cap`); when `hdr_out` is given a header is written there on the device.
static int scan_enqueue(ncd_ctx* ctx, const ncd_arena* ctrl, const ncd_arena* const* shards, int nshards,
                        uint32_t mode, ncd_diff* d, ncd_entry* entries_out, uint64_t entry_cap, uint32_t* hdr_out) {
  NCD_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  const uint32_t S = (uint32_t)nshards;
  const uint64_t nc = ctrl->nobj, nrows = ctrl->nrows;
  const uint64_t npairs = nc * S, nq = nrows * S;
  const uint32_t hm = mode & 0xff;
  uint32_t launches = 0, variant_used = 0;
  const int slot = (int)(ctx->scan_seq % NCD_EV_RING);
  cudaEvent_t* evs = ctx->evr[slot];

  // digests first (hash modes): every copy is hashed once, digests are kept with the arena
  NCD_CUDA(cudaEventRecord(evs[0], st));
  NCD_RANGE_PUSH("ncd:hash");
  if (hm != NCD_MODE_DIRECT) {
    std::vector<ncd_arena*> all;
    all.push_back(const_cast<ncd_arena*>(ctrl));
    for (uint32_t s = 0; s < S; ++s) all.push_back(const_cast<ncd_arena*>(shards[s]));
    if (mode & NCD_MODE_REHASH)
      for (ncd_arena* a : all) a->h64_valid = a->sha_valid = a->xxh_valid = false;
    NCD_TRY(ncd_hash_launch_many(ctx, all.data(), (int)all.size(), hm, &launches));
  }
  NCD_RANGE_POP();
  NCD_CUDA(cudaEventRecord(evs[1], st));
  NCD_RANGE_PUSH("ncd:join");

  // shard descriptors
  if ((int)S > ctx->d_shards_cap) {
    if (ctx->d_shards) cudaFree(ctx->d_shards);
    ctx->d_shards = nullptr;
    int cap = S < 16 ? 16 : (int)S;
    NCD_CUDA(cudaMalloc((void**)&ctx->d_shards, sizeof(DevArena) * cap));
    ctx->d_shards_cap = cap;
    ctx->h_shards_last_n = 0;  // nothing is on the device yet
  }
  DevArena dctrl;
  ncd_fill_dev_arena(ctrl, &dctrl);
  uint64_t algo = 0;
  std::vector<ncd_arena*> jc_filled;  // shard arenas whose join cache this scan fills
  std::vector<ncd_arena*> jl_filled;  // ... and whose line cache
  {
    DevArena* tmp = (DevArena*)calloc(S ? S : 1, sizeof(DevArena));  // zeroed: the copy is compared bytewise below
    if (!tmp) return NCD_E_NOMEM;
    for (uint32_t s = 0; s < S; ++s) {
      ncd_fill_dev_arena(shards[s], &tmp[s]);
      algo += shards[s]->len_sum;
      // join cache of this (controller, shard) pair
      ncd_arena* sa = const_cast<ncd_arena*>(shards[s]);
      bool listed_before = false;
      for (uint32_t q = 0; q < s; ++q) listed_before = listed_before || shards[q] == shards[s];
      const bool valid = sa->jc_idx && sa->jc_ctrl == ctrl && sa->jc_ctrl_gen == ctrl->generation &&
                         sa->jc_gen == sa->generation && sa->jc_n == nc;
      // the found objects' record offsets / lengths / tags, cached per pair as well (not when the
      // controller holds a key twice: two pairs would share one shard object and only one could be
      // kept current by k_pull_records)
      const bool lines = ctx->line_cache && ctrl->dup_keys == 0 && !listed_before && nc;
      if (valid && lines && sa->jc_line && sa->jc_line_ok) tmp[s].join_line = sa->jc_line;
      else if (lines) {
        if (!sa->jc_line || sa->jc_line_n < nc || sa->jc_rev_n < sa->nobj) {  // first use, or either side grew
          if (sa->jc_line) cudaFree(sa->jc_line);
          if (sa->jc_rev) cudaFree(sa->jc_rev);
          sa->jc_line = nullptr;
          sa->jc_rev = nullptr;
          sa->jc_rev_n = sa->jc_line_n = 0;
          const uint64_t rev_n = sa->nobj + sa->nobj / 8 + 1;
          if (cudaMalloc((void**)&sa->jc_line, nc * sizeof(uint4)) != cudaSuccess ||
              cudaMalloc((void**)&sa->jc_rev, rev_n * sizeof(uint32_t)) != cudaSuccess) {
            cudaGetLastError();
            if (sa->jc_line) cudaFree(sa->jc_line);
            sa->jc_line = nullptr;  // no cache: the scan reads the object lines
          } else {
            sa->jc_line_n = nc;
            sa->jc_rev_n = rev_n;
          }
        }
        sa->jc_line_ok = false;
        if (sa->jc_line) {
          NCD_CUDA(cudaMemsetAsync(sa->jc_rev, 0xFF, sa->jc_rev_n * sizeof(uint32_t), st));
          tmp[s].join_line_out = sa->jc_line;
          tmp[s].join_rev_out = sa->jc_rev;
          jl_filled.push_back(sa);
        }
      }
      if (valid) {
        tmp[s].join_cache = sa->jc_idx;
      } else if (!listed_before && nc) {
        if (sa->jc_n != nc || !sa->jc_idx) {
          if (sa->jc_idx) cudaFree(sa->jc_idx);
          sa->jc_idx = nullptr;
          if (cudaMalloc((void**)&sa->jc_idx, nc * sizeof(uint32_t)) != cudaSuccess) {
            cudaGetLastError();
            sa->jc_idx = nullptr;  // no cache: the scan just probes
          }
        }
        sa->jc_ctrl = nullptr;  // becomes valid once the join that fills it is enqueued (below)
        sa->jc_n = sa->jc_idx ? nc : 0;
        tmp[s].join_cache_out = sa->jc_idx;
        if (sa->jc_idx) jc_filled.push_back(sa);
      }
    }
    // The descriptors change only with a commit, a new cache or another shard list: a steady-state scan finds
    // them on the device already and skips the copy (a few microseconds, which a small scan notices).
    const size_t bytes = sizeof(DevArena) * S;
    if (ctx->h_shards_last && ctx->h_shards_last_n == S && memcmp(ctx->h_shards_last, tmp, bytes) == 0) {
      free(tmp);
    } else {
      // pageable source: the runtime stages it before returning
      cudaError_t e = cudaMemcpyAsync(ctx->d_shards, tmp, bytes, cudaMemcpyHostToDevice, st);
      free(ctx->h_shards_last);
      ctx->h_shards_last = tmp;
      ctx->h_shards_last_n = S;
      NCD_CUDA(e);
    }
  }
  algo += ctrl->len_sum;

  // scratch: sh_off u64[npairs] | sh_idx u32[npairs] | tag u32[npairs] | blk_cnt u32[nblk+1] | blk_off u32[nblk+1] | dv u8[npairs]
  const uint64_t nblk = (nrows + 255) / 256;  // k_classify / k_emit: a thread per row
  size_t o_off = 0;
  size_t o_idx = o_off + ncd_align_up(npairs * 8, 256);
  size_t o_tag = o_idx + ncd_align_up(npairs * 4, 256);
  size_t o_cnt = o_tag + ncd_align_up(npairs * 4, 256);
  size_t o_bo = o_cnt + ncd_align_up((nblk + 1) * 4, 256);
  size_t o_dv = o_bo + ncd_align_up((nblk + 2) * 4, 256);
  size_t o_codes = o_dv + ncd_align_up(npairs, 256);
  size_t total = o_codes + (d ? 0 : ncd_align_up(nq, 256));
  void* base = nullptr;
  NCD_TRY(ncd_ctx_scratch(ctx, total, &base));
  uint64_t* sh_off = (uint64_t*)((char*)base + o_off);
  uint32_t* sh_idx = (uint32_t*)((char*)base + o_idx);
  uint32_t* tag = (uint32_t*)((char*)base + o_tag);
  uint32_t* blk_cnt = (uint32_t*)((char*)base + o_cnt);
  uint32_t* blk_off = (uint32_t*)((char*)base + o_bo);
  uint8_t* dv = (uint8_t*)((char*)base + o_dv);

  uint8_t* codes = d ? d->d_codes : (uint8_t*)base + o_codes;
  uint32_t* counters = (uint32_t*)ctx->d_counters;
  NCD_CUDA(cudaMemsetAsync(counters, 0, 8 * sizeof(uint64_t), st));

  // Which byte-compare kernel this scan takes, from the mean tile fill (see below); small records
  // in direct mode take ONE kernel for join and compare together (k_scan_small).
  int variant = ctx->compare_variant;
  if (variant < 0 && ctrl->ntiles) {
    const uint64_t fill = ctrl->len_sum / ctrl->ntiles;
    variant = fill >= 3500 ? 0 : (fill >= 2500 ? 8 : (fill >= 700 ? 2 : 16));
    // The fused join + compare kernel walks a record with a handful of lanes: measured ahead of the tile
    // kernels up to records of ~1.5 KB (1 KiB: 1.62 vs 1.88 ms per 1M x 8), level at 2 KB — as long as no
    // record is long enough to keep one group of lanes busy after all others are done.
    // With five to eight shards (one group: nine loads per lane in flight) it is also ahead on larger records up
    // to 16 KB — it has no join launch and no tile lists to read (cfg 4: 12.12 vs 12.44 ms with k_compare_tma on
    // one box, cfg 2: 0.162 vs 0.183 ms); with more shards, few shards or long records the tile kernels stay ahead
    // (cfg 3: 64 KB x 16 shards, cfg 5: 1 MiB x 4).
    const bool fusable = hm == NCD_MODE_DIRECT && ctx->fuse_small && ctrl->len_max <= 16384;
    if (fusable && (fill < 1800 || (S >= 5 && S <= 8))) variant = 16;
    if (variant == 16 && ctrl->len_max > 16384) variant = 2;  // mostly small records and a few large ones: tiles spread those
  }
  const bool fused_small = npairs && variant == 16 && hm == NCD_MODE_DIRECT && ctx->fuse_small;
  if (npairs && !fused_small) {
    k_join<<<(unsigned)((npairs + 255) / 256), 256, 0, st>>>(dctrl, ctx->d_shards, S, mode, sh_idx,
                                                             sh_off, dv, tag);
    ++launches;
  }
  NCD_CUDA(cudaEventRecord(evs[2], st));
  NCD_RANGE_POP();
  NCD_RANGE_PUSH("ncd:compare");
  if (fused_small) {
    // as many lanes per object as the mean record has 16-byte vectors: 4 (up to ~100 bytes) to 32 (512 and more)
    const uint64_t mean = nc ? ctrl->len_sum / nc : 0;
    int lpo = mean <= 96 ? 4 : (mean <= 320 ? 8 : (mean <= 2048 ? 16 : 32));  // 32 no better than 16 at 512 and 1024 bytes, ahead at 4 KB
    if (ctx->ss_lpo) lpo = ctx->ss_lpo;
    const uint64_t want = (nc * (uint64_t)lpo + 255) / 256, cap = (uint64_t)ctx->sm_count * 16;
    const unsigned grid = (unsigned)(want < cap ? want : cap);
    const size_t smem = (size_t)S * 64;
#define NCD_SS_LAUNCH(G_, L_) k_scan_small<G_, L_><<<grid, 256, smem, st>>>(dctrl, ctx->d_shards, S, sh_idx, dv, tag)
    if (S <= 4) {
      if (lpo == 4) NCD_SS_LAUNCH(4, 4);
      else if (lpo == 8) NCD_SS_LAUNCH(4, 8);
      else if (lpo == 16) NCD_SS_LAUNCH(4, 16);
      else NCD_SS_LAUNCH(4, 32);
    } else {
      if (lpo == 4) NCD_SS_LAUNCH(8, 4);
      else if (lpo == 8) NCD_SS_LAUNCH(8, 8);
      else if (lpo == 16) NCD_SS_LAUNCH(8, 16);
      else NCD_SS_LAUNCH(8, 32);
    }
#undef NCD_SS_LAUNCH
    ++launches;
    variant_used = 32;
  }
  if (npairs) {
    for (ncd_arena* sa : jc_filled) {
      sa->jc_ctrl = ctrl;
      sa->jc_ctrl_gen = ctrl->generation;
      sa->jc_gen = sa->generation;
    }
    for (ncd_arena* sa : jl_filled) sa->jc_line_ok = true;
  }
  const bool do_compare = !fused_small && ((hm == NCD_MODE_DIRECT) || (mode & NCD_MODE_CONFIRM));
  if (do_compare && npairs && ctrl->ntiles) {
    // Three data-movement shapes for the same compare, picked from the mean tile fill
    // (measured on B200, profiles/r01_variants_and_shapes.txt):
    //   records made of full 4 KB tiles   k_compare: one shard copy at a time, 4 KB contiguous per
    //                                     copy per warp, registers as the staging buffer
    //   ragged records of a few tiles     k_compare_tma: TMA bulk copies into a shared-memory ring,
    //                                     producer / consumer warps, nothing staged in registers
    //   small records                     k_compare_grouped: controller + 8 shard copies in flight
    //                                     per warp, merged tail tiles
    //   records of a few hundred bytes    k_compare_small: the same, eight lanes per tile, four tiles per warp
    variant_used = (uint32_t)variant;
    // the merged tile list (tiles up to 6144 B) belongs to the kernels that walk a tile to its end;
    // k_compare and k_compare_tma compare at most 4096 B per tile and take the fixed 4 KB cut
    const bool fixed_tiles = !(variant == 1 || variant == 2 || variant == 4 || variant == 16);
    const uint64_t ntl = fixed_tiles ? ctrl->ntiles : ctrl->nctiles;
    const uint32_t* t_obj = fixed_tiles ? ctrl->tile_obj : ctrl->ctile_obj;
    const uint32_t* t_chunk = fixed_tiles ? ctrl->tile_chunk : ctrl->ctile_chunk;
    uint64_t want = (ntl + 7) / 8;
    uint64_t cap = (uint64_t)ctx->sm_count * ctx->compare_ctas_per_sm;
    unsigned grid = (unsigned)(want < cap ? want : cap);
    const uint32_t cc = hm != NCD_MODE_DIRECT ? 1u : 0u;
    const size_t smem = (size_t)S * sizeof(void*);
    switch (variant) {
      case 8:  // TMA bulk-copy pipeline, fixed 4 KB tiles
        NCD_TRY(ncd_launch_compare_tma(ctx, dctrl, S, ctrl->tile_obj, ctrl->tile_chunk, ctrl->ntiles, sh_off, dv,
                                       counters, cc));
        break;
      case 1:
        k_compare_grouped<1><<<grid, 256, smem, st>>>(dctrl, ctx->d_shards, S, t_obj, t_chunk, ntl, sh_off, dv,
                                                       counters, cc);
        break;
      case 2:
        k_compare_grouped<2><<<grid, 256, smem, st>>>(dctrl, ctx->d_shards, S, t_obj, t_chunk, ntl, sh_off, dv,
                                                       counters, cc);
        break;
      case 4:
        k_compare_grouped<4><<<grid, 256, smem, st>>>(dctrl, ctx->d_shards, S, t_obj, t_chunk, ntl, sh_off, dv,
                                                       counters, cc);
        break;
      case 16: {  // eight lanes per tile
        const uint64_t want4 = (ntl + 31) / 32;
        k_compare_small<<<(unsigned)(want4 < cap ? want4 : cap), 256, smem, st>>>(dctrl, ctx->d_shards, S, t_obj, t_chunk,
                                                                                  ntl, sh_off, dv, counters, cc);
        break;
      }
      default:
        k_compare<<<grid, 256, 0, st>>>(dctrl, ctx->d_shards, S, t_obj, t_chunk, ntl, sh_off, dv, counters, cc);
    }
    ++launches;
  }
  NCD_CUDA(cudaEventRecord(evs[3], st));
  NCD_RANGE_POP();
  NCD_RANGE(classify_range, "ncd:classify");
  if (nq) {
    const bool fuse = nblk <= 4096;
    k_classify<<<(unsigned)nblk, 256, 0, st>>>(dctrl, ctx->d_shards, S, nrows, ctrl->row_obj,
                                               ctrl->row_dep, ctrl->row_cidx, sh_idx, dv, tag, codes,
                                               blk_cnt, blk_off, counters, fuse ? 1u : 0u);
    if (!fuse) {
      NCD_TRY(ncd_exclusive_scan_u32(ctx, blk_cnt, blk_off, nblk, counters));
      launches += 3;
    }
    k_emit<<<(unsigned)nblk, 256, 0, st>>>(dctrl, S, nrows, ctrl->row_obj, ctrl->row_dep, codes, blk_off,
                                           entries_out, entry_cap, counters, hdr_out);
    launches += 2;
  } else if (hdr_out) {
    const uint32_t empty_hdr[4] = {0, S, (uint32_t)nrows, 0};  // nothing to decide: an empty list
    NCD_CUDA(cudaMemcpyAsync(hdr_out, empty_hdr, sizeof(empty_hdr), cudaMemcpyHostToDevice, st));  // pageable: staged now
  }
  NCD_CUDA(cudaEventRecord(evs[4], st));
  NCD_CUDA(cudaMemcpyAsync(ctx->h_counters + (size_t)slot * 8, counters, 8 * sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
  NCD_CUDA(cudaGetLastError());

  ncd_scan_stats& stt = ctx->stats_ring[slot];
  memset(&stt, 0, sizeof(stt));
  stt.rows = nrows;
  stt.pairs = nq;
  stt.algo_bytes = algo;
  stt.launches = launches;
  stt.compare_variant = variant_used;
  ++ctx->scan_seq;
  return NCD_OK;
}

// Phase timings + counters of the scan enqueued `back` scans ago (0 = the latest).  The scan must
// have completed on the GPU.
static int stats_of(ncd_ctx* ctx, uint64_t back, ncd_scan_stats* out) {
  if (back >= ctx->scan_seq || back >= NCD_EV_RING) {
    ncd_set_error("scan stats: only the last %d scans are kept (asked for %llu back of %llu)", NCD_EV_RING,
                  (unsigned long long)back, (unsigned long long)ctx->scan_seq);
    return NCD_E_INVALID;
  }
  const int slot = (int)((ctx->scan_seq - 1 - back) % NCD_EV_RING);
  cudaEvent_t* evs = ctx->evr[slot];
  ncd_scan_stats st = ctx->stats_ring[slot];
  cudaError_t e = cudaEventElapsedTime(&st.total_ms, evs[0], evs[4]);
  if (e != cudaSuccess) {
    cudaGetLastError();
    ncd_set_error("scan stats: that scan has not finished on the GPU (%s)", cudaGetErrorString(e));
    return NCD_E_STATE;
  }
  cudaEventElapsedTime(&st.hash_ms, evs[0], evs[1]);
  cudaEventElapsedTime(&st.join_ms, evs[1], evs[2]);
  cudaEventElapsedTime(&st.compare_ms, evs[2], evs[3]);
  cudaEventElapsedTime(&st.classify_ms, evs[3], evs[4]);
  const uint32_t* hc = (const uint32_t*)(ctx->h_counters + (size_t)slot * 8);
  st.ndiff = hc[0];
  st.hash_collisions = hc[1];
  *out = st;
  return NCD_OK;
}

int ncd_scan_launch(ncd_ctx* ctx, const ncd_arena* ctrl, const ncd_arena* const* shards, int nshards,
                    uint32_t mode) {
  NCD_TRY(check_scan_args(ctx, ctrl, shards, nshards, mode));
  NCD_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  const uint64_t nq = ctrl->nrows * (uint64_t)nshards;
  ncd_diff* d = (ncd_diff*)calloc(1, sizeof(ncd_diff));
  if (!d) return NCD_E_NOMEM;
  d->ctx = ctx;
  d->rows = ctrl->nrows;
  d->nshards = (uint64_t)nshards;
  cudaError_t e1 = cudaMallocAsync((void**)&d->d_codes, nq ? nq : 1, st);
  cudaError_t e2 = cudaMallocAsync((void**)&d->d_entries, (nq ? nq : 1) * sizeof(ncd_entry), st);
  if (e1 != cudaSuccess || e2 != cudaSuccess) {
    ncd_set_error("ncd_scan: diff buffers (%llu decisions): %s", (unsigned long long)nq,
                  cudaGetErrorString(e1 != cudaSuccess ? e1 : e2));
    if (e1 == cudaSuccess) cudaFreeAsync(d->d_codes, st);
    if (e2 == cudaSuccess) cudaFreeAsync(d->d_entries, st);
    free(d);
    return NCD_E_NOMEM;
  }
  int rc = scan_enqueue(ctx, ctrl, shards, nshards, mode, d, d->d_entries, nq, nullptr);
  if (rc != NCD_OK) {
    cudaFreeAsync(d->d_codes, st);
    cudaFreeAsync(d->d_entries, st);
    free(d);
    return rc;
  }
  ctx->pending = d;
  ctx->scan_pending = true;
  return NCD_OK;
}

int ncd_scan_into(ncd_ctx* ctx, const ncd_arena* ctrl, const ncd_arena* const* shards, int nshards, uint32_t mode,
                  void* dev_out, uint64_t cap_entries) {
  if (!dev_out) {
    ncd_set_error("ncd_scan_into: null output buffer");
    return NCD_E_INVALID;
  }
  NCD_TRY(check_scan_args(ctx, ctrl, shards, nshards, mode));
  NCD_CUDA(cudaSetDevice(ctx->device));
  return scan_enqueue(ctx, ctrl, shards, nshards, mode, nullptr, (ncd_entry*)((char*)dev_out + 16), cap_entries,
                      (uint32_t*)dev_out);
}

// (Re)allocates the context's list buffers for `cap` entries.
static int list_reserve(ncd_ctx* ctx, uint64_t cap) {
  if (cap <= ctx->list_cap) return NCD_OK;
  NCD_CUDA(cudaStreamSynchronize(ctx->stream));
  if (ctx->list_dev) cudaFree(ctx->list_dev);
  if (ctx->list_host) cudaFreeHost(ctx->list_host);
  ctx->list_dev = ctx->list_host = nullptr;
  ctx->list_cap = 0;
  const size_t bytes = 16 + (size_t)cap * sizeof(ncd_entry);
  cudaError_t e = cudaMalloc(&ctx->list_dev, bytes);
  if (e == cudaSuccess) e = cudaMallocHost(&ctx->list_host, bytes);
  if (e != cudaSuccess) {
    ncd_set_error("ncd_scan_host: diff list buffers of %zu bytes: %s", bytes, cudaGetErrorString(e));
    if (ctx->list_dev) cudaFree(ctx->list_dev);
    ctx->list_dev = nullptr;
    return NCD_E_NOMEM;
  }
  ctx->list_cap = cap;
  return NCD_OK;
}

int ncd_scan_host(ncd_ctx* ctx, const ncd_arena* ctrl, const ncd_arena* const* shards, int nshards, uint32_t mode,
                  const ncd_entry** entries, uint64_t* n) {
  if (!entries || !n) {
    ncd_set_error("ncd_scan_host: null output");
    return NCD_E_INVALID;
  }
  NCD_TRY(check_scan_args(ctx, ctrl, shards, nshards, mode));
  NCD_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  NCD_TRY(list_reserve(ctx, 1u << 16));
  // The list is usually a tiny part of the decisions.  One copy brings the header and the first
  // entries back; a longer list needs a second copy, a list beyond the buffer a larger buffer and
  // the scan again (the codes are recomputed, nothing is kept from the short run).
  constexpr uint64_t FIRST = 4096;
  for (int attempt = 0;; ++attempt) {
    NCD_TRY(scan_enqueue(ctx, ctrl, shards, nshards, mode, nullptr, (ncd_entry*)((char*)ctx->list_dev + 16), ctx->list_cap,
                         (uint32_t*)ctx->list_dev));
    const uint64_t first = ctx->list_cap < FIRST ? ctx->list_cap : FIRST;
    NCD_CUDA(cudaMemcpyAsync(ctx->list_host, ctx->list_dev, 16 + first * sizeof(ncd_entry), cudaMemcpyDeviceToHost, st));
    cudaError_t e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) {
      ncd_set_error("scan failed on the GPU: %s", cudaGetErrorString(e));
      return NCD_E_CUDA;
    }
    const uint64_t found = ((const uint32_t*)ctx->list_host)[0];
    if (found > ctx->list_cap) {
      if (attempt) {
        ncd_set_error("ncd_scan_host: the list grew between two scans of the same arenas");
        return NCD_E_STATE;
      }
      NCD_TRY(list_reserve(ctx, found + found / 2));
      continue;
    }
    if (found > first) {
      NCD_CUDA(cudaMemcpyAsync((char*)ctx->list_host + 16 + first * sizeof(ncd_entry),
                               (const char*)ctx->list_dev + 16 + first * sizeof(ncd_entry),
                               (found - first) * sizeof(ncd_entry), cudaMemcpyDeviceToHost, st));
      NCD_CUDA(cudaStreamSynchronize(st));
    }
    NCD_TRY(stats_of(ctx, 0, &ctx->stats));
    *entries = (const ncd_entry*)((const char*)ctx->list_host + 16);
    *n = found;
    return NCD_OK;
  }
}

int ncd_scan_finish(ncd_ctx* ctx, ncd_diff** out) {
  if (!ctx || !out) {
    ncd_set_error("ncd_scan_finish: null argument");
    return NCD_E_INVALID;
  }
  if (!ctx->scan_pending) {
    ncd_set_error("ncd_scan_finish: no scan pending");
    return NCD_E_STATE;
  }
  ncd_diff* d = ctx->pending;
  ctx->pending = nullptr;
  ctx->scan_pending = false;
  cudaError_t e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess) {
    ncd_set_error("scan failed on the GPU: %s", cudaGetErrorString(e));
    free(d);  // device buffers are lost with the context after a sticky error
    return NCD_E_CUDA;
  }
  NCD_TRY(stats_of(ctx, 0, &ctx->stats));
  d->n = ctx->stats.ndiff;
  *out = d;
  return NCD_OK;
}

int ncd_scan(ncd_ctx* ctx, const ncd_arena* ctrl, const ncd_arena* const* shards, int nshards,
             uint32_t mode, ncd_diff** out) {
  if (!out) {
    ncd_set_error("ncd_scan: null out");
    return NCD_E_INVALID;
  }
  NCD_TRY(ncd_scan_launch(ctx, ctrl, shards, nshards, mode));
  return ncd_scan_finish(ctx, out);
}

int ncd_scan_last_stats(ncd_ctx* ctx, ncd_scan_stats* out) { return ncd_scan_stats_back(ctx, 0, out); }

int ncd_scan_stats_back(ncd_ctx* ctx, uint64_t back, ncd_scan_stats* out) {
  if (!ctx || !out) {
    ncd_set_error("ncd_scan_stats_back: null argument");
    return NCD_E_INVALID;
  }
  return stats_of(ctx, back, out);
}

// ------------------------------------------------------------------ diff lists
int ncd_diff_view(ncd_diff* d, const ncd_entry** entries, uint64_t* n) {
  if (!d || !entries || !n) {
    ncd_set_error("ncd_diff_view: null argument");
    return NCD_E_INVALID;
  }
  if (!d->h_valid) {
    NCD_CUDA(cudaSetDevice(d->ctx->device));
    // a pinned buffer from the context's pool if one is large enough, else a new one (power of two)
    const size_t need = (size_t)(d->n ? d->n : 1) * sizeof(ncd_entry);
    ncd_ctx* ctx = d->ctx;
    int pick = -1;
    for (int i = 0; i < 4; ++i)
      if (ctx->hpool_ptr[i] && ctx->hpool_cap[i] >= need && (pick < 0 || ctx->hpool_cap[i] < ctx->hpool_cap[pick])) pick = i;
    if (pick >= 0) {
      d->h_entries = (ncd_entry*)ctx->hpool_ptr[pick];
      d->h_cap = ctx->hpool_cap[pick];
      ctx->hpool_ptr[pick] = nullptr;
    } else {
      size_t cap = 4096;
      while (cap < need) cap <<= 1;
      NCD_CUDA(cudaMallocHost((void**)&d->h_entries, cap));
      d->h_cap = cap;
    }
    if (d->n) {
      NCD_CUDA(cudaMemcpyAsync(d->h_entries, d->d_entries, d->n * sizeof(ncd_entry),
                               cudaMemcpyDeviceToHost, d->ctx->stream));
      NCD_CUDA(cudaStreamSynchronize(d->ctx->stream));
    }
    d->h_valid = true;
  }
  *entries = d->h_entries;
  *n = d->n;
  return NCD_OK;
}

int ncd_diff_device(ncd_diff* d, const void** dev_entries, uint64_t* n) {
  if (!d || !dev_entries || !n) {
    ncd_set_error("ncd_diff_device: null argument");
    return NCD_E_INVALID;
  }
  *dev_entries = d->d_entries;
  *n = d->n;
  return NCD_OK;
}

int ncd_diff_pack(ncd_diff* d, void* dev_buf, uint64_t cap_entries) {
  if (!d || !dev_buf) {
    ncd_set_error("ncd_diff_pack: null argument");
    return NCD_E_INVALID;
  }
  if (d->n > cap_entries) {
    ncd_set_error("ncd_diff_pack: %llu entries, buffer holds %llu", (unsigned long long)d->n,
                  (unsigned long long)cap_entries);
    return NCD_E_CAPACITY;
  }
  NCD_CUDA(cudaSetDevice(d->ctx->device));
  d->hdr[0] = (uint32_t)d->n;
  d->hdr[1] = (uint32_t)d->nshards;
  d->hdr[2] = (uint32_t)d->rows;
  d->hdr[3] = 0;
  // pageable source: staged by the runtime before the call returns
  NCD_CUDA(cudaMemcpyAsync(dev_buf, d->hdr, sizeof(d->hdr), cudaMemcpyHostToDevice, d->ctx->stream));
  if (d->n)
    NCD_CUDA(cudaMemcpyAsync((char*)dev_buf + sizeof(d->hdr), d->d_entries, d->n * sizeof(ncd_entry),
                             cudaMemcpyDeviceToDevice, d->ctx->stream));
  return NCD_OK;
}

int ncd_diff_dirty_objects(ncd_diff* d, uint32_t* obj_out, uint64_t cap, uint64_t* n) {
  if (!d || !n || (!obj_out && cap)) {
    ncd_set_error("ncd_diff_dirty_objects: null argument");
    return NCD_E_INVALID;
  }
  const ncd_entry* e = nullptr;
  uint64_t ne = 0;
  NCD_TRY(ncd_diff_view(d, &e, &ne));
  // the list is in (row, shard) order and rows are grouped by owning object: equal objs are adjacent
  uint64_t count = 0;
  for (uint64_t i = 0; i < ne; ++i) {
    if (i && e[i].obj == e[i - 1].obj) continue;
    if (count < cap) obj_out[count] = e[i].obj;
    ++count;
  }
  *n = count;
  return NCD_OK;
}

int ncd_diff_codes(ncd_diff* d, uint8_t* codes_out, uint64_t cap) {
  if (!d || (!codes_out && cap)) {
    ncd_set_error("ncd_diff_codes: null argument");
    return NCD_E_INVALID;
  }
  const uint64_t nq = d->rows * d->nshards;
  if (cap < nq) {
    ncd_set_error("ncd_diff_codes: buffer holds %llu codes, %llu needed", (unsigned long long)cap,
                  (unsigned long long)nq);
    return NCD_E_CAPACITY;
  }
  if (nq) {
    NCD_CUDA(cudaSetDevice(d->ctx->device));
    NCD_CUDA(cudaMemcpyAsync(codes_out, d->d_codes, nq, cudaMemcpyDeviceToHost, d->ctx->stream));
    NCD_CUDA(cudaStreamSynchronize(d->ctx->stream));
  }
  return NCD_OK;
}

void ncd_diff_destroy(ncd_diff* d) {
  if (!d) return;
  cudaSetDevice(d->ctx->device);
  if (d->d_entries) cudaFreeAsync(d->d_entries, d->ctx->stream);
  if (d->d_codes) cudaFreeAsync(d->d_codes, d->ctx->stream);
  if (d->h_entries) {
    // back to the pool: into a free slot, or in place of the smallest buffer if this one is larger
    ncd_ctx* ctx = d->ctx;
    int slot = -1;
    for (int i = 0; i < 4 && slot < 0; ++i)
      if (!ctx->hpool_ptr[i]) slot = i;
    if (slot < 0) {
      int smallest = 0;
      for (int i = 1; i < 4; ++i)
        if (ctx->hpool_cap[i] < ctx->hpool_cap[smallest]) smallest = i;
      if (ctx->hpool_cap[smallest] < d->h_cap) {
        cudaFreeHost(ctx->hpool_ptr[smallest]);
        slot = smallest;
      }
    }
    if (slot >= 0) {
      ctx->hpool_ptr[slot] = d->h_entries;
      ctx->hpool_cap[slot] = d->h_cap;
    } else {
      cudaFreeHost(d->h_entries);
    }
  }
  free(d);
}
