// ncd_compare_tma.cu — the byte-compare kernel as a TMA bulk-copy pipeline (sm_100a).
//
// Same result as k_compare / k_compare_grouped (reflect.DeepEqual of canonical records,
// controller.go:539,600,725,795), different data movement.  The register-staged kernels can keep
// at most ~128 KB per SM in flight (their load buffers are half the register file) and lose
// bytes-in-flight on ragged tiles, where lanes sit idle.  Here nothing is staged in registers:
//
//   4 producer warps look up the work items (tile, shard group) of this CTA 32 at a time — one
//                   lane per item, so the dependent index loads of 32 items overlap — then, item
//                   by item, wait for the item's stage to be free and issue one cp.async.bulk
//                   (UBLKCP) per copy: the controller tile and up to 8 shard tiles, 16-byte
//                   granular, straight into shared memory, completion counted in bytes on the
//                   stage's mbarrier
//   8 consumer warps wait on the stage's mbarrier, XOR/OR the controller tile against every shard
//                   tile from shared memory (one 16-byte vector per lane per copy), vote, flag
//                   drift with atomicOr, and hand the stage back
//
// As many stages of (1 + min(S, 8)) x 4 KB as fit ~216 KB of shared memory (5 at 8 shards, 10 at 4):
// up to ~200 KB per SM in flight, whatever the record shapes.
#include "ncd_common.cuh"
#include "ncd_tma.cuh"

namespace {

constexpr int TG = 8;                    // shard copies per work item
constexpr int MAX_STAGES = 24;           // stages actually used = what fits ~216 KB of shared memory
#ifndef NCD_TMA_NCONS
#define NCD_TMA_NCONS 8
#endif
#ifndef NCD_TMA_NPROD
#define NCD_TMA_NPROD 4
#endif
constexpr int NCONS = NCD_TMA_NCONS;     // consumer warps
constexpr int NPROD = NCD_TMA_NPROD;     // producer warps: a bulk copy is one warp-wide (uniform) instruction, so
                                         // issue bandwidth scales with warps, not lanes
constexpr uint32_t SENTINEL = 0xFFFFFFFFu;

struct StageDesc {
  uint32_t c;        // controller object
  uint32_t s0;       // first shard of the group
  uint32_t nbytes;   // bytes of the tile (not padded), SENTINEL = no more work
  uint32_t vmask;    // bit k: shard s0 + k was copied
};

struct Smem {
  uint64_t full[MAX_STAGES];
  uint64_t empty[MAX_STAGES];
  StageDesc desc[MAX_STAGES];
};

__global__ void __launch_bounds__(32 * (NPROD + NCONS), 1)
k_compare_tma(DevArena ctrl, const DevArena* __restrict__ shards, uint32_t S,
              const uint32_t* __restrict__ tile_obj, const uint32_t* __restrict__ tile_chunk, uint64_t ntiles,
              const uint64_t* __restrict__ sh_off, uint8_t* __restrict__ dv, uint32_t* __restrict__ counters,
              uint32_t count_collisions, uint32_t NSTAGE, uint32_t STAGE_BYTES) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  uint8_t* data = smem_raw;  // NSTAGE * STAGE_BYTES, 128-byte aligned
  Smem* sm = reinterpret_cast<Smem*>(smem_raw + (size_t)NSTAGE * STAGE_BYTES);
  const uint8_t** s_payload = reinterpret_cast<const uint8_t**>(sm + 1);  // [S]

  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (uint32_t i = 0; i < NSTAGE; ++i) {
      mbar_init(&sm->full[i], 1);
      mbar_init(&sm->empty[i], NCONS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (uint32_t s = threadIdx.x; s < S; s += blockDim.x) s_payload[s] = shards[s].payload;
  __syncthreads();

  const uint32_t ngroups = (S + TG - 1) / TG;
  const uint64_t nitems = ntiles * ngroups;

  // Items of this CTA, in the order the consumers will see them: seq 0, 1, 2, ... is item
  // blockIdx.x + seq * gridDim.x and lives in stage seq % NSTAGE.  Producer warp w owns the seqs
  // congruent to w modulo NPROD; every seq gets a stage (an item with nothing to compare is an
  // empty stage), so the mapping needs no coordination between producers.
  const uint64_t my_items = nitems > blockIdx.x ? (nitems - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

  if (warp < NPROD) {
    // ------------------------------------------------------------ producers
    const uint64_t pol = l2_evict_first_policy();
    // Raw index data of one item per lane.  Loading the next batch is started before the current
    // batch is issued and only consumed an iteration later, so its two dependent load levels
    // (tile -> object -> record / shard offsets) hide behind the copies in flight.
    struct Meta {
      bool in_range;
      uint32_t c, chunk, len, s0;
      uint64_t rec_off, off[TG];
    };
    auto seq_of = [&](uint64_t b) { return (b * 32 + lane) * NPROD + warp; };
    auto load_meta = [&](uint64_t b) {
      Meta m;
      const uint64_t seq = seq_of(b);
      m.in_range = seq < my_items;
      m.c = m.chunk = m.len = m.s0 = 0;
      m.rec_off = 0;
#pragma unroll
      for (int k = 0; k < TG; ++k) m.off[k] = NCD_NO_OFF;
      if (m.in_range) {
        const uint64_t item = (uint64_t)blockIdx.x + seq * gridDim.x;
        const uint64_t t = item / ngroups;
        m.s0 = (uint32_t)(item % ngroups) * TG;
        m.c = tile_obj[t];
        m.chunk = tile_chunk[t];
        m.len = ctrl.rec_len[m.c];
        m.rec_off = ctrl.rec_off[m.c];
#pragma unroll
        for (int k = 0; k < TG; ++k)
          if (m.s0 + k < S) m.off[k] = sh_off[(uint64_t)m.c * S + m.s0 + k];
      }
      return m;
    };
    Meta cur = load_meta(0);
    for (uint64_t b = 0;; ++b) {
      if ((b * 32) * NPROD + warp >= my_items) break;  // lane 0's seq of this batch
      const Meta nxt = load_meta(b + 1);
      const uint64_t toff = (uint64_t)cur.chunk * NCD_TILE_BYTES;
      uint32_t vmask = 0;
#pragma unroll
      for (int k = 0; k < TG; ++k)
        if (cur.off[k] != NCD_NO_OFF) vmask |= 1u << k;
      const uint32_t nbytes = (cur.in_range && vmask) ? min(NCD_TILE_BYTES, cur.len - (uint32_t)toff) : 0u;
      if (nbytes == 0) vmask = 0;
      const uint32_t todo = __ballot_sync(0xffffffffu, cur.in_range);
      for (uint32_t m = todo; m; m &= m - 1) {
        // lane l's item, broadcast: the copies are warp-wide instructions with uniform operands
        const int l = __ffs(m) - 1;
        const uint32_t i_c = __shfl_sync(0xffffffffu, cur.c, l), i_s0 = __shfl_sync(0xffffffffu, cur.s0, l);
        const uint32_t i_nbytes = __shfl_sync(0xffffffffu, nbytes, l), i_vmask = __shfl_sync(0xffffffffu, vmask, l);
        const uint64_t i_src = __shfl_sync(0xffffffffu, cur.rec_off + toff, l);
        const uint64_t i_toff = __shfl_sync(0xffffffffu, toff, l);
        uint64_t i_off[TG];
#pragma unroll
        for (int k = 0; k < TG; ++k) i_off[k] = __shfl_sync(0xffffffffu, cur.off[k], l);
        const uint64_t seq = (b * 32 + (uint64_t)l) * NPROD + warp;
        const uint32_t stage = (uint32_t)(seq % NSTAGE), phase = (uint32_t)((seq / NSTAGE) & 1);
        if (lane == 0) {
          mbar_wait(&sm->empty[stage], phase ^ 1);
          uint8_t* dst = data + (size_t)stage * STAGE_BYTES;
          const uint32_t padded = (i_nbytes + 15) & ~15u;
          sm->desc[stage] = StageDesc{i_c, i_s0, i_nbytes, i_vmask};
          if (i_vmask == 0) {
            mbar_arrive(&sm->full[stage]);  // nothing to compare: an empty stage keeps the order
          } else {
            mbar_arrive_expect_tx(&sm->full[stage], padded * (1 + __popc(i_vmask)));
            bulk_g2s(dst, ctrl.payload + i_src, padded, &sm->full[stage], pol);
#pragma unroll
            for (int k = 0; k < TG; ++k)
              if (i_vmask & (1u << k))
                bulk_g2s(dst + (size_t)(1 + k) * NCD_TILE_BYTES, s_payload[i_s0 + k] + i_off[k] + i_toff, padded,
                         &sm->full[stage], pol);
          }
        }
        __syncwarp();
      }
      cur = nxt;
    }
    if (lane == 0 && my_items % NPROD == warp) {  // seq my_items: tell the consumers there is nothing more
      const uint32_t stage = (uint32_t)(my_items % NSTAGE), phase = (uint32_t)((my_items / NSTAGE) & 1);
      mbar_wait(&sm->empty[stage], phase ^ 1);
      sm->desc[stage] = StageDesc{0, 0, SENTINEL, 0};
      mbar_arrive(&sm->full[stage]);
    }
  } else {
    // ------------------------------------------------------------ consumers
    const uint32_t v = (warp - NPROD) * 32 + lane;  // this lane's 16-byte vector of the tile
    uint32_t stage = 0, phase = 0;
    for (;;) {
      mbar_wait(&sm->full[stage], phase);
      const StageDesc d = sm->desc[stage];
      if (d.nbytes == SENTINEL) break;
      const uint4* base = reinterpret_cast<const uint4*>(data + (size_t)stage * STAGE_BYTES);
      const uint32_t nvec = (d.nbytes + 15) >> 4, tail = d.nbytes & 15;
      uint32_t bits = 0;
      for (uint32_t vv = v; vv < nvec; vv += 32 * NCONS) {  // one trip when 8 consumer warps cover the 256 vectors
        const uint4 a = base[vv];
        uint32_t keep[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};
        if (tail && vv == nvec - 1) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int nb = (int)tail - 4 * i;
            keep[i] = nb <= 0 ? 0u : (nb < 4 ? (1u << (8 * nb)) - 1u : 0xffffffffu);
          }
        }
#pragma unroll
        for (int k = 0; k < TG; ++k) {
          if (d.vmask & (1u << k)) {
            const uint4 b = base[(size_t)(1 + k) * NCD_TILE_VECS + vv];
            const uint32_t x = ((a.x ^ b.x) & keep[0]) | ((a.y ^ b.y) & keep[1]) | ((a.z ^ b.z) & keep[2]) |
                               ((a.w ^ b.w) & keep[3]);
            if (x) bits |= 1u << k;
          }
        }
      }
      bits = __reduce_or_sync(0xffffffffu, bits);
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&sm->empty[stage]);  // every read of this stage by this warp is done
        for (uint32_t m = bits; m; m &= m - 1) {
          const uint32_t k = __ffs(m) - 1;
          const uint64_t p = (uint64_t)d.c * S + d.s0 + k;
          uint32_t* word = reinterpret_cast<uint32_t*>(dv) + (p >> 2);
          const uint32_t sh8 = 8u * (uint32_t)(p & 3);
          const uint32_t old = atomicOr(word, (uint32_t)DV_DIFF << sh8);
          if (count_collisions && ((old >> sh8) & 0xffu) == DV_EQUAL) atomicAdd(&counters[1], 1u);
        }
      }
      if (++stage == NSTAGE) {
        stage = 0;
        phase ^= 1;
      }
    }
  }
}

}  // namespace

// Stage = the controller tile + one tile per shard of a group; as many stages as fit.
static void tma_shape(uint32_t S, uint32_t* nstage, uint32_t* stage_bytes, size_t* smem) {
  const uint32_t copies = 1 + (S < (uint32_t)TG ? S : (uint32_t)TG);
  *stage_bytes = copies * NCD_TILE_BYTES;
  const size_t fixed = sizeof(Smem) + (size_t)S * sizeof(void*) + 256;
  size_t n = (216 * 1024 - fixed) / *stage_bytes;
  if (n > MAX_STAGES) n = MAX_STAGES;
  *nstage = (uint32_t)n;
  *smem = n * *stage_bytes + fixed;
}

int ncd_launch_compare_tma(ncd_ctx* ctx, const DevArena& dctrl, uint32_t S, const uint32_t* tile_obj,
                           const uint32_t* tile_chunk, uint64_t ntiles, const uint64_t* sh_off, uint8_t* dv,
                           uint32_t* counters, uint32_t count_collisions) {
  uint32_t nstage, stage_bytes;
  size_t smem;
  tma_shape(S, &nstage, &stage_bytes, &smem);
  // A producer warp owns every NPROD-th item and waits on mbarrier phase PARITY, which cannot tell
  // two phases apart: with fewer stages than producers its next item would alias the stage of an
  // item still in flight.  (1 + 8) x 4 KB stages always give at least 5.
  if (nstage < (uint32_t)NPROD || smem > 227 * 1024) {
    ncd_set_error("compare: %u shards do not fit the shared-memory pipeline", S);
    return NCD_E_CAPACITY;
  }
  NCD_CUDA(cudaFuncSetAttribute(k_compare_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  const uint64_t ngroups = (S + TG - 1) / TG;
  const uint64_t nitems = ntiles * ngroups;
  const unsigned grid = (unsigned)(nitems < (uint64_t)ctx->sm_count ? nitems : (uint64_t)ctx->sm_count);
  k_compare_tma<<<grid, 32 * (NPROD + NCONS), smem, ctx->stream>>>(dctrl, ctx->d_shards, S, tile_obj, tile_chunk, ntiles,
                                                                   sh_off, dv, counters, count_collisions, nstage,
                                                                   stage_bytes);
  return NCD_OK;
}
