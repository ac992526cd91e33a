// ncd_orphans.cu — the reverse join: objects present in a shard's
// arena and absent from the controller's.  No counterpart in the reference's sync handlers: they
// only ask a shard about names the controller knows (controller.go:516,578,722,792), and a
// deletion reaches the shards only through the delete event (handleObject, :196-205).
#include <algorithm>

#include "ncd_common.cuh"

// One thread per shard object: probe the controller's table with the object's own line.
__global__ void __launch_bounds__(256)
k_orphans(DevArena ctrl, DevArena shard, uint32_t s, ncd_orphan* __restrict__ out, uint64_t cap,
          unsigned long long* __restrict__ count) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= shard.nobj) return;
  if (ncd_probe_meta(ctrl, shard.meta[i], shard.aux) != NCD_NOTFOUND) return;
  const unsigned long long pos = atomicAdd(count, 1ull);
  if (pos < cap) out[pos] = ncd_orphan{s, i};
}

int ncd_find_orphans(ncd_ctx* ctx, const ncd_arena* ctrl, const ncd_arena* const* shards, int nshards,
                     ncd_orphan* out, uint64_t cap, uint64_t* n) {
  if (!ctx || !ctrl || !n || (!shards && nshards > 0) || nshards < 0 || (!out && cap)) {
    ncd_set_error("ncd_find_orphans: null argument");
    return NCD_E_INVALID;
  }
  if (!ctrl->committed || ctrl->ctx != ctx) {
    ncd_set_error("ncd_find_orphans: controller arena not committed on this context");
    return NCD_E_STATE;
  }
  for (int s = 0; s < nshards; ++s)
    if (!shards[s] || !shards[s]->committed || shards[s]->ctx != ctx) {
      ncd_set_error("ncd_find_orphans: shard arena %d not committed on this context", s);
      return NCD_E_STATE;
    }
  NCD_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  ncd_orphan* d_out = nullptr;
  unsigned long long* d_count = nullptr;
  NCD_CUDA(cudaMallocAsync((void**)&d_out, (cap ? cap : 1) * sizeof(ncd_orphan), st));
  NCD_CUDA(cudaMallocAsync((void**)&d_count, 8, st));
  NCD_CUDA(cudaMemsetAsync(d_count, 0, 8, st));
  DevArena dc;
  ncd_fill_dev_arena(ctrl, &dc);
  for (int s = 0; s < nshards; ++s) {
    if (!shards[s]->nobj) continue;
    DevArena ds;
    ncd_fill_dev_arena(shards[s], &ds);
    k_orphans<<<(unsigned)((shards[s]->nobj + 255) / 256), 256, 0, st>>>(dc, ds, (uint32_t)s, d_out, cap, d_count);
  }
  unsigned long long found = 0;
  NCD_CUDA(cudaMemcpyAsync(&found, d_count, 8, cudaMemcpyDeviceToHost, st));
  NCD_CUDA(cudaStreamSynchronize(st));
  const uint64_t have = found < cap ? found : cap;
  if (have) NCD_CUDA(cudaMemcpyAsync(out, d_out, have * sizeof(ncd_orphan), cudaMemcpyDeviceToHost, st));
  NCD_CUDA(cudaFreeAsync(d_out, st));
  NCD_CUDA(cudaFreeAsync(d_count, st));
  NCD_CUDA(cudaStreamSynchronize(st));
  NCD_CUDA(cudaGetLastError());
  // appended in whatever order the threads ran: put the list in (shard, object) order
  std::sort(out, out + have, [](const ncd_orphan& a, const ncd_orphan& b) {
    return a.shard != b.shard ? a.shard < b.shard : a.obj < b.obj;
  });
  *n = found;
  return NCD_OK;
}
