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

// ------------------------------------------------------------------ controller-side adoption
// adoptReferences (controller.go:647-695) runs at the top of every templateSyncHandler call: a
// Secret / ConfigMap the template references on the CONTROLLER cluster that carries no
// ownerReference with the template's UID (isOwnedBy, :637-645 — UID only, any Kind) gets one.
// The per-shard decisions never see that, so a sweep feed built from them alone would never hand
// such a template to its handler again.  One thread per dependent row of the controller arena:
// the row's controller object (row_cidx) is scanned for an owner whose interned UID id equals the
// template's.  UIDs were interned at commit: equal ids <=> byte-equal strings.
__global__ void __launch_bounds__(256)
k_unadopted(DevArena ctrl, uint64_t nrows, const uint32_t* __restrict__ row_obj, const uint32_t* __restrict__ row_dep,
            const uint32_t* __restrict__ row_cidx, ncd_adoption* __restrict__ out, uint64_t cap,
            unsigned long long* __restrict__ count) {
  const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= nrows) return;
  const uint32_t dep = row_dep[r], cidx = row_cidx[r];
  if (dep == NCD_NOTFOUND || cidx == NCD_NOTFOUND) return;  // the object itself / a missing reference (:649-653 errors out)
  const uint32_t obj = row_obj[r];
  const uint32_t want = ctrl.uid_id[obj];
  for (uint32_t o = ctrl.own_start[cidx]; o < ctrl.own_start[cidx + 1]; ++o)
    if ((ctrl.own_packed[o] & 0x7FFFFFFFu) == want) return;
  const unsigned long long pos = atomicAdd(count, 1ull);
  if (pos < cap) out[pos] = ncd_adoption{obj, dep - ctrl.dep_start[obj] + 1u};
}

int ncd_find_unadopted(ncd_ctx* ctx, const ncd_arena* ctrl, ncd_adoption* out, uint64_t cap, uint64_t* n) {
  if (!ctx || !ctrl || !n || (!out && cap)) {
    ncd_set_error("ncd_find_unadopted: null argument");
    return NCD_E_INVALID;
  }
  if (!ctrl->committed || ctrl->ctx != ctx) {
    ncd_set_error("ncd_find_unadopted: controller arena not committed on this context");
    return NCD_E_STATE;
  }
  *n = 0;
  if (!ctrl->nrows) return NCD_OK;
  NCD_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  ncd_adoption* d_out = nullptr;
  unsigned long long* d_count = nullptr;
  NCD_CUDA(cudaMallocAsync((void**)&d_out, (cap ? cap : 1) * sizeof(ncd_adoption), st));
  NCD_CUDA(cudaMallocAsync((void**)&d_count, 8, st));
  NCD_CUDA(cudaMemsetAsync(d_count, 0, 8, st));
  DevArena dc;
  ncd_fill_dev_arena(ctrl, &dc);
  k_unadopted<<<(unsigned)((ctrl->nrows + 255) / 256), 256, 0, st>>>(dc, ctrl->nrows, ctrl->row_obj, ctrl->row_dep, ctrl->row_cidx,
                                                                      d_out, cap, d_count);
  unsigned long long found = 0;
  NCD_CUDA(cudaMemcpyAsync(&found, d_count, 8, cudaMemcpyDeviceToHost, st));
  NCD_CUDA(cudaStreamSynchronize(st));
  const uint64_t have = found < cap ? found : cap;
  if (have) NCD_CUDA(cudaMemcpyAsync(out, d_out, have * sizeof(ncd_adoption), cudaMemcpyDeviceToHost, st));
  NCD_CUDA(cudaFreeAsync(d_out, st));
  NCD_CUDA(cudaFreeAsync(d_count, st));
  NCD_CUDA(cudaStreamSynchronize(st));
  NCD_CUDA(cudaGetLastError());
  std::sort(out, out + have, [](const ncd_adoption& a, const ncd_adoption& b) { return a.obj != b.obj ? a.obj < b.obj : a.dep < b.dep; });
  *n = found;
  return NCD_OK;
}
