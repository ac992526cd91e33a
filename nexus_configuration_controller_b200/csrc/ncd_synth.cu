// ncd_synth.cu — on-device generator of the synthetic workloads of SURVEY.md section 8(d)
// (bench and test utility; 50-110 GB of input cannot be generated on the host and shipped).
//
// A config is `groups` templates, each owning n_secrets Secrets and n_configmaps ConfigMaps.
// Logical object id gid = g * per + j (per = 1 + n_secrets + n_configmaps; j = 0 template,
// 1..n_secrets secrets, then configmaps).  Everything about an object is a pure function of
// (cfg, gid) and, for a shard copy, of the shard index, so any sample can be regenerated.
// The partition filter keeps a group iff ncd_keyhash(template key) % part_mod == part_rem:
// the multi-GPU layout (a template's dependents travel with it).
#include "ncd_common.cuh"

int ncd_arena_commit_device(ncd_arena* a, uint64_t nobj, uint64_t payload_bytes, uint64_t aux_bytes,
                            uint64_t nowner, uint64_t ndep);

namespace {

constexpr uint32_t SLOT = 64;      // aux bytes per object: key at +0, uid at +32
constexpr uint32_t UID_AT = 32;
constexpr uint64_t PPM = 1000000;

struct SynthDev {
  ncd_synth_cfg c;
  int role;           // -1 controller, >= 0 shard index
  uint32_t per, nd;   // objects per group, dependents per group
  uint64_t nsel;      // kept groups
  uint64_t n;         // objects in this arena
  uint64_t perm_a, perm_b;  // stored position = (perm_a * i + perm_b) % n
};

__host__ __device__ inline uint32_t kind_of(const SynthDev& s, uint32_t j) {
  return j == 0 ? NCD_KIND_TEMPLATE : (j <= s.c.n_secrets ? NCD_KIND_SECRET : NCD_KIND_CONFIGMAP);
}

__host__ __device__ inline uint32_t put_dec(uint8_t* p, uint64_t v, int digits) {
  for (int i = digits - 1; i >= 0; --i) {
    p[i] = (uint8_t)('0' + v % 10);
    v /= 10;
  }
  return (uint32_t)digits;
}

// "nexus/t-0000000042", "nexus/s-0000000042-00", "nexus/c-0000000042-01"; 'x' when the
// object is absent from this shard (same slot, a name the controller never asks for).
__host__ __device__ inline uint32_t make_key(const SynthDev& s, uint64_t g, uint32_t j, bool gone, uint8_t* out) {
  const char pre[6] = {'n', 'e', 'x', 'u', 's', '/'};
  for (int i = 0; i < 6; ++i) out[i] = (uint8_t)pre[i];
  const uint32_t k = kind_of(s, j);
  out[6] = gone ? 'x' : (k == NCD_KIND_TEMPLATE ? 't' : (k == NCD_KIND_SECRET ? 's' : 'c'));
  out[7] = '-';
  uint32_t n = 8 + put_dec(out + 8, g, 10);
  if (j > 0) {
    out[n++] = '-';
    n += put_dec(out + n, k == NCD_KIND_SECRET ? j - 1 : j - 1 - s.c.n_secrets, 2);
  }
  return n;
}

__host__ __device__ inline uint32_t make_uid(const SynthDev& s, uint64_t g, uint32_t j, uint8_t* out) {
  out[0] = 'u'; out[1] = 'i'; out[2] = 'd'; out[3] = '-';
  uint32_t n = 4;
  if (s.role < 0) out[n++] = 'c';
  else n += put_dec(out + n, (uint64_t)s.role, 3);
  out[n++] = '-';
  n += put_dec(out + n, g, 10);
  out[n++] = '-';
  n += put_dec(out + n, j, 2);
  return n;
}

struct Drift {
  bool missing, flip, longer, rogue, foreign;
  uint64_t d;
};
__host__ __device__ inline Drift drift_of(const SynthDev& s, uint64_t gid, uint32_t j) {
  Drift r = {false, false, false, false, false, 0};
  if (s.role < 0) return r;
  const uint64_t d = ncd_splitmix64(s.c.drift_seed + gid * 0x9E3779B97F4A7C15ull +
                                    (uint64_t)s.role * 0xD1B54A32D192ED03ull);
  r.d = d;
  const uint64_t u = d % PPM;
  const uint64_t pm = s.c.p_missing_ppm, pf = s.c.p_flip_ppm, pl = s.c.p_len_ppm;
  if (u < pm) r.missing = true;
  else if (u < pm + pf) r.flip = true;
  else if (u < pm + pf + pl) r.longer = true;
  if (j > 0) {
    const uint64_t u2 = ncd_splitmix64(d ^ 0x5851F42D4C957F2Dull) % PPM;
    if (u2 < s.c.p_rogue_ppm) r.rogue = true;
    else if (u2 < (uint64_t)s.c.p_rogue_ppm + s.c.p_owner_ppm) r.foreign = true;
  }
  return r;
}

__host__ __device__ inline uint32_t base_len(const SynthDev& s, uint64_t gid, uint32_t j) {
  const uint32_t lo = j == 0 ? s.c.tmpl_len_lo : s.c.dep_len_lo;
  const uint32_t hi = j == 0 ? s.c.tmpl_len_hi : s.c.dep_len_hi;
  const uint64_t span = (uint64_t)hi - lo + 1;
  return lo + (uint32_t)(ncd_splitmix64(s.c.seed ^ 0xA5A5A5A55A5A5A5Aull ^ (gid * 0xC2B2AE3D27D4EB4Full)) % span);
}

__device__ inline uint64_t pos_of(const SynthDev& s, uint64_t i) {
  if (s.role < 0 || !s.c.shuffle) return i;
  return (uint64_t)(((unsigned __int128)s.perm_a * i + s.perm_b) % s.n);
}

// ---- kernels
__global__ void k_select(ncd_synth_cfg c, uint32_t* __restrict__ flags) {
  const uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= c.groups) return;
  uint32_t keep = 1;
  if (c.part_mod > 1) {
    SynthDev s;
    s.c = c;
    uint8_t key[32];
    const uint32_t n = make_key(s, g, 0, false, key);
    keep = (ncd_keyhash_impl(NCD_KIND_TEMPLATE, key, n) % c.part_mod) == c.part_rem;
  }
  flags[g] = keep;
}

__global__ void k_gather_sel(uint64_t groups, const uint32_t* __restrict__ flags,
                             const uint32_t* __restrict__ excl, uint32_t* __restrict__ gsel) {
  const uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= groups) return;
  if (flags[g]) gsel[excl[g]] = (uint32_t)g;
}

// one thread per logical object: metadata + padded length at its stored position
__global__ void k_meta(SynthDev s, const uint32_t* __restrict__ gsel, ncd_arena_view d,
                       uint32_t* __restrict__ lenpad, uint32_t* __restrict__ own_cnt,
                       uint32_t* __restrict__ logical) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= s.n) return;
  const uint64_t k = i / s.per;
  const uint32_t j = (uint32_t)(i % s.per);
  const uint64_t g = gsel[k], gid = g * s.per + j;
  const uint64_t pos = pos_of(s, i);
  const Drift dr = drift_of(s, gid, j);
  const uint32_t len = base_len(s, gid, j) + (dr.longer ? 1u : 0u);
  d.rec_len[pos] = len;
  lenpad[pos] = (uint32_t)((len + 15u) & ~15u);
  d.kind[pos] = (uint8_t)kind_of(s, j);
  uint8_t* slot = d.aux + pos * SLOT;
  d.key_off[pos] = (uint32_t)(pos * SLOT);
  d.key_len[pos] = make_key(s, g, j, dr.missing, slot);
  d.uid_off[pos] = (uint32_t)(pos * SLOT + UID_AT);
  d.uid_len[pos] = make_uid(s, g, j, slot + UID_AT);
  // dependents carry their template's ownerReference: on a shard unless the drift says rogue, on the
  // controller always (a controller cluster whose references adoptReferences, controller.go:647-695, has adopted)
  own_cnt[pos] = (j > 0 && (s.role < 0 || !dr.rogue)) ? 1u : 0u;
  logical[pos] = (uint32_t)i;
  if (s.role < 0) {
    // dependents of the template: secrets first, then configmaps (GetSecretNames /
    // GetConfigMapNames order, controller.go:505,567)
    d.dep_start[i] = (uint32_t)(j == 0 ? k * s.nd : (k + 1) * s.nd);
    if (j > 0) {
      const uint64_t e = k * s.nd + (j - 1);
      d.dep_kind[e] = (uint8_t)kind_of(s, j);
      d.dep_key_off[e] = (uint32_t)(pos * SLOT);
      d.dep_key_len[e] = d.key_len[pos];
    }
    if (i == s.n - 1) d.dep_start[s.n] = (uint32_t)(s.nsel * s.nd);
  } else {
    d.dep_start[pos] = 0;
    if (i == s.n - 1) d.dep_start[s.n] = 0;
  }
  if (i == 0) {
    const char f[12] = "uid-foreign";
    for (int b = 0; b < 11; ++b) d.aux[s.n * SLOT + b] = (uint8_t)f[b];
  }
}

__global__ void k_owners(SynthDev s, const uint32_t* __restrict__ gsel, ncd_arena_view d,
                         const uint32_t* __restrict__ logical) {
  const uint64_t pos = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (pos >= s.n) return;
  const uint32_t o0 = d.own_start[pos], o1 = d.own_start[pos + 1];
  if (o0 == o1) return;
  const uint64_t i = logical[pos], k = i / s.per;
  const uint32_t j = (uint32_t)(i % s.per);
  const uint64_t gid = (uint64_t)gsel[k] * s.per + j;
  const Drift dr = drift_of(s, gid, j);
  d.own_is_template[o0] = 1;
  if (dr.foreign) {
    d.own_uid_off[o0] = (uint32_t)(s.n * SLOT);
    d.own_uid_len[o0] = 11;
  } else {
    const uint64_t tpos = pos_of(s, k * s.per);
    d.own_uid_off[o0] = d.uid_off[tpos];
    d.own_uid_len[o0] = d.uid_len[tpos];
  }
}

// one CTA per stored object writes its canonical bytes (a counter-based stream) as 16-byte
// vectors; a flipped copy differs in exactly one byte, a longer copy has one more byte
__global__ void __launch_bounds__(128)
k_payload(SynthDev s, const uint32_t* __restrict__ gsel, ncd_arena_view d,
          const uint32_t* __restrict__ logical) {
  const uint64_t pos = blockIdx.x;
  const uint64_t i = logical[pos], k = i / s.per;
  const uint32_t j = (uint32_t)(i % s.per);
  const uint64_t gid = (uint64_t)gsel[k] * s.per + j;
  const Drift dr = drift_of(s, gid, j);
  const uint32_t len = d.rec_len[pos];
  const uint32_t nvec = (len + 15) >> 4;
  const uint64_t h = ncd_splitmix64(s.c.seed * 0x2545F4914F6CDD1Dull + gid);
  uint64_t flip_at = ~0ull;
  uint8_t flip_mask = 0;
  if (dr.flip) {
    flip_at = ncd_splitmix64(dr.d ^ 0x1234567ull) % len;
    flip_mask = (uint8_t)(((dr.d >> 8) & 0xff) | 1);
  }
  uint4* out = reinterpret_cast<uint4*>(d.payload + d.rec_off[pos]);
  for (uint32_t v = threadIdx.x; v < nvec; v += blockDim.x) {
    uint64_t w[2] = {ncd_splitmix64(h + 2ull * v), ncd_splitmix64(h + 2ull * v + 1)};
    uint8_t* b = reinterpret_cast<uint8_t*>(w);
    const uint64_t base = 16ull * v;
    if (flip_at >= base && flip_at < base + 16) b[flip_at - base] ^= flip_mask;
    if (base + 16 > len)
      for (uint32_t q = (uint32_t)(len - base); q < 16; ++q) b[q] = 0;
    out[v] = make_uint4((uint32_t)w[0], (uint32_t)(w[0] >> 32), (uint32_t)w[1], (uint32_t)(w[1] >> 32));
  }
}

// exact payload bytes of an arena: sum of the padded record lengths of the kept objects
__global__ void k_size_sum(SynthDev s, const uint32_t* __restrict__ gsel, unsigned long long* __restrict__ out) {
  unsigned long long acc = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < s.n; i += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t k = i / s.per;
    const uint32_t j = (uint32_t)(i % s.per);
    const uint64_t gid = (uint64_t)gsel[k] * s.per + j;
    const Drift dr = drift_of(s, gid, j);
    acc += (base_len(s, gid, j) + (dr.longer ? 1u : 0u) + 15u) & ~15u;
  }
  for (int dlt = 16; dlt > 0; dlt >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, dlt);
  if ((threadIdx.x & 31) == 0 && acc) atomicAdd(out, acc);
}

uint64_t gcd64(uint64_t a, uint64_t b) {
  while (b) {
    uint64_t t = a % b;
    a = b;
    b = t;
  }
  return a;
}

int check_cfg(const ncd_synth_cfg* c, int role) {
  if (!c) {
    ncd_set_error("synth: null cfg");
    return NCD_E_INVALID;
  }
  if (c->tmpl_len_lo == 0 || c->tmpl_len_hi < c->tmpl_len_lo ||
      ((c->n_secrets || c->n_configmaps) && (c->dep_len_lo == 0 || c->dep_len_hi < c->dep_len_lo))) {
    ncd_set_error("synth: length ranges must be 1 <= lo <= hi");
    return NCD_E_INVALID;
  }
  if ((uint64_t)c->p_flip_ppm + c->p_missing_ppm + c->p_len_ppm > PPM ||
      (uint64_t)c->p_owner_ppm + c->p_rogue_ppm > PPM) {
    ncd_set_error("synth: drift probabilities exceed 1");
    return NCD_E_INVALID;
  }
  if (c->n_secrets + c->n_configmaps > 99 || role < -1 || role > 999 ||
      (c->part_mod > 1 && c->part_rem >= c->part_mod) || c->groups >= 0xFFFFFFFFull) {
    ncd_set_error("synth: unsupported shape (<= 99 dependents per template, role -1..999, "
                  "part_rem < part_mod, groups < 2^32)");
    return NCD_E_INVALID;
  }
  return NCD_OK;
}

// Runs the selection; returns the kept group list (device, caller frees with cudaFreeAsync).
int select_groups(ncd_ctx* ctx, const ncd_synth_cfg* c, uint32_t** gsel_out, uint64_t* nsel_out) {
  cudaStream_t st = ctx->stream;
  const uint64_t G = c->groups;
  uint32_t *flags = nullptr, *excl = nullptr, *gsel = nullptr;
  NCD_CUDA(cudaMallocAsync((void**)&flags, (G + 1) * 4, st));
  NCD_CUDA(cudaMallocAsync((void**)&excl, (G + 1) * 4, st));
  if (G) k_select<<<(unsigned)((G + 255) / 256), 256, 0, st>>>(*c, flags);
  NCD_TRY(ncd_exclusive_scan_u32(ctx, flags, excl, G, nullptr));
  uint32_t nsel = 0;
  NCD_CUDA(cudaMemcpyAsync(&nsel, excl + G, 4, cudaMemcpyDeviceToHost, st));
  NCD_CUDA(cudaStreamSynchronize(st));
  NCD_CUDA(cudaMallocAsync((void**)&gsel, ((uint64_t)nsel + 1) * 4, st));
  if (G) k_gather_sel<<<(unsigned)((G + 255) / 256), 256, 0, st>>>(G, flags, excl, gsel);
  NCD_CUDA(cudaFreeAsync(flags, st));
  NCD_CUDA(cudaFreeAsync(excl, st));
  *gsel_out = gsel;
  *nsel_out = nsel;
  return NCD_OK;
}

void make_dev(const ncd_synth_cfg* c, int role, uint64_t nsel, SynthDev* s) {
  memset(s, 0, sizeof(*s));
  s->c = *c;
  s->role = role;
  s->nd = c->n_secrets + c->n_configmaps;
  s->per = 1 + s->nd;
  s->nsel = nsel;
  s->n = nsel * s->per;
  static const uint64_t primes[] = {1000003ull, 7919ull, 104729ull, 15485863ull, 32452843ull,
                                    49979687ull, 67867967ull, 86028121ull, 2147483647ull, 13ull};
  uint64_t a = 1;
  if (s->n > 1)
    for (int t = 0; t < 10; ++t) {
      a = primes[((role < 0 ? 0 : role) + t) % 10] % s->n;
      if (a > 1 && gcd64(a, s->n) == 1) break;
      a = 1;
    }
  s->perm_a = a;
  s->perm_b = s->n ? ((uint64_t)(role + 2) * 7919ull) % s->n : 0;
}

}  // namespace

extern "C" int ncd_synth_sizes(ncd_ctx* ctx, const ncd_synth_cfg* cfg, int role, uint64_t* payload_bytes,
                               uint64_t* nobj, uint64_t* aux_bytes, uint64_t* nowner, uint64_t* ndep) {
  if (!ctx) {
    ncd_set_error("ncd_synth_sizes: null ctx");
    return NCD_E_INVALID;
  }
  NCD_TRY(check_cfg(cfg, role));
  NCD_CUDA(cudaSetDevice(ctx->device));
  uint32_t* gsel = nullptr;
  uint64_t nsel = 0;
  NCD_TRY(select_groups(ctx, cfg, &gsel, &nsel));
  SynthDev s;
  make_dev(cfg, role, nsel, &s);
  unsigned long long* d_sum = nullptr;
  unsigned long long h_sum = 0;
  NCD_CUDA(cudaMallocAsync((void**)&d_sum, 8, ctx->stream));
  NCD_CUDA(cudaMemsetAsync(d_sum, 0, 8, ctx->stream));
  if (s.n) {
    uint64_t blocks = (s.n + 255) / 256;
    if (blocks > 4096) blocks = 4096;
    k_size_sum<<<(unsigned)blocks, 256, 0, ctx->stream>>>(s, gsel, d_sum);
  }
  NCD_CUDA(cudaMemcpyAsync(&h_sum, d_sum, 8, cudaMemcpyDeviceToHost, ctx->stream));
  NCD_CUDA(cudaStreamSynchronize(ctx->stream));
  NCD_CUDA(cudaFreeAsync(d_sum, ctx->stream));
  NCD_CUDA(cudaFreeAsync(gsel, ctx->stream));
  if (payload_bytes) *payload_bytes = h_sum;
  if (nobj) *nobj = s.n;
  if (aux_bytes) *aux_bytes = (s.n + 1) * SLOT;
  if (nowner) *nowner = nsel * s.nd;
  if (ndep) *ndep = role < 0 ? nsel * s.nd : 0;
  return NCD_OK;
}

extern "C" int ncd_synth_fill(ncd_ctx* ctx, ncd_arena* a, const ncd_synth_cfg* cfg, int role) {
  if (!ctx || !a || a->ctx != ctx) {
    ncd_set_error("ncd_synth_fill: null argument or arena of another context");
    return NCD_E_INVALID;
  }
  NCD_TRY(check_cfg(cfg, role));
  NCD_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  uint32_t* gsel = nullptr;
  uint64_t nsel = 0;
  NCD_TRY(select_groups(ctx, cfg, &gsel, &nsel));
  SynthDev s;
  make_dev(cfg, role, nsel, &s);
  const uint64_t n = s.n;
  if (n > a->nobj_cap || (n + 1) * SLOT > a->aux_cap || nsel * s.nd > a->nowner_cap ||
      (role < 0 && nsel * s.nd > a->ndep_cap)) {
    cudaFreeAsync(gsel, st);
    ncd_set_error("ncd_synth_fill: arena too small for %llu objects (size it with ncd_synth_sizes)",
                  (unsigned long long)n);
    return NCD_E_CAPACITY;
  }
  a->committed = false;
  uint32_t *lenpad = nullptr, *own_cnt = nullptr, *logical = nullptr;
  unsigned long long* d_tot = nullptr;
  NCD_CUDA(cudaMallocAsync((void**)&lenpad, (n + 1) * 4, st));
  NCD_CUDA(cudaMallocAsync((void**)&own_cnt, (n + 1) * 4, st));
  NCD_CUDA(cudaMallocAsync((void**)&logical, (n + 1) * 4, st));
  NCD_CUDA(cudaMallocAsync((void**)&d_tot, 16, st));
  NCD_CUDA(cudaMemsetAsync(d_tot, 0, 16, st));
  const unsigned blocks = (unsigned)((n + 255) / 256);
  if (n) k_meta<<<blocks, 256, 0, st>>>(s, gsel, a->d, lenpad, own_cnt, logical);
  else NCD_CUDA(cudaMemsetAsync(a->d.dep_start, 0, 4, st));
  // rec_off = exclusive scan of the padded lengths (u64), own_start likewise (u32)
  uint64_t* off_tmp = nullptr;
  NCD_CUDA(cudaMallocAsync((void**)&off_tmp, (n + 1) * 8, st));
  NCD_TRY(ncd_exclusive_scan_u64(ctx, lenpad, off_tmp, n, (uint64_t*)d_tot));
  NCD_TRY(ncd_exclusive_scan_u32(ctx, own_cnt, a->d.own_start, n, (uint32_t*)(d_tot + 1)));
  unsigned long long h_tot[2] = {0, 0};
  NCD_CUDA(cudaMemcpyAsync(h_tot, d_tot, 16, cudaMemcpyDeviceToHost, st));
  NCD_CUDA(cudaStreamSynchronize(st));
  const uint64_t payload_bytes = h_tot[0];
  const uint64_t nowner = (uint32_t)h_tot[1];
  int rc = NCD_OK;
  if (payload_bytes > a->payload_cap || nowner > a->nowner_cap) {
    ncd_set_error("ncd_synth_fill: arena too small (%llu payload bytes, %llu owners needed)",
                  (unsigned long long)payload_bytes, (unsigned long long)nowner);
    rc = NCD_E_CAPACITY;
  }
  if (rc == NCD_OK && n) {
    cudaMemcpyAsync(a->d.rec_off, off_tmp, n * 8, cudaMemcpyDeviceToDevice, st);
    k_owners<<<blocks, 256, 0, st>>>(s, gsel, a->d, logical);
    k_payload<<<(unsigned)n, 128, 0, st>>>(s, gsel, a->d, logical);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
      ncd_set_error("ncd_synth_fill: launch failed: %s", cudaGetErrorString(e));
      rc = NCD_E_CUDA;
    }
  }
  cudaFreeAsync(off_tmp, st);
  cudaFreeAsync(lenpad, st);
  cudaFreeAsync(own_cnt, st);
  cudaFreeAsync(logical, st);
  cudaFreeAsync(d_tot, st);
  cudaFreeAsync(gsel, st);
  if (rc != NCD_OK) return rc;
  return ncd_arena_commit_device(a, n, payload_bytes, (n + 1) * SLOT, nowner, role < 0 ? nsel * s.nd : 0);
}
