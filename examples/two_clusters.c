/* two_clusters.c — the C ABI (include/ncdrift.h) used from plain C, no Python, no C++.
 *
 * A controller cache with one NexusAlgorithmTemplate that references one Secret and one ConfigMap,
 * and two shard caches: shard 0 holds everything but its Secret has drifted and its ConfigMap is
 * owned by nobody; shard 1 holds nothing.  The program packs the three caches into arenas the way
 * a packer would, scans, and prints the decisions — what the handlers
 * of controller.go:504-626, 790-831 would do for these caches.
 *
 *   gcc -std=c99 -Iinclude examples/two_clusters.c -Lnexus_configuration_controller_b200/lib \
 *       -lncdrift -Wl,-rpath,$PWD/nexus_configuration_controller_b200/lib -o two_clusters
 *
 * Needs a GPU to run (the library has no CPU path); exit code 0 = the decisions are the expected ones.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "ncdrift.h"

#define MAX_OWNERS 2
#define MAX_DEPS 2

typedef struct {
  uint32_t kind;
  const char* key;   /* "namespace/name" */
  const char* value; /* canonical record (here: any bytes; equal bytes <=> DeepEqual) */
  const char* uid;
  int nowners;
  struct { int is_template; const char* uid; } owners[MAX_OWNERS];
  int ndeps;
  struct { uint32_t kind; const char* key; } deps[MAX_DEPS];
} object;

static void die(const char* what) {
  fprintf(stderr, "%s: %s\n", what, ncd_last_error());
  exit(2);
}

static uint32_t put(ncd_arena_view* v, uint64_t* used, const char* s, uint32_t* len) {
  const uint32_t off = (uint32_t)*used;
  *len = (uint32_t)strlen(s);
  memcpy(v->aux + off, s, *len);
  *used += *len;
  return off;
}

/* Packs objects into a fresh arena and commits it: records 16-byte aligned and zero padded,
 * keys / uids in the aux pool, ownerReferences and dependents as CSR lists. */
static ncd_arena* pack(ncd_ctx* ctx, const object* objs, int n) {
  uint64_t payload = 0, aux = 0, nowner = 0, ndep = 0;
  for (int i = 0; i < n; ++i) {
    payload += (strlen(objs[i].value) + 15) / 16 * 16;
    aux += strlen(objs[i].key) + strlen(objs[i].uid);
    for (int k = 0; k < objs[i].nowners; ++k) aux += strlen(objs[i].owners[k].uid);
    for (int k = 0; k < objs[i].ndeps; ++k) aux += strlen(objs[i].deps[k].key);
    nowner += (uint64_t)objs[i].nowners;
    ndep += (uint64_t)objs[i].ndeps;
  }
  ncd_arena* a = NULL;
  if (ncd_arena_create(ctx, payload + 16, (uint64_t)n + 1, aux + 16, nowner + 1, ndep + 1, NCD_ARENA_HOST_PINNED, &a) != NCD_OK)
    die("ncd_arena_create");
  ncd_arena_view v;
  if (ncd_arena_host_view(a, &v) != NCD_OK) die("ncd_arena_host_view");
  uint64_t off = 0, aux_used = 0;
  uint32_t oi = 0, di = 0;
  for (int i = 0; i < n; ++i) {
    const size_t len = strlen(objs[i].value), padded = (len + 15) / 16 * 16;
    v.rec_off[i] = off;
    v.rec_len[i] = (uint32_t)len;
    v.kind[i] = (uint8_t)objs[i].kind;
    memcpy(v.payload + off, objs[i].value, len);
    memset(v.payload + off + len, 0, padded - len);
    off += padded;
    v.key_off[i] = put(&v, &aux_used, objs[i].key, &v.key_len[i]);
    v.uid_off[i] = put(&v, &aux_used, objs[i].uid, &v.uid_len[i]);
    v.own_start[i] = oi;
    for (int k = 0; k < objs[i].nowners; ++k, ++oi) {
      v.own_is_template[oi] = (uint8_t)objs[i].owners[k].is_template;
      v.own_uid_off[oi] = put(&v, &aux_used, objs[i].owners[k].uid, &v.own_uid_len[oi]);
    }
    v.dep_start[i] = di;
    for (int k = 0; k < objs[i].ndeps; ++k, ++di) {
      v.dep_kind[di] = (uint8_t)objs[i].deps[k].kind;
      v.dep_key_off[di] = put(&v, &aux_used, objs[i].deps[k].key, &v.dep_key_len[di]);
    }
  }
  v.own_start[n] = oi;
  v.dep_start[n] = di;
  if (ncd_arena_commit(a, (uint64_t)n, off, aux_used, oi, di) != NCD_OK) die("ncd_arena_commit");
  return a;
}

static const char* code_name(uint32_t c) {
  switch (c) {
    case NCD_CREATE: return "CREATE";
    case NCD_UPDATE_DATA: return "UPDATE_DATA";
    case NCD_UPDATE_OWNER: return "UPDATE_OWNER";
    case NCD_UPDATE_DATA | NCD_UPDATE_OWNER: return "UPDATE_DATA|UPDATE_OWNER";
    case NCD_ERR_ROGUE: return "ERR_ROGUE";
    case NCD_ERR_MISSING_REF: return "ERR_MISSING_REF";
    default: return "?";
  }
}

int main(void) {
  ncd_ctx* ctx = NULL;
  if (ncd_init(0, &ctx) != NCD_OK) die("ncd_init");

  const object controller[] = {
      {NCD_KIND_TEMPLATE, "nexus/algo", "spec: image=algo:1.2 cpu=2", "uid-ctrl-algo", 0, {{0, ""}}, 2,
       {{NCD_KIND_SECRET, "nexus/algo-secret"}, {NCD_KIND_CONFIGMAP, "nexus/algo-config"}}},
      {NCD_KIND_SECRET, "nexus/algo-secret", "token=alpha", "uid-ctrl-secret", 0, {{0, ""}}, 0, {{0, ""}}},
      {NCD_KIND_CONFIGMAP, "nexus/algo-config", "mode=fast", "uid-ctrl-config", 0, {{0, ""}}, 0, {{0, ""}}},
  };
  /* shard 0: template equal; secret drifted but owned by the shard's template; configmap equal but
   * without any ownerReference (the reference refuses to touch it: controller.go:494-499) */
  const object shard0[] = {
      {NCD_KIND_CONFIGMAP, "nexus/algo-config", "mode=fast", "uid-s0-config", 0, {{0, ""}}, 0, {{0, ""}}},
      {NCD_KIND_TEMPLATE, "nexus/algo", "spec: image=algo:1.2 cpu=2", "uid-s0-algo", 0, {{0, ""}}, 0, {{0, ""}}},
      {NCD_KIND_SECRET, "nexus/algo-secret", "token=bravo", "uid-s0-secret", 1, {{1, "uid-s0-algo"}}, 0, {{0, ""}}},
  };
  ncd_arena* ca = pack(ctx, controller, 3);
  ncd_arena* s0 = pack(ctx, shard0, 3);
  ncd_arena* s1 = pack(ctx, shard0, 0); /* an empty cluster: everything is to be created */
  const ncd_arena* shards[2];
  shards[0] = s0;
  shards[1] = s1;

  ncd_diff* d = NULL;
  if (ncd_scan(ctx, ca, shards, 2, NCD_MODE_DIRECT, &d) != NCD_OK) die("ncd_scan");
  const ncd_entry* e = NULL;
  uint64_t n = 0;
  if (ncd_diff_view(d, &e, &n) != NCD_OK) die("ncd_diff_view");
  static const char* what[] = {"template nexus/algo", "secret nexus/algo-secret", "configmap nexus/algo-config"};
  for (uint64_t i = 0; i < n; ++i)
    printf("shard %u: %-28s %s\n", e[i].shard, what[e[i].dep], code_name(e[i].code));

  /* rows are (template, its secret, its configmap); shard 0 first within each row */
  const uint32_t want[][3] = {/* {dep, shard, code} */
                              {0, 1, NCD_CREATE},      {1, 0, NCD_UPDATE_DATA}, {1, 1, NCD_CREATE},
                              {2, 0, NCD_ERR_ROGUE},   {2, 1, NCD_CREATE}};
  int ok = n == 5;
  for (uint64_t i = 0; ok && i < n; ++i)
    ok = e[i].obj == 0 && e[i].dep == want[i][0] && e[i].shard == want[i][1] && e[i].code == want[i][2];
  ncd_scan_stats st;
  if (ncd_scan_last_stats(ctx, &st) != NCD_OK) die("ncd_scan_last_stats");
  printf("%llu decisions, %llu entries, %u kernels, %.3f ms on the GPU: %s\n", (unsigned long long)st.pairs,
         (unsigned long long)st.ndiff, st.launches, st.total_ms, ok ? "as expected" : "UNEXPECTED");

  ncd_diff_destroy(d);
  ncd_arena_destroy(s1);
  ncd_arena_destroy(s0);
  ncd_arena_destroy(ca);
  ncd_destroy(ctx);
  return ok ? 0 : 1;
}
