/*
 * ncdhost.h — C ABI of libncdhost.so: the host side of the drift check, above libncdrift.so.
 *
 * In a deployment this layer is Go: the canonicaliser / packer that turns
 * lister contents into arenas, the DriftChecker the patched handlers consult, and the ordered
 * replay of a handler's shard loop.  This image has no Go toolchain, so the layer is C++
 * (the reference is compiled code) with this C surface for the test harness.  It computes no
 * drift itself: every decision comes from ncd_scan on the GPU.
 */
#ifndef NCDHOST_H
#define NCDHOST_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ncdh_checker ncdh_checker;

/* Text of the last error on this thread. */
const char* ncdh_last_error(void);

/* Canonical encoding of one Go value given in the transport form of tests/govalues.py:
 * the bytes the packer appends to an arena for a .Spec / .Data value.  Byte equality of two
 * encodings <=> reflect.DeepEqual of the values (controller.go:539,600,725,795).  Returns the
 * encoded length (written only if it fits cap), or -1 on malformed input / NaN. */
int64_t ncdh_canon(const uint8_t* value, uint64_t value_len, uint8_t* out, uint64_t cap);

/* The DriftChecker of INTEGRATION.md section 2.  device = CUDA ordinal, mode = NCD_MODE_*,
 * nshards = len(c.nexusShards) (controller.go:266). */
int ncdh_checker_create(int device, uint32_t mode, int nshards, ncdh_checker** out);
void ncdh_checker_destroy(ncdh_checker* c);

/* Puts one object into a cluster's cache image — what Informer().GetIndexer().Add(obj) does in
 * the reference's fixture (controller_test.go:546-576).  cluster = -1 for the controller
 * cluster, s >= 0 for shard s.  kind = NCD_KIND_*.  value = the compared value (.Spec or .Data)
 * in transport form.  owner_kinds/owner_uids: metadata.ownerReferences.  secret_names /
 * configmap_names: GetSecretNames() / GetConfigMapNames() of a template (controller.go:505,567).
 * Adding the same (kind, namespace, name) again replaces the object, as a store would. */
int ncdh_cluster_add(ncdh_checker* c, int cluster, uint32_t kind, const char* ns, const char* name,
                     const char* uid, const uint8_t* value, uint64_t value_len,
                     const char* const* owner_kinds, const char* const* owner_uids, int nowners,
                     const char* const* secret_names, int nsecrets,
                     const char* const* configmap_names, int nconfigmaps);
int ncdh_cluster_remove(ncdh_checker* c, int cluster, uint32_t kind, const char* ns, const char* name);

/* The same as a batch: informer events that arrived since the last call (AddFunc / UpdateFunc /
 * DeleteFunc of the four informers, controller.go:286-359), applied in order.  The values are
 * canonicalised on `threads` threads (<= 0: one per host core), then the cache images are updated
 * under the checker's lock.  An event that fails (malformed value, unknown cluster) is skipped:
 * every other event of the batch is applied, the call returns -1 and ncdh_last_error names the
 * first skipped event.  A new value that fits the place its record already has in the arena's
 * pinned payload is written there directly (unchanged values — a resync — write nothing) and, in a
 * large batch, pulled to the device while the rest of the batch is still canonicalised; others
 * wait in pooled memory until the next ncdh_scan lays them out. */
enum { NCDH_EV_PUT = 0, NCDH_EV_DELETE = 1 };
typedef struct ncdh_event {
  int32_t cluster;  /* -1 controller, s >= 0 shard s                         */
  uint32_t kind;    /* NCD_KIND_*                                            */
  uint32_t op;      /* NCDH_EV_PUT | NCDH_EV_DELETE                          */
  int32_t nowners, nsecrets, nconfigmaps;
  const char* ns;
  const char* name;
  const char* uid;
  const uint8_t* value; /* transport form of .Spec / .Data (PUT only)        */
  uint64_t value_len;
  const char* const* owner_kinds;
  const char* const* owner_uids;
  const char* const* secret_names;
  const char* const* configmap_names;
} ncdh_event;
int ncdh_apply_events(ncdh_checker* c, const ncdh_event* events, uint64_t n, int threads);

/* Threading and freshness (the reference's handlers run on `workers` goroutines at once,
 * controller.go:875-877, and read LIVE caches, :766,792,507,516):
 *   - every function here may be called from any thread.  ncdh_cluster_add / _remove /
 *     ncdh_apply_events take the checker's cache lock for the update only; ncdh_scan runs one at a
 *     time and holds that lock only while it lays changed records out in pinned memory, not while
 *     the GPU works; ncdh_lookup / ncdh_plan_* / ncdh_sweep_queue read the snapshot the last
 *     finished scan published (swapped in atomically) and never wait for a scan in flight.
 *   - a snapshot is only as fresh as its scan.  Every put / remove is stamped with a counter; a
 *     lookup or plan whose template, any of its dependents, or any shard's copy of those was put or
 *     removed AFTER the snapshot's scan began answers NCDH_STALE instead of a decision the live
 *     caches no longer support (a Secret rotated one second after a scan is never answered SKIP).
 *     The caller then runs ncdh_scan (cheap: only the changed records cross PCIe) and asks again,
 *     or falls back to its own lister.Get + reflect.DeepEqual for that object. */
#define NCDH_NOT_FOUND 1
#define NCDH_STALE 2

/* DriftChecker.Scan: bring every cluster's resident arena up to date with what was added /
 * removed since the last Scan (only changed records or the small per-object arrays cross PCIe,
 * a full upload happens on the first Scan or when an arena runs out of room), then ncd_scan.
 * Replaces recomputing the six lister.Get and four reflect.DeepEqual per handler call. */
int ncdh_scan(ncdh_checker* c);

/* DriftChecker.Lookup: the decision of the last Scan for (template or workgroup kind/ns/name,
 * dep = 0 the object itself | j+1 its j-th dependent, shard).  Returns 0 and *code,
 * NCDH_NOT_FOUND when the object is not in the controller cache (controller.go:769-772),
 * NCDH_STALE when the caches changed under the snapshot (see above), -1 on a bad argument. */
int ncdh_lookup(ncdh_checker* c, uint32_t kind, const char* ns, const char* name, int dep, int shard,
                uint32_t* code);

/* Ordered replay of one handler call from the last Scan: the shard writes templateSyncHandler
 * (controller.go:790-831, with :504-626) or workgroupSyncHandler (:720-745) would issue, in
 * order, stopping where the handler would return an error.  Output, one line per action:
 *   "<shard>\t<nexus|kube>\t<verb>\t<resource>\t<name>\n"   then   "error\t<text>\n" if it failed.
 * Returns the text length (written only if it fits cap), -1 on error, or -NCDH_STALE (-2) when
 * anything the replay reads was put or removed after the snapshot's scan began. */
int64_t ncdh_plan_template(ncdh_checker* c, const char* ns, const char* name, char* out, uint64_t cap);
int64_t ncdh_plan_workgroup(ncdh_checker* c, const char* ns, const char* name, char* out, uint64_t cap);

/* The batch sweep (row f-3): after one Scan, the work-queue feed — only the templates and
 * workgroups with something to do, in controller-cache order.  Replaces re-enqueueing every
 * object on every resync (controller.go:288-290, 301-303; main.go:70-71) and letting the 50/s
 * limiter (controller.go:257-260) pace objects that did not drift.  Output, one line per item:
 *   "<template|workgroup>\t<namespace/name>\n"   (Element{objRef, objTypeName}, controller.go:88-91).
 * Returns the text length (written only if it fits cap) or -1. */
int64_t ncdh_sweep_queue(ncdh_checker* c, char* out, uint64_t cap);

/* adoptReferences (controller.go:647-695) as a plan: the controller-cluster Secrets / ConfigMaps
 * the template references that do not carry an ownerReference with its UID yet (decided on the GPU
 * by ncd_find_unadopted at the last Scan).  One line per Update the handler would issue on the
 * CONTROLLER cluster: "kube\tupdate-owner\t<secrets|configmaps>\t<name>\n".  Same return values as
 * ncdh_plan_template.  Templates with such references are part of ncdh_sweep_queue. */
int64_t ncdh_plan_adoption(ncdh_checker* c, const char* ns, const char* name, char* out, uint64_t cap);

/* Where the time and the bytes of the last Scan went (host clock, milliseconds). */
typedef struct ncdh_scan_timing {
  double canon_ms;    /* canonicalisation done by ncdh_cluster_add / ncdh_apply_events since the previous Scan */
  double pack_ms;     /* laying changed records out in the pinned arenas (under the cache lock)                */
  double upload_ms;   /* ncd_arena_commit / _recommit / ncd_arenas_sync_records                                */
  double scan_ms;     /* ncd_scan_host: kernels + diff list to the host                                        */
  double publish_ms;  /* building and swapping in the snapshot                                                 */
  double total_ms;    /* the whole ncdh_scan call                                                              */
  double gpu_scan_ms; /* device time of the scan kernels (CUDA events)                                         */
  uint64_t h2d_bytes; /* bytes this Scan moved host -> device                                                  */
  uint64_t d2h_bytes; /* bytes it read back                                                                    */
  uint64_t records_synced, full_uploads, structural_uploads;
  uint64_t events;    /* puts / removes that changed something since the previous Scan                         */
  uint64_t compactions; /* arenas compacted in place so far (ncd_arena_compact: dead payload reclaimed without PCIe) */
  uint64_t records_in_place; /* of records_synced: written into the pinned payload by their batch, not copied by the packer */
} ncdh_scan_timing;
int ncdh_last_timing(ncdh_checker* c, ncdh_scan_timing* out);

/* How a cluster's data has reached the GPU so far: scans that needed a full arena upload
 * (ncd_arena_commit), a structural one (ncd_arena_recommit: per-object arrays + new records) or
 * only rewritten records (ncd_arena_sync_records).  The arenas stay resident between scans. */
int ncdh_upload_counts(ncdh_checker* c, int cluster, uint64_t* full, uint64_t* structural, uint64_t* values);

/* Number of diff-list entries and decisions of the last Scan. */
int ncdh_last_counts(ncdh_checker* c, uint64_t* ndiff, uint64_t* decisions);

/* A checksum of the last Scan's diff list that does not depend on arena layout: the sum over the
 * entries of mix(fnv1a(kind digit + "namespace/name") ^ dep * 0x9E3779B185EBCA87 ^ shard *
 * 0xC2B2AE3D27D4EB4F ^ code * 0x165667B19E3779F9) (mix = the MurmurHash3 64-bit finaliser).  Two
 * implementations fed the same events can be compared by it (bench.py compares the GPU arm with the
 * CPU arm's sweep over parsed values).  dirty_objects (optional) = objects with at least one entry. */
int ncdh_diff_digest(ncdh_checker* c, uint64_t* digest, uint64_t* dirty_objects);

/* ------------------------------------------------------------------ synthetic events
 * Bench / test utility, not part of the drop-in surface (like ncd_synth_fill): informer events made
 * from objects laid out like an arena (struct ncd_arena_view of ncdrift.h with HOST pointers, e.g.
 * the synthetic generator's output downloaded with ncd_arena_download).  Names, UIDs,
 * ownerReferences and dependent names come from `meta`; the value of object i is record i of
 * `values` (NULL = meta; another cluster's arrays model "the shard received the controller's
 * value") wrapped as a Go value in transport form: map[string][]byte{"data": rec} for a Secret,
 * map[string]string{"data": rec} for a ConfigMap, struct{[]byte; string} for a template or
 * workgroup spec.  salt != 0 flips one byte of every value at a position derived from (salt,
 * kind, key): the same salt gives the same bytes for every cluster's copy of an object.
 * obj_idx = NULL takes every object of meta.  The batch owns everything its events point to. */
struct ncd_arena_view;
typedef struct ncdh_event_batch ncdh_event_batch;
int ncdh_batch_from_arrays(const struct ncd_arena_view* meta, const struct ncd_arena_view* values, int cluster,
                           const uint32_t* obj_idx, uint64_t n, uint64_t salt, int threads, ncdh_event_batch** out);
const ncdh_event* ncdh_batch_events(const ncdh_event_batch* b, uint64_t* n);
uint64_t ncdh_batch_value_bytes(const ncdh_event_batch* b);
void ncdh_batch_free(ncdh_event_batch* b);

#ifdef __cplusplus
}
#endif
#endif /* NCDHOST_H */
