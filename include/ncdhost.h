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

/* DriftChecker.Scan: bring every cluster's resident arena up to date with what was added /
 * removed since the last Scan (only changed records or the small per-object arrays cross PCIe,
 * a full upload happens on the first Scan or when an arena runs out of room), then ncd_scan.
 * Replaces recomputing the six lister.Get and four reflect.DeepEqual per handler call. */
int ncdh_scan(ncdh_checker* c);

/* DriftChecker.Lookup: the decision of the last Scan for (template or workgroup kind/ns/name,
 * dep = 0 the object itself | j+1 its j-th dependent, shard).  Returns 0 and *code, or 1 when
 * the object is not in the controller cache (controller.go:769-772). */
int ncdh_lookup(ncdh_checker* c, uint32_t kind, const char* ns, const char* name, int dep, int shard,
                uint32_t* code);

/* Ordered replay of one handler call from the last Scan: the shard writes templateSyncHandler
 * (controller.go:790-831, with :504-626) or workgroupSyncHandler (:720-745) would issue, in
 * order, stopping where the handler would return an error.  Output, one line per action:
 *   "<shard>\t<nexus|kube>\t<verb>\t<resource>\t<name>\n"   then   "error\t<text>\n" if it failed.
 * Returns the text length (written only if it fits cap) or -1. */
int64_t ncdh_plan_template(ncdh_checker* c, const char* ns, const char* name, char* out, uint64_t cap);
int64_t ncdh_plan_workgroup(ncdh_checker* c, const char* ns, const char* name, char* out, uint64_t cap);

/* The batch sweep (row f-3): after one Scan, the work-queue feed — only the templates and
 * workgroups with something to do, in controller-cache order.  Replaces re-enqueueing every
 * object on every resync (controller.go:288-290, 301-303; main.go:70-71) and letting the 50/s
 * limiter (controller.go:257-260) pace objects that did not drift.  Output, one line per item:
 *   "<template|workgroup>\t<namespace/name>\n"   (Element{objRef, objTypeName}, controller.go:88-91).
 * Returns the text length (written only if it fits cap) or -1. */
int64_t ncdh_sweep_queue(ncdh_checker* c, char* out, uint64_t cap);

/* How a cluster's data has reached the GPU so far: scans that needed a full arena upload
 * (ncd_arena_commit), a structural one (ncd_arena_recommit: per-object arrays + new records) or
 * only rewritten records (ncd_arena_sync_records).  The arenas stay resident between scans. */
int ncdh_upload_counts(ncdh_checker* c, int cluster, uint64_t* full, uint64_t* structural, uint64_t* values);

/* Number of diff-list entries and decisions of the last Scan. */
int ncdh_last_counts(ncdh_checker* c, uint64_t* ndiff, uint64_t* decisions);

#ifdef __cplusplus
}
#endif
#endif /* NCDHOST_H */
