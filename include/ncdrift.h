/*
 * ncdrift.h — C ABI of libncdrift.so, the B200 (sm_100a) drift checker.
 *
 * One hot path of SneaksAndData/nexus-configuration-controller lives behind this
 * header: the reconcile drift check.  In the reference it is not factored out; it
 * is six inlined lister.Get calls and four inlined reflect.DeepEqual calls inside
 * the sync handlers.  Each entry point below names the reference lines it stands
 * in for (paths relative to the reference checkout).
 *
 * Conventions
 *   - plain pointers and sizes only; every handle is opaque and owned by the library
 *   - every call returns NCD_OK (0) or a negative ncd_status; the text of the last
 *     failure on the calling thread is in ncd_last_error()
 *   - there is no CPU fallback: with no usable CUDA device ncd_init fails
 *   - one ncd_ctx per GPU per process; calls on one ctx must be serialised by the
 *     caller (the Go side holds a mutex)
 *
 * The reference binding for every function here (cgo) is shown in INTEGRATION.md.
 */
#ifndef NCDRIFT_H
#define NCDRIFT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ status */
typedef enum ncd_status {
  NCD_OK = 0,
  NCD_E_INVALID = -1,   /* bad argument / inconsistent arena contents            */
  NCD_E_CUDA = -2,      /* a CUDA runtime call failed (message has the detail)   */
  NCD_E_NOMEM = -3,     /* host or device allocation failed                      */
  NCD_E_CAPACITY = -4,  /* contents exceed the capacity the arena was made with  */
  NCD_E_STATE = -5      /* call made in the wrong state (e.g. scan before commit)*/
} ncd_status;

/* Object kinds held in an arena (the four resource types the sync handlers walk,
 * controller.go:697-845). */
enum {
  NCD_KIND_TEMPLATE = 0,  /* NexusAlgorithmTemplate : compared value = .Spec  (controller.go:795) */
  NCD_KIND_WORKGROUP = 1, /* NexusAlgorithmWorkgroup: compared value = .Spec  (controller.go:725) */
  NCD_KIND_SECRET = 2,    /* corev1.Secret          : compared value = .Data  (controller.go:539) */
  NCD_KIND_CONFIGMAP = 3  /* corev1.ConfigMap       : compared value = .Data  (controller.go:600) */
};

/* Decision codes, one per (row, shard).  They are the branches of the reference's
 * sync handlers; the diff list is the codes != NCD_SKIP in (row, shard) order. */
enum {
  NCD_SKIP = 0,            /* present, DeepEqual, ownership fine: fall-through of :539/:600/:725/:795 */
  NCD_CREATE = 1,          /* NotFound in the shard cache: :518-520 :580-582 :735-738 :805-808        */
  NCD_UPDATE_DATA = 2,     /* present, not DeepEqual: :539-541 :600-602 :725-727 :795-797             */
  NCD_UPDATE_OWNER = 4,    /* dependent has ownerRefs, none is {NexusAlgorithmTemplate, UID of the
                              shard template}: :489-501 -> :550-552 :612-614.  May be OR-ed with
                              NCD_UPDATE_DATA (the reference then writes data first, owner second)   */
  NCD_ERR_ROGUE = 8,       /* dependent on the shard has zero ownerRefs: :494-499                     */
  NCD_ERR_MISSING_REF = 16 /* dependent name absent from the CONTROLLER cache: :509-514 :571-576
                              (and earlier :649-653 :672-676)                                         */
};

/* Scan modes. */
enum {
  NCD_MODE_DIRECT = 0, /* byte compare of the canonical records: exact (default)                  */
  NCD_MODE_HASH64 = 1, /* striped 64-bit hash of every record, verdict from hash equality          */
  NCD_MODE_SHA256 = 2, /* FIPS 180-4 SHA-256 of every record, verdict from digest equality         */
  NCD_MODE_XXH64 = 3,  /* standard XXH64 (seed 0) of every record, verdict from digest equality    */
  NCD_MODE_CONFIRM = 0x100, /* OR with a hash mode: hash-equal pairs are re-checked by byte compare */
  NCD_MODE_REHASH = 0x200,  /* OR with a hash mode: recompute every digest (by default digests are
                               kept with a committed arena and reused by later scans; a value sync
                               re-hashes only the records it rewrote)                              */
  NCD_MODE_CACHED = 0x400   /* ncd_hash_arena only: return the digests the arena holds when they
                               are valid instead of recomputing them                               */
};

typedef struct ncd_ctx ncd_ctx;
typedef struct ncd_arena ncd_arena;
typedef struct ncd_diff ncd_diff;
typedef struct ncd_xchg ncd_xchg;

/* ------------------------------------------------------------------ arenas
 * An arena is one cluster's informer cache in canonical form: the controller
 * cluster's, or one shard's.  It replaces what the listers hold in the reference
 * (controller.go:268-278 and shards.Shard's four listers) for the purposes of the
 * drift check only.
 *
 * Layout (structure of arrays, all little-endian):
 *   payload   canonical bytes of the compared value of every object, each record
 *             starting on a 16-byte boundary, zero padded to the next boundary
 *   rec_off   byte offset of record i in payload (multiple of 16)
 *   rec_len   canonical length of record i in bytes (without padding)
 *   kind      NCD_KIND_*
 *   key       "namespace/name" (cache.ObjectName.String()) of object i, in aux
 *   uid       metadata.uid of object i, in aux (used for templates only)
 *   own_*     ownerReferences of object i are entries own_start[i]..own_start[i+1]:
 *             own_is_template = (Kind == "NexusAlgorithmTemplate"), own_uid in aux
 *   dep_*     for templates: the dependents named by GetSecretNames() followed by
 *             GetConfigMapNames() (controller.go:505,567), entries
 *             dep_start[i]..dep_start[i+1]; dep key = "namespace/name" in aux
 *   aux       byte pool for keys and uids
 */
typedef struct ncd_arena_view {
  uint64_t nobj;           /* objects                        */
  uint64_t payload_bytes;  /* used bytes of payload          */
  uint64_t aux_bytes;      /* used bytes of aux              */
  uint64_t nowner;         /* owner reference entries        */
  uint64_t ndep;           /* dependent reference entries    */
  uint8_t* payload;
  uint64_t* rec_off;
  uint32_t* rec_len;
  uint8_t* kind;
  uint32_t* key_off;
  uint32_t* key_len;
  uint32_t* uid_off;
  uint32_t* uid_len;
  uint32_t* own_start;     /* nobj + 1 entries */
  uint8_t* own_is_template;
  uint32_t* own_uid_off;
  uint32_t* own_uid_len;
  uint32_t* dep_start;     /* nobj + 1 entries */
  uint8_t* dep_kind;
  uint32_t* dep_key_off;
  uint32_t* dep_key_len;
  uint8_t* aux;
} ncd_arena_view;

enum {
  NCD_ARENA_HOST_PINNED = 0, /* pinned host arrays the caller fills + a device mirror */
  NCD_ARENA_DEVICE_ONLY = 1  /* device arrays only (filled by ncd_synth_fill)         */
};

/* One diff-list entry: what the host-side replay needs to issue the shard write. */
typedef struct ncd_entry {
  uint32_t obj;   /* controller index of the template / workgroup the row belongs to */
  uint32_t dep;   /* 0 = the object itself, j+1 = its j-th dependent reference        */
  uint32_t shard; /* index into the shard list passed to ncd_scan                     */
  uint32_t code;  /* NCD_* decision code (never NCD_SKIP in the list)                 */
} ncd_entry;

/* Device-side timings of the last scan (CUDA events on the ctx stream), ms. */
typedef struct ncd_scan_stats {
  float join_ms;      /* cache join (hash-table probes)              */
  float hash_ms;      /* hashing kernels (hash modes only)           */
  float compare_ms;   /* byte-compare kernel (direct / confirm)      */
  float classify_ms;  /* decision codes + ordered compaction         */
  float total_ms;     /* first launch to last kernel end             */
  uint64_t rows;      /* rows of the controller arena                */
  uint64_t pairs;     /* rows x shards = decisions made              */
  uint64_t algo_bytes;/* canonical bytes of every compared copy, each counted once */
  uint64_t ndiff;     /* entries in the diff list                    */
  uint64_t hash_collisions; /* CONFIRM only: hash-equal pairs whose bytes differ */
  uint32_t launches;  /* kernels launched by the scan                */
  uint32_t compare_variant; /* which byte-compare kernel ran: 0 k_compare, 2 k_compare_grouped, 8 k_compare_tma, 16 k_compare_small,
                               32 k_scan_small (join and compare fused, small records in direct mode) */
} ncd_scan_stats;

/* ------------------------------------------------------------------ context */
/* Library version string. */
const char* ncd_version(void);
/* Text of the last error on this thread ("" if none). */
const char* ncd_last_error(void);

/* Creates the context for one CUDA device (stream, scratch, events).  Replaces
 * nothing in the reference — it is the set-up NewController (controller.go:227-243)
 * would do once the controller is given a DriftChecker. */
int ncd_init(int device, ncd_ctx** out);
void ncd_destroy(ncd_ctx* ctx);
/* The CUDA stream (cudaStream_t) every launch of this ctx goes to, for callers that
 * time with their own events. */
void* ncd_ctx_stream(ncd_ctx* ctx);
int ncd_ctx_device(ncd_ctx* ctx);
int ncd_ctx_sync(ncd_ctx* ctx);

/* 64-bit hash of ("kind", "namespace/name") used by the join and by the multi-GPU
 * partition (gpu = ncd_keyhash % n_gpus).  Same function on host and device. */
uint64_t ncd_keyhash(uint32_t kind, const void* key, uint32_t key_len);

/* ------------------------------------------------------------------ arena API */
/* Allocates an arena with the given capacities.  Stands in for the informer store
 * behind a lister (controller.go:268-278; shards.Shard.*Lister). */
int ncd_arena_create(ncd_ctx* ctx, uint64_t payload_cap, uint64_t nobj_cap, uint64_t aux_cap,
                     uint64_t nowner_cap, uint64_t ndep_cap, uint32_t flags, ncd_arena** out);
void ncd_arena_destroy(ncd_arena* a);
/* Host pointers (pinned) the packer writes into; counts in the view are capacities.
 * NCD_ARENA_HOST_PINNED arenas only. */
int ncd_arena_host_view(ncd_arena* a, ncd_arena_view* out);
/* Uploads the first nobj objects and builds the arena's index on the GPU: key
 * hashes, the open-addressing table that answers lister.Get (controller.go:507,516,
 * 569,578,722,792), and for a controller arena the row and tile lists.  The arena
 * is immutable afterwards until the next commit. */
int ncd_arena_commit(ncd_arena* a, uint64_t nobj, uint64_t payload_bytes, uint64_t aux_bytes,
                     uint64_t nowner, uint64_t ndep);
/* The same for several arenas of one context in one call (the controller cluster and every shard
 * before a sweep): all uploads are enqueued up front on a copy stream and the index of arena k is
 * built while arena k+1 is still crossing PCIe, instead of copy and index build taking turns.
 * Arrays of `count` entries; an arena may appear once.  On an error the arenas before it are
 * committed, it and the ones after it are not. */
int ncd_arenas_commit(int count, ncd_arena* const* arenas, const uint64_t* nobj, const uint64_t* payload_bytes,
                      const uint64_t* aux_bytes, const uint64_t* nowner, const uint64_t* ndep);
/* Incremental update of a committed NCD_ARENA_HOST_PINNED arena: the records of the listed
 * objects were rewritten in the pinned host arrays (payload bytes at a new or the same rec_off,
 * rec_len; a record that grew is appended behind the old payload end) and the device copy must
 * follow.  Only those records cross PCIe (a kernel pulls them from the pinned arrays), then the
 * tile lists are rebuilt; keys, ownerReferences and dependents are NOT re-read — an object that
 * is added, deleted or re-owned needs ncd_arena_commit.  This is what an informer UpdateFunc
 * (controller.go:288-290, 301-303, 321-330, 344-353) turns into, instead of a full re-upload
 * on every 30 s resync (main.go:70-71).  obj_idx: host array of n distinct object indices;
 * new_payload_bytes: the used payload size after the rewrite (<= capacity).  Digests the arena
 * holds (from a hash-mode scan or ncd_hash_arena) are brought up to date for the listed records
 * and stay valid, so the next hash-mode scan hashes nothing. */
int ncd_arena_sync_records(ncd_arena* a, uint64_t n, const uint32_t* obj_idx, uint64_t new_payload_bytes);
/* The same for several arenas of one context in one call (a sweep syncs the controller and every
 * shard cluster before it scans): every pull is enqueued first, then one host synchronisation
 * reads all results back — one round trip instead of one per arena.  Arrays of `count` entries;
 * an arena may appear once.  On a refused record the call still finishes the other arenas and
 * returns the first error. */
int ncd_arenas_sync_records(int count, ncd_arena* const* arenas, const uint64_t* n,
                            const uint32_t* const* obj_idx, const uint64_t* new_payload_bytes);
/* Commit after a structural change (objects added, deleted or re-owned: informer AddFunc /
 * DeleteFunc, controller.go:286-305, 319-355) WITHOUT re-uploading the payload that did not
 * change.  The packer has rewritten the small per-object arrays in the pinned host arena (any
 * object order; a surviving object keeps its rec_off, so its bytes stay where they are on the
 * device; a deleted object's bytes simply become unreferenced) and appended the records of new
 * objects behind payload_dirty_from.  This call uploads every array except payload, uploads
 * payload[payload_dirty_from, payload_bytes) only, and rebuilds the index — per-object metadata
 * is ~40 bytes, so a 1M-object arena re-commits in tens of MB instead of its 8.7 GB. */
int ncd_arena_recommit(ncd_arena* a, uint64_t nobj, uint64_t payload_bytes, uint64_t aux_bytes,
                       uint64_t nowner, uint64_t ndep, uint64_t payload_dirty_from);
/* Reclaims payload no record references any more (objects that were deleted, the old places of
 * records that grew and were appended) WITHOUT a re-upload: the live records move to the front of
 * the payload, on the device (a gather through scratch memory) and in the pinned host arrays (in
 * place), and rec_off changes identically on both sides; keys, digests and the index stay as they
 * are.  A long-lived controller's arena (informer events for days: controller.go:286-359) no longer
 * needs a full ncd_arena_commit over PCIe to stop growing.  *new_payload_bytes = used payload
 * afterwards; the caller re-reads rec_off from the host view.  A record rewritten in the host arrays
 * and not synced yet may have a new LENGTH there (the slots are laid out for the host's rec_len);
 * its rec_off must still be the one the device holds.  No reference counterpart. */
int ncd_arena_compact(ncd_arena* a, uint64_t* new_payload_bytes);
/* Copies the committed device arrays into caller buffers (sized from
 * ncd_arena_counts); any pointer in dst may be NULL to skip that array. */
int ncd_arena_download(ncd_arena* a, ncd_arena_view* dst);
int ncd_arena_counts(ncd_arena* a, ncd_arena_view* counts_only);
/* Reads a few records of a committed arena back: record obj_idx[k] lands at out + out_off[k]
 * (16-byte padded; out_off has n + 1 entries, the last one = bytes written), lengths are the
 * arena's rec_len.  A device-side gather and one copy, so sampling a 100 GB arena costs the sample.
 * Fails with NCD_E_CAPACITY (nothing written) when out_cap is too small; out_off is filled in
 * either way, so a caller can size the buffer from a first call with out_cap = 0.  No reference
 * counterpart (used by the full-size parity tests and offline tools). */
int ncd_arena_fetch_records(ncd_arena* a, uint64_t n, const uint32_t* obj_idx, uint8_t* out, uint64_t out_cap,
                            uint64_t* out_off);
/* Snapshot of a committed arena as a file, and back (offline
 * diffing — compare what a cluster's caches held at some moment against another snapshot or a
 * live arena).  The file carries every array of ncd_arena_view plus a checksum; it streams
 * through a bounded pinned buffer.  ncd_arena_load creates the arena (flags as ncd_arena_create;
 * payload_slack = spare payload capacity for later ncd_arena_sync_records on a HOST_PINNED one),
 * verifies the checksum, validates the contents on the GPU like a commit does and builds the
 * index.  No reference counterpart (the reference keeps no record of its caches). */
int ncd_arena_save(ncd_arena* a, const char* path);
int ncd_arena_load(ncd_ctx* ctx, const char* path, uint32_t flags, uint64_t payload_slack, ncd_arena** out);
/* Rows of a committed arena when used as the controller side: one per template,
 * one per dependent reference of a template, one per workgroup. */
int ncd_arena_rows(ncd_arena* a, uint64_t* rows);

/* ------------------------------------------------------------------ scan */
/* The drift check: for every row of the controller arena and every shard, the
 * decision the reference's handlers would take (controller.go:504-626, 720-745,
 * 790-831), as a compacted, ordered diff list.  mode = NCD_MODE_*. */
int ncd_scan(ncd_ctx* ctx, const ncd_arena* ctrl, const ncd_arena* const* shards, int nshards,
             uint32_t mode, ncd_diff** out);
/* Same, split so that a caller can time the GPU work with its own events:
 * ncd_scan_launch only enqueues on the ctx stream; ncd_scan_finish synchronises
 * and publishes the diff list. */
int ncd_scan_launch(ncd_ctx* ctx, const ncd_arena* ctrl, const ncd_arena* const* shards,
                    int nshards, uint32_t mode);
int ncd_scan_finish(ncd_ctx* ctx, ncd_diff** out);
int ncd_scan_last_stats(ncd_ctx* ctx, ncd_scan_stats* out);
/* Stats of the scan enqueued `back` scans before the latest one (0 = latest; the last 64 are
 * kept).  The scan must have completed on the GPU (synchronise first). */
int ncd_scan_stats_back(ncd_ctx* ctx, uint64_t back, ncd_scan_stats* out);
/* The same drift check, fully asynchronous and without any allocation or host round trip: the
 * diff list is written by the compaction kernel straight into a caller-owned DEVICE buffer laid
 * out like ncd_diff_pack's (uint32 header[4] = {entries found, nshards, rows, 1 if the buffer
 * was too small} + entries).  This is the form the multi-GPU path uses: the buffer is the send
 * buffer of the all-gather, so scan and collective queue back to back on the GPU.  The call
 * returns once everything is enqueued on the ctx stream; no ncd_scan_finish is needed. */
int ncd_scan_into(ncd_ctx* ctx, const ncd_arena* ctrl, const ncd_arena* const* shards, int nshards,
                  uint32_t mode, void* dev_out, uint64_t cap_entries);

/* The same drift check for a host-side caller that scans again and again (the DriftChecker of
 * INTEGRATION.md): the diff list is delivered in pinned host memory the CONTEXT keeps, so a scan in
 * steady state allocates nothing — ncd_scan's per-scan buffers (rows x shards x 17 bytes) and its
 * ncd_diff are not used.  *entries (ordered like ncd_diff_view's) stays valid until the next scan on
 * this context; SKIP decisions are simply absent from it.  The context's list buffer starts at
 * 65 536 entries and is re-allocated, and the scan run again, only when a scan finds more. */
int ncd_scan_host(ncd_ctx* ctx, const ncd_arena* ctrl, const ncd_arena* const* shards, int nshards,
                  uint32_t mode, const ncd_entry** entries, uint64_t* n);

/* ---- exchange of diff lists between GPUs through peer memory (ncd_exchange.cu) ----------------
 * The multi-GPU drift check partitions objects by key hash (ncd_keyhash % world) and ends with
 * every rank holding every rank's diff list.  An ncd_xchg is one rank's end of that exchange:
 * a device buffer of [3 generations][world slots][header + cap_entries entries] that peers write
 * into directly over NVLink.  No reference counterpart: the reference is a single CPU process
 * (controller.go:790-831 loops over the shards); this replaces a library all-gather.
 *
 * Set-up, once:   ncd_xchg_create on every rank
 *   ranks in separate processes:  ncd_xchg_handle -> ship the 64 bytes to every rank by any
 *                                 channel -> ncd_xchg_open(all handles, rank order)
 *   ranks in one process:         ncd_xchg_buffer -> ncd_xchg_connect(all buffers, rank order;
 *                                 needs peer access between the devices, or the same device)
 * Per step:       ncd_scan_exchange (asynchronous: scan, compaction into this rank's slot, one
 *                 kernel that stores the slot into every peer's buffer, flags it, and waits for
 *                 the peers' lists of the PREVIOUS step: ranks run up to one step apart instead of
 *                 meeting in every step), then, to use the lists of THIS step: ncd_xchg_fetch (waits
 *                 for the peers' flags of this step, then copies the headers and the entries each
 *                 slot really holds to the host; the rest of host_out is left as it was), or
 *                 ncd_xchg_wait + ncd_xchg_result (device view, ordered on the ctx stream): world x
 *                 (16 + 16 * cap_entries) bytes, slot r = rank r's list in ncd_diff_pack layout.
 * Rules: every rank calls ncd_scan_exchange the same number of times; a gathered list stays
 * intact until two more ncd_scan_exchange have been enqueued on that context; a peer that does
 * not arrive within the timeout (default 10 s) makes ncd_xchg_fetch fail with NCD_E_STATE
 * instead of hanging the GPU. */
int ncd_xchg_create(ncd_ctx* ctx, int world, int rank, uint64_t cap_entries, ncd_xchg** out);
int ncd_xchg_buffer(ncd_xchg* x, void** dev_ptr, uint64_t* bytes);
int ncd_xchg_handle(ncd_xchg* x, void* handle64);
int ncd_xchg_open(ncd_xchg* x, const void* handles /* world x 64 bytes; own entry ignored */);
int ncd_xchg_connect(ncd_xchg* x, void* const* peer_bufs /* world pointers; own entry ignored */);
int ncd_scan_exchange(ncd_ctx* ctx, const ncd_arena* ctrl, const ncd_arena* const* shards, int nshards,
                      uint32_t mode, ncd_xchg* x);
int ncd_xchg_wait(ncd_xchg* x);
int ncd_xchg_result(ncd_xchg* x, const void** dev_ptr, uint64_t* bytes);
int ncd_xchg_fetch(ncd_xchg* x, void* host_out, uint64_t bytes);
int ncd_xchg_set_timeout(ncd_xchg* x, double seconds);
void ncd_xchg_destroy(ncd_xchg* x);

/* Host view of a diff list (copied device -> pinned host on first use). */
int ncd_diff_view(ncd_diff* d, const ncd_entry** entries, uint64_t* n);
/* Device pointer + count, for a caller that hands the list to a collective
 * (the all-gather of per-GPU diff lists) without a host round trip. */
int ncd_diff_device(ncd_diff* d, const void** dev_entries, uint64_t* n);
/* Enqueues, on the ctx stream, a copy of the list into a caller-owned DEVICE buffer laid out
 * as uint32 header[4] = {n, nshards, rows (low 32 bits), 0} followed by n ncd_entry: the
 * send buffer of the all-gather (e.g. an NCCL-registered tensor).  cap_entries = entries the
 * buffer can hold after the header. */
int ncd_diff_pack(ncd_diff* d, void* dev_buf, uint64_t cap_entries);
/* The sweep feed: controller indices, ascending and without
 * repeats, of the templates / workgroups that have at least one decision != NCD_SKIP — the only
 * objects a resync sweep still has to hand to templateSyncHandler / workgroupSyncHandler.  In
 * the reference every template and workgroup is re-enqueued on every 30 s resync
 * (controller.go:288-290, 301-303 with main.go:70-71) and then waits for the 50 items/s limiter
 * (controller.go:257-260) even when nothing drifted.  Writes at most cap indices, *n = the count. */
int ncd_diff_dirty_objects(ncd_diff* d, uint32_t* obj_out, uint64_t cap, uint64_t* n);
/* All codes, rows x nshards, row-major, SKIP included (tests and Lookup()). */
int ncd_diff_codes(ncd_diff* d, uint8_t* codes_out, uint64_t cap);
void ncd_diff_destroy(ncd_diff* d);

/* ------------------------------------------------------------------ orphans (row f-4)
 * Objects a shard holds that the controller cluster does not: the reverse join.  The reference
 * has no such sweep — a template deleted while the controller was down stays on the shards
 * forever, because deletion is only propagated from the delete EVENT (handleObject,
 * controller.go:196-205) and the sync handlers never look at what a shard has beyond what the
 * controller asks about.  The caller decides which shard objects it manages (the packer puts
 * only those carrying the controller-app label into a shard arena, controller_test.go:183-188).
 * Writes at most cap entries, ordered by (shard, object index in that shard's arena); *n = the
 * number found. */
typedef struct ncd_orphan {
  uint32_t shard; /* index into the shard list            */
  uint32_t obj;   /* object index in that shard's arena   */
} ncd_orphan;
int ncd_find_orphans(ncd_ctx* ctx, const ncd_arena* ctrl, const ncd_arena* const* shards, int nshards,
                     ncd_orphan* out, uint64_t cap, uint64_t* n);

/* Controller-side adoption (the part of a resync sweep the per-shard decisions do not cover):
 * adoptReferences (controller.go:647-695) gives every Secret / ConfigMap a template references on
 * the CONTROLLER cluster an ownerReference to that template unless one with the template's UID is
 * there already (isOwnedBy, controller.go:637-645: UID only).  This lists the (template,
 * dependent) references that still lack it, ordered by (obj, dep): templates the sweep feed must
 * still hand to templateSyncHandler although no shard drifted.  A reference to an object the
 * controller does not hold is not listed (it is NCD_ERR_MISSING_REF in the diff list).  Writes at
 * most cap entries; *n = the number found. */
typedef struct ncd_adoption {
  uint32_t obj; /* controller index of the template                    */
  uint32_t dep; /* j+1 = its j-th dependent reference (as in ncd_entry) */
} ncd_adoption;
int ncd_find_unadopted(ncd_ctx* ctx, const ncd_arena* ctrl, ncd_adoption* out, uint64_t cap, uint64_t* n);

/* ------------------------------------------------------------------ hashing */
/* Hashes every record of a committed arena on the GPU and copies the digests out.
 * mode NCD_MODE_HASH64 / NCD_MODE_XXH64: out is nobj x 8 bytes; NCD_MODE_SHA256: nobj x 32 bytes.
 * Recomputes by default (it is also how the hash kernels are timed: kernel_ms); with
 * | NCD_MODE_CACHED digests already valid for the arena are returned as they are (kernel_ms 0).
 * No counterpart in the reference (it never hashes, controller.go:539,600,725,795
 * call reflect.DeepEqual); this is the hash path BASELINE.json's north_star asks for. */
int ncd_hash_arena(ncd_ctx* ctx, ncd_arena* a, uint32_t mode, void* out, uint64_t out_bytes,
                   float* kernel_ms);

/* ------------------------------------------------------------------ hashing without allocation
 * No counterpart in the reference, and the one entry point here that needs no ncd_ctx.
 *
 * For callers that may not allocate GPU or pinned memory once they are running: a controller
 * sharing its GPU under a fixed memory reservation, a process whose device memory all belongs
 * to its own pool, a handler that cannot afford cudaMalloc's device-wide synchronisation or
 * an out-of-memory failure mid-reconcile.  The caller owns every buffer it passes (all DEVICE
 * pointers, on the device that is current when the call is made):
 *   dev_payload  records, each starting on a 16-byte boundary (arena payload layout); the kernels
 *                read whole 16-byte vectors, so the buffer must be readable up to the 16-byte
 *                padded end of every record (the bytes past dev_rec_len are ignored)
 *   dev_rec_off  byte offset of record i (multiple of 16), dev_rec_len its length
 *   dev_out      nrec x uint64, receives the NCD-H64 digests, identical to
 *                ncd_hash_arena(NCD_MODE_HASH64) on the same records
 *   stream       cudaStream_t to enqueue on (NULL = the default stream); the call returns
 *                once the work is enqueued
 * The working memory of the kernels is a fixed 768.25 KiB array inside the library image
 * (every GPU that loads libncdrift.so holds it until the process exits), so the call performs
 * no cudaMalloc / cudaFree / pinned allocation and calls no routine that does.
 * Limits and rules:
 *   - at most NCD_STATIC_MAX_RECORDS records per call; more is refused with NCD_E_CAPACITY
 *   - the first call on each device looks the array's device address up
 *     (cudaGetSymbolAddress); make that first call before any stream capture begins
 *   - all calls on a device share the array; a call re-initialises what it reads, but calls
 *     that overlap on the GPU (two streams, or unordered work) race and corrupt each other's
 *     digests: serialise them */
#define NCD_STATIC_MAX_RECORDS 65536u
int ncd_hash64_static(const void* dev_payload, const uint64_t* dev_rec_off, const uint32_t* dev_rec_len,
                      uint64_t nrec, uint64_t* dev_out, void* stream);

/* ------------------------------------------------------------------ synthetic input
 * Bench / test utility: fills a DEVICE_ONLY arena on the GPU with the synthetic
 * workload of SURVEY.md section 8(d).  role -1 = controller copy, s >= 0 = shard s.
 * The same generator is restated on the CPU in oracle/ so that samples can be
 * checked.  Not part of the drop-in surface. */
typedef struct ncd_synth_cfg {
  uint64_t groups;        /* templates; each owns n_secrets + n_configmaps dependents */
  uint32_t n_secrets;
  uint32_t n_configmaps;
  uint32_t tmpl_len_lo, tmpl_len_hi; /* canonical length range of a template spec (inclusive) */
  uint32_t dep_len_lo, dep_len_hi;   /* canonical length range of a dependent's data          */
  uint64_t seed;
  uint64_t drift_seed;
  uint32_t p_flip_ppm;    /* per (object, shard): one byte flipped (same length)      */
  uint32_t p_missing_ppm; /* per (object, shard): absent from the shard               */
  uint32_t p_owner_ppm;   /* per (dependent, shard): owned by a foreign template only */
  uint32_t p_rogue_ppm;   /* per (dependent, shard): no ownerRefs at all              */
  uint32_t p_len_ppm;     /* per (object, shard): shard copy one byte longer          */
  uint32_t part_mod;      /* keep only groups with keyhash(template) % part_mod == part_rem */
  uint32_t part_rem;      /* (part_mod = 0 or 1 keeps everything)                     */
  uint32_t shuffle;       /* 1 = shard copies stored in a permuted object order        */
} ncd_synth_cfg;

/* Capacities an arena needs for cfg (so the caller can size ncd_arena_create). */
int ncd_synth_sizes(ncd_ctx* ctx, const ncd_synth_cfg* cfg, int role, uint64_t* payload_bytes,
                    uint64_t* nobj, uint64_t* aux_bytes, uint64_t* nowner, uint64_t* ndep);
/* Generates and commits. */
int ncd_synth_fill(ncd_ctx* ctx, ncd_arena* a, const ncd_synth_cfg* cfg, int role);

#ifdef __cplusplus
}
#endif
#endif /* NCDRIFT_H */
