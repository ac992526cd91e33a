// drift_checker.cpp — packer, DriftChecker (Scan / Lookup) and handler replay over libncdrift.so.
//
// This is the C++ stand-in for the Go layer of INTEGRATION.md.  It decides nothing about drift:
// it lays the informer caches out as arenas (canonicalising the compared values), calls
// ncd_scan, and reads decision codes back.  The replay functions walk a template's / workgroup's
// slice of those codes in the order the reference's handlers walk the shards
// (controller.go:790-831 with :504-626; :720-745) and stop where the handler would return.
#include <stdint.h>
#include <string.h>

#include <memory>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../../include/ncdhost.h"
#include "../../../include/ncdrift.h"

namespace ncdh {
std::string canonicalise(const uint8_t* value, uint64_t len);
}

namespace {

thread_local std::string g_err;
int fail(const std::string& m) {
  g_err = m;
  return -1;
}
std::string ncd_err(const char* op) { return std::string(op) + ": " + ncd_last_error(); }

struct OwnerRef {
  std::string kind, uid;
  bool operator==(const OwnerRef& o) const { return kind == o.kind && uid == o.uid; }
};
struct CachedObject {
  uint32_t kind = 0;
  std::string key;  // "namespace/name"
  std::string name, uid;
  std::string canon;  // canonical bytes of .Spec / .Data
  std::vector<OwnerRef> owners;
  std::vector<std::string> secrets, configmaps;  // dependent names (templates)
  bool live = true;
  // where the record lives in the resident arena's payload (-1: not placed yet)
  int64_t rec_off = -1;
  uint64_t rec_cap = 0;     // padded bytes reserved at rec_off
  bool value_dirty = false; // canon changed since the device copy was written
};

std::string store_key(uint32_t kind, const std::string& key) { return std::string(1, (char)('0' + kind)) + key; }
inline uint64_t pad16(uint64_t n) { return (n + 15) & ~15ull; }

struct ArenaHandle {
  ncd_arena* a = nullptr;
  ArenaHandle() = default;
  ArenaHandle(const ArenaHandle&) = delete;
  ArenaHandle& operator=(const ArenaHandle&) = delete;
  ArenaHandle(ArenaHandle&& o) noexcept : a(o.a) { o.a = nullptr; }
  void reset() {
    if (a) ncd_arena_destroy(a);
    a = nullptr;
  }
  ~ArenaHandle() { reset(); }
};

// One cluster's informer caches (all four listers), insertion ordered, plus its resident arena.
// Informer events map onto three costs (include/ncdrift.h):
//   value of an existing object changed (UpdateFunc)   -> ncd_arena_sync_records: that record only
//   object added / deleted / re-owned (Add/DeleteFunc) -> ncd_arena_recommit: small arrays + new records
//   first scan, or the arena ran out of room           -> ncd_arena_commit of a fresh, larger arena
struct ClusterImage {
  std::vector<CachedObject> objs;
  std::unordered_map<std::string, size_t> index;
  ArenaHandle arena;
  uint64_t payload_used = 0, payload_cap = 0, nobj_cap = 0, aux_cap = 0, nowner_cap = 0, ndep_cap = 0;
  bool structural = true;  // keys / owners / dependents changed since the last commit
  uint64_t uploads_full = 0, uploads_structural = 0, uploads_values = 0;  // how each scan got its data across

  void put(CachedObject&& o) {
    const std::string k = store_key(o.kind, o.key);
    auto it = index.find(k);
    if (it == index.end()) {
      index.emplace(k, objs.size());
      objs.push_back(std::move(o));
      structural = true;
      return;
    }
    CachedObject& old = objs[it->second];
    if (old.uid != o.uid || !(old.owners == o.owners) || old.secrets != o.secrets || old.configmaps != o.configmaps)
      structural = true;
    if (old.canon != o.canon) {
      old.canon = std::move(o.canon);
      old.value_dirty = true;
    }
    old.name = std::move(o.name);
    old.uid = std::move(o.uid);
    old.owners = std::move(o.owners);
    old.secrets = std::move(o.secrets);
    old.configmaps = std::move(o.configmaps);
  }
  bool drop(uint32_t kind, const std::string& key) {
    auto it = index.find(store_key(kind, key));
    if (it == index.end()) return false;
    objs[it->second].live = false;
    index.erase(it);
    structural = true;
    return true;
  }
  const CachedObject* get(uint32_t kind, const std::string& key) const {
    auto it = index.find(store_key(kind, key));
    return it == index.end() ? nullptr : &objs[it->second];
  }

  // Brings the resident arena up to date with the image.  `order` (controller image only)
  // receives the live objects in arena order.
  // deferred: if given, a pure value sync is not issued here but left in pending_changed for the
  // caller, which pushes the rewritten records of every cluster in one ncd_arenas_sync_records
  void sync(ncd_ctx* ctx, std::vector<const CachedObject*>* order, std::vector<ClusterImage*>* deferred = nullptr) {
    // forget dead objects (their payload becomes unreferenced until the next full commit)
    if (structural) {
      std::vector<CachedObject> keep;
      for (CachedObject& o : objs)
        if (o.live) keep.push_back(std::move(o));
      objs.swap(keep);
      index.clear();
      for (size_t i = 0; i < objs.size(); ++i) index.emplace(store_key(objs[i].kind, objs[i].key), i);
    }
    if (order)
      for (const CachedObject& o : objs) order->push_back(&o);

    if (!arena.a || structural) {
      uint64_t aux = 0, nowner = 0, ndep = 0, append = 0, compact = 0;
      for (const CachedObject& o : objs) {
        aux += o.key.size() + o.uid.size();
        for (const OwnerRef& r : o.owners) aux += r.uid.size();
        nowner += o.owners.size();
        for (const auto& s : o.secrets) aux += o.key.size() + s.size();  // "namespace/" + name <= key + name
        for (const auto& s : o.configmaps) aux += o.key.size() + s.size();
        ndep += o.secrets.size() + o.configmaps.size();
        compact += pad16(o.canon.size());
        if (o.rec_off < 0 || o.value_dirty) append += pad16(o.canon.size());
      }
      const bool fits = arena.a && objs.size() <= nobj_cap && aux <= aux_cap && nowner <= nowner_cap &&
                        ndep <= ndep_cap && payload_used + append <= payload_cap;
      if (!fits) {
        // a fresh arena with room to grow; every record is placed again, compactly
        arena.reset();
        payload_cap = compact + compact / 2 + (64 << 10);
        nobj_cap = objs.size() + objs.size() / 2 + 16;
        aux_cap = aux + aux / 2 + 1024;
        nowner_cap = nowner + nowner / 2 + 16;
        ndep_cap = ndep + ndep / 2 + 16;
        if (ncd_arena_create(ctx, payload_cap, nobj_cap, aux_cap, nowner_cap, ndep_cap, NCD_ARENA_HOST_PINNED,
                             &arena.a) != NCD_OK)
          throw std::runtime_error(ncd_err("ncd_arena_create"));
        payload_used = 0;
        for (CachedObject& o : objs) o.rec_off = -1;
      }
      ncd_arena_view v;
      if (ncd_arena_host_view(arena.a, &v) != NCD_OK) throw std::runtime_error(ncd_err("ncd_arena_host_view"));
      const uint64_t dirty_from = payload_used;
      write_all(v);
      const uint64_t n = objs.size();
      const int rc = fits ? ncd_arena_recommit(arena.a, n, payload_used, aux_used, own_used, dep_used, dirty_from)
                          : ncd_arena_commit(arena.a, n, payload_used, aux_used, own_used, dep_used);
      if (rc != NCD_OK) throw std::runtime_error(ncd_err(fits ? "ncd_arena_recommit" : "ncd_arena_commit"));
      ++(fits ? uploads_structural : uploads_full);
      structural = false;
      return;
    }

    // same objects, some values changed: rewrite those records and push only them
    std::vector<uint32_t> changed;
    uint64_t append = 0;
    for (const CachedObject& o : objs)
      if (o.value_dirty && pad16(o.canon.size()) > o.rec_cap) append += pad16(o.canon.size());
    if (payload_used + append > payload_cap) {  // no room to append: start over with a larger arena
      structural = true;
      arena.reset();
      if (order) order->clear();
      sync(ctx, order, deferred);
      return;
    }
    ncd_arena_view v;
    if (ncd_arena_host_view(arena.a, &v) != NCD_OK) throw std::runtime_error(ncd_err("ncd_arena_host_view"));
    for (size_t i = 0; i < objs.size(); ++i) {
      CachedObject& o = objs[i];
      if (!o.value_dirty) continue;
      place(v, o, /*force_append=*/false);
      v.rec_off[i] = (uint64_t)o.rec_off;
      v.rec_len[i] = (uint32_t)o.canon.size();
      changed.push_back((uint32_t)i);
    }
    if (!changed.empty()) {
      ++uploads_values;
      if (deferred) {
        pending_changed.swap(changed);
        deferred->push_back(this);
      } else if (ncd_arena_sync_records(arena.a, changed.size(), changed.data(), payload_used) != NCD_OK) {
        throw std::runtime_error(ncd_err("ncd_arena_sync_records"));
      }
    }
  }
  std::vector<uint32_t> pending_changed;  // objects whose records a deferred value sync still has to push

 private:
  uint64_t aux_used = 0, own_used = 0, dep_used = 0;

  // Writes o's record into the pinned payload: where it already is if it fits, else at the end.
  void place(ncd_arena_view& v, CachedObject& o, bool force_append) {
    const uint64_t need = pad16(o.canon.size());
    if (o.rec_off < 0 || need > o.rec_cap || force_append) {
      o.rec_off = (int64_t)payload_used;
      o.rec_cap = need;
      payload_used += need;
    }
    memcpy(v.payload + o.rec_off, o.canon.data(), o.canon.size());
    memset(v.payload + o.rec_off + o.canon.size(), 0, need - o.canon.size());
    o.value_dirty = false;
  }

  // Every per-object array from scratch (arena layout: ncdrift.h "arenas"); records that are
  // already on the device and did not change keep their place, the others are appended.
  void write_all(ncd_arena_view& v) {
    uint64_t ax = 0, oi = 0, di = 0, i = 0;
    auto put_aux = [&](const std::string& s, uint32_t* o, uint32_t* l) {
      memcpy(v.aux + ax, s.data(), s.size());
      *o = (uint32_t)ax;
      *l = (uint32_t)s.size();
      ax += s.size();
    };
    for (CachedObject& o : objs) {
      if (o.rec_off < 0 || o.value_dirty) place(v, o, /*force_append=*/true);  // only appended payload is re-uploaded
      v.rec_off[i] = (uint64_t)o.rec_off;
      v.rec_len[i] = (uint32_t)o.canon.size();
      v.kind[i] = (uint8_t)o.kind;
      put_aux(o.key, &v.key_off[i], &v.key_len[i]);
      put_aux(o.uid, &v.uid_off[i], &v.uid_len[i]);
      v.own_start[i] = (uint32_t)oi;
      for (const OwnerRef& r : o.owners) {
        v.own_is_template[oi] = r.kind == "NexusAlgorithmTemplate";  // the Kind test of controller.go:490
        put_aux(r.uid, &v.own_uid_off[oi], &v.own_uid_len[oi]);
        ++oi;
      }
      v.dep_start[i] = (uint32_t)di;
      const std::string ns = o.key.substr(0, o.key.find('/') + 1);
      for (int pass = 0; pass < 2; ++pass)  // secrets first, then configmaps (:817 before :825)
        for (const std::string& name : pass == 0 ? o.secrets : o.configmaps) {
          v.dep_kind[di] = pass == 0 ? NCD_KIND_SECRET : NCD_KIND_CONFIGMAP;
          put_aux(ns + name, &v.dep_key_off[di], &v.dep_key_len[di]);
          ++di;
        }
      ++i;
    }
    v.own_start[i] = (uint32_t)oi;
    v.dep_start[i] = (uint32_t)di;
    aux_used = ax;
    own_used = oi;
    dep_used = di;
  }
};

}  // namespace

struct ncdh_checker {
  ncd_ctx* ctx = nullptr;
  uint32_t mode = 0;
  ClusterImage controller;
  std::vector<ClusterImage> shards;
  // snapshot of the last Scan
  bool scanned = false;
  std::vector<uint8_t> codes;                          // rows x nshards
  std::unordered_map<std::string, uint64_t> first_row;  // store_key -> first row of a template / workgroup
  std::unordered_map<std::string, CachedObject> snap;   // the controller objects the rows were made from
  std::vector<std::string> dirty;                       // sweep feed of the last Scan: "<type>\t<ns/name>"
  uint64_t rows = 0, ndiff = 0;

  ClusterImage* image(int cluster) {
    if (cluster == -1) return &controller;
    if (cluster >= 0 && (size_t)cluster < shards.size()) return &shards[cluster];
    return nullptr;
  }
};

namespace {

struct Action {
  int shard;
  const char* client;  // "nexus" (templates, workgroups) or "kube" (secrets, configmaps)
  const char* verb;
  const char* resource;
  std::string name;
};

int64_t render(const std::vector<Action>& acts, const std::string& error, char* out, uint64_t cap) {
  std::string s;
  for (const Action& a : acts)
    s += std::to_string(a.shard) + "\t" + a.client + "\t" + a.verb + "\t" + a.resource + "\t" + a.name + "\n";
  if (!error.empty()) s += "error\t" + error + "\n";
  if (out && s.size() <= cap) memcpy(out, s.data(), s.size());
  return (int64_t)s.size();
}

}  // namespace

extern "C" {

const char* ncdh_last_error(void) { return g_err.c_str(); }

int64_t ncdh_canon(const uint8_t* value, uint64_t value_len, uint8_t* out, uint64_t cap) {
  try {
    const std::string c = ncdh::canonicalise(value, value_len);
    if (out && c.size() <= cap) memcpy(out, c.data(), c.size());
    return (int64_t)c.size();
  } catch (const std::exception& e) {
    return fail(e.what());
  }
}

int ncdh_checker_create(int device, uint32_t mode, int nshards, ncdh_checker** out) {
  if (!out || nshards < 0) return fail("ncdh_checker_create: bad argument");
  std::unique_ptr<ncdh_checker> c(new ncdh_checker);
  if (ncd_init(device, &c->ctx) != NCD_OK) return fail(ncd_err("ncd_init"));  // no GPU: an error, not a fallback
  c->mode = mode;
  c->shards.resize(nshards);
  *out = c.release();
  return 0;
}

void ncdh_checker_destroy(ncdh_checker* c) {
  if (!c) return;
  c->controller.arena.reset();  // arenas before the context they live in
  for (ClusterImage& sh : c->shards) sh.arena.reset();
  if (c->ctx) ncd_destroy(c->ctx);
  delete c;
}

int ncdh_cluster_add(ncdh_checker* c, int cluster, uint32_t kind, const char* ns, const char* name, const char* uid,
                     const uint8_t* value, uint64_t value_len, const char* const* owner_kinds,
                     const char* const* owner_uids, int nowners, const char* const* secret_names, int nsecrets,
                     const char* const* configmap_names, int nconfigmaps) {
  if (!c || !ns || !name || !uid || (!value && value_len)) return fail("ncdh_cluster_add: null argument");
  ClusterImage* img = c->image(cluster);
  if (!img) return fail("ncdh_cluster_add: no such cluster");
  if (kind > NCD_KIND_CONFIGMAP) return fail("ncdh_cluster_add: unknown kind");
  try {
    CachedObject o;
    o.kind = kind;
    o.key = std::string(ns) + "/" + name;  // cache.ObjectName.String()
    o.name = name;
    o.uid = uid;
    o.canon = ncdh::canonicalise(value, value_len);
    for (int i = 0; i < nowners; ++i) o.owners.push_back({owner_kinds[i], owner_uids[i]});
    for (int i = 0; i < nsecrets; ++i) o.secrets.push_back(secret_names[i]);
    for (int i = 0; i < nconfigmaps; ++i) o.configmaps.push_back(configmap_names[i]);
    img->put(std::move(o));
  } catch (const std::exception& e) {
    return fail(std::string("ncdh_cluster_add: ") + e.what());
  }
  return 0;
}

int ncdh_cluster_remove(ncdh_checker* c, int cluster, uint32_t kind, const char* ns, const char* name) {
  ClusterImage* img = c ? c->image(cluster) : nullptr;
  if (!img || !ns || !name) return fail("ncdh_cluster_remove: bad argument");
  return img->drop(kind, std::string(ns) + "/" + name) ? 0 : 1;
}

int ncdh_scan(ncdh_checker* c) {
  if (!c) return fail("ncdh_scan: null checker");
  try {
    std::vector<const CachedObject*> order;
    std::vector<ClusterImage*> deferred;
    c->controller.sync(c->ctx, &order, &deferred);
    std::vector<const ncd_arena*> ptrs;
    for (ClusterImage& sh : c->shards) {
      sh.sync(c->ctx, nullptr, &deferred);
      ptrs.push_back(sh.arena.a);
    }
    if (!deferred.empty()) {  // the value rewrites of every cluster: one round trip
      std::vector<ncd_arena*> as;
      std::vector<uint64_t> ns, pbs;
      std::vector<const uint32_t*> idx;
      for (ClusterImage* img : deferred) {
        as.push_back(img->arena.a);
        ns.push_back(img->pending_changed.size());
        idx.push_back(img->pending_changed.data());
        pbs.push_back(img->payload_used);
      }
      const int rc = ncd_arenas_sync_records((int)as.size(), as.data(), ns.data(), idx.data(), pbs.data());
      for (ClusterImage* img : deferred) img->pending_changed.clear();
      if (rc != NCD_OK) return fail(ncd_err("ncd_arenas_sync_records"));
    }
    ncd_diff* d = nullptr;
    if (ncd_scan(c->ctx, c->controller.arena.a, ptrs.data(), (int)ptrs.size(), c->mode, &d) != NCD_OK)
      return fail(ncd_err("ncd_scan"));
    ncd_scan_stats st;
    ncd_scan_last_stats(c->ctx, &st);
    c->rows = st.rows;
    c->ndiff = st.ndiff;
    c->codes.assign(st.rows * ptrs.size(), 0);
    int rc = ncd_diff_codes(d, c->codes.data(), c->codes.size());
    std::vector<uint32_t> dirty_idx;
    if (rc == NCD_OK) {
      uint64_t nd = 0;
      rc = ncd_diff_dirty_objects(d, nullptr, 0, &nd);
      dirty_idx.resize(nd);
      if (rc == NCD_OK && nd) rc = ncd_diff_dirty_objects(d, dirty_idx.data(), nd, &nd);
    }
    ncd_diff_destroy(d);
    if (rc != NCD_OK) return fail(ncd_err("ncd_diff_codes"));
    // row numbering: arena order, one row per template + one per dependent, one per workgroup
    c->first_row.clear();
    c->snap.clear();
    uint64_t row = 0;
    for (const CachedObject* o : order) {
      if (o->kind == NCD_KIND_TEMPLATE || o->kind == NCD_KIND_WORKGROUP) {
        c->first_row[store_key(o->kind, o->key)] = row;
        c->snap[store_key(o->kind, o->key)] = *o;
        row += 1 + (o->kind == NCD_KIND_TEMPLATE ? o->secrets.size() + o->configmaps.size() : 0);
      }
    }
    if (row != st.rows) return fail("ncdh_scan: row count disagrees with the library");
    c->dirty.clear();
    for (uint32_t i : dirty_idx) {
      if (i >= order.size()) return fail("ncdh_scan: dirty object index out of range");
      c->dirty.push_back(std::string(order[i]->kind == NCD_KIND_TEMPLATE ? "template" : "workgroup") + "\t" + order[i]->key);
    }
    c->scanned = true;
  } catch (const std::exception& e) {
    return fail(std::string("ncdh_scan: ") + e.what());
  }
  return 0;
}

int ncdh_lookup(ncdh_checker* c, uint32_t kind, const char* ns, const char* name, int dep, int shard, uint32_t* code) {
  if (!c || !ns || !name || !code) return fail("ncdh_lookup: null argument");
  if (!c->scanned) return fail("ncdh_lookup: no scan yet");
  if (shard < 0 || (size_t)shard >= c->shards.size()) return fail("ncdh_lookup: no such shard");
  const std::string k = store_key(kind, std::string(ns) + "/" + name);
  auto it = c->first_row.find(k);
  if (it == c->first_row.end()) return 1;
  const CachedObject& o = c->snap[k];
  const size_t ndeps = o.kind == NCD_KIND_TEMPLATE ? o.secrets.size() + o.configmaps.size() : 0;
  if (dep < 0 || (size_t)dep > ndeps) return fail("ncdh_lookup: no such dependent");
  *code = c->codes[(it->second + dep) * c->shards.size() + shard];
  return 0;
}

int64_t ncdh_plan_template(ncdh_checker* c, const char* ns, const char* name, char* out, uint64_t cap) {
  if (!c || !ns || !name) return fail("ncdh_plan_template: null argument");
  if (!c->scanned) return fail("ncdh_plan_template: no scan yet");
  std::vector<Action> acts;
  std::string error;
  const std::string k = store_key(NCD_KIND_TEMPLATE, std::string(ns) + "/" + name);
  auto it = c->snap.find(k);
  if (it == c->snap.end()) return render(acts, error, out, cap);  // :769-772: gone from the cache, nil
  const CachedObject& t = it->second;
  const size_t S = c->shards.size(), nsec = t.secrets.size(), ndeps = nsec + t.configmaps.size();
  const uint64_t row0 = c->first_row[k];
  auto dep_name = [&](size_t j) { return j < nsec ? t.secrets[j] : t.configmaps[j - nsec]; };
  auto dep_res = [&](size_t j) { return j < nsec ? "secrets" : "configmaps"; };
  // adoptReferences (:783 -> :647-676) fails the handler before any shard when a referenced
  // object is missing on the controller; ERR_MISSING_REF does not depend on the shard
  if (S > 0) {
    for (size_t j = 0; j < ndeps; ++j)
      if (c->codes[(row0 + 1 + j) * S] == NCD_ERR_MISSING_REF) {
        error = std::string("missing ") + dep_res(j) + " " + dep_name(j);
        return render(acts, error, out, cap);
      }
  } else {
    for (size_t j = 0; j < ndeps; ++j)
      if (!c->controller.get(j < nsec ? NCD_KIND_SECRET : NCD_KIND_CONFIGMAP, std::string(ns) + "/" + dep_name(j))) {
        error = std::string("missing ") + dep_res(j) + " " + dep_name(j);
        return render(acts, error, out, cap);
      }
  }
  for (size_t s = 0; s < S && error.empty(); ++s) {  // :790
    const uint32_t tc = c->codes[row0 * S + s];
    if (tc & NCD_UPDATE_DATA) acts.push_back({(int)s, "nexus", "update", "NexusAlgorithmTemplates", t.name});  // :795-797
    if (tc == NCD_CREATE) acts.push_back({(int)s, "nexus", "create", "NexusAlgorithmTemplates", t.name});      // :805-808
    for (size_t j = 0; j < ndeps; ++j) {  // :817 then :825
      const uint32_t code = c->codes[(row0 + 1 + j) * S + s];
      if (code == NCD_ERR_ROGUE) {  // :529-535 / :591-597
        error = "Resource \"" + dep_name(j) + "\" already exists and is not managed by any Algorithm";
        break;
      }
      if (code == NCD_CREATE) {  // :518-520 / :580-582
        acts.push_back({(int)s, "kube", "create", dep_res(j), dep_name(j)});
        continue;
      }
      if (code & NCD_UPDATE_DATA) acts.push_back({(int)s, "kube", "update-data", dep_res(j), dep_name(j)});    // :539-541 / :600-602
      if (code & NCD_UPDATE_OWNER) acts.push_back({(int)s, "kube", "update-owner", dep_res(j), dep_name(j)});  // :550-552 / :612-614
    }
  }
  return render(acts, error, out, cap);
}

int64_t ncdh_plan_workgroup(ncdh_checker* c, const char* ns, const char* name, char* out, uint64_t cap) {
  if (!c || !ns || !name) return fail("ncdh_plan_workgroup: null argument");
  if (!c->scanned) return fail("ncdh_plan_workgroup: no scan yet");
  std::vector<Action> acts;
  const std::string k = store_key(NCD_KIND_WORKGROUP, std::string(ns) + "/" + name);
  auto it = c->snap.find(k);
  if (it == c->snap.end()) return render(acts, "", out, cap);  // :705-708
  const size_t S = c->shards.size();
  const uint64_t row0 = c->first_row[k];
  for (size_t s = 0; s < S; ++s) {  // :720
    const uint32_t code = c->codes[row0 * S + s];
    if (code & NCD_UPDATE_DATA) acts.push_back({(int)s, "nexus", "update", "NexusAlgorithmWorkgroups", it->second.name});  // :725-727
    if (code == NCD_CREATE) acts.push_back({(int)s, "nexus", "create", "NexusAlgorithmWorkgroups", it->second.name});      // :735-738
  }
  return render(acts, "", out, cap);
}

int64_t ncdh_sweep_queue(ncdh_checker* c, char* out, uint64_t cap) {
  if (!c || !c->scanned) return fail("ncdh_sweep_queue: no scan yet");
  std::string s;
  for (const std::string& line : c->dirty) s += line + "\n";
  if (out && s.size() <= cap) memcpy(out, s.data(), s.size());
  return (int64_t)s.size();
}

int ncdh_upload_counts(ncdh_checker* c, int cluster, uint64_t* full, uint64_t* structural, uint64_t* values) {
  ClusterImage* img = c ? c->image(cluster) : nullptr;
  if (!img) return fail("ncdh_upload_counts: no such cluster");
  if (full) *full = img->uploads_full;
  if (structural) *structural = img->uploads_structural;
  if (values) *values = img->uploads_values;
  return 0;
}

int ncdh_last_counts(ncdh_checker* c, uint64_t* ndiff, uint64_t* decisions) {
  if (!c || !c->scanned) return fail("ncdh_last_counts: no scan yet");
  if (ndiff) *ndiff = c->ndiff;
  if (decisions) *decisions = c->rows * c->shards.size();
  return 0;
}

}  // extern "C"
